"""the facade's VariableBlockSizeAdapter (host-side buffering around a fixed-block
process function, libear_b200/include/ear_b200/dsp.hpp) against libear's own adapter
(ref src/dsp/variable_block_size_impl.cpp:44-81) on ragged call sizes -- the sizes of libear's
test (ref tests/variable_block_size_tests.cpp:38: 0, 512, 1024, 300, 500) and harder ones.  The
inner process function is stateful (it stamps the index of the inner call into its output), so a
block processed at the wrong moment or twice shows.  Bit-exact.  CPU-only: the adapter does no
device work."""
import subprocess

import numpy as np
import pytest

from oracle import ref
from test_facade import build_exe

CASES = [
    (512, 2, 4, [0, 512, 1024, 300, 500]),               # libear's own test
    (512, 2, 4, [1, 511, 512, 513, 1, 0, 2047, 3]),
    (128, 3, 3, [1 + (7 * k) % 301 for k in range(40)]),
    (64, 1, 5, [64] * 9),
    (256, 4, 2, [1000, 24, 0, 0, 1, 255, 256, 257, 5000]),
    (16, 2, 2, [3] * 50 + [16, 32, 5]),
]


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libear_ref.so not built")
@pytest.mark.parametrize("block,n_in,n_out,sizes", CASES)
def test_adapter_matches_libear(block, n_in, n_out, sizes):
    T = int(sum(sizes))
    rng = np.random.default_rng(block + n_in + len(sizes))
    x = rng.standard_normal((n_in, T)).astype(np.float32)
    exp, delay = ref.vbs_process(block, x, n_out, sizes)
    assert delay == block
    r = subprocess.run([build_exe(), "vbs", str(block), str(n_in), str(n_out)] + [str(s) for s in sizes],
                       input=x.tobytes(), capture_output=True)
    assert r.returncode == 0, r.stderr[-2000:]
    got = np.frombuffer(r.stdout, dtype=np.float32).reshape(n_out, T)
    assert np.array_equal(got.view(np.uint32), exp.view(np.uint32))
    # and the property libear's own test asserts: the output is the processed input, block_size late
    first = exp[:, :min(block, T)]
    assert (first == 0).all()
