"""ear::dsp::DelayBuffer on the device against the compiled reference's golden vector
(tests/golden/delay.npz) and a numpy shift at the size of BASELINE config 4's direct bus.
A delay is a copy: bit-exact."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_delay_golden_in_ragged_calls(eb):
    g = np.load(os.path.join(GOLDEN, "delay.npz"))
    x, y, delay = g["x"], g["y"], int(g["delay"])
    for sizes in ([x.shape[1]], [128] * 7 + [104], [1, 2, 3, 250, 255, 256, 233], [100, 900]):
        d = eb.DelayBuffer(x.shape[0], delay)
        assert d.get_delay() == delay
        out, t = [], 0
        for n in sizes:
            out.append(d.process(x[:, t:t + n]))
            t += n
        assert t == x.shape[1]
        assert np.array_equal(np.concatenate(out, axis=1), y)


@pytest.mark.parametrize("delay", [0, 1, 255, 4096])
def test_delay_large(eb, delay):
    rng = np.random.default_rng(3)
    C, T = 24, 1 << 20
    x = rng.normal(0, 0.25, (C, T)).astype(np.float32)
    d = eb.DelayBuffer(C, delay)
    y = np.concatenate([d.process(x[:, :T // 2 + 3]), d.process(x[:, T // 2 + 3:])], axis=1)
    exp = np.concatenate([np.zeros((C, delay), np.float32), x], axis=1)[:, :T]
    assert np.array_equal(y, exp)


def test_delay_no_channels_and_empty_calls(eb):
    d = eb.DelayBuffer(2, 5)
    assert d.process(np.zeros((2, 0), np.float32)).shape == (2, 0)
    x = np.arange(12, dtype=np.float32).reshape(2, 6)
    assert np.array_equal(d.process(x), np.concatenate([np.zeros((2, 5), np.float32), x[:, :1]], 1))
