"""K4 parity: BlockConvolver bank vs the compiled libear reference (KISS float FFT).

Mirrors the scenarios of tests/block_convolver_tests.cpp :208-356 (multi-partition
filters, crossfades, fades to/from silence, different num_blocks, zero and NULL input
blocks) with the reference convolver itself as the oracle, plus an exact float64
direct convolution.  Tolerance: 1e-6 of peak (the reference's own test uses 1e-6
absolute on unit-scale signals, block_convolver_tests.cpp:77,192).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def peak_err(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)


def run_script(eb, ref, B, nb, filters, script, n_channels=1, chunk=None):
    """script: list of ('set'|'xfade', fid|None) or ('run', x[n_blocks*B]) or ('silence', n_blocks)."""
    r = [ref.Convolver(B, nb) for _ in range(n_channels)]
    g = eb.BlockConvolver(B, n_channels, nb)
    for taps in filters:
        ids = [rc.add_filter(taps)[0] for rc in r]
        gid, gnb = g.add_filter(taps)
        assert gid == ids[0] and gnb == -(-len(taps) // B)
    out_r, out_g = [], []
    for op, arg in script:
        if op == "set":
            [rc.set_filter(arg) for rc in r]
            g.set_filter(arg)
        elif op == "xfade":
            [rc.crossfade_filter(arg) for rc in r]
            g.crossfade_filter(arg)
        elif op == "silence":
            out_r.append(np.stack([rc.process(None, arg) for rc in r]))
            out_g.append(g.process(None, arg))
        else:
            x = arg.reshape(n_channels, -1)
            out_r.append(np.stack([rc.process(x[c]) for c, rc in enumerate(r)]))
            step = chunk or x.shape[1]
            out_g.append(np.concatenate([g.process(x[:, t:t + step]) for t in range(0, x.shape[1], step)], 1))
    return np.concatenate(out_g, 1), np.concatenate(out_r, 1)


@pytest.mark.parametrize("B,nb,taps", [(512, 1, 512), (128, 4, 512), (64, 2, 100), (32, 7, 200), (16, 3, 40),
                                       (1024, 1, 700), (256, 2, 256)])
def test_single_filter(eb, ref, B, nb, taps):
    rng = np.random.default_rng(B + nb)
    h = rng.normal(size=taps).astype(np.float32)
    x = rng.normal(0, 0.25, size=B * 40).astype(np.float32)
    out, exp = run_script(eb, ref, B, nb, [h], [("set", 0), ("run", x)])
    assert peak_err(out, exp) < 1e-6
    direct = np.convolve(x.astype(np.float64), h.astype(np.float64))[:len(x)]
    assert peak_err(out[0], direct) < 1e-6


def test_streaming_calls_match_one_call(eb, ref):
    rng = np.random.default_rng(1)
    B, nb = 64, 3
    h = rng.normal(size=150).astype(np.float32)
    x = rng.normal(size=B * 50).astype(np.float32)
    for chunk in [B, 2 * B, 7 * B]:
        out, exp = run_script(eb, ref, B, nb, [h], [("set", 0), ("run", x)], chunk=chunk)
        assert peak_err(out, exp) < 1e-6, chunk


def test_crossfades_and_filter_changes(eb, ref):
    rng = np.random.default_rng(2)
    B, nb = 64, 3
    f = [rng.normal(size=n).astype(np.float32) for n in (150, 64, 192, 10)]
    blk = lambda k: rng.normal(size=B * k).astype(np.float32)
    script = [("set", 0), ("run", blk(5)), ("xfade", 1), ("run", blk(6)), ("xfade", 2), ("run", blk(1)),
              ("xfade", 0), ("run", blk(8)), ("xfade", None), ("run", blk(6)), ("xfade", 3), ("run", blk(6)),
              ("set", 1), ("run", blk(4)), ("xfade", 2), ("run", blk(1)), ("set", 0), ("run", blk(5)),
              ("set", None), ("run", blk(3)), ("set", 2), ("silence", 4), ("run", blk(3)), ("silence", 5)]
    for chunk in [None, B]:
        rng = np.random.default_rng(2)
        out, exp = run_script(eb, ref, B, nb, f, script, chunk=chunk)
        assert peak_err(out, exp) < 1e-6, chunk


def test_too_many_blocks_rejected(eb):
    g = eb.BlockConvolver(64, 1, 2)
    fid, nb = g.add_filter(np.ones(200, dtype=np.float32))
    assert nb == 4
    with pytest.raises(eb.InvalidArgument, match="too many blocks"):
        g.set_filter(fid)
    with pytest.raises(eb.InvalidArgument):
        eb.BlockConvolver(100, 1, 1)  # not a power of two


def test_decorrelator_bank_9_10_3(eb, ref):
    # BASELINE config 3 shape (24 channels, B = 512, 512-tap decorrelators) on 30 blocks
    rng = np.random.default_rng(3)
    filt = eb.design_decorrelators("9+10+3")
    assert np.array_equal(filt, ref.decorrelators("9+10+3"))
    C, B, K = 24, 512, 30
    x = np.clip(rng.normal(0, 0.25, size=(C, B * K)), -1, 1).astype(np.float32)
    g = eb.BlockConvolver(B, C, 1)
    for c in range(C):
        fid, _ = g.add_filter(filt[c])
        g.set_filter(fid, channel=c)
    out = g.process(x)
    for c in range(C):
        r = ref.Convolver(B, 1)
        fid, _ = r.add_filter(filt[c])
        r.set_filter(fid)
        assert peak_err(out[c], r.process(x[c])) < 1e-6, c
    # all-pass property at full length scale: output energy == input energy (after the tail)
    e_in, e_out = (x.astype(np.float64) ** 2).sum(1), (out.astype(np.float64) ** 2).sum(1)
    assert np.abs(e_out / e_in - 1).max() < 0.05


def test_fast_path_streaming_odd_counts_and_crossfades(eb, ref):
    """B = 512, one partition: the warp-per-pair kernel (convolver_fast_kernels.cu) handles calls
    without a filter change in flight, the generic kernel the rest; state (last tail, last block)
    must carry across both, for odd and even block counts."""
    rng = np.random.default_rng(4)
    B = 512
    f = [rng.normal(size=n).astype(np.float32) for n in (512, 300, 17)]
    blk = lambda k: rng.normal(0, 0.25, size=B * k).astype(np.float32)
    script = [("set", 0), ("run", blk(1)), ("run", blk(2)), ("run", blk(7)), ("run", blk(70)), ("xfade", 1),
              ("run", blk(5)), ("run", blk(33)), ("silence", 2), ("run", blk(3)), ("set", 2), ("run", blk(64)),
              ("xfade", None), ("run", blk(4)), ("run", blk(9)), ("set", 0), ("run", blk(301))]
    out, exp = run_script(eb, ref, B, 1, f, script)
    assert peak_err(out, exp) < 1e-6


def test_fast_path_matches_generic_kernel(eb, monkeypatch):
    rng = np.random.default_rng(5)
    C, B, K = 5, 512, 257
    x = rng.normal(0, 0.25, size=(C, B * K)).astype(np.float32)
    filt = rng.normal(size=(C, 512)).astype(np.float32)
    outs = []
    for path in ("fast", "generic"):
        monkeypatch.setenv("EARB200_CONV_PATH", path)
        g = eb.BlockConvolver(B, C, 1)
        for c in range(C - 1):          # the last channel has no filter: silence out
            fid, _ = g.add_filter(filt[c])
            g.set_filter(fid, channel=c)
        outs.append(np.concatenate([g.process(x[:, :B * 100]), g.process(x[:, B * 100:])], axis=1))
    assert peak_err(outs[0], outs[1]) < 1e-6
    assert np.abs(outs[0][C - 1]).max() == 0.0
    direct = np.stack([np.convolve(x[c].astype(np.float64), filt[c].astype(np.float64))[:B * K] for c in range(C - 1)])
    assert peak_err(outs[0][:C - 1], direct) < 1e-6


@pytest.mark.parametrize("B,nb", [(128, 4), (64, 8), (256, 2), (16, 32), (32, 5), (256, 1)])
def test_fast_path_for_any_partitioning_up_to_512_taps(eb, ref, B, nb):
    """B * num_blocks <= 512: the call runs on the fast kernel as ONE 512-tap partition over
    super-blocks of 512 samples; libear's state (num_blocks input blocks, last_tail of B samples) is
    what is carried, so calls may alternate with the generic kernel -- block counts that are not whole
    super-blocks, filter changes (set: the old filter's tail is still due), crossfades, silence."""
    rng = np.random.default_rng(100 * B + nb)
    q = 512 // B
    n_ch = 3
    L = B * nb
    f = [rng.normal(size=n).astype(np.float32) for n in (L, max(1, L // 2 + 3), 7, L)]
    blk = lambda k: rng.normal(0, 0.25, size=n_ch * B * k).astype(np.float32)
    script = [("set", 0), ("run", blk(4 * q)), ("run", blk(q)), ("run", blk(3 * q + 1)), ("run", blk(17 * q)),
              ("set", 1), ("run", blk(6 * q)), ("run", blk(2 * q)), ("xfade", 2), ("run", blk(5 * q + nb + 2)),
              ("run", blk(8 * q)), ("silence", 3), ("run", blk(2 * q)), ("set", None), ("run", blk(2 * q)),
              ("set", 3), ("run", blk(1)), ("run", blk(9 * q)), ("xfade", None), ("run", blk(3 * q)),
              ("set", 0), ("run", blk(40 * q + q // 2))]
    out, exp = run_script(eb, ref, B, nb, f, script, n_channels=n_ch)
    assert peak_err(out, exp) < 1e-6


def test_fast_path_other_partitionings_match_generic_kernel(eb, monkeypatch):
    rng = np.random.default_rng(6)
    C, B, nb, K = 4, 128, 4, 4 * 130
    x = rng.normal(0, 0.25, size=(C, B * K)).astype(np.float32)
    filt = rng.normal(size=(C, 512)).astype(np.float32)
    outs = []
    for path in ("fast", "generic"):
        monkeypatch.setenv("EARB200_CONV_PATH", path)
        g = eb.BlockConvolver(B, C, nb)
        for c in range(C - 1):          # the last channel has no filter: silence out
            fid, _ = g.add_filter(filt[c, :512 - 100 * c])
            g.set_filter(fid, channel=c)
        outs.append(np.concatenate([g.process(x[:, :B * 200]), g.process(x[:, B * 200:])], axis=1))
    assert peak_err(outs[0], outs[1]) < 1e-6
    assert np.abs(outs[0][C - 1]).max() == 0.0
    direct = np.stack([np.convolve(x[c].astype(np.float64), filt[c, :512 - 100 * c].astype(np.float64))[:B * K]
                       for c in range(C - 1)])
    assert peak_err(outs[0][:C - 1], direct) < 1e-6


@pytest.mark.parametrize("K", [96, 97, 101, 196, 293, 600])
def test_wide_kernel_long_calls(eb, ref, K, monkeypatch):
    """With EARB200_CONV_PATH=wide calls of at least 96 super-blocks run on the 2048-point kernel
    (by default only calls that fill the GPU with long runs do: units of 3,072 samples; the
    halo of a run and the super-blocks that do not fill a unit on 1024-point pairs): every residue
    of the block count modulo 6, odd counts, several runs per channel, carried tails across calls
    that alternate with the 1024-point kernel (short calls) -- against libear."""
    monkeypatch.setenv("EARB200_CONV_PATH", "wide")
    rng = np.random.default_rng(K)
    B, n_ch = 512, 3
    f = [rng.normal(size=n).astype(np.float32) for n in (512, 200)]
    blk = lambda k: rng.normal(0, 0.25, size=n_ch * B * k).astype(np.float32)
    script = [("set", 0), ("run", blk(3)), ("run", blk(K)), ("run", blk(5)), ("run", blk(K + 2)), ("set", 1),
              ("run", blk(K + 3)), ("xfade", 0), ("run", blk(K + 4))]
    out, exp = run_script(eb, ref, B, 1, f, script, n_channels=n_ch)
    assert peak_err(out, exp) < 1e-6


@pytest.mark.parametrize("B,nb", [(128, 4), (64, 5), (16, 32)])
def test_wide_kernel_other_partitionings(eb, ref, B, nb, monkeypatch):
    monkeypatch.setenv("EARB200_CONV_PATH", "wide")
    rng = np.random.default_rng(7 * B + nb)
    q, n_ch = 512 // B, 2
    f = [rng.normal(size=n).astype(np.float32) for n in (B * nb, 9)]
    blk = lambda k: rng.normal(0, 0.25, size=n_ch * B * k).astype(np.float32)
    script = [("set", 0), ("run", blk(100 * q)), ("run", blk(2 * q)), ("run", blk(131 * q + 1)), ("set", 1),
              ("run", blk(99 * q + q // 2)), ("xfade", 0), ("run", blk(120 * q))]
    out, exp = run_script(eb, ref, B, nb, f, script, n_channels=n_ch)
    assert peak_err(out, exp) < 1e-6


def test_wide_kernel_matches_pair_kernel(eb, monkeypatch):
    rng = np.random.default_rng(8)
    C, B, K = 4, 512, 1000
    x = rng.normal(0, 0.25, size=(C, B * K)).astype(np.float32)
    filt = rng.normal(size=(C, 512)).astype(np.float32)
    outs = []
    for path in ("wide", "pair"):
        monkeypatch.setenv("EARB200_CONV_PATH", path)
        g = eb.BlockConvolver(B, C, 1)
        for c in range(C - 1):          # the last channel has no filter: silence out
            fid, _ = g.add_filter(filt[c])
            g.set_filter(fid, channel=c)
        outs.append(np.concatenate([g.process(x[:, :B * 400]), g.process(x[:, B * 400:])], axis=1))
    assert peak_err(outs[0], outs[1]) < 1e-6
    assert np.abs(outs[0][C - 1]).max() == 0.0
    direct = np.stack([np.convolve(x[c].astype(np.float64), filt[c].astype(np.float64))[:B * K] for c in range(C - 1)])
    assert peak_err(outs[0][:C - 1], direct) < 1e-6


def test_wide_kernel_one_wave_plan(eb, ref, monkeypatch):
    """Default kernel choice on a call that fills the GPU about once (24 channels x 1,807 blocks): the
    2048-point kernel with the run length of the wave model (runs of 5 units, one wave of CTAs),
    not the 16-unit runs of long calls -- against the pair kernel on every channel and against
    libear on two of them."""
    rng = np.random.default_rng(9)
    C, B, K = 24, 512, 1807
    x = rng.normal(0, 0.25, size=(C, B * K)).astype(np.float32)
    filt = rng.normal(size=(C, 512)).astype(np.float32)
    outs = []
    for path in (None, "pair"):
        if path:
            monkeypatch.setenv("EARB200_CONV_PATH", path)
        else:
            monkeypatch.delenv("EARB200_CONV_PATH", raising=False)
        g = eb.BlockConvolver(B, C, 1)
        for c in range(C):
            fid, _ = g.add_filter(filt[c])
            g.set_filter(fid, channel=c)
        outs.append(np.concatenate([g.process(x[:, :B * 3]), g.process(x[:, B * 3:])], axis=1))
    assert peak_err(outs[0], outs[1]) < 1e-6
    for c in (0, C - 1):
        r = ref.Convolver(B, 1)
        fid, _ = r.add_filter(filt[c])
        r.set_filter(fid)
        assert peak_err(outs[0][c], r.process(x[c])) < 1e-6, c
