"""The FFT plug-in seam (ref include/ear/fft.hpp:27-62, src/fft_kiss.cpp:61-88)
as batched device transforms, against numpy's double-precision rfft / irfft: unpacked n/2+1 bins,
unnormalised inverse.  Tolerance: 1e-6 of the output peak (float32 butterflies)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n_fft", [32, 64, 128, 512, 1024, 2048, 8192])
def test_fft_forward_reverse(eb, n_fft):
    rng = np.random.default_rng(n_fft)
    batch = 37
    x = rng.normal(0, 0.25, (batch, n_fft)).astype(np.float32)
    fft = eb.FFT(n_fft)
    X = fft.forward(x)
    ref = np.fft.rfft(x.astype(np.float64), axis=1)
    assert np.abs(X - ref).max() / np.abs(ref).max() < 1e-6
    back = fft.reverse(ref.astype(np.complex64))
    exp = np.fft.irfft(ref, n=n_fft, axis=1) * n_fft
    assert np.abs(back - exp).max() / np.abs(exp).max() < 1e-6
    assert fft.launch_count == 2


def test_fft_known_answers(eb):
    # tests/fft_tests.cpp style: an impulse has a flat spectrum, a cosine one bin
    n = 1024
    fft = eb.FFT(n)
    x = np.zeros((2, n), np.float32)
    x[0, 0] = 1.0
    x[1] = np.cos(2 * np.pi * 7 * np.arange(n) / n)
    X = fft.forward(x)
    assert np.abs(X[0] - 1.0).max() < 1e-6
    assert abs(X[1, 7] - n / 2) < 1e-2 and np.abs(np.delete(X[1], 7)).max() < 1e-2


def test_fft_rejects_bad_sizes(eb):
    for n in (0, 16, 48, 1000, 16384):
        with pytest.raises(eb.InvalidArgument):
            eb.FFT(n)
