// Decorrelation filter design (setup-time, host): one 512-tap random-phase all-pass FIR
// per loudspeaker.  Replaces designDecorrelatorBasic / designDecorrelator /
// designDecorrelators / decorrelatorCompensationDelay (ref src/decorrelate.cpp:31-97).
// The filters are *data* for the K4 convolver bank; their design is not on the
// per-sample path.
#include <cmath>
#include <complex>
#include <random>
#include <vector>

#include "capi_common.hpp"

using namespace earb;

namespace {
  const int kDecorrelatorSize = 512;

  // unnormalised inverse DFT (sign +) of a power-of-two length sequence, in double
  void ifft_double(std::vector<std::complex<double>> &a) {
    const size_t n = a.size();
    for (size_t i = 1, j = 0; i < n; i++) {
      size_t bit = n >> 1;
      for (; j & bit; bit >>= 1) j ^= bit;
      j ^= bit;
      if (i < j) std::swap(a[i], a[j]);
    }
    const double pi = 3.141592653589793238462643383279502884;
    for (size_t len = 2; len <= n; len <<= 1)
      for (size_t i = 0; i < n; i += len)
        for (size_t k = 0; k < len / 2; k++) {
          const double ang = 2.0 * pi * (double)k / (double)len;
          const std::complex<double> w(std::cos(ang), std::sin(ang));
          const std::complex<double> u = a[i + k], v = a[i + k + len / 2] * w;
          a[i + k] = u + v;
          a[i + k + len / 2] = u - v;
        }
  }

  std::vector<double> design_basic(int decorrelator_id, int size) {
    const double pi = 3.141592653589793238462643383279502884;
    std::mt19937 engine(decorrelator_id);
    std::vector<std::complex<double>> freq(size);
    freq[0] = {1.0, 0.0};
    for (int i = 0; i < size / 2 - 1; i++) {
      const double u = engine() / 4294967296.0;  // [0, 1)
      const double phase = 2.0 * pi * u;
      freq[i + 1] = {std::cos(phase), std::sin(phase)};
    }
    freq[size / 2] = {1.0, 0.0};
    for (int i = 0; i < size / 2; i++) freq[size / 2 + i] = std::conj(freq[size / 2 - i]);
    ifft_double(freq);
    std::vector<double> out(size);
    for (int i = 0; i < size; i++) out[i] = freq[i].real() / size;
    return out;
  }
}  // namespace

extern "C" {

int earb200_decorrelator_size(void) { return kDecorrelatorSize; }
int earb200_decorrelator_compensation_delay(void) { return (kDecorrelatorSize - 1) / 2; }

int earb200_design_decorrelators(const earb200_layout *layout, float *float_out, double *double_out) {
  return guarded([&] {
    if (!layout) throw Error(INVALID_ARGUMENT, "null argument");
    const auto &ch = layout->layout.channels;
    for (size_t c = 0; c < ch.size(); c++) {
      // the seed is the rank of the channel name among all channel names (ref :56-68)
      int id = 0;
      for (const auto &other : ch)
        if (other.name < ch[c].name) id++;
      const std::vector<double> taps = design_basic(id, kDecorrelatorSize);
      for (int i = 0; i < kDecorrelatorSize; i++) {
        if (double_out) double_out[c * kDecorrelatorSize + i] = taps[i];
        if (float_out) float_out[c * kDecorrelatorSize + i] = (float)taps[i];
      }
    }
  });
}

}  // extern "C"
