// C ABI: block convolver bank (K4).  The host keeps, per channel, exactly the filter
// queue libear's BlockConvolver keeps (ref src/dsp/block_convolver_impl.hpp:157-170,
// block_convolver_impl.cpp:64-122) and turns it into per-block filter tables for the
// kernel; audio state (input history, last tail) stays on the device.
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdlib>
#include <string>
#include <memory>
#include <vector>

#include "capi_common.hpp"
#include "kernels.hpp"

using namespace earb;

namespace {
  // in-place iterative radix-2 FFT in double (setup-time: filter spectra)
  void fft_double(std::vector<std::complex<double>> &a) {
    const size_t n = a.size();
    for (size_t i = 1, j = 0; i < n; i++) {
      size_t bit = n >> 1;
      for (; j & bit; bit >>= 1) j ^= bit;
      j ^= bit;
      if (i < j) std::swap(a[i], a[j]);
    }
    const double pi = 3.141592653589793238462643383279502884;
    for (size_t len = 2; len <= n; len <<= 1) {
      for (size_t i = 0; i < n; i += len)
        for (size_t k = 0; k < len / 2; k++) {
          double ang = -2.0 * pi * (double)k / (double)len;
          std::complex<double> w(std::cos(ang), std::sin(ang));
          std::complex<double> u = a[i + k], v = a[i + k + len / 2] * w;
          a[i + k] = u + v;
          a[i + k + len / 2] = u - v;
        }
    }
  }

  struct ChannelQueue {
    std::vector<int> queue;  // nb + 1 filter ids, -1 = none; filters(i) = queue[(ofs+i) % (nb+1)]
    size_t ofs = 0;
    std::vector<unsigned char> hist_faded;  // nb flags of the blocks before the next call
    int &at(size_t i) { return queue[(ofs + i) % queue.size()]; }
    bool uniform() const {
      for (int f : queue)
        if (f != queue[0]) return false;
      return true;
    }
  };
}  // namespace

struct earb200_convolver {
  int device = 0;
  int n_sms = 0;
  int B = 0, nb = 0, C = 0;
  std::vector<ChannelQueue> chan;
  std::vector<int> filter_blocks;    // partitions per filter
  std::vector<float> filt_host;      // [n_filters][nb][B+1][2]
  bool filt_dirty = false;
  DeviceBuffer<float> d_filt, d_tw2, d_tw_fast;
  bool use_fast = false;  // B * nb <= 512 taps (EARB200_CONV_PATH=generic turns it off)
  // fast path with B != 512: the whole filter as one 512-tap partition (1024-point spectrum, 513
  // bins) and its taps (for the carried tail), per filter
  std::vector<float> filt_fast_host, taps_host;
  DeviceBuffer<float> d_filt_fast, d_taps;
  // fast path, long calls: the 2048-point spectrum of the whole filter (all 2,048 bins), per filter
  // (EARB200_CONV_PATH=pair keeps long calls on the 1024-point kernel)
  bool use_wide = false;
  int wide_mode = 1;  // EARB200_CONV_PATH=wide: 2 (every call long enough for one run; tests)
  std::vector<float> filt_wide_host;
  DeviceBuffer<float> d_filt_wide;
  // B != 512: the carried last_tail was made with a filter that set_filter has replaced since; the
  // next block must run on the generic kernel (which adds the carried tail) before the fast one,
  // whose start-of-call overlap is recomputed with the CURRENT filter, may take over
  bool tail_stale = false;
  bool force_generic = false;  // set around the head of a split call
  DeviceBuffer<float> d_hist[2], d_tail[2];
  int cur = 0;  // which of the ping-pong state buffers is current
  DeviceBuffer<short> d_ftab;
  DeviceBuffer<int> d_uniform;
  DeviceBuffer<unsigned char> d_faded;
  DeviceBuffer<float> d_in, d_out;
  int64_t launches = 0;
  KernelTimer timer;
};

// `add`: {direct bus, its row pitch, delay, delay memory} to be joined in the store (fast path
// only).  Returns true when it was (the caller otherwise adds the delayed bus itself).
bool earb_conv_process_device(earb200_convolver &h, int64_t K, const float *d_in, int64_t in_stride,
                              float *d_out, int64_t out_stride, cudaStream_t stream, const float *add,
                              int64_t add_stride, int64_t add_delay, const float *add_mem) {
    if (K < 0) throw Error(INVALID_ARGUMENT, "negative block count");
    if (K == 0) return false;
    if (!d_out) throw Error(INVALID_ARGUMENT, "null output buffer");
    use_device(h.device);
    const int nb = h.nb, C = h.C;
    bool fused = false;
    // A filter change travels through a channel's queue for nb + 1 blocks
    // (ref block_convolver_impl.cpp:114-122); only those blocks need the generic kernel (crossfades,
    // per-partition filter tables).  A long call that starts with a change in flight is therefore cut
    // in two: the head on the generic kernel, the rest -- uniform again -- on the fast one, instead of
    // the whole call at a tenth of the speed.  (Not with a fused add: the caller falls back as a whole.)
    // The same cut serves the other reasons a block cannot go to the fast kernel: with B != 512 it
    // works on super-blocks of q = 512 / B blocks (the remainder goes first), and after set_filter
    // the first block must add the tail the OLD filter left (tail_stale).
    const int64_t q = h.use_fast ? 512 / h.B : 1;
    int64_t head = 0;
    if (h.use_fast && d_in && !h.force_generic) {
      bool uniform = true;
      for (const ChannelQueue &cq : h.chan) uniform = uniform && cq.uniform();
      if (!uniform)
        head = (int64_t)nb + 1;
      else if (h.B != 512 && h.tail_stale)
        head = 1;
      if (head < K) head += (K - head) % q;
      head = std::min(head, K);
      if (head > 0 && head < K && !add) {
        h.force_generic = true;
        try {
          earb_conv_process_device(h, head, d_in, in_stride, d_out, out_stride, stream, nullptr, 0, 0, nullptr);
        } catch (...) {
          h.force_generic = false;
          throw;
        }
        h.force_generic = false;
        earb_conv_process_device(h, K - head, d_in + head * h.B, in_stride, d_out + head * h.B, out_stride, stream,
                                 nullptr, 0, 0, nullptr);
        return false;
      }
    }
    const bool fast_ok = h.use_fast && d_in && !h.force_generic && head == 0;
    if (h.filt_dirty) {
      h.d_filt.reserve(std::max<size_t>(h.filt_host.size(), 2));
      if (!h.filt_host.empty())
        EARB_CUDA(cudaMemcpyAsync(h.d_filt.ptr, h.filt_host.data(), sizeof(float) * h.filt_host.size(),
                                  cudaMemcpyHostToDevice, stream));
      if (!h.filt_fast_host.empty()) {
        h.d_filt_fast.reserve(h.filt_fast_host.size());
        h.d_taps.reserve(h.taps_host.size());
        EARB_CUDA(cudaMemcpyAsync(h.d_filt_fast.ptr, h.filt_fast_host.data(), sizeof(float) * h.filt_fast_host.size(),
                                  cudaMemcpyHostToDevice, stream));
        EARB_CUDA(cudaMemcpyAsync(h.d_taps.ptr, h.taps_host.data(), sizeof(float) * h.taps_host.size(),
                                  cudaMemcpyHostToDevice, stream));
      }
      if (!h.filt_wide_host.empty()) {
        h.d_filt_wide.reserve(h.filt_wide_host.size());
        EARB_CUDA(cudaMemcpyAsync(h.d_filt_wide.ptr, h.filt_wide_host.data(), sizeof(float) * h.filt_wide_host.size(),
                                  cudaMemcpyHostToDevice, stream));
      }
      h.filt_dirty = false;
    }

    // simulate the filter queues until every channel's queue is uniform again
    // (process() rotates the queue and copies filters(1) into filters(0), :114-122)
    std::vector<std::vector<short>> tabs(C);   // per channel [k][nb][2]
    std::vector<std::vector<unsigned char>> fades(C);
    int Kg = 0;
    for (int c = 0; c < C; c++) {
      ChannelQueue &q = h.chan[c];
      fades[c] = q.hist_faded;
      int64_t k = 0;
      while (k < K && !q.uniform()) {
        fades[c].push_back(q.at(1) != q.at(0));
        for (int i = 0; i < nb; i++) {
          tabs[c].push_back((short)q.at(i + 1));
          tabs[c].push_back((short)q.at(i));
        }
        q.ofs = (q.ofs + q.queue.size() - 1) % q.queue.size();
        q.at(0) = q.at(1);
        k++;
      }
      Kg = std::max<int>(Kg, (int)k);
    }
    // pad every channel's tables to Kg blocks (uniform blocks: filter f for old and new)
    std::vector<short> ftab((size_t)C * Kg * nb * 2);
    std::vector<unsigned char> faded((size_t)C * (nb + Kg));
    std::vector<int> uniform(C);
    for (int c = 0; c < C; c++) {
      ChannelQueue &q = h.chan[c];
      const int f = q.queue[0];  // meaningful once uniform (k >= Kg)
      uniform[c] = f;
      const int have = (int)(tabs[c].size() / ((size_t)nb * 2));
      for (int k = 0; k < Kg; k++)
        for (int i = 0; i < nb; i++)
          for (int o = 0; o < 2; o++)
            ftab[(((size_t)c * Kg + k) * nb + i) * 2 + o] =
                k < have ? tabs[c][((size_t)k * nb + i) * 2 + o] : (short)f;
      for (int j = 0; j < nb + Kg; j++)
        faded[(size_t)c * (nb + Kg) + j] = j < (int)fades[c].size() ? fades[c][j] : 0;
      // flags of the last nb blocks become the history flags for the next call
      std::vector<unsigned char> all = fades[c];
      all.resize((size_t)nb + (size_t)K, 0);
      q.hist_faded.assign(all.end() - nb, all.end());
    }
    h.d_ftab.reserve(std::max<size_t>(ftab.size(), 1));
    h.d_faded.reserve(faded.size());
    h.d_uniform.reserve(C);
    if (!ftab.empty())
      EARB_CUDA(cudaMemcpyAsync(h.d_ftab.ptr, ftab.data(), sizeof(short) * ftab.size(),
                                cudaMemcpyHostToDevice, stream));
    EARB_CUDA(cudaMemcpyAsync(h.d_faded.ptr, faded.data(), faded.size(), cudaMemcpyHostToDevice, stream));
    EARB_CUDA(cudaMemcpyAsync(h.d_uniform.ptr, uniform.data(), sizeof(int) * C, cudaMemcpyHostToDevice,
                              stream));
    // the tables above live in pageable host memory: the async copies have staged them
    // before returning, so the vectors may go out of scope

    ConvLaunch L;
    L.B = h.B;
    L.nb = nb;
    L.C = C;
    L.Kg = Kg;
    L.K = (long)K;
    // run length: long enough to amortise the (nb + 1)-block halo, short enough to fill the GPU
    int R = 32;
    while (R > 4 && (K + R - 1) / R * C < 4L * h.n_sms) R /= 2;
    R = std::max(R, std::min<int>(4 * nb, 64));
    L.R = R;
    L.in = d_in;
    L.in_stride = (long)in_stride;
    L.out = d_out;
    L.out_stride = (long)out_stride;
    L.hist = h.d_hist[h.cur].ptr;
    L.hist_new = h.d_hist[1 - h.cur].ptr;
    L.tail = h.d_tail[h.cur].ptr;
    L.tail_new = h.d_tail[1 - h.cur].ptr;
    L.filt = h.d_filt.ptr;
    L.ftab = h.d_ftab.ptr;
    L.uniform = h.d_uniform.ptr;
    L.faded = h.d_faded.ptr;
    L.tw2 = h.d_tw2.ptr;
    h.timer.start(stream);
    if (fast_ok && Kg == 0) {
      if (add) {
        L.add = add, L.add_stride = (long)add_stride, L.add_delay = (long)add_delay, L.add_mem = add_mem;
        fused = true;
      }
      L.filt_fast = h.d_filt_fast.ptr;
      L.taps = h.d_taps.ptr;
      L.filt_wide = h.use_wide ? h.d_filt_wide.ptr : nullptr;
      L.wide = h.wide_mode;
      launch_convolver_fast(L, h.d_uniform.ptr, h.d_tw_fast.ptr, h.n_sms, stream);
      launch_conv_history(L, stream);
      if (h.B != 512) {
        launch_conv_generic_tail(L, h.d_uniform.ptr, stream);
        h.launches++;
      }
    } else {
      launch_convolver(L, stream);
      h.tail_stale = false;  // the block(s) just processed left a tail made with the current filters
    }
    h.timer.stop(stream);
    EARB_CUDA(cudaGetLastError());
    h.cur = 1 - h.cur;
    h.launches += 2;
    return fused;
}

namespace {
  void conv_process_device(earb200_convolver &h, int64_t K, const float *d_in, int64_t in_stride,
                           float *d_out, int64_t out_stride, cudaStream_t stream) {
    earb_conv_process_device(h, K, d_in, in_stride, d_out, out_stride, stream, nullptr, 0, 0, nullptr);
  }
}  // namespace

int64_t earb200_convolver_block_size_internal(const earb200_convolver *conv) { return conv->B; }

extern "C" {

int earb200_convolver_create(int64_t block_size, int n_channels, int64_t num_blocks, int device,
                             earb200_convolver **out) {
  return guarded([&] {
    if (!out) throw Error(INVALID_ARGUMENT, "null argument");
    if (block_size < 16 || block_size > 4096 || (block_size & (block_size - 1)))
      throw Error(INVALID_ARGUMENT, "block size must be a power of two in [16, 4096]");
    if (n_channels < 1 || n_channels > 65535) throw Error(INVALID_ARGUMENT, "invalid channel count");
    if (num_blocks < 1 || num_blocks > 64) throw Error(INVALID_ARGUMENT, "num_blocks must be in [1, 64]");
    if (conv_smem_bytes((int)block_size, (int)num_blocks) > 220 * 1024)
      throw Error(INVALID_ARGUMENT, "block_size * num_blocks too large for the device convolver");
    use_device(device);
    auto h = std::make_unique<earb200_convolver>();
    h->device = device;
    cudaDeviceProp prop;
    EARB_CUDA(cudaGetDeviceProperties(&prop, device));
    h->n_sms = prop.multiProcessorCount;
    h->B = (int)block_size;
    h->nb = (int)num_blocks;
    h->C = n_channels;
    h->chan.resize(n_channels);
    for (auto &q : h->chan) {
      q.queue.assign(h->nb + 1, -1);
      q.hist_faded.assign(h->nb, 0);
    }
    const size_t B = block_size;
    std::vector<float> tw(4 * B);
    const double pi = 3.141592653589793238462643383279502884;
    for (size_t k = 0; k < 2 * B; k++) {
      double ang = -2.0 * pi * (double)k / (double)(2 * B);
      tw[2 * k] = (float)std::cos(ang);
      tw[2 * k + 1] = (float)std::sin(ang);
    }
    EARB_CUDA(cudaMemcpy(h->d_tw2.reserve(tw.size()), tw.data(), sizeof(float) * tw.size(),
                         cudaMemcpyHostToDevice));
    for (int i = 0; i < 2; i++) {
      EARB_CUDA(cudaMemset(h->d_hist[i].reserve((size_t)n_channels * h->nb * B), 0,
                           sizeof(float) * n_channels * h->nb * B));
      EARB_CUDA(cudaMemset(h->d_tail[i].reserve((size_t)n_channels * B), 0, sizeof(float) * n_channels * B));
    }
    const char *path = std::getenv("EARB200_CONV_PATH");
    h->use_fast = conv_fast_supported(h->B, h->nb) && !(path && std::string(path) == "generic");
    h->use_wide = h->use_fast && !(path && std::string(path) == "pair");
    h->wide_mode = path && std::string(path) == "wide" ? 2 : 1;
    if (h->use_fast) {
      std::vector<float> twf(conv_fast_tw_floats());
      conv_fast_build_tw(twf.data());
      EARB_CUDA(cudaMemcpy(h->d_tw_fast.reserve(twf.size()), twf.data(), sizeof(float) * twf.size(),
                           cudaMemcpyHostToDevice));
    }
    h->d_filt.reserve(2);
    *out = h.release();
  });
}
void earb200_convolver_destroy(earb200_convolver *conv) { delete conv; }
int64_t earb200_convolver_launch_count(const earb200_convolver *conv) {
  return conv ? conv->launches : 0;
}

int earb200_convolver_set_profiling(earb200_convolver *conv, int enable) {
  return guarded([&] {
    if (!conv) throw Error(INVALID_ARGUMENT, "null argument");
    conv->timer.enabled = enable != 0;
  });
}
int earb200_convolver_kernel_time(earb200_convolver *conv, double *ms, int64_t *launches) {
  return guarded([&] {
    if (!conv || !ms) throw Error(INVALID_ARGUMENT, "null argument");
    use_device(conv->device);
    *ms = conv->timer.collect(launches);
  });
}

int earb200_convolver_add_filter(earb200_convolver *conv, int64_t n_taps, const float *taps,
                                 int *filter_id, int *num_blocks) {
  return guarded([&] {
    if (!conv || !filter_id || n_taps < 0 || (n_taps > 0 && !taps))
      throw Error(INVALID_ARGUMENT, "null argument");
    earb200_convolver &h = *conv;
    const size_t B = h.B;
    const int parts = (int)((n_taps + (int64_t)B - 1) / (int64_t)B);  // ref :21-36
    const size_t per_filter = (size_t)h.nb * (B + 1) * 2;
    const size_t base = h.filt_host.size();
    h.filt_host.resize(base + per_filter, 0.0f);
    for (int p = 0; p < parts && p < h.nb; p++) {
      std::vector<std::complex<double>> td(2 * B);
      const size_t len = std::min<size_t>(B, (size_t)n_taps - p * B);
      for (size_t n = 0; n < len; n++) td[n] = (double)taps[p * B + n];
      fft_double(td);
      for (size_t k = 0; k <= B; k++) {
        h.filt_host[base + ((size_t)p * (B + 1) + k) * 2] = (float)td[k].real();
        h.filt_host[base + ((size_t)p * (B + 1) + k) * 2 + 1] = (float)td[k].imag();
      }
    }
    if (h.use_fast && h.B != 512) {
      // the same taps as one partition of 512 (truncated to nb * B like the partitions above)
      const size_t L = (size_t)h.nb * B, kept = std::min<size_t>((size_t)n_taps, L);
      const size_t tbase = h.taps_host.size(), fbase = h.filt_fast_host.size();
      h.taps_host.resize(tbase + L, 0.0f);
      std::vector<std::complex<double>> td(1024);
      for (size_t n = 0; n < kept; n++) {
        h.taps_host[tbase + n] = taps[n];
        td[n] = (double)taps[n];
      }
      fft_double(td);
      h.filt_fast_host.resize(fbase + 2 * 513);
      for (size_t k = 0; k <= 512; k++) {
        h.filt_fast_host[fbase + 2 * k] = (float)td[k].real();
        h.filt_fast_host[fbase + 2 * k + 1] = (float)td[k].imag();
      }
    }
    if (h.use_wide) {
      // the same taps (truncated to nb * B like the partitions above) as ONE 2048-point spectrum
      const size_t kept = std::min<size_t>((size_t)n_taps, (size_t)h.nb * B), fbase = h.filt_wide_host.size();
      std::vector<std::complex<double>> td(2048);
      for (size_t n = 0; n < kept; n++) td[n] = (double)taps[n];
      fft_double(td);
      h.filt_wide_host.resize(fbase + 2 * 2048);
      for (size_t k = 0; k < 2048; k++) {
        h.filt_wide_host[fbase + 2 * k] = (float)td[k].real();
        h.filt_wide_host[fbase + 2 * k + 1] = (float)td[k].imag();
      }
    }
    h.filter_blocks.push_back(parts);
    h.filt_dirty = true;
    *filter_id = (int)h.filter_blocks.size() - 1;
    if (num_blocks) *num_blocks = parts;
  });
}

static void check_filter(earb200_convolver &h, int channel, int filter_id) {
  if (channel >= h.C) throw Error(INVALID_ARGUMENT, "channel index out of range");
  if (filter_id >= (int)h.filter_blocks.size()) throw Error(INVALID_ARGUMENT, "unknown filter id");
  if (filter_id >= 0 && h.filter_blocks[filter_id] > h.nb)
    throw Error(INVALID_ARGUMENT, "too many blocks in given Filter");  // ref :94-95
}

int earb200_convolver_set_filter(earb200_convolver *conv, int channel, int filter_id) {
  return guarded([&] {
    if (!conv) throw Error(INVALID_ARGUMENT, "null argument");
    check_filter(*conv, channel, filter_id);
    const int f = filter_id < 0 ? -1 : filter_id;
    for (int c = 0; c < conv->C; c++)
      if (channel < 0 || c == channel)
        for (int &e : conv->chan[c].queue) e = f;  // ref :78-83
    conv->tail_stale = true;
  });
}

int earb200_convolver_crossfade_filter(earb200_convolver *conv, int channel, int filter_id) {
  return guarded([&] {
    if (!conv) throw Error(INVALID_ARGUMENT, "null argument");
    check_filter(*conv, channel, filter_id);
    const int f = filter_id < 0 ? -1 : filter_id;
    for (int c = 0; c < conv->C; c++)
      if (channel < 0 || c == channel) conv->chan[c].at(0) = f;  // ref :71-76
  });
}

int earb200_convolver_process_device(earb200_convolver *conv, int64_t n_blocks, const float *d_in,
                                     int64_t in_stride, float *d_out, int64_t out_stride,
                                     void *stream) {
  return guarded([&] {
    if (!conv) throw Error(INVALID_ARGUMENT, "null argument");
    conv_process_device(*conv, n_blocks, d_in, in_stride, d_out, out_stride, (cudaStream_t)stream);
  });
}

int earb200_convolver_process_host(earb200_convolver *conv, int64_t n_blocks, const float *in,
                                   int64_t in_stride, float *out, int64_t out_stride) {
  return guarded([&] {
    if (!conv) throw Error(INVALID_ARGUMENT, "null argument");
    if (n_blocks < 0) throw Error(INVALID_ARGUMENT, "negative block count");
    if (n_blocks == 0) return;
    if (!out) throw Error(INVALID_ARGUMENT, "null output buffer");
    earb200_convolver &h = *conv;
    use_device(h.device);
    const int64_t T = n_blocks * h.B;
    float *d_in = nullptr;
    if (in) {
      d_in = h.d_in.reserve((size_t)T * h.C);
      EARB_CUDA(cudaMemcpy2DAsync(d_in, sizeof(float) * T, in, sizeof(float) * in_stride,
                                  sizeof(float) * T, h.C, cudaMemcpyHostToDevice, 0));
    }
    float *d_out = h.d_out.reserve((size_t)T * h.C);
    conv_process_device(h, n_blocks, d_in, T, d_out, T, 0);
    EARB_CUDA(cudaMemcpy2DAsync(out, sizeof(float) * out_stride, d_out, sizeof(float) * T,
                                sizeof(float) * T, h.C, cudaMemcpyDeviceToHost, 0));
    EARB_CUDA(cudaStreamSynchronize(0));
  });
}

}  // extern "C"
