// libear-b200 C++14 facade, part 4: the DSP classes with libear's signatures
// (ref include/ear/dsp/gain_interpolator.hpp, block_convolver.hpp, delay_buffer.hpp,
// variable_block_size.hpp, ptr_adapter.hpp, include/ear/fft.hpp) on top of the C ABI.
// Audio buffers are the caller's host `float**` rows, as in libear; every process() call copies
// to the device, runs the kernel and copies back, so call with large blocks.  Applications that
// keep audio in HBM use the *_device entry points of the C ABI through handle().
#pragma once
#include <complex>
#include <cstring>

#include "layout.hpp"

namespace ear {

  // ------------------------------------------------------------------ FFT plug-in
  // The abstract interface is libear's (ref include/ear/fft.hpp:14-67).
  class FFTWorkBuf {
   public:
    virtual ~FFTWorkBuf() {}
  };
  template <typename Real>
  class FFTPlan {
   public:
    using Complex = std::complex<Real>;
    virtual void transform_forward(Real *input, Complex *output, FFTWorkBuf &workbuf) const = 0;
    virtual void transform_reverse(Complex *input, Real *output, FFTWorkBuf &workbuf) const = 0;
    virtual std::unique_ptr<FFTWorkBuf> alloc_workbuf() const = 0;
    virtual ~FFTPlan() {}
  };
  template <typename Real>
  class FFTImpl {
   public:
    virtual std::shared_ptr<FFTPlan<Real>> plan(size_t n_fft) const = 0;
    virtual ~FFTImpl() {}
  };

  namespace b200 {
    // FFTPlan on the device; transform_*_batch are the efficient entry points
    class DeviceFFTPlan : public FFTPlan<float> {
     public:
      explicit DeviceFFTPlan(size_t n_fft) {
        earb200_fft *h = nullptr;
        check(earb200_fft_create((int64_t)n_fft, default_device(), &h));
        _h.reset(h);
      }
      void transform_forward(float *input, Complex *output, FFTWorkBuf &) const override {
        transform_forward_batch(1, input, output);
      }
      void transform_reverse(Complex *input, float *output, FFTWorkBuf &) const override {
        transform_reverse_batch(1, input, output);
      }
      void transform_forward_batch(size_t batch, const float *input, Complex *output) const {
        check(earb200_fft_forward_host(_h.get(), (int64_t)batch, input, reinterpret_cast<float *>(output)));
      }
      void transform_reverse_batch(size_t batch, const Complex *input, float *output) const {
        check(earb200_fft_reverse_host(_h.get(), (int64_t)batch, reinterpret_cast<const float *>(input), output));
      }
      std::unique_ptr<FFTWorkBuf> alloc_workbuf() const override {
        return std::unique_ptr<FFTWorkBuf>(new FFTWorkBuf());  // scratch lives in the handle
      }

     private:
      struct Deleter {
        void operator()(earb200_fft *h) const { earb200_fft_destroy(h); }
      };
      std::unique_ptr<earb200_fft, Deleter> _h;
    };
    class DeviceFFT : public FFTImpl<float> {
     public:
      std::shared_ptr<FFTPlan<float>> plan(size_t n_fft) const override {
        return std::make_shared<DeviceFFTPlan>(n_fft);
      }
    };
  }  // namespace b200

  /// the device FFT implementation
  template <typename Real>
  FFTImpl<Real> &get_fft_b200();
  template <>
  inline FFTImpl<float> &get_fft_b200<float>() {
    static b200::DeviceFFT fft;
    return fft;
  }
  /// libear's name for "the FFT that is always available" (ref include/ear/fft.hpp:66-67);
  /// here it is the device implementation -- there is no CPU FFT in this library.
  template <typename Real>
  FFTImpl<Real> &get_fft_kiss() {
    return get_fft_b200<Real>();
  }

  namespace dsp {
    using SampleIndex = long int;

    namespace detail {
      // stages host `float**` rows for the C ABI's (base, stride) form: rows that already lie
      // at a constant stride are used in place, anything else is packed
      struct Rows {
        const float *base = nullptr;
        int64_t stride = 0;
        std::vector<float> packed;
        void gather(const float *const *rows, size_t n_rows, size_t n) {
          if (n_rows == 0) return;
          bool strided = true;
          const std::ptrdiff_t step = n_rows > 1 ? rows[1] - rows[0] : (std::ptrdiff_t)n;
          for (size_t c = 1; c < n_rows && strided; c++) strided = rows[c] - rows[c - 1] == step;
          if (strided && step >= (std::ptrdiff_t)n) {
            base = rows[0];
            stride = (int64_t)step;
            return;
          }
          packed.resize(n_rows * n);
          for (size_t c = 0; c < n_rows; c++) std::memcpy(packed.data() + c * n, rows[c], sizeof(float) * n);
          base = packed.data();
          stride = (int64_t)n;
        }
      };
      struct OutRows {
        float *base = nullptr;
        int64_t stride = 0;
        std::vector<float> packed;
        bool in_place = false;
        void prepare(float *const *rows, size_t n_rows, size_t n) {
          if (n_rows == 0) return;
          in_place = true;
          const std::ptrdiff_t step = n_rows > 1 ? rows[1] - rows[0] : (std::ptrdiff_t)n;
          for (size_t c = 1; c < n_rows && in_place; c++) in_place = rows[c] - rows[c - 1] == step;
          in_place = in_place && step >= (std::ptrdiff_t)n;
          if (in_place) {
            base = rows[0];
            stride = (int64_t)step;
          } else {
            packed.resize(n_rows * n);
            base = packed.data();
            stride = (int64_t)n;
          }
        }
        void scatter(float *const *rows, size_t n_rows, size_t n) const {
          if (in_place) return;
          for (size_t c = 0; c < n_rows; c++) std::memcpy(rows[c], packed.data() + c * n, sizeof(float) * n);
        }
      };
    }  // namespace detail

    // --------------------------------------------------------- interpolation types
    // ref include/ear/dsp/gain_interpolator.hpp:133-300.  The arithmetic of apply_interp /
    // apply_constant runs in the render kernel; the types keep their names, their Point type
    // and constant_interp, and describe how a Point maps onto the [n_in][n_out] gain matrix.
    template <typename PointT>
    struct InterpType {
      using Point = PointT;
      static bool constant_interp(const Point &a, const Point &b) { return a == b; }
    };
    struct LinearInterpSingle : public InterpType<float> {
      static size_t n_in(const Point &) { return 1; }
      static size_t n_out(const Point &) { return 1; }
      static void flatten(const Point &p, float *dst) { dst[0] = p; }
    };
    struct LinearInterpVector : public InterpType<std::vector<float>> {
      static size_t n_in(const Point &) { return 1; }
      static size_t n_out(const Point &p) { return p.size(); }
      static void flatten(const Point &p, float *dst) { std::copy(p.begin(), p.end(), dst); }
    };
    struct LinearInterpMatrix : public InterpType<std::vector<std::vector<float>>> {
      static size_t n_in(const Point &p) { return p.size(); }
      static size_t n_out(const Point &p) { return p.empty() ? 0 : p[0].size(); }
      static void flatten(const Point &p, float *dst) {
        const size_t M = n_out(p);
        for (size_t i = 0; i < p.size(); i++) {
          if (p[i].size() != M) throw invalid_argument("interpolation point rows differ in size");
          std::copy(p[i].begin(), p[i].end(), dst + i * M);
        }
      }
    };

    /// Gain interpolator (ref include/ear/dsp/gain_interpolator.hpp:40-130): `interp_points` is
    /// the public, mutable curve; it is re-sent to the device when it differs from what the
    /// device holds.
    template <typename InterpT>
    class GainInterpolator {
     public:
      std::vector<std::pair<SampleIndex, typename InterpT::Point>> interp_points;

      void process(SampleIndex block_start, size_t nsamples, const float *const *in, float *const *out) {
        if (nsamples == 0) return;
        if (interp_points.empty()) throw invalid_argument("no interpolation points");
        sync_points();
        _in.gather(in, _n_in, nsamples);
        _out.prepare(out, _n_out, nsamples);
        b200::check(earb200_interp_process_host(_h.get(), (int64_t)block_start, (int64_t)nsamples, _in.base,
                                                _in.stride, _out.base, _out.stride));
        _out.scatter(out, _n_out, nsamples);
      }

      /// device handle (valid after the first process() or sync_points())
      earb200_interp *handle() { return _h.get(); }

      void sync_points() {
        if (_h && interp_points == _synced) return;
        if (interp_points.empty()) throw invalid_argument("no interpolation points");
        const size_t N = InterpT::n_in(interp_points[0].second), M = InterpT::n_out(interp_points[0].second);
        if (!_h || N != _n_in || M != _n_out) {
          earb200_interp *h = nullptr;
          b200::check(earb200_interp_create((int)N, (int)M, b200::default_device(), &h));
          _h.reset(h);
          _n_in = N;
          _n_out = M;
        }
        std::vector<int64_t> idx(interp_points.size());
        std::vector<float> gains(interp_points.size() * N * M);
        for (size_t p = 0; p < interp_points.size(); p++) {
          if (InterpT::n_in(interp_points[p].second) != N || InterpT::n_out(interp_points[p].second) != M)
            throw invalid_argument("interpolation points differ in size");
          idx[p] = (int64_t)interp_points[p].first;
          InterpT::flatten(interp_points[p].second, gains.data() + p * N * M);
        }
        b200::check(earb200_interp_set_points(_h.get(), (int64_t)idx.size(), idx.data(), gains.data()));
        _synced = interp_points;
      }

     private:
      struct Deleter {
        void operator()(earb200_interp *h) const { earb200_interp_destroy(h); }
      };
      std::unique_ptr<earb200_interp, Deleter> _h;
      size_t _n_in = 0, _n_out = 0;
      std::vector<std::pair<SampleIndex, typename InterpT::Point>> _synced;
      detail::Rows _in;
      detail::OutRows _out;
    };

    // ------------------------------------------------------------ block convolver
    // ref include/ear/dsp/block_convolver.hpp:28-112
    namespace block_convolver {
      using real_t = float;
      using complex_t = std::complex<real_t>;

      class Context {
       public:
        /// fft_impl is accepted for source compatibility; the transforms run inside the
        /// convolver kernel.
        Context(size_t block_size, FFTImpl<real_t> &) : _block_size(block_size) {}
        size_t block_size() const { return _block_size; }

       private:
        size_t _block_size;
      };

      class Filter {
       public:
        Filter(const Context &ctx, size_t n, const real_t *filter)
            : _impl(std::make_shared<Impl>(Impl{ctx.block_size(), std::vector<real_t>(filter, filter + n)})) {}
        size_t num_blocks() const { return (_impl->taps.size() + _impl->block_size - 1) / _impl->block_size; }

       private:
        struct Impl {
          size_t block_size;
          std::vector<real_t> taps;
        };
        std::shared_ptr<const Impl> _impl;
        friend class BlockConvolver;
      };

      class BlockConvolver {
       public:
        BlockConvolver(const Context &ctx, size_t num_blocks) { init(ctx, num_blocks); }
        BlockConvolver(const Context &ctx, const Filter &filter, size_t num_blocks = 0) {
          init(ctx, num_blocks > 0 ? num_blocks : filter.num_blocks());
          set_filter(filter);
        }

        /// one block of block_size samples; in == nullptr is silence
        void process(const float *in, float *out) { process_blocks(1, in, out); }
        /// n consecutive blocks in one launch
        void process_blocks(size_t n_blocks, const float *in, float *out) {
          const int64_t n = (int64_t)(n_blocks * _block_size);
          b200::check(earb200_convolver_process_host(_h.get(), (int64_t)n_blocks, in, n, out, n));
        }
        void crossfade_filter(const Filter &filter) {
          b200::check(earb200_convolver_crossfade_filter(_h.get(), 0, filter_id(filter)));
        }
        void fade_down() { b200::check(earb200_convolver_crossfade_filter(_h.get(), 0, -1)); }
        void set_filter(const Filter &filter) { b200::check(earb200_convolver_set_filter(_h.get(), 0, filter_id(filter))); }
        void unset_filter() { b200::check(earb200_convolver_set_filter(_h.get(), 0, -1)); }
        earb200_convolver *handle() { return _h.get(); }

       private:
        void init(const Context &ctx, size_t num_blocks) {
          _block_size = ctx.block_size();
          _num_blocks = num_blocks;
          earb200_convolver *h = nullptr;
          b200::check(earb200_convolver_create((int64_t)_block_size, 1, (int64_t)num_blocks, b200::default_device(), &h));
          _h.reset(h);
        }
        // ref src/dsp/block_convolver_impl.cpp:83-96 (check_filter)
        int filter_id(const Filter &filter) {
          if (filter._impl->block_size != _block_size)
            throw invalid_argument(
                "Filter block size is not equal to BlockConvolver block size; was this created using the same "
                "context?");
          if (filter.num_blocks() > _num_blocks) throw invalid_argument("too many blocks in given Filter");
          auto it = _ids.find(filter._impl.get());
          if (it != _ids.end()) return it->second;
          int id = -1, blocks = 0;
          b200::check(earb200_convolver_add_filter(_h.get(), (int64_t)filter._impl->taps.size(),
                                                   filter._impl->taps.data(), &id, &blocks));
          _ids[filter._impl.get()] = id;
          _keep.push_back(filter._impl);
          return id;
        }
        struct Deleter {
          void operator()(earb200_convolver *h) const { earb200_convolver_destroy(h); }
        };
        std::unique_ptr<earb200_convolver, Deleter> _h;
        size_t _block_size = 0, _num_blocks = 0;
        std::map<const void *, int> _ids;
        std::vector<std::shared_ptr<const void>> _keep;
      };
    }  // namespace block_convolver

    // ------------------------------------------------------------------ delay buffer
    // ref include/ear/dsp/delay_buffer.hpp:12-29
    class DelayBuffer {
     public:
      DelayBuffer(size_t nchannels, size_t nsamples) : _n(nchannels) {
        earb200_delay *h = nullptr;
        b200::check(earb200_delay_create((int)nchannels, (int64_t)nsamples, b200::default_device(), &h));
        _h.reset(h);
      }
      void process(size_t nsamples, const float *const *input, float *const *output) {
        if (nsamples == 0 || _n == 0) return;
        _in.gather(input, _n, nsamples);
        _out.prepare(output, _n, nsamples);
        b200::check(earb200_delay_process_host(_h.get(), (int64_t)nsamples, _in.base, _in.stride, _out.base, _out.stride));
        _out.scatter(output, _n, nsamples);
      }
      int get_delay() const { return (int)earb200_delay_get_delay(_h.get()); }
      earb200_delay *handle() { return _h.get(); }

     private:
      struct Deleter {
        void operator()(earb200_delay *h) const { earb200_delay_destroy(h); }
      };
      std::unique_ptr<earb200_delay, Deleter> _h;
      size_t _n;
      detail::Rows _in;
      detail::OutRows _out;
    };

    // ------------------------------------------------------------------- ptr adapter
    // ref include/ear/dsp/ptr_adapter.hpp:11-40: `float**` over the columns of a column-major
    // matrix expression (anything with cols() and col(i).data())
    template <typename PtrT = float *>
    class PtrAdapterT {
     public:
      explicit PtrAdapterT(size_t nchannels) : _ptrs(nchannels) {}
      template <typename T>
      void set_eigen(T &&mat, size_t offset = 0) {
        if ((size_t)mat.cols() != _ptrs.size()) throw internal_error("wrong number of channels");
        for (size_t i = 0; i < _ptrs.size(); i++) _ptrs[i] = mat.col(i).data() + offset;
      }
      /// point channel i at base + i * stride + offset
      void set_strided(PtrT base, size_t stride, size_t offset = 0) {
        for (size_t i = 0; i < _ptrs.size(); i++) _ptrs[i] = base + i * stride + offset;
      }
      PtrT *ptrs() { return _ptrs.data(); }
      PtrAdapterT(const PtrAdapterT &) = delete;
      PtrAdapterT &operator=(const PtrAdapterT &) = delete;

     private:
      std::vector<PtrT> _ptrs;
    };
    using PtrAdapter = PtrAdapterT<float *>;
    using PtrAdapterConst = PtrAdapterT<const float *>;

    // ------------------------------------------------------ variable block size adapter
    // ref include/ear/dsp/variable_block_size.hpp:16-40, src/dsp/variable_block_size_impl.cpp:44-81.
    // Pure host-side buffering around a fixed-block process function (which is where the device
    // work happens); adds block_size samples of delay.
    class VariableBlockSizeAdapter {
     public:
      using ProcessFunc = void(const float *const *in, float *const *out);

      VariableBlockSizeAdapter(size_t block_size, size_t num_channels_in, size_t num_channels_out,
                               std::function<ProcessFunc> process_func)
          : _process(std::move(process_func)),
            _block(block_size),
            _n_in(num_channels_in),
            _n_out(num_channels_out),
            _inbuf(block_size * num_channels_in, 0.0f),
            _outbuf(block_size * num_channels_out, 0.0f),
            _in_ptrs(num_channels_in),
            _out_ptrs(num_channels_out) {
        for (size_t c = 0; c < _n_in; c++) _in_ptrs[c] = _inbuf.data() + c * _block;
        for (size_t c = 0; c < _n_out; c++) _out_ptrs[c] = _outbuf.data() + c * _block;
      }

      void process(size_t nsamples, const float *const *in, float *const *out) {
        size_t sample = 0;
        while (sample < nsamples) {
          // fill the input block and drain the output block in step; the output block holds
          // block_size - fill samples aligned to its end
          const size_t n = std::min(nsamples - sample, _block - _fill);
          for (size_t c = 0; c < _n_in; c++) std::memcpy(_inbuf.data() + c * _block + _fill, in[c] + sample, sizeof(float) * n);
          for (size_t c = 0; c < _n_out; c++)
            std::memcpy(out[c] + sample, _outbuf.data() + c * _block + _fill, sizeof(float) * n);
          sample += n;
          _fill += n;
          const bool run = _fill == _block;
          if (run) {
            _process(_in_ptrs.data(), _out_ptrs.data());
            _fill = 0;
          }
          if (!run && n == 0) throw internal_error("no progress made");
        }
      }
      int get_delay() const { return (int)_block; }

     private:
      std::function<ProcessFunc> _process;
      size_t _block, _n_in, _n_out, _fill = 0;
      std::vector<float> _inbuf, _outbuf;
      std::vector<const float *> _in_ptrs;
      std::vector<float *> _out_ptrs;
    };

  }  // namespace dsp
}  // namespace ear
