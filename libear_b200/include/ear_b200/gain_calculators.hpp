// libear-b200 C++14 facade, part 3: the three gain calculators with libear's signatures
// (ref include/ear/gain_calculators.hpp:18-82) on top of the C ABI.  Each object owns a device
// handle; like libear's, an instance is not thread-safe.  calculate_batch() is the addition a
// GPU is for: many metadata blocks per call.
#pragma once
#include <limits>
#include <cstdint>

#include "layout.hpp"

namespace ear {

  namespace b200 {
    // ObjectsTypeMetadata -> one row of the SoA + the flags libear refuses
    // (ref src/object_based/gain_calculator_objects.cpp:37-43)
    inline int32_t objectFlags(const ObjectsTypeMetadata &m) {
      int32_t flags = 0;
      if (m.cartesian) flags |= EARB200_OBJ_CARTESIAN;
      if (m.position.which() == 0) flags |= EARB200_OBJ_CARTESIAN_POSITION;
      const double divergence = m.objectDivergence.which() == 0
                                    ? compat::get<PolarObjectDivergence>(m.objectDivergence).divergence
                                    : compat::get<CartesianObjectDivergence>(m.objectDivergence).divergence;
      if (divergence != 0.0) flags |= EARB200_OBJ_DIVERGENCE;
      if (m.channelLock.flag) flags |= EARB200_OBJ_CHANNEL_LOCK;
      if (!m.zoneExclusion.zones.empty()) flags |= EARB200_OBJ_ZONE_EXCLUSION;
      if (m.screenRef) flags |= EARB200_OBJ_SCREEN_REF;
      return flags;
    }
    template <typename T>
    struct is_double {
      static const int value = 0;
    };
    template <>
    struct is_double<double> {
      static const int value = 1;
    };
  }  // namespace b200

  /// Gain calculator for typeDefinition == "Objects"
  class GainCalculatorObjects {
   public:
    explicit GainCalculatorObjects(const Layout &layout) {
      b200::LayoutHandle l = b200::toHandle(layout);
      earb200_objects *h = nullptr;
      b200::check(earb200_objects_create(l.get(), b200::default_device(), &h));
      _h.reset(h);
      _n_out = (size_t)earb200_objects_num_outputs(h);
    }

    template <typename T>
    void calculate(const ObjectsTypeMetadata &metadata, std::vector<T> &directGains, std::vector<T> &diffuseGains,
                   const WarningCB & = default_warning_cb) {
      static_assert(sizeof(T) == 4 || sizeof(T) == 8, "T is float or double");
      std::vector<T> direct(_n_out), diffuse(_n_out);
      calculate_batch(&metadata, 1, direct.data(), diffuse.data());
      // libear computes first and checks the caller's vectors when writing
      // (ref include/ear/helpers/output_gains.hpp:23-44)
      if (directGains.size() != _n_out || diffuseGains.size() != _n_out)
        throw invalid_argument("incorrect size for output vector");
      directGains = direct;
      diffuseGains = diffuse;
    }

    /// n metadata blocks in one call: direct / diffuse are [n][numOutputs()] row-major.
    template <typename T>
    void calculate_batch(const ObjectsTypeMetadata *metadata, size_t n, T *direct, T *diffuse) {
      _soa.resize(8 * n);
      _flags.resize(n);
      for (size_t i = 0; i < n; i++) {
        const ObjectsTypeMetadata &m = metadata[i];
        _flags[i] = b200::objectFlags(m);
        const PolarPosition p =
            m.position.which() == 1 ? compat::get<PolarPosition>(m.position) : PolarPosition();
        const double row[8] = {p.azimuth, p.elevation, p.distance, m.width, m.height, m.depth, m.gain, m.diffuse};
        for (int k = 0; k < 8; k++) _soa[k * n + i] = row[k];
      }
      const double *c = _soa.data();
      earb200_objects_soa in = {c, c + n, c + 2 * n, c + 3 * n, c + 4 * n, c + 5 * n, c + 6 * n, c + 7 * n,
                                _flags.data()};
      if (_extensions) {
        // opt-in: polar objectDivergence / channelLock are computed instead of refused (see
        // enableExtensions); the flags of what is enabled are cleared, everything else still throws
        _ext.assign(3 * n, 0.0);
        bool any_div = false, any_lock = false;
        for (size_t i = 0; i < n; i++) {
          const ObjectsTypeMetadata &m = metadata[i];
          _ext[2 * n + i] = -1.0;
          if ((_extensions & EARB200_EXT_DIVERGENCE) && m.objectDivergence.which() == 0) {
            const PolarObjectDivergence &d = compat::get<PolarObjectDivergence>(m.objectDivergence);
            _ext[i] = d.divergence;
            _ext[n + i] = d.azimuthRange;
            any_div = any_div || d.divergence != 0.0;
            _flags[i] &= ~(int32_t)EARB200_OBJ_DIVERGENCE;
          }
          if ((_extensions & EARB200_EXT_CHANNEL_LOCK) && m.channelLock.flag) {
            _ext[2 * n + i] = m.channelLock.maxDistance ? m.channelLock.maxDistance.get()
                                                        : std::numeric_limits<double>::infinity();
            any_lock = true;
            _flags[i] &= ~(int32_t)EARB200_OBJ_CHANNEL_LOCK;
          }
        }
        if (any_div || any_lock) {
          const double *e = _ext.data();
          earb200_objects_ext ext = {any_div ? e : nullptr, any_div ? e + n : nullptr, any_lock ? e + 2 * n : nullptr};
          b200::check(earb200_objects_calculate_ext_host(_h.get(), (int64_t)n, &in, &ext, b200::is_double<T>::value,
                                                         direct, diffuse));
          return;
        }
      }
      b200::check(earb200_objects_calculate_host(_h.get(), (int64_t)n, &in, b200::is_double<T>::value, direct,
                                                 diffuse, nullptr));
    }

    /// OPT-IN, not in libear (which throws not_implemented for both, and so does this class by
    /// default): compute polar objectDivergence and / or channelLock.  PARITY UNPINNED: there is no
    /// libear output to compare with (see earb200.h, csrc/objects_ext_kernels.cu).
    void enableExtensions(bool divergence, bool channelLock) {
      _extensions = (divergence ? (uint32_t)EARB200_EXT_DIVERGENCE : 0u) |
                    (channelLock ? (uint32_t)EARB200_EXT_CHANNEL_LOCK : 0u);
      b200::check(earb200_objects_enable_extensions(_h.get(), _extensions));
    }
    template <typename T>
    void calculate_batch(const std::vector<ObjectsTypeMetadata> &metadata, std::vector<T> &direct,
                         std::vector<T> &diffuse) {
      direct.resize(metadata.size() * _n_out);
      diffuse.resize(metadata.size() * _n_out);
      calculate_batch(metadata.data(), metadata.size(), direct.data(), diffuse.data());
    }

    size_t numOutputs() const { return _n_out; }
    earb200_objects *handle() { return _h.get(); }  // for the *_device entry points of the C ABI

   private:
    struct Deleter {
      void operator()(earb200_objects *h) const { earb200_objects_destroy(h); }
    };
    std::unique_ptr<earb200_objects, Deleter> _h;
    size_t _n_out = 0;
    std::vector<double> _soa, _ext;
    std::vector<int32_t> _flags;
    uint32_t _extensions = 0;
  };

  /// Gain calculator for typeDefinition == "DirectSpeakers"
  class GainCalculatorDirectSpeakers {
   public:
    explicit GainCalculatorDirectSpeakers(const Layout &layout,
                                          std::map<std::string, std::string> additionalSubstitutions = {}) {
      b200::LayoutHandle l = b200::toHandle(layout);
      std::vector<const char *> from, to;
      for (const auto &kv : additionalSubstitutions) {
        from.push_back(kv.first.c_str());
        to.push_back(kv.second.c_str());
      }
      earb200_direct_speakers *h = nullptr;
      b200::check(earb200_direct_speakers_create(l.get(), b200::default_device(), (int)from.size(), from.data(),
                                                 to.data(), &h));
      _h.reset(h);
      _n_out = (size_t)earb200_direct_speakers_num_outputs(h);
    }

    template <typename T>
    void calculate(const DirectSpeakersTypeMetadata &metadata, std::vector<T> &gains,
                   const WarningCB &warning_cb = default_warning_cb) {
      // ref src/direct_speakers/gain_calculator_direct_speakers.cpp:247-252: the pack/label
      // consistency check comes before the size check
      if (metadata.audioPackFormatID && metadata.speakerLabels.empty())
        throw adm_error(
            "common definitions audioPackFormatID specified without any speakerLabels as specified in the common "
            "definitions file");
      if (gains.size() != _n_out) throw invalid_argument("incorrect size for output vector");
      std::vector<const char *> labels;
      for (const std::string &s : metadata.speakerLabels) labels.push_back(s.c_str());
      earb200_direct_speakers_metadata md = {};
      md.speaker_labels = labels.data();
      md.n_speaker_labels = (int32_t)labels.size();
      if (metadata.position.which() == 1) {
        md.cartesian_position = 1;
      } else {
        const PolarSpeakerPosition &p = compat::get<PolarSpeakerPosition>(metadata.position);
        md.position[0] = p.azimuth;
        md.position[1] = p.elevation;
        md.position[2] = p.distance;
        const compat::optional<double> *bounds[6] = {&p.azimuthMin,   &p.azimuthMax,  &p.elevationMin,
                                                     &p.elevationMax, &p.distanceMin, &p.distanceMax};
        for (int i = 0; i < 6; i++)
          if (*bounds[i]) {
            md.bounds[i] = bounds[i]->get();
            md.has_bounds |= 1u << i;
          }
        if (p.screenEdgeLock.horizontal) md.screen_edge_lock |= 1;
        if (p.screenEdgeLock.vertical) md.screen_edge_lock |= 2;
      }
      if (metadata.channelFrequency.lowPass) {
        md.low_pass = metadata.channelFrequency.lowPass.get();
        md.has_frequency |= 1;
      }
      if (metadata.channelFrequency.highPass) {
        md.high_pass = metadata.channelFrequency.highPass.get();
        md.has_frequency |= 2;
      }
      std::string pack;
      if (metadata.audioPackFormatID) {
        pack = metadata.audioPackFormatID.get();
        md.audio_pack_format_id = pack.c_str();
      }
      std::vector<double> g(_n_out);
      uint32_t warnings = 0;
      b200::check(earb200_direct_speakers_calculate(_h.get(), 1, &md, g.data(), &warnings));
      // libear raises these two synchronously and in this order (ref :112-133)
      if (warnings & (1u << EARB200_WARN_FREQ_NOT_LFE))
        warning_cb({Warning::Code::FREQ_NOT_LFE, "frequency indication present but does not indicate an LFE channel"});
      if (warnings & (1u << EARB200_WARN_FREQ_SPEAKERLABEL_LFE_MISMATCH))
        warning_cb({Warning::Code::FREQ_SPEAKERLABEL_LFE_MISMATCH,
                    "LFE indication from frequency element does not match speakerLabel"});
      for (size_t i = 0; i < _n_out; i++) gains[i] = (T)g[i];
    }

   private:
    struct Deleter {
      void operator()(earb200_direct_speakers *h) const { earb200_direct_speakers_destroy(h); }
    };
    std::unique_ptr<earb200_direct_speakers, Deleter> _h;
    size_t _n_out = 0;
  };

  /// Gain calculator for typeDefinition == "HOA"
  class GainCalculatorHOA {
   public:
    explicit GainCalculatorHOA(const Layout &layout) {
      b200::LayoutHandle l = b200::toHandle(layout);
      earb200_hoa *h = nullptr;
      b200::check(earb200_hoa_create(l.get(), b200::default_device(), &h));
      _h.reset(h);
      _n_out = layout.channels().size();
    }

    /// gains: one vector of per-loudspeaker gains per input channel, pre-sized by the caller
    template <typename T>
    void calculate(const HOATypeMetadata &metadata, std::vector<std::vector<T>> &gains,
                   const WarningCB &warning_cb = default_warning_cb) {
      if (metadata.orders.size() != metadata.degrees.size())
        throw invalid_argument("orders and degrees must be the same size");
      const size_t C = metadata.orders.size();
      std::vector<double> D(std::max<size_t>(1, C * _n_out));
      uint32_t warnings = 0;
      b200::check(earb200_hoa_calculate(_h.get(), (int)C, metadata.orders.data(), metadata.degrees.data(),
                                        metadata.normalization.c_str(), metadata.nfcRefDist, metadata.screenRef ? 1 : 0,
                                        D.data(), &warnings));
      if (warnings & (1u << EARB200_WARN_HOA_SCREENREF))
        warning_cb({Warning::Code::HOA_SCREENREF_NOT_IMPLEMENTED, "screenRef for HOA is not implemented; ignoring"});
      if (warnings & (1u << EARB200_WARN_HOA_NFCREFDIST))
        warning_cb({Warning::Code::HOA_NFCREFDIST_NOT_IMPLEMENTED, "nfcRefDist is not implemented; ignoring"});
      // ref include/ear/helpers/output_gains.hpp:85-92
      if (gains.size() != C) throw invalid_argument("incorrect number of cols in output matrix");
      for (const std::vector<T> &col : gains)
        if (col.size() != _n_out) throw invalid_argument("incorrect number of rows in output matrix column");
      for (size_t c = 0; c < C; c++)
        for (size_t o = 0; o < _n_out; o++) gains[c][o] = (T)D[c * _n_out + o];
    }

   private:
    struct Deleter {
      void operator()(earb200_hoa *h) const { earb200_hoa_destroy(h); }
    };
    std::unique_ptr<earb200_hoa, Deleter> _h;
    size_t _n_out = 0;
  };

  // ---------------------------------------------------------------- decorrelators
  // ref include/ear/decorrelate.hpp:17-34
  template <typename T = float>
  std::vector<std::vector<T>> designDecorrelators(const Layout &layout) {
    b200::LayoutHandle l = b200::toHandle(layout);
    const size_t n = layout.channels().size(), size = (size_t)earb200_decorrelator_size();
    std::vector<T> flat(std::max<size_t>(1, n * size));
    b200::check(earb200_design_decorrelators(l.get(), sizeof(T) == 4 ? (float *)(void *)flat.data() : nullptr,
                                             sizeof(T) == 8 ? (double *)(void *)flat.data() : nullptr));
    std::vector<std::vector<T>> out(n);
    for (size_t c = 0; c < n; c++) out[c].assign(flat.begin() + c * size, flat.begin() + (c + 1) * size);
    return out;
  }
  template <typename T = float>
  std::vector<T> designDecorrelator(const Layout &layout, size_t channelIdx) {
    return designDecorrelators<T>(layout).at(channelIdx);
  }
  inline int decorrelatorCompensationDelay() { return earb200_decorrelator_compensation_delay(); }

}  // namespace ear
