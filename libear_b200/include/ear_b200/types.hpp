// libear-b200 C++14 facade, part 1: value types.
//
// The names, members and construction syntax are libear's (ref include/ear/common_types.hpp,
// exceptions.hpp, warnings.hpp, screen.hpp, layout.hpp, metadata.hpp) so that code written
// against libear compiles against these headers; nothing here computes anything.  libear's
// public structs use boost::optional / boost::variant; Boost is not required here: when
// <boost/optional.hpp> is on the include path (or EAR_B200_USE_BOOST is defined) the real
// Boost types are used, otherwise the two small stand-ins below (same spelling of the
// operations libear's users need: none, get(), operator bool, which(), get<T>(v)).
#pragma once
#include <earb200.h>

#include <functional>
#include <iostream>
#include <map>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#if !defined(EAR_B200_USE_BOOST) && defined(__has_include)
#if __has_include(<boost/optional.hpp>) && __has_include(<boost/variant.hpp>)
#define EAR_B200_USE_BOOST 1
#endif
#endif

#ifdef EAR_B200_USE_BOOST
#include <boost/optional.hpp>
#include <boost/variant.hpp>
namespace ear {
  namespace compat {
    using boost::get;
    using boost::none;
    using boost::optional;
    template <typename A, typename B>
    using variant2 = boost::variant<A, B>;
  }  // namespace compat
}  // namespace ear
#else
namespace ear {
  namespace compat {
    struct none_t {};
    constexpr none_t none{};

    template <typename T>
    class optional {
     public:
      optional() : _set(false), _value() {}
      optional(none_t) : _set(false), _value() {}
      optional(const T &v) : _set(true), _value(v) {}
      template <typename U, typename = decltype(T(std::declval<U>()))>
      optional(const U &v) : _set(true), _value(v) {}
      optional &operator=(none_t) {
        _set = false;
        _value = T();
        return *this;
      }
      explicit operator bool() const { return _set; }
      bool operator!() const { return !_set; }
      const T &get() const { return checked(); }
      T &get() {
        checked();
        return _value;
      }
      const T &operator*() const { return checked(); }
      const T *operator->() const { return &checked(); }
      const T &value_or(const T &other) const { return _set ? _value : other; }
      friend bool operator==(const optional &a, const optional &b) {
        return a._set == b._set && (!a._set || a._value == b._value);
      }
      friend bool operator!=(const optional &a, const optional &b) { return !(a == b); }
      friend bool operator==(const optional &a, none_t) { return !a._set; }
      friend bool operator!=(const optional &a, none_t) { return a._set; }

     private:
      const T &checked() const {
        if (!_set) throw std::logic_error("ear::compat::optional: no value");
        return _value;
      }
      bool _set;
      T _value;
    };

    // a two-alternative variant; like boost::variant it default-constructs its first type
    template <typename A, typename B>
    class variant2 {
     public:
      variant2() : _which(0), _a(), _b() {}
      variant2(const A &a) : _which(0), _a(a), _b() {}
      variant2(const B &b) : _which(1), _a(), _b(b) {}
      int which() const { return _which; }
      template <typename T>
      const T &as() const {
        return pick(static_cast<const T *>(nullptr));
      }

     private:
      const A &pick(const A *) const {
        if (_which != 0) throw std::logic_error("ear::compat::variant2: bad get");
        return _a;
      }
      const B &pick(const B *) const {
        if (_which != 1) throw std::logic_error("ear::compat::variant2: bad get");
        return _b;
      }
      int _which;
      A _a;
      B _b;
    };

    template <typename T, typename A, typename B>
    const T &get(const variant2<A, B> &v) {
      return v.template as<T>();
    }
  }  // namespace compat
}  // namespace ear
#endif

namespace ear {

  // ------------------------------------------------------------------ exceptions
  // ref include/ear/exceptions.hpp:8-43 (same classes, bases and message prefixes)
  class not_implemented : public std::runtime_error {
   public:
    explicit not_implemented(const std::string &what) : std::runtime_error("not implemented: " + what) {}
  };
  class internal_error : public std::runtime_error {
   public:
    explicit internal_error(const std::string &what) : std::runtime_error("internal error: " + what) {}
  };
  class adm_error : public std::invalid_argument {
   public:
    explicit adm_error(const std::string &what) : std::invalid_argument("ADM error: " + what) {}
  };
  class unknown_layout : public std::invalid_argument {
   public:
    explicit unknown_layout(const std::string &what) : std::invalid_argument("unknown layout: " + what) {}
  };
  class invalid_argument : public std::invalid_argument {
   public:
    explicit invalid_argument(const std::string &what) : std::invalid_argument(what) {}
  };
  // not in libear: the device is missing or a CUDA call failed (there is no CPU fallback)
  class cuda_error : public std::runtime_error {
   public:
    explicit cuda_error(const std::string &what) : std::runtime_error(what) {}
  };

  namespace b200 {
    // status code of the C ABI -> the libear exception it stands for
    inline void check(int rc) {
      if (rc == EARB200_OK) return;
      // earb200_last_error() holds the what() string libear would produce; the classes
      // below add their prefix again, so drop it here
      std::string msg = earb200_last_error();
      for (const char *prefix : {"not implemented: ", "internal error: ", "ADM error: ", "unknown layout: "})
        if (msg.compare(0, std::string(prefix).size(), prefix) == 0) msg = msg.substr(std::string(prefix).size());
      switch (rc) {
        case EARB200_NOT_IMPLEMENTED: throw not_implemented(msg);
        case EARB200_ADM_ERROR: throw adm_error(msg);
        case EARB200_UNKNOWN_LAYOUT: throw unknown_layout(msg);
        case EARB200_INVALID_ARGUMENT: throw invalid_argument(msg);
        case EARB200_CUDA_ERROR: throw cuda_error(msg);
        default: throw internal_error(msg);
      }
    }
    // CUDA device new handles are created on (per thread of the caller)
    inline int &default_device() {
      static thread_local int device = 0;
      return device;
    }
  }  // namespace b200

  // -------------------------------------------------------------------- warnings
  // ref include/ear/warnings.hpp:12-37
  struct Warning {
    enum class Code {
      FREQ_SPEAKERLABEL_LFE_MISMATCH = 1,
      FREQ_NOT_LFE,
      FREQ_IGNORED,
      HOA_SCREENREF_NOT_IMPLEMENTED,
      HOA_NFCREFDIST_NOT_IMPLEMENTED,
    };
    Code code;
    std::string message;
  };
  using WarningCB = std::function<void(const Warning &warning)>;
  // prints "libear: warning: <message>" to stderr (ref src/warnings.cpp)
  inline void print_warning(const Warning &warning) { std::cerr << "libear: warning: " << warning.message << std::endl; }
  static const WarningCB default_warning_cb = print_warning;

  // ---------------------------------------------------------------- common types
  // ref include/ear/common_types.hpp:8-25
  struct CartesianPosition {
    CartesianPosition(double X = 0.0, double Y = 0.0, double Z = 0.0) : X(X), Y(Y), Z(Z) {}
    double X, Y, Z;
  };
  struct PolarPosition {
    PolarPosition(double azimuth = 0.0, double elevation = 0.0, double distance = 1.0)
        : azimuth(azimuth), elevation(elevation), distance(distance) {}
    double azimuth, elevation, distance;
  };
  using Position = compat::variant2<CartesianPosition, PolarPosition>;

  // ---------------------------------------------------------------------- screen
  // ref include/ear/screen.hpp:8-22, src/screen.cpp
  struct CartesianScreen {
    double aspectRatio;
    CartesianPosition centrePosition;
    double widthX;
  };
  struct PolarScreen {
    double aspectRatio;
    PolarPosition centrePosition;
    double widthAzimuth;
  };
  using Screen = compat::variant2<PolarScreen, CartesianScreen>;
  inline Screen getDefaultScreen() { return PolarScreen{1.78, PolarPosition(0.0, 0.0, 1.0), 58.0}; }

}  // namespace ear
