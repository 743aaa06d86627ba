// Boost→std shim (test infrastructure only): lets the unmodified libear sources
// compile in an image without Boost.  boost::optional → std::optional.
#pragma once
#include <algorithm>
#include <functional>
#include <optional>
#include <ostream>
#include <sstream>
#include <string>

namespace boost {
  struct none_t {
    constexpr operator std::nullopt_t() const { return std::nullopt; }
  };
  constexpr none_t none{};

  template <typename T>
  class optional : public std::optional<T> {
   public:
    using base_t = std::optional<T>;
    using base_t::base_t;
    optional() = default;
    optional(const optional &) = default;
    optional(optional &&) = default;
    optional &operator=(const optional &) = default;
    optional &operator=(optional &&) = default;
    optional(none_t) : base_t(std::nullopt) {}
    optional &operator=(none_t) {
      this->reset();
      return *this;
    }
    // std::optional's converting assignment is disabled for scalar T = decay<U>
    // (it goes through the implicit constructor instead), which is ambiguous in a
    // derived class; provide the assignment directly.
    template <typename U,
              typename = std::enable_if_t<
                  std::is_constructible<T, U>::value &&
                  !std::is_same<std::decay_t<U>, optional>::value &&
                  !std::is_same<std::decay_t<U>, none_t>::value &&
                  !std::is_same<std::decay_t<U>, base_t>::value>>
    optional &operator=(U &&v) {
      this->emplace(std::forward<U>(v));
      return *this;
    }
    T &get() { return this->value(); }
    const T &get() const { return this->value(); }
    bool operator!() const { return !this->has_value(); }
  };

  template <typename T>
  bool operator==(const optional<T> &o, none_t) { return !o.has_value(); }
  template <typename T>
  bool operator==(none_t, const optional<T> &o) { return !o.has_value(); }
  template <typename T>
  bool operator!=(const optional<T> &o, none_t) { return o.has_value(); }
  template <typename T>
  bool operator!=(none_t, const optional<T> &o) { return o.has_value(); }

  template <typename T>
  T get_optional_value_or(const optional<T> &o, const T &fallback) {
    return o.has_value() ? *o : fallback;
  }
  template <typename T>
  T &get(optional<T> &o) { return o.value(); }
  template <typename T>
  const T &get(const optional<T> &o) { return o.value(); }
}  // namespace boost
