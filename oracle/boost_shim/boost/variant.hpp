// Boost→std shim: boost::variant → std::variant.
#pragma once
#include <utility>
#include <variant>

namespace boost {
  template <typename... Ts>
  class variant : public std::variant<Ts...> {
   public:
    using base_t = std::variant<Ts...>;
    using base_t::base_t;
    using base_t::operator=;
    variant() = default;
    base_t &base() { return *this; }
    const base_t &base() const { return *this; }
  };

  template <typename R = void>
  struct static_visitor {
    using result_type = R;
  };

  template <typename Visitor, typename... Ts>
  decltype(auto) apply_visitor(Visitor &&v, variant<Ts...> &x) {
    return std::visit(std::forward<Visitor>(v), x.base());
  }
  template <typename Visitor, typename... Ts>
  decltype(auto) apply_visitor(Visitor &&v, const variant<Ts...> &x) {
    return std::visit(std::forward<Visitor>(v), x.base());
  }

  template <typename T, typename... Ts>
  T &get(variant<Ts...> &x) { return std::get<T>(x.base()); }
  template <typename T, typename... Ts>
  const T &get(const variant<Ts...> &x) { return std::get<T>(x.base()); }
}  // namespace boost
