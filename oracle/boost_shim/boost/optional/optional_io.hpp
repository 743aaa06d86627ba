#pragma once
#include "../optional.hpp"
namespace boost {
  template <typename T>
  std::ostream &operator<<(std::ostream &os, const optional<T> &o) {
    if (o.has_value()) return os << ' ' << *o;
    return os << "--";
  }
}  // namespace boost
