#pragma once
#include <algorithm>
namespace boost { namespace algorithm {
  template <typename T>
  const T &clamp(const T &v, const T &lo, const T &hi) { return std::clamp(v, lo, hi); }
}}
