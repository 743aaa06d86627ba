#pragma once
namespace boost { namespace math {
  template <typename T>
  T factorial(unsigned n) { T r = 1; for (unsigned i = 2; i <= n; ++i) r *= static_cast<T>(i); return r; }
}}
