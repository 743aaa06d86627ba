#pragma once
namespace boost { namespace math { namespace constants {
  template <typename T>
  constexpr T pi() { return static_cast<T>(3.141592653589793238462643383279502884L); }
}}}
