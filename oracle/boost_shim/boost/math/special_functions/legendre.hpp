#pragma once
#include <cmath>
namespace boost { namespace math {
  // Boost's legendre_p(l, m, x) includes the Condon-Shortley phase (-1)^m;
  // std::assoc_legendre does not.
  inline double legendre_p(int l, int m, double x) {
    return ((m & 1) ? -1.0 : 1.0) * std::assoc_legendre(static_cast<unsigned>(l), static_cast<unsigned>(m), x);
  }
}}
