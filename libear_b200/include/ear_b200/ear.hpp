// libear-b200: header-only C++14 facade with libear's API (namespace ear) over the C ABI in
// <earb200.h>.  Link with -learb200.  See INTEGRATION.md.
#pragma once
#include "conversion.hpp"
#include "dsp.hpp"
#include "gain_calculators.hpp"
#include "layout.hpp"
#include "types.hpp"
