// libear-compatible include path; the implementation is in ear_b200/ 
#pragma once
#include "../ear_b200/conversion.hpp"
