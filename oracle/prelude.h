/* Force-included portability prelude for building the UNMODIFIED reference
 * sources with g++ 13: std::sqrtf is missing from libstdc++ < 14
 * (src/Ray.h:68) and several headers are only reached transitively on MSVC
 * (FLT_EPSILON at src/SettingData.cpp:244, std::function at src/SettingData.h:20).
 * Test infrastructure only. */
#pragma once
#include <cmath>
#include <cfloat>
#include <functional>
#include <algorithm>
#include <limits>
#include <vector>
namespace std { using ::sqrtf; }
