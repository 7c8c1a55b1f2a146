// Pre-included (-include) into every reference TU.  MSVC pulls these in transitively; g++ does
// not (memcpy vec.cpp:24, memset matrix3d.cpp:5 / RadianceRGB.cpp:8, pow/floor
// RadianceRGB.cpp:59,64, fmin/fmax/DBL_MAX Myobj.cpp:101-133, isinf/isnan Mylight.cpp:250,410).
#pragma once
#include <cstring>
#include <cmath>
#include <cfloat>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
using std::isinf;
using std::isnan;
#include "rng_shim.h"
