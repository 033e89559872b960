// Every system header the library uses, included FIRST by every translation unit, followed — in the single-precision
// build (-DMHH_F32: the `_f32` symbol set, the reference's USESP / FLOAT_SINGLE build, main/microhh.cxx:40-46) — by the
// switch of the working type: the sources are written once, in `double`, and compiled twice.
//   mhh_f64 : always IEEE double (time steps — the reference's rk3_g / rk4_g take `double dt` for TF = float too,
//             src/timeloop.cu:35 — and the bit patterns of the NaN-propagating maxima)
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cudaTypedefs.h>

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

typedef double mhh_f64;
typedef long double mhh_f80;      // host-side twiddle factors
#ifndef MHH_DOUBLE_DEFINED
#define MHH_DOUBLE_DEFINED
typedef double mhh_double;      // the ABI's always-double scalar (include/mhh_b200.h)
#endif

#ifdef MHH_F32
#include "../../include/mhh_b200_f32_names.h"   // mhh_xxx -> mhh_xxx_f32 for every export, mhh_ctx -> mhh_ctx_f32
#define mhh mhh_f32                              // the C++ namespace of the single-precision objects
#define double float
#define double2 float2
#define make_double2 make_float2
#define MHH_REAL_BYTES 4
#else
#define MHH_REAL_BYTES 8
#endif
