// Host stand-in for <cuda_fp8.h>: the one conversion the EP dispatch kernel uses (float -> e4m3, saturating, round to
// nearest even), for the protocol simulator.
#pragma once
#include <cuda_runtime.h>

enum { __NV_SATFINITE = 1 };
enum { __NV_E4M3 = 0 };
typedef unsigned char __nv_fp8_storage_t;

inline __nv_fp8_storage_t __nv_cvt_float_to_fp8(float f, int, int) {
  const unsigned sign = f < 0.f || (f == 0.f && signbit(f)) ? 0x80u : 0u;
  float a = fabsf(f);
  if (a != a) return __nv_fp8_storage_t(sign | 0x7f);
  if (a >= 448.f) return __nv_fp8_storage_t(sign | 0x7e);  // saturate to the largest finite value
  if (a < ldexpf(1.f, -10)) return __nv_fp8_storage_t(sign);  // below half the smallest subnormal
  int e;
  frexpf(a, &e);  // a = m * 2^e, m in [0.5, 1)
  int exp = e - 1;  // a = 1.xxx * 2^exp
  if (exp < -6) exp = -6;  // subnormal range: fixed exponent
  const float q = ldexpf(a, 3 - exp);  // value in units of 2^(exp-3): 8..15.x for normals, 0..7.x for subnormals
  float r = nearbyintf(q);  // round to nearest even (default rounding mode)
  int mant = int(r);
  if (mant >= 16) {
    mant = 8;
    ++exp;
  }
  if (exp > 8 || (exp == 8 && mant > 14)) return __nv_fp8_storage_t(sign | 0x7e);
  if (mant < 8) return __nv_fp8_storage_t(sign | unsigned(mant));  // subnormal (exp == -6)
  return __nv_fp8_storage_t(sign | (unsigned(exp + 7) << 3) | unsigned(mant - 8));
}
