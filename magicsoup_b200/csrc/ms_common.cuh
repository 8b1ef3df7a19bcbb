// Shared device helpers for the magicsoup_b200 kernels (sm_100a).
// Arithmetic contract ("spec order"): float32, no FMA contraction
// (the .cu files are compiled with -fmad=false), sums in ascending index order,
// integer powers by repeated multiplication.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/magicsoup_b200.h"

namespace ms {

// reference: python/magicsoup/kinetics.py:11-13 (_EPS, _MAX, _MIN as float32)
constexpr float V_EPS = 1e-36f;
constexpr float V_MAX = 1e36f;
constexpr float V_MIN = -1e36f;
// kinetics.py:822-823, python floats compared against float32 tensors
constexpr float UPPER = 1.5f;
constexpr float LOWER = (float)(1.0 / 1.5);
constexpr float IMPACT = 0.1f;  // kinetics.py:825

__host__ __device__ inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// step gate (ms_step_gate): a committing kernel leaves without writing when the gate word is 0
__device__ __forceinline__ bool gate_closed(const unsigned *gate) { return gate != nullptr && *gate == 0u; }

// x^n, n >= 1, right-to-left binary exponentiation
__device__ __forceinline__ float powi(float x, int n) {
  float r = 1.0f, b = x;
  while (n) {
    if (n & 1) r = r * b;
    n >>= 1;
    if (n) b = b * b;
  }
  return r;
}

// kinetics.py:913-915
__device__ __forceinline__ float fix_product(float xx) {
  if (isnan(xx)) xx = 0.0f;
  if (xx < 0.0f) xx = 0.0f;
  if (isinf(xx)) xx = V_MAX;
  return xx;
}

// min with torch semantics: NaN propagates
__device__ __forceinline__ float nanmin(float a, float b) {
  if (isnan(a)) return a;
  if (isnan(b)) return b;
  return b < a ? b : a;
}

__device__ __forceinline__ float clamp0(float x) { return x < 0.0f ? 0.0f : x; }

// kinetics.py:881-892 from the two products (already masked for non-involved)
__device__ __forceinline__ float quotient_from(float xb, float xf) {
  float q = xb / xf;
  if (q < V_EPS) q = V_EPS;
  if (q > V_MAX) q = V_MAX;
  if (isnan(q)) q = 1.0f;
  return q;
}

// Same value as quotient_from() for products that went through fix_product (finite,
// >= 0, never NaN), but without dividing by or into zero: x/0 = inf -> MAX, 0/0 = NaN -> 1,
// 0/x = 0 -> EPS.  Saturated quotients are the common case and would
// otherwise take the slow path of the IEEE division.
__device__ __forceinline__ float quotient_fixed(float xb, float xf) {
  if (xf == 0.0f) return xb == 0.0f ? 1.0f : V_MAX;
  if (xb == 0.0f) return V_EPS;
  float q = xb / xf;
  if (q < V_EPS) q = V_EPS;
  if (q > V_MAX) q = V_MAX;
  return q;
}

// kinetics.py:786-806 from kf, kb (already masked / inf-fixed), a_reg, vmax
__device__ __forceinline__ float velocity_from(float kf, float kb, float a_reg, float vmax) {
  float a_cat = (kf - kb) / ((1.0f + kf) + kb);
  if (isinf(a_reg)) a_reg = V_MAX;
  float v = (a_cat * vmax) * a_reg;
  if (v < V_MIN) v = V_MIN;
  if (v > V_MAX) v = V_MAX;
  return v;
}

// one allosteric factor, kinetics.py:795-797
__device__ __forceinline__ float reg_factor(float x, int a, float kmr) {
  float t = a > 0 ? powi(x, a) : 1.0f / powi(x, -a);
  t = t / (t + kmr);
  if (isnan(t)) t = 1.0f;
  return t;
}

// kinetics.py:825-843 ; returns bit0 = too high (F -= inc), bit1 = too low (F += inc),
// bit2 = has impact
__device__ __forceinline__ unsigned eq_decide(float v, float q, float ke, float F) {
  const bool imp = fabsf(v) > IMPACT;
  const bool fwd = v > 0.0f;
  const float r = q / ke;
  bool lo = fwd ? (r < LOWER) : (r > UPPER);
  if (fwd && F == 1.0f) lo = false;
  bool hi = fwd ? (r > UPPER) : (r < LOWER);
  if (!fwd && F == 0.0f) hi = false;
  return (hi ? 1u : 0u) | (lo ? 2u : 0u) | (imp ? 4u : 0u);
}

// Same decisions as eq_decide() for q in [EPS, MAX] (a clamped quotient).  The value of
// r = q / ke only matters relative to 1/1.5 and 1.5: for a positive normal ke,
// q > 4 ke implies fl(q/ke) >= 4 > 1.5 and q < ke/4 implies fl(q/ke) <= 1/4 < 1/1.5 (the
// scalings by 4 are exact), so the division is only needed near the equilibrium.
__device__ __forceinline__ unsigned eq_decide_fast(float v, float q, float ke, float F) {
  const bool imp = fabsf(v) > IMPACT;
  const bool fwd = v > 0.0f;
  bool below, above;
  if (ke >= 1e-37f && ke <= 1e37f) {
    above = q > 4.0f * ke;
    below = q < 0.25f * ke;
    if (!above && !below) {
      const float r = q / ke;
      below = r < LOWER;
      above = r > UPPER;
    }
  } else {
    const float r = q / ke;
    below = r < LOWER;
    above = r > UPPER;
  }
  bool lo = fwd ? below : above;
  if (fwd && F == 1.0f) lo = false;
  bool hi = fwd ? above : below;
  if (!fwd && F == 0.0f) hi = false;
  return (hi ? 1u : 0u) | (lo ? 2u : 0u) | (imp ? 4u : 0u);
}

// kinetics.py:850-853
__device__ __forceinline__ float eq_update(float F, unsigned dec, float inc) {
  if (dec & 1u) F = F - inc;
  if (dec & 2u) F = F + inc;
  if (F > 1.0f) F = 1.0f;
  if (F < 0.0f) F = 0.0f;
  return F;
}

__device__ __forceinline__ float eq_increment(int k) {
  // 0.5, 0.25, 0.125, 0.0625, 0.0625, ... (kinetics.py:846: 0.5 ** (k + 1), at most 4 iterations)
  return __int_as_float((126 - (k < 3 ? k : 3)) << 23);
}

// streaming 32-bit loads that do not pollute L1 (every parameter is read once)
__device__ __forceinline__ int ldg_stream(const int *p) {
  int v;
  asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ float ldg_stream(const float *p) {
  float v;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}

}  // namespace ms

// host-side helpers ---------------------------------------------------------
extern "C" void ms_set_cuda_error_(cudaError_t e);

#define MS_CUDA_CHECK(expr)                \
  do {                                     \
    cudaError_t e__ = (expr);              \
    if (e__ != cudaSuccess) {              \
      ms_set_cuda_error_(e__);             \
      return MS_ERR_CUDA;                  \
    }                                      \
  } while (0)
