// Register value types of the per-node arithmetic.
//
//   float / double : one lattice node per value.
//   F2             : TWO fp32 nodes packed in a 64-bit register pair.  On sm_100a every
//                    +, -, * and fma of an F2 is ONE packed instruction (FADD2 / FMUL2 / FFMA2;
//                    operand negation and broadcast immediates are free), so a thread that owns
//                    two z-adjacent nodes issues half the floating-point instructions.  Each lane
//                    of a packed instruction rounds exactly like the scalar instruction
//                    (round-to-nearest-even, no flush), so packing does not change any result.
//
// The physics in lb_physics.cuh is written once against this small vocabulary:
//   R(c)  broadcast constant     fma_(a, b, c)  a * b + c with ONE rounding
//   div_(a, b) IEEE division     rcp_fast(a)    1-ulp reciprocal (MUFU.RCP for fp32)
//   lt_select(a, c, v)  a < c ? v : a      nan_select(a, v)  isnan(a) ? v : a
//   max_(a, b)          hsum(a) / hmax(a)  fold the lanes of a value into a scalar
#pragma once
#include "lb_stencil.cuh"

#if defined(__CUDACC__)
#include <cuda_runtime.h>
#else
#include <cmath>
#endif

namespace lb {

struct alignas(8) F2 {
  float x, y;
  F2() = default;
  __host__ __device__ constexpr F2(float a, float b) : x(a), y(b) {}
  template <class A, class = std::enable_if_t<std::is_arithmetic<A>::value>>
  __host__ __device__ constexpr explicit F2(A c) : x(float(c)), y(float(c)) {}
};

// scalar type underneath a value type, and the number of nodes it carries
template <typename R> struct ValueTraits { using scalar = R; static constexpr int lanes = 1; };
template <> struct ValueTraits<F2> { using scalar = float; static constexpr int lanes = 2; };

#if defined(__CUDACC__)
// -0.0 that ptxas cannot see through (constant memory may be rewritten by the host): operator* below
static __constant__ float2 LB_NEG_ZERO2 = {-0.0f, -0.0f};
#endif

#if defined(__CUDA_ARCH__)
__device__ __forceinline__ F2 operator+(F2 a, F2 b) {
  const float2 r = __fadd2_rn(make_float2(a.x, a.y), make_float2(b.x, b.y));
  return F2(r.x, r.y);
}
__device__ __forceinline__ F2 operator-(F2 a, F2 b) {
  const float2 r = __fadd2_rn(make_float2(a.x, a.y), make_float2(-b.x, -b.y));
  return F2(r.x, r.y);
}
__device__ __forceinline__ F2 operator*(F2 a, F2 b) {
  // a * b as fma(a, b, -0) with a -0 from constant memory: the same value as mul.rn (x * y + (-0) rounds x * y once
  // and keeps the sign of a zero product), but ptxas can neither fold it back into a multiply nor contract it with
  // a following add.  It DOES contract mul.rn.f32x2 + add.rn.f32x2 into FFMA2 -- despite the explicit rounding
  // modifiers and -fmad=false, and unlike the scalar mul.rn / add.rn -- wherever a product feeds additions, which
  // made the packed kernels round differently from the one-node-per-thread kernels (found at 8 GPUs, where 48 / 8 =
  // 6 planes per slab run one node per thread: scripts/dev/scalar_vs_packed.cu, scripts/check_vec_width.py,
  // profiles/r2_vector_width_dependence_*.log).
  const float2 r = __ffma2_rn(make_float2(a.x, a.y), make_float2(b.x, b.y), LB_NEG_ZERO2);
  return F2(r.x, r.y);
}
__device__ __forceinline__ F2 fma_(F2 a, F2 b, F2 c) {
  const float2 r = __ffma2_rn(make_float2(a.x, a.y), make_float2(b.x, b.y), make_float2(c.x, c.y));
  return F2(r.x, r.y);
}
#else
// host emulation of the packed instructions (tests/host_emul only)
__forceinline__ F2 operator+(F2 a, F2 b) { return F2(a.x + b.x, a.y + b.y); }
__forceinline__ F2 operator-(F2 a, F2 b) { return F2(a.x - b.x, a.y - b.y); }
__forceinline__ F2 operator*(F2 a, F2 b) { return F2(a.x * b.x, a.y * b.y); }
__forceinline__ F2 fma_(F2 a, F2 b, F2 c) { return F2(std::fma(a.x, b.x, c.x), std::fma(a.y, b.y, c.y)); }
#endif
__host__ __device__ __forceinline__ F2 operator-(F2 a) { return F2(-a.x, -a.y); }
__host__ __device__ __forceinline__ F2& operator+=(F2& a, F2 b) { a = a + b; return a; }
__host__ __device__ __forceinline__ F2& operator-=(F2& a, F2 b) { a = a - b; return a; }

// scalars: a * b + c is contracted to one FMA by nvcc (default -fmad=true); spelled out so that
// host emulation and the packed type round the same way
__host__ __device__ __forceinline__ float fma_(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
  return __fmaf_rn(a, b, c);
#else
  return std::fma(a, b, c);
#endif
}
__host__ __device__ __forceinline__ double fma_(double a, double b, double c) {
#if defined(__CUDA_ARCH__)
  return __fma_rn(a, b, c);
#else
  return std::fma(a, b, c);
#endif
}

__host__ __device__ __forceinline__ float div_(float a, float b) { return a / b; }
__host__ __device__ __forceinline__ double div_(double a, double b) { return a / b; }
__host__ __device__ __forceinline__ F2 div_(F2 a, F2 b) { return F2(a.x / b.x, a.y / b.y); }

// 1-ulp reciprocal (one MUFU.RCP, no range fix-up) for fp32, exact division for fp64
__host__ __device__ __forceinline__ float rcp_fast(float x) {
#if defined(__CUDA_ARCH__) && !defined(LB_EXACT_RCP)   // -DLB_EXACT_RCP: IEEE division (scripts/kbc_rcp_count.py, A/B only)
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
#else
  return 1.0f / x;
#endif
}
__host__ __device__ __forceinline__ double rcp_fast(double x) { return 1.0 / x; }
__host__ __device__ __forceinline__ F2 rcp_fast(F2 a) { return F2(rcp_fast(a.x), rcp_fast(a.y)); }

__host__ __device__ __forceinline__ float lt_select(float a, float c, float v) { return a < c ? v : a; }
__host__ __device__ __forceinline__ double lt_select(double a, double c, double v) { return a < c ? v : a; }
__host__ __device__ __forceinline__ F2 lt_select(F2 a, float c, float v) {
  return F2(a.x < c ? v : a.x, a.y < c ? v : a.y);
}
__host__ __device__ __forceinline__ float nan_select(float a, float v) { return a != a ? v : a; }
__host__ __device__ __forceinline__ double nan_select(double a, double v) { return a != a ? v : a; }
__host__ __device__ __forceinline__ F2 nan_select(F2 a, float v) {
  return F2(a.x != a.x ? v : a.x, a.y != a.y ? v : a.y);
}

__host__ __device__ __forceinline__ float max_(float a, float b) { return a > b ? a : b; }
__host__ __device__ __forceinline__ double max_(double a, double b) { return a > b ? a : b; }
__host__ __device__ __forceinline__ F2 max_(F2 a, F2 b) { return F2(max_(a.x, b.x), max_(a.y, b.y)); }

__host__ __device__ __forceinline__ float hsum(float a) { return a; }
__host__ __device__ __forceinline__ double hsum(double a) { return a; }
__host__ __device__ __forceinline__ float hsum(F2 a) { return a.x + a.y; }
__host__ __device__ __forceinline__ float hmax(float a) { return a; }
__host__ __device__ __forceinline__ double hmax(double a) { return a; }
__host__ __device__ __forceinline__ float hmax(F2 a) { return max_(a.x, a.y); }

// lane k of a value (k = 0 for scalars)
template <int K> __host__ __device__ __forceinline__ float lane_of(F2 a) { return K == 0 ? a.x : a.y; }
template <int K> __host__ __device__ __forceinline__ float lane_of(float a) { return a; }
template <int K> __host__ __device__ __forceinline__ double lane_of(double a) { return a; }

}  // namespace lb
