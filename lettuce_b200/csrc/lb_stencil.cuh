// Lattice constants as compile-time tables.
// Index order and values follow the reference's stencils (reference/lettuce/ext/_stencil/
// d2q9.py:8-10, d3q19.py:8-13, d3q27.py:8-12); cs = 1/np.sqrt(3.0) (lettuce/_stencil.py:17).
//
// Kernel coordinates are always (X, Y, Z) with Z the contiguous axis.  A 2-D lattice
// f[q][x][y] is run as X = 1, Y = x, Z = y, so D2Q9's (ex, ey) become (0, ex, ey).
#pragma once
#include <cstdint>
#include <type_traits>

// tests/host_emul compiles the arithmetic headers (lb_stencil.cuh, lb_values.cuh, lb_physics.cuh)
// with a plain C++ compiler to study round-off of formula variants on the CPU: TEST TOOLING ONLY,
// nothing in the product ever runs these functions on the host.
#if !defined(__CUDACC__)
#define __host__
#define __device__
#define __forceinline__ inline __attribute__((always_inline))
#endif

namespace lb {

// (1/np.sqrt(3.0))**2 in double: 0.3333333333333334 (NOT 1/3), SURVEY appendix A.7
#define LB_CS2 0.33333333333333337034

template <int N, class F>
__host__ __device__ __forceinline__ void static_for_impl(F&&, std::integral_constant<int, N>,
                                                         std::integral_constant<int, N>) {}
template <int I, int N, class F>
__host__ __device__ __forceinline__ void static_for_impl(F&& f, std::integral_constant<int, I>,
                                                         std::integral_constant<int, N>) {
  f(std::integral_constant<int, I>{});
  static_for_impl(f, std::integral_constant<int, I + 1>{}, std::integral_constant<int, N>{});
}
// static_for<N>([&](auto qc) { constexpr int q = decltype(qc)::value; ... });
template <int N, class F>
__host__ __device__ __forceinline__ void static_for(F&& f) {
  static_for_impl(f, std::integral_constant<int, 0>{}, std::integral_constant<int, N>{});
}

struct D2Q9 {
  static constexpr int ID = 0, Q = 9, D = 2;
  __host__ __device__ static constexpr int e(int q, int a) {  // a: 0=X 1=Y 2=Z (kernel axes)
    constexpr int t[9][3] = {{0, 0, 0}, {0, 1, 0}, {0, 0, 1}, {0, -1, 0}, {0, 0, -1},
                             {0, 1, 1}, {0, -1, 1}, {0, -1, -1}, {0, 1, -1}};
    return t[q][a];
  }
  // lattice axis c (0..D-1) -> kernel axis
  __host__ __device__ static constexpr int axis(int c) { return c + 1; }
  __host__ __device__ static constexpr double w(int q) {
    return q == 0 ? 4.0 / 9.0 : (q < 5 ? 1.0 / 9.0 : 1.0 / 36.0);
  }
  __host__ __device__ static constexpr int opp(int q) {
    constexpr int t[9] = {0, 3, 4, 1, 2, 7, 8, 5, 6};
    return t[q];
  }
};

struct D3Q27 {
  static constexpr int ID = 2, Q = 27, D = 3;
  __host__ __device__ static constexpr int e(int q, int a) {
    constexpr int t[27][3] = {
        {0, 0, 0},  {1, 0, 0},  {-1, 0, 0},  {0, 1, 0},   {0, -1, 0},  {0, 0, 1},   {0, 0, -1},
        {0, 1, 1},  {0, -1, -1}, {0, 1, -1}, {0, -1, 1},  {1, 0, 1},   {-1, 0, -1}, {1, 0, -1},
        {-1, 0, 1}, {1, 1, 0},  {-1, -1, 0}, {1, -1, 0},  {-1, 1, 0},  {1, 1, 1},   {-1, -1, -1},
        {1, 1, -1}, {-1, -1, 1}, {1, -1, 1}, {-1, 1, -1}, {1, -1, -1}, {-1, 1, 1}};
    return t[q][a];
  }
  __host__ __device__ static constexpr int axis(int c) { return c; }
  __host__ __device__ static constexpr double w(int q) {
    return q == 0 ? 8.0 / 27.0 : (q < 7 ? 2.0 / 27.0 : (q < 19 ? 1.0 / 54.0 : 1.0 / 216.0));
  }
  __host__ __device__ static constexpr int opp(int q) {
    return q == 0 ? 0 : ((q & 1) ? q + 1 : q - 1);
  }
};

struct D3Q19 {
  static constexpr int ID = 1, Q = 19, D = 3;
  // the first 19 directions of D3Q27 (d3q19.py:8-13 vs d3q27.py:8-10)
  __host__ __device__ static constexpr int e(int q, int a) { return D3Q27::e(q, a); }
  __host__ __device__ static constexpr int axis(int c) { return c; }
  __host__ __device__ static constexpr double w(int q) {
    return q == 0 ? 1.0 / 3.0 : (q < 7 ? 1.0 / 18.0 : 1.0 / 36.0);
  }
  __host__ __device__ static constexpr int opp(int q) { return D3Q27::opp(q); }
};

}  // namespace lb
