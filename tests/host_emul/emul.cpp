// TEST TOOLING ONLY -- never part of the product, never loaded by lettuce_b200.
//
// Compiles the device arithmetic headers (lettuce_b200/csrc/lb_physics.cuh and friends) with the
// host C++ compiler so that the round-off behaviour of the collision formulas can be studied
// against the reference's golden vectors without a GPU.  Because every fused multiply-add in
// those headers is explicit (fma_) and the CUDA build uses -fmad=false, the host build performs the
// same IEEE operations in the same order as the kernels (fp64: bit for bit; fp32: up to the 1-ulp
// MUFU.RCP inside KBC).  Build: g++ -O2 -std=c++17 -mfma -ffp-contract=off -shared -fPIC.
#include "../../lettuce_b200/csrc/lb_physics.cuh"

using namespace lb;

template <class S, typename T, typename R, int COLL>
static void run(T* f, long n, const double* cpd) {
  CollisionParams<T> cp;
  cp.omega = T(cpd[0]); cp.inv_2tau_p = T(cpd[1]); cp.inv_2tau_m = T(cpd[2]); cp.beta = T(cpd[3]); cp.inv_beta = T(cpd[4]);
  cp.reg_c = T(cpd[5]); cp.smag_tau = T(cpd[6]); cp.smag_nu = T(cpd[7]); cp.smag_c2 = T(cpd[8]);
  for (int c = 0; c < 3; ++c) { cp.force_a[c] = T(cpd[9 + c]); cp.force_ueq[c] = T(cpd[12 + c]); }
  cp.force_src = T(cpd[15]);
  constexpr int L = ValueTraits<R>::lanes;
  long i = 0;
  for (; i + L <= n; i += L) {
    R g[S::Q];
    for (int q = 0; q < S::Q; ++q) {
      if constexpr (L == 2) g[q] = R(f[q * n + i], f[q * n + i + 1]);
      else g[q] = f[q * n + i];
    }
    collide_node<S, R, COLL, T>(g, cp);
    for (int q = 0; q < S::Q; ++q) {
      if constexpr (L == 2) { f[q * n + i] = g[q].x; f[q * n + i + 1] = g[q].y; }
      else f[q * n + i] = g[q];
    }
  }
  for (; i < n; ++i) {   // odd tail of the packed path
    T g[S::Q];
    for (int q = 0; q < S::Q; ++q) g[q] = f[q * n + i];
    collide_node<S, T, COLL, T>(g, cp);
    for (int q = 0; q < S::Q; ++q) f[q * n + i] = g[q];
  }
}

template <class S, typename T, typename R>
static int by_coll(int coll, T* f, long n, const double* cp) {
  switch (coll) {
    case COLL_BGK: run<S, T, R, COLL_BGK>(f, n, cp); return 0;
    case COLL_TRT: run<S, T, R, COLL_TRT>(f, n, cp); return 0;
    case COLL_KBC:
      if constexpr (S::ID != 1) { run<S, T, R, COLL_KBC>(f, n, cp); return 0; }
      return 1;
    case COLL_REGULARIZED: run<S, T, R, COLL_REGULARIZED>(f, n, cp); return 0;
    case COLL_SMAGORINSKY: run<S, T, R, COLL_SMAGORINSKY>(f, n, cp); return 0;
    case COLL_BGK_FORCE: run<S, T, R, COLL_BGK_FORCE>(f, n, cp); return 0;
  }
  return 1;
}

template <class S>
static int by_type(int dtype, int packed, int coll, void* f, long n, const double* cp) {
  if (dtype == 1) return by_coll<S, double, double>(coll, static_cast<double*>(f), n, cp);
  if (packed) return by_coll<S, float, F2>(coll, static_cast<float*>(f), n, cp);
  return by_coll<S, float, float>(coll, static_cast<float*>(f), n, cp);
}

// stencil: 0 D2Q9, 1 D3Q19, 2 D3Q27; dtype: 0 f32, 1 f64; f: [Q][n] in kernel order, collided in place
extern "C" int emul_collide(int stencil, int dtype, int packed, int coll, void* f, long n, const double* cp) {
  switch (stencil) {
    case 0: return by_type<D2Q9>(dtype, packed, coll, f, n, cp);
    case 1: return by_type<D3Q19>(dtype, packed, coll, f, n, cp);
    case 2: return by_type<D3Q27>(dtype, packed, coll, f, n, cp);
  }
  return 1;
}

// moments of every node: rho [n], u [3][n] (kernel axes)
extern "C" int emul_moments(int stencil, int dtype, const void* fv, long n, void* rho_v, void* u_v) {
  auto go = [&](auto s, auto t) {
    using S = decltype(s); using T = decltype(t);
    const T* f = static_cast<const T*>(fv);
    T* rho = static_cast<T*>(rho_v); T* u = static_cast<T*>(u_v);
    for (long i = 0; i < n; ++i) {
      T g[S::Q], r, uu[3];
      for (int q = 0; q < S::Q; ++q) g[q] = f[q * n + i];
      moments<S, T>(g, r, uu);
      rho[i] = r; u[i] = uu[0]; u[n + i] = uu[1]; u[2 * n + i] = uu[2];
    }
  };
  auto st = [&](auto t) {
    switch (stencil) { case 0: go(D2Q9{}, t); return 0; case 1: go(D3Q19{}, t); return 0; case 2: go(D3Q27{}, t); return 0; }
    return 1;
  };
  return dtype == 1 ? st(double{}) : st(float{});
}

// equilibrium of (rho, u) on every node: rho [n], u [3][n] (kernel axes) -> f [Q][n]; the function the step
// kernel calls for LB_BND_EQUILIBRIUM nodes (lb_kernels.cuh node_update)
extern "C" int emul_equilibrium(int stencil, int dtype, long n, const void* rho_v, const void* u_v, void* f_v) {
  auto go = [&](auto s, auto t) {
    using S = decltype(s); using T = decltype(t);
    const T* rho = static_cast<const T*>(rho_v); const T* u = static_cast<const T*>(u_v);
    T* f = static_cast<T*>(f_v);
    for (long i = 0; i < n; ++i) {
      T g[S::Q];
      const T uu[3] = {u[i], u[n + i], u[2 * n + i]};
      equilibrium<S, T>(rho[i], uu, g);
      for (int q = 0; q < S::Q; ++q) f[q * n + i] = g[q];
    }
  };
  auto st = [&](auto t) {
    switch (stencil) { case 0: go(D2Q9{}, t); return 0; case 1: go(D3Q19{}, t); return 0; case 2: go(D3Q27{}, t); return 0; }
    return 1;
  };
  return dtype == 1 ? st(double{}) : st(float{});
}

// bounce-back of every node in place, f [Q][n] (lb_physics.cuh bounce_back_node)
extern "C" int emul_bounce_back(int stencil, int dtype, long n, void* f_v) {
  auto go = [&](auto s, auto t) {
    using S = decltype(s); using T = decltype(t);
    T* f = static_cast<T*>(f_v);
    for (long i = 0; i < n; ++i) {
      T g[S::Q];
      for (int q = 0; q < S::Q; ++q) g[q] = f[q * n + i];
      bounce_back_node<S, T>(g);
      for (int q = 0; q < S::Q; ++q) f[q * n + i] = g[q];
    }
  };
  auto st = [&](auto t) {
    switch (stencil) { case 0: go(D2Q9{}, t); return 0; case 1: go(D3Q19{}, t); return 0; case 2: go(D3Q27{}, t); return 0; }
    return 1;
  };
  return dtype == 1 ? st(double{}) : st(float{});
}
