// Per-node physics in registers: moments, quadratic equilibrium, BGK / TRT / KBC relaxation,
// bounce-back and equilibrium boundaries.  One call handles ONE lattice node whose Q
// populations live in a register array.  All loops are unrolled at compile time so the
// stencil constants fold into immediates.
//
// Arithmetic follows the reference's torch path op for op where it matters for parity
// (SURVEY appendix A.7-A.11); file:line citations are relative to the reference checkout.
#pragma once
#include "lb_stencil.cuh"

namespace lb {

template <typename T>
struct CollisionParams {
  T omega;        // BGK: 1/tau (bgk_collision.py:22: `1.0 / self.tau` is a double, demoted to dtype)
  T inv_2tau_p;   // TRT: 1/(2 tau_plus)   (trt_collision.py:20-22)
  T inv_2tau_m;   // TRT: 1/(2 tau_minus)  (trt_collision.py:23-25)
  T beta;         // KBC: 1/(2 tau)        (kbc_collision.py:98-99)
  T inv_beta;     // KBC: 1/beta as the reference evaluates `1. / self.beta` (:153)
};

// rho = sum_q f_q (_flow.py:136-138), j = sum_q e_q f_q (:152-155), u = j / rho (:161-162)
template <class S, typename T>
__device__ __forceinline__ void moments(const T (&f)[S::Q], T& rho, T (&u)[3]) {
  T r = f[0];
  T jx = T(0), jy = T(0), jz = T(0);
  static_for<S::Q>([&](auto qc) {
    constexpr int q = decltype(qc)::value;
    if constexpr (q > 0) r += f[q];
    if constexpr (S::e(q, 0) == 1) jx += f[q];
    if constexpr (S::e(q, 0) == -1) jx -= f[q];
    if constexpr (S::e(q, 1) == 1) jy += f[q];
    if constexpr (S::e(q, 1) == -1) jy -= f[q];
    if constexpr (S::e(q, 2) == 1) jz += f[q];
    if constexpr (S::e(q, 2) == -1) jz -= f[q];
  });
  rho = r;
  // true division, as torch does
  if constexpr (S::D == 3) u[0] = jx / r; else u[0] = T(0);  // 2-D runs with X == 1
  u[1] = jy / r;
  u[2] = jz / r;
}

// quadratic_equilibrium.py:11-25:
//   feq_q = w_q * (rho * ((2 exu - uxu) / (2 cs^2) + 0.5 (exu / cs^2)^2 + 1))
// Division by the two constants is done as multiplication by their (double-rounded-to-T)
// reciprocals: <= 1 ulp away from the reference's division, well inside the parity budget.
template <class S, typename T>
__device__ __forceinline__ void equilibrium(T rho, const T (&u)[3], T (&feq)[S::Q]) {
  constexpr T inv_cs2 = T(1.0 / LB_CS2);
  constexpr T inv_2cs2 = T(1.0 / (2.0 * LB_CS2));
  const T uxu = u[0] * u[0] + u[1] * u[1] + u[2] * u[2];
  static_for<S::Q>([&](auto qc) {
    constexpr int q = decltype(qc)::value;
    T exu = T(0);
    if constexpr (S::e(q, 0) == 1) exu += u[0];
    if constexpr (S::e(q, 0) == -1) exu -= u[0];
    if constexpr (S::e(q, 1) == 1) exu += u[1];
    if constexpr (S::e(q, 1) == -1) exu -= u[1];
    if constexpr (S::e(q, 2) == 1) exu += u[2];
    if constexpr (S::e(q, 2) == -1) exu -= u[2];
    const T a = exu * inv_cs2;
    const T t = (T(2) * exu - uxu) * inv_2cs2 + T(0.5) * (a * a) + T(1);
    feq[q] = T(S::w(q)) * (rho * t);
  });
}

// ---------------------------------------------------------------------------------------
// KBC helper: the `s` part of the moment representation (kbc_collision.py:25-94).
// The reference builds all 27 (9) raw moments with an einsum; only rho and the six (three)
// second moments enter s, so only those are formed here.  s has 10 (2-D: 5) distinct values:
//   3-D  s0 | sA (q 1,2) | sB (3,4) | sC (5,6) | +-pyz (7,8 | 9,10) | +-pxz (11,12 | 13,14)
//        | +-pxy (15,16 | 17,18) | 0 for the corner populations 19..26     (:54-72)
//   2-D  s0 | sA (q 1,3) | sB (2,4) | +-pxy (5,7 | 6,8)                     (:84-92)
// ---------------------------------------------------------------------------------------
template <typename T>
struct KbcS {
  T s0, sA, sB, sC, pyz, pxz, pxy;
};

template <typename T> __device__ __forceinline__ T fast_rcp(T x) { return T(1) / x; }
// 1 ulp reciprocal (MUFU.RCP): used only for the two sums that set gamma, where the reference's
// own result is no better conditioned than that (operands are O(w_q), never denormal)
template <> __device__ __forceinline__ float fast_rcp<float>(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));   // one MUFU.RCP, no range fix-up
  return r;
}

// s from rho and the second moments ALREADY divided by rho
template <class S, typename T>
__device__ __forceinline__ KbcS<T> kbc_s_from_moments(T r, T mxx, T myy, T mzz, T mxy, T mxz, T myz) {
  KbcS<T> s;
  if constexpr (S::D == 3) {
    // kbc_collision.py:47-72
    const T Tr = mxx + myy + mzz;
    const T Nxz = mxx - mzz;
    const T Nyz = myy - mzz;
    const T sixth = T(1. / 6.), quarter = T(1. / 4);
    s.s0 = r * -Tr;
    s.sA = sixth * r * (T(2) * Nxz - Nyz + Tr);
    s.sB = sixth * r * (T(2) * Nyz - Nxz + Tr);
    s.sC = sixth * r * (-Nxz - Nyz + Tr);
    s.pyz = quarter * r * myz;
    s.pxz = quarter * r * mxz;
    s.pxy = quarter * r * mxy;
  } else {
    // kbc_collision.py:79-92
    const T Tr = mxx + myy;
    const T N = mxx - myy;
    s.s0 = r * -Tr;
    s.sA = T(0.5) * r * (T(0.5) * (Tr + N));
    s.sB = T(0.5) * r * (T(0.5) * (Tr - N));
    s.sC = T(0);
    s.pyz = T(0);
    s.pxz = T(0);
    s.pxy = T(0.25) * r * mxy;
  }
  return s;
}

// s(f): second moments by summation, m = m / rho (kbc_collision.py:27-30, 36-39)
template <class S, typename T>
__device__ __forceinline__ KbcS<T> kbc_s(const T (&g)[S::Q]) {
  T r = T(0), mxx = T(0), myy = T(0), mzz = T(0), mxy = T(0), mxz = T(0), myz = T(0);
  static_for<S::Q>([&](auto qc) {
    constexpr int q = decltype(qc)::value;
    // lattice axes (x, y[, z]) -> kernel axes
    constexpr int ex = S::e(q, S::axis(0));
    constexpr int ey = S::e(q, S::axis(1));
    constexpr int ez = (S::D == 3) ? S::e(q, S::axis(S::D == 3 ? 2 : 0)) : 0;
    r += g[q];
    if constexpr (ex * ex == 1) mxx += g[q];
    if constexpr (ey * ey == 1) myy += g[q];
    if constexpr (ez * ez == 1) mzz += g[q];
    if constexpr (ex * ey == 1) mxy += g[q];
    if constexpr (ex * ey == -1) mxy -= g[q];
    if constexpr (ex * ez == 1) mxz += g[q];
    if constexpr (ex * ez == -1) mxz -= g[q];
    if constexpr (ey * ez == 1) myz += g[q];
    if constexpr (ey * ez == -1) myz -= g[q];
  });
  const T ri = T(1) / r;   // one reciprocal, six products
  return kbc_s_from_moments<S, T>(r, mxx * ri, myy * ri, mzz * ri, mxy * ri, mxz * ri, myz * ri);
}

// s(feq) in closed form.  The reference sums e_a e_b feq_q numerically (kbc_collision.py:146-147);
// for the quadratic equilibrium on D2Q9 / D3Q27 those sums are, identically,
//   sum_q feq_q = rho   and   sum_q e_a e_b feq_q / rho = cs^2 delta_ab + u_a u_b ,
// so the closed form differs from the reference's sums only by their round-off (a few ulp of
// rho).  NOT used by default: measured on B200 it did not make the (latency-bound) KBC kernel
// faster, and one fp32 parity case moved from 2e-6 to 1.4e-5 relative, outside the 1e-5 bar.
template <class S, typename T>
__device__ __forceinline__ KbcS<T> kbc_s_equilibrium(T rho, const T (&u)[3]) {
  constexpr T cs2 = T(LB_CS2);
  const T ux = u[S::axis(0)], uy = u[S::axis(1)];
  const T uz = (S::D == 3) ? u[S::axis(S::D == 3 ? 2 : 0)] : T(0);
  return kbc_s_from_moments<S, T>(rho, cs2 + ux * ux, cs2 + uy * uy, (S::D == 3) ? cs2 + uz * uz : T(0),
                                  ux * uy, ux * uz, uy * uz);
}

// delta_s of population q from the distinct differences
template <class S, typename T, int q>
__device__ __forceinline__ T kbc_ds(const KbcS<T>& d) {
  if constexpr (q == 0) return d.s0;
  if constexpr (S::D == 3) {
    if constexpr (q == 1 || q == 2) return d.sA;
    else if constexpr (q == 3 || q == 4) return d.sB;
    else if constexpr (q == 5 || q == 6) return d.sC;
    else if constexpr (q == 7 || q == 8) return d.pyz;
    else if constexpr (q == 9 || q == 10) return -d.pyz;
    else if constexpr (q == 11 || q == 12) return d.pxz;
    else if constexpr (q == 13 || q == 14) return -d.pxz;
    else if constexpr (q == 15 || q == 16) return d.pxy;
    else if constexpr (q == 17 || q == 18) return -d.pxy;
    else return T(0);
  } else {
    if constexpr (q == 1 || q == 3) return d.sA;
    else if constexpr (q == 2 || q == 4) return d.sB;
    else if constexpr (q == 5 || q == 7) return d.pxy;
    else return -d.pxy;
  }
}

enum { COLL_NONE = 0, COLL_BGK = 1, COLL_TRT = 2, COLL_KBC = 3 };

// f <- collision(f); rho, u are the moments of the incoming f (already computed).
template <class S, typename T, int COLL>
__device__ __forceinline__ void collide_node(T (&f)[S::Q], T rho, const T (&u)[3],
                                             const CollisionParams<T>& cp) {
  if constexpr (COLL == COLL_NONE) {
    return;  // no_collision.py:9-11
  } else {
    T feq[S::Q];
    equilibrium<S, T>(rho, u, feq);
    if constexpr (COLL == COLL_BGK) {
      // bgk_collision.py:22: f - 1.0/tau * (f - feq)
      static_for<S::Q>([&](auto qc) {
        constexpr int q = decltype(qc)::value;
        f[q] = f[q] - cp.omega * (f[q] - feq[q]);
      });
    } else if constexpr (COLL == COLL_TRT) {
      // trt_collision.py:20-26, evaluated once per opposite pair (q, o): with
      //   sp = ((f_q + f_o) - (feq_q + feq_o)) / (2 tau+),  sm = ((f_q - f_o) - (feq_q - feq_o)) / (2 tau-)
      // the reference's values are f_q - (sp + sm) and f_o - (sp - sm): identical bit for bit, since
      // the symmetric part is commutative and the antisymmetric part changes sign exactly.
      static_for<S::Q>([&](auto qc) {
        constexpr int q = decltype(qc)::value;
        constexpr int o = S::opp(q);
        if constexpr (q <= o) {
          const T sp = ((f[q] + f[o]) - (feq[q] + feq[o])) * cp.inv_2tau_p;
          const T sm = ((f[q] - f[o]) - (feq[q] - feq[o])) * cp.inv_2tau_m;
          const T fq = f[q], fo = f[o];
          f[q] = fq - (sp + sm);
          if constexpr (q < o) f[o] = fo - (sp - sm);
        }
      });
    } else {
      // kbc_collision.py:131-158
      static_assert(S::ID != 1, "KBC is defined for D2Q9 and D3Q27 only (kbc_collision.py:101,115)");
      const KbcS<T> sf = kbc_s<S, T>(f);
#ifdef LB_KBC_ANALYTIC_FEQ   // experiment only: 10 % fewer instructions, one parity outlier
      const KbcS<T> se = kbc_s_equilibrium<S, T>(rho, u);
#else
      const KbcS<T> se = kbc_s<S, T>(feq);
#endif
      KbcS<T> d;   // delta_s = s(f) - s(feq), per distinct value
      d.s0 = sf.s0 - se.s0; d.sA = sf.sA - se.sA; d.sB = sf.sB - se.sB; d.sC = sf.sC - se.sC;
      d.pyz = sf.pyz - se.pyz; d.pxz = sf.pxz - se.pxz; d.pxy = sf.pxy - se.pxy;
      T sum_s = T(0), sum_h = T(0);
      T dh[S::Q];
      static_for<S::Q>([&](auto qc) {
        constexpr int q = decltype(qc)::value;
        const T ds = kbc_ds<S, T, q>(d);
        dh[q] = f[q] - feq[q] - ds;
        const T w = dh[q] * fast_rcp<T>(feq[q]);   // (ds dh / feq, dh dh / feq) share 1 / feq
        sum_s += ds * w;
        sum_h += dh[q] * w;
      });
      T gamma = cp.inv_beta - (T(2) - cp.inv_beta) * sum_s / sum_h;
      if (gamma < T(1E-15)) gamma = T(2);   // :154
      if (gamma != gamma) gamma = T(2);     // NaN guard :156
      static_for<S::Q>([&](auto qc) {
        constexpr int q = decltype(qc)::value;
        f[q] = f[q] - cp.beta * (T(2) * kbc_ds<S, T, q>(d) + gamma * dh[q]);
      });
    }
  }
}

// bounce_back_boundary.py:17-18: f <- f[opposite]
template <class S, typename T>
__device__ __forceinline__ void bounce_back_node(T (&f)[S::Q]) {
  static_for<S::Q>([&](auto qc) {
    constexpr int q = decltype(qc)::value;
    constexpr int o = S::opp(q);
    if constexpr (q < o) {
      const T t = f[q];
      f[q] = f[o];
      f[o] = t;
    }
  });
}

}  // namespace lb
