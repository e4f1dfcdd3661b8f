// Per-node physics in registers: moments, quadratic equilibrium, BGK / TRT / KBC relaxation,
// bounce-back and equilibrium boundaries.  One call handles ONE value `R` per population:
// R = float / double (one node) or F2 (two fp32 nodes in packed FADD2/FMUL2/FFMA2 instructions,
// lb_values.cuh).  All loops are unrolled at compile time so stencil constants fold away.
//
// Opposite-pair formulation.  Every stencil here is symmetric, so the Q-1 moving populations
// form (Q-1)/2 opposite pairs (q, o).  With a_p = f_q + f_o and b_p = f_q - f_o
//   rho = f_0 + sum a_p,   j = sum e_q b_p,   sum e e f = sum e_q e_q a_p        (even / odd moments)
//   feq_q = S_p + D_p,  feq_o = S_p - D_p   with   S_p = w rho (1 - u.u/(2cs^2) + (e.u/cs^2)^2 / 2),
//                                                  D_p = w rho (e.u/cs^2)
// which is the reference's feq (quadratic_equilibrium.py:11-25) split into its even and odd part
// in e.  Half the additions of the moment sums and a third of the equilibrium arithmetic of the
// literal population-by-population evaluation; TRT's symmetric / antisymmetric parts
// (trt_collision.py:20-25) are a_p - 2 S_p and b_p - 2 D_p directly.
//
// Every multiply-add that is meant to be fused is written fma_(); the library is compiled with
// -fmad=false, so scalar and packed instantiations (and the host emulation in tests/host_emul) round
// identically: the same node gives the same bits whichever vector width the launcher picks.
//
// Differences to the reference's torch arithmetic are at round-off level (it sums in torch's
// order, divides by cs^2 where a reciprocal is multiplied here, ...) and inside the parity budget
// (1e-12 fp64 / 1e-5 fp32); file:line citations are relative to the reference checkout.
#pragma once
#include "lb_values.cuh"

namespace lb {

template <typename T>
struct CollisionParams {
  T omega;        // BGK: 1/tau (bgk_collision.py:22: `1.0 / self.tau` is a double, demoted to dtype)
  T inv_2tau_p;   // TRT: 1/(2 tau_plus)   (trt_collision.py:20-22)
  T inv_2tau_m;   // TRT: 1/(2 tau_minus)  (trt_collision.py:23-25)
  T beta;         // KBC: 1/(2 tau)        (kbc_collision.py:98-99)
  T inv_beta;     // KBC: 1/beta as the reference evaluates `1. / self.beta` (:153)
  T reg_c;        // Regularized: 1 - 1/tau  (regularized_collision.py:41)
  T smag_tau;     // Smagorinsky: tau, nu = (tau - 0.5) / 3.0, constant ** 2  (smagorinsky_collision.py:27-33)
  T smag_nu;
  T smag_c2;
  T force_a[3];   // BGK + force: acceleration (kernel axes), scale * acceleration of u_eq, source prefactor
  T force_ueq[3];
  T force_src;
};

// ---------------------------------------------------------------------------------------
// Opposite pairs, by direction class ("slot") in kernel axes (X, Y, Z):
//   0 X(1,0,0) 1 Y(0,1,0) 2 Z(0,0,1) | 3 (0,1,1) 4 (0,1,-1) | 5 (1,0,1) 6 (1,0,-1) |
//   7 (1,1,0) 8 (1,-1,0) | 9 (1,1,1) 10 (1,1,-1) 11 (1,-1,1) 12 (1,-1,-1)
// pair_q<S>(slot) = the population with that direction (its partner is S::opp), -1 if absent.
// ---------------------------------------------------------------------------------------
constexpr int N_SLOTS = 13;

__host__ __device__ constexpr int slot_e(int slot, int a) {
  constexpr int t[N_SLOTS][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {0, 1, 1},  {0, 1, -1}, {1, 0, 1},  {1, 0, -1},
                                 {1, 1, 0}, {1, -1, 0}, {1, 1, 1}, {1, 1, -1}, {1, -1, 1}, {1, -1, -1}};
  return t[slot][a];
}
template <class S>
__host__ __device__ constexpr int pair_q(int slot) {
  if (S::Q == 9) {
    constexpr int t[N_SLOTS] = {-1, 1, 2, 5, 8, -1, -1, -1, -1, -1, -1, -1, -1};
    return t[slot];
  }
  return (2 * slot + 1 < S::Q) ? 2 * slot + 1 : -1;
}
template <class S>
__host__ __device__ constexpr bool pair_table_ok() {
  int seen = 1;  // population 0
  for (int s = 0; s < N_SLOTS; ++s) {
    const int q = pair_q<S>(s);
    if (q < 0) continue;
    for (int a = 0; a < 3; ++a)
      if (S::e(q, a) != slot_e(s, a) || S::e(S::opp(q), a) != -slot_e(s, a)) return false;
    seen += 2;
  }
  return seen == S::Q;
}
static_assert(pair_table_ok<D2Q9>() && pair_table_ok<D3Q19>() && pair_table_ok<D3Q27>(), "pair tables");

template <class S, int SLOT> constexpr bool has_slot = pair_q<S>(SLOT) >= 0;

// a_p = f_q + f_o, b_p = f_q - f_o
template <class S, typename R>
__host__ __device__ __forceinline__ void pair_split(const R (&f)[S::Q], R (&a)[N_SLOTS], R (&b)[N_SLOTS]) {
  static_for<N_SLOTS>([&](auto sc) {
    constexpr int s = decltype(sc)::value;
    if constexpr (has_slot<S, s>) {
      constexpr int q = pair_q<S>(s), o = S::opp(q);
      a[s] = f[q] + f[o];
      b[s] = f[q] - f[o];
    }
  });
}

// Even moments of a pair-symmetric quantity v_p (one value per pair):
//   sum = sum_p v_p,  m_AB = sum_p e_A e_B v_p, shared partial sums by direction class.
template <typename R>
struct EvenMoments {
  R sum, xx, yy, zz, xy, xz, yz;   // 2-D (D2Q9 runs as X = 1): only sum, yy, zz, yz are set
};

template <class S, typename R>
__host__ __device__ __forceinline__ EvenMoments<R> even_moments(const R (&v)[N_SLOTS]) {
  EvenMoments<R> m;
  const R eyz = v[3] + v[4];
  m.yz = v[3] - v[4];
  if constexpr (S::D == 2) {
    m.yy = v[1] + eyz;
    m.zz = v[2] + eyz;
    m.sum = (v[1] + v[2]) + eyz;
    m.xx = m.xy = m.xz = R(0);
  } else {
    const R exz = v[5] + v[6];
    const R exy = v[7] + v[8];
    m.xz = v[5] - v[6];
    m.xy = v[7] - v[8];
    m.xx = (v[0] + exz) + exy;
    m.yy = (v[1] + eyz) + exy;
    m.zz = (v[2] + eyz) + exz;
    m.sum = ((v[0] + v[1]) + v[2]) + ((eyz + exz) + exy);
    if constexpr (S::Q == 27) {
      const R c1 = v[9] + v[10], c2 = v[11] + v[12];
      const R c3 = v[9] - v[10], c4 = v[11] - v[12];
      const R c = c1 + c2;
      m.xx += c; m.yy += c; m.zz += c; m.sum += c;
      m.xy += c1 - c2;
      m.xz += c3 + c4;
      m.yz += c3 - c4;
    }
  }
  return m;
}

// first moment of a pair-antisymmetric quantity: j_A = sum_p e_A b_p
template <class S, typename R>
__host__ __device__ __forceinline__ void odd_moments(const R (&b)[N_SLOTS], R (&j)[3]) {
  if constexpr (S::D == 2) {
    j[0] = R(0);
    j[1] = b[1] + (b[3] + b[4]);
    j[2] = b[2] + (b[3] - b[4]);
  } else {
    R jx = (b[0] + (b[5] + b[6])) + (b[7] + b[8]);
    R jy = (b[1] + (b[3] + b[4])) + (b[7] - b[8]);
    R jz = (b[2] + (b[3] - b[4])) + (b[5] - b[6]);
    if constexpr (S::Q == 27) {
      const R d1 = b[9] + b[10], d2 = b[11] + b[12];
      jx += d1 + d2;
      jy += d1 - d2;
      jz += (b[9] - b[10]) + (b[11] - b[12]);
    }
    j[0] = jx; j[1] = jy; j[2] = jz;
  }
}

// rho = sum_q f_q (_flow.py:136-138), j = sum_q e_q f_q (:152-155), u = j / rho (:161-162, true division)
template <class S, typename R>
__host__ __device__ __forceinline__ void moments_from_pairs(R f0, const R (&a)[N_SLOTS], const R (&b)[N_SLOTS],
                                                            R& rho, R (&u)[3]) {
  const EvenMoments<R> m = even_moments<S, R>(a);   // only .sum is used; the rest is dead code
  rho = f0 + m.sum;
  R j[3];
  odd_moments<S, R>(b, j);
  u[0] = (S::D == 3) ? div_(j[0], rho) : R(0);
  u[1] = div_(j[1], rho);
  u[2] = div_(j[2], rho);
}

template <class S, typename R>
__host__ __device__ __forceinline__ void moments(const R (&f)[S::Q], R& rho, R (&u)[3]) {
  R a[N_SLOTS], b[N_SLOTS];
  pair_split<S, R>(f, a, b);
  moments_from_pairs<S, R>(f[0], a, b, rho, u);
}

// e_slot . v  without a zero seed (0 + x is not a no-op in IEEE arithmetic); (vx +- vy) is shared by
// the edge and the corner directions through common-subexpression elimination
template <int SLOT, typename R>
__host__ __device__ __forceinline__ R slot_dot(const R (&v)[3]) {
  constexpr int ex = slot_e(SLOT, 0), ey = slot_e(SLOT, 1), ez = slot_e(SLOT, 2);
  if constexpr (ex != 0 && ey != 0 && ez != 0) return ez > 0 ? (ey > 0 ? (v[0] + v[1]) + v[2] : (v[0] - v[1]) + v[2])
                                                               : (ey > 0 ? (v[0] + v[1]) - v[2] : (v[0] - v[1]) - v[2]);
  else if constexpr (ex != 0 && ey != 0) return ey > 0 ? v[0] + v[1] : v[0] - v[1];
  else if constexpr (ex != 0 && ez != 0) return ez > 0 ? v[0] + v[2] : v[0] - v[2];
  else if constexpr (ey != 0 && ez != 0) return ez > 0 ? v[1] + v[2] : v[1] - v[2];
  else if constexpr (ex != 0) return v[0];
  else if constexpr (ey != 0) return v[1];
  else return v[2];
}

// quadratic_equilibrium.py:11-25 in even / odd form:
//   feq_0 = w_0 rho c0,  S_p = w rho (c0 + a^2 / 2),  D_p = w rho a,   a = e.u / cs^2,  c0 = 1 - u.u / (2 cs^2)
template <class S, typename R>
__host__ __device__ __forceinline__ void equilibrium_pairs(R rho, const R (&u)[3], R& feq0, R (&Sp)[N_SLOTS],
                                                           R (&Dp)[N_SLOTS]) {
  const R inv_cs2 = R(1.0 / LB_CS2);
  const R U[3] = {u[0] * inv_cs2, u[1] * inv_cs2, u[2] * inv_cs2};
  R uxu = fma_(u[2], u[2], u[1] * u[1]);
  if constexpr (S::D == 3) uxu = fma_(u[0], u[0], uxu);
  const R c0 = fma_(uxu, R(-1.0 / (2.0 * LB_CS2)), R(1));
  feq0 = (R(S::w(0)) * rho) * c0;
  static_for<N_SLOTS>([&](auto sc) {
    constexpr int s = decltype(sc)::value;
    if constexpr (has_slot<S, s>) {
      const R wr = R(S::w(pair_q<S>(s))) * rho;
      const R a = slot_dot<s, R>(U);
      const R sym = fma_(R(0.5) * a, a, c0);
      Sp[s] = wr * sym;
      Dp[s] = wr * a;
    }
  });
}

// full equilibrium, population by population (boundaries, initialisation)
template <class S, typename R>
__host__ __device__ __forceinline__ void equilibrium(R rho, const R (&u)[3], R (&feq)[S::Q]) {
  R Sp[N_SLOTS], Dp[N_SLOTS];
  equilibrium_pairs<S, R>(rho, u, feq[0], Sp, Dp);
  static_for<N_SLOTS>([&](auto sc) {
    constexpr int s = decltype(sc)::value;
    if constexpr (has_slot<S, s>) {
      constexpr int q = pair_q<S>(s), o = S::opp(q);
      feq[q] = Sp[s] + Dp[s];
      feq[o] = Sp[s] - Dp[s];
    }
  });
}

enum { COLL_NONE = 0, COLL_BGK = 1, COLL_TRT = 2, COLL_KBC = 3, COLL_REGULARIZED = 4, COLL_SMAGORINSKY = 5,
       COLL_BGK_FORCE = 6 };   // internal: LB_COLL_BGK with lb_desc.force set

// Second moments of the non-equilibrium part, Pi_AB = sum_q e_A e_B (f_q - feq_q)  (Flow.shear_tensor of
// f - feq, _flow.py:202-208): population 0 does not contribute and an opposite pair contributes through
// n_p = (f_q + f_o) - (feq_q + feq_o) = a_p - 2 S_p only.
template <class S, typename R>
__host__ __device__ __forceinline__ EvenMoments<R> nonequilibrium_shear(const R (&a)[N_SLOTS], const R (&Sp)[N_SLOTS]) {
  R n[N_SLOTS];
  static_for<N_SLOTS>([&](auto sc) {
    constexpr int s = decltype(sc)::value;
    if constexpr (has_slot<S, s>) n[s] = fma_(R(-2), Sp[s], a[s]);
  });
  return even_moments<S, R>(n);   // .sum unused
}

// e . Pi . e for the direction of a slot (D2Q9 has no slot with an x component)
template <int SLOT, typename R>
__host__ __device__ __forceinline__ R slot_quadratic_form(const EvenMoments<R>& m) {
  constexpr int ex = slot_e(SLOT, 0), ey = slot_e(SLOT, 1), ez = slot_e(SLOT, 2);
  if constexpr (ex != 0 && ey != 0 && ez != 0) {
    R off = (ex * ey > 0) ? m.xy : -m.xy;
    off = (ex * ez > 0) ? off + m.xz : off - m.xz;
    off = (ey * ez > 0) ? off + m.yz : off - m.yz;
    return fma_(R(2), off, (m.xx + m.yy) + m.zz);
  } else if constexpr (ex != 0 && ey != 0) {
    return fma_(R(2 * ex * ey), m.xy, m.xx + m.yy);
  } else if constexpr (ex != 0 && ez != 0) {
    return fma_(R(2 * ex * ez), m.xz, m.xx + m.zz);
  } else if constexpr (ey != 0 && ez != 0) {
    return fma_(R(2 * ey * ez), m.yz, m.yy + m.zz);
  } else if constexpr (ex != 0) {
    return m.xx;
  } else if constexpr (ey != 0) {
    return m.yy;
  } else {
    return m.zz;
  }
}

// ---------------------------------------------------------------------------------------
// Regularized LBM, Latt & Chopard 2006 (regularized_collision.py:17-44):
//   Pi = shear_tensor(f - feq);  fi1_q = w_q (Q_q : Pi) / (2 cs^4),  Q_q,ab = e_a e_b - cs^2 delta_ab;
//   f_q = feq_q + (1 - 1/tau) fi1_q.
// Q_q : Pi = e.Pi.e - cs^2 tr Pi is even in e, so both members of a pair get the same fi1.
// ---------------------------------------------------------------------------------------
template <class S, typename R>
__host__ __device__ __forceinline__ void regularized_relax(R (&f)[S::Q], R feq0, const R (&a)[N_SLOTS],
                                                           const R (&Sp)[N_SLOTS], const R (&Dp)[N_SLOTS], R c) {
  const EvenMoments<R> pi = nonequilibrium_shear<S, R>(a, Sp);
  const R trace = (S::D == 3) ? (pi.xx + pi.yy) + pi.zz : pi.yy + pi.zz;
  const R cs2_trace = R(LB_CS2) * trace;
  const R k = c * R(1.0 / (2.0 * LB_CS2 * LB_CS2));   // (1 - 1/tau) / (2 cs^4)
  f[0] = fma_(-(k * R(S::w(0))), cs2_trace, feq0);
  static_for<N_SLOTS>([&](auto sc) {
    constexpr int s = decltype(sc)::value;
    if constexpr (has_slot<S, s>) {
      constexpr int q = pair_q<S>(s), o = S::opp(q);
      const R fi1 = (k * R(S::w(q))) * (slot_quadratic_form<s, R>(pi) - cs2_trace);
      f[q] = (Sp[s] + Dp[s]) + fi1;
      f[o] = (Sp[s] - Dp[s]) + fi1;
    }
  });
}

// ---------------------------------------------------------------------------------------
// Smagorinsky LES with the BGK operator (smagorinsky_collision.py:19-36):
//   S_shear = shear_tensor(f - feq) / (2 rho cs^2);  tau_eff = tau;  nu = (tau - 0.5) / 3;
//   twice:  S = S_shear / tau_eff;  nu_t = C^2 sum_ab S_ab S_ab;  tau_eff = 3 (nu + nu_t) + 0.5;
//   f - (f - feq) / tau_eff.
// ---------------------------------------------------------------------------------------
template <class S, typename R>
__host__ __device__ __forceinline__ void smagorinsky_relax(R (&f)[S::Q], R feq0, R rho, const R (&a)[N_SLOTS],
                                                           const R (&Sp)[N_SLOTS], const R (&Dp)[N_SLOTS], R tau,
                                                           R nu, R c2) {
  const EvenMoments<R> pi = nonequilibrium_shear<S, R>(a, Sp);
  const R inv = div_(R(1), R(2.0 * LB_CS2) * rho);
  // sum_ab (Pi_ab / (2 rho cs^2))^2 over the full d x d tensor: diagonal once, off-diagonal twice
  const R yy = pi.yy * inv, zz = pi.zz * inv, yz = pi.yz * inv;
  R diag = fma_(yy, yy, zz * zz), off = yz * yz;
  if constexpr (S::D == 3) {
    const R xx = pi.xx * inv, xy = pi.xy * inv, xz = pi.xz * inv;
    diag = fma_(xx, xx, diag);
    off = fma_(xy, xy, fma_(xz, xz, off));
  }
  const R p2 = fma_(R(2), off, diag);
  R tau_eff = tau;
#pragma unroll
  for (int it = 0; it < 2; ++it) {   // self.iterations = 2 (:17)
    const R ss = div_(p2, tau_eff * tau_eff);
    tau_eff = fma_(fma_(c2, ss, nu), R(3), R(0.5));
  }
  const R om = div_(R(1), tau_eff);
  f[0] = fma_(om, feq0 - f[0], f[0]);
  static_for<N_SLOTS>([&](auto sc) {
    constexpr int s = decltype(sc)::value;
    if constexpr (has_slot<S, s>) {
      constexpr int q = pair_q<S>(s), o = S::opp(q);
      f[q] = fma_(om, (Sp[s] + Dp[s]) - f[q], f[q]);
      f[o] = fma_(om, (Sp[s] - Dp[s]) - f[o], f[o]);
    }
  });
}

// ---------------------------------------------------------------------------------------
// KBC (kbc_collision.py:25-160).  The reference builds all raw moments of f and of feq with an
// einsum, normalises by rho, forms  T, N_xz, N_yz, Pi_ab  and from them the "shear part" s of the
// populations (:44-94).  Written out, that shear part is simply
//     s_0 = -(M_xx + M_yy + M_zz),  s_{+-A} = M_AA / 2,  s_{edge AB, sign} = sign * M_AB / 4,  s_corner = 0
// with the UN-normalised second moments M_AB = sum_q e_A e_B f_q  (the 1/rho and the factor rho
// cancel; 2-D: s_{1,3} = M_xx/2, s_{2,4} = M_yy/2, s_{5,7} = -s_{6,8} = M_xy/4), and
// delta_s = s(f) - s(feq) is linear in f - feq:  it needs the second moments of the pair sums
// n_p = a_p - 2 S_p of the non-equilibrium part only.  Same quantity as the reference's, with less
// cancellation (the reference subtracts two O(rho) numbers).
// ---------------------------------------------------------------------------------------
template <class S, typename R>
__host__ __device__ __forceinline__ void kbc_relax(R (&f)[S::Q], R feq0, const R (&a)[N_SLOTS], const R (&Sp)[N_SLOTS],
                                                   const R (&Dp)[N_SLOTS], R beta, R inv_beta) {
  static_assert(S::ID != 1, "KBC is defined for D2Q9 and D3Q27 only (kbc_collision.py:101,115)");
  using scalar = typename ValueTraits<R>::scalar;
  R n[N_SLOTS];
  static_for<N_SLOTS>([&](auto sc) {
    constexpr int s = decltype(sc)::value;
    if constexpr (has_slot<S, s>) n[s] = fma_(R(-2), Sp[s], a[s]);
  });
  const EvenMoments<R> dm = even_moments<S, R>(n);   // .sum unused
  // delta_s per direction class (index = slot; corners 9..12 have none), and for population 0
  R ds[9];
  R ds0;
  if constexpr (S::D == 3) {
    ds0 = -((dm.xx + dm.yy) + dm.zz);
    ds[0] = R(0.5) * dm.xx;
  } else {
    ds0 = -(dm.yy + dm.zz);
  }
  ds[1] = R(0.5) * dm.yy;
  ds[2] = R(0.5) * dm.zz;
  ds[3] = R(0.25) * dm.yz;
  ds[4] = -ds[3];
  if constexpr (S::D == 3) {
    ds[5] = R(0.25) * dm.xz;
    ds[6] = -ds[5];
    ds[7] = R(0.25) * dm.xy;
    ds[8] = -ds[7];
  }
  // delta_h = f - feq - delta_s;  sum_s = sum ds dh / feq,  sum_h = sum dh dh / feq  (:148-152);
  // fn = f - feq is kept for the update, feq itself is dead after this pass
  R fn[S::Q];
  R sum_s, sum_h;
  {
    fn[0] = f[0] - feq0;
    const R dh = fn[0] - ds0;
    const R w = dh * rcp_fast(feq0);
    sum_s = ds0 * w;
    sum_h = dh * w;
  }
  static_for<N_SLOTS>([&](auto sc) {
    constexpr int s = decltype(sc)::value;
    if constexpr (has_slot<S, s>) {
      constexpr int q = pair_q<S>(s), o = S::opp(q);
      const R fq = Sp[s] + Dp[s], fo = Sp[s] - Dp[s];
      fn[q] = f[q] - fq;
      fn[o] = f[o] - fo;
      if constexpr (s < 9) {
        const R dhq = fn[q] - ds[s], dho = fn[o] - ds[s];
        const R wq = dhq * rcp_fast(fq), wo = dho * rcp_fast(fo);
        sum_s = fma_(ds[s], wq, sum_s);
        sum_s = fma_(ds[s], wo, sum_s);
        sum_h = fma_(dhq, wq, sum_h);
        sum_h = fma_(dho, wo, sum_h);
      } else {
        const R wq = fn[q] * rcp_fast(fq), wo = fn[o] * rcp_fast(fo);
        sum_h = fma_(fn[q], wq, sum_h);
        sum_h = fma_(fn[o], wo, sum_h);
      }
    }
  });
  // gamma = 1/beta - (2 - 1/beta) sum_s / sum_h, guards of :153-157
  R gamma = fma_(-(R(2) - inv_beta), div_(sum_s, sum_h), inv_beta);
  gamma = lt_select(gamma, scalar(1E-15), scalar(2));
  gamma = nan_select(gamma, scalar(2));
  // f - beta (2 ds + gamma dh)  =  f - beta gamma (f - feq) - beta (2 - gamma) ds   (:158)
  const R bg = beta * gamma;
  const R bs = beta * (R(2) - gamma);
  f[0] = fma_(-bg, fn[0], f[0] - bs * ds0);
  static_for<N_SLOTS>([&](auto sc) {
    constexpr int s = decltype(sc)::value;
    if constexpr (has_slot<S, s>) {
      constexpr int q = pair_q<S>(s), o = S::opp(q);
      if constexpr (s < 9) {
        const R k = bs * ds[s];
        f[q] = fma_(-bg, fn[q], f[q] - k);
        f[o] = fma_(-bg, fn[o], f[o] - k);
      } else {
        f[q] = fma_(-bg, fn[q], f[q]);
        f[o] = fma_(-bg, fn[o], f[o]);
      }
    }
  });
}

// ---------------------------------------------------------------------------------------
// BGK with a forcing scheme (bgk_collision.py:17-22; ext/_force/guo.py:16-37, shan_chen.py:13-23):
//   u = j / rho + scale a / rho;   f - (f - feq(rho, u)) / tau + S,
//   Guo:  S_q = c w_q ((e_q - u) / cs^2 + (e_q.u) e_q / cs^4) . a,  c = 1 - 1/(2 tau_force);  Shan-Chen: S = 0.
// Even / odd in e:  S_q = c (E_p + O_p),  S_o = c (E_p - O_p)  with  E_p = w ((e.U)(e.A) - u.A),
// O_p = w (e.A),  U = u / cs^2,  A = a / cs^2;  S_0 = -c w_0 (u.A).
// ---------------------------------------------------------------------------------------
template <class S, typename R, typename T>
__host__ __device__ __forceinline__ void bgk_force_relax(R (&f)[S::Q], R rho, const R (&u)[3],
                                                         const CollisionParams<T>& cp) {
  const R inv_cs2 = R(1.0 / LB_CS2);
  R uf[3], U[3], A[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    uf[c] = (S::D == 3 || c > 0) ? u[c] + div_(R(cp.force_ueq[c]), rho) : R(0);
    U[c] = uf[c] * inv_cs2;
    A[c] = R(cp.force_a[c]) * inv_cs2;
  }
  R feq0, Sp[N_SLOTS], Dp[N_SLOTS];
  equilibrium_pairs<S, R>(rho, uf, feq0, Sp, Dp);
  R ua = fma_(uf[2], A[2], uf[1] * A[1]);
  if constexpr (S::D == 3) ua = fma_(uf[0], A[0], ua);
  const R om = R(cp.omega), c = R(cp.force_src);
  f[0] = fma_(om, feq0 - f[0], f[0]) - c * (R(S::w(0)) * ua);
  static_for<N_SLOTS>([&](auto sc) {
    constexpr int s = decltype(sc)::value;
    if constexpr (has_slot<S, s>) {
      constexpr int q = pair_q<S>(s), o = S::opp(q);
      const R w = R(S::w(q));
      const R ea = slot_dot<s, R>(A);
      const R even = w * fma_(slot_dot<s, R>(U), ea, -ua);
      const R odd = w * ea;
      f[q] = fma_(om, (Sp[s] + Dp[s]) - f[q], f[q]) + c * (even + odd);
      f[o] = fma_(om, (Sp[s] - Dp[s]) - f[o], f[o]) + c * (even - odd);
    }
  });
}

// f <- collision(f), given the pair sums / differences and the moments of the incoming f
template <class S, typename R, int COLL, typename T>
__host__ __device__ __forceinline__ void collide_node(R (&f)[S::Q], const R (&a)[N_SLOTS], const R (&b)[N_SLOTS],
                                                      R rho, const R (&u)[3], const CollisionParams<T>& cp) {
  if constexpr (COLL == COLL_NONE) {
    return;  // no_collision.py:9-11
  } else if constexpr (COLL == COLL_BGK_FORCE) {
    bgk_force_relax<S, R, T>(f, rho, u, cp);
  } else {
    R feq0, Sp[N_SLOTS], Dp[N_SLOTS];
    equilibrium_pairs<S, R>(rho, u, feq0, Sp, Dp);
    if constexpr (COLL == COLL_BGK) {
      // bgk_collision.py:22: f - 1.0/tau * (f - feq)
      const R om = R(cp.omega);
      f[0] = fma_(om, feq0 - f[0], f[0]);
      static_for<N_SLOTS>([&](auto sc) {
        constexpr int s = decltype(sc)::value;
        if constexpr (has_slot<S, s>) {
          constexpr int q = pair_q<S>(s), o = S::opp(q);
          f[q] = fma_(om, (Sp[s] + Dp[s]) - f[q], f[q]);
          f[o] = fma_(om, (Sp[s] - Dp[s]) - f[o], f[o]);
        }
      });
    } else if constexpr (COLL == COLL_TRT) {
      // trt_collision.py:20-26 per opposite pair:  sp = ((f_q + f_o) - (feq_q + feq_o)) / (2 tau+),
      // sm = ((f_q - f_o) - (feq_q - feq_o)) / (2 tau-);  f_q - (sp + sm),  f_o - (sp - sm).
      const R itp = R(cp.inv_2tau_p), itm = R(cp.inv_2tau_m);
      f[0] = fma_(-(R(2) * itp), f[0] - feq0, f[0]);
      static_for<N_SLOTS>([&](auto sc) {
        constexpr int s = decltype(sc)::value;
        if constexpr (has_slot<S, s>) {
          constexpr int q = pair_q<S>(s), o = S::opp(q);
          const R sp = fma_(R(-2), Sp[s], a[s]) * itp;
          const R sm = fma_(R(-2), Dp[s], b[s]) * itm;
          f[q] = f[q] - (sp + sm);
          f[o] = f[o] - (sp - sm);
        }
      });
    } else if constexpr (COLL == COLL_REGULARIZED) {
      regularized_relax<S, R>(f, feq0, a, Sp, Dp, R(cp.reg_c));
    } else if constexpr (COLL == COLL_SMAGORINSKY) {
      smagorinsky_relax<S, R>(f, feq0, rho, a, Sp, Dp, R(cp.smag_tau), R(cp.smag_nu), R(cp.smag_c2));
    } else {
      kbc_relax<S, R>(f, feq0, a, Sp, Dp, R(cp.beta), R(cp.inv_beta));
    }
  }
}

// convenience: collision of one value from its populations alone
template <class S, typename R, int COLL, typename T>
__host__ __device__ __forceinline__ void collide_node(R (&f)[S::Q], const CollisionParams<T>& cp) {
  if constexpr (COLL != COLL_NONE) {
    R a[N_SLOTS], b[N_SLOTS], rho, u[3];
    pair_split<S, R>(f, a, b);
    moments_from_pairs<S, R>(f[0], a, b, rho, u);
    collide_node<S, R, COLL, T>(f, a, b, rho, u, cp);
  }
}

// bounce_back_boundary.py:17-18: f <- f[opposite]
template <class S, typename T>
__host__ __device__ __forceinline__ void bounce_back_node(T (&f)[S::Q]) {
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
