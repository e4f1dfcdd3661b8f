// Development probe: which operation of the TRT relaxation rounds differently in the packed instantiation?
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "lb_physics.cuh"
using namespace lb;
using S = D3Q19;

__global__ void probe(const float* f, int n, unsigned* counts, float* first) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float itp = float(1.0 / (2 * 0.56)), itm = 0.5f;
  float g[S::Q]; F2 h[S::Q];
  for (int q = 0; q < S::Q; ++q) { g[q] = f[q * n + i]; h[q] = F2(g[q], g[q]); }
  float a[N_SLOTS], b[N_SLOTS], rho, u[3], feq0, Sp[N_SLOTS], Dp[N_SLOTS];
  F2 a2[N_SLOTS], b2[N_SLOTS], rho2, u2[3], feq02, Sp2[N_SLOTS], Dp2[N_SLOTS];
  pair_split<S, float>(g, a, b); moments_from_pairs<S, float>(g[0], a, b, rho, u); equilibrium_pairs<S, float>(rho, u, feq0, Sp, Dp);
  pair_split<S, F2>(h, a2, b2); moments_from_pairs<S, F2>(h[0], a2, b2, rho2, u2); equilibrium_pairs<S, F2>(rho2, u2, feq02, Sp2, Dp2);
  static_for<N_SLOTS>([&](auto sc) {
    constexpr int s = decltype(sc)::value;
    if constexpr (has_slot<S, s>) {
      constexpr int q = pair_q<S>(s);
      const float t1 = fma_(-2.0f, Sp[s], a[s]);
      const F2 t1p = fma_(F2(-2), Sp2[s], a2[s]);
      const float sp = t1 * itp;
      const F2 spp = t1p * F2(itp);
      const float t2 = fma_(-2.0f, Dp[s], b[s]);
      const F2 t2p = fma_(F2(-2), Dp2[s], b2[s]);
      const float sm = t2 * itm;
      const F2 smp = t2p * F2(itm);
      const float sum = sp + sm;
      const F2 sump = spp + smp;
      const float r = g[q] - sum;
      const F2 rp = h[q] - sump;
      int stage = 0;
      if (t1 != t1p.x) stage = 1; else if (sp != spp.x) stage = 2; else if (t2 != t2p.x) stage = 3; else if (sm != smp.x) stage = 4;
      else if (sum != sump.x) stage = 5; else if (r != rp.x) stage = 6;
      if (stage) {
        atomicAdd(&counts[stage], 1u);
        if (atomicAdd(&counts[0], 1u) == 0) { first[0] = stage; first[1] = t1; first[2] = t1p.x; first[3] = sp; first[4] = spp.x; first[5] = sm; first[6] = smp.x;
          first[7] = sum; first[8] = sump.x; first[9] = r; first[10] = rp.x; first[11] = g[q]; first[12] = itp; }
      }
    }
  });
}

int main() {
  const int n = 1 << 16;
  std::vector<float> f(S::Q * n);
  srand(1);
  for (int i = 0; i < n; ++i) for (int q = 0; q < S::Q; ++q) f[q * n + i] = float(S::w(q)) * (1.0f + 0.05f * (rand() / float(RAND_MAX) - 0.5f));
  float *df, *dfirst; unsigned* dn;
  cudaMalloc(&df, f.size() * 4); cudaMalloc(&dn, 32); cudaMalloc(&dfirst, 64);
  cudaMemcpy(df, f.data(), f.size() * 4, cudaMemcpyHostToDevice); cudaMemset(dn, 0, 32); cudaMemset(dfirst, 0, 64);
  probe<<<n / 128, 128>>>(df, n, dn, dfirst);
  unsigned c[8]; float first[16];
  cudaMemcpy(c, dn, 32, cudaMemcpyDeviceToHost); cudaMemcpy(first, dfirst, 64, cudaMemcpyDeviceToHost);
  printf("differences by stage (1 t1=fma(-2,S,a)  2 sp=t1*itp  3 t2  4 sm=t2*itm  5 sp+sm  6 f-(sp+sm)): %u %u %u %u %u %u, total %u\n", c[1], c[2], c[3], c[4], c[5], c[6], c[0]);
  printf("first: stage %d  t1 %.9g/%.9g  sp %.9g/%.9g  sm %.9g/%.9g  sum %.9g/%.9g  r %.9g/%.9g  f %.9g itp %.9g\n", int(first[0]), first[1], first[2], first[3], first[4], first[5], first[6], first[7], first[8], first[9], first[10], first[11], first[12]);
  return 0;
}
