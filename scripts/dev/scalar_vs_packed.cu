// Development probe: does the packed fp32x2 arithmetic of lb_physics.cuh round like the scalar instantiation ON THE
// DEVICE?  (tests/host_emul shows it does on the host.)  Build + run:
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a --expt-relaxed-constexpr -fmad=false \
//        -Ilettuce_b200/csrc scripts/dev/scalar_vs_packed.cu -o /tmp/svp && /tmp/svp
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "lb_physics.cuh"
using namespace lb;

template <class S, int COLL>
__global__ void probe(const float* f, int n, unsigned* n_diff, float* first) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  CollisionParams<float> cp{};
  const double tau = 0.56;
  cp.omega = float(1.0 / tau); cp.inv_2tau_p = float(1.0 / (2 * tau)); cp.inv_2tau_m = 0.5f;
  cp.beta = float(1.0 / (2 * tau)); cp.inv_beta = float(2 * tau);
  float g[S::Q];
  F2 h[S::Q];
  for (int q = 0; q < S::Q; ++q) { g[q] = f[q * n + i]; h[q] = F2(g[q], g[q]); }
  // intermediate: moments
  float a[N_SLOTS], b[N_SLOTS], rho, u[3];
  F2 a2[N_SLOTS], b2[N_SLOTS], rho2, u2[3];
  pair_split<S, float>(g, a, b); moments_from_pairs<S, float>(g[0], a, b, rho, u);
  pair_split<S, F2>(h, a2, b2); moments_from_pairs<S, F2>(h[0], a2, b2, rho2, u2);
  float feq0, Sp[N_SLOTS], Dp[N_SLOTS];
  F2 feq02, Sp2[N_SLOTS], Dp2[N_SLOTS];
  equilibrium_pairs<S, float>(rho, u, feq0, Sp, Dp);
  equilibrium_pairs<S, F2>(rho2, u2, feq02, Sp2, Dp2);
  int stage = 0;
  if (rho != rho2.x || u[0] != u2[0].x || u[1] != u2[1].x || u[2] != u2[2].x) stage = 1;
  if (!stage && feq0 != feq02.x) stage = 2;
  for (int s = 0; s < N_SLOTS && !stage; ++s) if (Sp[s] != Sp2[s].x || Dp[s] != Dp2[s].x) stage = 3 + s;
  collide_node<S, float, COLL, float>(g, a, b, rho, u, cp);
  collide_node<S, F2, COLL, float>(h, a2, b2, rho2, u2, cp);
  bool diff = false;
  for (int q = 0; q < S::Q; ++q) diff |= (g[q] != h[q].x) || (h[q].x != h[q].y);
  if (diff || stage) {
    if (atomicAdd(n_diff, 1u) == 0) { first[0] = float(i); first[1] = float(stage); first[2] = rho; first[3] = rho2.x;
      first[4] = u[1]; first[5] = u2[1].x; first[6] = feq0; first[7] = feq02.x; }
  }
}

template <class S, int COLL>
void run(const char* name) {
  const int n = 1 << 16;
  std::vector<float> f(S::Q * n);
  srand(1);
  for (int i = 0; i < n; ++i) for (int q = 0; q < S::Q; ++q) f[q * n + i] = float(S::w(q)) * (1.0f + 0.05f * (rand() / float(RAND_MAX) - 0.5f));
  float *df, *dfirst; unsigned* dn;
  cudaMalloc(&df, f.size() * 4); cudaMalloc(&dn, 4); cudaMalloc(&dfirst, 32);
  cudaMemcpy(df, f.data(), f.size() * 4, cudaMemcpyHostToDevice); cudaMemset(dn, 0, 4); cudaMemset(dfirst, 0, 32);
  probe<S, COLL><<<n / 128, 128>>>(df, n, dn, dfirst);
  unsigned nd; float first[8];
  cudaMemcpy(&nd, dn, 4, cudaMemcpyDeviceToHost); cudaMemcpy(first, dfirst, 32, cudaMemcpyDeviceToHost);
  printf("%s: %u of %d nodes differ between the scalar and the packed instantiation", name, nd, n);
  if (nd) printf("  [first: node %d stage %d rho %.9g/%.9g uy %.9g/%.9g feq0 %.9g/%.9g]", int(first[0]), int(first[1]), first[2], first[3], first[4], first[5], first[6], first[7]);
  printf("\n");
}

int main() {
  run<D3Q19, COLL_BGK>("D3Q19 BGK");
  run<D3Q19, COLL_TRT>("D3Q19 TRT");
  run<D3Q27, COLL_KBC>("D3Q27 KBC");
  run<D2Q9, COLL_BGK>("D2Q9 BGK");
  run<D3Q27, COLL_BGK>("D3Q27 BGK");
  run<D3Q19, COLL_REGULARIZED>("D3Q19 regularized");
  return 0;
}
