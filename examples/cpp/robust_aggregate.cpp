// Byzantine-robust aggregation from a plain C++ host program through the C ABI of libddl25_kernels.so - no Python, no torch.
//
//   g++ -O2 -std=c++17 -I include examples/cpp/robust_aggregate.cpp -o robust_aggregate \
//       -L ddl25spring_b200/lib -lddl25_kernels -L /usr/local/cuda/lib64 -lcudart -Wl,-rpath,$PWD/ddl25spring_b200/lib
//   ./robust_aggregate [K] [D] [n_byzantine]
//
// K client updates of D coordinates sit in one [K, D] device matrix (the layout of core.arena / fl.aggregators).  The last
// n_byzantine rows are sign-flipped and scaled (fl/attacks.py).  The program runs Multi-Krum (tcgen05 Gram matrix in the
// library's static scratch - no allocation after set-up - then the on-device selection and the weighted sum driven by the
// DEVICE selection weights: no host round trip between the three launches), the coordinate-wise median and the trimmed mean,
// times each with CUDA events and checks the results against the honest mean on the host.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include "ddl25_kernels.h"

#define CK(x)                                                                                   \
  do {                                                                                          \
    cudaError_t e_ = (x);                                                                       \
    if (e_ != cudaSuccess) { std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); return 2; } \
  } while (0)
#define DDL(x)                                                                                  \
  do {                                                                                          \
    if ((x) != 0) { std::fprintf(stderr, "%s failed: %s\n", #x, ddl_last_error()); return 3; }  \
  } while (0)

template <typename F>
static float timed_ms(cudaStream_t st, int iters, F&& launch) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  launch();                                   // warm-up (one-time kernel attribute set-up happens here)
  cudaStreamSynchronize(st);
  cudaEventRecord(e0, st);
  for (int i = 0; i < iters; ++i) launch();
  cudaEventRecord(e1, st);
  cudaEventSynchronize(e1);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  return ms / iters;
}

int main(int argc, char** argv) {
  const int K = argc > 1 ? std::atoi(argv[1]) : 16;
  const int64_t D = argc > 2 ? std::atoll(argv[2]) : (1 << 20);
  const int F = argc > 3 ? std::atoi(argv[3]) : 3;
  if (K < 4 || K > 64 || D % 32 != 0 || F < 0 || 2 * F + 2 >= K) { std::fprintf(stderr, "need 4 <= K <= 64, D %% 32 == 0, 2f + 2 < K\n"); return 1; }
  std::printf("library arch %s; K = %d updates x D = %lld, %d sign-flip attackers\n", ddl_arch(), K, (long long)D, F);

  // honest update k = signal + small per-client noise; attackers send -4 x an honest update
  std::vector<float> signal(D), U((size_t)K * D);
  std::mt19937 rng(7);
  std::normal_distribution<float> nd(0.f, 1.f);
  for (auto& s : signal) s = nd(rng);
  for (int k = 0; k < K; ++k)
    for (int64_t d = 0; d < D; ++d) {
      const float honest = signal[d] + 0.1f * nd(rng);
      U[(size_t)k * D + d] = k >= K - F ? -4.f * honest : honest;
    }
  std::vector<double> honest_mean(D, 0.0);
  for (int k = 0; k < K - F; ++k)
    for (int64_t d = 0; d < D; ++d) honest_mean[d] += U[(size_t)k * D + d] / double(K - F);

  cudaStream_t st;
  CK(cudaStreamCreate(&st));
  float *dU, *dG, *dScores, *dSelW, *dOut;
  int* dSel;
  CK(cudaMalloc(&dU, sizeof(float) * K * D));
  CK(cudaMalloc(&dG, sizeof(float) * K * K));
  CK(cudaMalloc(&dScores, sizeof(float) * K));
  CK(cudaMalloc(&dSelW, sizeof(float) * K));
  CK(cudaMalloc(&dSel, sizeof(int) * K));
  CK(cudaMalloc(&dOut, sizeof(float) * D));
  CK(cudaMemcpy(dU, U.data(), sizeof(float) * K * D, cudaMemcpyHostToDevice));

  const int multi = K - F;
  auto krum = [&]() -> int {
    DDL(ddl_gram_static_scratch(dU, D, K, D, dG, K, /*n_split=*/148, /*precise=*/1, st));
    DDL(ddl_krum_select(dG, K, K, F, multi, dScores, dSel, dSelW, st));
    DDL(ddl_weighted_sum(dU, D, K, nullptr, dSelW, nullptr, 0.f, 1.f, dOut, D, st));
    return 0;
  };
  auto max_err = [&](const char* what, double tol) -> int {
    std::vector<float> out(D);
    cudaMemcpy(out.data(), dOut, sizeof(float) * D, cudaMemcpyDeviceToHost);
    double worst = 0.0;
    for (int64_t d = 0; d < D; ++d) worst = std::fmax(worst, std::fabs(out[d] - honest_mean[d]));
    std::printf("  %-14s max |out - honest mean| = %.4f (tolerance %.2f)\n", what, worst, tol);
    return worst <= tol ? 0 : 1;
  };

  int bad = 0;
  if (krum() != 0) return 3;
  const float t_krum = timed_ms(st, 10, [&] { krum(); });
  std::vector<int> sel(multi);
  CK(cudaMemcpy(sel.data(), dSel, sizeof(int) * multi, cudaMemcpyDeviceToHost));
  int attackers_selected = 0;
  for (int s : sel) attackers_selected += (s >= K - F);
  std::printf("multi-Krum (m = %d): %.3f ms per aggregation (Gram + selection + weighted sum), attackers selected: %d\n", multi, t_krum,
              attackers_selected);
  bad += attackers_selected != 0;
  bad += max_err("multi-Krum", 1e-3);

  const float t_med = timed_ms(st, 10, [&] { ddl_coordinate_median(dU, D, K, 1, dOut, D, st); });
  std::printf("coordinate-wise median: %.3f ms\n", t_med);
  bad += max_err("median", 0.5);
  const float t_trim = timed_ms(st, 10, [&] { ddl_trimmed_mean(dU, D, K, F, dOut, D, st); });
  std::printf("trimmed mean (trim %d): %.3f ms\n", F, t_trim);
  bad += max_err("trimmed mean", 0.5);

  std::printf("kernel launches issued by the library: %lld\n", (long long)ddl_launch_count());
  cudaFree(dU); cudaFree(dG); cudaFree(dScores); cudaFree(dSelW); cudaFree(dSel); cudaFree(dOut);
  cudaStreamDestroy(st);
  std::printf(bad ? "FAILED\n" : "OK\n");
  return bad ? 4 : 0;
}
