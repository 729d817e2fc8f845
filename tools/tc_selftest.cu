// Stand-alone check of the tcgen05 GEMM (csrc/gemm_tc.cu) against an fp64 CPU product.
//   tools/tc_selftest.bin [split: 2|3]  exit 0 = every case within tolerance
// Built by `make -C masters-thesis_b200/csrc selftest`; run on the GPU box only.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <vector>
#include <random>

#include "../masters-thesis_b200/csrc/tnt_internal.h"
void gemm_tc(tnt_ctx* c, int cacheA, int cacheB, bool ta, bool tb, int M, int N, int K, const float* A, int lda,
             const float* B, int ldb, float* C, int ldc, const float* bias, int accumulate, cudaStream_t st);
extern int g_tc_desc_swap, g_tc_split;

struct Case { int M, N, K, ta, tb, pad, bias, acc, splitk; };

static int run_case(const Case& cs, std::mt19937& rng) {
  const int M = cs.M, N = cs.N, K = cs.K;
  const int lda = (cs.ta ? M : K) + cs.pad, ldb = (cs.tb ? K : N) + cs.pad, ldc = N + cs.pad;
  const size_t na = (size_t)(cs.ta ? K : M) * lda, nb = (size_t)(cs.tb ? N : K) * ldb, nc = (size_t)M * ldc;
  std::vector<float> A(na), B(nb), C(nc), bias(N);
  std::normal_distribution<float> nd(0.f, 1.f);
  for (auto& x : A) x = nd(rng);
  for (auto& x : B) x = nd(rng);
  for (auto& x : C) x = nd(rng);
  for (auto& x : bias) x = nd(rng);
  std::vector<double> ref(nc, 0.0);
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) {
      double s = 0.0;
      for (int k = 0; k < K; ++k) {
        double a = cs.ta ? A[(size_t)k * lda + m] : A[(size_t)m * lda + k];
        double b = cs.tb ? B[(size_t)n * ldb + k] : B[(size_t)k * ldb + n];
        s += a * b;
      }
      if (cs.bias) s += bias[n];
      if (cs.acc) s += C[(size_t)m * ldc + n];
      ref[(size_t)m * ldc + n] = s;
    }
  float *dA, *dB, *dC, *dbias;
  cudaMalloc(&dA, na * 4); cudaMalloc(&dB, nb * 4); cudaMalloc(&dC, nc * 4); cudaMalloc(&dbias, N * 4);
  cudaMemcpy(dA, A.data(), na * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, B.data(), nb * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dC, C.data(), nc * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dbias, bias.data(), N * 4, cudaMemcpyHostToDevice);
  gemm_tc(nullptr, 0, 0, cs.ta, cs.tb, M, N, K, dA, lda, dB, ldb, dC, ldc, cs.bias ? dbias : nullptr, cs.acc, 0);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("CUDA error: %s\n", cudaGetErrorString(e));
    return 2;
  }
  std::vector<float> out(nc);
  cudaMemcpy(out.data(), dC, nc * 4, cudaMemcpyDeviceToHost);
  double num = 0.0, den = 0.0, maxabs = 0.0;
  int pad_touched = 0;
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < ldc; ++n) {
      size_t i = (size_t)m * ldc + n;
      if (n >= N) { if (out[i] != C[i]) pad_touched++; continue; }
      double d = (double)out[i] - ref[i];
      num += d * d; den += ref[i] * ref[i];
      if (fabs(d) > maxabs) maxabs = fabs(d);
    }
  double rel = sqrt(num / (den > 0 ? den : 1));
  int ok = rel < (g_tc_split == 2 ? 2e-5 : 5e-7) && pad_touched == 0;
  printf("%s M=%d N=%d K=%d ta=%d tb=%d pad=%d bias=%d acc=%d splitk=%d : rel=%.3e maxabs=%.3e pad_touched=%d\n",
         ok ? "ok  " : "FAIL", M, N, K, cs.ta, cs.tb, cs.pad, cs.bias, cs.acc, cs.splitk, rel, maxabs, pad_touched);
  cudaFree(dA); cudaFree(dB); cudaFree(dC); cudaFree(dbias);
  return ok ? 0 : 1;
}

// `tc_selftest.bin bench [split]`: device time of the train step's time-parallel GEMM shapes (warm L2)
static int run_bench(int split) {
  g_tc_split = split;
  struct Sh { const char* name; int ta, tb, M, N, K, splitk; };
  const Sh shapes[] = {
      {"Zx      NN", 0, 0, 960, 2048, 512, 1},  {"inter   NN", 0, 0, 960, 256, 512, 1},
      {"logits  NN", 0, 0, 960, 5001, 256, 1},  {"dinter  NT", 0, 1, 960, 256, 5001, 1},
      {"doW     TN", 1, 0, 256, 5001, 960, 1},  {"diW     TN", 1, 0, 512, 256, 960, 1},
      {"dhout   NT", 0, 1, 960, 512, 256, 1},   {"dU      TN", 1, 0, 512, 2048, 960, 1},
      {"dWc     TN", 1, 0, 32, 2048, 960, 1},   {"dWe     TN", 1, 0, 512, 2048, 960, 1},
      {"dW2     TN", 1, 0, 512, 32, 960, 1},    {"dxtext  NT", 0, 1, 960, 512, 2048, 1},
  };
  float *dA, *dB, *dC;
  const size_t cap = (size_t)5008 * 2048;
  cudaMalloc(&dA, cap * 4); cudaMalloc(&dB, cap * 4); cudaMalloc(&dC, cap * 4);
  cudaMemset(dA, 0, cap * 4); cudaMemset(dB, 0, cap * 4); cudaMemset(dC, 0, cap * 4);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  double total = 0.0, total_pack = 0.0;
  auto time_us = [&](auto&& fn, int reps) {
    for (int i = 0; i < 3; ++i) fn();
    cudaEventRecord(e0);
    for (int i = 0; i < reps; ++i) fn();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms * 1000.0 / reps;
  };
  for (const Sh& s : shapes) {
    const int lda = s.ta ? s.M : s.K, ldb = s.tb ? s.K : s.N;
    tc_arena_reset();
    TcOperand a = s.ta ? tc_pack(nullptr, dA, s.K, s.M, lda, true, 0) : tc_pack(nullptr, dA, s.M, s.K, lda, false, 0);
    TcOperand b = s.tb ? tc_pack(nullptr, dB, s.N, s.K, ldb, false, 0) : tc_pack(nullptr, dB, s.K, s.N, ldb, true, 0);
    const double us = time_us([&] { gemm_packed(nullptr, a, b, s.M, s.N, dC, s.N, nullptr, 0, 0); }, 20);
    const double pa = time_us([&] { tc_arena_reset(); if (s.ta) tc_pack(nullptr, dA, s.K, s.M, lda, true, 0); else tc_pack(nullptr, dA, s.M, s.K, lda, false, 0); }, 20);
    const double pb = time_us([&] { tc_arena_reset(); if (s.tb) tc_pack(nullptr, dB, s.N, s.K, ldb, false, 0); else tc_pack(nullptr, dB, s.K, s.N, ldb, true, 0); }, 20);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("CUDA error\n"); return 2; }
    const double gf = 2.0 * s.M * s.N * s.K * 1e-9;
    total += us; total_pack += pa + pb;
    printf("split%d %s M=%4d N=%4d K=%4d : gemm %7.2f us (%6.1f TFLOP/s fp32-equivalent)  pack A %6.2f us  pack B %6.2f us\n", split,
           s.name, s.M, s.N, s.K, us, gf / us * 1e-3, pa, pb);
  }
  printf("split%d total %.1f us for the 12 GEMMs of one train step (+ %.1f us if every operand were packed per GEMM)\n", split, total,
         total_pack);
  return 0;
}

int main(int argc, char** argv) {
  if (argc > 1 && !strcmp(argv[1], "bench")) return run_bench(argc > 2 ? atoi(argv[2]) : 3);
  g_tc_split = argc > 1 ? atoi(argv[1]) : 3;
  printf("tc_selftest split=%d\n", g_tc_split);
  std::mt19937 rng(123);
  std::vector<Case> cases = {
      {128, 128, 64, 0, 0, 0, 0, 0, 1},  {128, 128, 64, 0, 1, 0, 0, 0, 1},  {128, 128, 64, 1, 0, 0, 0, 0, 1},
      {128, 128, 64, 1, 1, 0, 0, 0, 1},  {128, 128, 256, 0, 0, 0, 1, 0, 1}, {256, 384, 192, 1, 0, 0, 0, 1, 1},
      {100, 77, 50, 0, 0, 0, 1, 0, 1},   {100, 77, 50, 0, 1, 1, 0, 0, 1},   {100, 77, 50, 1, 0, 1, 1, 1, 1},
      {60, 200, 333, 1, 1, 3, 0, 0, 1},  {192, 130, 700, 0, 1, 0, 0, 2, 4}, {64, 160, 129, 0, 0, 1, 0, 0, 3},
      {960, 512, 256, 0, 0, 0, 1, 0, 1}, {32, 2048, 960, 1, 0, 0, 0, 0, 1}, {300, 500, 257, 0, 0, 1, 1, 0, 1},
  };
  int bad = 0;
  for (auto& cs : cases) {
    int r = run_case(cs, rng);
    if (r == 2) return 2;
    bad += r;
  }
  printf("%s: %d failing case(s)\n", bad ? "SELFTEST FAILED" : "SELFTEST PASSED", bad);
  return bad ? 1 : 0;
}
