// Launchers of the generic fp32 GEMM / column-sum kernels (gemm.cuh).
#include "tnt_internal.h"
#include "gemm.cuh"

// splitk: 0/1 = one CTA per output tile (store, or read-add-store when accumulate != 0);
//         >1 or accumulate == 2 = atomicAdd into C (C must hold the value to be accumulated onto).
void sgemm(tnt_ctx* c, bool ta, bool tb, int M, int N, int K, const float* A, int lda, const float* B, int ldb,
           float* C, int ldc, const float* bias, int accumulate, int splitk, cudaStream_t st) {
  if (M <= 0 || N <= 0 || K <= 0) return;
  int splits = splitk > 1 ? splitk : 1;
  int kps = (K + splits - 1) / splits;
  kps = ((kps + GEMM_BK - 1) / GEMM_BK) * GEMM_BK;
  splits = (K + kps - 1) / kps;
  if (accumulate == 2 && splits == 1) { splits = 1; }
  dim3 grid((N + GEMM_BN - 1) / GEMM_BN, (M + GEMM_BM - 1) / GEMM_BM, splits);
  int acc_flag = accumulate;
#define LAUNCH(TA_, TB_) sgemm_kernel<TA_, TB_><<<grid, 256, 0, st>>>(M, N, K, A, lda, B, ldb, C, ldc, bias, acc_flag, kps)
  if (!ta && !tb) LAUNCH(false, false);
  else if (ta && !tb) LAUNCH(true, false);
  else if (!ta && tb) LAUNCH(false, true);
  else LAUNCH(true, true);
#undef LAUNCH
  c->launches += 1;
}

void colsum(tnt_ctx* c, int M, int N, const float* X, int ldx, float* out, cudaStream_t st) {
  if (N <= 0) return;
  colsum_kernel<<<(N + 31) / 32, 256, 0, st>>>(M, N, X, ldx, out);
  c->launches += 1;
}
