// Launchers of the generic fp32 GEMM / column-sum kernels (gemm.cuh).
#include "tnt_internal.h"
#include "gemm.cuh"
#include <cstdlib>

void gemm_tc(tnt_ctx* c, int cacheA, int cacheB, bool ta, bool tb, int M, int N, int K, const float* A, int lda,
             const float* B, int ldb, float* C, int ldc, const float* bias, int accumulate, cudaStream_t st);
extern int g_tc_desc_swap, g_tc_split;

// TNT_USE_TC=0 keeps the fp32 SIMT kernel (debugging aid); default is the tcgen05 path of gemm_tc.cu.
static int use_tc() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("TNT_USE_TC");
    v = e ? atoi(e) : 1;
    const char* s = getenv("TNT_TC_SWAP");
    if (s) g_tc_desc_swap = atoi(s);
    const char* sp = getenv("TNT_TC_SPLIT");
    if (sp) g_tc_split = atoi(sp) == 2 ? 2 : 3;
  }
  return v;
}

int gemm_uses_tc() { return use_tc(); }

// splitk: 0/1 = one CTA per output tile (store, or read-add-store when accumulate != 0);
//         >1 or accumulate == 2 = atomicAdd into C (C must hold the value to be accumulated onto).
void sgemm(tnt_ctx* c, bool ta, bool tb, int M, int N, int K, const float* A, int lda, const float* B, int ldb,
           float* C, int ldc, const float* bias, int accumulate, int splitk, cudaStream_t st) {
  sgemm_ex(c, 0, 0, ta, tb, M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate, splitk, st);
}

// cacheA / cacheB: the operand is not rewritten before the next tc_arena_reset(), so its packed form may be reused
void sgemm_ex(tnt_ctx* c, int cacheA, int cacheB, bool ta, bool tb, int M, int N, int K, const float* A, int lda,
              const float* B, int ldb, float* C, int ldc, const float* bias, int accumulate, int splitk, cudaStream_t st) {
  if (M <= 0 || N <= 0 || K <= 0) return;
  if (use_tc()) { gemm_tc(c, cacheA, cacheB, ta, tb, M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate, st); return; }
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
  colsum_kernel<<<(N + 31) / 32, COLSUM_WARPS * 32, 0, st>>>(M, N, X, ldx, out);
  c->launches += 1;
}

// up to COLSUM_MAXJOBS column sums with one launch (same arithmetic and order as colsum per job)
void colsum_multi(tnt_ctx* c, const ColsumReq* reqs, int n, cudaStream_t st) {
  ColsumJobs jobs; jobs.n = 0;
  int ctas = 0;
  for (int i = 0; i < n; ++i) {
    if (reqs[i].N <= 0) continue;
    if (jobs.n == COLSUM_MAXJOBS) { colsum(c, reqs[i].M, reqs[i].N, reqs[i].X, reqs[i].ldx, reqs[i].out, st); continue; }
    ColsumJob& J = jobs.j[jobs.n++];
    J.X = reqs[i].X; J.out = reqs[i].out; J.M = reqs[i].M; J.N = reqs[i].N; J.ldx = reqs[i].ldx; J.cta0 = ctas;
    ctas += (reqs[i].N + 31) / 32;
  }
  if (jobs.n == 0) return;
  colsum_multi_kernel<<<ctas, COLSUM_WARPS * 32, 0, st>>>(jobs);
  c->launches += 1;
}
