// Launchers of the column-sum kernels (gemm.cuh): bias gradients of the time-parallel products.
#include "tnt_internal.h"
#include "gemm.cuh"

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
