// Column sums out[n] = sum_m X[m][n]: the bias gradients of the time-parallel products of the decoder and the head
// (dense_inter / dense_out / LSTM bias of lc_NIC.py:118-159 and the attention biases).  Row-major storage.
#pragma once
#include "common.cuh"

// Column sums: out[n] = sum_m X[m][n]  (bias gradients).  One block of 32 warps per 32 columns; warp w owns rows
// w, w+32, ... and keeps four independent loads in flight; fixed summation order (no atomics).
#define COLSUM_WARPS 32
__device__ __forceinline__ void colsum_block(int M, int N, const float* __restrict__ X, int ldx, float* __restrict__ out,
                                             int blk) {
  __shared__ float red[COLSUM_WARPS][33];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int n = blk * 32 + lane;
  float s0 = 0.0f, s1 = 0.0f, s2 = 0.0f, s3 = 0.0f;
  if (n < N) {
    const float* p = X + n;
    int m = w;
    for (; m + 3 * COLSUM_WARPS < M; m += 4 * COLSUM_WARPS) {
      s0 += p[(size_t)m * ldx];
      s1 += p[(size_t)(m + COLSUM_WARPS) * ldx];
      s2 += p[(size_t)(m + 2 * COLSUM_WARPS) * ldx];
      s3 += p[(size_t)(m + 3 * COLSUM_WARPS) * ldx];
    }
    for (; m < M; m += COLSUM_WARPS) s0 += p[(size_t)m * ldx];
  }
  red[w][lane] = (s0 + s1) + (s2 + s3);
  __syncthreads();
  if (w == 0 && n < N) {
    float t = 0.0f;
#pragma unroll
    for (int i = 0; i < COLSUM_WARPS; ++i) t += red[i][lane];
    out[n] = t;
  }
}
__global__ void __launch_bounds__(COLSUM_WARPS * 32)
colsum_kernel(int M, int N, const float* __restrict__ X, int ldx, float* __restrict__ out) {
  colsum_block(M, N, X, ldx, out, blockIdx.x);
}
// several column sums in one launch (the bias gradients of a GEMM group are four 1..64-block kernels otherwise)
#define COLSUM_MAXJOBS 4
struct ColsumJob { const float* X; float* out; int M, N, ldx, cta0; };
struct ColsumJobs { ColsumJob j[COLSUM_MAXJOBS]; int n; };
__global__ void __launch_bounds__(COLSUM_WARPS * 32)
colsum_multi_kernel(const __grid_constant__ ColsumJobs jobs) {
  int k = 0;
#pragma unroll
  for (int i = 1; i < COLSUM_MAXJOBS; ++i)
    if (i < jobs.n && (int)blockIdx.x >= jobs.j[i].cta0) k = i;
  const ColsumJob& J = jobs.j[k];
  colsum_block(J.M, J.N, J.X, J.ldx, J.out, (int)blockIdx.x - J.cta0);
}
