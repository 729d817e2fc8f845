// Generic fp32 SIMT GEMM used for the time-parallel products of the decoder and the head
// (dense_inter / dense_out of lc_NIC.py:139-159,261 and their tape gradients).
// Row-major storage everywhere.  C[M,N] = opA(A) * opB(B) (+ bias[n]) (+ C if accumulate).
//   TA == false: A is [M,K] (lda)      TA == true: A is stored [K,M] (lda)  -> C = A^T B
//   TB == false: B is [K,N] (ldb)      TB == true: B is stored [N,K] (ldb)  -> C = A B^T
// grid.z > 1 (split K) or accumulate == 2: partial products are atomically added onto C.
#pragma once
#include "common.cuh"

#define GEMM_BM 64
#define GEMM_BN 64
#define GEMM_BK 16
#define GEMM_PAD 4

template <bool TA, bool TB>
__global__ void __launch_bounds__(256)
sgemm_kernel(int M, int N, int K, const float* __restrict__ A, int lda, const float* __restrict__ B, int ldb,
             float* __restrict__ C, int ldc, const float* __restrict__ bias, int accumulate, int k_per_split) {
  __shared__ __align__(16) float As[GEMM_BK][GEMM_BM + GEMM_PAD];
  __shared__ __align__(16) float Bs[GEMM_BK][GEMM_BN + GEMM_PAD];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;          // 16 x 16 threads, each a 4x4 micro-tile
  const int m0 = blockIdx.y * GEMM_BM, n0 = blockIdx.x * GEMM_BN;
  const int kbeg = blockIdx.z * k_per_split;
  const int kend = min(K, kbeg + k_per_split);
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

  for (int k0 = kbeg; k0 < kend; k0 += GEMM_BK) {
    // ---- stage A tile: As[k][m]
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int e = tid + i * 256;                       // 1024 elements
      int m, k;
      if (TA) { m = e & 63; k = e >> 6; } else { k = e & 15; m = e >> 4; }
      int gm = m0 + m, gk = k0 + k;
      float v = 0.0f;
      if (gm < M && gk < kend) v = TA ? A[(size_t)gk * lda + gm] : A[(size_t)gm * lda + gk];
      As[k][m] = v;
    }
    // ---- stage B tile: Bs[k][n]
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int e = tid + i * 256;
      int n, k;
      if (TB) { k = e & 15; n = e >> 4; } else { n = e & 63; k = e >> 6; }
      int gn = n0 + n, gk = k0 + k;
      float v = 0.0f;
      if (gn < N && gk < kend) v = TB ? B[(size_t)gn * ldb + gk] : B[(size_t)gk * ldb + gn];
      Bs[k][n] = v;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < GEMM_BK; ++k) {
      float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
  const bool split = gridDim.z > 1 || accumulate == 2;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int gm = m0 + ty * 4 + i;
    if (gm >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int gn = n0 + tx * 4 + j;
      if (gn >= N) continue;
      float v = acc[i][j];
      if (bias != nullptr && blockIdx.z == 0) v += bias[gn];
      float* c = &C[(size_t)gm * ldc + gn];
      if (split) atomicAdd(c, v);
      else *c = accumulate ? (*c + v) : v;
    }
  }
}

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
