// Shared device helpers for libtnt (sm_100a).  fp32 throughout; no fast-math.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define TNT_LRELU_ALPHA 0.2f
#define TNT_BN_EPS 1e-3f
#define TNT_BN_MOMENTUM 0.99f
#define TNT_CE_EPS 1e-7f

// dropout site ids -- must match oracle/philox.py
#define SITE_INPUT 1
#define SITE_FEAT 2
#define SITE_TEXT 3
#define SITE_ATTN 4
#define SITE_LSTM_IN 5
#define SITE_LSTM_OUT 6
#define SITE_OUT 7
#define SITE_FEAT2 8

// ---- Philox4x32-10 (counter = (block, site, step, 0), key = seed) ------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                               uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0;
    uint32_t n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}

struct DropSite {
  uint32_t thr;     // keep iff r >= thr; thr == 0 => dropout disabled (multiplier 1)
  float scale;      // 1/(1-rate)
  uint32_t k0, k1;  // seed
  uint32_t site, step;
};

__host__ __device__ inline DropSite make_drop(float rate, int training, uint64_t seed, uint32_t site,
                                              uint32_t step) {
  DropSite d;
  d.thr = 0; d.scale = 1.0f;
  if (training && rate > 0.0f) {
    double t = floor((double)rate * 4294967296.0);
    d.thr = (uint32_t)t;
    d.scale = 1.0f / (1.0f - rate);
  }
  d.k0 = (uint32_t)(seed & 0xFFFFFFFFull); d.k1 = (uint32_t)(seed >> 32);
  d.site = site; d.step = step;
  return d;
}

// multiplier (0 or scale) of linear element `idx` of the site
__device__ __forceinline__ float drop_mult(const DropSite& d, uint64_t idx) {
  if (d.thr == 0) return 1.0f;
  uint4 r = philox4x32_10((uint32_t)(idx >> 2), d.site, d.step, 0u, d.k0, d.k1);
  uint32_t lane = (uint32_t)(idx & 3);
  uint32_t v = lane == 0 ? r.x : lane == 1 ? r.y : lane == 2 ? r.z : r.w;
  return v >= d.thr ? d.scale : 0.0f;
}

// ---- math -----------------------------------------------------------------------------------------
__device__ __forceinline__ float lrelu(float x) { return x > 0.0f ? x : TNT_LRELU_ALPHA * x; }
__device__ __forceinline__ float lrelu_grad_from_out(float y) { return y > 0.0f ? 1.0f : TNT_LRELU_ALPHA; }
// attention projections: slope 0.2 (attention.py:25-44 as built by lc_NIC.py:94-102) or 1 (= no activation, the
// plain Bahdanau form of ThinkAndTell/att_model.py:11-29)
__device__ __forceinline__ float lrelu_s(float x, float s) { return x > 0.0f ? x : s * x; }
__device__ __forceinline__ float lrelu_grad_s(float y, float s) { return y > 0.0f ? 1.0f : s; }
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// block-wide sum; `red` is shared scratch of >= 32 floats; all threads get the result
__device__ __forceinline__ float block_sum(float v, float* red) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : 0.0f;
  r = warp_sum(r);
  return r;
}
__device__ __forceinline__ float block_max(float v, float* red) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : -INFINITY;
  r = warp_max(r);
  return r;
}
