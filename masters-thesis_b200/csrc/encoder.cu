// Parcel encoder: LocallyDense (AttemptFour/Model/layers.py:43-53) forward and tape gradient.
//   y_r = LeakyReLU(x[:, idx_r] @ W_r + b_r)   for every parcel r      (layers.py:45)
//   stack -> (B,R,G); BatchNormalization(axis=-1); Dropout               (layers.py:48-51)
// The image control (img_localDense.py:28-38) is the same computation with contiguous groups.
//
// HBM layout: x (B,N) row-major as the reference's generator delivers it
// (data_generator_guse.py:137); W packed (N_lab,G): parcel r owns rows off[r]..off[r+1]; Y/feats
// (B,R,G).  Kernels: enc_fwd_tf / enc_bwd_dw_t on a vertex-major copy of the batch (<= 64 rows per
// encoder, G == 32: the training configurations); enc_fwd_mma gathers straight from x for larger batches (decode);
// enc_fwd_kernel / enc_bwd_dw_kernel are the general form (any G, input dropout).
#include "tnt_internal.h"
#include <cstdlib>

#define ENC_BT 64      // batch rows per tile
#define ENC_KT 32      // gathered columns per smem stage (forward)
#define ENC_THREADS 128

// ------------------------------------------------------------------------------------------------
// forward: partial[slot][b][g] = sum_{k in work range} x[b, idx[k]] * W[k][g]
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(ENC_THREADS)
enc_fwd_kernel(const float* __restrict__ x, int N, int b_begin, int Bh, const int32_t* __restrict__ idx,
               const int64_t* __restrict__ off, const EncWork* __restrict__ work,
               const int32_t* __restrict__ slot_first, const float* __restrict__ W, int G,
               float* __restrict__ part, DropSite din) {
  __shared__ __align__(16) float Xs[ENC_BT][ENC_KT + 4];
  __shared__ __align__(16) float Ws[ENC_KT][32];
  const EncWork wk = work[blockIdx.x];
  const int bt0 = blockIdx.y * ENC_BT;
  const int64_t base = off[wk.r];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int bg = tid >> 3, gg = tid & 7;           // 16 b-groups x 8 g-groups, 4x4 each
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

  for (int k0 = wk.kbeg; k0 < wk.kend; k0 += ENC_KT) {
    const int kk = k0 + lane;
    const bool kok = kk < wk.kend;
    const int32_t vi = kok ? idx[base + kk] : 0;
    // gather X tile: warp w loads rows w, w+4, ... ; lanes run along the index list
#pragma unroll 4
    for (int i = 0; i < ENC_BT / 4; ++i) {
      int b = warp + i * 4;
      int gb = bt0 + b;
      float v = 0.0f;
      if (kok && gb < Bh) {
        size_t a = (size_t)(b_begin + gb) * N + vi;
        v = __ldg(&x[a]);
        if (din.thr) v *= drop_mult(din, a);
      }
      Xs[b][lane] = v;
    }
    // W tile: rows k0..k0+31, G columns (zero padded to 32)
#pragma unroll
    for (int i = 0; i < (ENC_KT * 32) / ENC_THREADS; ++i) {
      int e = tid + i * ENC_THREADS;
      int k = e >> 5, g = e & 31;
      float v = 0.0f;
      if (k0 + k < wk.kend && g < G) v = W[(size_t)(base + k0 + k) * G + g];
      Ws[k][g] = v;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < ENC_KT; k += 4) {
      float4 a[4], w4[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(&Xs[bg * 4 + i][k]);
#pragma unroll
      for (int i = 0; i < 4; ++i) w4[i] = *reinterpret_cast<const float4*>(&Ws[k + i][gg * 4]);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float av[4] = {a[i].x, a[i].y, a[i].z, a[i].w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          acc[i][0] = fmaf(av[q], w4[q].x, acc[i][0]);
          acc[i][1] = fmaf(av[q], w4[q].y, acc[i][1]);
          acc[i][2] = fmaf(av[q], w4[q].z, acc[i][2]);
          acc[i][3] = fmaf(av[q], w4[q].w, acc[i][3]);
        }
      }
    }
    __syncthreads();
  }
  const int slot = slot_first[wk.r] + wk.slot;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int gb = bt0 + bg * 4 + i;
    if (gb >= Bh) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int g = gg * 4 + j;
      if (g < G) part[((size_t)slot * Bh + gb) * G + g] = acc[i][j];
    }
  }
}

// ---- cp.async helpers for the pipelined kernels below: gathers go straight into a 3-stage shared-memory ring, so a
// CTA keeps two whole tiles of loads in flight while it multiplies the third.
__device__ __forceinline__ void cp_async4(void* dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N_> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N_) : "memory"); }

#define ENC_STAGES 3
#define ENC_XS_STRIDE (ENC_KT + 4)
#define ENC_STAGE_FLOATS (ENC_BT * ENC_XS_STRIDE + ENC_KT * 32)

// ---- row-major path for batches of more than 64 rows (greedy decode at large batch; no input dropout, G == 32):
// 4-byte cp.async gathers straight from x (B,N) into the ring; the 64 x 32 x 32 tile product runs as mma.sync m16n8k8
// TF32 with the 3xTF32 operand split (x = hi + lo, hi*hi + hi*lo + lo*hi: 22 mantissa bits, fp32 accumulate).  Each
// warp owns one k8 slice of every tile and the whole 64 x 32 output (16 accumulator fragments); the four warps are
// summed once at the end of the work item.  A per-slab transpose + the vertex-major kernel would move 2.7x the bytes.
__device__ __forceinline__ uint32_t tf32_hi(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

#define ENCM_WS_STRIDE 40                      // 32 + 8: conflict-free B-fragment reads
#define ENCM_STAGE_FLOATS (ENC_BT * ENC_XS_STRIDE + ENC_KT * ENCM_WS_STRIDE)
#define ENCM_SMEM (ENC_STAGES * ENCM_STAGE_FLOATS * sizeof(float))

__global__ void __launch_bounds__(ENC_THREADS, 5)
enc_fwd_mma_kernel(const float* __restrict__ x, int N, int b_begin, int Bh, const int32_t* __restrict__ idx,
                   const int64_t* __restrict__ off, const EncWork* __restrict__ work,
                   const int32_t* __restrict__ slot_first, const float* __restrict__ W, float* __restrict__ part) {
  extern __shared__ __align__(16) float enc_smem[];
  const EncWork wk = work[blockIdx.x];
  const int bt0 = blockIdx.y * ENC_BT;
  const int64_t base = off[wk.r];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int fg = lane >> 2, ft = lane & 3;         // fragment coordinates
  const int ntile = (wk.kend - wk.kbeg + ENC_KT - 1) / ENC_KT;
  const float* xrow = x + (size_t)(b_begin + bt0) * N;

  int32_t vi_next = (wk.kbeg + lane < wk.kend) ? __ldg(&idx[base + wk.kbeg + lane]) : -1;
  auto issue = [&](int tile) {
    float* Xs = enc_smem + (tile % ENC_STAGES) * ENCM_STAGE_FLOATS;
    float* Ws = Xs + ENC_BT * ENC_XS_STRIDE;
    const int k0 = wk.kbeg + tile * ENC_KT;
    const int32_t vi = vi_next;
    const int kn = k0 + ENC_KT + lane;               // index of the following tile: in flight across this one
    vi_next = (tile + 1 < ntile && kn < wk.kend) ? __ldg(&idx[base + kn]) : -1;
#pragma unroll
    for (int i = 0; i < ENC_BT / 4; ++i) {
      const int b = warp + i * 4;
      float* dst = Xs + b * ENC_XS_STRIDE + lane;
      if (vi >= 0 && bt0 + b < Bh) cp_async4(dst, xrow + (size_t)b * N + vi);
      else *dst = 0.0f;
    }
#pragma unroll
    for (int i = 0; i < (ENC_KT * 8) / ENC_THREADS; ++i) {
      const int e = tid + i * ENC_THREADS;           // float4 index: row k = e / 8
      const int k = e >> 3;
      float* dst = Ws + k * ENCM_WS_STRIDE + (e & 7) * 4;
      if (k0 + k < wk.kend) cp_async16(dst, W + (size_t)(base + k0 + k) * 32 + (e & 7) * 4);
      else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };

  float acc[4][4][4];                              // [m-tile][n-tile][fragment register]
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[i][j][q] = 0.0f;
#pragma unroll
  for (int s = 0; s < ENC_STAGES - 1; ++s) {
    if (s < ntile) issue(s);
    cp_async_commit();
  }
  for (int tile = 0; tile < ntile; ++tile) {
    if (tile + ENC_STAGES - 1 < ntile) issue(tile + ENC_STAGES - 1);
    cp_async_commit();
    cp_async_wait<ENC_STAGES - 1>();
    __syncthreads();
    const float* Xs = enc_smem + (tile % ENC_STAGES) * ENCM_STAGE_FLOATS;
    const float* Ws = Xs + ENC_BT * ENC_XS_STRIDE;
    const int kk = 8 * warp;                         // this warp's k8 slice of the tile
    uint32_t bh[4][2], bl[4][2];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const float v = Ws[(kk + ft + 4 * q) * ENCM_WS_STRIDE + 8 * nt + fg];
        bh[nt][q] = tf32_hi(v);
        bl[nt][q] = tf32_hi(v - __uint_as_float(bh[nt][q]));
      }
    }
#pragma unroll
    for (int mt = 0; mt < 4; ++mt) {
      uint32_t ah[4], al[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float v = Xs[(16 * mt + fg + 8 * (q & 1)) * ENC_XS_STRIDE + kk + ft + 4 * (q >> 1)];
        ah[q] = tf32_hi(v);
        al[q] = tf32_hi(v - __uint_as_float(ah[q]));
      }
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        mma_tf32(acc[mt][nt], al, bh[nt]);
        mma_tf32(acc[mt][nt], ah, bl[nt]);
        mma_tf32(acc[mt][nt], ah, bh[nt]);
      }
    }
    __syncthreads();
  }
  // sum the four warps' partial products through shared memory (the ring is free now): red[warp][64][32]
  cp_async_wait<0>();
  __syncthreads();
  float* red = enc_smem;
#pragma unroll
  for (int mt = 0; mt < 4; ++mt)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      float* r0 = red + (warp * 64 + 16 * mt + fg) * 32 + 8 * nt + 2 * ft;
      *reinterpret_cast<float2*>(r0) = make_float2(acc[mt][nt][0], acc[mt][nt][1]);
      *reinterpret_cast<float2*>(r0 + 8 * 32) = make_float2(acc[mt][nt][2], acc[mt][nt][3]);
    }
  __syncthreads();
  const int slot = slot_first[wk.r] + wk.slot;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int e = tid + i * ENC_THREADS;             // float4 index into [64][32]
    const int b = e >> 3;
    const float4 p0 = *reinterpret_cast<const float4*>(red + e * 4);
    const float4 p1 = *reinterpret_cast<const float4*>(red + 2048 + e * 4);
    const float4 p2 = *reinterpret_cast<const float4*>(red + 4096 + e * 4);
    const float4 p3 = *reinterpret_cast<const float4*>(red + 6144 + e * 4);
    if (bt0 + b < Bh)
      *reinterpret_cast<float4*>(part + ((size_t)slot * Bh + bt0 + b) * 32 + (e & 7) * 4) =
          make_float4((p0.x + p1.x) + (p2.x + p3.x), (p0.y + p1.y) + (p2.y + p3.y), (p0.z + p1.z) + (p2.z + p3.z),
                      (p0.w + p1.w) + (p2.w + p3.w));
  }
}

// ---- parcel-ordered path (Bh <= 64, no input dropout, G == 32) ---------------------------------------------------
// With x in its (B, N) layout a gathered column costs one 32-byte sector per 4-byte element (measured: 55 us per pass
// for the loads alone).  The batch is therefore re-laid ONCE per step: a coalesced read of x, a transpose through
// shared memory, and every vertex's 64 batch values written as one 256-byte row at EACH position the vertex has in the
// concatenated parcel index list (xP[n_lab][64]).  Both encoder passes then stream xP with whole-tile TMA bulk copies:
// no gather is left anywhere (encoder_tc.cu).
__global__ void __launch_bounds__(256)
enc_permute_kernel(const float* __restrict__ x, int N, int b_begin, int Bh, int vec, const int64_t* __restrict__ inv_off,
                   const int32_t* __restrict__ inv_pos, float* __restrict__ xP) {
  __shared__ float tile[64][65];
  const int v0 = blockIdx.x * 64;
  const int tid = threadIdx.x;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int b = (tid >> 4) + 16 * i, c4 = (tid & 15) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (b < Bh && v0 + c4 < N) {
      const float* p = x + (size_t)(b_begin + b) * N + v0 + c4;
      if (vec && v0 + c4 + 3 < N) v = __ldcs(reinterpret_cast<const float4*>(p));
      else {
        v.x = p[0];
        if (v0 + c4 + 1 < N) v.y = p[1];
        if (v0 + c4 + 2 < N) v.z = p[2];
        if (v0 + c4 + 3 < N) v.w = p[3];
      }
    }
    tile[b][c4] = v.x; tile[b][c4 + 1] = v.y; tile[b][c4 + 2] = v.z; tile[b][c4 + 3] = v.w;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int vl = (tid >> 4) + 16 * i, b4 = (tid & 15) * 4;
    if (v0 + vl < N) {
      const int64_t p0 = __ldg(inv_off + v0 + vl), p1 = __ldg(inv_off + v0 + vl + 1);
      const float4 row = make_float4(tile[b4][vl], tile[b4 + 1][vl], tile[b4 + 2][vl], tile[b4 + 3][vl]);
      for (int64_t q = p0; q < p1; ++q)
        *reinterpret_cast<float4*>(xP + (size_t)__ldg(inv_pos + q) * 64 + b4) = row;
    }
  }
}

// Y[b][r][g] = LeakyReLU(sum_slots part + bias); BN sums (fp64) per channel.
__global__ void __launch_bounds__(256)
enc_finalize_kernel(int b_begin, int Bh, int R, int G, const float* __restrict__ part,
                    const int32_t* __restrict__ slot_first, const float* __restrict__ bias,
                    float* __restrict__ Y, double* __restrict__ stats) {
  __shared__ double red[2][8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t rows = (int64_t)Bh * R;
  double s1 = 0.0, s2 = 0.0;
  for (int64_t row = (int64_t)blockIdx.x * 8 + warp; row < rows; row += (int64_t)gridDim.x * 8) {
    int b = (int)(row / R), r = (int)(row % R);
    if (lane < G) {
      float y = 0.0f;
      for (int s = slot_first[r]; s < slot_first[r + 1]; ++s) y += part[((size_t)s * Bh + b) * G + lane];
      y = lrelu(y + bias[r * G + lane]);
      Y[((size_t)(b_begin + b) * R + r) * G + lane] = y;
      s1 += (double)y; s2 += (double)y * (double)y;
    }
  }
  red[0][warp][lane] = s1; red[1][warp][lane] = s2;
  __syncthreads();
  if (warp == 0 && lane < G) {
    double a = 0.0, b2 = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) { a += red[0][i][lane]; b2 += red[1][i][lane]; }
    atomicAdd(&stats[lane], a);
    atomicAdd(&stats[32 + lane], b2);
  }
}

// coef[0]=inv (gamma*rsqrt(var+eps)), coef[1]=shift (beta-mean*inv), coef[2]=mean, coef[3]=rsqrt(var+eps)
__global__ void bn_prepare_kernel(int G, double n, int training, const double* __restrict__ stats,
                                  const float* __restrict__ gamma, const float* __restrict__ beta,
                                  float* __restrict__ mov_mean, float* __restrict__ mov_var,
                                  float* __restrict__ coef) {
  int g = threadIdx.x;
  if (g >= G) return;
  float mean, var;
  if (training) {
    double mu = stats[g] / n;
    double vv = stats[32 + g] / n - mu * mu;
    if (vv < 0.0) vv = 0.0;
    mean = (float)mu; var = (float)vv;
    mov_mean[g] = mov_mean[g] * TNT_BN_MOMENTUM + mean * (1.0f - TNT_BN_MOMENTUM);
    mov_var[g] = mov_var[g] * TNT_BN_MOMENTUM + var * (1.0f - TNT_BN_MOMENTUM);
  } else {
    mean = mov_mean[g]; var = mov_var[g];
  }
  float invstd = rsqrtf(var + TNT_BN_EPS);
  float inv = gamma[g] * invstd;
  coef[g] = inv; coef[32 + g] = beta[g] - mean * inv; coef[64 + g] = mean; coef[96 + g] = invstd;
}

// feats = dropout(BN(Y));  Pf = LeakyReLU(feats @ W1 + b1)  (attention.py:31 -- time-invariant, hoisted)
__global__ void __launch_bounds__(256)
bn_apply_kernel(int b_begin, int Bh, int R, int G, int A, const float* __restrict__ Y,
                const float* __restrict__ coef, const float* __restrict__ W1, const float* __restrict__ b1,
                float* __restrict__ feats, float* __restrict__ Pf, DropSite d1, DropSite d2, float att_slope) {
  __shared__ float W1s[32][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int e = threadIdx.x; e < 32 * 32; e += blockDim.x) {
    int g = e >> 5, a = e & 31;
    W1s[g][a] = (g < G && a < A) ? W1[g * A + a] : 0.0f;
  }
  __syncthreads();
  const float inv = lane < G ? coef[lane] : 0.0f, shift = lane < G ? coef[32 + lane] : 0.0f;
  const float bias = lane < A ? b1[lane] : 0.0f;
  const int64_t rows = (int64_t)Bh * R;
  for (int64_t row = (int64_t)blockIdx.x * 8 + warp; row < rows; row += (int64_t)gridDim.x * 8) {
    int b = (int)(row / R), r = (int)(row % R);
    size_t o = ((size_t)(b_begin + b) * R + r);
    float f = 0.0f;
    if (lane < G) {
      f = Y[o * G + lane] * inv + shift;
      uint64_t di = ((uint64_t)b * R + r) * G + lane;     // index local to this encoder's half
      f *= drop_mult(d1, di);
      f *= drop_mult(d2, di);
      feats[o * G + lane] = f;
    }
    float p = bias;
#pragma unroll 8
    for (int g = 0; g < 32; ++g) p = fmaf(__shfl_sync(0xffffffffu, f, g), W1s[g][lane], p);
    if (lane < A) Pf[o * A + lane] = lrelu_s(p, att_slope);
  }
}

// ------------------------------------------------------------------------------------------------
// backward
// ------------------------------------------------------------------------------------------------
// sums over (b,r) of dy and dy*xhat per channel, dy = dfeats * dropout multipliers
__global__ void __launch_bounds__(256)
bn_bwd_stats_kernel(int b_begin, int Bh, int R, int G, const float* __restrict__ Y,
                    const float* __restrict__ dfeats, const float* __restrict__ coef,
                    double* __restrict__ stats, DropSite d1, DropSite d2) {
  __shared__ double red[2][8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float mean = lane < G ? coef[64 + lane] : 0.0f, invstd = lane < G ? coef[96 + lane] : 0.0f;
  const int64_t rows = (int64_t)Bh * R;
  double s1 = 0.0, s2 = 0.0;
  for (int64_t row = (int64_t)blockIdx.x * 8 + warp; row < rows; row += (int64_t)gridDim.x * 8) {
    int b = (int)(row / R), r = (int)(row % R);
    if (lane < G) {
      size_t o = ((size_t)(b_begin + b) * R + r) * G + lane;
      uint64_t di = ((uint64_t)b * R + r) * G + lane;
      float dy = dfeats[o] * drop_mult(d1, di) * drop_mult(d2, di);
      float xh = (Y[o] - mean) * invstd;
      s1 += (double)dy; s2 += (double)dy * (double)xh;
    }
  }
  red[0][warp][lane] = s1; red[1][warp][lane] = s2;
  __syncthreads();
  if (warp == 0 && lane < G) {
    double a = 0.0, b2 = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) { a += red[0][i][lane]; b2 += red[1][i][lane]; }
    atomicAdd(&stats[lane], a);
    atomicAdd(&stats[32 + lane], b2);
  }
}

// dYpre = LeakyReLU'(Y) * gamma*invstd * (dy - mean(dy) - xhat*mean(dy*xhat));  dgamma, dbeta
__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(int b_begin, int Bh, int R, int G, double n_global, float local_frac,
                    const float* __restrict__ Y, const float* __restrict__ dfeats,
                    const float* __restrict__ coef, const double* __restrict__ stats,
                    float* __restrict__ dYpre, float* __restrict__ dgamma, float* __restrict__ dbeta,
                    DropSite d1, DropSite d2) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float inv = lane < G ? coef[lane] : 0.0f;
  const float mean = lane < G ? coef[64 + lane] : 0.0f, invstd = lane < G ? coef[96 + lane] : 0.0f;
  const float mdy = lane < G ? (float)(stats[lane] / n_global) : 0.0f;
  const float mdyx = lane < G ? (float)(stats[32 + lane] / n_global) : 0.0f;
  if (blockIdx.x == 0 && warp == 0 && lane < G) {
    dbeta[lane] = (float)(stats[lane]) * local_frac;        // all-reduced sums: each rank keeps its share
    dgamma[lane] = (float)(stats[32 + lane]) * local_frac;
  }
  const int64_t rows = (int64_t)Bh * R;
  for (int64_t row = (int64_t)blockIdx.x * 8 + warp; row < rows; row += (int64_t)gridDim.x * 8) {
    int b = (int)(row / R), r = (int)(row % R);
    if (lane < G) {
      size_t o = ((size_t)(b_begin + b) * R + r) * G + lane;
      uint64_t di = ((uint64_t)b * R + r) * G + lane;
      float y = Y[o];
      float dy = dfeats[o] * drop_mult(d1, di) * drop_mult(d2, di);
      float xh = (y - mean) * invstd;
      float dx = inv * (dy - mdy - xh * mdyx);
      dYpre[o] = dx * lrelu_grad_from_out(y);
    }
  }
}

// db_r[g] = sum_b dYpre[b][r][g]
__global__ void enc_bwd_bias_kernel(int b_begin, int Bh, int R, int G, const float* __restrict__ dYpre,
                                    float* __restrict__ db) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= R * G) return;
  float s = 0.0f;
  for (int b = 0; b < Bh; ++b) s += dYpre[(size_t)(b_begin + b) * R * G + e];
  db[e] = s;
}

// dW[off_r + k][g] = sum_b x[b, idx[k]] * dYpre[b][r][g]
#define ENCB_KT 64
__global__ void __launch_bounds__(ENC_THREADS)
enc_bwd_dw_kernel(const float* __restrict__ x, int N, int b_begin, int Bh, int R, const int32_t* __restrict__ idx,
                  const int64_t* __restrict__ off, const EncWork* __restrict__ work, int G,
                  const float* __restrict__ dYpre, float* __restrict__ dW, DropSite din) {
  __shared__ __align__(16) float Xs[ENC_BT][ENCB_KT + 4];
  __shared__ __align__(16) float Ds[ENC_BT][32];
  const EncWork wk = work[blockIdx.x];
  const int64_t base = off[wk.r];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int kg = tid >> 3, gg = tid & 7;           // 16 k-groups x 8 g-groups; 4 k x 4 g each

  for (int k0 = wk.kbeg; k0 < wk.kend; k0 += ENCB_KT) {
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
    const int kk0 = k0 + lane, kk1 = k0 + 32 + lane;
    const bool ok0 = kk0 < wk.kend, ok1 = kk1 < wk.kend;
    const int32_t v0 = ok0 ? idx[base + kk0] : 0, v1 = ok1 ? idx[base + kk1] : 0;
    for (int bt0 = 0; bt0 < Bh; bt0 += ENC_BT) {
#pragma unroll 4
      for (int i = 0; i < ENC_BT / 4; ++i) {
        int b = warp + i * 4;
        int gb = bt0 + b;
        float a0 = 0.0f, a1 = 0.0f;
        if (gb < Bh) {
          size_t rowo = (size_t)(b_begin + gb) * N;
          if (ok0) { a0 = __ldg(&x[rowo + v0]); if (din.thr) a0 *= drop_mult(din, rowo + v0); }
          if (ok1) { a1 = __ldg(&x[rowo + v1]); if (din.thr) a1 *= drop_mult(din, rowo + v1); }
        }
        Xs[b][lane] = a0; Xs[b][32 + lane] = a1;
      }
#pragma unroll
      for (int i = 0; i < (ENC_BT * 32) / ENC_THREADS; ++i) {
        int e = tid + i * ENC_THREADS;
        int b = e >> 5, g = e & 31;
        int gb = bt0 + b;
        float v = 0.0f;
        if (gb < Bh && g < G) v = dYpre[((size_t)(b_begin + gb) * R + wk.r) * G + g];
        Ds[b][g] = v;
      }
      __syncthreads();
#pragma unroll 8
      for (int b = 0; b < ENC_BT; ++b) {
        float4 a = *reinterpret_cast<const float4*>(&Xs[b][kg * 4]);
        float4 d = *reinterpret_cast<const float4*>(&Ds[b][gg * 4]);
        float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          acc[i][0] = fmaf(av[i], d.x, acc[i][0]);
          acc[i][1] = fmaf(av[i], d.y, acc[i][1]);
          acc[i][2] = fmaf(av[i], d.z, acc[i][2]);
          acc[i][3] = fmaf(av[i], d.w, acc[i][3]);
        }
      }
      __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int k = k0 + kg * 4 + i;
      if (k >= wk.kend) continue;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        int g = gg * 4 + j;
        if (g < G) dW[(size_t)(base + k) * G + g] = acc[i][j];
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
// parcel-ordered tensor-core path: no input dropout, G == 32, at most 64 rows per encoder
static bool enc_t_path(tnt_ctx* c, const DropSite& din, int Bh) {
  return din.thr == 0 && c->cfg.G == 32 && Bh <= ENC_BT && ws(c, "xT") != nullptr;
}
static float* enc_xT(tnt_ctx* c, int e, const float* betas, int b_begin, int Bh, cudaStream_t st) {
  float* xP = wsT<float>(c, "xT");
  for (int k = 0; k < e; ++k) xP += (size_t)(c->enc[k].n_lab + 128) * 64;
  if (c->xT_src[e] != betas || c->xT_b0[e] != b_begin || c->xT_Bh[e] != Bh) {
    TNT_TRACE(c, "enc_permute_kernel", st);
    const int vec = ((((uintptr_t)betas) & 15) == 0 && (c->cfg.N & 3) == 0) ? 1 : 0;
    enc_permute_kernel<<<(c->cfg.N + 63) / 64, 256, 0, st>>>(betas, c->cfg.N, b_begin, Bh, vec, c->enc[e].inv_off,
                                                             c->enc[e].inv_pos, xP);
    c->launches += 1;
    c->xT_src[e] = betas; c->xT_b0[e] = b_begin; c->xT_Bh[e] = Bh;
  }
  return xP;
}

static inline int rows_grid(tnt_ctx* c, int64_t rows) {
  int64_t need = (rows + 7) / 8;
  int64_t cap = (int64_t)c->sm_count * 8;
  return (int)(need < cap ? (need > 0 ? need : 1) : cap);
}

void enc_forward(tnt_ctx* c, int e, const float* betas, int b_begin, int Bh, int B_total, int training,
                 uint64_t seed, cudaStream_t st) {
  const tnt_config& f = c->cfg;
  EncTables& t = c->enc[e];
  double* stats = wsT<double>(c, "bn_fwd_stats") + e * 64;
  cudaMemsetAsync(stats, 0, 64 * sizeof(double), st);
  DropSite din = make_drop(f.dropout_input, training, seed, SITE_INPUT, 0);
  dim3 grid(t.n_work, (Bh + ENC_BT - 1) / ENC_BT);
  if (enc_t_path(c, din, Bh)) {
    c->xT_src[e] = nullptr;                                  // the caller may have rewritten the batch in place: transpose again
    const float* xT = enc_xT(c, e, betas, b_begin, Bh, st);
    TNT_TRACE(c, "enc_fwd_tc_kernel", st);
    enc_fwd_tc(c, e, xT, Bh, st);
    c->variant_enc_fwd = "enc_fwd_tc_kernel (tcgen05 + TMEM, TMA tile streams of the parcel-ordered batch)";
  } else if (din.thr == 0 && f.G == 32) {
    TNT_TRACE(c, "enc_fwd_mma_kernel", st);
    enc_fwd_mma_kernel<<<grid, ENC_THREADS, ENCM_SMEM, st>>>(betas, f.N, b_begin, Bh, t.idx, t.off, t.work, t.slot_first,
                                                             c->w + t.w_off, wsT<float>(c, "enc_part"));
    c->variant_enc_fwd = "enc_fwd_mma_kernel";
  } else {
    TNT_TRACE(c, "enc_fwd_kernel", st);
    enc_fwd_kernel<<<grid, ENC_THREADS, 0, st>>>(betas, f.N, b_begin, Bh, t.idx, t.off, t.work, t.slot_first,
                                                 c->w + t.w_off, f.G, wsT<float>(c, "enc_part"), din);
    c->variant_enc_fwd = "enc_fwd_kernel";
  }
  TNT_TRACE(c, "enc_finalize_kernel", st);
  enc_finalize_kernel<<<rows_grid(c, (int64_t)Bh * f.R), 256, 0, st>>>(
      b_begin, Bh, f.R, f.G, wsT<float>(c, "enc_part"), t.slot_first, c->w + t.b_off, wsT<float>(c, "Y"), stats);
  c->launches += 2;
}

void enc_bn_finish(tnt_ctx* c, int e, int b_begin, int Bh, int Bh_global, int training, uint64_t seed,
                   cudaStream_t st) {
  const tnt_config& f = c->cfg;
  EncTables& t = c->enc[e];
  double* stats = wsT<double>(c, "bn_fwd_stats") + e * 64;
  float* coef = wsT<float>(c, "bn_coef") + e * 128;
  side_join(c, st);            // first reader of a decoder variable (W1): a deferred optimizer part must be done
  TNT_TRACE(c, "bn_prepare + bn_apply (feats, Pf)", st);
  bn_prepare_kernel<<<1, 32, 0, st>>>(f.G, (double)Bh_global * f.R, training, stats, c->w + t.gamma_off,
                                      c->w + t.beta_off, c->state + t.mean_off, c->state + t.var_off, coef);
  DropSite d1 = make_drop(f.dropout_features, training, seed, SITE_FEAT, e);
  DropSite d2 = make_drop(f.second_feature_dropout ? f.dropout_features : 0.0f, training, seed, SITE_FEAT2, e);
  bn_apply_kernel<<<rows_grid(c, (int64_t)Bh * f.R), 256, 0, st>>>(
      b_begin, Bh, f.R, f.G, f.A, wsT<float>(c, "Y"), coef, c->w + c->o_W1, c->w + c->o_b1,
      wsT<float>(c, "feats"), wsT<float>(c, "Pf"), d1, d2, att_slope_of(f));
  c->launches += 2;
}

void enc_bwd_stats(tnt_ctx* c, int e, int b_begin, int Bh, uint64_t seed, cudaStream_t st) {
  const tnt_config& f = c->cfg;
  double* stats = wsT<double>(c, "bn_bwd_stats") + e * 64;
  float* coef = wsT<float>(c, "bn_coef") + e * 128;
  TNT_TRACE(c, "bn_bwd_stats_kernel", st);
  cudaMemsetAsync(stats, 0, 64 * sizeof(double), st);
  DropSite d1 = make_drop(f.dropout_features, 1, seed, SITE_FEAT, e);
  DropSite d2 = make_drop(f.second_feature_dropout ? f.dropout_features : 0.0f, 1, seed, SITE_FEAT2, e);
  bn_bwd_stats_kernel<<<rows_grid(c, (int64_t)Bh * f.R), 256, 0, st>>>(
      b_begin, Bh, f.R, f.G, wsT<float>(c, "Y"), wsT<float>(c, "dfeats"), coef, stats, d1, d2);
  c->launches += 1;
}

void enc_bwd_finish(tnt_ctx* c, int e, const float* betas, int b_begin, int Bh, int Bh_global, uint64_t seed,
                    cudaStream_t st) {
  const tnt_config& f = c->cfg;
  EncTables& t = c->enc[e];
  double* stats = wsT<double>(c, "bn_bwd_stats") + e * 64;
  float* coef = wsT<float>(c, "bn_coef") + e * 128;
  DropSite d1 = make_drop(f.dropout_features, 1, seed, SITE_FEAT, e);
  DropSite d2 = make_drop(f.second_feature_dropout ? f.dropout_features : 0.0f, 1, seed, SITE_FEAT2, e);
  DropSite din = make_drop(f.dropout_input, 1, seed, SITE_INPUT, 0);
  { TNT_TRACE(c, "bn_bwd_apply + enc_bwd_bias", st);
  bn_bwd_apply_kernel<<<rows_grid(c, (int64_t)Bh * f.R), 256, 0, st>>>(
      b_begin, Bh, f.R, f.G, (double)Bh_global * f.R, (float)Bh / (float)Bh_global, wsT<float>(c, "Y"),
      wsT<float>(c, "dfeats"), coef, stats, wsT<float>(c, "dYpre"), c->g + t.gamma_off, c->g + t.beta_off, d1, d2);
  int rg = f.R * f.G;
  enc_bwd_bias_kernel<<<(rg + 255) / 256, 256, 0, st>>>(b_begin, Bh, f.R, f.G, wsT<float>(c, "dYpre"),
                                                        c->g + t.b_off);
  }
  if (enc_t_path(c, din, Bh)) {
    const float* xT = enc_xT(c, e, betas, b_begin, Bh, st);   // normally still valid from the forward pass of this step
    TNT_TRACE(c, "enc_bwd_dw_tc_kernel", st);
    enc_bwd_dw_tc(c, e, xT, b_begin, Bh, st);
    c->launches += 2;
    c->variant_enc_bwd = "enc_bwd_dw_tc_kernel (tcgen05 + TMEM, TMA tile streams of the parcel-ordered batch)";
    return;
  }
  TNT_TRACE(c, "enc_bwd_dw_kernel", st);
  enc_bwd_dw_kernel<<<t.n_work, ENC_THREADS, 0, st>>>(betas, f.N, b_begin, Bh, f.R, t.idx, t.off, t.work, f.G,
                                                      wsT<float>(c, "dYpre"), c->g + t.w_off, din);
  c->variant_enc_bwd = "enc_bwd_dw_kernel";
  c->launches += 3;
}
