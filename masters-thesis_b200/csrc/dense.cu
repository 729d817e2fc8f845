// Dense voxel encoder of AttemptFour/Model/NIC.py (SURVEY 8a row A1''):
//   features = BatchNormalization()( expand_dims( Dropout( LeakyReLU_0.2( dropout_input(x) @ W + b ) ), 1 ) )
//   (NIC.py:64-69 layer, :122-128 call), W (K, E) with K up to 327 684 and E = 512: a skinny GEMM (M = batch <= 64)
// whose cost is reading W (671 MB) -- forward -- and writing dW (671 MB) -- backward; plus per-variable clip + Adam
// over 168 M parameters.  Standalone object (tnt_dense) next to tnt_ctx: the reference's NIC.py model feeds these
// features to a plain LSTM, which is outside the hot path of SURVEY 8.
//
// Layout: the batch is first written vertex-major, xT[K][64] (input dropout applied on the way), so that both GEMMs
// read 256-byte rows; W and dW stay in Keras' (K, E) row-major layout.
//   forward : grid (E/128, S): CTA = 64 x 128 output tile over one K slab, 3-stage cp.async ring, 8x8 fp32 register
//             tile per thread (FFMA, the fp32-exact path), partial tiles reduced by dense_act_kernel.
//   backward: grid (E/128, n): dY slab (64 x 128) resident in shared memory, CTA walks 64-row tiles of xT and
//             writes each 64 x 128 tile of dW exactly once.
#include "tnt_internal.h"
#include <string>

#define DN_THREADS 128
#define DN_KT 32          // K rows per forward stage
#define DN_NB 128         // output columns per CTA
#define DN_STAGES 3
#define DN_BT 64          // batch rows (padded)
#define DN_KB 64          // K rows per backward tile

__device__ __forceinline__ void dn_cp16(void* dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void dn_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N_> __device__ __forceinline__ void dn_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N_) : "memory"); }

// x (B, K) row-major -> xT[K][64]; rows b >= B are zero; dropout_input (NIC.py:122) applied here
__global__ void __launch_bounds__(256)
dense_transpose_kernel(const float* __restrict__ x, int K, int B, float* __restrict__ xT, DropSite din) {
  __shared__ float tile[64][65];
  const int k0 = blockIdx.x * 64;
  const int c = threadIdx.x & 63, r4 = threadIdx.x >> 6;
#pragma unroll 4
  for (int i = 0; i < 16; ++i) {
    const int b = r4 + 4 * i;
    float v = 0.0f;
    if (b < B && k0 + c < K) {
      v = __ldcs(x + (size_t)b * K + k0 + c);
      v *= drop_mult(din, (uint64_t)b * K + k0 + c);
    }
    tile[b][c] = v;
  }
  __syncthreads();
#pragma unroll 4
  for (int i = 0; i < 16; ++i) {
    const int kk = r4 + 4 * i;
    if (k0 + kk < K) xT[(size_t)(k0 + kk) * 64 + c] = tile[c][kk];
  }
}

// part[s][b][n] = sum over the K slab s of xT[k][b] * W[k][n]
#define DN_FWD_STAGE (DN_KT * DN_BT + DN_KT * DN_NB)
#define DN_FWD_SMEM (DN_STAGES * DN_FWD_STAGE * sizeof(float))
__global__ void __launch_bounds__(DN_THREADS, 3)
dense_fwd_kernel(const float* __restrict__ xT, const float* __restrict__ W, int K, int E, int kslab,
                 float* __restrict__ part) {
  extern __shared__ __align__(16) float dn_smem[];
  const int n0 = blockIdx.x * DN_NB;
  const int kbeg = blockIdx.y * kslab;
  const int kend = min(K, kbeg + kslab);
  const int ntile = (kend - kbeg + DN_KT - 1) / DN_KT;
  const int tid = threadIdx.x;
  const int tb = tid & 7, tn = tid >> 3;                 // b: 4tb.., 32+4tb..;  n: 4tn.., 64+4tn..
  auto issue = [&](int tile) {
    float* Xs = dn_smem + (tile % DN_STAGES) * DN_FWD_STAGE;
    float* Ws = Xs + DN_KT * DN_BT;
    const int k0 = kbeg + tile * DN_KT;
#pragma unroll
    for (int i = 0; i < (DN_KT * DN_BT / 4) / DN_THREADS; ++i) {
      const int e = tid + i * DN_THREADS;
      const int k = e >> 4, c4 = (e & 15) * 4;
      float* dst = Xs + k * DN_BT + c4;
      if (k0 + k < kend) dn_cp16(dst, xT + (size_t)(k0 + k) * 64 + c4);
      else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int i = 0; i < (DN_KT * DN_NB / 4) / DN_THREADS; ++i) {
      const int e = tid + i * DN_THREADS;
      const int k = e >> 5, c4 = (e & 31) * 4;
      float* dst = Ws + k * DN_NB + c4;
      if (k0 + k < kend) dn_cp16(dst, W + (size_t)(k0 + k) * E + n0 + c4);
      else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.0f;
#pragma unroll
  for (int s = 0; s < DN_STAGES - 1; ++s) {
    if (s < ntile) issue(s);
    dn_commit();
  }
  for (int tile = 0; tile < ntile; ++tile) {
    if (tile + DN_STAGES - 1 < ntile) issue(tile + DN_STAGES - 1);
    dn_commit();
    dn_wait<DN_STAGES - 1>();
    __syncthreads();
    const float* Xs = dn_smem + (tile % DN_STAGES) * DN_FWD_STAGE;
    const float* Ws = Xs + DN_KT * DN_BT;
#pragma unroll 8
    for (int k = 0; k < DN_KT; ++k) {
      const float4 xa = *reinterpret_cast<const float4*>(Xs + k * DN_BT + 4 * tb);
      const float4 xb = *reinterpret_cast<const float4*>(Xs + k * DN_BT + 32 + 4 * tb);
      const float4 wa = *reinterpret_cast<const float4*>(Ws + k * DN_NB + 4 * tn);
      const float4 wb = *reinterpret_cast<const float4*>(Ws + k * DN_NB + 64 + 4 * tn);
      const float xv[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
      const float wv[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(xv[i], wv[j], acc[i][j]);
    }
    __syncthreads();
  }
  float* out = part + (size_t)blockIdx.y * DN_BT * E;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int b = (i < 4 ? 4 * tb + i : 32 + 4 * tb + (i - 4));
    float* row = out + (size_t)b * E + n0;
    *reinterpret_cast<float4*>(row + 4 * tn) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    *reinterpret_cast<float4*>(row + 64 + 4 * tn) = make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]);
  }
}

// act[b][e] = Dropout( LeakyReLU( sum_s part[s][b][e] + bias[e] ) )          (NIC.py:125-126)
__global__ void __launch_bounds__(256)
dense_act_kernel(const float* __restrict__ part, int S, int B, int E, const float* __restrict__ bias,
                 float* __restrict__ act, DropSite dfeat) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * E) return;
  const int b = i / E, e = i % E;
  float s = 0.0f;
  for (int k = 0; k < S; ++k) s += part[((size_t)k * DN_BT + b) * E + e];
  act[i] = lrelu(s + bias[e]) * drop_mult(dfeat, (uint64_t)b * E + e);
}

// BatchNormalization(axis=-1) on (B, 1, E): per channel over the batch (NIC.py:128).  One thread per channel.
__global__ void __launch_bounds__(128)
dense_bn_fwd_kernel(const float* __restrict__ act, int B, int E, const float* __restrict__ gamma,
                    const float* __restrict__ beta, float* __restrict__ mov_mean, float* __restrict__ mov_var,
                    int training, float* __restrict__ feats, float* __restrict__ save /* [2][E]: mean, invstd */) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  float mean, var;
  if (training) {
    double s = 0.0;
    for (int b = 0; b < B; ++b) s += (double)act[(size_t)b * E + e];
    const double mu = s / B;
    double q = 0.0;
    for (int b = 0; b < B; ++b) { const double d = (double)act[(size_t)b * E + e] - mu; q += d * d; }
    mean = (float)mu; var = (float)(q / B);                       // biased, as tf.nn.moments
    mov_mean[e] = mov_mean[e] * TNT_BN_MOMENTUM + mean * (1.0f - TNT_BN_MOMENTUM);
    mov_var[e] = mov_var[e] * TNT_BN_MOMENTUM + var * (1.0f - TNT_BN_MOMENTUM);
  } else {
    mean = mov_mean[e]; var = mov_var[e];
  }
  const float invstd = rsqrtf(var + TNT_BN_EPS);
  const float inv = gamma[e] * invstd;
  const float shift = beta[e] - mean * inv;
  for (int b = 0; b < B; ++b) feats[(size_t)b * E + e] = act[(size_t)b * E + e] * inv + shift;
  save[e] = mean; save[E + e] = invstd;
}

// d features -> BatchNorm backward -> dropout mask -> LeakyReLU' -> dpre (B, E); d gamma, d beta, d bias
__global__ void __launch_bounds__(128)
dense_bn_bwd_kernel(const float* __restrict__ dfeats, const float* __restrict__ act, int B, int E,
                    const float* __restrict__ gamma, const float* __restrict__ save, float* __restrict__ dpre,
                    float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dbias, DropSite dfeat) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  const float mean = save[e], invstd = save[E + e];
  double sdy = 0.0, sdyx = 0.0;
  for (int b = 0; b < B; ++b) {
    const float dy = dfeats[(size_t)b * E + e];
    const float xh = (act[(size_t)b * E + e] - mean) * invstd;
    sdy += (double)dy; sdyx += (double)dy * (double)xh;
  }
  dgamma[e] = (float)sdyx; dbeta[e] = (float)sdy;
  const float m1 = (float)(sdy / B), m2 = (float)(sdyx / B);
  const float gi = gamma[e] * invstd;
  float sb = 0.0f;
  for (int b = 0; b < B; ++b) {
    const float a = act[(size_t)b * E + e];
    const float dy = dfeats[(size_t)b * E + e];
    const float xh = (a - mean) * invstd;
    float d = gi * (dy - m1 - xh * m2);
    d *= drop_mult(dfeat, (uint64_t)b * E + e);              // a dropped unit has act == 0 and gets no gradient
    d *= lrelu_grad_from_out(a);                             // dropout scales by a positive factor: sign(act) = sign(pre)
    dpre[(size_t)b * E + e] = d;
    sb += d;
  }
  for (int b = B; b < DN_BT; ++b) dpre[(size_t)b * E + e] = 0.0f;   // padded rows of the dW GEMM
  dbias[e] = sb;
}

// dW[k][n] = sum_b xT[k][b] * dpre[b][n]: every 64 x 128 tile of dW is written exactly once
#define DN_BWD_SMEM ((DN_BT * DN_NB + 2 * DN_KB * DN_BT) * sizeof(float))
__global__ void __launch_bounds__(DN_THREADS, 3)
dense_dw_kernel(const float* __restrict__ xT, const float* __restrict__ dpre, int K, int E, float* __restrict__ dW) {
  extern __shared__ __align__(16) float dn_smem[];
  float* Ys = dn_smem;                                   // [64 b][128 n]
  float* Xs = dn_smem + DN_BT * DN_NB;                   // [2][64 k][64 b]
  const int n0 = blockIdx.x * DN_NB;
  const int tid = threadIdx.x;
  const int tk = tid >> 4, te = tid & 15;                // k rows 8tk..8tk+7;  n: 4te.., 64+4te..
  const int ntiles = (K + DN_KB - 1) / DN_KB;
  auto issue = [&](int t, int buf) {
    float* dst0 = Xs + buf * DN_KB * DN_BT;
    const int k0 = t * DN_KB;
#pragma unroll
    for (int i = 0; i < (DN_KB * DN_BT / 4) / DN_THREADS; ++i) {
      const int e = tid + i * DN_THREADS;
      const int k = e >> 4, c4 = (e & 15) * 4;
      float* dst = dst0 + k * DN_BT + c4;
      if (k0 + k < K) dn_cp16(dst, xT + (size_t)(k0 + k) * 64 + c4);
      else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
#pragma unroll
  for (int i = 0; i < (DN_BT * DN_NB / 4) / DN_THREADS; ++i) {
    const int e = tid + i * DN_THREADS;
    const int b = e >> 5, c4 = (e & 31) * 4;
    dn_cp16(Ys + b * DN_NB + c4, dpre + (size_t)b * E + n0 + c4);
  }
  int t = blockIdx.y;
  if (t < ntiles) issue(t, 0);
  dn_commit();
  int buf = 0;
  for (; t < ntiles; t += gridDim.y, buf ^= 1) {
    if (t + (int)gridDim.y < ntiles) issue(t + gridDim.y, buf ^ 1);
    dn_commit();
    dn_wait<1>();
    __syncthreads();
    const float* X = Xs + buf * DN_KB * DN_BT + (8 * tk) * DN_BT;
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = 0.0f;
#pragma unroll 2
    for (int b4 = 0; b4 < DN_BT; b4 += 4) {
      float4 xq[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) xq[i] = *reinterpret_cast<const float4*>(X + i * DN_BT + b4);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 ya = *reinterpret_cast<const float4*>(Ys + (b4 + q) * DN_NB + 4 * te);
        const float4 yb = *reinterpret_cast<const float4*>(Ys + (b4 + q) * DN_NB + 64 + 4 * te);
        const float yv[8] = {ya.x, ya.y, ya.z, ya.w, yb.x, yb.y, yb.z, yb.w};
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float xv = q == 0 ? xq[i].x : (q == 1 ? xq[i].y : (q == 2 ? xq[i].z : xq[i].w));
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(xv, yv[j], acc[i][j]);
        }
      }
    }
    const int k0 = t * DN_KB + 8 * tk;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (k0 + i < K) {
        float* row = dW + (size_t)(k0 + i) * E + n0;
        __stcs(reinterpret_cast<float4*>(row + 4 * te), make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]));
        __stcs(reinterpret_cast<float4*>(row + 64 + 4 * te), make_float4(acc[i][4], acc[i][5], acc[i][6], acc[i][7]));
      }
    }
    __syncthreads();
  }
  dn_wait<0>();
}

// ------------------------------------------------------------------------------------------------
// host side / C ABI
// ------------------------------------------------------------------------------------------------
#define DN_OPT_CHUNK 8192

// ---- scratch of tnt_dense_forward_noalloc: lives in the module image, never cudaMalloc'ed ------------------------
// Every GPU that loads libtnt.so holds these 32.1 MiB until the process exits.  Layout (floats):
//   [0, XT)            xT     DN_NA_KMAX x 64   transposed input
//   [XT, XT + PART)    part   split-K partial tiles, S x 64 x E
//   then               act    64 x DN_NA_EMAX,   save 2 x DN_NA_EMAX
// Every region a call reads is rewritten in full by that call first (xT rows < K incl. the zero rows b >= B, all
// S x 64 x E partial tiles, act, save), so nothing has to start at zero and nothing is cleared; stale contents of the
// previous call beyond those ranges are never read.  ALL calls on a GPU share the array: two calls that overlap on the
// GPU (different streams, or concurrent host threads) race.
#define DN_NA_KMAX 65536
#define DN_NA_EMAX 512
#define DN_NA_XT ((size_t)DN_NA_KMAX * 64)
#define DN_NA_PART ((size_t)4 << 20)
#define DN_NA_FLOATS (DN_NA_XT + DN_NA_PART + (size_t)DN_BT * DN_NA_EMAX + 2 * DN_NA_EMAX)
static __device__ __align__(256) float dn_na_scratch[DN_NA_FLOATS];
// In host code `dn_na_scratch` is only a handle for the runtime, not a device address: the address on the CURRENT
// device comes from cudaGetSymbolAddress (no allocation; the module image of this device already holds the array).
static float* dn_na_resolve(int* dev_out, std::string* err) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  float* p = nullptr;
  if (e == cudaSuccess) e = cudaGetSymbolAddress((void**)&p, dn_na_scratch);
  if (e != cudaSuccess || !p) {
    cudaGetLastError();
    if (err) *err = std::string("cudaGetSymbolAddress(dn_na_scratch): ") + cudaGetErrorString(e);
    return nullptr;
  }
  if (dev_out) *dev_out = dev;
  return p;
}

struct tnt_dense {
  tnt_dense_config cfg;
  int sm_count = 0, S = 0, kslab = 0;
  float *w = nullptr, *m = nullptr, *v = nullptr, *g = nullptr, *state = nullptr;   // caller-owned
  int64_t o_W = 0, o_b = 0, o_gamma = 0, o_beta = 0, n_param = 0;
  float *xT = nullptr, *part = nullptr, *act = nullptr, *dpre = nullptr, *save = nullptr;
  Seg* d_segs = nullptr; OptChunk* d_chunks = nullptr; int n_chunks = 0;
  double *seg_gsq = nullptr, *seg_wsq = nullptr;
  int64_t adam_t = 0, launches = 0;
  int last_B = 0; uint64_t last_seed = 0;
  float* na_scratch = nullptr;       // device address of dn_na_scratch on the device this object was created on
  int na_dev = -1;
  std::string err;
};
static std::string g_dense_err;

static int dfail(tnt_dense* d, int code, const std::string& msg) {
  if (d) d->err = msg; else g_dense_err = msg;
  return code;
}
static int dcheck(tnt_dense* d, const char* where) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return dfail(d, TNT_ERR_CUDA, std::string(where) + ": " + cudaGetErrorString(e));
  return TNT_OK;
}

extern "C" const char* tnt_dense_last_error(tnt_dense* d) { return d ? d->err.c_str() : g_dense_err.c_str(); }
extern "C" int64_t tnt_dense_param_floats(tnt_dense* d) { return d ? d->n_param : 0; }
extern "C" int64_t tnt_dense_launch_count(tnt_dense* d) { return d ? d->launches : 0; }

extern "C" int tnt_dense_create(const tnt_dense_config* cfg, tnt_dense** out) {
  if (!cfg || !out) return dfail(nullptr, TNT_ERR_ARG, "tnt_dense_create: null argument");
  if (cfg->K < 1 || cfg->E < DN_NB || cfg->E % DN_NB || cfg->B_max < 1 || cfg->B_max > DN_BT)
    return dfail(nullptr, TNT_ERR_ARG, "tnt_dense_create: need K >= 1, E a multiple of 128, 1 <= B_max <= 64");
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess ||
      major < 10) {
    cudaGetLastError();
    return dfail(nullptr, TNT_ERR_NO_DEVICE, "tnt_dense_create: no sm_100 device (there is no CPU fallback)");
  }
  tnt_dense* d = new tnt_dense();
  d->cfg = *cfg;
  cudaDeviceGetAttribute(&d->sm_count, cudaDevAttrMultiProcessorCount, dev);
  const int64_t K = cfg->K, E = cfg->E;
  d->o_W = 0; d->o_b = K * E; d->o_gamma = d->o_b + E; d->o_beta = d->o_gamma + E; d->n_param = d->o_beta + E;
  // forward split: ~3 CTAs per SM in total
  const int colblk = (int)(E / DN_NB);
  int S = (3 * d->sm_count + colblk - 1) / colblk;
  int kslab = (int)((K + S - 1) / S);
  kslab = (kslab + DN_KT - 1) / DN_KT * DN_KT;
  S = (int)((K + kslab - 1) / kslab);
  d->S = S; d->kslab = kslab;
  bool ok = cudaMalloc(&d->xT, (size_t)K * 64 * sizeof(float)) == cudaSuccess &&
            cudaMalloc(&d->part, (size_t)S * DN_BT * E * sizeof(float)) == cudaSuccess &&
            cudaMalloc(&d->act, (size_t)DN_BT * E * sizeof(float)) == cudaSuccess &&
            cudaMalloc(&d->dpre, (size_t)DN_BT * E * sizeof(float)) == cudaSuccess &&
            cudaMalloc(&d->save, (size_t)2 * E * sizeof(float)) == cudaSuccess;
  // optimizer tables: one segment per Keras variable (dense_img/kernel, bias, batch_norm/gamma, beta)
  std::vector<Seg> segs(4);
  segs[0] = {d->o_W, K * E, cfg->input_reg, 0};
  segs[1] = {d->o_b, E, 0.0f, 0}; segs[2] = {d->o_gamma, E, 0.0f, 0}; segs[3] = {d->o_beta, E, 0.0f, 0};
  std::vector<OptChunk> chunks;
  for (int s = 0; s < 4; ++s)
    for (int64_t o = 0; o < segs[s].len; o += DN_OPT_CHUNK) {
      OptChunk ch; ch.seg = s; ch.off = segs[s].off + o; ch.len = (int32_t)std::min<int64_t>(DN_OPT_CHUNK, segs[s].len - o);
      chunks.push_back(ch);
    }
  d->n_chunks = (int)chunks.size();
  ok = ok && cudaMalloc(&d->d_segs, sizeof(Seg) * 4) == cudaSuccess &&
       cudaMalloc(&d->d_chunks, sizeof(OptChunk) * chunks.size()) == cudaSuccess &&
       cudaMalloc(&d->seg_gsq, sizeof(double) * 4) == cudaSuccess && cudaMalloc(&d->seg_wsq, sizeof(double) * 4) == cudaSuccess;
  if (ok) {
    cudaMemcpy(d->d_segs, segs.data(), sizeof(Seg) * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(d->d_chunks, chunks.data(), sizeof(OptChunk) * chunks.size(), cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(dense_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DN_FWD_SMEM);
    cudaFuncSetAttribute(dense_dw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DN_BWD_SMEM);
  }
  // set-up, once per object, with its device current: where this device keeps the no-alloc scratch (failure is
  // reported by tnt_dense_forward_noalloc, the only user)
  if (ok) d->na_scratch = dn_na_resolve(&d->na_dev, &d->err);
  if (!ok || cudaGetLastError() != cudaSuccess) {
    std::string msg = "tnt_dense_create: device allocation failed";
    extern int tnt_dense_destroy(tnt_dense*);
    tnt_dense_destroy(d);
    return dfail(nullptr, TNT_ERR_CUDA, msg);
  }
  *out = d;
  return TNT_OK;
}

extern "C" int tnt_dense_destroy(tnt_dense* d) {
  if (!d) return TNT_OK;
  for (void* p : {(void*)d->xT, (void*)d->part, (void*)d->act, (void*)d->dpre, (void*)d->save, (void*)d->d_segs,
                  (void*)d->d_chunks, (void*)d->seg_gsq, (void*)d->seg_wsq})
    if (p) cudaFree(p);
  delete d;
  return TNT_OK;
}

extern "C" int tnt_dense_bind_params(tnt_dense* d, float* w, float* m, float* v, float* g, float* state) {
  if (!d || !w || !state) return dfail(d, TNT_ERR_ARG, "tnt_dense_bind_params: w and state are required");
  d->w = w; d->m = m; d->v = v; d->g = g; d->state = state;
  return TNT_OK;
}
extern "C" int tnt_dense_set_adam_step(tnt_dense* d, int64_t t) { if (!d) return TNT_ERR_ARG; d->adam_t = t; return TNT_OK; }
extern "C" int tnt_dense_set_lr(tnt_dense* d, float lr) { if (!d) return TNT_ERR_ARG; d->cfg.lr = lr; return TNT_OK; }

// features (B, E) [= (B, 1, E)] = NIC.py:122-128.  training: dropout + batch statistics (moving statistics updated).
extern "C" int tnt_dense_forward(tnt_dense* d, const float* x, int B, int training, uint64_t seed, float* features,
                                 void* stream) {
  if (!d || !x || !features) return dfail(d, TNT_ERR_ARG, "tnt_dense_forward: null argument");
  if (!d->w || !d->state) return dfail(d, TNT_ERR_STATE, "tnt_dense_forward: parameters not bound");
  if (B < 1 || B > d->cfg.B_max) return dfail(d, TNT_ERR_ARG, "tnt_dense_forward: batch outside [1, B_max]");
  cudaStream_t st = (cudaStream_t)stream;
  const int K = d->cfg.K, E = d->cfg.E;
  DropSite din = make_drop(d->cfg.dropout_input, training, seed, SITE_INPUT, 0);
  DropSite dft = make_drop(d->cfg.dropout, training, seed, SITE_FEAT, 0);
  dense_transpose_kernel<<<(K + 63) / 64, 256, 0, st>>>(x, K, B, d->xT, din);
  dense_fwd_kernel<<<dim3(E / DN_NB, d->S), DN_THREADS, DN_FWD_SMEM, st>>>(d->xT, d->w + d->o_W, K, E, d->kslab, d->part);
  dense_act_kernel<<<(B * E + 255) / 256, 256, 0, st>>>(d->part, d->S, B, E, d->w + d->o_b, d->act, dft);
  dense_bn_fwd_kernel<<<(E + 127) / 128, 128, 0, st>>>(d->act, B, E, d->w + d->o_gamma, d->w + d->o_beta, d->state,
                                                      d->state + E, training, features, d->save);
  d->launches += 4;
  d->last_B = B; d->last_seed = seed;
  return dcheck(d, "tnt_dense_forward");
}

// Inference forward that allocates nothing: tnt_dense_forward(d, x, B, training = 0, ., features, stream) with the
// scratch (xT, partial tiles, act, save) in dn_na_scratch instead of the object's cudaMalloc'ed buffers.  Same kernels,
// same launch geometry, same summation order: bit-identical features.  Does not touch the object's training state
// (save / last_B / last_seed), so it may be interleaved with training steps on the same stream.
extern "C" int tnt_dense_forward_noalloc(tnt_dense* d, const float* x, int B, float* features, void* stream) {
  if (!d || !x || !features) return dfail(d, TNT_ERR_ARG, "tnt_dense_forward_noalloc: null argument");
  if (!d->w || !d->state) return dfail(d, TNT_ERR_STATE, "tnt_dense_forward_noalloc: parameters not bound");
  if (B < 1 || B > d->cfg.B_max) return dfail(d, TNT_ERR_ARG, "tnt_dense_forward_noalloc: batch outside [1, B_max]");
  const int K = d->cfg.K, E = d->cfg.E;
  if (K > DN_NA_KMAX || E > DN_NA_EMAX || (size_t)d->S * DN_BT * E > DN_NA_PART)
    return dfail(d, TNT_ERR_ARG, "tnt_dense_forward_noalloc: static scratch holds K <= 65536, E <= 512 and S * 64 * E <= "
                                 "4 Mi partial sums (K = " + std::to_string(K) + ", E = " + std::to_string(E) + ", S = " +
                                 std::to_string(d->S) + "); use tnt_dense_forward");
  if (!d->na_scratch)
    return dfail(d, TNT_ERR_STATE, "tnt_dense_forward_noalloc: scratch address was not resolved at tnt_dense_create (" +
                                   d->err + ")");
  cudaStream_t st = (cudaStream_t)stream;
  float* xT = d->na_scratch;
  float* part = xT + DN_NA_XT;
  float* act = part + DN_NA_PART;
  float* save = act + (size_t)DN_BT * DN_NA_EMAX;
  DropSite none = make_drop(0.0f, 0, 0, SITE_INPUT, 0);
  dense_transpose_kernel<<<(K + 63) / 64, 256, 0, st>>>(x, K, B, xT, none);
  dense_fwd_kernel<<<dim3(E / DN_NB, d->S), DN_THREADS, DN_FWD_SMEM, st>>>(xT, d->w + d->o_W, K, E, d->kslab, part);
  dense_act_kernel<<<(B * E + 255) / 256, 256, 0, st>>>(part, d->S, B, E, d->w + d->o_b, act, none);
  dense_bn_fwd_kernel<<<(E + 127) / 128, 128, 0, st>>>(act, B, E, d->w + d->o_gamma, d->w + d->o_beta, d->state,
                                                      d->state + E, 0, features, save);
  d->launches += 4;
  return dcheck(d, "tnt_dense_forward_noalloc");
}

// tape gradient of the encoder variables given d(loss)/d(features) (B, E) of the last training forward -> g
// (the L2 term's gradient 2*lambda*W is added by the optimizer, as in optimizer.cu)
extern "C" int tnt_dense_backward(tnt_dense* d, const float* dfeatures, void* stream) {
  if (!d || !dfeatures) return dfail(d, TNT_ERR_ARG, "tnt_dense_backward: null argument");
  if (!d->g) return dfail(d, TNT_ERR_STATE, "tnt_dense_backward: gradient buffer not bound");
  if (d->last_B < 1) return dfail(d, TNT_ERR_STATE, "tnt_dense_backward: no training forward to differentiate");
  cudaStream_t st = (cudaStream_t)stream;
  const int K = d->cfg.K, E = d->cfg.E, B = d->last_B;
  DropSite dft = make_drop(d->cfg.dropout, 1, d->last_seed, SITE_FEAT, 0);
  dense_bn_bwd_kernel<<<(E + 127) / 128, 128, 0, st>>>(dfeatures, d->act, B, E, d->w + d->o_gamma, d->save, d->dpre,
                                                      d->g + d->o_gamma, d->g + d->o_beta, d->g + d->o_b, dft);
  const int nt = (K + DN_KB - 1) / DN_KB;
  int gy = (3 * d->sm_count) / (E / DN_NB);
  if (gy > nt) gy = nt;
  if (gy < 1) gy = 1;
  dense_dw_kernel<<<dim3(E / DN_NB, gy), DN_THREADS, DN_BWD_SMEM, st>>>(d->xT, d->dpre, K, E, d->g + d->o_W);
  d->launches += 2;
  return dcheck(d, "tnt_dense_backward");
}

// Adam(clipnorm).apply_gradients over the four variables (per-variable tf.clip_by_norm; L2(input_reg) on the kernel)
extern "C" int tnt_dense_apply_gradients(tnt_dense* d, void* stream) {
  if (!d || !d->w || !d->m || !d->v || !d->g) return dfail(d, TNT_ERR_STATE, "tnt_dense_apply_gradients: w/m/v/g not bound");
  cudaStream_t st = (cudaStream_t)stream;
  const tnt_dense_config& f = d->cfg;
  cudaMemsetAsync(d->seg_gsq, 0, sizeof(double) * 4, st);
  cudaMemsetAsync(d->seg_wsq, 0, sizeof(double) * 4, st);
  opt_launch_norm(d->d_chunks, d->n_chunks, d->d_segs, d->w, d->g, d->seg_gsq, d->seg_wsq, st);
  d->adam_t += 1;
  const double t = (double)d->adam_t;
  const double alpha = (double)f.lr * sqrt(1.0 - pow((double)f.beta2, t)) / (1.0 - pow((double)f.beta1, t));
  opt_launch_adam(d->d_chunks, d->n_chunks, d->d_segs, d->w, d->m, d->v, d->g, d->seg_gsq, f.clipnorm, (float)alpha,
                  f.beta1, f.beta2, f.adam_eps, st);
  d->launches += 2;
  return dcheck(d, "tnt_dense_apply_gradients");
}
