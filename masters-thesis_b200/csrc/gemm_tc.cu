// Tensor-core GEMM for every time-parallel product of the path (LSTM input projection, dense_inter,
// dense_out and all their tape gradients: lc_NIC.py:118-159,255,261,387): fp32 in, fp32 out,
// tcgen05.mma kind::f16 on a bf16x3 operand split with fp32 accumulation in TMEM (see tc.cuh).
//   C[M,N] = opA(A) * opB(B) (+ bias[n]) (+ C);  row-major storage, same conventions as gemm.cuh:
//   TA: A stored [K,M];  TB: B stored [N,K].
// CTA tile 128 x 128, K step 64, two smem stages (2 x 64 KB: A_hi, A_lo, B_hi, B_lo of 16 KB each).
// All 8 warps stage global fp32 -> bf16 hi/lo core-matrix tiles; thread 0 issues the MMAs; completion
// comes back through mbarriers (tcgen05.commit), so the loads of step i+1 overlap the MMAs of step i.
#include "tnt_internal.h"
#include "tc.cuh"

#define TC_BM 128
#define TC_BN 128
#define TC_BK 64
#define TC_TILE_BYTES (128 * 64 * 2)          // one bf16 operand tile (hi or lo)
// NS = operand split: 2 -> bf16x3 (16 mantissa bits, 3 MMAs per k-slice), 3 -> bf16x6 (24 bits, 6 MMAs)
#define TC_STAGE_BYTES(NS) (2 * (NS) * TC_TILE_BYTES)    // A parts, then B parts
#define TC_SMEM_BYTES(NS) (2 * TC_STAGE_BYTES(NS) + 1024)

int g_tc_desc_swap = 0;   // debug: swap LBO/SBO (tools/tc_selftest)
int g_tc_split = 3;       // 3: bf16x6 (fp32-equivalent operands, TNT_FP32); 2: bf16x3 (TNT_BF16X3 fast mode)

// One operand tile: OUTER x INNER elements, `inner` contiguous in global memory (stride ld between
// outer indices).  A warp task covers 8 outer x 16 inner; lane l -> outer l%8, inner 4*(l/8)..+3.
template <int OUTER, int INNER>
struct Stager {
  static constexpr int TI = INNER / 16;                  // inner tasks
  static constexpr int TASKS = (OUTER / 8) * TI;         // 64
  static constexpr int PER_WARP = TASKS / 16;            // 4 (16 warps)

  __device__ static __forceinline__ void load(float4 (&r)[PER_WARP], const float* __restrict__ src, int64_t ld,
                                              int outer0, int outer_end, int inner0, int inner_end, bool vec,
                                              int warp, int lane) {
#pragma unroll
    for (int j = 0; j < PER_WARP; ++j) {
      const int tk = warp + 16 * j;
      const int og = tk / TI, it = tk % TI;
      const int o = outer0 + og * 8 + (lane & 7);
      const int i = inner0 + it * 16 + (lane >> 3) * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (o < outer_end && i < inner_end) {
        const float* p = src + (int64_t)o * ld + i;
        if (vec && i + 3 < inner_end) {
          v = __ldg(reinterpret_cast<const float4*>(p));
        } else {
          v.x = __ldg(p);
          if (i + 1 < inner_end) v.y = __ldg(p + 1);
          if (i + 2 < inner_end) v.z = __ldg(p + 2);
          if (i + 3 < inner_end) v.w = __ldg(p + 3);
        }
      }
      r[j] = v;
    }
  }

  template <int NS>
  __device__ static __forceinline__ void store(const float4 (&r)[PER_WARP], uint8_t* tiles, int warp, int lane) {
#pragma unroll
    for (int j = 0; j < PER_WARP; ++j) {
      const int tk = warp + 16 * j;
      const int og = tk / TI, it = tk % TI;
      const int ic = lane >> 3;
      const uint32_t off = (uint32_t)((og * (INNER / 8) + it * 2 + (ic >> 1)) * 128 + (lane & 7) * 16 + (ic & 1) * 8);
      if (NS == 2) {
        uint2 hi, lo;
        tc::split4(r[j], hi, lo);
        *reinterpret_cast<uint2*>(tiles + off) = hi;
        *reinterpret_cast<uint2*>(tiles + TC_TILE_BYTES + off) = lo;
      } else {
        uint2 hi, mid, lo;
        tc::split4x3(r[j], hi, mid, lo);
        *reinterpret_cast<uint2*>(tiles + off) = hi;
        *reinterpret_cast<uint2*>(tiles + TC_TILE_BYTES + off) = mid;
        *reinterpret_cast<uint2*>(tiles + 2 * TC_TILE_BYTES + off) = lo;
      }
    }
  }
};

// Accumulation: the tensor core adds into TMEM with truncation, a BIASED rounding whose error grows
// linearly with the number of accumulate steps (measured: 4e-7 at K=64 -> 6e-6 at K=960) and that
// biased error is what breaks parity on gradients with heavy cancellation (BatchNorm backward means).
// So (1) the leading hi*hi products go to their own accumulator ("main") and all cross terms, 2^-8 and
// smaller, to a second one ("corr") whose truncation error is correspondingly smaller, and (2) both are
// drained every TC_DRAIN k-steps into fp32 registers, where the long sum runs with round-to-nearest.
// TMEM: 2 ping-pong buffers x (main 128 + corr 128 columns) = all 512 columns.
#define TC_THREADS 512
#define TC_DRAIN 2        // stages (of TC_BK) per TMEM accumulation chain: 8 truncating adds on `main`

template <bool TA, bool TB, int NS>
__global__ void __launch_bounds__(TC_THREADS, 1)
gemm_tc_kernel(int M, int N, int K, const float* __restrict__ A, int lda, const float* __restrict__ B, int ldb,
               float* __restrict__ C, int ldc, const float* __restrict__ bias, int accumulate, int k_per_split,
               int vecA, int vecB, int vecC, int desc_swap) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t mma_bar[2];
  __shared__ __align__(8) uint64_t grp_bar[2];
  __shared__ uint32_t tmem_slot;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int m0 = blockIdx.y * TC_BM, n0 = blockIdx.x * TC_BN;
  const int kbeg = blockIdx.z * k_per_split;
  const int kend = min(K, kbeg + k_per_split);
  const int n_iter = (kend - kbeg + TC_BK - 1) / TC_BK;

  if (tid == 0) {
    tc::mbar_init(&mma_bar[0], 1);
    tc::mbar_init(&mma_bar[1], 1);
    tc::mbar_init(&grp_bar[0], 1);
    tc::mbar_init(&grp_bar[1], 1);
    tc::mbar_fence_init();
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  using StK = Stager<128, 64>;
  using StMN = Stager<64, 128>;
  float4 ra[4], rb[4];
  auto load_stage = [&](int it) {
    const int k0 = kbeg + it * TC_BK;
    if (TA) StMN::load(ra, A, lda, k0, kend, m0, M, vecA, warp, lane);
    else StK::load(ra, A, lda, m0, M, k0, kend, vecA, warp, lane);
    if (TB) StK::load(rb, B, ldb, n0, N, k0, kend, vecB, warp, lane);
    else StMN::load(rb, B, ldb, k0, kend, n0, N, vecB, warp, lane);
  };
  constexpr uint32_t idesc = tc::idesc_bf16(TC_BM, TC_BN, TA ? 1 : 0, TB ? 0 : 1);
  constexpr uint32_t A_LBO = TA ? 2048u : 128u, A_SBO = TA ? 128u : 1024u, A_KSTEP = TA ? 4096u : 256u;
  constexpr uint32_t B_LBO = TB ? 128u : 2048u, B_SBO = TB ? 1024u : 128u, B_KSTEP = TB ? 256u : 4096u;

  // this thread's 32 fp32 accumulators: row q*32+lane of the tile, columns cg*32 .. cg*32+31
  const int q = warp & 3, cg = warp >> 2;
  float acc[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) acc[i] = 0.0f;
  auto drain = [&](int g) {
    const int tb = g & 1;
    tc::mbar_wait(&grp_bar[tb], (uint32_t)((g >> 1) & 1));
    tc::tc_fence_after();
    const uint32_t ta = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(tb * 256 + cg * 32);
    float vm[32], vc[32];
    tc::tmem_ld32(ta, vm);
    tc::tmem_ld32(ta + 128, vc);
    tc::tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 32; ++i) acc[i] += vm[i] + vc[i];
    tc::tc_fence_before();
  };

  if (n_iter > 0) load_stage(0);
  for (int it = 0; it < n_iter; ++it) {
    const int s = it & 1;
    const int g = it / TC_DRAIN, tb = g & 1;
    const bool grp_first = (it % TC_DRAIN) == 0;
    const bool grp_last = (it % TC_DRAIN) == TC_DRAIN - 1 || it + 1 == n_iter;
    uint8_t* st = smem + s * TC_STAGE_BYTES(NS);
    if (it >= 2) tc::mbar_wait(&mma_bar[s], (uint32_t)(((it >> 1) - 1) & 1));   // MMAs that read stage s are done
    if (TA) StMN::template store<NS>(ra, st, warp, lane);
    else StK::template store<NS>(ra, st, warp, lane);
    if (TB) StK::template store<NS>(rb, st + NS * TC_TILE_BYTES, warp, lane);
    else StMN::template store<NS>(rb, st + NS * TC_TILE_BYTES, warp, lane);
    if (it + 1 < n_iter) load_stage(it + 1);                                    // in flight across the MMA issue
    tc::fence_async_smem();
    __syncthreads();
    if (tid == 0) {
      tc::tc_fence_after();
      const uint32_t a0 = tc::smem_u32(st), b0 = a0 + NS * TC_TILE_BYTES;
      const uint32_t al = desc_swap ? A_SBO : A_LBO, as = desc_swap ? A_LBO : A_SBO;
      const uint32_t bl = desc_swap ? B_SBO : B_LBO, bs = desc_swap ? B_LBO : B_SBO;
      const uint32_t d_main = tmem_base + (uint32_t)(tb * 256), d_corr = d_main + 128;
#pragma unroll
      for (int j = 0; j < TC_BK / 16; ++j) {
        uint64_t da[NS], db[NS];
#pragma unroll
        for (int p = 0; p < NS; ++p) {
          da[p] = tc::smem_desc(a0 + p * TC_TILE_BYTES + j * A_KSTEP, al, as);
          db[p] = tc::smem_desc(b0 + p * TC_TILE_BYTES + j * B_KSTEP, bl, bs);
        }
        const bool first = grp_first && j == 0;
        tc::mma_bf16(d_main, da[0], db[0], idesc, first ? 0u : 1u);
        bool cfirst = first;
        // cross terms of order 1 .. NS-1 (part p of A with part ord-p of B), smallest first
#pragma unroll
        for (int ord = NS - 1; ord >= 1; --ord)
#pragma unroll
          for (int p = 0; p <= ord; ++p) {
            tc::mma_bf16(d_corr, da[p], db[ord - p], idesc, cfirst ? 0u : 1u);
            cfirst = false;
          }
      }
      tc::mma_commit(&mma_bar[s]);
      if (grp_last) tc::mma_commit(&grp_bar[tb]);
    }
    // drain the previous chain while this one runs on the tensor pipe
    if (grp_first && g > 0) drain(g - 1);
  }

  // ---- epilogue: last chain, then registers -> global ----------------------------------------------
  if (n_iter > 0) {
    drain((n_iter - 1) / TC_DRAIN);
    const int gm = m0 + q * 32 + lane;
    const bool split = gridDim.z > 1 || accumulate == 2;
    if (gm < M) {
      float* crow = C + (int64_t)gm * ldc;
      const int gn0 = n0 + cg * 32;
      if (bias != nullptr && blockIdx.z == 0) {
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (gn0 + i < N) acc[i] += __ldg(&bias[gn0 + i]);
      }
      if (split) {
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (gn0 + i < N) atomicAdd(&crow[gn0 + i], acc[i]);
      } else if (vecC && gn0 + 31 < N) {
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          float4 o = make_float4(acc[i], acc[i + 1], acc[i + 2], acc[i + 3]);
          float4* p = reinterpret_cast<float4*>(&crow[gn0 + i]);
          if (accumulate) { float4 c = *p; o.x += c.x; o.y += c.y; o.z += c.z; o.w += c.w; }
          *p = o;
        }
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (gn0 + i < N) crow[gn0 + i] = accumulate ? crow[gn0 + i] + acc[i] : acc[i];
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem_base, 512);
}

static bool aligned16(const void* p, int ld) { return (((uintptr_t)p) & 15) == 0 && (ld & 3) == 0; }

// Same contract as sgemm() in gemm.cu.
void gemm_tc(tnt_ctx* c, bool ta, bool tb, int M, int N, int K, const float* A, int lda, const float* B, int ldb,
             float* C, int ldc, const float* bias, int accumulate, int splitk, cudaStream_t st) {
  if (M <= 0 || N <= 0 || K <= 0) return;
  static bool attr_done = false;
  if (!attr_done) {
#define SETATTR(TA_, TB_, NS_) \
  cudaFuncSetAttribute(gemm_tc_kernel<TA_, TB_, NS_>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES(NS_))
    SETATTR(false, false, 2); SETATTR(true, false, 2); SETATTR(false, true, 2); SETATTR(true, true, 2);
    SETATTR(false, false, 3); SETATTR(true, false, 3); SETATTR(false, true, 3); SETATTR(true, true, 3);
#undef SETATTR
    attr_done = true;
  }
  const int ns = (c && c->cfg.precision == TNT_BF16X3) ? 2 : g_tc_split;
  int splits = splitk > 1 ? splitk : 1;
  int kps = (K + splits - 1) / splits;
  kps = ((kps + TC_BK - 1) / TC_BK) * TC_BK;
  splits = (K + kps - 1) / kps;
  dim3 grid((N + TC_BN - 1) / TC_BN, (M + TC_BM - 1) / TC_BM, splits);
  const int vA = aligned16(A, lda), vB = aligned16(B, ldb), vC = aligned16(C, ldc);
#define LAUNCH(TA_, TB_)                                                                                       \
  do {                                                                                                         \
    if (ns == 2)                                                                                               \
      gemm_tc_kernel<TA_, TB_, 2><<<grid, TC_THREADS, TC_SMEM_BYTES(2), st>>>(M, N, K, A, lda, B, ldb, C, ldc, bias,  \
                                                                        accumulate, kps, vA, vB, vC,           \
                                                                        g_tc_desc_swap);                       \
    else                                                                                                       \
      gemm_tc_kernel<TA_, TB_, 3><<<grid, TC_THREADS, TC_SMEM_BYTES(3), st>>>(M, N, K, A, lda, B, ldb, C, ldc, bias,  \
                                                                        accumulate, kps, vA, vB, vC,           \
                                                                        g_tc_desc_swap);                       \
  } while (0)
  if (!ta && !tb) LAUNCH(false, false);
  else if (ta && !tb) LAUNCH(true, false);
  else if (!ta && tb) LAUNCH(false, true);
  else LAUNCH(true, true);
#undef LAUNCH
  if (c) c->launches += 1;
}
