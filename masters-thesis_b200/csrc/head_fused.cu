// Fused vocabulary projection + softmax + cross-entropy (+ its gradient): TimeDistributed(Dense(V, softmax))
// (AttemptFour/Model/lc_NIC.py:150-159,261) feeding loss_function / accuracy_calculation (lc_NIC.py:370-375,461-486;
// loss object main.py:107-110; ThinkAndTell variant att_model.py:306-321) in ONE kernel: the (T*B) x V logits are
// accumulated in tensor memory and never written to HBM (north_star item 4).
//
// Grid: (n_chunks, m_groups).  A CTA owns 128 rows (one m-tile) x 384 columns (three 128-column sub-tiles = 384 of the
// 512 TMEM columns, fp32) per pass and walks the m-tiles mt = blockIdx.y, + gridDim.y, ...  Per pass:
//   warp 16  streams the packed operand tiles (inter_d rows, dense_out kernel columns) with 1-D bulk copies (TMA)
//   warp 17  issues tcgen05.mma kind::f16 on the bf16 operand planes (x = hi + mid + lo: six products, fp32 in TMEM)
//   warps 0-15 read the accumulators back (one TMEM lane = one row; four warps per lane quadrant share the columns):
//     sweep 1: row max, first arg-max, target logit; sweep 2: e = exp(x - max) ONCE per logit, written back to tensor
//     memory, and its sum -> a 4-float record per row and warp; the n_chunks CTAs of the m-tile meet at a counter
//     (cooperative launch: all resident), every thread merges the records of its row (log-sum-exp merge) -> loss,
//     correct-flag (chunk 0 writes them);
//     sweep 3 (training): d logits = scale (p - onehot) from the stored e, emitted ONLY as the packed bf16-plane
//     operands the two gradient GEMMs consume: [row][v] K-major over v (A of d inter = d logits . W_out^T) and, through
//     8 x 8 register transposes, [v][row] K-major over rows (B of d W_out = inter^T . d logits); bias gradient = column
//     sums by warp butterflies.
// Keras renormalises the probabilities (p / sum p) before the clip; sum p differs from 1 by fp32 rounding (~1e-7) and is
// taken as 1 here -- the per-row three-pass softmax_ce_reg_kernel (head.cu) keeps the exact form for tnt_forward.
#include "tnt_internal.h"
#include "tc.cuh"
#include <climits>
#include <algorithm>

#define HF_SUB 3
#define HF_BLOCK (128 * 64 * 2)
#define HF_THREADS 576                // warps 0-15 epilogue, warp 16 TMA producer, warp 17 MMA issuer
#define HF_EPI 512
#define HF_PARTS 4                     // epilogue warps per TMEM lane quadrant: they share the 32-column chunks
#define HF_SPIN_LIMIT (1ll << 31)

struct HeadFusedParams {
  TcOperand opA;            // inter_d packed: MN = rows (T*B), K = H
  TcOperand opB;            // dense_out kernel packed: MN = V, K = H
  int M, V, B, T, m_tiles, n_chunks;
  const float* bias;        // [V]
  const int32_t* target;    // (B,T) ids
  float scale;
  int backward, sparse_masked;
  float* stats;             // [m_tiles][4 n_chunks][128][4]: one record per (chunk, column-share warp)
  unsigned* counters;       // [m_tiles], zeroed before the launch
  unsigned* abort_flag;
  float* rowloss; float* rowcorrect;
  uint8_t* outA; int64_t outA_stride; int outA_mnb, outA_kt;     // d logits as A operand: MN = rows, K = V
  uint8_t* outB; int64_t outB_stride; int outB_mnb, outB_kt;     // d logits^T as B operand: MN = V, K = rows
  float* colpart;           // [m_tiles][V] column sums of d logits over the m-tile's rows
};

template <int NS> struct HfCfg {
  static constexpr int STAGE = NS * 2 * HF_BLOCK;
  static constexpr int S = (200 * 1024) / STAGE > 6 ? 6 : (200 * 1024) / STAGE;
  static constexpr int SMEM = S * STAGE + 1024;
};

template <int NS>
__global__ void __launch_bounds__(HF_THREADS, 1)
head_fused_kernel(const __grid_constant__ HeadFusedParams p) {
  constexpr int S = HfCfg<NS>::S;
  constexpr int STAGE = HfCfg<NS>::STAGE;
  extern __shared__ uint8_t hf_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)hf_smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t full_bar[S], empty_bar[S], acc_full, acc_empty;
  __shared__ uint32_t tmem_slot;
  __shared__ float sbias[HF_SUB * 128];
  __shared__ float scol[HF_SUB * 128];
  __shared__ int s_ok;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ch = blockIdx.x;
  const int nb0 = ch * HF_SUB;
  const int n_sub = min(HF_SUB, p.opB.mn_blocks - nb0);
  const int KT = p.opA.k_tiles;
  const int n_iter = n_sub * KT;
  const int n_pass = (p.m_tiles - (int)blockIdx.y + (int)gridDim.y - 1) / (int)gridDim.y;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < S; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
    tc::mbar_init(&acc_full, 1);
    tc::mbar_init(&acc_empty, HF_EPI);
    tc::mbar_fence_init();
  }
  if (warp == 17) tc::tmem_alloc(&tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 16) {
    // ================= producer =========================================================================
    if (lane == 0) {
      int g = 0;
      for (int ps = 0; ps < n_pass; ++ps) {
        const int mt = blockIdx.y + ps * gridDim.y;
        for (int it = 0; it < n_iter; ++it, ++g) {
          const int s = g % S, sub = it / KT, kt = it % KT;
          tc::mbar_wait(&empty_bar[s], (uint32_t)(((g / S) & 1) ^ 1));
          uint8_t* stg = smem + s * STAGE;
          tc::mbar_expect_tx(&full_bar[s], NS * 2 * HF_BLOCK);
#pragma unroll
          for (int q = 0; q < NS; ++q) {
            tc::bulk_g2s(stg + q * HF_BLOCK,
                         p.opA.planes + (size_t)q * p.opA.plane_stride + ((size_t)kt * p.opA.mn_blocks + mt) * HF_BLOCK,
                         HF_BLOCK, &full_bar[s]);
            tc::bulk_g2s(stg + (NS + q) * HF_BLOCK,
                         p.opB.planes + (size_t)q * p.opB.plane_stride + ((size_t)kt * p.opB.mn_blocks + nb0 + sub) * HF_BLOCK,
                         HF_BLOCK, &full_bar[s]);
          }
        }
      }
    }
  } else if (warp == 17) {
    // ================= MMA issuer =======================================================================
    if (lane == 0) {
      constexpr uint32_t idesc = tc::idesc_bf16(128, 128, 0, 0);
      const uint32_t lbo = 128u, sbo = 1024u;
      int g = 0;
      for (int ps = 0; ps < n_pass; ++ps) {
        if (ps > 0) { tc::mbar_wait(&acc_empty, (uint32_t)((ps - 1) & 1)); tc::tc_fence_after(); }   // epilogue drained TMEM
        for (int it = 0; it < n_iter; ++it, ++g) {
          const int s = g % S, sub = it / KT, kt = it % KT;
          tc::mbar_wait(&full_bar[s], (uint32_t)((g / S) & 1));
          tc::tc_fence_after();
          const uint32_t a0 = tc::smem_u32(smem + s * STAGE), b0 = a0 + NS * HF_BLOCK;
          const uint32_t d = tmem_base + (uint32_t)(sub * 128);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint64_t da[NS], db[NS];
#pragma unroll
            for (int q = 0; q < NS; ++q) {
              da[q] = tc::smem_desc(a0 + q * HF_BLOCK + j * 256u, lbo, sbo);
              db[q] = tc::smem_desc(b0 + q * HF_BLOCK + j * 256u, lbo, sbo);
            }
            bool first = (kt == 0 && j == 0);
            // smallest cross terms first, the leading hi*hi product last
#pragma unroll
            for (int ord = NS - 1; ord >= 0; --ord)
#pragma unroll
              for (int q = 0; q <= ord; ++q) {
                tc::mma_bf16(d, da[q], db[ord - q], idesc, first ? 0u : 1u);
                first = false;
              }
          }
          tc::mma_commit(&empty_bar[s]);
        }
        tc::mma_commit(&acc_full);
      }
    }
  } else {
    // ================= epilogue: 16 warps; warp w reads TMEM lanes 32 (w & 3) .. +31 (one lane = one row) and the
    // 32-column chunks sc with sc % 4 == w >> 2 (four warps per scheduler: the sweeps are latency bound) =========
    const int quad = warp & 3, half = warp >> 2;
    const int row = quad * 32 + lane;
    const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16);
    const int n_rec = p.n_chunks * HF_PARTS;
    for (int i = tid; i < HF_SUB * 128; i += HF_EPI) {
      const int v = nb0 * 128 + i;
      sbias[i] = v < p.V ? __ldg(p.bias + v) : 0.0f;
    }
    asm volatile("bar.sync 1, 512;" ::: "memory");
    for (int ps = 0; ps < n_pass; ++ps) {
      const int mt = blockIdx.y + ps * gridDim.y;
      const int gm = mt * 128 + row;
      const bool valid = gm < p.M;
      int y = -1;
      if (valid) {
        const int t = gm / p.B, b = gm % p.B;
        y = p.target[(size_t)b * p.T + t];
        y = y < 0 ? 0 : (y >= p.V ? p.V - 1 : y);
      }
      for (int i = tid; i < HF_SUB * 128; i += HF_EPI) scol[i] = 0.0f;
      tc::mbar_wait(&acc_full, (uint32_t)(ps & 1));
      tc::tc_fence_after();
      // ---- sweep 1: max, first arg-max, target logit over the warp's chunks -------------------------------
      float mx = -INFINITY, zy = -INFINITY;
      int am = INT_MAX;
      float v32[32];
#pragma unroll 1
      for (int sc = half; sc < n_sub * 4; sc += HF_PARTS) {
        const int col0 = sc * 32, gv0 = nb0 * 128 + col0;
        tc::tmem_ld32(taddr + col0, v32);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const int gv = gv0 + i;
          if (gv < p.V) {
            const float x = v32[i] + sbias[col0 + i];
            if (x > mx) { mx = x; am = gv; }
            if (gv == y) zy = x;
          }
        }
      }
      // ---- sweep 2: e = exp(x - max) once per logit, kept in tensor memory; running sum ------------------
      float se = 0.0f;
#pragma unroll 1
      for (int sc = half; sc < n_sub * 4; sc += HF_PARTS) {
        const int col0 = sc * 32, gv0 = nb0 * 128 + col0;
        tc::tmem_ld32(taddr + col0, v32);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float e = (gv0 + i < p.V) ? __expf(v32[i] + sbias[col0 + i] - mx) : 0.0f;
          v32[i] = e;
          se += e;
        }
        tc::tmem_st32(taddr + col0, v32);
      }
      tc::tmem_st_wait();
      {
        float* st = p.stats + (((size_t)mt * n_rec + ch * HF_PARTS + half) * 128 + row) * 4;
        __stcg(st, mx); __stcg(st + 1, se); __stcg(st + 2, __int_as_float(am)); __stcg(st + 3, zy);
      }
      // ---- the chunks of this m-tile meet -----------------------------------------------------------------
      __threadfence();
      asm volatile("bar.sync 1, 512;" ::: "memory");
      if (tid == 0) {
        tc::red_release(p.counters + mt, 1u);
        int ok = 1;
        const long long t0 = clock64();
        while (tc::ld_acquire(p.counters + mt) < (unsigned)p.n_chunks) {
          if (clock64() - t0 > HF_SPIN_LIMIT || tc::ld_acquire(p.abort_flag) != 0u) { atomicExch(p.abort_flag, 1u); ok = 0; break; }
        }
        s_ok = ok;
      }
      asm volatile("bar.sync 1, 512;" ::: "memory");
      const bool alive = s_ok != 0;          // a timed-out meeting: skip the work, keep the barrier protocol going
      if (alive) {
        float Mx = -INFINITY;
        for (int c2 = 0; c2 < n_rec; ++c2)
          Mx = fmaxf(Mx, __ldcg(p.stats + (((size_t)mt * n_rec + c2) * 128 + row) * 4));
        float SUM = 0.0f, ZY = -INFINITY, bestm = -INFINITY;
        int ARG = INT_MAX;
        for (int c2 = 0; c2 < n_rec; ++c2) {
          const float4 r = __ldcg(reinterpret_cast<const float4*>(p.stats + (((size_t)mt * n_rec + c2) * 128 + row) * 4));
          if (r.y > 0.0f) SUM += r.y * __expf(r.x - Mx);
          // first maximal index: records interleave the column ranges, so equal maxima are broken by the index itself
          const int a2 = __float_as_int(r.z);
          if (r.x > bestm || (r.x == bestm && a2 < ARG)) { bestm = r.x; ARG = a2; }
          if (r.w > ZY) ZY = r.w;
        }
        const float py = __expf(ZY - Mx) / SUM;
        const float qc = fminf(fmaxf(py, TNT_CE_EPS), 1.0f - TNT_CE_EPS);
        const bool inside = (py >= TNT_CE_EPS) && (py <= 1.0f - TNT_CE_EPS);   // clip_by_value passes gradient inside only
        const float keep = (y != 0) ? 1.0f : 0.0f;
        if (ch == 0 && half == 0 && valid) {
          p.rowloss[gm] = p.sparse_masked ? keep * (logf(SUM) - (ZY - Mx)) : -logf(qc);
          p.rowcorrect[gm] = (ARG == y) ? 1.0f : 0.0f;
        }
        // ---- sweep 3: d logits = scale (p - onehot), p = e * exp(max_local - max) / sum, in packed operand form ----
        if (p.backward) {
          const float rs = !valid ? 0.0f : (p.sparse_masked ? keep * p.scale : (inside ? p.scale : 0.0f));
          const float fac = rs * __expf(mx - Mx) / SUM;
          const int ktB = gm >> 6;                                   // k-tile of this row in the transposed operand
          const int kgB = (row & 63) >> 3;
#pragma unroll 1
          for (int sc = half; sc < n_sub * 4; sc += HF_PARTS) {
            const int sub = sc >> 2, c32 = sc & 3;
            const int col0 = sc * 32, gv0 = nb0 * 128 + col0;
            tc::tmem_ld32(taddr + col0, v32);
            tc::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const int gv = gv0 + i;
              float d = valid ? v32[i] * fac : 0.0f;                 // e is 0 beyond V
              if (gv == y) d -= rs;
              v32[i] = d;
            }
            // A operand: [row][v], 8 consecutive v per 16-byte chunk
            const int ktA = gv0 >> 6, kg0 = (gv0 & 63) >> 3;
            if (ktA < p.outA_kt) {
              uint8_t* blk = p.outA + ((size_t)ktA * p.outA_mnb + mt) * HF_BLOCK;
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                uint4 pl[NS];
                tc::split8<NS>(v32 + 8 * g, pl);
                const size_t off = (size_t)((((row >> 3) * 8 + kg0 + g) * 8 + (row & 7)) * 16);
#pragma unroll
                for (int q = 0; q < NS; ++q) *reinterpret_cast<uint4*>(blk + (size_t)q * p.outA_stride + off) = pl[q];
              }
            }
            // B operand: [v][row], 8 consecutive rows per chunk: 8 x 8 transposes inside each lane octet
            if (ktB < p.outB_kt) {
              uint8_t* blk = p.outB + ((size_t)ktB * p.outB_mnb + nb0 + sub) * HF_BLOCK;
#pragma unroll
              for (int g = 0; g < 4; ++g) {
                float a[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) a[i] = v32[8 * g + i];
#pragma unroll
                for (int d = 4; d >= 1; d >>= 1)
#pragma unroll
                  for (int i = 0; i < 8; ++i)
                    if ((i & d) == 0) {
                      const float send = (lane & d) ? a[i] : a[i + d];
                      const float recv = __shfl_xor_sync(0xffffffffu, send, d);
                      if (lane & d) a[i] = recv; else a[i + d] = recv;
                    }
                // lane j of the octet now holds column (8 g + j) of the octet's 8 rows
                uint4 pl[NS];
                tc::split8<NS>(a, pl);
                const int mg = c32 * 4 + g;
                const size_t off = (size_t)(((mg * 8 + kgB) * 8 + (lane & 7)) * 16);
#pragma unroll
                for (int q = 0; q < NS; ++q) *reinterpret_cast<uint4*>(blk + (size_t)q * p.outB_stride + off) = pl[q];
              }
            }
            // bias gradient: column sums over the warp's 32 rows by a butterfly; lane l ends with column col0 + l
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1)
#pragma unroll
              for (int i = 0; i < off; ++i) {
                const float send = (lane & off) ? v32[i] : v32[i + off];
                const float keepv = (lane & off) ? v32[i + off] : v32[i];
                v32[i] = keepv + __shfl_xor_sync(0xffffffffu, send, off);
              }
            atomicAdd(&scol[col0 + lane], v32[0]);
          }
        }
      }   // alive
      tc::tc_fence_before();
      tc::mbar_arrive(&acc_empty);                                 // TMEM may be overwritten by the next pass
      if (p.backward) {
        asm volatile("bar.sync 1, 512;" ::: "memory");
        for (int i = tid; alive && i < n_sub * 128; i += HF_EPI) {
          const int v = nb0 * 128 + i;
          if (v < p.V) p.colpart[(size_t)mt * p.V + v] = scol[i];
        }
        asm volatile("bar.sync 1, 512;" ::: "memory");
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 17) tc::tmem_dealloc(tmem_base, 512);
}

// logits never materialised: inter_d (after head_forward(..., want_logits = 0)) -> loss, correct flags, and with
// `backward` the packed d logits operands (cached under the key of the "logits" workspace) + "dob_part" column sums.
// false: the context is marked failed.
bool head_fused_loss(tnt_ctx* c, const int32_t* target, int B, int B_global, int backward, cudaStream_t st) {
  const tnt_config& f = c->cfg;
  const int TB = f.T * B, V = f.V, H = f.H;
  const int ns = tc_planes(c);
  HeadFusedParams p;
  p.opA = tc_pack_cached(c, wsT<float>(c, "inter_d"), TB, H, H, false, st);
  p.opB = tc_pack_cached(c, c->w + c->o_oW, H, V, V, true, st);
  if (!p.opA.planes || !p.opB.planes) { ctx_fail(c, TNT_ERR_CUDA, "head_fused_loss: operand arena exhausted"); return false; }
  p.M = TB; p.V = V; p.B = B; p.T = f.T;
  p.m_tiles = (TB + 127) / 128;
  p.n_chunks = (p.opB.mn_blocks + HF_SUB - 1) / HF_SUB;
  p.bias = c->w + c->o_ob; p.target = target;
  const int sparse_masked = (f.variant & TNT_VARIANT_MASKED_SPARSE_CE) ? 1 : 0;
  p.scale = 1.0f / ((float)B_global * (float)(sparse_masked ? f.T + 1 : f.T));
  p.backward = backward; p.sparse_masked = sparse_masked;
  p.stats = wsT<float>(c, "hf_stats"); p.counters = wsT<unsigned>(c, "hf_counters"); p.abort_flag = pd_abort_flag(c);
  p.rowloss = wsT<float>(c, "rowloss"); p.rowcorrect = wsT<float>(c, "rowcorrect");
  p.outA = nullptr; p.outB = nullptr; p.outA_stride = p.outB_stride = 0; p.outA_mnb = p.outB_mnb = p.outA_kt = p.outB_kt = 0;
  p.colpart = wsT<float>(c, "dob_part");
  if (backward) {
    const float* key = wsT<float>(c, "logits");
    TcOperand oa = tc_reserve_operand(c, key, TB, V, V, false);
    TcOperand ob = tc_reserve_operand(c, key, TB, V, V, true);
    if (!oa.planes || !ob.planes) { ctx_fail(c, TNT_ERR_CUDA, "head_fused_loss: operand arena exhausted"); return false; }
    p.outA = const_cast<uint8_t*>(oa.planes); p.outA_stride = oa.plane_stride; p.outA_mnb = oa.mn_blocks; p.outA_kt = oa.k_tiles;
    p.outB = const_cast<uint8_t*>(ob.planes); p.outB_stride = ob.plane_stride; p.outB_mnb = ob.mn_blocks; p.outB_kt = ob.k_tiles;
    // rows of the last k-tile of the transposed operand that no m-tile covers do not exist: k_tiles = ceil(TB / 64) and
    // m-tiles cover ceil(TB / 128) * 128 >= that; columns v >= V inside the last blocks are written as zeros
  }
  cudaMemsetAsync(p.counters, 0, sizeof(unsigned) * p.m_tiles, st);
  const int groups = std::max(1, std::min(p.m_tiles, c->sm_count / p.n_chunks));
  dim3 grid(p.n_chunks, groups);
  void* args[] = {&p};
  cudaError_t e;
#define HF_LAUNCH(NS_)                                                                                                  \
  do {                                                                                                                  \
    if (!(c->hf_attr & (1 << NS_))) {                                                                                   \
      cudaFuncSetAttribute(head_fused_kernel<NS_>, cudaFuncAttributeMaxDynamicSharedMemorySize, HfCfg<NS_>::SMEM);      \
      c->hf_attr |= (1 << NS_);                                                                                         \
    }                                                                                                                   \
    e = cudaLaunchCooperativeKernel((const void*)head_fused_kernel<NS_>, grid, dim3(HF_THREADS), args, HfCfg<NS_>::SMEM, st); \
  } while (0)
  if (ns == 1) HF_LAUNCH(1); else if (ns == 2) HF_LAUNCH(2); else HF_LAUNCH(3);
#undef HF_LAUNCH
  if (e != cudaSuccess) {
    cudaGetLastError();
    ctx_fail(c, TNT_ERR_CUDA, std::string("head_fused_loss: cooperative launch failed: ") + cudaGetErrorString(e));
    return false;
  }
  c->launches += 1;
  c->variant_head = "head_fused_kernel (dense_out + softmax + CE + d logits in TMEM, logits never in HBM)";
  return true;
}
