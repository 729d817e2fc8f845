// Parcel encoder on the Blackwell tensor path (north_star item 1): the tape gradient of LocallyDense
// (AttemptFour/Model/layers.py:43-53)  dW_r[k][g] = sum_b x[b, idx_r[k]] * dYpre[b][r][g]  as tcgen05.mma tiles with the
// accumulator in tensor memory, fed by asynchronous copies that complete on mbarriers.
//
// Per tile of 128 gathered vertices (M = 128), N = 32 channels, K = 64 batch rows:
//   warps 1-15  issue the row gathers two tiles ahead: asynchronous 16-byte copies (16 threads per 256-byte vertex row of
//               the vertex-major batch copy xT[N][64]) into a 3-slot staging ring, completion signalled on the slot's
//               mbarrier by the copy unit; then convert a landed tile from fp32 to three bf16 planes (x = hi + mid + lo, 24 mantissa bits) in the
//               tensor core's K-major core-matrix layout (A[m = vertex][k = batch row]: 8 batch values = one 16-byte
//               chunk, read as 32 contiguous bytes of the staged row), likewise dYpre of the parcel once per parcel
//               (B[n = channel][k = batch row])
//   warp 0      issues the 24 MMAs of the tile (4 k-steps x 6 plane products, kind::f16, fp32 accumulate in TMEM);
//               tcgen05.commit hands the operand buffer back and the accumulator on
//   warps 12-15 read the accumulator (one TMEM lane = one vertex = one 128-byte row of dW) and store it
// Persistent: one CTA per SM walks the work list (largest items first); operands and accumulators are double buffered so
// gather, conversion, MMA and the store of consecutive tiles overlap.
#include "tnt_internal.h"
#include "tc.cuh"

#define ET_THREADS 512
#define ET_TILE 128
#define ET_ROWF 68                       // staged row: 64 floats + 4 of padding (conflict-free 16-byte reads down a column)
#define ET_STAGES 3
#define ET_NS 3
#define ET_A_BYTES (128 * 64 * 2)        // one plane of the A operand tile
#define ET_B_BYTES (32 * 64 * 2)
#define ET_NW 480                        // worker threads (warps 1-15)
#define ET_MAXT 256                      // tiles one CTA may own
#define ET_MAXI 64                       // work items one CTA may own
#define ET_STAGE_BYTES (ET_TILE * ET_ROWF * 4)
struct EtTile { int64_t idx_off; int64_t dw_row; int32_t nv; int32_t r; };
#define ET_SMEM (ET_STAGES * ET_STAGE_BYTES + 2 * ET_NS * ET_A_BYTES + ET_NS * ET_B_BYTES + ET_MAXT * (int)sizeof(EtTile) + 1024)

__global__ void __launch_bounds__(ET_THREADS, 1)
enc_bwd_dw_tc_kernel(const float* __restrict__ xT, int b_begin, int Bh, int R, const int32_t* __restrict__ idx,
                     const int64_t* __restrict__ off, const EncWork* __restrict__ work, int n_work,
                     const float* __restrict__ dYpre, float* __restrict__ dW) {
  extern __shared__ uint8_t et_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)et_smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* stage_base = smem;                                             // [ET_STAGES][128][ET_ROWF] fp32
  uint8_t* a_base = smem + ET_STAGES * ET_STAGE_BYTES;                    // [2][NS][16 KB]
  uint8_t* b_base = a_base + 2 * ET_NS * ET_A_BYTES;                      // [NS][4 KB]
  EtTile* tiles = reinterpret_cast<EtTile*>(b_base + ET_NS * ET_B_BYTES);
  __shared__ __align__(8) uint64_t stage_full[ET_STAGES], stage_empty[ET_STAGES], a_full[2], a_empty[2], acc_full[2], acc_empty[2];
  __shared__ uint32_t tmem_slot;
  __shared__ int s_cnt[ET_MAXI + 1];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < ET_STAGES; ++s) { tc::mbar_init(&stage_full[s], ET_NW); tc::mbar_init(&stage_empty[s], ET_NW); }
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      tc::mbar_init(&a_full[b], ET_NW); tc::mbar_init(&a_empty[b], 1);
      tc::mbar_init(&acc_full[b], 1); tc::mbar_init(&acc_empty[b], 128);
    }
    tc::mbar_fence_init();
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 64);
  // ---- the CTA's tile list: work items blockIdx.x, + gridDim.x, ... cut into 128-vertex tiles ---------------------
  const int n_items = (n_work - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  EncWork my_wk; my_wk.r = 0; my_wk.kbeg = 0; my_wk.kend = 0; my_wk.slot = 0;
  int64_t my_base = 0;
  if (tid < n_items) {
    my_wk = work[blockIdx.x + tid * gridDim.x];
    my_base = off[my_wk.r];
    s_cnt[tid] = (my_wk.kend - my_wk.kbeg + ET_TILE - 1) / ET_TILE;
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  if (tid < n_items) {
    int o = 0;
    for (int i = 0; i < tid; ++i) o += s_cnt[i];
    for (int k0 = my_wk.kbeg; k0 < my_wk.kend; k0 += ET_TILE, ++o)
      if (o < ET_MAXT) {
        EtTile t; t.idx_off = my_base + k0; t.dw_row = my_base + k0; t.nv = min(ET_TILE, my_wk.kend - k0); t.r = my_wk.r;
        tiles[o] = t;
      }
    if (tid == n_items - 1) s_cnt[ET_MAXI] = min(o, ET_MAXT);
  }
  __syncthreads();
  const int n_tiles = n_items > 0 ? s_cnt[ET_MAXI] : 0;
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    // ================= MMA issuer =========================================================================
    if (lane == 0) {
      constexpr uint32_t idesc = tc::idesc_bf16(128, 32, 0, 0);
      const uint32_t lbo = 128u, sbo = 1024u;
      for (int j = 0; j < n_tiles; ++j) {
        const int buf = j & 1, u = j >> 1;
        tc::mbar_wait(&a_full[buf], (uint32_t)(u & 1));
        if (u >= 1) tc::mbar_wait(&acc_empty[buf], (uint32_t)((u - 1) & 1));
        tc::tc_fence_after();
        const uint32_t a0 = tc::smem_u32(a_base + (size_t)buf * ET_NS * ET_A_BYTES), b0 = tc::smem_u32(b_base);
        const uint32_t d = tmem_base + (uint32_t)(buf * 32);
        bool first = true;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          uint64_t da[ET_NS], db[ET_NS];
#pragma unroll
          for (int q = 0; q < ET_NS; ++q) {
            da[q] = tc::smem_desc(a0 + q * ET_A_BYTES + ks * 256u, lbo, sbo);
            db[q] = tc::smem_desc(b0 + q * ET_B_BYTES + ks * 256u, lbo, sbo);
          }
#pragma unroll
          for (int ord = ET_NS - 1; ord >= 0; --ord)
#pragma unroll
            for (int q = 0; q <= ord; ++q) {
              tc::mma_bf16(d, da[q], db[ord - q], idesc, first ? 0u : 1u);
              first = false;
            }
        }
        tc::mma_commit(&a_empty[buf]);
        tc::mma_commit(&acc_full[buf]);
      }
    }
  } else {
    // ================= workers (warps 1-15): gather issue + conversion; warps 12-15 also drain the accumulators ====
    const int wt = tid - 32;
    const int quad = warp & 3;                                            // TMEM lane quadrant of warps 12-15
    // gathers of tile jt: 16 threads x 16 bytes cover one 256-byte vertex row of xT; asynchronous (LDGSTS), the stage's
    // mbarrier trips when every worker's copies have landed (cp.async.mbarrier.arrive.noinc, ET_NW expected arrivals).
    // Every worker issues its share: a single producer warp could not issue a tile's 2048 copies fast enough, and one
    // 256-byte cp.async.bulk (TMA) per row costs ~85 cycles inside the TMA unit (profiles/r2f: 89.6 us for the pass).
    auto issue = [&](int jt) {
      if (jt >= n_tiles) return;
      const int s = jt % ET_STAGES;
      if (jt >= ET_STAGES) tc::mbar_wait(&stage_empty[s], (uint32_t)((jt / ET_STAGES - 1) & 1));
      const EtTile t = tiles[jt];
      uint8_t* dst = stage_base + (size_t)s * ET_STAGE_BYTES;
#pragma unroll
      for (int i = 0; i < (2048 + ET_NW - 1) / ET_NW; ++i) {
        const int c = wt + ET_NW * i;
        const int row = c >> 4, piece = c & 15;
        if (c < 2048 && row < t.nv) {
          const int32_t v = __ldg(&idx[t.idx_off + row]);
          const uint32_t d = tc::smem_u32(dst + (size_t)row * ET_ROWF * 4 + piece * 16);
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(xT + (size_t)v * 64 + piece * 4) : "memory");
        }
      }
      asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(tc::smem_u32(&stage_full[s])) : "memory");
    };
    auto drain = [&](int jd) {
      const int buf = jd & 1, u = jd >> 1;
      const EtTile t = tiles[jd];
      tc::mbar_wait(&acc_full[buf], (uint32_t)(u & 1));
      tc::tc_fence_after();
      float v[32];
      tc::tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(buf * 32), v);
      tc::tmem_ld_wait();
      tc::tc_fence_before();
      tc::mbar_arrive(&acc_empty[buf]);
      const int k = quad * 32 + lane;
      if (k < t.nv) {
        float4* dst = reinterpret_cast<float4*>(dW + (size_t)(t.dw_row + k) * 32);
#pragma unroll
        for (int i = 0; i < 8; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
      }
    };
    issue(0);
    issue(1);
    int last_r = -1;
    for (int j = 0; j < n_tiles; ++j) {
      const int s = j % ET_STAGES, buf = j & 1, u = j >> 1;
      issue(j + 2);
      const EtTile t = tiles[j];
      if (t.r != last_r) {
        // B operand of the parcel: dYpre[b][r][:] as [n = g][k = b]; the MMAs of the previous tile still read the old one
        if (j >= 1) tc::mbar_wait(&a_empty[(j - 1) & 1], (uint32_t)(((j - 1) >> 1) & 1));
        for (int c = wt; c < 256; c += ET_NW) {
          const int g = c & 31, bg = c >> 5;
          float x[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int b = bg * 8 + i;
            x[i] = b < Bh ? __ldg(dYpre + ((size_t)(b_begin + b) * R + t.r) * 32 + g) : 0.0f;
          }
          uint4 pl[ET_NS];
          tc::split8<ET_NS>(x, pl);
          const size_t o = (size_t)((g >> 3) * 1024 + bg * 128 + (g & 7) * 16);
#pragma unroll
          for (int q = 0; q < ET_NS; ++q) *reinterpret_cast<uint4*>(b_base + q * ET_B_BYTES + o) = pl[q];
        }
        last_r = t.r;
      }
      tc::mbar_wait(&stage_full[s], (uint32_t)((j / ET_STAGES) & 1));
      if (u >= 1) tc::mbar_wait(&a_empty[buf], (uint32_t)((u - 1) & 1));
      const float* st = reinterpret_cast<const float*>(stage_base + (size_t)s * ET_STAGE_BYTES);
      uint8_t* ab = a_base + (size_t)buf * ET_NS * ET_A_BYTES;
      for (int c = wt; c < 1024; c += ET_NW) {
        const int vl = c & 7, bg = (c >> 3) & 7, vg = c >> 6;
        const float* src = st + (size_t)(vg * 8 + vl) * ET_ROWF + bg * 8;
        const float4 lo4 = *reinterpret_cast<const float4*>(src), hi4 = *reinterpret_cast<const float4*>(src + 4);
        float x[8] = {lo4.x, lo4.y, lo4.z, lo4.w, hi4.x, hi4.y, hi4.z, hi4.w};
        if (Bh < 64) {
#pragma unroll
          for (int i = 0; i < 8; ++i) if (bg * 8 + i >= Bh) x[i] = 0.0f;   // rows of xT beyond the batch are stale
        }
        uint4 pl[ET_NS];
        tc::split8<ET_NS>(x, pl);
        const size_t o = (size_t)(vg * 1024 + bg * 128 + vl * 16);
#pragma unroll
        for (int q = 0; q < ET_NS; ++q) *reinterpret_cast<uint4*>(ab + q * ET_A_BYTES + o) = pl[q];
      }
      tc::fence_async_smem();                                           // generic-proxy writes -> visible to the MMA reads
      tc::mbar_arrive(&a_full[buf]);
      tc::mbar_arrive(&stage_empty[s]);
      if (warp >= 12 && j >= 1) drain(j - 1);
    }
    if (warp >= 12 && n_tiles >= 1) drain(n_tiles - 1);
    asm volatile("cp.async.wait_all;" ::: "memory");
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem_base, 64);
}

// dW of encoder e from the vertex-major batch copy (enc_xT); false: the context is marked failed
bool enc_bwd_dw_tc(tnt_ctx* c, int e, const float* xT, int b_begin, int Bh, cudaStream_t st) {
  EncTables& t = c->enc[e];
  if (!c->et_attr) {
    if (cudaFuncSetAttribute(enc_bwd_dw_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ET_SMEM) != cudaSuccess) {
      cudaGetLastError();
      ctx_fail(c, TNT_ERR_CUDA, "enc_bwd_dw_tc: cannot reserve shared memory");
      return false;
    }
    c->et_attr = 1;
  }
  int grid = t.n_work < c->sm_count ? t.n_work : c->sm_count;
  if (t.n_work > grid * 64) {      // a CTA owns <= 64 work items (<= 256 tiles of 128 vertices: items hold <= 512 vertices)
    ctx_fail(c, TNT_ERR_ARG, "enc_bwd_dw_tc: more than 64 work items per SM (raise ET_MAXI / ET_MAXT)");
    return false;
  }
  enc_bwd_dw_tc_kernel<<<grid, ET_THREADS, ET_SMEM, st>>>(xT, b_begin, Bh, c->cfg.R, t.idx, t.off, t.work, t.n_work,
                                                          wsT<float>(c, "dYpre"), c->g + t.w_off);
  c->launches += 1;
  return true;
}
