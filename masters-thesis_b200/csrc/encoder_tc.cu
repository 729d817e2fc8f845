// Parcel encoder on the Blackwell tensor path (north_star item 1): LocallyDense (AttemptFour/Model/layers.py:43-53)
//   forward   partial_r[b][g] = sum_k x[b, idx_r[k]] * W_r[k][g]                (layers.py:45, before bias / LeakyReLU)
//   gradient  dW_r[k][g]      = sum_b x[b, idx_r[k]] * dYpre[b][r][g]           (tape gradient of the same line)
// as tcgen05.mma tiles with the accumulator in tensor memory, fed by the TMA engine.
//
// Both kernels stream the parcel-ordered copy of the batch xP[n_lab][64] (encoder.cu: enc_permute_kernel; row = one
// position of the concatenated parcel index list, 64 batch values = 256 bytes) and the packed kernel W (n_lab, 32) in
// the same order: a tile of positions is a contiguous range of rows -- no gather, no LSU instruction and no register on
// the load path.  (Gathering the rows of a vertex-major copy was measured first: one 256-byte bulk copy per row costs
// ~85 cycles inside the TMA unit, 89.6 us per pass; 16-byte LDGSTS gathers 52.8 us; the old mma.sync kernel 40.4 us.)
//
// gradient: the contraction runs over the batch, so a staged tile [position][batch] IS the K-major A operand.  A
//   tensor-map TMA lands it with the 128-byte swizzle the tensor core expects and kind::tf32 MMAs read the fp32 words
//   in place.  Three tf32 products (hi*hi + lo*hi + hi*lo, hi = nearest tf32, lo = nearest tf32 of the exact remainder)
//   keep ~2^-23 per product; the only per-element work left is that rounding, element-wise at the swizzled offsets.
// forward: the contraction runs over positions, the slow index of both xP and W, and kind::tf32 does not honour
//   MN-major operands (tools/tf32_probe.cu, profiles/r2v_tf32_probe.txt) -- so the tile is transposed on its way into
//   bf16 x 3 operand planes (x = hi + mid + lo, 24 mantissa bits, six products): 4 x 8 register blocks, 128-bit shared
//   loads and stores, padded plane pitch so that both directions are bank-conflict free.  Two converter groups work on
//   alternate tiles into two operand buffers, so conversion, MMA and accumulator drain of neighbouring tiles overlap.
// Persistent: one CTA per SM walks the work list (items of <= 512 positions, largest first), cut into tiles.
#include <cuda.h>
#include "tnt_internal.h"
#include "tc.cuh"

#define ET_NS 3
#define ET_MAXT 512                      // tiles one CTA may own
#define ET_MAXI 64                       // work items one CTA may own
struct EtTile { int32_t pos0; int16_t nv; uint8_t first, last; int32_t r; int32_t slot; };   // 16 bytes

// the CTA's tile list: work items blockIdx.x, + gridDim.x, ... cut into tiles of TILE positions
template <int TILE, int MAXT>
__device__ __forceinline__ int et_build_tiles(EtTile* tiles, int* s_cnt, const EncWork* __restrict__ work, int n_work,
                                              const int64_t* __restrict__ off, const int32_t* __restrict__ slot_first) {
  const int tid = threadIdx.x;
  const int n_items = (n_work - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  EncWork wk; wk.r = 0; wk.kbeg = 0; wk.kend = 0; wk.slot = 0;
  int64_t base = 0;
  int slot = 0;
  if (tid < n_items) {
    wk = work[blockIdx.x + tid * gridDim.x];
    base = off[wk.r];
    slot = slot_first ? slot_first[wk.r] + wk.slot : 0;
    s_cnt[tid] = (wk.kend - wk.kbeg + TILE - 1) / TILE;
  }
  __syncthreads();
  if (tid < n_items) {
    int o = 0;
    for (int i = 0; i < tid; ++i) o += s_cnt[i];
    for (int k0 = wk.kbeg; k0 < wk.kend; k0 += TILE, ++o)
      if (o < MAXT) {
        EtTile t; t.pos0 = (int32_t)(base + k0); t.nv = (int16_t)min(TILE, wk.kend - k0); t.r = wk.r; t.slot = slot;
        t.first = k0 == wk.kbeg; t.last = k0 + TILE >= wk.kend;
        tiles[o] = t;
      }
    if (tid == n_items - 1) s_cnt[ET_MAXI] = min(o, MAXT);
  }
  __syncthreads();
  return n_items > 0 ? s_cnt[ET_MAXI] : 0;
}

// A group of converter threads waits for an mbarrier phase through ONE poller and a hardware named barrier: a dozen
// warps spinning on try_wait starve the single MMA-issuing thread of issue slots.
// diagnostic time line of CTA 0 (tools/enc_trace.py): eight clock64() stamps per tile; trace == nullptr in production
#define ET_TR(slot) do { if (trace && blockIdx.x == 0) trace[(size_t)j * 8 + (slot)] = clock64(); } while (0)

template <int ID, int N>
__device__ __forceinline__ void et_group_sync() { asm volatile("bar.sync %0, %1;" ::"n"(ID), "n"(N) : "memory"); }

// =====================================================================================================================
// gradient: per tile M = 128 positions, N = 32 channels, K = 64 batch rows -- kind::tf32 on the staged tile itself
// =====================================================================================================================
//   warp 0       one thread issues the tensor-map TMA loads, EB_STAGES tiles ahead (two 128 x 32 boxes of xP; with the first
//                tile of a parcel also the 64 x 32 box of dYpre)
//   warp 1       one thread issues the MMAs of a tile and commits (stage free, remainder buffer free, accumulator ready)
//   warps 2-11   write the tile's remainder lo = tf32(x - trunc_tf32(x)) next to it (the tensor core reads the staged
//                words themselves as hi: it ignores the low 13 mantissa bits, measured); on a parcel change transpose
//                dYpre[b][g] into the K-major B operand [hi | lo] (nearest tf32 and nearest tf32 of the remainder)
//   warps 12-15  read finished accumulators back (tcgen05.ld) and store the dW rows
// Three products per 8 batch rows, two instructions: hi x [B_hi | B_lo] (N = 64: the planes of B are adjacent row groups
// of one operand) and lo x B_hi (N = 32) -- every product type has its own accumulator columns and they are summed
// small-to-large in fp32 by the drain.  Dropped: lo x lo (2^-21 relative).  Remainder buffer, B operand and accumulator
// are double buffered: rounding of tile j + 1, MMAs of tile j and the drain of tile j - 1 overlap.
#define EB_THREADS 512
#define EB_NW 320
#define EB_TILE 128
#define EB_STAGES 3
#define EB_MAXT 256
#define EB_X_BYTES (EB_TILE * 64 * 4)                            // 32 KB: two swizzled boxes [128 pos][32 b]
#define EB_DY_BYTES (64 * 32 * 4)                                // 8 KB: dYpre of a parcel, [b][g]
#define EB_STAGE_BYTES (EB_X_BYTES + EB_DY_BYTES)
#define EB_BPL 8192                                              // one plane of B: [32 g][64 b] fp32, LBO 128, SBO 2048
#define EB_ACC 96                                                // accumulator columns of one tile: hi*hi | hi*lo | lo*hi
#define EB_SMEM (EB_STAGES * EB_STAGE_BYTES + 2 * EB_X_BYTES + 4 * EB_BPL + EB_MAXT * (int)sizeof(EtTile) + 4096 + 1024)

__global__ void __launch_bounds__(EB_THREADS, 1)
enc_bwd_dw_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_dy, int b_begin,
                     int Bh, const int64_t* __restrict__ off, const EncWork* __restrict__ work, int n_work,
                     float* __restrict__ dW, long long* __restrict__ trace, const DpPush push, int64_t g_off) {
  extern __shared__ uint8_t et_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)et_smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* stage_base = smem;                                             // [EB_STAGES] x (tile | dYpre)
  uint8_t* lo_base = smem + EB_STAGES * EB_STAGE_BYTES;                   // [2] remainders
  uint8_t* b_base = lo_base + 2 * EB_X_BYTES;                             // [2] x (B hi | B lo)
  EtTile* tiles = reinterpret_cast<EtTile*>(b_base + 4 * EB_BPL);
  uint8_t* xpose = reinterpret_cast<uint8_t*>(tiles + EB_MAXT);           // [4 drain warps] x 1 KB (push mode)
  __shared__ __align__(8) uint64_t stage_full[EB_STAGES], stage_empty[EB_STAGES], lo_full[2], lo_empty[2], acc_full[2],
      acc_empty[2];
  __shared__ uint32_t tmem_slot;
  __shared__ int s_cnt[ET_MAXI + 1];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < EB_STAGES; ++s) { tc::mbar_init(&stage_full[s], 1); tc::mbar_init(&stage_empty[s], 1); }
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      tc::mbar_init(&lo_full[b], EB_NW); tc::mbar_init(&lo_empty[b], 1);
      tc::mbar_init(&acc_full[b], 1); tc::mbar_init(&acc_empty[b], 128);
    }
    tc::mbar_fence_init();
  }
  if (warp == 1) tc::tmem_alloc(&tmem_slot, 256);
  tc::tc_fence_before();
  const int n_tiles = et_build_tiles<EB_TILE, EB_MAXT>(tiles, s_cnt, work, n_work, off, nullptr);
  tc::tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    // ================= producer ==========================================================================
    if (lane == 0) {
      int last_r = -1;
      for (int j = 0; j < n_tiles; ++j) {
        const int s = j % EB_STAGES;
        const EtTile t = tiles[j];
        tc::mbar_wait(&stage_empty[s], (uint32_t)(((j / EB_STAGES) & 1) ^ 1));
        ET_TR(0);
        uint8_t* dst = stage_base + (size_t)s * EB_STAGE_BYTES;
        const bool new_parcel = t.r != last_r;
        tc::mbar_expect_tx(&stage_full[s], (uint32_t)(EB_X_BYTES + (new_parcel ? EB_DY_BYTES : 0)));
        tc::tma_2d(dst, &map_x, 0, t.pos0, &stage_full[s]);                 // batch rows 0..31 of 128 positions
        tc::tma_2d(dst + EB_X_BYTES / 2, &map_x, 32, t.pos0, &stage_full[s]);
        if (new_parcel) tc::tma_2d(dst + EB_X_BYTES, &map_dy, t.r * 32, b_begin, &stage_full[s]);
        last_r = t.r;
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =========================================================================
    if (lane == 0) {
      constexpr uint32_t idesc64 = tc::idesc_tf32(128, 64), idesc32 = tc::idesc_tf32(128, 32);
      // descriptors differ in the start address only (bits 0-13, address >> 4; the window is < 256 KB): plain adds
      const uint64_t hi_d0 = tc::smem_desc_sw128(tc::smem_u32(stage_base), 1024);
      const uint64_t lo_d0 = tc::smem_desc_sw128(tc::smem_u32(lo_base), 1024);
      const uint64_t b_d0 = tc::smem_desc(tc::smem_u32(b_base), 128, 2048);
      int last_r = -1, pc = -1;
      for (int j = 0; j < n_tiles; ++j) {
        const int s = j % EB_STAGES, q = j & 1, u = j >> 1;
        if (tiles[j].r != last_r) { last_r = tiles[j].r; ++pc; }
        tc::mbar_wait(&lo_full[q], (uint32_t)(u & 1));
        if (u >= 1) tc::mbar_wait(&acc_empty[q], (uint32_t)((u - 1) & 1));
        ET_TR(4);
        tc::tc_fence_after();
        const uint64_t ah0 = hi_d0 + (uint64_t)((s * EB_STAGE_BYTES) >> 4), al0 = lo_d0 + (uint64_t)((q * EB_X_BYTES) >> 4);
        const uint64_t bb0 = b_d0 + (uint64_t)(((pc & 1) * 2 * EB_BPL) >> 4);
        const uint32_t d0 = tmem_base + (uint32_t)(q * EB_ACC);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {                                    // 8 batch rows per instruction
          const uint64_t ao = (uint64_t)(((ks >> 2) * (EB_X_BYTES / 2) + (ks & 3) * 32) >> 4), bo = (uint64_t)((ks * 256) >> 4);
          tc::mma_tf32(d0, ah0 + ao, bb0 + bo, idesc64, ks == 0 ? 0u : 1u);            // hi x [B_hi | B_lo]
          tc::mma_tf32(d0 + 64, al0 + ao, bb0 + bo, idesc32, ks == 0 ? 0u : 1u);       // lo x B_hi
        }
        tc::mma_commit(&lo_empty[q]);                                       // remainder buffer (and the older B) may be rewritten
        tc::mma_commit(&stage_empty[s]);                                    // the stage may be refilled
        tc::mma_commit(&acc_full[q]);
        ET_TR(5);
      }
    }
  } else if (warp < 12) {
    // ================= remainders / B operand (warps 2-11) ====================================================
    const int wt = tid - 64;
    int last_r = -1, pc = -1;
    for (int j = 0; j < n_tiles; ++j) {
      const int s = j % EB_STAGES, q = j & 1, u = j >> 1;
      const EtTile t = tiles[j];
      if (wt == 0) {
        ET_TR(1);
        if (u >= 1) tc::mbar_wait(&lo_empty[q], (uint32_t)((u - 1) & 1)); // the MMAs of tile j - 2 (and all before) are done
        tc::mbar_wait(&stage_full[s], (uint32_t)((j / EB_STAGES) & 1));   // the tile has landed
        ET_TR(2);
      }
      et_group_sync<2, EB_NW>();
      const uint32_t st0 = tc::smem_u32(stage_base + (size_t)s * EB_STAGE_BYTES);
      if (t.r != last_r) {
        // B operand of the parcel: dYpre[b][r][:] as [n = g][k = b], K-major core matrices of fp32 words.  Buffer pc & 1
        // was last read by the MMAs of the parcel before the previous one: those tiles are <= j - 2.
        last_r = t.r; ++pc;
        const uint32_t b0 = tc::smem_u32(b_base + (size_t)(pc & 1) * 2 * EB_BPL);
        const float* dy = reinterpret_cast<const float*>(stage_base + (size_t)s * EB_STAGE_BYTES + EB_X_BYTES);
        for (int c = wt; c < 512; c += EB_NW) {
          const int g = c & 31, b4 = c >> 5;
          float4 h, l;
          float x;
          x = 4 * b4 + 0 < Bh ? dy[(4 * b4 + 0) * 32 + g] : 0.0f; h.x = tc::tf32_rn(x); l.x = tc::tf32_rn(x - h.x);
          x = 4 * b4 + 1 < Bh ? dy[(4 * b4 + 1) * 32 + g] : 0.0f; h.y = tc::tf32_rn(x); l.y = tc::tf32_rn(x - h.y);
          x = 4 * b4 + 2 < Bh ? dy[(4 * b4 + 2) * 32 + g] : 0.0f; h.z = tc::tf32_rn(x); l.z = tc::tf32_rn(x - h.z);
          x = 4 * b4 + 3 < Bh ? dy[(4 * b4 + 3) * 32 + g] : 0.0f; h.w = tc::tf32_rn(x); l.w = tc::tf32_rn(x - h.w);
          const uint32_t o = (uint32_t)((g >> 3) * 2048 + b4 * 128 + (g & 7) * 16);
          tc::sts128f(b0 + o, h);
          tc::sts128f(b0 + EB_BPL + o, l);
        }
      }
      // the tile's remainder at the same (swizzled) offsets; loads first, then arithmetic, then stores
      const uint32_t lo0 = tc::smem_u32(lo_base + (size_t)q * EB_X_BYTES);
      constexpr int NIT = (EB_X_BYTES / 16 + EB_NW - 1) / EB_NW;             // 7
      float4 x[NIT];
#pragma unroll
      for (int i = 0; i < NIT; ++i) {
        const int c = wt + i * EB_NW;
        if (c < EB_X_BYTES / 16) x[i] = tc::lds128(st0 + c * 16);
      }
#pragma unroll
      for (int i = 0; i < NIT; ++i) {
        const int c = wt + i * EB_NW;
        if (c < EB_X_BYTES / 16) {
          float4 l;
          l.x = tc::tf32_rn(x[i].x - __uint_as_float(__float_as_uint(x[i].x) & 0xffffe000u));
          l.y = tc::tf32_rn(x[i].y - __uint_as_float(__float_as_uint(x[i].y) & 0xffffe000u));
          l.z = tc::tf32_rn(x[i].z - __uint_as_float(__float_as_uint(x[i].z) & 0xffffe000u));
          l.w = tc::tf32_rn(x[i].w - __uint_as_float(__float_as_uint(x[i].w) & 0xffffe000u));
          tc::sts128f(lo0 + c * 16, l);
        }
      }
      tc::fence_async_smem();                                               // generic-proxy writes -> visible to the MMA reads
      tc::mbar_arrive(&lo_full[q]);
      if (wt == 0) ET_TR(3);
    }
  } else {
    // ================= drain (warps 12-15) ==================================================================
    const int quad = warp & 3;
    for (int j = 0; j < n_tiles; ++j) {
      const int q = j & 1, u = j >> 1;
      const EtTile t = tiles[j];
      if (lane == 0) tc::mbar_wait(&acc_full[q], (uint32_t)(u & 1));
      __syncwarp();
      if (warp == 12 && lane == 0) ET_TR(6);
      tc::tc_fence_after();
      float v[32], v1[32], v2[32];
      const uint32_t ta = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(q * EB_ACC);
      tc::tmem_ld32(ta, v);
      tc::tmem_ld32(ta + 32, v1);
      tc::tmem_ld32(ta + 64, v2);
      tc::tmem_ld_wait();
      tc::tc_fence_before();
      tc::mbar_arrive(&acc_empty[q]);
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] += v1[i] + v2[i];
      const int k = quad * 32 + lane;
      if (push.world == 0) {
        if (k < t.nv) {
          float4* dst = reinterpret_cast<float4*>(dW + (size_t)(t.pos0 + k) * 32);
#pragma unroll
          for (int i = 0; i < 8; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
      } else {
        // Data parallel: the row goes to the rank that owns its slice of the flat gradient, over NVLink.  Remote stores
        // must be wide: the warp's 32 rows pass through a 1 KB shared tile eight at a time (16-byte chunks XOR-swizzled
        // by row: both directions bank-conflict free) so that one store instruction writes 4 whole rows = 512
        // contiguous bytes instead of 32 separate 16-byte pieces.
        const uint32_t xp = tc::smem_u32(xpose) + (uint32_t)(quad * 1024);
#pragma unroll
        for (int pass = 0; pass < 4; ++pass) {
          if ((lane >> 3) == pass) {
            const int rr = lane & 7;
#pragma unroll
            for (int i = 0; i < 8; ++i)
              tc::sts128f(xp + (uint32_t)(rr * 128 + ((i ^ rr) << 4)), make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]));
          }
          __syncwarp();
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int rr = (lane >> 3) + 4 * h, ch = lane & 7;
            const float4 val = tc::lds128(xp + (uint32_t)(rr * 128 + ((ch ^ rr) << 4)));
            const int kk = quad * 32 + pass * 8 + rr;
            if (kk < t.nv) {
              const int64_t o = g_off + (int64_t)(t.pos0 + kk) * 32 + ch * 4;      // flat offset in g
              int q = 0;
#pragma unroll
              for (int z = 1; z < TNT_DP_MAX_WORLD; ++z) q += (z < push.world && o >= (int64_t)z * push.SL) ? 1 : 0;
              *reinterpret_cast<float4*>(push.in[q] + (size_t)push.rank * push.SL + (o - (int64_t)q * push.SL)) = val;
            }
          }
          __syncwarp();
        }
      }
      if (warp == 12 && lane == 0) ET_TR(7);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem_base, 256);
}

// =====================================================================================================================
// forward: per tile M = 64 batch rows, N = 32 channels, K = 64 positions; a work item's tiles are summed in registers
// =====================================================================================================================
//   warp 0        one thread issues the bulk copies (the tile's rows of xP and of W), EF_STAGES tiles ahead
//   warp 1        one thread issues the 12 MMAs of a tile (kind::f16 on bf16 planes) and commits
//   warps 2-7     converter group 0: even tiles -> operand buffer 0
//   warps 8-13    converter group 1: odd tiles  -> operand buffer 1
//   warps 14-17   drain: tcgen05.ld of every finished tile, summed over the work item, stored at its last tile
// A[m = batch row][k = position]: a thread owns a block of 4 batch rows x 8 positions -- eight 128-bit loads along the
// staged rows (a quarter warp reads 128 contiguous bytes), four operand chunks (8 consecutive positions of one batch
// row) per plane out.  Row groups of a plane sit 1040 bytes apart (8 chunks of 128 + 16): the eight lanes of a quarter
// warp then store to eight different 16-byte bank groups.  B[n = channel][k = position] likewise from the staged W rows.
// M = 64: accumulator row i lives in TMEM lane 32 (i / 16) + i % 16 (tools/tmem_probe.cu, profiles/r2e_tmem_probe.txt).
#define EF_THREADS 576
#define EF_GROUP 192
#define EF_TILE 64
#define EF_STAGES 5
#define EF_X_BYTES (EF_TILE * 64 * 4)                           // 16 KB of xP
#define EF_W_BYTES (EF_TILE * 32 * 4)                           // 8 KB of W
#define EF_STAGE_BYTES (EF_X_BYTES + EF_W_BYTES)
#define EF_SBO 1040
#define EF_A_BYTES (8 * EF_SBO)                                 // one plane of A: 8 row groups x 8 position groups
#define EF_B_BYTES (4 * EF_SBO)
#define EF_OP_BYTES (ET_NS * (EF_A_BYTES + EF_B_BYTES))         // one operand buffer
#define EF_ACC 192                                               // accumulator columns of one tile: 96 + 64 + 32
#define EF_SMEM (EF_STAGES * EF_STAGE_BYTES + 2 * EF_OP_BYTES + ET_MAXT * (int)sizeof(EtTile) + 1024)

__global__ void __launch_bounds__(EF_THREADS, 1)
enc_fwd_tc_kernel(const float* __restrict__ xP, int Bh, const int64_t* __restrict__ off, const EncWork* __restrict__ work,
                  int n_work, const int32_t* __restrict__ slot_first, const float* __restrict__ W,
                  float* __restrict__ part, long long* __restrict__ trace) {
  extern __shared__ uint8_t et_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)et_smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* stage_base = smem;
  uint8_t* op_base = smem + EF_STAGES * EF_STAGE_BYTES;                   // [2] x (A planes | B planes)
  EtTile* tiles = reinterpret_cast<EtTile*>(op_base + 2 * EF_OP_BYTES);
  __shared__ __align__(8) uint64_t stage_full[EF_STAGES], stage_empty[EF_STAGES], op_full[2], op_empty[2], acc_full[2],
      acc_empty[2];
  __shared__ uint32_t tmem_slot;
  __shared__ int s_cnt[ET_MAXI + 1];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < EF_STAGES; ++s) { tc::mbar_init(&stage_full[s], 1); tc::mbar_init(&stage_empty[s], EF_GROUP); }
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      tc::mbar_init(&op_full[b], EF_GROUP); tc::mbar_init(&op_empty[b], 1);
      tc::mbar_init(&acc_full[b], 1); tc::mbar_init(&acc_empty[b], 128);
    }
    tc::mbar_fence_init();
  }
  if (warp == 1) tc::tmem_alloc(&tmem_slot, 512);
  tc::tc_fence_before();
  const int n_tiles = et_build_tiles<EF_TILE, ET_MAXT>(tiles, s_cnt, work, n_work, off, slot_first);
  tc::tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    // ================= producer ==========================================================================
    if (lane == 0)
      for (int j = 0; j < n_tiles; ++j) {
        const int s = j % EF_STAGES;
        const EtTile t = tiles[j];
        tc::mbar_wait(&stage_empty[s], (uint32_t)(((j / EF_STAGES) & 1) ^ 1));
        ET_TR(0);
        tc::mbar_expect_tx(&stage_full[s], (uint32_t)((256 + 128) * t.nv));
        uint8_t* dst = stage_base + (size_t)s * EF_STAGE_BYTES;
        tc::bulk_g2s(dst, xP + (size_t)t.pos0 * 64, (uint32_t)(256 * t.nv), &stage_full[s]);
        tc::bulk_g2s(dst + EF_X_BYTES, W + (size_t)t.pos0 * 32, (uint32_t)(128 * t.nv), &stage_full[s]);
      }
  } else if (warp == 1) {
    // ================= MMA issuer: one accumulation chain per tile (K = 64) =================================
    // Six bf16 products per 16 positions in three instructions: the planes of B are adjacent row groups of ONE operand,
    // so A_hi x [B_hi | B_mid | B_lo] is a single N = 96 MMA, A_mid x [B_hi | B_mid] N = 64, A_lo x B_hi N = 32 -- same
    // tensor-pipe time as six N = 32 instructions at half the issue cost (the issuing thread was the limiter: ~70 cycles
    // per instruction, tools/enc_trace.py), and every product type keeps its own accumulator columns.
    if (lane == 0) {
      constexpr uint32_t id96 = tc::idesc_bf16(64, 96, 0, 0), id64 = tc::idesc_bf16(64, 64, 0, 0), id32 = tc::idesc_bf16(64, 32, 0, 0);
      const uint64_t op_d0 = tc::smem_desc(tc::smem_u32(op_base), 128u, EF_SBO);
      for (int j = 0; j < n_tiles; ++j) {
        const int q = j & 1, u = j >> 1;
        tc::mbar_wait(&op_full[q], (uint32_t)(u & 1));
        if (u >= 1) tc::mbar_wait(&acc_empty[q], (uint32_t)((u - 1) & 1));
        ET_TR(4);
        tc::tc_fence_after();
        const uint64_t a0 = op_d0 + (uint64_t)((q * EF_OP_BYTES) >> 4), b0 = a0 + (uint64_t)((ET_NS * EF_A_BYTES) >> 4);
        const uint32_t d0 = tmem_base + (uint32_t)(q * EF_ACC);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t ko = (uint64_t)((ks * 256) >> 4);
          const uint32_t acc = ks == 0 ? 0u : 1u;
          tc::mma_bf16(d0, a0 + ko, b0 + ko, id96, acc);                                          // hi x [hi | mid | lo]
          tc::mma_bf16(d0 + 96, a0 + (uint64_t)(EF_A_BYTES >> 4) + ko, b0 + ko, id64, acc);        // mid x [hi | mid]
          tc::mma_bf16(d0 + 160, a0 + (uint64_t)((2 * EF_A_BYTES) >> 4) + ko, b0 + ko, id32, acc); // lo x hi
        }
        tc::mma_commit(&op_empty[q]);
        tc::mma_commit(&acc_full[q]);
        ET_TR(5);
      }
    }
  } else if (warp < 14) {
    // ================= converter groups ====================================================================
    const int q = warp >= 8 ? 1 : 0;
    const int gt = tid - 64 - q * EF_GROUP;
    const uint32_t op0 = tc::smem_u32(op_base + (size_t)q * EF_OP_BYTES);
    // block of this thread: A (threads 0..127): 4 batch rows x 8 positions; B (threads 128..191): 4 channels x 8 positions
    const bool is_a = gt < 128;
    const int cq = is_a ? (gt & 15) : ((gt - 128) & 7);                   // which group of 4 rows / channels
    const int vg = is_a ? (gt >> 4) : ((gt - 128) >> 3);                  // which group of 8 positions
    const uint32_t src_off = is_a ? (uint32_t)(vg * 8 * 256 + cq * 16) : (uint32_t)(EF_X_BYTES + vg * 8 * 128 + cq * 16);
    const uint32_t src_pitch = is_a ? 256u : 128u;
    const uint32_t dst_off = (is_a ? 0u : (uint32_t)(ET_NS * EF_A_BYTES)) +
                             (uint32_t)((cq >> 1) * EF_SBO + vg * 128 + (cq & 1) * 64);
    const uint32_t plane = is_a ? (uint32_t)EF_A_BYTES : (uint32_t)EF_B_BYTES;
    for (int j = q; j < n_tiles; j += 2) {
      const int s = j % EF_STAGES, u = j >> 1;
      const EtTile t = tiles[j];
      if (gt == 0) {
        ET_TR(1);
        if (u >= 1) tc::mbar_wait(&op_empty[q], (uint32_t)((u - 1) & 1)); // the MMAs of this buffer's previous tile are done
        tc::mbar_wait(&stage_full[s], (uint32_t)((j / EF_STAGES) & 1));   // the tile has landed
        ET_TR(2);
      }
      if (q == 0) et_group_sync<2, EF_GROUP>(); else et_group_sync<3, EF_GROUP>();
      const uint32_t src = tc::smem_u32(stage_base + (size_t)s * EF_STAGE_BYTES) + src_off;
      float4 r[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        r[i] = tc::lds128(src + i * src_pitch);
        if (vg * 8 + i >= t.nv) r[i] = make_float4(0.f, 0.f, 0.f, 0.f);    // positions beyond the tile's end were not copied
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float x[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = e == 0 ? r[i].x : e == 1 ? r[i].y : e == 2 ? r[i].z : r[i].w;
        uint4 pl[ET_NS];
        tc::split8<ET_NS>(x, pl);
#pragma unroll
        for (int p = 0; p < ET_NS; ++p) tc::sts128(op0 + dst_off + p * plane + e * 16, pl[p]);
      }
      tc::fence_async_smem();
      tc::mbar_arrive(&op_full[q]);
      tc::mbar_arrive(&stage_empty[s]);
      if (gt == 0) ET_TR(3);
    }
  } else {
    // ================= drain (warps 14-17) ================================================================
    const int quad = warp & 3;
    float accr[32];                                                          // the work item's sum over its tiles
#pragma unroll
    for (int i = 0; i < 32; ++i) accr[i] = 0.0f;
    for (int j = 0; j < n_tiles; ++j) {
      const int q = j & 1, u = j >> 1;
      const EtTile t = tiles[j];
      if (lane == 0) tc::mbar_wait(&acc_full[q], (uint32_t)(u & 1));
      __syncwarp();
      if (warp == 14 && lane == 0) ET_TR(6);
      tc::tc_fence_after();
      // columns: hi*hi | hi*mid | hi*lo | mid*hi | mid*mid | lo*hi -- summed small to large
      const uint32_t ta = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(q * EF_ACC);
      float v[32], w[32];
      tc::tmem_ld32(ta + 160, v);                                             // lo*hi
      tc::tmem_ld32(ta + 128, w);                                             // mid*mid
      tc::tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] += w[i];
      tc::tmem_ld32(ta + 64, w);                                              // hi*lo
      tc::tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] += w[i];
      tc::tmem_ld32(ta + 96, w);                                              // mid*hi
      tc::tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] += w[i];
      tc::tmem_ld32(ta + 32, w);                                              // hi*mid
      tc::tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] += w[i];
      tc::tmem_ld32(ta, w);                                                   // hi*hi
      tc::tmem_ld_wait();
      tc::tc_fence_before();
      tc::mbar_arrive(&acc_empty[q]);
#pragma unroll
      for (int i = 0; i < 32; ++i) accr[i] += v[i] + w[i];
      if (t.last) {
        const int b = quad * 16 + lane;                                      // M = 64: row i -> lane 32 (i / 16) + i % 16
        if (lane < 16 && b < Bh) {
          float4* dst = reinterpret_cast<float4*>(part + ((size_t)t.slot * Bh + b) * 32);
#pragma unroll
          for (int i = 0; i < 8; ++i) dst[i] = make_float4(accr[4 * i], accr[4 * i + 1], accr[4 * i + 2], accr[4 * i + 3]);
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) accr[i] = 0.0f;
      }
      if (warp == 14 && lane == 0) ET_TR(7);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem_base, 512);
}

// ---------------------------------------------------------------------------------------------------------------------
typedef CUresult (*EtEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled through the runtime (libcuda is not linked); the pointer is process-wide and immutable
static EtEncodeFn et_encode_fn() {
  static EtEncodeFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    cudaGetLastError();
    return reinterpret_cast<EtEncodeFn>(p);
  }();
  return fn;
}

// fp32 matrix (rows, cols) row-major -> tensor map with a (box_rows x 32 columns = 128 bytes) box
static bool et_make_map(void* out, const void* base, uint64_t cols, uint64_t rows, uint32_t box_rows, bool swizzle) {
  EtEncodeFn enc = et_encode_fn();
  if (!enc) return false;
  cuuint64_t dims[2] = {cols, rows}, strides[1] = {cols * 4};
  cuuint32_t box[2] = {32, box_rows}, es[2] = {1, 1};
  CUtensorMap m;
  if (enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, es,
          CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
          CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return false;
  static_assert(sizeof(CUtensorMap) == 128, "CUtensorMap is 128 bytes");
  memcpy(out, &m, sizeof(m));
  return true;
}

static bool et_prepare(tnt_ctx* c, const EncTables& t, int* grid) {
  if (!c->et_attr) {
    if (cudaFuncSetAttribute(enc_bwd_dw_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, EB_SMEM) != cudaSuccess ||
        cudaFuncSetAttribute(enc_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, EF_SMEM) != cudaSuccess) {
      cudaGetLastError();
      ctx_fail(c, TNT_ERR_CUDA, "encoder_tc: cannot reserve shared memory");
      return false;
    }
    c->et_attr = 1;
  }
  *grid = t.n_work < c->sm_count ? t.n_work : c->sm_count;
  if (t.n_work > *grid * ET_MAXI) {   // a CTA owns <= 64 work items (<= 512 tiles: items hold <= 512 positions)
    ctx_fail(c, TNT_ERR_ARG, "encoder_tc: more than 64 work items per SM (raise ET_MAXI / ET_MAXT)");
    return false;
  }
  return true;
}

// partial sums of encoder e (-> enc_part, one slot per work item) from the parcel-ordered batch copy
bool enc_fwd_tc(tnt_ctx* c, int e, const float* xP, int Bh, cudaStream_t st) {
  EncTables& t = c->enc[e];
  int grid;
  if (!et_prepare(c, t, &grid)) return false;
  enc_fwd_tc_kernel<<<grid, EF_THREADS, EF_SMEM, st>>>(xP, Bh, t.off, t.work, t.n_work, t.slot_first, c->w + t.w_off,
                                                       wsT<float>(c, "enc_part"), c->et_trace);
  c->launches += 1;
  return true;
}

// dW of encoder e from the parcel-ordered batch copy; false: the context is marked failed
bool enc_bwd_dw_tc(tnt_ctx* c, int e, const float* xP, int b_begin, int Bh, cudaStream_t st) {
  EncTables& t = c->enc[e];
  int grid;
  if (!et_prepare(c, t, &grid)) return false;
  const float* dY = wsT<float>(c, "dYpre");
  if (t.tmap_x_base != xP || t.tmap_dy_base != dY) {
    // xP: (n_lab + 128 rows of padding, 64); dYpre: (B_max, R * 32).  Rows past the end read as zeros.
    if (!et_make_map(t.tmap_x, xP, 64, (uint64_t)t.n_lab + 128, EB_TILE, true) ||
        !et_make_map(t.tmap_dy, dY, (uint64_t)c->cfg.R * 32, (uint64_t)c->cfg.B_max, 64, false)) {
      ctx_fail(c, TNT_ERR_CUDA, "encoder_tc: cuTensorMapEncodeTiled failed");
      return false;
    }
    t.tmap_x_base = xP; t.tmap_dy_base = dY;
  }
  CUtensorMap mx, my;
  memcpy(&mx, t.tmap_x, sizeof(mx));
  memcpy(&my, t.tmap_dy, sizeof(my));
  DpPush push = {};
  if (c->dp_push_on && dp_push_ready(c) && t.w_off % 4 == 0 && c->dp_n_pushed < 4) {
    push = dp_push_params(c);                              // the rows go to their owners' staging, not to c->g
    c->dp_pushed[2 * c->dp_n_pushed] = t.w_off;
    c->dp_pushed[2 * c->dp_n_pushed + 1] = t.w_off + t.n_lab * 32;
    c->dp_n_pushed += 1;
  }
  enc_bwd_dw_tc_kernel<<<grid, EB_THREADS, EB_SMEM, st>>>(mx, my, b_begin, Bh, t.off, t.work, t.n_work, c->g + t.w_off,
                                                          c->et_trace ? c->et_trace + 8 * ET_MAXT : nullptr, push, t.w_off);
  c->launches += 1;
  return true;
}

// diagnostic (tools/enc_trace.py): device buffer of 2 x 8 x ET_MAXT clock64() stamps (forward | gradient), nullptr = off
extern "C" void tnt_debug_encoder_trace(tnt_ctx* c, long long* dev_buf) { c->et_trace = dev_buf; }
