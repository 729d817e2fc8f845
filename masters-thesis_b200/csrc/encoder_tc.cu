// Parcel encoder on the Blackwell tensor path (north_star item 1): LocallyDense (AttemptFour/Model/layers.py:43-53)
//   forward   partial_r[b][g] = sum_k x[b, idx_r[k]] * W_r[k][g]                (layers.py:45, before bias / LeakyReLU)
//   gradient  dW_r[k][g]      = sum_b x[b, idx_r[k]] * dYpre[b][r][g]           (tape gradient of the same line)
// as tcgen05.mma tiles with the accumulator in tensor memory, fed by the TMA engine.
//
// Both kernels stream the parcel-ordered copy of the batch xP[n_lab][64] (encoder.cu: enc_permute_kernel; row = one
// position of the concatenated parcel index list, 64 batch values = 256 bytes) and the packed kernel W (n_lab, 32) in
// the same order: a tile of 128 positions is ONE contiguous 32 KB (xP) / 16 KB (W) bulk copy (cp.async.bulk, completion
// by mbarrier transaction count) into a staging ring -- no gather, no LSU instruction, no register on the load path.
// (Gathering the rows of a vertex-major copy was measured first: one 256-byte bulk copy per row costs ~85 cycles inside
// the TMA unit, 89.6 us per pass; 16-byte LDGSTS gathers 52.8 us; the old mma.sync kernel 40.4 us -- profiles/r2*.)
//   warp 0      one thread issues the bulk copies, STAGES tiles ahead
//   warp 1      one thread issues the MMAs of a tile (kind::f16 on bf16 operand planes, fp32 accumulate in TMEM)
//   warps 2-15  convert a landed fp32 tile into three bf16 planes (x = hi + mid + lo, 24 mantissa bits: fp32-equivalent
//               products) in the tensor core's K-major core-matrix layout; warps 12-15 also read finished accumulators
//               back (tcgen05.ld) and store them
// Persistent: one CTA per SM walks the work list (items of <= 512 positions, largest first), cut into 128-position tiles.
#include "tnt_internal.h"
#include "tc.cuh"

#define ET_THREADS 512
#define ET_TILE 128
#define ET_NS 3
#define ET_NW 448                        // converter threads (warps 2-15)
#define ET_MAXT 256                      // tiles one CTA may own
#define ET_MAXI 64                       // work items one CTA may own
struct EtTile { int64_t pos0; int32_t nv; int32_t r; int32_t slot; int32_t first, last, pad; };

// the CTA's tile list: work items blockIdx.x, + gridDim.x, ... cut into tiles of ET_TILE positions
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
    s_cnt[tid] = (wk.kend - wk.kbeg + ET_TILE - 1) / ET_TILE;
  }
  __syncthreads();
  if (tid < n_items) {
    int o = 0;
    for (int i = 0; i < tid; ++i) o += s_cnt[i];
    for (int k0 = wk.kbeg; k0 < wk.kend; k0 += ET_TILE, ++o)
      if (o < ET_MAXT) {
        EtTile t; t.pos0 = base + k0; t.nv = min(ET_TILE, wk.kend - k0); t.r = wk.r; t.slot = slot;
        t.first = k0 == wk.kbeg; t.last = k0 + ET_TILE >= wk.kend; t.pad = 0;
        tiles[o] = t;
      }
    if (tid == n_items - 1) s_cnt[ET_MAXI] = min(o, ET_MAXT);
  }
  __syncthreads();
  return n_items > 0 ? s_cnt[ET_MAXI] : 0;
}

// The 448 converter threads wait for an mbarrier phase through ONE poller and a hardware named barrier: fourteen warps
// spinning on try_wait starve the single MMA-issuing thread of issue slots (measured: 6-10 k cycles per tile).
__device__ __forceinline__ void et_conv_sync() { asm volatile("bar.sync 2, 448;" ::: "memory"); }

// =====================================================================================================================
// gradient: per tile M = 128 positions, N = 32 channels, K = 64 batch rows
// =====================================================================================================================
// A[m = position][k = batch row] is the staged tile itself (row = 256 contiguous bytes): a 16-byte operand chunk is 8
// consecutive batch values of one position.  Chunks of one position are read by neighbouring lanes (contiguous, conflict
// free); to keep the operand WRITES conflict free too the core matrices along K sit 144 bytes apart instead of 128
// (descriptor LBO = 144, SBO = 8 * 144): lanes bg = 0..7 land in eight different bank groups.
#define EB_STAGES 4
#define EB_STAGE_BYTES (ET_TILE * 64 * 4)                       // 32 KB
#define EB_LBO 144
#define EB_SBO (8 * EB_LBO)
#define EB_A_BYTES (16 * EB_SBO)                                // one plane of A: 16 row groups
#define EB_B_BYTES (4 * EB_SBO)                                 // one plane of B: 32 channels
#define EB_SMEM (EB_STAGES * EB_STAGE_BYTES + ET_NS * EB_A_BYTES + ET_NS * EB_B_BYTES + ET_MAXT * (int)sizeof(EtTile) + 1024)

__global__ void __launch_bounds__(ET_THREADS, 1)
enc_bwd_dw_tc_kernel(const float* __restrict__ xP, int b_begin, int Bh, int R, const int64_t* __restrict__ off,
                     const EncWork* __restrict__ work, int n_work, const float* __restrict__ dYpre,
                     float* __restrict__ dW) {
  extern __shared__ uint8_t et_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)et_smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* stage_base = smem;                                             // [EB_STAGES][128][64] fp32
  uint8_t* a_base = smem + EB_STAGES * EB_STAGE_BYTES;                    // [NS] planes
  uint8_t* b_base = a_base + ET_NS * EB_A_BYTES;
  EtTile* tiles = reinterpret_cast<EtTile*>(b_base + ET_NS * EB_B_BYTES);
  __shared__ __align__(8) uint64_t stage_full[EB_STAGES], stage_empty[EB_STAGES], a_full, a_empty, acc_full[2], acc_empty[2];
  __shared__ uint32_t tmem_slot;
  __shared__ int s_cnt[ET_MAXI + 1];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < EB_STAGES; ++s) { tc::mbar_init(&stage_full[s], 1); tc::mbar_init(&stage_empty[s], ET_NW); }
    tc::mbar_init(&a_full, ET_NW); tc::mbar_init(&a_empty, 1);
#pragma unroll
    for (int b = 0; b < 2; ++b) { tc::mbar_init(&acc_full[b], 1); tc::mbar_init(&acc_empty[b], 128); }
    tc::mbar_fence_init();
  }
  if (warp == 1) tc::tmem_alloc(&tmem_slot, 128);
  tc::tc_fence_before();
  const int n_tiles = et_build_tiles(tiles, s_cnt, work, n_work, off, nullptr);
  tc::tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    // ================= producer: one 32 KB bulk copy per tile ==============================================
    if (lane == 0)
      for (int j = 0; j < n_tiles; ++j) {
        const int s = j % EB_STAGES;
        const EtTile t = tiles[j];
        tc::mbar_wait(&stage_empty[s], (uint32_t)(((j / EB_STAGES) & 1) ^ 1));
        tc::mbar_expect_tx(&stage_full[s], (uint32_t)(256 * t.nv));
        tc::bulk_g2s(stage_base + (size_t)s * EB_STAGE_BYTES, xP + (size_t)t.pos0 * 64, (uint32_t)(256 * t.nv), &stage_full[s]);
      }
  } else if (warp == 1) {
    // ================= MMA issuer =========================================================================
    if (lane == 0) {
      constexpr uint32_t idesc = tc::idesc_bf16(128, 32, 0, 0);
      for (int j = 0; j < n_tiles; ++j) {
        const int buf = j & 1, u = j >> 1;
        tc::mbar_wait(&a_full, (uint32_t)(j & 1));
        if (u >= 1) tc::mbar_wait(&acc_empty[buf], (uint32_t)((u - 1) & 1));
        tc::tc_fence_after();
        const uint32_t a0 = tc::smem_u32(a_base), b0 = tc::smem_u32(b_base);
        // the leading hi*hi products go to their own accumulator, the 2^-8 and smaller cross terms to a second one: the
        // tensor core adds into TMEM with truncation, and this keeps the biased part of that error 2^-8 smaller (gemm_tc.cu)
        const uint32_t d_main = tmem_base + (uint32_t)(buf * 64), d_corr = d_main + 32;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          uint64_t da[ET_NS], db[ET_NS];
#pragma unroll
          for (int q = 0; q < ET_NS; ++q) {
            da[q] = tc::smem_desc(a0 + q * EB_A_BYTES + ks * 2 * EB_LBO, EB_LBO, EB_SBO);
            db[q] = tc::smem_desc(b0 + q * EB_B_BYTES + ks * 2 * EB_LBO, EB_LBO, EB_SBO);
          }
          bool cfirst = ks == 0;
#pragma unroll
          for (int ord = ET_NS - 1; ord >= 1; --ord)
#pragma unroll
            for (int q = 0; q <= ord; ++q) {
              tc::mma_bf16(d_corr, da[q], db[ord - q], idesc, cfirst ? 0u : 1u);
              cfirst = false;
            }
          tc::mma_bf16(d_main, da[0], db[0], idesc, ks == 0 ? 0u : 1u);
        }
        tc::mma_commit(&a_empty);
        tc::mma_commit(&acc_full[buf]);
      }
    }
  } else {
    // ================= converters (warps 2-15); warps 12-15 also drain the accumulators ========================
    const int wt = tid - 64;
    const int quad = warp & 3;
    auto drain = [&](int jd) {
      const int buf = jd & 1, u = jd >> 1;
      const EtTile t = tiles[jd];
      if (lane == 0) tc::mbar_wait(&acc_full[buf], (uint32_t)(u & 1));
      __syncwarp();
      tc::tc_fence_after();
      float v[32], vc[32];
      tc::tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(buf * 64), v);
      tc::tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(buf * 64 + 32), vc);
      tc::tmem_ld_wait();
      tc::tc_fence_before();
      tc::mbar_arrive(&acc_empty[buf]);
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] += vc[i];
      const int k = quad * 32 + lane;
      if (k < t.nv) {
        float4* dst = reinterpret_cast<float4*>(dW + (size_t)(t.pos0 + k) * 32);
#pragma unroll
        for (int i = 0; i < 8; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
      }
    };
    int last_r = -1;
    for (int j = 0; j < n_tiles; ++j) {
      const int s = j % EB_STAGES;
      const EtTile t = tiles[j];
      if (wt == 0) {
        if (j >= 1) tc::mbar_wait(&a_empty, (uint32_t)((j - 1) & 1));     // the MMAs of the previous tile have read A (and B)
        tc::mbar_wait(&stage_full[s], (uint32_t)((j / EB_STAGES) & 1));   // the tile has landed
      }
      et_conv_sync();
      if (t.r != last_r) {
        // B operand of the parcel: dYpre[b][r][:] as [n = g][k = b]
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
          const size_t o = (size_t)((g >> 3) * EB_SBO + bg * EB_LBO + (g & 7) * 16);
#pragma unroll
          for (int q = 0; q < ET_NS; ++q) *reinterpret_cast<uint4*>(b_base + q * EB_B_BYTES + o) = pl[q];
        }
        last_r = t.r;
      }
      const float* st = reinterpret_cast<const float*>(stage_base + (size_t)s * EB_STAGE_BYTES);
      for (int c = wt; c < 1024; c += ET_NW) {
        const int bg = c & 7, v = c >> 3;                                   // 8 lanes read one position's 256 bytes
        const float* src = st + (size_t)v * 64 + bg * 8;
        const float4 lo4 = *reinterpret_cast<const float4*>(src), hi4 = *reinterpret_cast<const float4*>(src + 4);
        float x[8] = {lo4.x, lo4.y, lo4.z, lo4.w, hi4.x, hi4.y, hi4.z, hi4.w};
        if (v >= t.nv) {
#pragma unroll
          for (int i = 0; i < 8; ++i) x[i] = 0.0f;                           // positions beyond the tile's end were not copied
        }
        uint4 pl[ET_NS];
        tc::split8<ET_NS>(x, pl);
        const size_t o = (size_t)((v >> 3) * EB_SBO + bg * EB_LBO + (v & 7) * 16);
#pragma unroll
        for (int q = 0; q < ET_NS; ++q) *reinterpret_cast<uint4*>(a_base + q * EB_A_BYTES + o) = pl[q];
      }
      tc::fence_async_smem();                                               // generic-proxy writes -> visible to the MMA reads
      tc::mbar_arrive(&a_full);
      tc::mbar_arrive(&stage_empty[s]);
      if (warp >= 12 && j >= 1) drain(j - 1);
    }
    if (warp >= 12 && n_tiles >= 1) drain(n_tiles - 1);
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem_base, 128);
}

// =====================================================================================================================
// forward: per work item M = 64 batch rows, N = 32 channels, K = its positions (tiles of 128)
// =====================================================================================================================
// A[m = batch row][k = position]: an operand chunk is 8 consecutive positions of one batch row = a column walk of the
// staged tile (stride 256 bytes, neighbouring lanes = neighbouring batch rows: conflict free), written 128 contiguous
// bytes per 8 lanes.  B[n = channel][k = position] likewise from the staged W rows.  M = 64: the accumulator row i lives
// in TMEM lane 32 (i / 16) + i % 16 (measured with tools/tmem_probe.cu, profiles/r2e_tmem_probe.txt).
#define EF_STAGES 3
#define EF_X_BYTES (ET_TILE * 64 * 4)                           // 32 KB of xP
#define EF_W_BYTES (ET_TILE * 32 * 4)                           // 16 KB of W
#define EF_STAGE_BYTES (EF_X_BYTES + EF_W_BYTES)
#define EF_A_BYTES (64 * 128 * 2)                               // [64 rows][128 k] bf16: LBO 128, SBO 16 * 128
#define EF_B_BYTES (32 * 128 * 2)
#define EF_SMEM (EF_STAGES * EF_STAGE_BYTES + ET_NS * EF_A_BYTES + ET_NS * EF_B_BYTES + ET_MAXT * (int)sizeof(EtTile) + 1024)

__global__ void __launch_bounds__(ET_THREADS, 1)
enc_fwd_tc_kernel(const float* __restrict__ xP, int Bh, const int64_t* __restrict__ off, const EncWork* __restrict__ work,
                  int n_work, const int32_t* __restrict__ slot_first, const float* __restrict__ W,
                  float* __restrict__ part) {
  extern __shared__ uint8_t et_smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)et_smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* stage_base = smem;
  uint8_t* a_base = smem + EF_STAGES * EF_STAGE_BYTES;
  uint8_t* b_base = a_base + ET_NS * EF_A_BYTES;
  EtTile* tiles = reinterpret_cast<EtTile*>(b_base + ET_NS * EF_B_BYTES);
  __shared__ __align__(8) uint64_t stage_full[EF_STAGES], stage_empty[EF_STAGES], a_full, a_empty, acc_full[2], acc_empty[2];
  __shared__ uint32_t tmem_slot;
  __shared__ int s_cnt[ET_MAXI + 1];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < EF_STAGES; ++s) { tc::mbar_init(&stage_full[s], 1); tc::mbar_init(&stage_empty[s], ET_NW); }
    tc::mbar_init(&a_full, ET_NW); tc::mbar_init(&a_empty, 1);
#pragma unroll
    for (int b = 0; b < 2; ++b) { tc::mbar_init(&acc_full[b], 1); tc::mbar_init(&acc_empty[b], 128); }
    tc::mbar_fence_init();
  }
  if (warp == 1) tc::tmem_alloc(&tmem_slot, 128);
  tc::tc_fence_before();
  const int n_tiles = et_build_tiles(tiles, s_cnt, work, n_work, off, slot_first);
  tc::tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    // ================= producer: the tile's rows of xP and of W, one bulk copy each ==========================
    if (lane == 0)
      for (int j = 0; j < n_tiles; ++j) {
        const int s = j % EF_STAGES;
        const EtTile t = tiles[j];
        tc::mbar_wait(&stage_empty[s], (uint32_t)(((j / EF_STAGES) & 1) ^ 1));
        tc::mbar_expect_tx(&stage_full[s], (uint32_t)((256 + 128) * t.nv));
        uint8_t* dst = stage_base + (size_t)s * EF_STAGE_BYTES;
        tc::bulk_g2s(dst, xP + (size_t)t.pos0 * 64, (uint32_t)(256 * t.nv), &stage_full[s]);
        tc::bulk_g2s(dst + EF_X_BYTES, W + (size_t)t.pos0 * 32, (uint32_t)(128 * t.nv), &stage_full[s]);
      }
  } else if (warp == 1) {
    // ================= MMA issuer: one accumulation chain per tile (K = 128) ================================
    if (lane == 0) {
      constexpr uint32_t idesc = tc::idesc_bf16(64, 32, 0, 0);
      for (int j = 0; j < n_tiles; ++j) {
        const int buf = j & 1, u = j >> 1;
        tc::mbar_wait(&a_full, (uint32_t)(j & 1));
        if (u >= 1) tc::mbar_wait(&acc_empty[buf], (uint32_t)((u - 1) & 1));
        tc::tc_fence_after();
        const uint32_t a0 = tc::smem_u32(a_base), b0 = tc::smem_u32(b_base);
        const uint32_t d_main = tmem_base + (uint32_t)(buf * 64), d_corr = d_main + 32;   // hi*hi | cross terms (see above)
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          uint64_t da[ET_NS], db[ET_NS];
#pragma unroll
          for (int q = 0; q < ET_NS; ++q) {
            da[q] = tc::smem_desc(a0 + q * EF_A_BYTES + ks * 256u, 128u, 2048u);
            db[q] = tc::smem_desc(b0 + q * EF_B_BYTES + ks * 256u, 128u, 2048u);
          }
          bool cfirst = ks == 0;
#pragma unroll
          for (int ord = ET_NS - 1; ord >= 1; --ord)
#pragma unroll
            for (int q = 0; q <= ord; ++q) {
              tc::mma_bf16(d_corr, da[q], db[ord - q], idesc, cfirst ? 0u : 1u);
              cfirst = false;
            }
          tc::mma_bf16(d_main, da[0], db[0], idesc, ks == 0 ? 0u : 1u);
        }
        tc::mma_commit(&a_empty);
        tc::mma_commit(&acc_full[buf]);
      }
    }
  } else {
    // ================= converters (warps 2-15); warps 12-15 also drain every tile into registers ============
    const int wt = tid - 64;
    const int quad = warp & 3;
    float accr[32];                                                          // the work item's sum over its tiles (drain warps)
#pragma unroll
    for (int i = 0; i < 32; ++i) accr[i] = 0.0f;
    auto drain = [&](int jd) {
      const int buf = jd & 1, u = jd >> 1;
      const EtTile t = tiles[jd];
      if (lane == 0) tc::mbar_wait(&acc_full[buf], (uint32_t)(u & 1));
      __syncwarp();
      tc::tc_fence_after();
      float v[32], vc[32];
      tc::tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(buf * 64), v);
      tc::tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(buf * 64 + 32), vc);
      tc::tmem_ld_wait();
      tc::tc_fence_before();
      tc::mbar_arrive(&acc_empty[buf]);
#pragma unroll
      for (int i = 0; i < 32; ++i) accr[i] += v[i] + vc[i];
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
    };
    for (int j = 0; j < n_tiles; ++j) {
      const int s = j % EF_STAGES;
      const EtTile t = tiles[j];
      if (wt == 0) {
        if (j >= 1) tc::mbar_wait(&a_empty, (uint32_t)((j - 1) & 1));
        tc::mbar_wait(&stage_full[s], (uint32_t)((j / EF_STAGES) & 1));
      }
      et_conv_sync();
      const float* xs = reinterpret_cast<const float*>(stage_base + (size_t)s * EF_STAGE_BYTES);
      const float* wsm = reinterpret_cast<const float*>(stage_base + (size_t)s * EF_STAGE_BYTES + EF_X_BYTES);
      // A: 64 rows x 16 position groups = 1024 chunks
      for (int c = wt; c < 1024; c += ET_NW) {
        const int b = c & 63, vg = c >> 6;
        float x[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = (vg * 8 + i < t.nv) ? xs[(size_t)(vg * 8 + i) * 64 + b] : 0.0f;
        uint4 pl[ET_NS];
        tc::split8<ET_NS>(x, pl);
        const size_t o = (size_t)((b >> 3) * 2048 + vg * 128 + (b & 7) * 16);
#pragma unroll
        for (int q = 0; q < ET_NS; ++q) *reinterpret_cast<uint4*>(a_base + q * EF_A_BYTES + o) = pl[q];
      }
      // B: 32 channels x 16 position groups = 512 chunks
      for (int c = wt; c < 512; c += ET_NW) {
        const int g = c & 31, vg = c >> 5;
        float x[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = (vg * 8 + i < t.nv) ? wsm[(size_t)(vg * 8 + i) * 32 + g] : 0.0f;
        uint4 pl[ET_NS];
        tc::split8<ET_NS>(x, pl);
        const size_t o = (size_t)((g >> 3) * 2048 + vg * 128 + (g & 7) * 16);
#pragma unroll
        for (int q = 0; q < ET_NS; ++q) *reinterpret_cast<uint4*>(b_base + q * EF_B_BYTES + o) = pl[q];
      }
      tc::fence_async_smem();
      tc::mbar_arrive(&a_full);
      tc::mbar_arrive(&stage_empty[s]);
      if (warp >= 12 && j >= 1) drain(j - 1);
    }
    if (warp >= 12 && n_tiles >= 1) drain(n_tiles - 1);
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem_base, 128);
}

// ---------------------------------------------------------------------------------------------------------------------
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
  if (t.n_work > *grid * ET_MAXI) {   // a CTA owns <= 64 work items (<= 256 tiles: items hold <= 512 positions)
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
  enc_fwd_tc_kernel<<<grid, ET_THREADS, EF_SMEM, st>>>(xP, Bh, t.off, t.work, t.n_work, t.slot_first, c->w + t.w_off,
                                                       wsT<float>(c, "enc_part"));
  c->launches += 1;
  return true;
}

// dW of encoder e from the parcel-ordered batch copy; false: the context is marked failed
bool enc_bwd_dw_tc(tnt_ctx* c, int e, const float* xP, int b_begin, int Bh, cudaStream_t st) {
  EncTables& t = c->enc[e];
  int grid;
  if (!et_prepare(c, t, &grid)) return false;
  enc_bwd_dw_tc_kernel<<<grid, ET_THREADS, EB_SMEM, st>>>(xP, b_begin, Bh, c->cfg.R, t.off, t.work, t.n_work,
                                                          wsT<float>(c, "dYpre"), c->g + t.w_off);
  c->launches += 1;
  return true;
}
