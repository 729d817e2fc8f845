// Tensor-core GEMM for every time-parallel product of the path (LSTM input projection, dense_inter,
// dense_out and all their tape gradients: lc_NIC.py:118-159,255,261,387): fp32 in, fp32 out,
// tcgen05.mma kind::f16 on a bf16 operand split (x = hi + mid + lo) with fp32 accumulation in TMEM (tc.cuh).
//
// Two kernels:
//  * tc_pack_kernel  -- splits an fp32 matrix ONCE into NS bf16 planes and writes them in the tensor core's
//    no-swizzle core-matrix layout, K-major, in blocks of 128 (MN) x 64 (K) elements = 16 KB, zero padded:
//        plane[p][k_tile][mn_block][mn_group 16][k_group 8][8 mn][8 k]
//    so an operand tile is one contiguous chunk of global memory.  `transpose` picks which matrix dimension
//    is the contraction dimension, so every GEMM sees K-major operands.  An operand used by several GEMMs
//    (dZ: 4x, every weight: 2x) is converted once instead of once per output tile.
//  * gemm_pk_kernel  -- CTA tile 128 x BN (64 or 128), K step 64, warp-specialised:
//      warp 4  producer: one thread streams the packed tiles with 1-D bulk copies (TMA engine) into an
//              mbarrier ring (expect_tx / complete_tx), no registers involved
//      warp 5  one thread issues the MMAs of a stage as soon as it lands; tcgen05.commit frees the stage and,
//              at the end of an accumulation chain, hands the TMEM buffer to the drain
//      warps 0..3  drain finished chains from TMEM into fp32 registers (lane quadrant = warp), final epilogue
//              (bias / accumulate / split-K fix-up) and the store of C
// Split K (grid.z): partial tiles go to a workspace and the LAST CTA of a tile sums them in split order, so the
// result does not depend on scheduling (no float atomics) unless the caller asks for accumulate == 2.
#include "tnt_internal.h"
#include "tc.cuh"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define TC_BM 128
#define TC_BK 64
#define TC_BLOCK_BYTES (128 * 64 * 2)         // one packed block of one plane: 16 KB
#define TC_STAGE_BYTES(NS, BN) ((NS) * (TC_BLOCK_BYTES + (BN) * TC_BK * 2))
#define TC_STAGES(NS, BN) ((int)((220 * 1024) / TC_STAGE_BYTES(NS, BN)) > 6 ? 6 : (int)((220 * 1024) / TC_STAGE_BYTES(NS, BN)))
#define TC_SMEM_BYTES(NS, BN) (TC_STAGES(NS, BN) * TC_STAGE_BYTES(NS, BN) + 1024)
#define TC_THREADS 192


// ------------------------------------------------------------------------------------------------
// pack: fp32 row-major -> NS bf16 planes in packed K-major blocks
// ------------------------------------------------------------------------------------------------
// Logical operand Op[mn][k]:  transpose == 0: Op[mn][k] = X[mn * ld + k]   (mn < rows, k < cols)
//                             transpose == 1: Op[mn][k] = X[k * ld + mn]   (mn < cols, k < rows)
template <int NS>
__device__ __forceinline__ void
tc_pack_block(const float* __restrict__ X, int rows, int cols, int64_t ld, int transpose, int vec,
              uint8_t* __restrict__ out, int mn_blocks, int64_t plane_stride, int mb, int kt) {
  __shared__ float tile[128][65];
  const int mn0 = mb * 128, k0 = kt * 64;
  const int tid = threadIdx.x;
  if (!transpose) {
    // 128 rows x 64 cols of X; 16 float4 per row
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int e = tid + 256 * i;
      const int r = e >> 4, c4 = (e & 15) * 4;
      const int gr = mn0 + r, gc = k0 + c4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (gr < rows && gc < cols) {
        const float* p = X + (int64_t)gr * ld + gc;
        if (vec && gc + 3 < cols) v = __ldg(reinterpret_cast<const float4*>(p));
        else {
          v.x = __ldg(p);
          if (gc + 1 < cols) v.y = __ldg(p + 1);
          if (gc + 2 < cols) v.z = __ldg(p + 2);
          if (gc + 3 < cols) v.w = __ldg(p + 3);
        }
      }
      tile[r][c4] = v.x; tile[r][c4 + 1] = v.y; tile[r][c4 + 2] = v.z; tile[r][c4 + 3] = v.w;
    }
  } else {
    // 64 rows (k) x 128 cols (mn) of X; 32 float4 per row
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int e = tid + 256 * i;
      const int r = e >> 5, c4 = (e & 31) * 4;
      const int gr = k0 + r, gc = mn0 + c4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (gr < rows && gc < cols) {
        const float* p = X + (int64_t)gr * ld + gc;
        if (vec && gc + 3 < cols) v = __ldg(reinterpret_cast<const float4*>(p));
        else {
          v.x = __ldg(p);
          if (gc + 1 < cols) v.y = __ldg(p + 1);
          if (gc + 2 < cols) v.z = __ldg(p + 2);
          if (gc + 3 < cols) v.w = __ldg(p + 3);
        }
      }
      tile[c4][r] = v.x; tile[c4 + 1][r] = v.y; tile[c4 + 2][r] = v.z; tile[c4 + 3][r] = v.w;
    }
  }
  __syncthreads();
  uint8_t* blk = out + ((int64_t)kt * mn_blocks + mb) * TC_BLOCK_BYTES;
  // 1024 chunks of 16 bytes: chunk q = ((mn_group * 8 + k_group) * 8 + r)
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int q = tid + 256 * i;
    const int r = q & 7, kg = (q >> 3) & 7, mg = q >> 6;
    const float* src = &tile[mg * 8 + r][kg * 8];
    float x8[8];
#pragma unroll
    for (int h = 0; h < 8; ++h) x8[h] = src[h];
    uint4 pl[NS];
    tc::split8<NS>(x8, pl);
#pragma unroll
    for (int p = 0; p < NS; ++p)
      *reinterpret_cast<uint4*>(blk + (int64_t)p * plane_stride + (size_t)q * 16) = pl[p];
  }
}

template <int NS>
__global__ void __launch_bounds__(256)
tc_pack_kernel(const float* __restrict__ X, int rows, int cols, int64_t ld, int transpose, int vec,
               uint8_t* __restrict__ out, int mn_blocks, int64_t plane_stride) {
  tc_pack_block<NS>(X, rows, cols, ld, transpose, vec, out, mn_blocks, plane_stride, blockIdx.x, blockIdx.y);
}

// Several operands in ONE launch: the packs of a GEMM group are 15..256-CTA kernels of ~8 us each, launch-latency
// sized; the operands of a whole group (all final before the group starts) are converted together instead.
#define TC_MAXJOBS 8
struct TcPackJob {
  const float* X; uint8_t* out; int64_t ld, plane_stride;
  int rows, cols, transpose, vec, mn_blocks, cta0;
};
struct TcPackJobs { TcPackJob j[TC_MAXJOBS]; int n; };
template <int NS>
__global__ void __launch_bounds__(256)
tc_pack_multi_kernel(const __grid_constant__ TcPackJobs jobs) {
  int k = 0;
#pragma unroll
  for (int i = 1; i < TC_MAXJOBS; ++i)
    if (i < jobs.n && (int)blockIdx.x >= jobs.j[i].cta0) k = i;
  const TcPackJob& J = jobs.j[k];
  const int local = (int)blockIdx.x - J.cta0;
  tc_pack_block<NS>(J.X, J.rows, J.cols, J.ld, J.transpose, J.vec, J.out, J.mn_blocks, J.plane_stride,
                    local % J.mn_blocks, local / J.mn_blocks);
}

// Accumulation: the tensor core adds into TMEM with truncation, a BIASED rounding whose error grows
// linearly with the number of accumulate steps (measured: 4e-7 at K=64 -> 6e-6 at K=960) and that
// biased error is what breaks parity on gradients with heavy cancellation (BatchNorm backward means).
// So (1) the leading hi*hi products go to their own accumulator ("main") and all cross terms, 2^-8 and
// smaller, to a second one ("corr") whose truncation error is correspondingly smaller, and (2) both are
// drained every TC_DRAIN k-steps into fp32 registers, where the long sum runs with round-to-nearest.
// TMEM: 2 ping-pong buffers x (main BN + corr BN columns).
#define TC_DRAIN 2        // stages (of TC_BK) per TMEM accumulation chain: 8 truncating adds on `main`

// dropout multipliers of elements 4*idx4 .. 4*idx4+3 of a site: one Philox call (common.cuh: drop_mult)
__device__ __forceinline__ void tc_drop4(const DropSite& d, uint64_t idx4, float m[4]) {
  if (d.thr == 0) { m[0] = m[1] = m[2] = m[3] = 1.0f; return; }
  const uint4 r = philox4x32_10((uint32_t)idx4, d.site, d.step, 0u, d.k0, d.k1);
  m[0] = r.x >= d.thr ? d.scale : 0.0f;
  m[1] = r.y >= d.thr ? d.scale : 0.0f;
  m[2] = r.z >= d.thr ? d.scale : 0.0f;
  m[3] = r.w >= d.thr ? d.scale : 0.0f;
}

template <int BN, int NS>
__global__ void __launch_bounds__(TC_THREADS, 1)
gemm_pk_kernel(TcOperand opA, TcOperand opB, int M, int N, float* __restrict__ C, int ldc,
               const float* __restrict__ bias, int accumulate, int kt_per_split, int vecC, TcSplitWs sws,
               const TcEpi epi) {
  constexpr int S = TC_STAGES(NS, BN);
  constexpr uint32_t B_BYTES = BN * TC_BK * 2;
  constexpr uint32_t TMEM_COLS = 4 * BN;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t full_bar[S], empty_bar[S], acc_full[2], acc_empty[2];
  __shared__ uint32_t tmem_slot;
  __shared__ int s_last;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int mt = blockIdx.y, nt = blockIdx.x;
  const int m0 = mt * TC_BM, n0 = nt * BN;
  const int kt_beg = blockIdx.z * kt_per_split;
  const int kt_end = min(opA.k_tiles, kt_beg + kt_per_split);
  const int n_iter = max(0, kt_end - kt_beg);
  const int n_chain = (n_iter + TC_DRAIN - 1) / TC_DRAIN;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < S; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
    tc::mbar_init(&acc_full[0], 1); tc::mbar_init(&acc_full[1], 1);
    tc::mbar_init(&acc_empty[0], 128); tc::mbar_init(&acc_empty[1], 128);
    tc::mbar_fence_init();
  }
  if (warp == 5) tc::tmem_alloc(&tmem_slot, TMEM_COLS);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 4) {
    // ================= producer: bulk copies of packed tiles ==========================================
    if (lane == 0) {
      const uint8_t* a_src = opA.planes + (size_t)mt * TC_BLOCK_BYTES;
      const uint8_t* b_src = opB.planes + (BN == 128 ? (size_t)nt * TC_BLOCK_BYTES
                                                     : (size_t)(nt >> 1) * TC_BLOCK_BYTES + (size_t)(nt & 1) * B_BYTES);
      for (int it = 0; it < n_iter; ++it) {
        const int s = it % S, kt = kt_beg + it;
        tc::mbar_wait(&empty_bar[s], (uint32_t)(((it / S) & 1) ^ 1));             // MMAs that read this slot are done
        uint8_t* st = smem + s * TC_STAGE_BYTES(NS, BN);
        tc::mbar_expect_tx(&full_bar[s], NS * (TC_BLOCK_BYTES + B_BYTES));
#pragma unroll
        for (int p = 0; p < NS; ++p) {
          tc::bulk_g2s(st + p * TC_BLOCK_BYTES, a_src + (size_t)p * opA.plane_stride + (size_t)kt * opA.mn_blocks * TC_BLOCK_BYTES,
                       TC_BLOCK_BYTES, &full_bar[s]);
          tc::bulk_g2s(st + NS * TC_BLOCK_BYTES + p * B_BYTES,
                       b_src + (size_t)p * opB.plane_stride + (size_t)kt * opB.mn_blocks * TC_BLOCK_BYTES, B_BYTES, &full_bar[s]);
        }
      }
    }
  } else if (warp == 5) {
    // ================= MMA issuer ===================================================================
    if (lane == 0) {
      constexpr uint32_t idesc = tc::idesc_bf16(TC_BM, BN, 0, 0);       // both operands K-major
      const uint32_t lbo = 128u, sbo = 1024u;
      for (int it = 0; it < n_iter; ++it) {
        const int s = it % S;
        const int g = it / TC_DRAIN, tb = g & 1;
        const bool grp_first = (it % TC_DRAIN) == 0;
        const bool grp_last = (it % TC_DRAIN) == TC_DRAIN - 1 || it + 1 == n_iter;
        if (grp_first) tc::mbar_wait(&acc_empty[tb], (uint32_t)(((g >> 1) & 1) ^ 1));   // the drain freed this buffer
        tc::mbar_wait(&full_bar[s], (uint32_t)((it / S) & 1));
        tc::tc_fence_after();
        const uint32_t a0 = tc::smem_u32(smem + s * TC_STAGE_BYTES(NS, BN)), b0 = a0 + NS * TC_BLOCK_BYTES;
        const uint32_t d_main = tmem_base + (uint32_t)(tb * 2 * BN), d_corr = d_main + BN;
#pragma unroll
        for (int j = 0; j < TC_BK / 16; ++j) {
          uint64_t da[NS], db[NS];
#pragma unroll
          for (int p = 0; p < NS; ++p) {
            da[p] = tc::smem_desc(a0 + p * TC_BLOCK_BYTES + j * 256u, lbo, sbo);
            db[p] = tc::smem_desc(b0 + p * B_BYTES + j * 256u, lbo, sbo);
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
        tc::mma_commit(&empty_bar[s]);
        if (grp_last) tc::mma_commit(&acc_full[tb]);
      }
    }
  } else {
    // ================= drain + epilogue (warp = TMEM lane quadrant) ====================================
    float acc[BN];
#pragma unroll
    for (int i = 0; i < BN; ++i) acc[i] = 0.0f;
    for (int g = 0; g < n_chain; ++g) {
      const int tb = g & 1;
      tc::mbar_wait(&acc_full[tb], (uint32_t)((g >> 1) & 1));
      tc::tc_fence_after();
      const uint32_t ta = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(tb * 2 * BN);
      float vm[32];
#pragma unroll
      for (int h = (NS > 1 ? 0 : BN / 32); h < 2 * (BN / 32); ++h) {   // corr first (small terms), then main
        const int c32 = h % (BN / 32);
        tc::tmem_ld32(ta + c32 * 32 + (h < BN / 32 ? BN : 0), vm);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) acc[c32 * 32 + i] += vm[i];
      }
      tc::tc_fence_before();
      tc::mbar_arrive(&acc_empty[tb]);
    }
    const int row = warp * 32 + lane, gm = m0 + row;
    const bool atomic_split = accumulate == 2;
    bool do_store = true;
    if (gridDim.z > 1 && !atomic_split) {
      // deterministic split K: park the partial tile, the last CTA of the tile sums all of them in split order
      const int tile = blockIdx.y * gridDim.x + blockIdx.x, ntile = gridDim.x * gridDim.y;
      float* mine = sws.part + ((size_t)blockIdx.z * ntile + tile) * (TC_BM * BN);
#pragma unroll
      for (int i = 0; i < BN; ++i) __stcg(mine + i * TC_BM + row, acc[i]);
      __threadfence();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (tid == 0) {
        const unsigned old = atomicAdd(sws.count + tile, 1u);
        s_last = (old == gridDim.z - 1) ? 1 : 0;
        if (s_last) sws.count[tile] = 0u;
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      do_store = s_last != 0;
      if (do_store) {
        __threadfence();
#pragma unroll
        for (int i = 0; i < BN; ++i) acc[i] = 0.0f;
        for (unsigned z = 0; z < gridDim.z; ++z) {
          const float* src = sws.part + ((size_t)z * ntile + tile) * (TC_BM * BN);
#pragma unroll
          for (int i = 0; i < BN; ++i) acc[i] += __ldcg(src + i * TC_BM + row);
        }
      }
    }
    // fp32 tensors go through a warp-private shared-memory tile ([32 rows][BN + 4], the operand ring is idle now) so
    // that global accesses are whole 512-byte row segments instead of one 16-byte piece per lane and row
    float* stg = reinterpret_cast<float*>(smem) + (size_t)warp * 32 * (BN + 4);
    const bool vec = vecC && (n0 + BN <= N);
    auto tile_out = [&](float* dst) {
#pragma unroll
      for (int i = 0; i < BN; i += 4)
        *reinterpret_cast<float4*>(stg + lane * (BN + 4) + i) = make_float4(acc[i], acc[i + 1], acc[i + 2], acc[i + 3]);
      __syncwarp();
      for (int r = 0; r < 32; ++r) {
        const int gr = m0 + warp * 32 + r;
        if (gr >= M) break;
        for (int c4 = lane * 4; c4 < BN; c4 += 128) {
          const float4 v = *reinterpret_cast<const float4*>(stg + r * (BN + 4) + c4);
          float* o = dst + (int64_t)gr * ldc + n0 + c4;
          if (vec) *reinterpret_cast<float4*>(o) = v;
          else {
            if (n0 + c4 < N) o[0] = v.x;
            if (n0 + c4 + 1 < N) o[1] = v.y;
            if (n0 + c4 + 2 < N) o[2] = v.z;
            if (n0 + c4 + 3 < N) o[3] = v.w;
          }
        }
      }
      __syncwarp();
    };
    if (do_store && epi.mode != TC_EPI_NONE) {
      // Fused point-wise epilogue of the dense_inter layer (lc_NIC.py:139-148,261): the product leaves the kernel as
      // the fp32 tensors the backward pass needs AND as the packed bf16-plane A operand of the GEMM that consumes it,
      // so no activation kernel and no pack kernel sit between two GEMMs of the critical path.
      // The point-wise part runs as a compact loop over the warp's 32 rows with lanes along the columns (whole 512-byte
      // row segments per access), not as straight-line code over the thread's own 128 accumulators: unrolled, the 32
      // Philox calls alone are ~30 k instructions executed once -- instruction-fetch bound (measured: +57 us).
#pragma unroll
      for (int i = 0; i < BN; i += 4)
        *reinterpret_cast<float4*>(stg + lane * (BN + 4) + i) = make_float4(acc[i], acc[i + 1], acc[i + 2], acc[i + 3]);
      __syncwarp();
      const int c4 = lane * 4;
      const bool lane_on = c4 < BN;
      const int nc = n0 + c4;
      float bias4[4] = {0.f, 0.f, 0.f, 0.f};
      if (lane_on && bias != nullptr) {
#pragma unroll
        for (int q = 0; q < 4; ++q) if (nc + q < N) bias4[q] = __ldg(&bias[nc + q]);
      }
      // four rows per trip: with one warp per scheduler every instruction pays its full latency, so the four
      // independent Philox chains are what fills the pipeline (measured: 13.6 cycles per issued instruction unrolled 1)
#pragma unroll 4
      for (int r = 0; r < 32; ++r) {
        const int gr = m0 + warp * 32 + r;
        if (!lane_on) continue;
        float* sp = stg + r * (BN + 4) + c4;
        float4 v = *reinterpret_cast<const float4*>(sp);
        float o[4] = {0.f, 0.f, 0.f, 0.f};
        if (gr < M) {
          const int t = gr / epi.B, b = gr % epi.B;
          const uint64_t e0 = ((uint64_t)b * epi.T + t) * (uint64_t)N + (uint64_t)nc;
          float m4[4];
          if ((N & 3) == 0) tc_drop4(epi.drop, e0 >> 2, m4);                           // nc is a multiple of 4: one Philox call
          else {
#pragma unroll
            for (int q = 0; q < 4; ++q) m4[q] = drop_mult(epi.drop, e0 + q);
          }
          const float x[4] = {v.x, v.y, v.z, v.w};
          float* crow = C + (int64_t)gr * ldc + nc;
          {
            float pre[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) { pre[q] = x[q] + bias4[q]; o[q] = (nc + q < N) ? lrelu(pre[q]) * m4[q] : 0.0f; }
            float* drow = epi.C2 + (int64_t)gr * ldc + nc;
            if (vecC && nc + 3 < N) {
              *reinterpret_cast<float4*>(crow) = make_float4(pre[0], pre[1], pre[2], pre[3]);   // inter_pre
              *reinterpret_cast<float4*>(drow) = make_float4(o[0], o[1], o[2], o[3]);           // inter_d
            } else {
#pragma unroll
              for (int q = 0; q < 4; ++q) if (nc + q < N) { crow[q] = pre[q]; drow[q] = o[q]; }
            }
          }
        }
        *reinterpret_cast<float4*>(sp) = make_float4(o[0], o[1], o[2], o[3]);
      }
      __syncwarp();
#pragma unroll
      for (int i = 0; i < BN; i += 4) {
        const float4 v = *reinterpret_cast<const float4*>(stg + lane * (BN + 4) + i);
        acc[i] = v.x; acc[i + 1] = v.y; acc[i + 2] = v.z; acc[i + 3] = v.w;
      }
      // packed A operand [row][n], K-major over n: rows beyond M and columns beyond N are written as zeros
#pragma unroll
      for (int g8 = 0; g8 < BN / 8; ++g8) {
        const int nn = n0 + 8 * g8, kt = nn >> 6, kg = (nn & 63) >> 3;
        if (kt < epi.pk_kt) {
          uint8_t* blk = epi.pk + ((size_t)kt * epi.pk_mnb + mt) * TC_BLOCK_BYTES +
                         (size_t)((((row >> 3) * 8 + kg) * 8 + (row & 7)) * 16);
          uint4 pl[NS];
          tc::split8<NS>(acc + 8 * g8, pl);
#pragma unroll
          for (int q = 0; q < NS; ++q) *reinterpret_cast<uint4*>(blk + (size_t)q * epi.pk_stride) = pl[q];
        }
      }
    } else if (do_store && accumulate == 0) {
      if (bias != nullptr) {
#pragma unroll
        for (int i = 0; i < BN; ++i)
          if (n0 + i < N) acc[i] += __ldg(&bias[n0 + i]);
      }
      tile_out(C);
    } else if (do_store && gm < M) {
      float* crow = C + (int64_t)gm * ldc;
      if (bias != nullptr && (blockIdx.z == 0 || !atomic_split)) {
#pragma unroll
        for (int i = 0; i < BN; ++i)
          if (n0 + i < N) acc[i] += __ldg(&bias[n0 + i]);
      }
      if (atomic_split) {
#pragma unroll
        for (int i = 0; i < BN; ++i)
          if (n0 + i < N) atomicAdd(&crow[n0 + i], acc[i]);
      } else if (vecC && n0 + BN - 1 < N) {
#pragma unroll
        for (int i = 0; i < BN; i += 4) {
          float4 o = make_float4(acc[i], acc[i + 1], acc[i + 2], acc[i + 3]);
          float4* p = reinterpret_cast<float4*>(&crow[n0 + i]);
          if (accumulate) { float4 c = *p; o.x += c.x; o.y += c.y; o.z += c.z; o.w += c.w; }
          *p = o;
        }
      } else {
#pragma unroll
        for (int i = 0; i < BN; ++i)
          if (n0 + i < N) crow[n0 + i] = accumulate ? crow[n0 + i] + acc[i] : acc[i];
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 5) tc::tmem_dealloc(tmem_base, TMEM_COLS);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
// Arena for packed operands: a stack per context (TcState in tnt_internal.h).  tc_arena_reset() at the start of a
// phase; tc_pack() pushes and the operand stays valid until the next reset; gemm_tc() pops what it pushed for operands
// it does not cache.  Nothing here is shared between contexts: two models in one process (a training and a decode
// model, the tests' twins) each own their arena, operand cache, lane flag and split-K workspaces.
#define TC_WS_FLOATS ((size_t)320 * 128 * 64)

void tc_set_lane(tnt_ctx* c, int lane) { c->tc.lane = lane ? 1 : 0; }

void tc_arena_reserve(tnt_ctx* c, size_t bytes) {
  // whole MiB: lane 1 allocates downwards from the END of the arena, and packed tiles are read with 16-byte bulk
  // copies -- a capacity that is not a multiple of 16 would misalign every side-stream operand
  TcState& t = c->tc;
  bytes = (bytes + ((size_t)1 << 20) - 1) & ~(((size_t)1 << 20) - 1);
  if (bytes <= t.cap) return;
  t.want = std::max(t.want, bytes);
  if (t.arena) { cudaDeviceSynchronize(); cudaFree(t.arena); t.arena = nullptr; t.cap = 0; t.off = 0; t.cache.clear(); }
}
void tc_arena_reset(tnt_ctx* c) { c->tc.off = 0; c->tc.top = c->tc.cap; c->tc.cache.clear(); }

void tc_destroy(tnt_ctx* c) {
  TcState& t = c->tc;
  if (t.arena) cudaFree(t.arena);
  for (int i = 0; i < 2; ++i) { if (t.ws[i].part) cudaFree(t.ws[i].part); if (t.ws[i].count) cudaFree(t.ws[i].count); }
  t = TcState();
}

static bool tc_ready(tnt_ctx* c) {
  TcState& t = c->tc;
  if (!t.arena) {
    if (cudaMalloc(&t.arena, t.want) != cudaSuccess) { cudaGetLastError(); t.arena = nullptr; return false; }
    t.cap = t.want; t.off = 0; t.top = t.cap;
  }
  for (int i = 0; i < 2; ++i)
    if (!t.ws[i].part) {
      if (cudaMalloc(&t.ws[i].part, TC_WS_FLOATS * sizeof(float)) != cudaSuccess ||
          cudaMalloc(&t.ws[i].count, 4096 * sizeof(unsigned)) != cudaSuccess) {
        cudaGetLastError(); t.ws[i].part = nullptr; return false;
      }
      cudaMemset(t.ws[i].count, 0, 4096 * sizeof(unsigned));
    }
  return true;
}

// planes per operand: TNT_FP32 -> 3 (six MMAs, fp32-equivalent), TNT_BF16X3 -> 2 (three MMAs), TNT_BF16 -> 1
static int tc_ns(tnt_ctx* c) {
  return c->cfg.precision == TNT_BF16X3 ? 2 : (c->cfg.precision == TNT_BF16 ? 1 : 3);
}

size_t tc_packed_bytes(int64_t mn, int64_t k, int ns) {
  return (size_t)ns * (size_t)((mn + 127) / 128) * (size_t)((k + 63) / 64) * TC_BLOCK_BYTES;
}

static uint8_t* tc_alloc(tnt_ctx* c, size_t bytes) {
  TcState& t = c->tc;
  const size_t rounded = (bytes + 1023) & ~(size_t)1023;
  if (t.off + rounded > t.top) return nullptr;
  if (t.lane == 0) { uint8_t* p = t.arena + t.off; t.off += rounded; return p; }
  t.top -= rounded;
  return t.arena + t.top;
}

template <int NS> static void launch_pack(dim3 grid, cudaStream_t st, const float* X, int rows, int cols, int ld, int tr,
                                          int vec, uint8_t* out, int mnb, int64_t ps) {
  tc_pack_kernel<NS><<<grid, 256, 0, st>>>(X, rows, cols, ld, tr, vec, out, mnb, ps);
}

// Pack X (rows x cols, row stride ld) as an operand whose contraction dimension is `cols` (transpose = false)
// or `rows` (transpose = true).  planes == nullptr on failure (arena exhausted / no memory).
TcOperand tc_pack(tnt_ctx* c, const float* X, int rows, int cols, int ld, bool transpose, cudaStream_t st) {
  TcOperand op; op.planes = nullptr; op.plane_stride = 0; op.mn_blocks = 0; op.k_tiles = 0;
  if (!tc_ready(c)) { c->err = "tc_pack: cannot allocate the operand arena"; return op; }
  const int ns = tc_ns(c);
  const int mn = transpose ? cols : rows, k = transpose ? rows : cols;
  op.mn_blocks = (mn + 127) / 128; op.k_tiles = (k + 63) / 64;
  op.plane_stride = (int64_t)op.mn_blocks * op.k_tiles * TC_BLOCK_BYTES;
  uint8_t* out = tc_alloc(c, (size_t)ns * op.plane_stride);
  if (!out) { c->err = "tc_pack: operand arena exhausted"; return op; }
  const int vec = ((((uintptr_t)X) & 15) == 0 && (ld & 3) == 0) ? 1 : 0;
  dim3 grid(op.mn_blocks, op.k_tiles);
  if (ns == 1) launch_pack<1>(grid, st, X, rows, cols, ld, transpose ? 1 : 0, vec, out, op.mn_blocks, op.plane_stride);
  else if (ns == 2) launch_pack<2>(grid, st, X, rows, cols, ld, transpose ? 1 : 0, vec, out, op.mn_blocks, op.plane_stride);
  else launch_pack<3>(grid, st, X, rows, cols, ld, transpose ? 1 : 0, vec, out, op.mn_blocks, op.plane_stride);
  op.planes = out;
  c->launches += 1;
  return op;
}

static const TcCacheEntry* tc_cache_find(tnt_ctx* c, const float* X, int rows, int cols, int ld, int tr, int ns) {
  for (const TcCacheEntry& e : c->tc.cache)
    if (e.X == X && e.rows == rows && e.cols == cols && e.ld == ld && e.transpose == tr && e.ns == ns) return &e;
  return nullptr;
}

// An operand a kernel writes in packed form itself (head_fused.cu): arena space + a cache entry under the key of the
// fp32 matrix it stands for, so the GEMMs that consume it find it with tc_pack_cached.  planes == nullptr: no room.
TcOperand tc_reserve_operand(tnt_ctx* c, const float* key, int rows, int cols, int ld, bool transpose) {
  TcOperand op; op.planes = nullptr; op.plane_stride = 0; op.mn_blocks = 0; op.k_tiles = 0;
  if (!tc_ready(c)) { c->err = "tc_reserve_operand: cannot allocate the operand arena"; return op; }
  const int ns = tc_ns(c), tr = transpose ? 1 : 0;
  const int mn = transpose ? cols : rows, k = transpose ? rows : cols;
  op.mn_blocks = (mn + 127) / 128; op.k_tiles = (k + 63) / 64;
  op.plane_stride = (int64_t)op.mn_blocks * op.k_tiles * TC_BLOCK_BYTES;
  uint8_t* out = tc_alloc(c, (size_t)ns * op.plane_stride);
  if (!out) { c->err = "tc_reserve_operand: operand arena exhausted"; return op; }
  op.planes = out;
  TcCacheEntry e;
  e.X = key; e.rows = rows; e.cols = cols; e.ld = ld; e.transpose = tr; e.ns = ns; e.op = op;
  c->tc.cache.push_back(e);
  return op;
}
int tc_planes(tnt_ctx* c) { return tc_ns(c); }

// Same as tc_pack, but remembers the packed operand until the next tc_arena_reset(): only for matrices that are not
// rewritten inside a phase (weights, and activations once their producer has run).
TcOperand tc_pack_cached(tnt_ctx* c, const float* X, int rows, int cols, int ld, bool transpose, cudaStream_t st) {
  const int ns = tc_ns(c), tr = transpose ? 1 : 0;
  if (const TcCacheEntry* hit = tc_cache_find(c, X, rows, cols, ld, tr, ns)) return hit->op;
  TcCacheEntry e;
  e.X = X; e.rows = rows; e.cols = cols; e.ld = ld; e.transpose = tr; e.ns = ns;
  e.op = tc_pack(c, X, rows, cols, ld, transpose, st);
  if (e.op.planes) c->tc.cache.push_back(e);
  return e.op;
}

// Packs every listed operand that is not cached yet with ONE launch and remembers them until the next
// tc_arena_reset(): the GEMMs that follow find their operands with tc_pack_cached.  Same rule as tc_pack_cached: only
// matrices that are final when the call is enqueued and not rewritten before the reset.
void tc_prepack(tnt_ctx* c, const TcPackReq* reqs, int n, cudaStream_t st) {
  if (n <= 0 || !tc_ready(c)) return;
  const int ns = tc_ns(c);
  TcPackJobs jobs; jobs.n = 0;
  TcCacheEntry ent[TC_MAXJOBS];
  int ctas = 0;
  for (int i = 0; i < n && jobs.n < TC_MAXJOBS; ++i) {
    const TcPackReq& r = reqs[i];
    const int tr = r.transpose ? 1 : 0;
    bool have = tc_cache_find(c, r.X, r.rows, r.cols, r.ld, tr, ns) != nullptr;
    for (int q = 0; q < jobs.n; ++q)
      if (ent[q].X == r.X && ent[q].rows == r.rows && ent[q].cols == r.cols && ent[q].ld == r.ld && ent[q].transpose == tr) have = true;
    if (have || r.rows <= 0 || r.cols <= 0) continue;
    const int mn = tr ? r.cols : r.rows, k = tr ? r.rows : r.cols;
    TcOperand op;
    op.mn_blocks = (mn + 127) / 128; op.k_tiles = (k + 63) / 64;
    op.plane_stride = (int64_t)op.mn_blocks * op.k_tiles * TC_BLOCK_BYTES;
    uint8_t* out = tc_alloc(c, (size_t)ns * op.plane_stride);
    if (!out) break;                                          // no room: the GEMM's own tc_pack reports the error
    op.planes = out;
    TcPackJob& J = jobs.j[jobs.n];
    J.X = r.X; J.out = out; J.ld = r.ld; J.plane_stride = op.plane_stride;
    J.rows = r.rows; J.cols = r.cols; J.transpose = tr;
    J.vec = ((((uintptr_t)r.X) & 15) == 0 && (r.ld & 3) == 0) ? 1 : 0;
    J.mn_blocks = op.mn_blocks; J.cta0 = ctas;
    ctas += op.mn_blocks * op.k_tiles;
    TcCacheEntry& e = ent[jobs.n];
    e.X = r.X; e.rows = r.rows; e.cols = r.cols; e.ld = r.ld; e.transpose = tr; e.ns = ns; e.op = op;
    jobs.n += 1;
  }
  if (jobs.n == 0) return;
  if (ns == 1) tc_pack_multi_kernel<1><<<ctas, 256, 0, st>>>(jobs);
  else if (ns == 2) tc_pack_multi_kernel<2><<<ctas, 256, 0, st>>>(jobs);
  else tc_pack_multi_kernel<3><<<ctas, 256, 0, st>>>(jobs);
  for (int q = 0; q < jobs.n; ++q) c->tc.cache.push_back(ent[q]);
  c->launches += 1;
}

// C[M,N] = A * B^T-in-packed-form (+ bias) (+ C): A packed with MN = M, B packed with MN = N, same K.
void gemm_packed(tnt_ctx* c, const TcOperand& A, const TcOperand& B, int M, int N, float* C, int ldc, const float* bias,
                 int accumulate, cudaStream_t st, const TcEpi* epi_in) {
  TcEpi epi;
  if (epi_in) epi = *epi_in; else { epi = TcEpi(); epi.mode = TC_EPI_NONE; }
  if (M <= 0 || N <= 0 || !A.planes || !B.planes || A.k_tiles != B.k_tiles || !tc_ready(c)) {
    if (c->err.empty()) c->err = "gemm_packed: bad operands";
    return;
  }
  if (!c->tc.attr_done) {
#define SETATTR(BN_, NS_) \
  cudaFuncSetAttribute(gemm_pk_kernel<BN_, NS_>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES(NS_, BN_))
    SETATTR(64, 1); SETATTR(128, 1); SETATTR(64, 2); SETATTR(128, 2); SETATTR(64, 3); SETATTR(128, 3);
#undef SETATTR
    c->tc.attr_done = 1;
  }
  const int ns = tc_ns(c);
  const int bn = N > 64 ? 128 : 64;
  const int tm = (M + TC_BM - 1) / TC_BM, tn = (N + bn - 1) / bn, tiles = tm * tn;
  const int iters = A.k_tiles;
  const int sms = c->sm_count;
  int splits = 1;
  if (accumulate != 2) {
    // fill the SMs, keep >= 2 k-steps per CTA, stay inside the workspace
    splits = std::max(1, std::min(sms / tiles, iters / 2));
    while (splits > 1 && ((size_t)tiles * splits * TC_BM * bn > TC_WS_FLOATS || tiles > 4096)) --splits;
  }
  int kps = (iters + splits - 1) / splits;
  splits = (iters + kps - 1) / kps;
  dim3 grid(tn, tm, splits);
  const int vC = ((((uintptr_t)C) & 15) == 0 && (ldc & 3) == 0) ? 1 : 0;
  const TcSplitWs sws = c->tc.ws[c->tc.lane];
#define LAUNCH(BN_, NS_) \
  gemm_pk_kernel<BN_, NS_><<<grid, TC_THREADS, TC_SMEM_BYTES(NS_, BN_), st>>>(A, B, M, N, C, ldc, bias, accumulate, kps, vC, sws, epi)
  if (bn == 128) { if (ns == 1) LAUNCH(128, 1); else if (ns == 2) LAUNCH(128, 2); else LAUNCH(128, 3); }
  else { if (ns == 1) LAUNCH(64, 1); else if (ns == 2) LAUNCH(64, 2); else LAUNCH(64, 3); }
#undef LAUNCH
  c->launches += 1;
}

// Row-major fp32 operands (TA: A stored [K,M]; TB: B stored [N,K]): packs both operands, multiplies, releases the arena
// space of the operands that are not cached.
void gemm_tc(tnt_ctx* c, int cacheA, int cacheB, bool ta, bool tb, int M, int N, int K, const float* A, int lda,
             const float* B, int ldb, float* C, int ldc, const float* bias, int accumulate, cudaStream_t st) {
  if (M <= 0 || N <= 0 || K <= 0) return;
  if (!tc_ready(c)) { c->err = "gemm_tc: cannot allocate workspaces"; return; }
  // A: logical [M][K]; stored [M][K] (ta = false) or [K][M] (ta = true)
  // B: logical [N][K]; stored [N][K] (tb = true) or [K][N] (tb = false)
  TcOperand a, b;
  if (cacheA) a = ta ? tc_pack_cached(c, A, K, M, lda, true, st) : tc_pack_cached(c, A, M, K, lda, false, st);
  if (cacheB) b = tb ? tc_pack_cached(c, B, N, K, ldb, false, st) : tc_pack_cached(c, B, K, N, ldb, true, st);
  const size_t mark = c->tc.off;
  if (!cacheA) a = ta ? tc_pack(c, A, K, M, lda, true, st) : tc_pack(c, A, M, K, lda, false, st);
  if (!cacheB) b = tb ? tc_pack(c, B, N, K, ldb, false, st) : tc_pack(c, B, K, N, ldb, true, st);
  gemm_packed(c, a, b, M, N, C, ldc, bias, accumulate, st);
  if (c->tc.lane == 0) c->tc.off = mark;
}

// dense_inter with its point-wise tail in the GEMM epilogue (TcEpi): C[M,N] = A[M,K] . W[K,N] where the stored result is
//   C = pre-activation (+ bias), C2 = dropout(LeakyReLU(pre)), and the packed operand of C2
// The packed operand is cached under (pk_key, M, N, ldc, not transposed): the next GEMM finds it with tc_pack_cached.
// Atb: A is given as a row-major [M,K] matrix (cached or packed here); tb: W stored [N,K].
void gemm_tc_epi(tnt_ctx* c, int cacheA, bool tb, int M, int N, int K, const float* A, int lda, const float* W, int ldw,
                 float* C, int ldc, const float* bias, int mode, float* C2, const float* pre, const DropSite& drop, int Bsz,
                 int T, const float* pk_key, cudaStream_t st) {
  if (M <= 0 || N <= 0 || K <= 0) return;
  if (!tc_ready(c)) { c->err = "gemm_tc_epi: cannot allocate workspaces"; return; }
  TcOperand b = tb ? tc_pack_cached(c, W, N, K, ldw, false, st) : tc_pack_cached(c, W, K, N, ldw, true, st);
  TcOperand a = cacheA ? tc_pack_cached(c, A, M, K, lda, false, st) : TcOperand();
  TcOperand pk = tc_reserve_operand(c, pk_key, M, N, ldc, false);        // before any pop-able pack: it must survive
  const size_t mark = c->tc.off;
  if (!cacheA) a = tc_pack(c, A, M, K, lda, false, st);
  if (!pk.planes || !a.planes || !b.planes) { ctx_fail(c, TNT_ERR_CUDA, "gemm_tc_epi: operand arena exhausted"); return; }
  TcEpi epi;
  epi.mode = mode; epi.C2 = C2; (void)pre; epi.drop = drop; epi.B = Bsz; epi.T = T;
  epi.pk = const_cast<uint8_t*>(pk.planes); epi.pk_stride = pk.plane_stride; epi.pk_mnb = pk.mn_blocks; epi.pk_kt = pk.k_tiles;
  gemm_packed(c, a, b, M, N, C, ldc, bias, 0, st, &epi);
  if (c->tc.lane == 0) c->tc.off = mark;
}
