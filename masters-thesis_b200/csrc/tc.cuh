// sm_100a tensor-core plumbing for libtnt: tcgen05.mma with accumulators in TMEM, mbarrier completion,
// shared-memory matrix descriptors for the NO-SWIZZLE canonical layouts, and the bf16x3 operand split.
//
// Why bf16x3: the reference computes in fp32 and the parity bar is 1e-3 relative on every gradient.
// One bf16 (or tf32) pass misses that bar on the decoder gradients (7e-3 / 2e-3); writing each fp32
// operand as hi + lo with hi = bf16(x), lo = bf16(x - hi) and issuing hi*hi + hi*lo + lo*hi keeps 16
// mantissa bits (~4e-5) while all three products run on the tensor pipe with fp32 accumulation in TMEM.
//
// Shared-memory operand tile ("core matrix" = 8 x 16 bytes, stored as 128 contiguous bytes):
//     tile[outer_group][inner_group][8 outer][8 inner elements (bf16)]
//   inner = the dimension that is contiguous in global memory, outer = the strided one.
//   K-major operand  (global [MN][K], K contiguous): outer = MN, inner = K
//        descriptor: LBO = 128 (next core matrix along K), SBO = (K_tile/8)*128 (next 8 rows of MN)
//   MN-major operand (global [K][MN], MN contiguous): outer = K, inner = MN
//        descriptor: SBO = 128 (next 8 elements of MN), LBO = (MN_tile/8)*128 (next 8 rows of K)
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier ---------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
}

// transaction barrier: expect `bytes` from bulk copies, count one arrival
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
// 1-D bulk copy global -> shared (TMA engine, no tensor map); completes on `bar` with complete_tx
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// orders generic-proxy accesses (plain ld/st, any state space) against async-proxy ones (bulk copies, MMA reads)
__device__ __forceinline__ void fence_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

// ---- grid-wide counters (cooperative launch: every CTA is resident) -------------------------------
__device__ __forceinline__ unsigned ld_acquire(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release(unsigned* p, unsigned v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void spin_until(const unsigned* p, unsigned target) {
  while (ld_acquire(p) < target) { }
}

// generic-proxy smem writes -> visible to the async proxy (tensor core reads of smem operands)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// ---- TMEM -------------------------------------------------------------------------------------
// one full warp; ncols power of two >= 32; the base address (lane<<16 | column) lands in *slot
__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(ncols) : "memory");
}
// 32 lanes x 32 consecutive columns: thread `lane` of the warp receives row (lane base + lane)
__device__ __forceinline__ void tmem_ld32(uint32_t addr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(addr)
      : "memory");
}
// 32 lanes x 16 consecutive columns
__device__ __forceinline__ void tmem_ld16(uint32_t addr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(addr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// 32 lanes x 32 consecutive columns back into tensor memory (thread `lane` writes row lane base + lane)
__device__ __forceinline__ void tmem_st32(uint32_t addr, const float* v) {
  const uint32_t* r = reinterpret_cast<const uint32_t*>(v);
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(addr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]),
        "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]),
        "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ---- descriptors --------------------------------------------------------------------------------
// Shared-memory matrix descriptor, SWIZZLE_NONE, version 1 (Blackwell).  Offsets in bytes.
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;                        // descriptor version
  return d;
}
// Instruction descriptor for kind::f16 with bf16 operands and an fp32 accumulator.
__host__ __device__ constexpr uint32_t idesc_bf16(int M, int N, int a_mn_major, int b_mn_major) {
  return (1u << 4)                               // D format: f32
         | (1u << 7) | (1u << 10)                // A, B format: bf16
         | ((uint32_t)(a_mn_major & 1) << 15) | ((uint32_t)(b_mn_major & 1) << 16)
         | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// SWIZZLE_128B variant: the operand is exactly what a tensor-map TMA with CU_TENSOR_MAP_SWIZZLE_128B lands (rows of 128
// bytes, 16-byte chunk index XOR row % 8; 1024-byte aligned atoms of 8 rows).  K-major only: measured with
// tools/tf32_probe.cu, kind::tf32 does not honour MN-major operands (profiles/r2v_tf32_probe.txt).
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr, uint32_t sbo_bytes) {
  return smem_desc(saddr, 16, sbo_bytes) | ((uint64_t)2 << 61);
}
// Instruction descriptor for kind::tf32 (fp32 words in shared memory, the low 13 mantissa bits are ignored -- measured:
// truncation), fp32 accumulator; K = 8 per instruction.
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// fp32 -> nearest tf32 (10 explicit mantissa bits), still an fp32 word
__device__ __forceinline__ float tf32_rn(float x) {
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
}
// 2-D tensor-map TMA load (box of the map at column c0, row c1) into shared memory, completion on an mbarrier
__device__ __forceinline__ void tma_2d(void* dst_smem, const void* tensor_map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
          smem_u32(dst_smem)),
      "l"(reinterpret_cast<uint64_t>(tensor_map)), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ float4 lds128(uint32_t saddr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(saddr));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t saddr, const uint4& v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void sts128f(uint32_t saddr, const float4& v) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]; one thread issues on behalf of the CTA
__device__ __forceinline__ void mma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// mbarrier arrives once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

// ---- bf16x3 split -------------------------------------------------------------------------------
// four fp32 -> four bf16 "hi" and four bf16 "lo" (x ~= hi + lo), packed as 2 x 32-bit words each
__device__ __forceinline__ void split4(const float4& x, uint2& hi, uint2& lo) {
  __nv_bfloat16 h0 = __float2bfloat16_rn(x.x), h1 = __float2bfloat16_rn(x.y);
  __nv_bfloat16 h2 = __float2bfloat16_rn(x.z), h3 = __float2bfloat16_rn(x.w);
  __nv_bfloat16 l0 = __float2bfloat16_rn(x.x - __bfloat162float(h0));
  __nv_bfloat16 l1 = __float2bfloat16_rn(x.y - __bfloat162float(h1));
  __nv_bfloat16 l2 = __float2bfloat16_rn(x.z - __bfloat162float(h2));
  __nv_bfloat16 l3 = __float2bfloat16_rn(x.w - __bfloat162float(h3));
  hi.x = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
  hi.y = (uint32_t)__bfloat16_as_ushort(h2) | ((uint32_t)__bfloat16_as_ushort(h3) << 16);
  lo.x = (uint32_t)__bfloat16_as_ushort(l0) | ((uint32_t)__bfloat16_as_ushort(l1) << 16);
  lo.y = (uint32_t)__bfloat16_as_ushort(l2) | ((uint32_t)__bfloat16_as_ushort(l3) << 16);
}

// three-way split: x ~= hi + mid + lo carries 24 mantissa bits (fp32-equivalent operands); with the six
// products hh, hm, mh, hl, lh, mm the dropped terms are ~2^-27 relative
__device__ __forceinline__ uint32_t pack_bf16x2(__nv_bfloat16 a, __nv_bfloat16 b) {
  return (uint32_t)__bfloat16_as_ushort(a) | ((uint32_t)__bfloat16_as_ushort(b) << 16);
}
// Eight fp32 values -> NS planes of eight bf16 (one 16-byte operand chunk per plane): x ~= p0 + p1 + ... with each plane the
// bf16 rounding of what the previous planes left.  Two elements per F2FP (cvt.rn.bf16x2.f32); the bf16 -> fp32 way back is a
// shift / mask of the packed word.
template <int NS>
__device__ __forceinline__ void split8(const float* x, uint4* out) {
  float r[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = x[i];
#pragma unroll
  for (int p = 0; p < NS; ++p) {
    uint32_t w[4];
#pragma unroll
    for (int h = 0; h < 4; ++h) {
      const __nv_bfloat162 pk = __floats2bfloat162_rn(r[2 * h], r[2 * h + 1]);     // .x (low half) = first element
      w[h] = *reinterpret_cast<const uint32_t*>(&pk);
      if (p + 1 < NS) {
        r[2 * h] -= __uint_as_float(w[h] << 16);
        r[2 * h + 1] -= __uint_as_float(w[h] & 0xFFFF0000u);
      }
    }
    out[p] = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

__device__ __forceinline__ void split4x3(const float4& x, uint2& hi, uint2& mid, uint2& lo) {
  const float v[4] = {x.x, x.y, x.z, x.w};
  __nv_bfloat16 h[4], m[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    h[i] = __float2bfloat16_rn(v[i]);
    float r1 = v[i] - __bfloat162float(h[i]);
    m[i] = __float2bfloat16_rn(r1);
    float r2 = r1 - __bfloat162float(m[i]);
    l[i] = __float2bfloat16_rn(r2);
  }
  hi.x = pack_bf16x2(h[0], h[1]); hi.y = pack_bf16x2(h[2], h[3]);
  mid.x = pack_bf16x2(m[0], m[1]); mid.y = pack_bf16x2(m[2], m[3]);
  lo.x = pack_bf16x2(l[0], l[1]); lo.y = pack_bf16x2(l[2], l[3]);
}

}  // namespace tc
