// Probe (GPU box): can tcgen05 kind::tf32 consume fp32 tiles exactly as a SWIZZLE_128B tensor-map TMA lands them?
//   forward shape : D[64 b][32 g]   = sum_pos X[pos][b] W[pos][g]   -- A and B both MN-major (pos = K is the row index)
//   gradient shape: D[128 pos][32 g] = sum_b  X[pos][b] dY[b][g]    -- A K-major straight from the same tile, B packed
// and does the tensor core truncate fp32 -> tf32 (so that lo = x - (x & 0xffffe000) is the exact remainder)?
// Prints max |error| against fp64 for hi*hi alone and for the 3-product split, with and without masking hi in place.
// Build: nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/tf32_probe.bin tools/tf32_probe.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>
#include "../masters-thesis_b200/csrc/tc.cuh"

__device__ __forceinline__ uint64_t desc_sw128(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = tc::smem_desc(saddr, lbo_bytes, sbo_bytes);
  d |= (uint64_t)2 << 61;                                    // layout type: SWIZZLE_128B
  return d;
}
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N, int a_mn, int b_mn) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(a_mn & 1) << 15) | ((uint32_t)(b_mn & 1) << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void mma_tf32(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d),
      "l"(a), "l"(b), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void tma_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
          tc::smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(tc::smem_u32(bar))
      : "memory");
}

constexpr int XH = 0, XL = 32768, WH = 65536, WL = 81920, BH = 98304, BL = 106496, BS = 114688, SMEM = 122880 + 1024;

struct OpDesc { int mn, off, lbo, sbo, kstep, kstep_hi, swz; };   // operand: major, smem offset, LBO, SBO, bytes per k-step (+ per 4 k-steps)
struct Case { int M, N, ksteps; OpDesc a, b; };

__global__ void __launch_bounds__(128)
probe(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapW, const float* __restrict__ dY,
      int mask_hi, float* __restrict__ out, const Case cs, int generic) {
  extern __shared__ unsigned char raw[];
  unsigned char* sm = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t bar_load, bar_mma;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) { tc::mbar_init(&bar_load, 1); tc::mbar_init(&bar_mma, 1); tc::mbar_fence_init(); }
  if (warp == 0) tc::tmem_alloc(&slot, 128);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tm = slot;
  if (tid == 0) {
    tc::mbar_expect_tx(&bar_load, 49152);
    tma_2d(sm + XH, &mapX, 0, 0, &bar_load);                 // batch columns 0..31 of 128 positions
    tma_2d(sm + XH + 16384, &mapX, 32, 0, &bar_load);        // batch columns 32..63
    tma_2d(sm + WH, &mapW, 0, 0, &bar_load);
  }
  tc::mbar_wait(&bar_load, 0);
  // remainders at the same (swizzled) offsets
  for (int i = tid; i < 3072; i += 128) {                    // 2048 float4 = 32 KB of X, then 1024 float4 of W
    const int region = i < 2048 ? 0 : 1;
    float4* hp = reinterpret_cast<float4*>(sm + (region == 0 ? XH : WH)) + (region == 0 ? i : i - 2048);
    float4* lp = reinterpret_cast<float4*>(sm + (region == 0 ? XL : WL)) + (region == 0 ? i : i - 2048);
    float4 v = *hp, h, l;
    h.x = __uint_as_float(__float_as_uint(v.x) & 0xffffe000u); l.x = v.x - h.x;
    h.y = __uint_as_float(__float_as_uint(v.y) & 0xffffe000u); l.y = v.y - h.y;
    h.z = __uint_as_float(__float_as_uint(v.z) & 0xffffe000u); l.z = v.z - h.z;
    h.w = __uint_as_float(__float_as_uint(v.w) & 0xffffe000u); l.w = v.w - h.w;
    *lp = l;
    if (mask_hi) *hp = h;
  }
  // dY^T as K-major no-swizzle core matrices of 4-byte elements: (g, b) -> (g/8)*2048 + (b/4)*128 + (g%8)*16 + (b%4)*4
  for (int i = tid; i < 2048; i += 128) {
    const int b = i >> 5, g = i & 31;
    const float v = dY[b * 32 + g];
    const float h = __uint_as_float(__float_as_uint(v) & 0xffffe000u);
    const int o = (g >> 3) * 2048 + (b >> 2) * 128 + (g & 7) * 16 + (b & 3) * 4;
    *reinterpret_cast<float*>(sm + BH + o) = mask_hi ? h : v;
    *reinterpret_cast<float*>(sm + BL + o) = v - h;
    // the same matrix as an MN-major SWIZZLE_128B image: row b (= K) of 128 bytes, 16-byte chunk index XOR (b % 8)
    *reinterpret_cast<float*>(sm + BS + b * 128 + (((g >> 2) ^ (b & 7)) << 4) + (g & 3) * 4) = v;
  }
  tc::fence_async_smem();
  __syncthreads();
  if (tid == 0 && generic) {
    tc::tc_fence_after();
    const uint32_t s0 = tc::smem_u32(sm);
    const uint32_t id = idesc_tf32(cs.M, cs.N, cs.a.mn, cs.b.mn);
    for (int ks = 0; ks < cs.ksteps; ++ks) {
      const uint32_t ao = cs.a.off + (ks & 3) * cs.a.kstep + (ks >> 2) * cs.a.kstep_hi;
      const uint32_t bo = cs.b.off + (ks & 3) * cs.b.kstep + (ks >> 2) * cs.b.kstep_hi;
      uint64_t da = tc::smem_desc(s0 + ao, cs.a.lbo, cs.a.sbo) | ((uint64_t)cs.a.swz << 61);
      uint64_t db = tc::smem_desc(s0 + bo, cs.b.lbo, cs.b.sbo) | ((uint64_t)cs.b.swz << 61);
      mma_tf32(tm + 0, da, db, id, ks > 0);
    }
    tc::mma_commit(&bar_mma);
  }
  if (tid == 0 && !generic) {
    tc::tc_fence_after();
    const uint32_t s0 = tc::smem_u32(sm);
    // ---- forward shape: M = 64 (b), N = 32 (g), K = 128 positions, 8 per instruction; both operands MN-major
    constexpr uint32_t idf = idesc_tf32(64, 32, 1, 1);
    for (int ks = 0; ks < 16; ++ks) {
      const uint64_t ah = desc_sw128(s0 + XH + ks * 1024, 16384, 1024), al = desc_sw128(s0 + XL + ks * 1024, 16384, 1024);
      const uint64_t bh = desc_sw128(s0 + WH + ks * 1024, 16384, 1024), bl = desc_sw128(s0 + WL + ks * 1024, 16384, 1024);
      mma_tf32(tm + 0, ah, bh, idf, ks > 0);
      mma_tf32(tm + 32, al, bh, idf, ks > 0);
      mma_tf32(tm + 32, ah, bl, idf, 1);
      mma_tf32(tm + 32, ah, bh, idf, 1);
    }
    // ---- gradient shape: M = 128 (pos), N = 32 (g), K = 64 batch rows; A K-major from the tile, B packed
    constexpr uint32_t idg = idesc_tf32(128, 32, 0, 0);
    for (int ks = 0; ks < 8; ++ks) {
      const uint32_t ao = (ks >> 2) * 16384 + (ks & 3) * 32;
      const uint64_t ah = desc_sw128(s0 + XH + ao, 16, 1024), al = desc_sw128(s0 + XL + ao, 16, 1024);
      const uint64_t bh = tc::smem_desc(s0 + BH + ks * 256, 128, 2048), bl = tc::smem_desc(s0 + BL + ks * 256, 128, 2048);
      mma_tf32(tm + 64, ah, bh, idg, ks > 0);
      mma_tf32(tm + 96, al, bh, idg, ks > 0);
      mma_tf32(tm + 96, ah, bl, idg, 1);
      mma_tf32(tm + 96, ah, bh, idg, 1);
    }
    tc::mma_commit(&bar_mma);
  }
  tc::mbar_wait(&bar_mma, 0);
  tc::tc_fence_after();
  float v[32];
  for (int blk = 0; blk < 4; ++blk) {
    tc::tmem_ld32(tm + ((uint32_t)(warp * 32) << 16) + blk * 32, v);
    tc::tmem_ld_wait();
    // out[blk][lane_global][32]
    for (int i = 0; i < 32; ++i) out[((size_t)blk * 128 + warp * 32 + lane) * 32 + i] = v[i];
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tm, 128);
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static CUtensorMap make_map(EncodeFn enc, void* base, uint64_t cols, uint64_t rows, uint32_t box_cols, uint32_t box_rows) {
  CUtensorMap m;
  cuuint64_t dims[2] = {cols, rows}, strides[1] = {cols * 4};
  cuuint32_t box[2] = {box_cols, box_rows}, es[2] = {1, 1};
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("cuTensorMapEncodeTiled failed: %d\n", (int)r); exit(1); }
  return m;
}

int main() {
  EncodeFn enc = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&enc, cudaEnableDefault, &q) != cudaSuccess || !enc) {
    printf("no cuTensorMapEncodeTiled\n");
    return 1;
  }
  std::vector<float> X(128 * 64), W(128 * 32), dY(64 * 32);
  srand(7);
  auto rnd = [] { return (float)((double)rand() / RAND_MAX * 2.0 - 1.0); };
  for (auto& v : X) v = rnd();
  for (auto& v : W) v = rnd();
  for (auto& v : dY) v = rnd();
  float *dX, *dW, *ddY, *dout;
  cudaMalloc(&dX, X.size() * 4); cudaMalloc(&dW, W.size() * 4); cudaMalloc(&ddY, dY.size() * 4);
  cudaMalloc(&dout, 4 * 128 * 32 * 4);
  cudaMemcpy(dX, X.data(), X.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dW, W.data(), W.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(ddY, dY.data(), dY.size() * 4, cudaMemcpyHostToDevice);
  CUtensorMap mx = make_map(enc, dX, 64, 128, 32, 128), mw = make_map(enc, dW, 32, 128, 32, 128);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
  std::vector<double> rf(64 * 32), rg(128 * 32);
  for (int b = 0; b < 64; ++b)
    for (int g = 0; g < 32; ++g) {
      double s = 0;
      for (int p = 0; p < 128; ++p) s += (double)X[p * 64 + b] * W[p * 32 + g];
      rf[b * 32 + g] = s;
    }
  for (int p = 0; p < 128; ++p)
    for (int g = 0; g < 32; ++g) {
      double s = 0;
      for (int b = 0; b < 64; ++b) s += (double)X[p * 64 + b] * dY[b * 32 + g];
      rg[p * 32 + g] = s;
    }
  for (int mask = 0; mask < 2; ++mask) {
    cudaMemset(dout, 0, 4 * 128 * 32 * 4);
    probe<<<1, 128, SMEM>>>(mx, mw, ddY, mask, dout, Case{}, 0);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("mask %d: kernel failed: %s\n", mask, cudaGetErrorString(e)); return 1; }
    std::vector<float> o(4 * 128 * 32);
    cudaMemcpy(o.data(), dout, o.size() * 4, cudaMemcpyDeviceToHost);
    double e1 = 0, e3 = 0, g1 = 0, g3 = 0;
    for (int b = 0; b < 64; ++b) {
      const int ln = 32 * (b / 16) + b % 16;                 // M = 64: row i -> lane 32 (i / 16) + i % 16
      for (int g = 0; g < 32; ++g) {
        e1 = fmax(e1, fabs(o[(0 * 128 + ln) * 32 + g] - rf[b * 32 + g]));
        e3 = fmax(e3, fabs(o[(1 * 128 + ln) * 32 + g] - rf[b * 32 + g]));
      }
    }
    for (int p = 0; p < 128; ++p)
      for (int g = 0; g < 32; ++g) {
        g1 = fmax(g1, fabs(o[(2 * 128 + p) * 32 + g] - rg[p * 32 + g]));
        g3 = fmax(g3, fabs(o[(3 * 128 + p) * 32 + g] - rg[p * 32 + g]));
      }
    printf("mask_hi=%d  forward (MN-major A,B; M=64): hi*hi max err %.3e, 3-product %.3e | gradient (K-major A from the tile; "
           "M=128): hi*hi %.3e, 3-product %.3e   [|D| ~ %.2f]\n", mask, e1, e3, g1, g3, fabs(rf[0]));
  }
  // ---- layout search: hi*hi only (an error ~1e-2 means the layout is right, ~10 means garbage)
  struct Named { const char* name; int shape; Case c; };
  const OpDesc A_K = {0, XH, 16, 1024, 32, 16384, 2};                       // known-good K-major A (gradient shape)
  const OpDesc B_K = {0, BH, 128, 2048, 256, 1024, 0};                      // known-good packed B (gradient shape)
  std::vector<Named> cases = {
      {"grad  A K-major sw128 | B packed K-major (control)", 1, {128, 32, 8, A_K, B_K}},
      {"grad  A K-major sw128 | B MN-major sw128 lbo=16384 sbo=1024", 1, {128, 32, 8, A_K, {1, BS, 16384, 1024, 1024, 4096, 2}}},
      {"grad  A K-major sw128 | B MN-major sw128 lbo=1024 sbo=16384", 1, {128, 32, 8, A_K, {1, BS, 1024, 16384, 1024, 4096, 2}}},
      {"grad  A K-major sw128 | B MN-major sw128 lbo=16 sbo=1024", 1, {128, 32, 8, A_K, {1, BS, 16, 1024, 1024, 4096, 2}}},
      {"fwd   A MN sw128 lbo=16384 sbo=1024 | B MN sw128 lbo=16384 sbo=1024", 0,
       {64, 32, 16, {1, XH, 16384, 1024, 1024, 4096, 2}, {1, WH, 16384, 1024, 1024, 4096, 2}}},
      {"fwd   A MN sw128 lbo=1024 sbo=16384 | B MN sw128 lbo=1024 sbo=16384", 0,
       {64, 32, 16, {1, XH, 1024, 16384, 1024, 4096, 2}, {1, WH, 1024, 16384, 1024, 4096, 2}}},
      {"fwd   A MN sw128 lbo=16384 sbo=1024 | B MN sw128 lbo=1024 sbo=16384", 0,
       {64, 32, 16, {1, XH, 16384, 1024, 1024, 4096, 2}, {1, WH, 1024, 16384, 1024, 4096, 2}}},
      {"fwd   A MN sw128 lbo=1024 sbo=16384 | B MN sw128 lbo=16384 sbo=1024", 0,
       {64, 32, 16, {1, XH, 1024, 16384, 1024, 4096, 2}, {1, WH, 16384, 1024, 1024, 4096, 2}}},
      {"fwd32 A MN sw128 (M=64 from ONE atom repeated: lbo=0) | B MN sw128", 2,
       {64, 32, 16, {1, XH, 0, 1024, 1024, 4096, 2}, {1, WH, 16384, 1024, 1024, 4096, 2}}},
  };
  for (auto& n : cases) {
    cudaMemset(dout, 0, 4 * 128 * 32 * 4);
    probe<<<1, 128, SMEM>>>(mx, mw, ddY, 0, dout, n.c, 1);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: kernel failed: %s\n", n.name, cudaGetErrorString(e)); return 1; }
    std::vector<float> o(128 * 32);
    cudaMemcpy(o.data(), dout, o.size() * 4, cudaMemcpyDeviceToHost);
    double err = 0, err_lo = 0;
    if (n.shape == 1) {
      for (int p = 0; p < 128; ++p)
        for (int g = 0; g < 32; ++g) err = fmax(err, fabs(o[p * 32 + g] - rg[p * 32 + g]));
    } else {
      for (int b = 0; b < 64; ++b) {
        const int ln = 32 * (b / 16) + b % 16;
        const int bref = n.shape == 2 ? b % 32 : b;
        for (int g = 0; g < 32; ++g) {
          double d = fabs(o[ln * 32 + g] - rf[bref * 32 + g]);
          err = fmax(err, d);
          if (b < 32) err_lo = fmax(err_lo, d);
        }
      }
    }
    printf("%-78s max err %.3e%s\n", n.name, err, n.shape == 1 ? "" : (std::string("  (rows 0..31 only: ") + std::to_string(err_lo) + ")").c_str());
  }
  return 0;
}
