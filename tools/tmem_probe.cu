// Which TMEM lanes hold the rows of a cta_group::1 tcgen05.mma accumulator with M = 64 (and M = 128 as a control)?
// D[m][n] = (m + 1) * (n + 1) from rank-1 operands; every (lane, column) of the 128 x 32 window is read back and the row
// it belongs to is printed.  Build: nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/tmem_probe.bin
// tools/tmem_probe.cu ; run on the GPU box.
#include <cuda_runtime.h>
#include <cstdio>
#include <vector>
#include "../masters-thesis_b200/csrc/tc.cuh"

template <int M>
__global__ void probe_kernel(float* out) {
  __shared__ __align__(1024) unsigned char sm[8192];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  __nv_bfloat16* A = reinterpret_cast<__nv_bfloat16*>(sm);            // [M][16] K-major core matrices: (m/8, k/8)
  __nv_bfloat16* Bm = reinterpret_cast<__nv_bfloat16*>(sm + 4096);    // [32][16]
  for (int i = tid; i < 2048; i += blockDim.x) { A[i] = __float2bfloat16(0.f); Bm[i] = __float2bfloat16(0.f); }
  __syncthreads();
  // element (m, k): byte offset (m/8)*256 + (k/8)*128 + (m%8)*16 + (k%8)*2
  if (tid < M) A[((tid >> 3) * 256 + (tid & 7) * 16) / 2] = __float2bfloat16((float)(tid + 1));
  if (tid < 32) Bm[((tid >> 3) * 256 + (tid & 7) * 16) / 2] = __float2bfloat16((float)(tid + 1));
  if (tid == 0) { tc::mbar_init(&bar, 1); tc::mbar_fence_init(); }
  if (warp == 0) tc::tmem_alloc(&slot, 32);
  tc::fence_async_smem();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tm = slot;
  // pre-fill the window with a marker so untouched lanes show up
  {
    uint32_t addr = tm + ((uint32_t)(warp * 32) << 16);
    asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(addr), "r"(0xFFC00000u) : "memory");
    for (int c = 1; c < 32; ++c)
      asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(addr + c), "r"(0xFFC00000u) : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  if (tid == 0) {
    const uint32_t idesc = tc::idesc_bf16(M, 32, 0, 0);
    tc::mma_bf16(tm, tc::smem_desc(tc::smem_u32(A), 128, 256), tc::smem_desc(tc::smem_u32(Bm), 128, 256), idesc, 0u);
    tc::mma_commit(&bar);
  }
  tc::mbar_wait(&bar, 0);
  tc::tc_fence_after();
  float v[32];
  tc::tmem_ld32(tm + ((uint32_t)(warp * 32) << 16), v);
  tc::tmem_ld_wait();
  for (int c = 0; c < 32; ++c) out[(size_t)tid * 32 + c] = v[c];
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tm, 32);
}

template <int M> static void run() {
  float* d;
  cudaMalloc(&d, 128 * 32 * 4);
  probe_kernel<M><<<1, 128>>>(d);
  cudaError_t e = cudaDeviceSynchronize();
  std::vector<float> h(128 * 32);
  cudaMemcpy(h.data(), d, h.size() * 4, cudaMemcpyDeviceToHost);
  printf("M=%d: %s\n", M, cudaGetErrorString(e));
  for (int lane = 0; lane < 128; ++lane) {
    const float c0 = h[lane * 32], c1 = h[lane * 32 + 1];
    int row = -1;
    if (c0 == c0 && c0 > 0) row = (int)(c0 + 0.5f) - 1;                     // column 0 holds (m + 1)
    const bool ok = row >= 0 && c1 == 2.0f * (row + 1);
    printf("%s%d:%d%s", lane % 16 ? " " : "\n  ", lane, row, ok ? "" : "?");
  }
  printf("\n");
  cudaFree(d);
}

int main() {
  run<128>();
  run<64>();
  return 0;
}
