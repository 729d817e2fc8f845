// Batch assembly of the reference's data generators from tables resident in HBM.
//   data_generator_guse.py:129-171 (`__data_generation`): betas_batch[i] = np.load(key_i) (one 327 684-float row per
//   sample), cap_vector = pad_sequences(...), target[:, :-1] = cap_vector[:, 1:] -> to_categorical;
//   data_generator_multisub.py:81-120: the same with subject A's rows first, subject B's rows behind them.
// Here a subject's whole beta matrix (n_keys, N) and the tokenised, padded caption table (n_pairs, T) live in HBM;
// a batch is an index gather:
//   ds_gather_rows_kernel : out[j, :] = table[rows[j], :]       (HBM copy: 2 * B * N * 4 bytes per launch)
//   ds_captions_kernel    : cap[j, :] = caps[idx[j], :];  target[j, t] = caps[idx[j], t+1] (0 at t = T-1);
//                           optionally the dense one-hot (B, T, V) the reference hands to train_step.
// Stateless entry points (no tnt_ctx): plain device pointers and sizes, asynchronous on `stream`.
#include "../../include/tnt.h"
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

static thread_local char g_data_err[256] = "";
static long long g_data_launches = 0;

static int data_fail(int code, const char* msg) {
  snprintf(g_data_err, sizeof(g_data_err), "%s", msg);
  return code;
}
static int data_check(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    snprintf(g_data_err, sizeof(g_data_err), "%s: %s", what, cudaGetErrorString(e));
    return TNT_ERR_CUDA;
  }
  return TNT_OK;
}

#define DS_THREADS 256
#define DS_UNROLL 8            // 16-byte loads in flight per thread
#define DS_CHUNK (DS_THREADS * DS_UNROLL)   // float4 per CTA iteration = 32 KB

// One work item = (batch row j, 32 KB chunk c) of the output.  Rows are whole-row contiguous in both table and batch,
// so every warp instruction moves 512 contiguous bytes; loads are streaming (read once), 8 in flight per thread.
__global__ void __launch_bounds__(DS_THREADS)
ds_gather_rows_v4_kernel(const float4* __restrict__ table, int64_t n_rows, int64_t width4, const int32_t* __restrict__ rows,
                         int B, float4* __restrict__ out, int* __restrict__ bad) {
  const int64_t chunks = (width4 + DS_CHUNK - 1) / DS_CHUNK;
  const int64_t items = chunks * B;
  for (int64_t it = blockIdx.x; it < items; it += gridDim.x) {
    const int j = (int)(it / chunks);
    const int64_t c0 = (it - (int64_t)j * chunks) * DS_CHUNK + threadIdx.x;
    const int64_t r = rows[j];
    if (r < 0 || r >= n_rows) {            // never read outside the table: flag it, leave the row zero
      if (threadIdx.x == 0) atomicExch(bad, 1);
      for (int u = 0; u < DS_UNROLL; ++u) {
        const int64_t c = c0 + (int64_t)u * DS_THREADS;
        if (c < width4) out[(int64_t)j * width4 + c] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      continue;
    }
    const float4* src = table + r * width4;
    float4* dst = out + (int64_t)j * width4;
    float4 v[DS_UNROLL];
#pragma unroll
    for (int u = 0; u < DS_UNROLL; ++u) {
      const int64_t c = c0 + (int64_t)u * DS_THREADS;
      if (c < width4) v[u] = __ldcs(src + c);
    }
#pragma unroll
    for (int u = 0; u < DS_UNROLL; ++u) {
      const int64_t c = c0 + (int64_t)u * DS_THREADS;
      if (c < width4) dst[c] = v[u];
    }
  }
}

// widths that are not a multiple of 4 floats (or unaligned bases): scalar path, same item decomposition
__global__ void __launch_bounds__(DS_THREADS)
ds_gather_rows_s_kernel(const float* __restrict__ table, int64_t n_rows, int64_t width, const int32_t* __restrict__ rows,
                        int B, float* __restrict__ out, int* __restrict__ bad) {
  const int64_t chunks = (width + DS_CHUNK - 1) / DS_CHUNK;
  const int64_t items = chunks * B;
  for (int64_t it = blockIdx.x; it < items; it += gridDim.x) {
    const int j = (int)(it / chunks);
    const int64_t c0 = (it - (int64_t)j * chunks) * DS_CHUNK + threadIdx.x;
    const int64_t r = rows[j];
    const bool ok = r >= 0 && r < n_rows;
    if (!ok && threadIdx.x == 0) atomicExch(bad, 1);
#pragma unroll
    for (int u = 0; u < DS_UNROLL; ++u) {
      const int64_t c = c0 + (int64_t)u * DS_THREADS;
      if (c < width) out[(int64_t)j * width + c] = ok ? __ldcs(table + r * width + c) : 0.0f;
    }
  }
}

// grid (B, T): CTA (j, 0) writes the caption row and the shifted target ids of sample j; CTA (j, t) writes the V
// floats of one-hot row (j, t) when the dense target is asked for (coalesced 4-byte stores, 19 MB at B=64)
__global__ void __launch_bounds__(256)
ds_captions_kernel(const int32_t* __restrict__ caps, int64_t n_caps, int T, const int32_t* __restrict__ idx,
                   int32_t* __restrict__ cap_out, int32_t* __restrict__ tgt_out, float* __restrict__ onehot, int V,
                   int* __restrict__ bad) {
  const int j = blockIdx.x, t0 = blockIdx.y;
  const int64_t r = idx[j];
  const bool ok = r >= 0 && r < n_caps;
  const int32_t* src = caps + (ok ? r : 0) * (int64_t)T;
  if (t0 == 0) {
    if (!ok && threadIdx.x == 0) atomicExch(bad, 1);
    for (int t = threadIdx.x; t < T; t += blockDim.x) {
      const int32_t w = ok ? src[t] : 0;
      const int32_t nx = (ok && t + 1 < T) ? src[t + 1] : 0;      // target[:, :-1] = cap[:, 1:], trailing 0
      if (cap_out) cap_out[(int64_t)j * T + t] = w;
      if (tgt_out) tgt_out[(int64_t)j * T + t] = nx;
    }
  }
  if (onehot) {
    const int32_t y = (ok && t0 + 1 < T) ? src[t0 + 1] : 0;
    // a token id outside [0, V) has no one-hot column (to_categorical raises): flag it, row stays all-zero
    if ((y < 0 || y >= V) && threadIdx.x == 0) atomicExch(bad, 1);
    float* o = onehot + ((int64_t)j * T + t0) * V;
    for (int v = threadIdx.x; v < V; v += blockDim.x) o[v] = (v == y) ? 1.0f : 0.0f;
  }
}

static int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
  }
  return n;
}

extern "C" const char* tnt_data_last_error(void) { return g_data_err; }
extern "C" int64_t tnt_data_launch_count(void) { return g_data_launches; }

extern "C" int tnt_assemble_betas(const float* table, int64_t n_rows, int64_t width, const int32_t* rows, int B,
                                  float* out, int32_t* bad_flag, void* stream) {
  if (!table || !rows || !out || !bad_flag || n_rows <= 0 || width <= 0 || B < 0)
    return data_fail(TNT_ERR_ARG, "tnt_assemble_betas: bad argument");
  if (B == 0) return TNT_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const bool vec = (width % 4 == 0) && (((uintptr_t)table | (uintptr_t)out) % 16 == 0);
  const int64_t w = vec ? width / 4 : width;
  const int64_t items = ((w + DS_CHUNK - 1) / DS_CHUNK) * B;
  const int64_t cap = (int64_t)sm_count() * 8;                 // 8 resident CTAs of 256 threads per SM
  const unsigned grid = (unsigned)(items < cap ? items : cap);
  if (vec)
    ds_gather_rows_v4_kernel<<<grid, DS_THREADS, 0, st>>>((const float4*)table, n_rows, w, rows, B, (float4*)out, bad_flag);
  else
    ds_gather_rows_s_kernel<<<grid, DS_THREADS, 0, st>>>(table, n_rows, w, rows, B, out, bad_flag);
  g_data_launches += 1;
  return data_check("tnt_assemble_betas");
}

extern "C" int tnt_assemble_captions(const int32_t* cap_table, int64_t n_caps, int T, const int32_t* idx, int B,
                                     int32_t* cap_out, int32_t* target_ids_out, float* onehot_out, int V,
                                     int32_t* bad_flag, void* stream) {
  if (!cap_table || !idx || !bad_flag || n_caps <= 0 || T <= 0 || B < 0 || (onehot_out && V <= 0))
    return data_fail(TNT_ERR_ARG, "tnt_assemble_captions: bad argument");
  if (B == 0) return TNT_OK;
  ds_captions_kernel<<<dim3((unsigned)B, onehot_out ? (unsigned)T : 1u), 256, 0, (cudaStream_t)stream>>>(cap_table, n_caps, T, idx, cap_out, target_ids_out,
                                                                     onehot_out, V, bad_flag);
  g_data_launches += 1;
  return data_check("tnt_assemble_captions");
}
