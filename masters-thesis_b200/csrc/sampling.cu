// Decoders beyond arg-max: the word choice of lc_NIC.sample_choice (tf.random.categorical,
// AttemptFour/Model/lc_NIC.py:571-575), top-k / nucleus selection (AttemptFour/Model/img_NIC.py:509-533) and beam
// search (what lc_NIC.beam_search, lc_NIC.py:640-692, set out to do) on top of the decode step of decoder.cu
// (attention -> LSTM -> dense_inter -> dense_out, lc_NIC.py:607-632).  Inference mode: no dropout anywhere.
#include "tnt_internal.h"
#include <climits>

#define SITE_SAMPLE 9          // Philox site of the sampling draw (oracle/philox.py)

// ------------------------------------------------------------------------------------------------
// random word choice: categorical / top-k / nucleus in one rule (include/tnt.h: tnt_sample)
// ------------------------------------------------------------------------------------------------
// before(a, b): a precedes b in the descending order (larger probability first, ties: lower word id first)
__device__ __forceinline__ bool samp_before(float ka, int ia, float kb, int ib) { return ka > kb || (ka == kb && ia < ib); }

__global__ void __launch_bounds__(256)
sample_kernel(int V, int NP, const float* __restrict__ logits, uint32_t step, int top_k, float top_p, uint32_t k0,
              uint32_t k1, int32_t* __restrict__ word, int64_t* __restrict__ out_ids, int out_stride,
              float* __restrict__ probs, int64_t probs_stride) {
  extern __shared__ __align__(16) unsigned char samp_smem[];
  float* key = reinterpret_cast<float*>(samp_smem);          // [NP]
  int* idx = reinterpret_cast<int*>(key + NP);               // [NP]
  __shared__ float rs[32];
  __shared__ double seg[256];
  __shared__ int s_first;
  __shared__ double s_total;
  const int b = blockIdx.x, tid = threadIdx.x;
  const float* row = logits + (size_t)b * V;
  // probabilities exactly as the greedy path forms them: exp(z - max) / sum
  float best = -INFINITY;
  for (int v = tid; v < V; v += 256) best = fmaxf(best, row[v]);
  best = block_max(best, rs);
  float sum = 0.0f;
  for (int v = tid; v < V; v += 256) sum += expf(row[v] - best);
  sum = block_sum(sum, rs);
  float* pr = probs ? probs + (size_t)b * probs_stride : nullptr;
  for (int v = tid; v < NP; v += 256) {
    float p = -1.0f;
    if (v < V) { p = expf(row[v] - best) / sum; if (pr) pr[v] = p; }
    key[v] = p; idx[v] = v < V ? v : INT_MAX;
  }
  __syncthreads();
  // bitonic sort, descending by (probability, -id)
  for (int k = 2; k <= NP; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < NP; i += 256) {
        const int x = i ^ j;
        if (x > i) {
          const float ka = key[i], kb = key[x];
          const int ia = idx[i], ib = idx[x];
          const bool up = (i & k) == 0;
          const bool swap = up ? samp_before(kb, ib, ka, ia) : samp_before(ka, ia, kb, ib);
          if (swap) { key[i] = kb; key[x] = ka; idx[i] = ib; idx[x] = ia; }
        }
      }
      __syncthreads();
    }
  // running sums in fp64 over the sorted order: thread t owns entries [t*L, (t+1)*L)
  const int L = NP / 256;
  double loc = 0.0;
  for (int i = 0; i < L; ++i) { const int e = tid * L + i; if (e < V) loc += (double)key[e]; }
  seg[tid] = loc;
  __syncthreads();
  if (tid == 0) {
    double run = 0.0;
    for (int t = 0; t < 256; ++t) { const double v = seg[t]; seg[t] = run; run += v; }   // exclusive
    s_total = run;
    s_first = V;
  }
  __syncthreads();
  // how many words are kept
  int n_keep = V;
  if (top_k > 0 && top_k < n_keep) n_keep = top_k;
  if (top_p > 0.0f && top_p < 1.0f) {
    double run = seg[tid];
    for (int i = 0; i < L; ++i) {
      const int e = tid * L + i;
      if (e >= V) break;
      run += (double)key[e];
      if (run > (double)top_p) { atomicMin(&s_first, e + 1); break; }      // shortest prefix whose sum exceeds top_p
    }
    __syncthreads();
    if (s_first < n_keep) n_keep = s_first;
    __syncthreads();
  }
  // total of the kept prefix
  if (tid == 0) s_first = n_keep - 1;
  {
    // thread owning entry n_keep-1 publishes the inclusive sum there
    const int owner = (n_keep - 1) / L;
    if (tid == owner) {
      double run = seg[tid];
      for (int i = 0; i <= (n_keep - 1) - tid * L; ++i) run += (double)key[tid * L + i];
      s_total = run;
    }
  }
  __syncthreads();
  const uint4 r = philox4x32_10((uint32_t)b, SITE_SAMPLE, step, 0u, k0, k1);
  const double u = ((double)r.x + 0.5) * (1.0 / 4294967296.0);
  const double thr = u * s_total;
  {
    double run = seg[tid];
    for (int i = 0; i < L; ++i) {
      const int e = tid * L + i;
      if (e >= n_keep) break;
      run += (double)key[e];
      if (run > thr) { atomicMin(&s_first, e); break; }
    }
  }
  __syncthreads();
  if (tid == 0) {
    const int w = idx[s_first];
    word[b] = w;
    out_ids[(size_t)b * out_stride] = (int64_t)w;
  }
}

void dec_sample_words(tnt_ctx* c, int B, int step, int top_k, float top_p, uint64_t seed, const float* logits,
                      int32_t* word, int64_t* out_ids, int out_stride, float* probs, int64_t probs_stride,
                      cudaStream_t st) {
  const int V = c->cfg.V;
  int NP = 256;
  while (NP < V) NP <<= 1;
  const size_t smem = (size_t)NP * 8;
  static bool attr = false;
  if (!attr) { cudaFuncSetAttribute(sample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192 * 8); attr = true; }
  sample_kernel<<<B, 256, smem, st>>>(V, NP, logits, (uint32_t)step, top_k, top_p, (uint32_t)(seed & 0xFFFFFFFFull),
                                      (uint32_t)(seed >> 32), word, out_ids, out_stride, probs, probs_stride);
  c->launches += 1;
}

// ------------------------------------------------------------------------------------------------
// beam search
// ------------------------------------------------------------------------------------------------
#define BEAM_MAX 8
#define BEAM_NEG (-1.0e30f)

// rows b*W + w of dst = row b of src
__global__ void beam_expand_kernel(int B, int W, int n4, const float4* __restrict__ src, float4* __restrict__ dst) {
  const int row = blockIdx.x;
  const float4* s = src + (size_t)(row / W) * n4;
  float4* d = dst + (size_t)row * n4;
  for (int i = threadIdx.x; i < n4; i += blockDim.x) d[i] = s[i];
}

__global__ void beam_init_kernel(int B, int W, const int32_t* __restrict__ start, int32_t* __restrict__ word,
                                 float* __restrict__ score, int32_t* __restrict__ fin) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= B * W) return;
  word[row] = start[row / W];
  score[row] = (row % W) == 0 ? 0.0f : BEAM_NEG;       // every beam starts from the same prefix: only beam 0 is live
  fin[row] = 0;
}

// per row: the W best continuations, score + log p(v), ordered by (score desc, v asc); a finished row continues with
// id 0 at no cost
__global__ void __launch_bounds__(256)
beam_row_topk_kernel(int V, int W, const float* __restrict__ logits, const float* __restrict__ score,
                     const int32_t* __restrict__ fin, float* __restrict__ cand_score, int32_t* __restrict__ cand_word) {
  __shared__ float rs[32];
  __shared__ float rv[8];
  __shared__ int ri[8];
  __shared__ int chosen[BEAM_MAX];
  const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float* z = logits + (size_t)row * V;
  const float sc = score[row];
  if (fin[row] || sc <= BEAM_NEG) {
    if (tid < W) {
      cand_score[(size_t)row * W + tid] = (tid == 0 && fin[row]) ? sc : BEAM_NEG;
      cand_word[(size_t)row * W + tid] = 0;
    }
    return;
  }
  float best = -INFINITY;
  for (int v = tid; v < V; v += 256) best = fmaxf(best, z[v]);
  best = block_max(best, rs);
  float sum = 0.0f;
  for (int v = tid; v < V; v += 256) sum += expf(z[v] - best);
  sum = block_sum(sum, rs);
  const float lse = best + logf(sum);
  for (int k = 0; k < W; ++k) {
    float bv = -INFINITY; int bi = INT_MAX;
    for (int v = tid; v < V; v += 256) {
      bool taken = false;
      for (int q = 0; q < k; ++q) taken |= (chosen[q] == v);
      const float x = z[v];
      if (!taken && (x > bv || (x == bv && v < bi))) { bv = x; bi = v; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ob > bv || (ob == bv && oi < bi)) { bv = ob; bi = oi; }
    }
    if (lane == 0) { rv[warp] = bv; ri[warp] = bi; }
    __syncthreads();
    if (tid == 0) {
      bv = rv[0]; bi = ri[0];
      for (int i = 1; i < 8; ++i)
        if (rv[i] > bv || (rv[i] == bv && ri[i] < bi)) { bv = rv[i]; bi = ri[i]; }
      chosen[k] = bi;
      cand_score[(size_t)row * W + k] = sc + (bv - lse);
      cand_word[(size_t)row * W + k] = bi;
    }
    __syncthreads();
  }
}

// per sample: the W best of the W*W row candidates, ordered by (score desc, flat candidate index asc)
__global__ void beam_select_kernel(int W, int end_id, const float* __restrict__ cand_score,
                                   const int32_t* __restrict__ cand_word, const int32_t* __restrict__ fin,
                                   int32_t* __restrict__ parent, int32_t* __restrict__ new_word,
                                   float* __restrict__ new_score, int32_t* __restrict__ new_fin) {
  const int b = blockIdx.x;
  if (threadIdx.x != 0) return;
  const float* cs = cand_score + (size_t)b * W * W;
  const int32_t* cw = cand_word + (size_t)b * W * W;
  bool used[BEAM_MAX * BEAM_MAX];
  for (int i = 0; i < W * W; ++i) used[i] = false;
  for (int k = 0; k < W; ++k) {
    int bi = -1; float bv = -INFINITY;
    for (int i = 0; i < W * W; ++i)
      if (!used[i] && (bi < 0 || cs[i] > bv)) { bv = cs[i]; bi = i; }
    used[bi] = true;
    const int row = b * W + k, prow = b * W + bi / W;
    parent[row] = prow;
    new_word[row] = cw[bi];
    new_score[row] = bv;
    new_fin[row] = (fin[prow] || cw[bi] == end_id) ? 1 : 0;
  }
}

// state and history follow their parent beam; the new word is appended
__global__ void beam_reorder_kernel(int U, int max_len, int step, const int32_t* __restrict__ parent,
                                    const int32_t* __restrict__ new_word, const float* __restrict__ h_src,
                                    const float* __restrict__ c_src, float* __restrict__ h_dst, float* __restrict__ c_dst,
                                    const int32_t* __restrict__ ids_src, int32_t* __restrict__ ids_dst) {
  const int row = blockIdx.x, p = parent[row];
  for (int u = threadIdx.x; u < U; u += blockDim.x) {
    h_dst[(size_t)row * U + u] = h_src[(size_t)p * U + u];
    c_dst[(size_t)row * U + u] = c_src[(size_t)p * U + u];
  }
  for (int t = threadIdx.x; t <= step; t += blockDim.x)
    ids_dst[(size_t)row * max_len + t] = t == step ? new_word[row] : ids_src[(size_t)p * max_len + t];
}

__global__ void beam_finish_kernel(int n, int max_len, const int32_t* __restrict__ ids, const float* __restrict__ score,
                                   int64_t* __restrict__ out_ids, float* __restrict__ out_scores) {
  const int row = blockIdx.x;
  for (int t = threadIdx.x; t < max_len; t += blockDim.x) out_ids[(size_t)row * max_len + t] = (int64_t)ids[(size_t)row * max_len + t];
  if (threadIdx.x == 0) out_scores[row] = score[row];
}

__global__ void __launch_bounds__(256)
beam_gather_rows_kernel(int n4, const int32_t* __restrict__ word, int V, const float4* __restrict__ table, float4* __restrict__ out) {
  int w = word[blockIdx.x];
  w = w < 0 ? 0 : (w >= V ? V - 1 : w);
  const float4* src = table + (size_t)w * n4;
  float4* dst = out + (size_t)blockIdx.x * n4;
  for (int i = threadIdx.x; i < n4; i += blockDim.x) dst[i] = __ldg(src + i);
}

static void* beam_ws(tnt_ctx* c, const char* name, size_t bytes) {
  auto it = c->bufs.find(name);
  if (it != c->bufs.end() && (size_t)it->second.second >= bytes) return it->second.first;
  if (it != c->bufs.end()) { cudaFree(it->second.first); c->bufs.erase(it); }
  void* p = nullptr;
  if (cudaMalloc(&p, bytes < 16 ? 16 : bytes) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  c->bufs[name] = std::make_pair(p, (int64_t)bytes);
  return p;
}

// features / Pf of rows 0..B-1 are in place (enc_forward + enc_bn_finish); h0, c0 (B,U); start (B)
int dec_beam_search(tnt_ctx* c, const float* h0, const float* c0, const int32_t* start, int B, int max_len, int W,
                    int end_id, int64_t* out_ids, float* out_scores, cudaStream_t st) {
  const tnt_config& f = c->cfg;
  const int U = f.U, E = f.E, G = f.G, V = f.V, R = f.R, A = f.A;
  const int rows = B * W;
  const size_t bu = (size_t)rows * U;
  float* score = (float*)beam_ws(c, "beam_score", 2 * rows * sizeof(float));
  int32_t* fin = (int32_t*)beam_ws(c, "beam_fin", 2 * rows * sizeof(int32_t));
  int32_t* ids = (int32_t*)beam_ws(c, "beam_ids", 2 * (size_t)rows * max_len * sizeof(int32_t));
  float* cand_score = (float*)beam_ws(c, "beam_cand_score", (size_t)rows * W * sizeof(float));
  int32_t* cand_word = (int32_t*)beam_ws(c, "beam_cand_word", (size_t)rows * W * sizeof(int32_t));
  int32_t* parent = (int32_t*)beam_ws(c, "beam_parent", rows * sizeof(int32_t));
  if (!score || !fin || !ids || !cand_score || !cand_word || !parent) return TNT_ERR_CUDA;
  float* hA = wsT<float>(c, "Hall");            // slot 0: state before the step, slot 1: after the step (T >= 1)
  float* cA = wsT<float>(c, "Call");
  float* Zx = wsT<float>(c, "Zx");
  float* inter = wsT<float>(c, "inter_d");
  float* logits = wsT<float>(c, "logits");
  int32_t* word = wsT<int32_t>(c, "word");
  // every beam of a sample sees the sample's features: rows b*W + w <- row b (through a scratch copy)
  float* feats = wsT<float>(c, "feats"); float* Pf = wsT<float>(c, "Pf");
  float* tmpF = wsT<float>(c, "Y"); float* tmpP = wsT<float>(c, "dPf");
  cudaMemcpyAsync(tmpF, feats, (size_t)B * R * G * sizeof(float), cudaMemcpyDeviceToDevice, st);
  cudaMemcpyAsync(tmpP, Pf, (size_t)B * R * A * sizeof(float), cudaMemcpyDeviceToDevice, st);
  beam_expand_kernel<<<rows, 256, 0, st>>>(B, W, R * G / 4, reinterpret_cast<const float4*>(tmpF), reinterpret_cast<float4*>(feats));
  beam_expand_kernel<<<rows, 256, 0, st>>>(B, W, R * A / 4, reinterpret_cast<const float4*>(tmpP), reinterpret_cast<float4*>(Pf));
  beam_expand_kernel<<<rows, 128, 0, st>>>(B, W, U / 4, reinterpret_cast<const float4*>(h0), reinterpret_cast<float4*>(hA));
  beam_expand_kernel<<<rows, 128, 0, st>>>(B, W, U / 4, reinterpret_cast<const float4*>(c0), reinterpret_cast<float4*>(cA));
  beam_init_kernel<<<(rows + 127) / 128, 128, 0, st>>>(B, W, start, word, score, fin);
  c->launches += 5;
  // text half of the LSTM input projection for every word, once per call (as dec_greedy)
  float* emb_proj = wsT<float>(c, "emb_proj");
  gemm_tc(c, 0, 1, false, false, V, 4 * U, E, c->w + c->o_emb, E, c->w + c->o_lW + (size_t)G * 4 * U, 4 * U, emb_proj,
          4 * U, c->w + c->o_lb, 0, st);
  for (int i = 0; i < max_len; ++i) {
    const int cur = i & 1, nxt = cur ^ 1;
    beam_gather_rows_kernel<<<rows, 256, 0, st>>>(U, word, V, reinterpret_cast<const float4*>(emb_proj), reinterpret_cast<float4*>(Zx));
    dec_step(c, rows, i, 0, 0, hA, cA, Zx, wsT<float>(c, "alpha"), wsT<float>(c, "qact"), wsT<float>(c, "xctx"), nullptr,
             cA + bu, hA + bu, nullptr, nullptr, st);
    head_decode_step(c, rows, hA + bu, inter, logits, st);
    beam_row_topk_kernel<<<rows, 256, 0, st>>>(V, W, logits, score + cur * rows, fin + cur * rows, cand_score, cand_word);
    beam_select_kernel<<<B, 32, 0, st>>>(W, end_id, cand_score, cand_word, fin + cur * rows, parent, word,
                                         score + nxt * rows, fin + nxt * rows);
    beam_reorder_kernel<<<rows, 128, 0, st>>>(U, max_len, i, parent, word, hA + bu, cA + bu, hA, cA,
                                              ids + (size_t)cur * rows * max_len, ids + (size_t)nxt * rows * max_len);
    c->launches += 4;
  }
  const int last = max_len & 1;
  beam_finish_kernel<<<rows, 32, 0, st>>>(rows, max_len, ids + (size_t)last * rows * max_len, score + last * rows, out_ids, out_scores);
  c->launches += 1;
  return TNT_OK;
}
