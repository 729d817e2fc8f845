// optimizer.apply_gradients (AttemptFour/Model/lc_NIC.py:389) for
// tf.keras.optimizers.Adam(1e-4, beta_1=.9, beta_2=.98, epsilon=1e-8, clipnorm=.1) (main.py:97):
// per-VARIABLE tf.clip_by_norm, then ResourceApplyAdam; and the L2 regulariser terms
// (`tf.add_n(self.losses)`, lc_NIC.py:379) whose gradient 2*lambda*W is part of the tape gradient.
// The flat w/m/v/g buffers are cut into chunks of <= OPT_CHUNK floats that never straddle a variable.
#include "tnt_internal.h"

__global__ void __launch_bounds__(256)
opt_norm_kernel(const OptChunk* __restrict__ chunks, const Seg* __restrict__ segs, const float* __restrict__ w,
                const float* __restrict__ g, int with_grad, double* __restrict__ seg_gsq,
                double* __restrict__ seg_wsq) {
  __shared__ float rs[32];
  const OptChunk ch = chunks[blockIdx.x];
  const Seg sg = segs[ch.seg];
  const float two_l = 2.0f * sg.l2;
  float gs = 0.0f, wsq = 0.0f;
  // chunks start 16-byte aligned (tensors are placed on 4-float boundaries, chunk size is a multiple of 4)
  const int n4 = (ch.off & 3) ? 0 : (ch.len >> 2);
  const float4* w4 = reinterpret_cast<const float4*>(w + ch.off);
  const float4* g4 = reinterpret_cast<const float4*>(g + ch.off);
#pragma unroll 2
  for (int i = threadIdx.x; i < n4; i += blockDim.x) {
    const float4 wv = __ldg(w4 + i);
    wsq = fmaf(wv.x, wv.x, fmaf(wv.y, wv.y, fmaf(wv.z, wv.z, fmaf(wv.w, wv.w, wsq))));
    if (with_grad) {
      const float4 gr = __ldg(g4 + i);
      const float a = gr.x + two_l * wv.x, b = gr.y + two_l * wv.y, c2 = gr.z + two_l * wv.z, d = gr.w + two_l * wv.w;
      gs = fmaf(a, a, fmaf(b, b, fmaf(c2, c2, fmaf(d, d, gs))));
    }
  }
  for (int i = 4 * n4 + threadIdx.x; i < ch.len; i += blockDim.x) {
    float wv = w[ch.off + i];
    wsq = fmaf(wv, wv, wsq);
    if (with_grad) {
      float gv = g[ch.off + i] + two_l * wv;
      gs = fmaf(gv, gv, gs);
    }
  }
  if (with_grad) {
    float t = block_sum(gs, rs);
    if (threadIdx.x == 0) atomicAdd(&seg_gsq[ch.seg], (double)t);
  }
  if (sg.l2 > 0.0f) {
    float t = block_sum(wsq, rs);
    if (threadIdx.x == 0) atomicAdd(&seg_wsq[ch.seg], (double)t);
  }
}

__global__ void __launch_bounds__(256)
opt_l2_reduce_kernel(int nseg, const Seg* __restrict__ segs, const double* __restrict__ seg_wsq,
                     double* __restrict__ out) {
  __shared__ double red[256];
  double s = 0.0;
  for (int i = threadIdx.x; i < nseg; i += blockDim.x) s += (double)segs[i].l2 * seg_wsq[i];
  red[threadIdx.x] = s;
  __syncthreads();
  for (int k = 128; k > 0; k >>= 1) {
    if ((int)threadIdx.x < k) red[threadIdx.x] += red[threadIdx.x + k];
    __syncthreads();
  }
  if (threadIdx.x == 0) *out = red[0];
}

__global__ void __launch_bounds__(256)
opt_adam_kernel(const OptChunk* __restrict__ chunks, const Seg* __restrict__ segs, float* __restrict__ w,
                float* __restrict__ m, float* __restrict__ v, const float* __restrict__ g,
                const double* __restrict__ seg_gsq, const double* __restrict__ emb_sq, int emb_unagg,
                float clipnorm, float alpha, float one_m_b1, float one_m_b2, float eps,
                const float* __restrict__ agc_factor, const unsigned* __restrict__ abort_flag) {
  // a persistent recurrence kernel whose barrier timed out leaves invalid gradients behind: never apply them
  if (abort_flag != nullptr && *abort_flag != 0u) return;
  const OptChunk ch = chunks[blockIdx.x];
  const Seg sg = segs[ch.seg];
  const float two_l = 2.0f * sg.l2;
  float denom = 1.0f, cn = 1.0f;
  if (clipnorm > 0.0f) {
    double sq = (sg.is_emb && emb_unagg) ? *emb_sq : seg_gsq[ch.seg];
    float nrm = (float)sqrt(sq);
    denom = fmaxf(nrm, clipnorm);
    cn = clipnorm;
  }
  const float scale = cn / denom;
  auto upd = [&](float& wv, float& mv, float& vv, float gr, float fac) {
    float gv = (gr + two_l * wv) * fac;         // fac: adaptive clipping factor of the element's unit (1 when off)
    gv = gv * cn / denom;                       // tf.clip_by_norm: t * clip_norm / max(||t||, clip_norm)
    mv += (gv - mv) * one_m_b1;
    vv += (gv * gv - vv) * one_m_b2;
    wv = wv - alpha * mv / (sqrtf(vv) + eps);
  };
  (void)scale;
  const int n4 = ((ch.off & 3) || agc_factor != nullptr) ? 0 : (ch.len >> 2);
  float4* w4 = reinterpret_cast<float4*>(w + ch.off);
  float4* m4 = reinterpret_cast<float4*>(m + ch.off);
  float4* v4 = reinterpret_cast<float4*>(v + ch.off);
  const float4* g4 = reinterpret_cast<const float4*>(g + ch.off);
#pragma unroll 2
  for (int i = threadIdx.x; i < n4; i += blockDim.x) {
    float4 wv = w4[i], mv = m4[i], vv = v4[i];
    const float4 gr = __ldg(g4 + i);
    upd(wv.x, mv.x, vv.x, gr.x, 1.0f); upd(wv.y, mv.y, vv.y, gr.y, 1.0f);
    upd(wv.z, mv.z, vv.z, gr.z, 1.0f); upd(wv.w, mv.w, vv.w, gr.w, 1.0f);
    m4[i] = mv; v4[i] = vv; w4[i] = wv;
  }
  for (int i = 4 * n4 + threadIdx.x; i < ch.len; i += blockDim.x) {
    const int64_t o = ch.off + i;
    float wv = w[o], mv = m[o], vv = v[o];
    float fac = 1.0f;
    if (agc_factor != nullptr) fac = agc_factor[sg.unit0 + (sg.cols > 0 ? (int)((o - sg.off) % sg.cols) : 0)];
    upd(wv, mv, vv, g[o], fac);
    m[o] = mv; v[o] = vv; w[o] = wv;
  }
}

// part: -1 = every variable; 0 = encoder variables (chunks [0, n_chunks_enc)); 1 = decoder variables.
// Data-parallel callers update the decoder part while the encoder gradients are still being all-reduced.
static void chunk_range(tnt_ctx* c, int part, int* first, int* count) {
  *first = part == 1 ? c->n_chunks_enc : 0;
  *count = part == 0 ? c->n_chunks_enc : (part == 1 ? c->n_chunks - c->n_chunks_enc : c->n_chunks);
}

void opt_norms(tnt_ctx* c, int with_grad, cudaStream_t st, int part, int first_part) {
  int first, count;
  chunk_range(c, part, &first, &count);
  // each part clears the sums of its own variables, so the two parts may run on different streams in any order
  (void)first_part;
  const int s0 = part == 1 ? c->n_segs_enc : 0;
  const int sn = part == 0 ? c->n_segs_enc : (part == 1 ? c->n_segs - c->n_segs_enc : c->n_segs);
  if (sn > 0) {
    cudaMemsetAsync(c->seg_gsq + s0, 0, sizeof(double) * sn, st);
    cudaMemsetAsync(c->seg_wsq + s0, 0, sizeof(double) * sn, st);
  }
  if (count > 0)
    opt_norm_kernel<<<count, 256, 0, st>>>(c->d_chunks + first, c->d_segs, c->w, (part == 0 && c->g_enc_src) ? c->g_enc_src : c->g,
                                           with_grad, c->seg_gsq, c->seg_wsq);
  c->launches += 1;
}

// L2 = sum_k lambda_k ||W_k||^2 -> metrics_raw[6]; after the norms of ALL parts
void opt_l2(tnt_ctx* c, cudaStream_t st) {
  if (c->l2_frozen) { c->l2_frozen = 0; return; }      // tnt_sam_restore already wrote the value this step reports
  double* mraw = wsT<double>(c, "metrics_raw");
  opt_l2_reduce_kernel<<<1, 256, 0, st>>>(c->n_segs, c->d_segs, c->seg_wsq, mraw + 6);
  c->launches += 1;
}

void opt_adam(tnt_ctx* c, cudaStream_t st, int part, int first_part) {
  const tnt_config& f = c->cfg;
  int first, count;
  chunk_range(c, part, &first, &count);
  if (first_part) c->adam_t += 1;                // one optimizer iteration per step, whatever the number of parts
  double t = (double)c->adam_t;
  double alpha = (double)f.lr * sqrt(1.0 - pow((double)f.beta2, t)) / (1.0 - pow((double)f.beta1, t));
  double* mraw = wsT<double>(c, "metrics_raw");
  const bool agc = f.agc_clip > 0.0f;
  if (count > 0)
    opt_adam_kernel<<<count, 256, 0, st>>>(c->d_chunks + first, c->d_segs, c->w, c->m, c->v,
                                           (part == 0 && c->g_enc_src) ? c->g_enc_src : c->g, c->seg_gsq,
                                           agc ? wsT<double>(c, "agc_emb_sq") : mraw + 7, f.emb_unaggregated_norm,
                                           f.clipnorm, (float)alpha, 1.0f - f.beta1, 1.0f - f.beta2, f.adam_eps,
                                           agc ? wsT<float>(c, "agc_factor") : nullptr, pd_abort_flag(c));
  c->launches += 1;
}

// Raw launchers for callers with their own chunk / segment tables (dense.cu)
void opt_launch_norm(const OptChunk* chunks, int n_chunks, const Seg* segs, const float* w, const float* g,
                     double* seg_gsq, double* seg_wsq, cudaStream_t st) {
  if (n_chunks > 0) opt_norm_kernel<<<n_chunks, 256, 0, st>>>(chunks, segs, w, g, 1, seg_gsq, seg_wsq);
}
void opt_launch_adam(const OptChunk* chunks, int n_chunks, const Seg* segs, float* w, float* m, float* v,
                     const float* g, const double* seg_gsq, float clipnorm, float alpha, float beta1, float beta2,
                     float eps, cudaStream_t st) {
  if (n_chunks > 0)
    opt_adam_kernel<<<n_chunks, 256, 0, st>>>(chunks, segs, w, m, v, g, seg_gsq, seg_gsq, 0, clipnorm, alpha,
                                              1.0f - beta1, 1.0f - beta2, eps, nullptr, nullptr);
}
