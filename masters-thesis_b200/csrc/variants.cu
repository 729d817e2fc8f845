// Step variants that live in the reference's model files next to the default train_step:
//   * the attention loss as part of the differentiated loss (AttemptFour/Model/lc_NIC.py:365-367 with :384 enabled;
//     element-wise form of train_step_sam, lc_NIC.py:751-752,774);
//   * adaptive gradient clipping (AttemptFour/Model/agc.py:20-38, call site lc_NIC.py:388), unit-wise norms including
//     the IndexedSlices branch of the embedding;
//   * the two weight moves of the sharpness-aware step (train_step_sam, lc_NIC.py:713-838).
#include "tnt_internal.h"

// ------------------------------------------------------------------------------------------------
// attention loss gradient on alpha
// ------------------------------------------------------------------------------------------------
// mode 1: L = mean_{t,r} (1 - sum_b alpha[t,b,r])^2          -> dL/d alpha[t,b,r] = -2 (1 - colsum[t,r]) / (T R)
// mode 2: L = mean_{t,b,r} (1 - alpha[t,b,r])^2              -> dL/d alpha[t,b,r] = -2 (1 - alpha) / (T B_global R)
// Se[t][b] = sum_r alpha dL/d alpha (the softmax tape gradient needs it; decoder_persist.cu).  One block per (t, b).
__global__ void __launch_bounds__(128)
attn_loss_grad_kernel(int mode, int T, int B, int R, int B_global, const float* __restrict__ alpha,
                      const float* __restrict__ colsum, float* __restrict__ e, float* __restrict__ Se,
                      double* __restrict__ elem_sq) {
  __shared__ float rs[32];
  const int t = blockIdx.x / B, b = blockIdx.x % B;
  const float* al = alpha + ((size_t)t * B + b) * R;
  float* er = e + ((size_t)t * B + b) * R;
  const float k1 = -2.0f / ((float)T * (float)R);
  const float k2 = -2.0f / ((float)T * (float)B_global * (float)R);
  float s = 0.0f, sq = 0.0f;
  for (int r = threadIdx.x; r < R; r += blockDim.x) {
    const float a = al[r];
    const float d = 1.0f - a;
    const float g = mode == 1 ? k1 * (1.0f - colsum[(size_t)t * R + r]) : k2 * d;
    er[r] = g;
    s = fmaf(a, g, s);
    sq = fmaf(d, d, sq);
  }
  s = block_sum(s, rs);
  sq = block_sum(sq, rs);
  if (threadIdx.x == 0) {
    Se[(size_t)t * B + b] = s;
    if (elem_sq != nullptr) atomicAdd(elem_sq, (double)sq);
  }
}

// alpha[T][B][R] is final (recurrence forward done); colsum is computed here on `st` when the batch form needs it
void dec_attn_loss_grad(tnt_ctx* c, int B, int B_global, cudaStream_t st) {
  const tnt_config& f = c->cfg;
  if (!f.attn_loss) return;
  double* mraw = wsT<double>(c, "metrics_raw");
  if (f.attn_loss == TNT_ATTN_LOSS_BATCH) dec_attn_colsum(c, B, st);
  attn_loss_grad_kernel<<<f.T * B, 128, 0, st>>>(f.attn_loss, f.T, B, f.R, B_global, wsT<float>(c, "alpha"),
                                                 wsT<float>(c, "attn_colsum"), wsT<float>(c, "attn_e"),
                                                 wsT<float>(c, "attn_Se"),
                                                 f.attn_loss == TNT_ATTN_LOSS_ELEM ? mraw + 8 : nullptr);
  c->launches += 1;
}

// ------------------------------------------------------------------------------------------------
// adaptive gradient clipping (agc.py:20-38)
// ------------------------------------------------------------------------------------------------
// A "unit" is what unitwise_norm reduces over: one column of a 2-D variable (axis 0, keepdims) or a whole 1-D
// variable.  One thread per unit walks its rows (neighbouring threads = neighbouring columns: coalesced).
__global__ void __launch_bounds__(128)
agc_unit_norm_kernel(int n_units, const AgcUnit* __restrict__ units, const Seg* __restrict__ segs,
                     const float* __restrict__ w, const float* __restrict__ g, float* __restrict__ unit_gsq,
                     float* __restrict__ unit_wsq) {
  const int u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= n_units) return;
  const AgcUnit un = units[u];
  const float two_l = 2.0f * segs[un.seg].l2;
  float gs = 0.0f, wsq = 0.0f;
  const float* wp = w + un.base;
  const float* gp = g + un.base;
  for (int64_t r = 0; r < un.rows; ++r) {
    const float wv = wp[r * un.stride];
    const float gv = gp[r * un.stride] + two_l * wv;         // the tape gradient carries the L2 term
    wsq = fmaf(wv, wv, wsq);
    gs = fmaf(gv, gv, gs);
  }
  unit_gsq[u] = gs;
  unit_wsq[u] = wsq;
}

// factor[u] = 1 if ||g_u|| < max(||w_u||, eps) * clip else max_norm / max(||g_u||, 1e-6)   (agc.py:31-35)
// For the embedding (IndexedSlices) the gradient norm is taken over the un-aggregated values (agc.py:24-29):
// emb_colsq[col] = sum over the (B*T) slices of value^2.  Also the clipnorm inputs AFTER this clipping: the
// per-variable squared norm of the clipped gradient, and the embedding's un-aggregated one.
__global__ void __launch_bounds__(128)
agc_factor_kernel(int n_units, const AgcUnit* __restrict__ units, const Seg* __restrict__ segs,
                  const float* __restrict__ unit_gsq, const float* __restrict__ unit_wsq,
                  const float* __restrict__ emb_colsq, float clip, float eps, float* __restrict__ factor,
                  double* __restrict__ seg_gsq, double* __restrict__ emb_sq_clipped) {
  const int u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= n_units) return;
  const AgcUnit un = units[u];
  const Seg sg = segs[un.seg];
  const float pn = sqrtf(unit_wsq[u]);
  const float max_norm = fmaxf(pn, eps) * clip;
  const float gsq_for_norm = sg.is_emb ? emb_colsq[u - sg.unit0] : unit_gsq[u];
  const float gn = sqrtf(gsq_for_norm);
  const float fac = gn < max_norm ? 1.0f : max_norm / fmaxf(gn, 1e-6f);
  factor[u] = fac;
  atomicAdd(&seg_gsq[un.seg], (double)(fac * fac * unit_gsq[u]));
  if (sg.is_emb) atomicAdd(emb_sq_clipped, (double)(fac * fac * emb_colsq[u - sg.unit0]));
}

void opt_agc(tnt_ctx* c, cudaStream_t st) {
  const tnt_config& f = c->cfg;
  const int nb = (c->n_units + 127) / 128;
  // after opt_norms (which leaves the L2 sums in seg_wsq): seg_gsq is rewritten with the norms of the CLIPPED gradients
  cudaMemsetAsync(c->seg_gsq, 0, sizeof(double) * c->n_segs, st);
  cudaMemsetAsync(wsT<double>(c, "agc_emb_sq"), 0, sizeof(double), st);
  agc_unit_norm_kernel<<<nb, 128, 0, st>>>(c->n_units, c->d_units, c->d_segs, c->w, c->g, wsT<float>(c, "agc_gsq"),
                                           wsT<float>(c, "agc_wsq"));
  agc_factor_kernel<<<nb, 128, 0, st>>>(c->n_units, c->d_units, c->d_segs, wsT<float>(c, "agc_gsq"),
                                        wsT<float>(c, "agc_wsq"), wsT<float>(c, "emb_colsq"), f.agc_clip, f.agc_eps,
                                        wsT<float>(c, "agc_factor"), c->seg_gsq, wsT<double>(c, "agc_emb_sq"));
  c->launches += 2;
}

// ------------------------------------------------------------------------------------------------
// sharpness-aware step: w += e, e = (g + 2 lambda w) * rho / (global_norm + eps); later w -= e
// ------------------------------------------------------------------------------------------------
__global__ void sam_scale_kernel(int nseg, const Seg* __restrict__ segs, const double* __restrict__ seg_gsq,
                                 const double* __restrict__ emb_sq, float rho, float eps, float* __restrict__ scale) {
  __shared__ double red[256];
  double s = 0.0;
  for (int i = threadIdx.x; i < nseg; i += blockDim.x) s += segs[i].is_emb ? *emb_sq : seg_gsq[i];
  red[threadIdx.x] = s;
  __syncthreads();
  for (int k = 128; k > 0; k >>= 1) {
    if ((int)threadIdx.x < k) red[threadIdx.x] += red[threadIdx.x + k];
    __syncthreads();
  }
  if (threadIdx.x == 0) *scale = rho / ((float)sqrt(red[0]) + eps);
}

__global__ void __launch_bounds__(256)
sam_move_kernel(const OptChunk* __restrict__ chunks, const Seg* __restrict__ segs, float* __restrict__ w,
                float* __restrict__ g, float* __restrict__ e, const float* __restrict__ scale, int restore) {
  const OptChunk ch = chunks[blockIdx.x];
  const float two_l = 2.0f * segs[ch.seg].l2;
  const float s = restore ? 0.0f : *scale;
  for (int i = threadIdx.x; i < ch.len; i += blockDim.x) {
    const int64_t o = ch.off + i;
    // the second tape gradient was taken at w + e, L2 term 2 lambda (w + e) included; the optimizer adds 2 lambda w
    // of the RESTORED weights, so the remainder 2 lambda e goes into g here
    if (restore) { const float ev = e[o]; w[o] -= ev; g[o] += two_l * ev; }
    else {
      const float wv = w[o];
      const float ev = (g[o] + two_l * wv) * s;
      e[o] = ev;
      w[o] = wv + ev;
    }
  }
}

int sam_perturb(tnt_ctx* c, float rho, float eps, cudaStream_t st) {
  if (!ws(c, "sam_e")) {
    void* p = nullptr;
    if (cudaMalloc(&p, sizeof(float) * (size_t)c->n_param) != cudaSuccess) { cudaGetLastError(); return TNT_ERR_CUDA; }
    c->bufs["sam_e"] = std::make_pair(p, (int64_t)sizeof(float) * c->n_param);
    if (cudaMalloc(&p, 16) != cudaSuccess) { cudaGetLastError(); return TNT_ERR_CUDA; }
    c->bufs["sam_scale"] = std::make_pair(p, (int64_t)16);
  }
  side_join(c, st);                                  // every gradient of the first pass is final
  opt_norms(c, 1, st);                               // per-variable ||g + 2 lambda w||^2
  sam_scale_kernel<<<1, 256, 0, st>>>(c->n_segs, c->d_segs, c->seg_gsq, wsT<double>(c, "metrics_raw") + 7, rho, eps,
                                      wsT<float>(c, "sam_scale"));
  sam_move_kernel<<<c->n_chunks, 256, 0, st>>>(c->d_chunks, c->d_segs, c->w, c->g, wsT<float>(c, "sam_e"),
                                               wsT<float>(c, "sam_scale"), 0);
  c->launches += 2;
  return TNT_OK;
}

int sam_restore(tnt_ctx* c, cudaStream_t st) {
  if (!ws(c, "sam_e")) return TNT_ERR_STATE;
  side_join(c, st);
  // the step reports the L2 term of the SECOND tape, taken at the perturbed weights (lc_NIC.py:818): compute it now
  // and keep the optimizer phase from overwriting it with the restored weights' value
  opt_norms(c, 0, st);
  opt_l2(c, st);
  c->l2_frozen = 1;
  sam_move_kernel<<<c->n_chunks, 256, 0, st>>>(c->d_chunks, c->d_segs, c->w, c->g, wsT<float>(c, "sam_e"),
                                               wsT<float>(c, "sam_scale"), 1);
  c->launches += 1;
  return TNT_OK;
}
