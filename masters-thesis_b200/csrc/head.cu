// Output head: TimeDistributed(Dense(256, LeakyReLU)) -> dropout -> TimeDistributed(Dense(V, softmax))
// (AttemptFour/Model/lc_NIC.py:139-159,261) and the loss / accuracy of train_step
// (lc_NIC.py:370-375,461-486; loss object main.py:107-110), plus their tape gradients.
// Rows are time-major: row = t*B + b.
#include "tnt_internal.h"

// d(pre) = d(inter_d) * mask * LeakyReLU'(pre)
__global__ void __launch_bounds__(256)
inter_bwd_kernel(int B, int T, int H, const float* __restrict__ pre, float* __restrict__ d, DropSite dout) {
  int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= (int64_t)B * T * H) return;
  int row = (int)(e / H), h = (int)(e % H);
  int t = row / B, b = row % B;
  d[e] = d[e] * drop_mult(dout, ((uint64_t)b * T + t) * H + h) * lrelu_grad_from_out(pre[e]);
}

// One block per row (t,b): softmax over V, Keras categorical_crossentropy on the probabilities
// (renormalise, clip to [1e-7, 1-1e-7], -log), argmax for accuracy, and d(logits) in place.
//   rowstat[row] = {loss, correct}
__global__ void __launch_bounds__(256)
softmax_ce_kernel(int B, int T, int V, float* __restrict__ logits, const int32_t* __restrict__ target,
                  float scale, int backward, float* __restrict__ probs_out, float* __restrict__ rowloss,
                  float* __restrict__ rowcorrect, int sparse_masked) {
  __shared__ float rs[32];
  __shared__ float rv[8];
  __shared__ int ri[8];
  const int row = blockIdx.x;
  const int t = row / B, b = row % B;
  float* z = logits + (size_t)row * V;
  int y = target ? target[(size_t)b * T + t] : 0;
  y = y < 0 ? 0 : (y >= V ? V - 1 : y);
  // max + first argmax
  float best = -INFINITY; int bi = 0x7fffffff;
  for (int v = threadIdx.x; v < V; v += blockDim.x) {
    float x = z[v];
    if (x > best || (x == best && v < bi)) { best = x; bi = v; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float ob = __shfl_xor_sync(0xffffffffu, best, o);
    int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { rv[warp] = best; ri[warp] = bi; }
  __syncthreads();
  best = rv[0]; bi = ri[0];
  for (int i = 1; i < 8; ++i)
    if (rv[i] > best || (rv[i] == best && ri[i] < bi)) { best = rv[i]; bi = ri[i]; }
  // softmax denominator
  float sum = 0.0f;
  for (int v = threadIdx.x; v < V; v += blockDim.x) sum += expf(z[v] - best);
  sum = block_sum(sum, rs);
  // S = sum of the fp32 probabilities (Keras renormalises: p / reduce_sum(p))
  float S = 0.0f;
  for (int v = threadIdx.x; v < V; v += blockDim.x) S += expf(z[v] - best) / sum;
  S = block_sum(S, rs);
  const float py = expf(z[y] - best) / sum;
  const float qy = py / S;
  const float qc = fminf(fmaxf(qy, TNT_CE_EPS), 1.0f - TNT_CE_EPS);
  // variant (ThinkAndTell/att_model.py:306-321): sparse CE on the LOGITS, times the pad mask (y != 0)
  const float keep = (y != 0) ? 1.0f : 0.0f;
  const float zy = z[y];
  if (threadIdx.x == 0 && target) {
    rowloss[row] = sparse_masked ? keep * (logf(sum) - (zy - best)) : -logf(qc);
    rowcorrect[row] = (bi == y) ? 1.0f : 0.0f;
  }
  const bool inside = (qy >= TNT_CE_EPS) && (qy <= 1.0f - TNT_CE_EPS);   // clip_by_value passes gradient
  float* pr = probs_out ? probs_out + ((size_t)b * T + t) * V : nullptr;   // caller layout (B,T,V)
  __syncthreads();                                                         // zy read before the row is overwritten
  for (int v = threadIdx.x; v < V; v += blockDim.x) {
    float p = expf(z[v] - best) / sum;
    if (pr) pr[v] = p;
    if (backward && sparse_masked) z[v] = keep * scale * (p - (v == y ? 1.0f : 0.0f));
    else
    if (backward) z[v] = inside ? scale * (p / S - (v == y ? 1.0f : 0.0f)) : 0.0f;
  }
}

// Same computation with the row held in registers (V <= 256 * CE_EPT): one read of the logits, one expf per
// element, one write of d(logits) -- the kernel above walks the row four times through L2.
#define CE_EPT 20
__global__ void __launch_bounds__(256)
softmax_ce_reg_kernel(int B, int T, int V, float* __restrict__ logits, const int32_t* __restrict__ target,
                      float scale, int backward, float* __restrict__ probs_out, float* __restrict__ rowloss,
                      float* __restrict__ rowcorrect, int sparse_masked) {
  __shared__ float rs[32];
  __shared__ float rv[8];
  __shared__ int ri[8];
  __shared__ float s_py, s_zy;
  const int row = blockIdx.x;
  const int t = row / B, b = row % B;
  float* z = logits + (size_t)row * V;
  int y = target ? target[(size_t)b * T + t] : 0;
  y = y < 0 ? 0 : (y >= V ? V - 1 : y);
  float e[CE_EPT];
  float best = -INFINITY; int bi = 0x7fffffff;
#pragma unroll
  for (int i = 0; i < CE_EPT; ++i) {
    const int v = threadIdx.x + 256 * i;
    e[i] = v < V ? z[v] : -INFINITY;
    if (v == y) s_zy = e[i];
    if (e[i] > best) { best = e[i]; bi = v; }          // v ascends per thread: first maximal index kept
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float ob = __shfl_xor_sync(0xffffffffu, best, o);
    int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { rv[warp] = best; ri[warp] = bi; }
  __syncthreads();
  best = rv[0]; bi = ri[0];
  for (int i = 1; i < 8; ++i)
    if (rv[i] > best || (rv[i] == best && ri[i] < bi)) { best = rv[i]; bi = ri[i]; }
  float sum = 0.0f;
#pragma unroll
  for (int i = 0; i < CE_EPT; ++i) {
    const int v = threadIdx.x + 256 * i;
    e[i] = v < V ? expf(e[i] - best) : 0.0f;
    sum += e[i];
  }
  sum = block_sum(sum, rs);
  // S = sum of the fp32 probabilities (Keras renormalises: p / reduce_sum(p))
  float S = 0.0f;
#pragma unroll
  for (int i = 0; i < CE_EPT; ++i) {
    e[i] = e[i] / sum;
    S += e[i];
    if (threadIdx.x + 256 * i == y) s_py = e[i];
  }
  S = block_sum(S, rs);                                   // (its barriers also publish s_py)
  const float qy = s_py / S;
  const float qc = fminf(fmaxf(qy, TNT_CE_EPS), 1.0f - TNT_CE_EPS);
  // variant (ThinkAndTell/att_model.py:306-321): sparse CE on the LOGITS, times the pad mask (y != 0)
  const float keep = (y != 0) ? 1.0f : 0.0f;
  if (threadIdx.x == 0 && target) {
    rowloss[row] = sparse_masked ? keep * (logf(sum) - (s_zy - best)) : -logf(qc);
    rowcorrect[row] = (bi == y) ? 1.0f : 0.0f;
  }
  const bool inside = (qy >= TNT_CE_EPS) && (qy <= 1.0f - TNT_CE_EPS);   // clip_by_value passes gradient
  float* pr = probs_out ? probs_out + ((size_t)b * T + t) * V : nullptr;   // caller layout (B,T,V)
#pragma unroll
  for (int i = 0; i < CE_EPT; ++i) {
    const int v = threadIdx.x + 256 * i;
    if (v < V) {
      if (pr) pr[v] = e[i];
      if (backward && sparse_masked) z[v] = keep * scale * (e[i] - (v == y ? 1.0f : 0.0f));
      else if (backward) z[v] = inside ? scale * (e[i] / S - (v == y ? 1.0f : 0.0f)) : 0.0f;
    }
  }
}

// sums of rowloss / rowcorrect over each half of the batch -> metrics_raw[0..3]; attention MSE -> [4..5]
__global__ void __launch_bounds__(256)
metrics_reduce_kernel(int B, int T, int R, int halves, const float* __restrict__ rowloss,
                      const float* __restrict__ rowcorrect, const float* __restrict__ attn_colsum,
                      double* __restrict__ mraw) {
  __shared__ double red[3][256];
  for (int h = 0; h < halves; ++h) {
    int nb = halves == 2 ? B / 2 : B, b0 = h * nb;
    double l = 0.0, cr = 0.0, am = 0.0;
    if (rowloss)
      for (int i = threadIdx.x; i < T * nb; i += blockDim.x) {
        int t = i / nb, b = b0 + i % nb;
        l += (double)rowloss[t * B + b]; cr += (double)rowcorrect[t * B + b];
      }
    for (int i = threadIdx.x; i < T * R; i += blockDim.x) {
      double d = 1.0 - (double)attn_colsum[(size_t)h * T * R + i];
      am += d * d;
    }
    red[0][threadIdx.x] = l; red[1][threadIdx.x] = cr; red[2][threadIdx.x] = am;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
      if ((int)threadIdx.x < s)
        for (int k = 0; k < 3; ++k) red[k][threadIdx.x] += red[k][threadIdx.x + s];
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      mraw[0 + h] = red[0][0]; mraw[2 + h] = red[1][0]; mraw[4 + h] = red[2][0] / (double)(T * R);
    }
    __syncthreads();
  }
}

// metrics_raw[4 + h] = mean over (t, r) of (1 - colsum[h][t][r])^2 (lc_NIC.py:365-367) -- data parallel: after the column
// sums have been all-reduced over the ranks
__global__ void __launch_bounds__(256)
attn_mse_kernel(int TR, int halves, const float* __restrict__ attn_colsum, double* __restrict__ mraw) {
  __shared__ double red[256];
  for (int h = 0; h < halves; ++h) {
    double am = 0.0;
    for (int i = threadIdx.x; i < TR; i += blockDim.x) {
      const double d = 1.0 - (double)attn_colsum[(size_t)h * TR + i];
      am += d * d;
    }
    red[threadIdx.x] = am;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
      if ((int)threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
      __syncthreads();
    }
    if (threadIdx.x == 0) mraw[4 + h] = red[0] / (double)TR;
    __syncthreads();
  }
}

// data parallel: global metric sums and attention column sums, then the attention metric from the global sums
void dec_reduce_metrics(tnt_ctx* c, cudaStream_t st) {
  if (dp_world(c) == 1) return;
  const tnt_config& f = c->cfg;
  const int halves = f.n_enc == 2 ? 2 : 1;
  // one-shot peer-memory reductions (dp.cu): a ring all-reduce of 16 doubles costs 68 us at 8 GPUs, in the middle of the
  // side stream's chain of weight-gradient GEMMs
  dp_allreduce_small(c, wsT<double>(c, "metrics_raw"), 16, st);
  if ((int64_t)halves * f.T * f.R <= 8192 && c->side && st == c->side)
    dp_allreduce_f32_small(c, wsT<float>(c, "attn_colsum"), halves * f.T * f.R, st);
  else
    dp_allreduce(c, wsT<float>(c, "attn_colsum"), (int64_t)halves * f.T * f.R, 0, 1, st);
  attn_mse_kernel<<<1, 256, 0, st>>>(f.T * f.R, halves, wsT<float>(c, "attn_colsum"), wsT<double>(c, "metrics_raw"));
  c->launches += 1;
}

__global__ void lrelu_inplace_kernel(int64_t n, float* x) {
  int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e < n) x[e] = lrelu(x[e]);
}

// one-hot (rows,V) -> ids, first maximal index (tf.math.argmax on the target, lc_NIC.py:481)
__global__ void __launch_bounds__(128)
onehot_to_ids_kernel(int V, const float* __restrict__ onehot, int32_t* __restrict__ ids) {
  __shared__ float rv[4];
  __shared__ int ri[4];
  const float* row = onehot + (size_t)blockIdx.x * V;
  float best = -INFINITY; int bi = 0x7fffffff;
  for (int v = threadIdx.x; v < V; v += blockDim.x) {
    float x = row[v];
    if (x > best || (x == best && v < bi)) { best = x; bi = v; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float ob = __shfl_xor_sync(0xffffffffu, best, o);
    int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
  }
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { rv[warp] = best; ri[warp] = bi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    best = rv[0]; bi = ri[0];
    for (int i = 1; i < 4; ++i)
      if (rv[i] > best || (rv[i] == best && ri[i] < bi)) { best = rv[i]; bi = ri[i]; }
    ids[blockIdx.x] = bi;
  }
}

// ------------------------------------------------------------------------------------------------
static inline int ceil_div64(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

void head_forward(tnt_ctx* c, int B, int training, uint64_t seed, cudaStream_t st, int want_logits) {
  const tnt_config& f = c->cfg;
  const int TB = f.T * B;
  float* pre = wsT<float>(c, "inter_pre");
  float* d = wsT<float>(c, "inter_d");
  TNT_TRACE(c, want_logits ? "head forward: dense_inter gemm_pk (fused act + pack) + dense_out gemm_pk" : "head forward: dense_inter gemm_pk (fused LeakyReLU + dropout + pack)", st);
  DropSite dout = make_drop(f.dropout_out, training, seed, SITE_OUT, 0);
  // inter_pre = hout . W + b ; inter_d = dropout(LeakyReLU(inter_pre)) and its packed operand, all from one epilogue
  gemm_tc_epi(c, 0, false, TB, f.H, f.U, wsT<float>(c, "hout"), f.U, c->w + c->o_iW, f.H, pre, f.H, c->w + c->o_ib,
              TC_EPI_INTER_FWD, d, nullptr, dout, B, f.T, d, st);
  if (want_logits)      // tnt_forward hands the probabilities to the caller; the steps use head_fused_loss instead
    gemm_tc(c, 1, 1, false, false, TB, f.V, f.H, d, f.H, c->w + c->o_oW, f.V, wsT<float>(c, "logits"), f.V, c->w + c->o_ob, 0, st);
}

void head_decode_step(tnt_ctx* c, int B, const float* h, float* inter, float* logits, cudaStream_t st) {
  const tnt_config& f = c->cfg;
  gemm_tc(c, 0, 1, false, false, B, f.H, f.U, h, f.U, c->w + c->o_iW, f.H, inter, f.H, c->w + c->o_ib, 0, st);
  lrelu_inplace_kernel<<<ceil_div64((int64_t)B * f.H, 256), 256, 0, st>>>((int64_t)B * f.H, inter);
  c->launches += 1;
  gemm_tc(c, 0, 1, false, false, B, f.V, f.H, inter, f.H, c->w + c->o_oW, f.V, logits, f.V, c->w + c->o_ob, 0, st);
}

void head_loss(tnt_ctx* c, const int32_t* target, int B, int B_global, int backward, float* probs_out,
               cudaStream_t st) {
  const tnt_config& f = c->cfg;
  const int TB = f.T * B;
  // loss = mean over (b,t) of the GLOBAL batch; ms2's (CE_A + CE_B)/2 with equal halves is the same mean
  const int sparse_masked = (f.variant & TNT_VARIANT_MASKED_SPARSE_CE) ? 1 : 0;
  // mean over the batch, then sum over steps / T (lc_NIC.py:375); the variant divides by the T+1 tokens of its
  // target (att_model.py:262: 15 loss terms / target.shape[1] = 16)
  float scale = 1.0f / ((float)B_global * (float)(sparse_masked ? f.T + 1 : f.T));
  { TNT_TRACE(c, "softmax_ce_reg_kernel", st);
  if (f.V <= 256 * CE_EPT)
    softmax_ce_reg_kernel<<<TB, 256, 0, st>>>(B, f.T, f.V, wsT<float>(c, "logits"), target, scale, backward, probs_out,
                                              wsT<float>(c, "rowloss"), wsT<float>(c, "rowcorrect"), sparse_masked);
  else
    softmax_ce_kernel<<<TB, 256, 0, st>>>(B, f.T, f.V, wsT<float>(c, "logits"), target, scale, backward, probs_out,
                                          wsT<float>(c, "rowloss"), wsT<float>(c, "rowcorrect"), sparse_masked);
  }
  c->variant_head = "gemm_pk + softmax_ce_reg_kernel (probabilities handed to the caller)";
  head_metrics(c, target != nullptr, B, backward, st);
}

// the reported metrics are off the critical path: side stream
void head_metrics(tnt_ctx* c, int have_target, int B, int backward, cudaStream_t st) {
  const tnt_config& f = c->cfg;
  cudaStream_t ss = c->side ? c->side : st;
  if (c->side) {
    cudaEventRecord(c->ev_fork, st);
    cudaStreamWaitEvent(ss, c->ev_fork, 0);
  }
  dec_attn_colsum(c, B, ss);
  metrics_reduce_kernel<<<1, 256, 0, ss>>>(B, f.T, f.R, f.n_enc == 2 ? 2 : 1,
                                           have_target ? wsT<float>(c, "rowloss") : nullptr, wsT<float>(c, "rowcorrect"),
                                           wsT<float>(c, "attn_colsum"), wsT<double>(c, "metrics_raw"));
  c->launches += 2;
  if (!backward) dec_reduce_metrics(c, ss);       // with a backward pass the embedding's clip norm joins the sums first
  if (c->side) {
    cudaEventRecord(c->ev_join, ss);
    c->side_pending = 1;
  }
}

void side_join(tnt_ctx* c, cudaStream_t st) {
  if (c->side_pending) {
    cudaStreamWaitEvent(st, c->ev_join, 0);
    c->side_pending = 0;
  }
}

void head_backward(tnt_ctx* c, int B, uint64_t seed, cudaStream_t st) {
  const tnt_config& f = c->cfg;
  const int TB = f.T * B;
  float* dl = wsT<float>(c, "logits");          // holds d(logits) after head_loss(backward=1)
  float* d = wsT<float>(c, "inter_d");
  float* pre = wsT<float>(c, "inter_pre");
  float* dinter = wsT<float>(c, "dinter");
  // critical path: d(inter) -> d(hout), which the recurrence gradient waits for
  { TNT_TRACE(c, "head backward (critical): 2 gemm_pk + mask/LeakyReLU' + pack", st);
  // A = d logits: the packed operand head_fused_loss wrote (cached under the key of the "logits" workspace).  The
  // point-wise tail is NOT fused into this GEMM's epilogue: with nine K-splits the last CTA of each of the 16 tiles would
  // run it alone, one warp per scheduler (measured 56.9 us against 31.8 + 5.2 + 7.3 us for the three kernels).
  gemm_tc(c, 1, 1, false, true, TB, f.H, f.V, dl, f.V, c->w + c->o_oW, f.V, dinter, f.H, nullptr, 0, st);
  DropSite dout = make_drop(f.dropout_out, 1, seed, SITE_OUT, 0);
  inter_bwd_kernel<<<ceil_div64((int64_t)TB * f.H, 256), 256, 0, st>>>(B, f.T, f.H, pre, dinter, dout);
  c->launches += 1;
  gemm_tc(c, 0, 1, false, true, TB, f.U, f.H, dinter, f.H, c->w + c->o_iW, f.H, wsT<float>(c, "dhout"), f.U, nullptr, 0, st);
  }
  // off the critical path: the weight gradients of dense_out / dense_inter run on the side stream, next to the
  // persistent recurrence kernel (which leaves 20 SMs free)
  cudaStream_t ss = c->side ? c->side : st;
  if (c->side) {
    cudaEventRecord(c->ev_fork, st);
    cudaStreamWaitEvent(ss, c->ev_fork, 0);
    tc_set_lane(c, 1);
  }
  TraceScope* ts = new TraceScope(c, "head weight gradients: 2 gemm_pk + packs + colsums (side stream)", ss);
  {
    const TcPackReq reqs[] = {
        {d, TB, f.H, f.H, true},                                // A of d(oW) = inter^T @ d(logits)
        {dl, TB, f.V, f.V, true},                               // B of d(oW)
        {wsT<float>(c, "hout"), TB, f.U, f.U, true},            // A of d(iW) = hout^T @ d(inter)
        {dinter, TB, f.H, f.H, true},                           // B of d(iW)
    };
    tc_prepack(c, reqs, 4, ss);
  }
  gemm_tc(c, 1, 1, true, false, f.H, f.V, TB, d, f.H, dl, f.V, c->g + c->o_oW, f.V, nullptr, 0, ss);
  gemm_tc(c, 1, 1, true, false, f.U, f.H, TB, wsT<float>(c, "hout"), f.U, dinter, f.H, c->g + c->o_iW, f.H, nullptr, 0, ss);
  // d(bias of dense_out) = column sums of d logits: the fused kernel left one partial row per m-tile
  const ColsumReq cs[] = {{wsT<float>(c, "dob_part"), c->g + c->o_ob, (TB + 127) / 128, f.V, f.V}, {dinter, c->g + c->o_ib, TB, f.H, f.H}};
  colsum_multi(c, cs, 2, ss);
  // data parallel: the head's gradients are final here, long before the recurrence gradient is: their all-reduce
  // (dense_inter + dense_out kernels and biases, contiguous at the end of g) hides completely behind it
  dp_allreduce(c, c->g + c->o_iW, c->n_param - c->o_iW, 0, 1, ss);
  delete ts;
  if (c->side) {
    tc_set_lane(c, 0);
    cudaEventRecord(c->ev_head, ss);          // head gradients final: a data-parallel caller may reduce them now
    cudaEventRecord(c->ev_join, ss);
    c->side_pending = 1;
  }
}

void launch_onehot_to_ids(const float* onehot, int64_t rows, int V, int32_t* ids, cudaStream_t st) {
  if (rows > 0) onehot_to_ids_kernel<<<(unsigned)rows, 128, 0, st>>>(V, onehot, ids);
}
