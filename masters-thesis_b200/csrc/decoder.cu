// Attention-LSTM decoder loop: lc_NIC.call_attention (AttemptFour/Model/lc_NIC.py:233-258),
// Attention.call (AttemptFour/Model/attention.py:25-44), the Keras LSTM step (lc_NIC.py:118-124,255)
// and their tape gradients; greedy_predict_attention's loop (lc_NIC.py:596-632).
//
// Layout: every recurrent buffer is time-major: xtext[T][B][E], Zx[T][B][4U] (gate order i,f,c~,o as
// Keras stores the LSTM kernel), xctx[T][B][G], alpha[T][B][R], qact[T][B][A], Hall/Call[T+1][B][U]
// (slot 0 = initial state), gates[T][B][4U] (post-activation), hout[T][B][U] (after dropout_lstm).
#include "tnt_internal.h"

// ------------------------------------------------------------------------------------------------
// embedding gather (+ dropout_text, + the text part of the LSTM input-dropout mask)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
emb_fwd_kernel(int B, int T, int E, int V, int GE, int G, const int32_t* __restrict__ cap, int cap_stride,
               int t_offset, const float* __restrict__ emb, float* __restrict__ xtext, DropSite dtext,
               DropSite dlin) {
  // grid: (T_local * B) rows; row = tl*B + b; global step t = t_offset + tl
  int row = blockIdx.x;
  int tl = row / B, b = row % B;
  int t = t_offset + tl;
  int id = cap[(size_t)b * cap_stride + tl];
  id = id < 0 ? 0 : (id >= V ? V - 1 : id);
  DropSite dl = dlin; dl.step = (uint32_t)t;
  for (int e = threadIdx.x; e < E; e += blockDim.x) {
    float v = emb[(size_t)id * E + e];
    v *= drop_mult(dtext, ((uint64_t)b * T + t) * E + e);
    v *= drop_mult(dl, (uint64_t)b * GE + G + e);
    xtext[(size_t)row * E + e] = v;
  }
}

// ------------------------------------------------------------------------------------------------
// one attention step for sample b = blockIdx.x
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
att_fwd_kernel(int B, int R, int G, int A, int U, int GE, const float* __restrict__ hprev,
               const float* __restrict__ Pf, const float* __restrict__ feats, const float* __restrict__ W2,
               const float* __restrict__ b2, const float* __restrict__ Vw, const float* __restrict__ bV,
               float* __restrict__ alpha_t, float* __restrict__ qact_t, float* __restrict__ xctx_t,
               float* __restrict__ s_out, DropSite dattn, DropSite dlin, float att_slope) {
  extern __shared__ float sc[];                 // R scores / weights
  __shared__ float red[8][33];
  __shared__ float rs[32];
  const int b = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // q = LeakyReLU(h @ W2 + b2)                                      (attention.py:32, lc_NIC.py:98)
  // The three loops below are chains of dependent L2 round trips when left rolled (one CTA per sample, ~150 of them
  // in a row: 75 us per CTA, 248 us per decode step at B = 4096).  Each issues the loads of 4 (8) iterations together
  // and then does the arithmetic in the original order, so results are bit-identical.
  float acc = 0.0f;
  if (lane < A) {
    const float* hp = hprev + (size_t)b * U;
    int u = warp;
    for (; u + 56 < U; u += 64) {
      float hv[8], wv[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) { hv[i] = hp[u + 8 * i]; wv[i] = W2[(size_t)(u + 8 * i) * A + lane]; }
#pragma unroll
      for (int i = 0; i < 8; ++i) acc = fmaf(hv[i], wv[i], acc);
    }
    for (; u < U; u += 8) acc = fmaf(hp[u], W2[(size_t)u * A + lane], acc);
  }
  red[warp][lane] = acc;
  __syncthreads();
  float q = 0.0f;
  if (lane < A) {
    q = b2[lane];
#pragma unroll
    for (int i = 0; i < 8; ++i) q += red[i][lane];
    q = lrelu_s(q, att_slope);
  }
  if (warp == 0 && lane < A) qact_t[(size_t)b * A + lane] = q;
  const float vv = lane < A ? Vw[lane] : 0.0f;
  const float bv = bV[0];
  // score_r = tanh(Pf_r + q) (dropout) . V + bV                     (attention.py:30-38)
  for (int r0 = warp; r0 < R; r0 += 32) {
    float pf[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = r0 + 8 * i;
      pf[i] = (lane < A && r < R) ? Pf[((size_t)b * R + r) * A + lane] : 0.0f;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = r0 + 8 * i;
      if (r >= R) break;
      size_t o = ((size_t)b * R + r) * A + lane;
      float s = 0.0f;
      if (lane < A) {
        s = tanhf(pf[i] + q);
        s *= drop_mult(dattn, (uint64_t)o);
        if (s_out) s_out[o] = s;
      }
      float e = warp_sum(s * vv) + bv;
      if (lane == 0) sc[r] = e;
    }
  }
  __syncthreads();
  // softmax over regions                                            (attention.py:40)
  float mx = -INFINITY;
  for (int r = threadIdx.x; r < R; r += blockDim.x) mx = fmaxf(mx, sc[r]);
  mx = block_max(mx, rs);
  float sum = 0.0f;
  for (int r = threadIdx.x; r < R; r += blockDim.x) { float ex = expf(sc[r] - mx); sc[r] = ex; sum += ex; }
  sum = block_sum(sum, rs);
  for (int r = threadIdx.x; r < R; r += blockDim.x) {
    float a = sc[r] / sum;
    sc[r] = a;
    alpha_t[(size_t)b * R + r] = a;
  }
  __syncthreads();
  // context = sum_r alpha_r * f_r                                   (attention.py:42)
  float cx = 0.0f;
  if (lane < G) {
    const float* fp = feats + (size_t)b * R * G + lane;
    int r = warp;
    for (; r + 24 < R; r += 32) {
      float fv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) fv[i] = fp[(size_t)(r + 8 * i) * G];
#pragma unroll
      for (int i = 0; i < 4; ++i) cx = fmaf(sc[r + 8 * i], fv[i], cx);
    }
    for (; r < R; r += 8) cx = fmaf(sc[r], fp[(size_t)r * G], cx);
  }
  __syncthreads();
  red[warp][lane] = cx;
  __syncthreads();
  if (warp == 0 && lane < G) {
    float t = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][lane];
    xctx_t[(size_t)b * G + lane] = t * drop_mult(dlin, (uint64_t)b * GE + lane);
  }
}

// ------------------------------------------------------------------------------------------------
// LSTM cell pointwise part: z (already x@W + h@U + b) -> gates, c, h        (Keras LSTMCell.call)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
lstm_pointwise_kernel(int B, int U, const float* __restrict__ z, const float* __restrict__ cprev,
                      float* __restrict__ gates, float* __restrict__ cnew, float* __restrict__ hnew,
                      float* __restrict__ hout, DropSite dout) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= B * U) return;
  int b = e / U, u = e % U;
  const float* zr = z + (size_t)b * 4 * U;
  float i = sigmoidf_(zr[u]), f = sigmoidf_(zr[U + u]), g = tanhf(zr[2 * U + u]), o = sigmoidf_(zr[3 * U + u]);
  float c = f * cprev[e] + i * g;
  float h = o * tanhf(c);
  if (gates) {
    float* gr = gates + (size_t)b * 4 * U;
    gr[u] = i; gr[U + u] = f; gr[2 * U + u] = g; gr[3 * U + u] = o;
  }
  cnew[e] = c; hnew[e] = h;
  if (hout) hout[e] = h * drop_mult(dout, (uint64_t)e);
}

// attn_colsum[t][r] = sum_b alpha[t][b][r] over rows [b0, b0+nb)             (lc_NIC.py:365)
__global__ void attn_colsum_kernel(int B, int R, int b0, int nb, const float* __restrict__ alpha,
                                   float* __restrict__ out) {
  int t = blockIdx.x;
  for (int r = threadIdx.x; r < R; r += blockDim.x) {
    float s = 0.0f;
    for (int b = b0; b < b0 + nb; ++b) s += alpha[((size_t)t * B + b) * R + r];
    out[(size_t)t * R + r] = s;
  }
}

// (T,B,R) time-major -> caller layout is the same (T,B,R,1): plain copy done with cudaMemcpyAsync.

// ------------------------------------------------------------------------------------------------
// backward kernels
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
lstm_bwd_pointwise_kernel(int B, int U, const float* __restrict__ dhout_t, const float* __restrict__ dh_rec,
                          float* __restrict__ dc, const float* __restrict__ gates, const float* __restrict__ c_t,
                          const float* __restrict__ c_prev, float* __restrict__ dZ, DropSite dout) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= B * U) return;
  int b = e / U, u = e % U;
  const float* gr = gates + (size_t)b * 4 * U;
  float i = gr[u], f = gr[U + u], g = gr[2 * U + u], o = gr[3 * U + u];
  float dh = dhout_t[e] * drop_mult(dout, (uint64_t)e) + dh_rec[e];
  float tc = tanhf(c_t[e]);
  float dcc = dc[e] + dh * o * (1.0f - tc * tc);
  float* dz = dZ + (size_t)b * 4 * U;
  dz[u] = dcc * g * i * (1.0f - i);
  dz[U + u] = dcc * c_prev[e] * f * (1.0f - f);
  dz[2 * U + u] = dcc * i * (1.0f - g * g);
  dz[3 * U + u] = dh * tc * o * (1.0f - o);
  dc[e] = dcc * f;
}

__global__ void __launch_bounds__(256)
att_bwd_kernel(int B, int R, int G, int A, int U, int GE, const float* __restrict__ dxctx,
               const float* __restrict__ alpha_t, const float* __restrict__ qact_t,
               const float* __restrict__ Pf, const float* __restrict__ feats, const float* __restrict__ W2,
               const float* __restrict__ Vw, float* __restrict__ dfeats, float* __restrict__ dPf,
               float* __restrict__ dqpre_t, float* __restrict__ dVpart_t, float* __restrict__ dbVpart_t,
               float* __restrict__ dh_rec, DropSite dattn, DropSite dlin, float att_slope,
               const float* __restrict__ attn_e_t /* dL/d alpha of the attention loss [B][R] or nullptr */) {
  extern __shared__ float sm[];                 // alpha[R], dalpha[R]
  float* sal = sm;
  float* sda = sm + R;
  __shared__ float red[2][8][33];
  __shared__ float rs[32];
  __shared__ float dqs[32];
  __shared__ float rb[8];
  const int b = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float dctx = lane < G ? dxctx[(size_t)b * G + lane] * drop_mult(dlin, (uint64_t)b * GE + lane) : 0.0f;
  const float q = lane < A ? qact_t[(size_t)b * A + lane] : 0.0f;
  const float vv = lane < A ? Vw[lane] : 0.0f;
  for (int r = warp; r < R; r += 8) {
    size_t o = ((size_t)b * R + r) * G + lane;
    float f = lane < G ? feats[o] : 0.0f;
    float da = warp_sum(f * dctx);
    if (attn_e_t != nullptr) da += attn_e_t[(size_t)b * R + r];
    float al = alpha_t[(size_t)b * R + r];
    if (lane == 0) { sal[r] = al; sda[r] = da; }
    if (lane < G) dfeats[o] += al * dctx;
  }
  __syncthreads();
  float part = 0.0f;
  for (int r = threadIdx.x; r < R; r += blockDim.x) part += sal[r] * sda[r];
  const float S = block_sum(part, rs);
  float dq = 0.0f, dV = 0.0f, dbv = 0.0f;
  for (int r = warp; r < R; r += 8) {
    float de = sal[r] * (sda[r] - S);
    dbv += de;
    if (lane < A) {
      size_t o = ((size_t)b * R + r) * A + lane;
      float s = tanhf(Pf[o] + q);
      float m = drop_mult(dattn, (uint64_t)o);
      dV = fmaf(de, s * m, dV);
      float dpre = de * vv * m * (1.0f - s * s);
      dPf[o] += dpre;
      dq += dpre;
    }
  }
  red[0][warp][lane] = dq; red[1][warp][lane] = dV;
  if (lane == 0) rb[warp] = dbv;
  __syncthreads();
  if (warp == 0) {
    float tq = 0.0f, tv = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { tq += red[0][i][lane]; tv += red[1][i][lane]; }
    float dqp = tq * lrelu_grad_s(q, att_slope);
    if (lane < A) {
      dqpre_t[(size_t)b * A + lane] = dqp;
      dVpart_t[(size_t)b * A + lane] = tv;
    }
    dqs[lane] = lane < A ? dqp : 0.0f;
    if (lane == 0) {
      float t = 0.0f;
#pragma unroll
      for (int i = 0; i < 8; ++i) t += rb[i];
      dbVpart_t[b] = t;
    }
  }
  __syncthreads();
  for (int u = threadIdx.x; u < U; u += blockDim.x) {
    float a = 0.0f;
    for (int k = 0; k < A; ++k) a = fmaf(dqs[k], W2[(size_t)u * A + k], a);
    dh_rec[(size_t)b * U + u] += a;
  }
}

// After the time loop: gradient through W1 (attention.py:31): dW1, db1 partials and dfeats += dPfpre @ W1^T
__global__ void __launch_bounds__(256)
att_w1_bwd_kernel(int64_t rows, int G, int A, const float* __restrict__ Pf, const float* __restrict__ dPf,
                  const float* __restrict__ feats, const float* __restrict__ W1, float* __restrict__ dfeats,
                  float* __restrict__ partial /* [grid][32*32+32] */,
                  // persistent path: d feats = sum_t alpha_t d ctx_t is folded in here (attention.py:42)
                  int T, int B, int R, const float* __restrict__ alpha, const float* __restrict__ dxctx,
                  float att_slope) {
  __shared__ float W1s[32][33];
  __shared__ float red[8][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int e = threadIdx.x; e < 32 * 32; e += blockDim.x) {
    int g = e >> 5, a = e & 31;
    W1s[g][a] = (g < G && a < A) ? W1[g * A + a] : 0.0f;
  }
  __syncthreads();
  float acc[32];
#pragma unroll
  for (int g = 0; g < 32; ++g) acc[g] = 0.0f;
  float db = 0.0f;
  for (int64_t row = (int64_t)blockIdx.x * 8 + warp; row < rows; row += (int64_t)gridDim.x * 8) {
    float dp = 0.0f, f = 0.0f;
    if (lane < A) dp = dPf[row * A + lane] * lrelu_grad_s(Pf[row * A + lane], att_slope);
    if (lane < G) f = feats[row * G + lane];
    db += dp;
    float df = 0.0f;
#pragma unroll
    for (int k = 0; k < 32; ++k) {
      acc[k] = fmaf(__shfl_sync(0xffffffffu, f, k), dp, acc[k]);        // dW1[k][lane]
      df = fmaf(__shfl_sync(0xffffffffu, dp, k), W1s[lane][k], df);     // dfeats[lane] += dp[k] W1[lane][k]
    }
    if (alpha != nullptr) {
      const int b = (int)(row / R), r = (int)(row % R);
      // 15 dependent L2 round trips per row if left rolled: issue the loads of five steps together (same fma order)
      float s = 0.0f;
      const float* ap = alpha + (size_t)b * R + r;
      const float* dp_ = dxctx + (size_t)b * G + lane;
      const size_t as = (size_t)B * R, ds = (size_t)B * G;
      int t = 0;
      for (; t + 4 < T; t += 5) {
        float av[5], dv[5];
#pragma unroll
        for (int u = 0; u < 5; ++u) {
          av[u] = __ldg(ap + (size_t)(t + u) * as);
          dv[u] = lane < G ? __ldg(dp_ + (size_t)(t + u) * ds) : 0.0f;
        }
#pragma unroll
        for (int u = 0; u < 5; ++u) s = fmaf(av[u], dv[u], s);
      }
      for (; t < T; ++t) s = fmaf(__ldg(ap + (size_t)t * as), lane < G ? __ldg(dp_ + (size_t)t * ds) : 0.0f, s);
      if (lane < G) dfeats[row * G + lane] = s + df;
    } else if (lane < G) {
      dfeats[row * G + lane] += df;
    }
  }
  float* out = partial + (size_t)blockIdx.x * (32 * 32 + 32);
#pragma unroll
  for (int g = 0; g < 32; ++g) {
    __syncthreads();
    red[warp][lane] = acc[g];
    __syncthreads();
    if (warp == 0) {
      float t = 0.0f;
#pragma unroll
      for (int i = 0; i < 8; ++i) t += red[i][lane];
      out[g * 32 + lane] = t;
    }
  }
  __syncthreads();
  red[warp][lane] = db;
  __syncthreads();
  if (warp == 0) {
    float t = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][lane];
    out[32 * 32 + lane] = t;
  }
}

// one block per 32 outputs; warp w sums partials w, w+8, ... (coalesced), then the 8 warps are combined in order
__global__ void __launch_bounds__(256)
att_w1_reduce_kernel(int nblk, int G, int A, const float* __restrict__ partial, float* __restrict__ dW1,
                     float* __restrict__ db1) {
  __shared__ float red[8][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int e = blockIdx.x * 32 + lane;               // < 33 * 32 = 32*32 + 32
  float t = 0.0f;
  for (int i = warp; i < nblk; i += 8) t += partial[(size_t)i * (32 * 32 + 32) + e];
  red[warp][lane] = t;
  __syncthreads();
  if (warp != 0) return;
  t = 0.0f;
#pragma unroll
  for (int i = 0; i < 8; ++i) t += red[i][lane];
  if (e < 32 * 32) {
    int g = e >> 5, a = e & 31;
    if (g < G && a < A) dW1[g * A + a] = t;
  } else {
    int a = e - 32 * 32;
    if (a < A) db1[a] = t;
  }
}

// embedding tape gradient: IndexedSlices values (B*T,E) -> scatter-add; sum of squares of the values
__global__ void __launch_bounds__(128)
emb_bwd_kernel(int B, int T, int E, int V, int GE, int G, const int32_t* __restrict__ cap,
               const float* __restrict__ dxtext, float* __restrict__ dEmb, double* __restrict__ sq_out,
               DropSite dtext, DropSite dlin, float* __restrict__ colsq /* per column sum of value^2 (agc.py) or nullptr */) {
  __shared__ float rs[32];
  int row = blockIdx.x;
  int t = row / B, b = row % B;
  int id = cap[(size_t)b * T + t];
  id = id < 0 ? 0 : (id >= V ? V - 1 : id);
  DropSite dl = dlin; dl.step = (uint32_t)t;
  float sq = 0.0f;
  for (int e = threadIdx.x; e < E; e += blockDim.x) {
    float v = dxtext[(size_t)row * E + e];
    v *= drop_mult(dtext, ((uint64_t)b * T + t) * E + e);
    v *= drop_mult(dl, (uint64_t)b * GE + G + e);
    sq = fmaf(v, v, sq);
    if (colsq != nullptr) atomicAdd(&colsq[e], v * v);
    atomicAdd(&dEmb[(size_t)id * E + e], v);
  }
  float tot = block_sum(sq, rs);
  if (threadIdx.x == 0) atomicAdd(sq_out, (double)tot);
}

// greedy decode: argmax over a logits row, first maximal index (np.argmax, lc_NIC.py:627); optional softmax
__global__ void __launch_bounds__(256)
argmax_softmax_kernel(int V, const float* __restrict__ logits, int32_t* __restrict__ word,
                      int64_t* __restrict__ out_ids, int out_stride, float* __restrict__ probs, int64_t probs_stride) {
  __shared__ float rv[8];
  __shared__ int ri[8];
  __shared__ float rs[32];
  const int b = blockIdx.x;
  const float* row = logits + (size_t)b * V;
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
  best = rv[0]; bi = ri[0];
  for (int i = 1; i < 8; ++i)
    if (rv[i] > best || (rv[i] == best && ri[i] < bi)) { best = rv[i]; bi = ri[i]; }
  if (threadIdx.x == 0) { word[b] = bi; out_ids[(size_t)b * out_stride] = (int64_t)bi; }
  if (probs) {
    float sum = 0.0f;
    for (int v = threadIdx.x; v < V; v += blockDim.x) sum += expf(row[v] - best);
    sum = block_sum(sum, rs);
    float* pr = probs + (size_t)b * probs_stride;
    for (int v = threadIdx.x; v < V; v += blockDim.x) pr[v] = expf(row[v] - best) / sum;
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static inline int ceil_div(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// One decoder step given xtext-projection already in Zx_t:  attention -> z += [ctx|h] @ [Wc;U] -> cell
void dec_step(tnt_ctx* c, int B, int t, int training, uint64_t seed, const float* hprev, const float* cprev,
                     float* Zx_t, float* alpha_t, float* qact_t, float* xctx_t, float* gates_t, float* cnew,
                     float* hnew, float* hout_t, float* s_out, cudaStream_t st) {
  const tnt_config& f = c->cfg;
  const int GE = f.G + f.E;
  DropSite dattn = make_drop(f.dropout_attn, training, seed, SITE_ATTN, t);
  DropSite dlin = make_drop(f.dropout_lstm, training, seed, SITE_LSTM_IN, t);
  DropSite dout = make_drop(f.dropout_lstm, training, seed, SITE_LSTM_OUT, t);
  att_fwd_kernel<<<B, 256, f.R * sizeof(float), st>>>(B, f.R, f.G, f.A, f.U, GE, hprev, wsT<float>(c, "Pf"),
                                                      wsT<float>(c, "feats"), c->w + c->o_W2, c->w + c->o_b2,
                                                      c->w + c->o_V, c->w + c->o_bV, alpha_t, qact_t, xctx_t,
                                                      s_out, dattn, dlin, att_slope_of(f));
  c->launches += 1;
  // z = Zx_t + xctx @ W[0:G] + hprev @ U
  gemm_tc(c, 0, 1, false, false, B, 4 * f.U, f.G, xctx_t, f.G, c->w + c->o_lW, 4 * f.U, Zx_t, 4 * f.U, nullptr, 2, st);
  gemm_tc(c, 0, 1, false, false, B, 4 * f.U, f.U, hprev, f.U, c->w + c->o_lU, 4 * f.U, Zx_t, 4 * f.U, nullptr, 2, st);
  lstm_pointwise_kernel<<<ceil_div((int64_t)B * f.U, 256), 256, 0, st>>>(B, f.U, Zx_t, cprev, gates_t, cnew, hnew,
                                                                        hout_t, dout);
  c->launches += 1;
}

// The part of the decoder that depends on neither the encoder nor the recurrence: embedding gather (+ dropout_text,
// + the text part of the LSTM input-dropout mask), Zx = xtext @ W[G:] + b, and the packed forms of the weights the
// later GEMMs of the phase use.  Runs on the side stream next to the encoder when the caller announces the step.
void dec_prologue(tnt_ctx* c, const int32_t* cap, int B, int training, int backward, uint64_t seed, cudaStream_t st) {
  const tnt_config& f = c->cfg;
  const int T = f.T, U = f.U, E = f.E, G = f.G, GE = f.G + f.E;
  float* xtext = wsT<float>(c, "xtext");
  float* Zx = wsT<float>(c, "Zx");
  const float* We = c->w + c->o_lW + (size_t)G * 4 * U;
  DropSite dtext = make_drop(f.dropout_text, training, seed, SITE_TEXT, 0);
  DropSite dlin = make_drop(f.dropout_lstm, training, seed, SITE_LSTM_IN, 0);
  {
    // every weight operand of the phase in one launch (the tc_pack_cached calls below then hit the cache)
    TcPackReq reqs[6];
    int n = 0;
    reqs[n++] = {We, E, 4 * U, 4 * U, true};                         // B of Zx
    reqs[n++] = {c->w + c->o_iW, U, f.H, f.H, true};                 // dense_inter forward
    reqs[n++] = {c->w + c->o_oW, f.H, f.V, f.V, true};               // dense_out forward
    if (backward) {
      reqs[n++] = {c->w + c->o_oW, f.H, f.V, f.V, false};            // d inter
      reqs[n++] = {c->w + c->o_iW, U, f.H, f.H, false};              // d hout
      reqs[n++] = {We, E, 4 * U, 4 * U, false};                      // d xtext
    }
    tc_prepack(c, reqs, n, st);
  }
  emb_fwd_kernel<<<T * B, 128, 0, st>>>(B, T, E, f.V, GE, G, cap, T, 0, c->w + c->o_emb, xtext, dtext, dlin);
  c->launches += 1;
  // Zx = xtext @ W[G:] + b   (the time-parallel part of x_t @ W)
  gemm_tc(c, 1, 1, false, false, T * B, 4 * U, E, xtext, E, We, 4 * U, Zx, 4 * U, c->w + c->o_lb, 0, st);
}

// attn_colsum[t][r] per half of the batch                                              (lc_NIC.py:365)
void dec_attn_colsum(tnt_ctx* c, int B, cudaStream_t st) {
  const tnt_config& f = c->cfg;
  int halves = f.n_enc == 2 ? 2 : 1;
  for (int h = 0; h < halves; ++h) {
    int nb = halves == 2 ? B / 2 : B, b0 = h * nb;
    attn_colsum_kernel<<<f.T, 128, 0, st>>>(B, f.R, b0, nb, wsT<float>(c, "alpha"),
                                            wsT<float>(c, "attn_colsum") + (size_t)h * f.T * f.R);
    c->launches += 1;
  }
}

void dec_forward(tnt_ctx* c, const int32_t* cap, const float* h0, const float* c0, int B, int training,
                 uint64_t seed, cudaStream_t st) {
  const tnt_config& f = c->cfg;
  const int T = f.T, U = f.U, E = f.E, G = f.G, GE = f.G + f.E;
  float* xtext = wsT<float>(c, "xtext");
  float* Zx = wsT<float>(c, "Zx");
  float* Hall = wsT<float>(c, "Hall");
  float* Call = wsT<float>(c, "Call");
  if (!(c->pro_valid && c->pro_cap == cap && c->pro_B == B && c->pro_training == training && c->pro_seed == seed))
    dec_prologue(c, cap, B, training, 0, seed, st);
  c->pro_valid = 0;
  cudaMemcpyAsync(Hall, h0, (size_t)B * U * sizeof(float), cudaMemcpyDeviceToDevice, st);
  cudaMemcpyAsync(Call, c0, (size_t)B * U * sizeof(float), cudaMemcpyDeviceToDevice, st);
  TNT_TRACE(c, "pd_fwd_kernel (15-step attention-LSTM recurrence)", st);
  const bool persistent = pd_enabled(c);
  c->variant_rec = persistent ? "pd_fwd_kernel/pd_bwd_kernel (persistent)" : "per-step kernels (shape outside the persistent kernels)";
  if (persistent) {
    pd_forward(c, B, T, training, seed, wsT<float>(c, "gates"), wsT<float>(c, "hout"), wsT<float>(c, "alpha"), nullptr, st);
    return;
  }
  for (int t = 0; t < T; ++t) {
    size_t bu = (size_t)B * U;
    dec_step(c, B, t, training, seed, Hall + t * bu, Call + t * bu, Zx + (size_t)t * B * 4 * U,
             wsT<float>(c, "alpha") + (size_t)t * B * f.R, wsT<float>(c, "qact") + (size_t)t * B * f.A,
             wsT<float>(c, "xctx") + (size_t)t * B * G, wsT<float>(c, "gates") + (size_t)t * B * 4 * U,
             Call + (t + 1) * bu, Hall + (t + 1) * bu, wsT<float>(c, "hout") + t * bu, nullptr, st);
  }
}

void dec_backward(tnt_ctx* c, const int32_t* cap, int B, int B_global, uint64_t seed, cudaStream_t st) {
  const tnt_config& f = c->cfg;
  const int T = f.T, U = f.U, E = f.E, G = f.G, A = f.A, R = f.R, GE = f.G + f.E;
  const size_t bu = (size_t)B * U;
  float* Hall = wsT<float>(c, "Hall");
  float* Call = wsT<float>(c, "Call");
  float* dZ = wsT<float>(c, "dZ");
  float* dhx = wsT<float>(c, "dh_rec");          // [B][U] followed by dxctx [B][G]
  float* dxctx = dhx + bu;
  float* dc = wsT<float>(c, "dc");
  float* dfeats = wsT<float>(c, "dfeats");
  float* dPf = wsT<float>(c, "dPf");
  const bool persistent = pd_enabled(c);
  if (persistent) {
    TNT_TRACE(c, "pd_bwd_kernel (recurrence gradient)", st);
    pd_backward(c, B, T, seed, st);
  }
  if (!persistent) {
    cudaMemsetAsync(dhx, 0, (bu + (size_t)B * G) * sizeof(float), st);
    cudaMemsetAsync(dc, 0, bu * sizeof(float), st);
    cudaMemsetAsync(dfeats, 0, (size_t)B * R * G * sizeof(float), st);
    cudaMemsetAsync(dPf, 0, (size_t)B * R * A * sizeof(float), st);
  }
  for (int t = T - 1; t >= 0 && !persistent; --t) {
    DropSite dattn = make_drop(f.dropout_attn, 1, seed, SITE_ATTN, t);
    DropSite dlin = make_drop(f.dropout_lstm, 1, seed, SITE_LSTM_IN, t);
    DropSite dout = make_drop(f.dropout_lstm, 1, seed, SITE_LSTM_OUT, t);
    float* dZt = dZ + (size_t)t * B * 4 * U;
    lstm_bwd_pointwise_kernel<<<ceil_div((int64_t)bu, 256), 256, 0, st>>>(
        B, U, wsT<float>(c, "dhout") + t * bu, dhx, dc, wsT<float>(c, "gates") + (size_t)t * B * 4 * U,
        Call + (t + 1) * bu, Call + t * bu, dZt, dout);
    c->launches += 1;
    cudaMemsetAsync(dhx, 0, (bu + (size_t)B * G) * sizeof(float), st);
    // dh_{t-1} = dZ_t @ U^T ; d(xctx_t) = dZ_t @ W[0:G]^T
    gemm_tc(c, 0, 1, false, true, B, U, 4 * U, dZt, 4 * U, c->w + c->o_lU, 4 * U, dhx, U, nullptr, 2, st);
    gemm_tc(c, 0, 1, false, true, B, G, 4 * U, dZt, 4 * U, c->w + c->o_lW, 4 * U, dxctx, G, nullptr, 2, st);
    att_bwd_kernel<<<B, 256, 2 * R * sizeof(float), st>>>(
        B, R, G, A, U, GE, dxctx, wsT<float>(c, "alpha") + (size_t)t * B * R,
        wsT<float>(c, "qact") + (size_t)t * B * A, wsT<float>(c, "Pf"), wsT<float>(c, "feats"), c->w + c->o_W2,
        c->w + c->o_V, dfeats, dPf, wsT<float>(c, "dqpre") + (size_t)t * B * A,
        wsT<float>(c, "dVpart") + (size_t)t * B * A, wsT<float>(c, "dbVpart") + (size_t)t * B, dhx, dattn, dlin, att_slope_of(f),
        f.attn_loss ? wsT<float>(c, "attn_e") + (size_t)t * B * R : nullptr);
    c->launches += 1;
  }
  const int TB = T * B;
  // Everything below that only produces decoder weight gradients runs on the side stream; the main stream carries
  // on with the chain that leads back into the encoder (d feats -> W1 path -> BatchNorm backward -> encoder dW).
  cudaStream_t main_st = st;
  if (c->side) {
    cudaEventRecord(c->ev_fork, main_st);
    cudaStreamWaitEvent(c->side, c->ev_fork, 0);
    tc_set_lane(c, 1);
    st = c->side;
  }
  TraceScope* ts_side = new TraceScope(c, "decoder weight gradients: 5 gemm_pk + pack + emb_bwd (side stream)", st);
  // every operand of the five weight-gradient products is final here: convert them with one launch
  {
    const TcPackReq reqs[] = {
        {dZ, TB, 4 * U, 4 * U, false},                          // A of d(xtext) = dZ @ W[G:]^T
        {dZ, TB, 4 * U, 4 * U, true},                           // B of dU, dW[0:G], dW[G:]
        {Hall, TB, U, U, true},                                 // A of dU and dW2
        {wsT<float>(c, "xctx"), TB, G, G, true},                // A of dW[0:G]
        {wsT<float>(c, "xtext"), TB, E, E, true},               // A of dW[G:]
        {wsT<float>(c, "dqpre"), TB, A, A, true},               // B of dW2
    };
    tc_prepack(c, reqs, 6, st);
  }
  // embedding: d(xtext) = dZ @ W[G:]^T -> masks -> IndexedSlices values -> scatter-add
  float* dxtext = wsT<float>(c, "dxtext");
  gemm_tc(c, 1, 1, false, true, TB, E, 4 * U, dZ, 4 * U, c->w + c->o_lW + (size_t)G * 4 * U, 4 * U, dxtext, E, nullptr, 0, st);
  cudaMemsetAsync(c->g + c->o_emb, 0, (size_t)f.V * E * sizeof(float), st);
  DropSite dtext = make_drop(f.dropout_text, 1, seed, SITE_TEXT, 0);
  DropSite dlin0 = make_drop(f.dropout_lstm, 1, seed, SITE_LSTM_IN, 0);
  double* mraw = wsT<double>(c, "metrics_raw");
  float* colsq = f.agc_clip > 0.0f ? wsT<float>(c, "emb_colsq") : nullptr;
  if (colsq) cudaMemsetAsync(colsq, 0, (size_t)E * sizeof(float), st);
  emb_bwd_kernel<<<TB, 128, 0, st>>>(B, T, E, f.V, GE, G, cap, dxtext, c->g + c->o_emb, mraw + 7, dtext, dlin0, colsq);
  c->launches += 1;
  // metrics_raw is complete here (the un-aggregated embedding norm was its last entry): a data-parallel caller
  // may reduce it now, long before the rest of the decoder bucket is ready
  dec_reduce_metrics(c, st);                       // data parallel: metric sums, embedding clip norm, attention column sums
  if (c->side) cudaEventRecord(c->ev_metrics, st);
  // time-parallel weight gradients
  gemm_tc(c, 1, 1, true, false, U, 4 * U, TB, Hall, U, dZ, 4 * U, c->g + c->o_lU, 4 * U, nullptr, 0, st);
  gemm_tc(c, 1, 1, true, false, G, 4 * U, TB, wsT<float>(c, "xctx"), G, dZ, 4 * U, c->g + c->o_lW, 4 * U, nullptr, 0, st);
  gemm_tc(c, 1, 1, true, false, E, 4 * U, TB, wsT<float>(c, "xtext"), E, dZ, 4 * U, c->g + c->o_lW + (size_t)G * 4 * U,
        4 * U, nullptr, 0, st);
  gemm_tc(c, 1, 1, true, false, U, A, TB, Hall, U, wsT<float>(c, "dqpre"), A, c->g + c->o_W2, A, nullptr, 0, st);
  // W1 path (+ dfeats contribution): main stream
  { TNT_TRACE(c, "att_w1_bwd + reduce (d feats)", main_st);
  int nblk = 4 * c->sm_count;                       // ~5 rows per warp: the kernel is one L2 round trip per row
  att_w1_bwd_kernel<<<nblk, 256, 0, main_st>>>((int64_t)B * R, G, A, wsT<float>(c, "Pf"), dPf, wsT<float>(c, "feats"),
                                               c->w + c->o_W1, dfeats, wsT<float>(c, "w1part"), T, B, R,
                                               persistent ? wsT<float>(c, "alpha") : nullptr,
                                               persistent ? wsT<float>(c, "pd_dxctx") : nullptr, att_slope_of(c->cfg));
  att_w1_reduce_kernel<<<33, 256, 0, main_st>>>(nblk, G, A, wsT<float>(c, "w1part"), c->g + c->o_W1, c->g + c->o_b1);
  c->launches += 2; }
  // the four bias / V column sums read buffers that are final since the recurrence gradient: they run on the main
  // stream (which has slack before it joins the side stream), keeping the side stream's chain to the five GEMMs
  { TNT_TRACE(c, "decoder bias gradients: colsum_multi", main_st);
  const ColsumReq cs[] = {
      {dZ, c->g + c->o_lb, TB, 4 * U, 4 * U},
      {wsT<float>(c, "dqpre"), c->g + c->o_b2, TB, A, A},
      {wsT<float>(c, "dVpart"), c->g + c->o_V, persistent ? 2 * TB : TB, A, A},
      {wsT<float>(c, "dbVpart"), c->g + c->o_bV, persistent ? 2 * TB : TB, 1, 1},
  };
  colsum_multi(c, cs, 4, main_st); }
  if (c->side) cudaEventRecord(c->ev_w1, main_st);   // dW1/db1 belong to the decoder bucket but are made on the main stream
  delete ts_side;
  if (c->side) {
    tc_set_lane(c, 0);
    cudaStreamWaitEvent(c->side, c->ev_w1, 0);         // side-stream order now covers every decoder gradient
    // data parallel: attention, embedding and LSTM gradients (the decoder bucket without the head) -- overlaps the
    // encoder backward on the main stream
    dp_allreduce(c, c->g + c->dec_off, c->o_iW - c->dec_off, 0, 1, c->side);
    cudaEventRecord(c->ev_join, c->side);
    c->side_pending = 1;
  }
  (void)B_global;
}

// Zx[b][:] = table[word[b]][:]  (4U floats per row)
__global__ void __launch_bounds__(256)
gather_rows_kernel(int n4, const int32_t* __restrict__ word, int V, const float4* __restrict__ table, float4* __restrict__ out) {
  int w = word[blockIdx.x];
  w = w < 0 ? 0 : (w >= V ? V - 1 : w);
  const float4* src = table + (size_t)w * n4;
  float4* dst = out + (size_t)blockIdx.x * n4;
  for (int i = threadIdx.x; i < n4; i += blockDim.x) dst[i] = __ldg(src + i);
}

void dec_greedy(tnt_ctx* c, const float* h0, const float* c0, const int32_t* start, int B, int max_len,
                int64_t* out_ids, float* out_probs, float* out_attn, float* out_attn_hidden, cudaStream_t st,
                int sample, int top_k, float top_p, uint64_t seed) {
  // greedy_predict_attention loop (lc_NIC.py:607-632); inference mode: no dropout anywhere.
  const tnt_config& f = c->cfg;
  const int U = f.U, E = f.E, G = f.G, GE = f.G + f.E, H = f.H, V = f.V;
  const size_t bu = (size_t)B * U;
  float* hA = wsT<float>(c, "Hall");            // ping-pong in the first two slots
  float* cA = wsT<float>(c, "Call");
  float* xtext = wsT<float>(c, "xtext");
  float* Zx = wsT<float>(c, "Zx");
  float* inter = wsT<float>(c, "inter_d");
  float* logits = wsT<float>(c, "logits");
  int32_t* word = wsT<int32_t>(c, "word");
  cudaMemcpyAsync(hA, h0, bu * sizeof(float), cudaMemcpyDeviceToDevice, st);
  cudaMemcpyAsync(cA, c0, bu * sizeof(float), cudaMemcpyDeviceToDevice, st);
  cudaMemcpyAsync(word, start, (size_t)B * sizeof(int32_t), cudaMemcpyDeviceToDevice, st);
  DropSite none = make_drop(0.0f, 0, 0, 0, 0);
  // Without dropout the text half of the LSTM input projection depends on the word only: project the whole
  // embedding table once per call (V x 4U) and gather rows per step, instead of gather + GEMM per step.
  float* emb_proj = (max_len > 2 && (U & 0) == 0) ? wsT<float>(c, "emb_proj") : nullptr;
  if (emb_proj)
    gemm_tc(c, 0, 1, false, false, V, 4 * U, E, c->w + c->o_emb, E, c->w + c->o_lW + (size_t)G * 4 * U, 4 * U, emb_proj,
             4 * U, c->w + c->o_lb, 0, st);
  for (int i = 0; i < max_len; ++i) {
    float* hp = hA + (i & 1) * bu; float* hn = hA + ((i + 1) & 1) * bu;
    float* cp = cA + (i & 1) * bu; float* cn = cA + ((i + 1) & 1) * bu;
    if (emb_proj) {
      gather_rows_kernel<<<B, 256, 0, st>>>(U, word, V, reinterpret_cast<const float4*>(emb_proj), reinterpret_cast<float4*>(Zx));
      c->launches += 1;
    } else {
      emb_fwd_kernel<<<B, 128, 0, st>>>(B, 1, E, V, GE, G, word, 1, i, c->w + c->o_emb, xtext, none, none);
      c->launches += 1;
      gemm_tc(c, 0, 1, false, false, B, 4 * U, E, xtext, E, c->w + c->o_lW + (size_t)G * 4 * U, 4 * U, Zx, 4 * U,
            c->w + c->o_lb, 0, st);
    }
    dec_step(c, B, i, 0, 0, hp, cp, Zx, out_attn ? out_attn + (size_t)i * B * f.R : wsT<float>(c, "alpha"),
             wsT<float>(c, "qact"), wsT<float>(c, "xctx"), nullptr, cn, hn, nullptr,
             out_attn_hidden ? out_attn_hidden + (size_t)i * B * f.R * f.A : nullptr, st);
    // dense_inter (LeakyReLU) -> dense_out                              (lc_NIC.py:620-623)
    head_decode_step(c, B, hn, inter, logits, st);
    if (sample) {
      // sample_choice / top-k / nucleus (lc_NIC.py:571-575, img_NIC.py:509-533): sampling.cu
      dec_sample_words(c, B, i, top_k, top_p, seed, logits, word, out_ids + i, max_len,
                       out_probs ? out_probs + (size_t)i * V : nullptr, (int64_t)max_len * V, st);
    } else {
      argmax_softmax_kernel<<<B, 256, 0, st>>>(V, logits, word, out_ids + i, max_len,
                                               out_probs ? out_probs + (size_t)i * V : nullptr, (int64_t)max_len * V);
      c->launches += 1;
    }
  }
  (void)H;
}
