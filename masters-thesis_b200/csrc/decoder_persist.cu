// Persistent attention-LSTM recurrence: the whole `for i in range(max_len)` loop of lc_NIC.call_attention
// (AttemptFour/Model/lc_NIC.py:244-256) -- Attention.call (attention.py:25-44) + the Keras LSTM step
// (lc_NIC.py:118-124,255) -- and its tape gradient, each in ONE cooperative kernel.  The T dependent steps
// cost two group-scoped counter barriers each instead of ~8 kernel launches.
//
// Decomposition (U = 512, G = A = 32): the batch is cut into row groups of 16 samples; a group is served by
// NCG = U/16 = 32 CTAs (one per SM), so 64 rows occupy 128 SMs.  CTA j of a group owns hidden units
// 16j .. 16j+15, i.e. 64 gate columns (i, f, c~, o).  Its slice of [W_ctx ; U] (544 x 64 fp32 = 136 KB) stays
// resident in shared memory for all steps -- the recurrent weights never leave the SM.  CTA j < 16 also runs
// the attention of sample 16g + j with that sample's features / hoisted projection Pf and its slice of W2 in
// registers.  All arithmetic is fp32 FMA (the reference is fp32); reductions are warp shuffles + shared memory.
//
//   forward step t:   wait H(t-1) -> stage h_{t-1}[16 rows] -> attention (q, tanh score, softmax, context)
//                     -> publish ctx_t, arrive A -> wait A -> z = Zx_t + [ctx_t | h_{t-1}] W_slice -> cell
//                     -> publish h_t, arrive H
//   backward step t:  dz_t (pointwise, own units) -> partial dx = dz_t[:, own 64] W_slice^T (16 x 544)
//                     -> publish partial, arrive B1 -> wait B1 -> attention backward for sample j (needs
//                     d ctx_t = sum of partials) and dh_{t-1}[own units] = sum of partials -> publish dq, arrive
//                     B2 -> wait B2 -> dh_{t-1} += dq W2[own units]^T
// Partials are summed by their consumer in a fixed order: no atomics, bit-reproducible run to run.
#include "tnt_internal.h"
#include "tc.cuh"
#include <cstdlib>
#include <cstring>
#include <cstdio>
#include <algorithm>

#define PD_THREADS 512
#define PD_QI 6                  // region quads per attention warp: R <= 16 warps * 4 * PD_QI = 384
#define PD_MAXR (16 * 4 * PD_QI)
#define PD_HQ (PD_QI / 2)        // forward: the regions of a sample are split over two CTAs
#define PD_AXW 36                // floats per half-attention record: context (32), max, sum, pad
#define PD_SPIN_LIMIT (1ll << 31) // ~1 s of SM clocks: a barrier that waits longer aborts the kernel

struct PdFwd {
  int B, row0, nrows, T, R, GE, NGtot;
  const float *W2, *b2, *Vw, *bV, *lW, *lU;
  const float *Zx, *Pf, *feats;
  float *Hall, *Call, *gates, *alpha, *qact, *xctx, *hout, *XT, *s_out, *AX, *ctxraw;
  unsigned* bar;                 // [group][4]: 0 = H, 1 = A; bar[4*NG] = abort flag
  unsigned* abort_flag;
  DropSite dattn, dlin, dout;    // .step is set per step
  float att_slope;               // LeakyReLU slope of the attention projections (1 = none)
  long long* trace;              // debug: clock64 stamps of CTAs (0,0) and (0,NCG-1): [2][T][16]
};

struct PdBwd {
  int B, row0, nrows, T, R, GE, NGtot;
  const float *W2, *Vw, *lW, *lU, *Pf, *feats, *gates, *Call, *dhout, *alpha, *qact;
  float *dZ, *dPf, *dqpre, *dVpart, *dbVpart, *dxctx, *P, *DQ;
  const float* ctxraw;
  const float *attn_e, *attn_Se;  // attention loss in the differentiated loss (lc_NIC.py:384): dL/d alpha [T][B][R] and
                                  // sum_r alpha dL/d alpha [T][B]; nullptr = reported only
  unsigned* bar;                 // [group][4]: 2 = B1, 3 = B2
  unsigned* abort_flag;
  DropSite dattn, dlin, dout;
  float att_slope;
  long long* trace;
};

// Fast transcendental forms (MUFU ex2 / rcp): absolute error ~1e-7 on tanh and sigmoid, far inside the 1e-3 parity bar;
// the accurate tanhf/expf cost 4-6x the issue slots and the attention phase is issue bound.
__device__ __forceinline__ float pd_tanh(float x) {
  const float e = __expf(-2.0f * fabsf(x));
  return copysignf(__fdividef(1.0f - e, 1.0f + e), x);
}
__device__ __forceinline__ float pd_sigmoid(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }

// four dropout multipliers of elements 4*idx4 .. 4*idx4+3 of a site: one Philox call (common.cuh: drop_mult)
__device__ __forceinline__ void drop_mult4(const DropSite& d, uint64_t idx4, float m[4]) {
  if (d.thr == 0) { m[0] = m[1] = m[2] = m[3] = 1.0f; return; }
  const uint4 r = philox4x32_10((uint32_t)idx4, d.site, d.step, 0u, d.k0, d.k1);
  m[0] = r.x >= d.thr ? d.scale : 0.0f;
  m[1] = r.y >= d.thr ? d.scale : 0.0f;
  m[2] = r.z >= d.thr ? d.scale : 0.0f;
  m[3] = r.w >= d.thr ? d.scale : 0.0f;
}

#define PD_T(ev) do { if (p.trace && g == 0 && (j == 0 || j == NCG - 1) && tid == 0) p.trace[((j ? 1 : 0) * T + t) * 16 + (ev)] = clock64(); } while (0)

// thread 0 waits until *ctr >= target; everybody learns whether to go on
__device__ __forceinline__ bool pd_wait(const unsigned* ctr, unsigned target, unsigned* abort_flag, int* s_ok) {
  if (threadIdx.x == 0) {
    int ok = 1;
    if (tc::ld_acquire(ctr) < target) {
      const long long t0 = clock64();
      while (tc::ld_acquire(ctr) < target) {
        if (clock64() - t0 > PD_SPIN_LIMIT || tc::ld_acquire(abort_flag) != 0u) {
          atomicExch(abort_flag, 1u);
          ok = 0;
          break;
        }
      }
    }
    *s_ok = ok;
  }
  __syncthreads();
  return *s_ok != 0;
}
// all threads' global writes so far become visible to the CTAs that observe the counter
__device__ __forceinline__ void pd_arrive(unsigned* ctr) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    tc::red_release(ctr, 1u);
  }
}

// ------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------
template <int U>
__global__ void __launch_bounds__(PD_THREADS, 1) pd_fwd_kernel(PdFwd p) {
  constexpr int G = 32, A = 32, K = G + U, NCG = U / 16, UPW = U / 16, KS = K / 8;
  static_assert(K % 8 == 0 && UPW <= 32 && NCG == 32, "unsupported U");
  extern __shared__ __align__(16) float pd_smem[];
  float* Ws = pd_smem;                 // [K][64]   resident weight slice
  float* xs = Ws + K * 64;             // [K][16]   [ctx_t ; h_{t-1}] of the 16 rows, k-major
  float* red = xs + K * 16;            // [8][16*64] split-K partials
  float* sc = red + 8 * 1024;          // [PD_MAXR] attention scores (own half of the regions)
  float* part = sc + PD_MAXR;          // [16][33]
  __shared__ int s_ok;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = blockIdx.x / NCG, j = blockIdx.x % NCG;
  const int gg = p.row0 / 16 + g;
  const int nrg = min(16, p.nrows - 16 * g);              // valid rows of this group
  // attention: sample jj of the group is shared by CTAs jj (regions [0, Rh)) and 16 + jj (regions [Rh, R))
  const int jj = j & 15, hf = j >> 4;
  const bool has_sample = jj < nrg;
  const int b = p.row0 + 16 * g + jj;
  const int B = p.B, T = p.T, R = p.R;
  const int Rh = (((R + 1) >> 1) + 3) & ~3;               // regions of half 0 (multiple of 4, <= PD_MAXR / 2)
  const int r0 = hf * Rh;
  const int nloc = max(0, min(Rh, R - r0));               // regions of this CTA
  const int sub = lane >> 3, a4 = lane & 7;
  // cell ownership: thread (crow, cul), tid < 256
  const int crow = (tid >> 4) & 15, cul = tid & 15;
  const bool cvalid = tid < 256 && crow < nrg;
  const int cb = p.row0 + 16 * g + crow, cu = 16 * j + cul;
  unsigned* barH = p.bar + 4 * g;
  unsigned* barA = p.bar + 4 * g + 1;
  const size_t BU = (size_t)B * U;

  // ---- resident operands ---------------------------------------------------------------------------
  for (int idx = tid; idx < K * 64; idx += PD_THREADS) {
    const int k = idx >> 6, c = idx & 63;
    const float* src = (k < G ? p.lW + (size_t)k * 4 * U : p.lU + (size_t)(k - G) * 4 * U) + (c >> 4) * U + 16 * j + (c & 15);
    Ws[idx] = __ldg(src);
  }
  float pf[PD_HQ][4], ft[PD_HQ][4], w2r[UPW];
#pragma unroll
  for (int i = 0; i < PD_HQ; ++i) {
    const int rl = 4 * (warp + 16 * i) + sub;
#pragma unroll
    for (int c = 0; c < 4; ++c) { pf[i][c] = 0.0f; ft[i][c] = 0.0f; }
    if (has_sample && rl < nloc) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(p.Pf + ((size_t)b * R + r0 + rl) * A) + a4);
      const float4 f = __ldg(reinterpret_cast<const float4*>(p.feats + ((size_t)b * R + r0 + rl) * G) + a4);
      pf[i][0] = a.x; pf[i][1] = a.y; pf[i][2] = a.z; pf[i][3] = a.w;
      ft[i][0] = f.x; ft[i][1] = f.y; ft[i][2] = f.z; ft[i][3] = f.w;
    }
  }
#pragma unroll
  for (int i = 0; i < UPW; ++i) w2r[i] = has_sample ? __ldg(p.W2 + (size_t)(UPW * warp + i) * A + lane) : 0.0f;
  const float b2v = __ldg(p.b2 + lane);
  const float bv = __ldg(p.bV);
  float v4[4];
#pragma unroll
  for (int c = 0; c < 4; ++c) v4[c] = __ldg(p.Vw + 4 * a4 + c);

  // initial state: h_{-1} into XT[0], c_{-1} into a register
  float creg = 0.0f;
  if (tid < 256) {
    float h = 0.0f;
    if (cvalid) { h = __ldg(p.Hall + (size_t)cb * U + cu); creg = __ldg(p.Call + (size_t)cb * U + cu); }
    p.XT[((size_t)gg * K + G + cu) * 16 + crow] = h;
  }
  pd_arrive(barH);

  for (int t = 0; t < T; ++t) {
    DropSite dattn = p.dattn, dlin = p.dlin, dout = p.dout;
    dattn.step = (uint32_t)t; dlin.step = (uint32_t)t; dout.step = (uint32_t)t;
    float* XTt = p.XT + ((size_t)t * p.NGtot + gg) * K * 16;
    // exchange of the two half-attentions: [half][sample][PD_AXW] = unnormalised context (32), max, sum
    float* AXt = p.AX + ((size_t)t * p.NGtot + gg) * 2 * 16 * PD_AXW;
    float zx[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    if (cvalid) {
      const float* zr = p.Zx + ((size_t)t * B + cb) * 4 * U + cu;
#pragma unroll
      for (int q = 0; q < 4; ++q) zx[q] = __ldg(zr + q * U);
    }
    PD_T(0);
    if (!pd_wait(barH, (unsigned)(t + 1) * NCG, p.abort_flag, &s_ok)) return;
    PD_T(1);
    {
      const float4* src = reinterpret_cast<const float4*>(XTt + G * 16);
      float4* dst = reinterpret_cast<float4*>(xs + G * 16);
      for (int idx = tid; idx < U * 4; idx += PD_THREADS) dst[idx] = __ldcg(src + idx);
    }
    __syncthreads();
    PD_T(2);

    float ev[PD_MAXR / 64];              // warp 0 keeps its exp() values across barrier A to normalise alpha
    float mx = -INFINITY;
    if (has_sample) {
      // ---- q = LeakyReLU(h W2 + b2)                                    (attention.py:32, lc_NIC.py:98)
      float acc = 0.0f;
#pragma unroll
      for (int i = 0; i < UPW; ++i) acc = fmaf(xs[(G + UPW * warp + i) * 16 + jj], w2r[i], acc);
      part[warp * 33 + lane] = acc;
      __syncthreads();
      float q = b2v;
#pragma unroll
      for (int i = 0; i < 16; ++i) q += part[i * 33 + lane];
      q = lrelu_s(q, p.att_slope);
      if (warp == 0 && hf == 0) p.qact[((size_t)t * B + b) * A + lane] = q;
      float q4[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) q4[c] = __shfl_sync(0xffffffffu, q, 4 * a4 + c);
      PD_T(3);
      // ---- score_r = tanh(Pf_r + q) (dropout) . V + bV                  (attention.py:30-38)
#pragma unroll
      for (int i = 0; i < PD_HQ; ++i) {
        const int rl = 4 * (warp + 16 * i) + sub;
        const int r = r0 + rl;
        float e = 0.0f;
        if (rl < nloc) {
          float m[4], s[4];
          drop_mult4(dattn, (((uint64_t)b * R + r) * A) / 4 + a4, m);
#pragma unroll
          for (int c = 0; c < 4; ++c) { s[c] = pd_tanh(pf[i][c] + q4[c]) * m[c]; e = fmaf(s[c], v4[c], e); }
          if (p.s_out) *(reinterpret_cast<float4*>(p.s_out + (((size_t)t * B + b) * R + r) * A) + a4) = make_float4(s[0], s[1], s[2], s[3]);
        }
        e += __shfl_xor_sync(0xffffffffu, e, 1);
        e += __shfl_xor_sync(0xffffffffu, e, 2);
        e += __shfl_xor_sync(0xffffffffu, e, 4);
        if (a4 == 0 && rl < nloc) sc[rl] = e + bv;
      }
      __syncthreads();
      PD_T(4);
      // ---- softmax pieces of the own half, every warp redundantly (same order => same bits)  (attention.py:40)
#pragma unroll
      for (int k = 0; k < PD_MAXR / 64; ++k) {
        const int rl = lane + 32 * k;
        ev[k] = rl < nloc ? sc[rl] : -INFINITY;
        mx = fmaxf(mx, ev[k]);
      }
      mx = warp_max(mx);
      float sum = 0.0f;
#pragma unroll
      for (int k = 0; k < PD_MAXR / 64; ++k) {
        const int rl = lane + 32 * k;
        ev[k] = rl < nloc ? __expf(ev[k] - mx) : 0.0f;
        sum += ev[k];
      }
      sum = warp_sum(sum);
      // ---- unnormalised context of the half = sum_r exp(e_r - max) f_r       (attention.py:42)
      float cx[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
      for (int i = 0; i < PD_HQ; ++i) {
        const int rl = 4 * (warp + 16 * i) + sub;
        if (rl < nloc) {
          const float al = __expf(sc[rl] - mx);
#pragma unroll
          for (int c = 0; c < 4; ++c) cx[c] = fmaf(al, ft[i][c], cx[c]);
        }
      }
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        cx[c] += __shfl_xor_sync(0xffffffffu, cx[c], 8);
        cx[c] += __shfl_xor_sync(0xffffffffu, cx[c], 16);
      }
      if (lane < 8) {
#pragma unroll
        for (int c = 0; c < 4; ++c) part[warp * 33 + 4 * lane + c] = cx[c];
      }
      __syncthreads();
      if (warp == 0) {
        float tot = 0.0f;
#pragma unroll
        for (int i = 0; i < 16; ++i) tot += part[i * 33 + lane];
        float* ax = AXt + (size_t)(hf * 16 + jj) * PD_AXW;
        ax[lane] = tot;
        if (lane == 0) { ax[32] = mx; ax[33] = sum; }
      }
      PD_T(5);
      pd_arrive(barA);
    }
    // ---- z[16 rows, own 64 columns] = [ctx_t | h_{t-1}] W_slice, split K over 8 warp pairs: the h rows first, they
    //      do not depend on this step's attention ------------------------------------------------------------------
    const int ch = warp & 1, ks = warp >> 1, rg = lane >> 3, cg = lane & 7;
    float2 acc[4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i) { acc[i][0] = make_float2(0.0f, 0.0f); acc[i][1] = make_float2(0.0f, 0.0f); }
    auto gemm_rows = [&](int kb, int ke) {
      const float* xp = xs + (size_t)kb * 16 + 4 * rg;
      const float* wp = Ws + (size_t)kb * 64 + 32 * ch + 4 * cg;
#pragma unroll 4
      for (int k = 0; k < ke - kb; ++k) {
        const float4 x = *reinterpret_cast<const float4*>(xp + k * 16);
        const float4 w = *reinterpret_cast<const float4*>(wp + k * 64);
        const float xv[4] = {x.x, x.y, x.z, x.w};
        const float2 w01 = make_float2(w.x, w.y), w23 = make_float2(w.z, w.w);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 xx = make_float2(xv[i], xv[i]);
          acc[i][0] = __ffma2_rn(xx, w01, acc[i][0]);
          acc[i][1] = __ffma2_rn(xx, w23, acc[i][1]);
        }
      }
    };
    gemm_rows(G + ks * (U / 8), G + (ks + 1) * (U / 8));
    // LSTM input-dropout multiplier of (sample = warp, context channel = lane): does not depend on the barrier
    const float lin_mult = drop_mult(dlin, (uint64_t)(p.row0 + 16 * g + warp) * p.GE + lane);
    if (!pd_wait(barA, (unsigned)(t + 1) * 2 * nrg, p.abort_flag, &s_ok)) return;
    PD_T(6);
    // ---- merge the two halves: thread (sample = warp, g = lane)            softmax over all regions ----------------
    {
      const float* a0 = AXt + (size_t)warp * PD_AXW;
      const float* a1 = AXt + (size_t)(16 + warp) * PD_AXW;
      float x = 0.0f;
      if (warp < nrg) {
        const float m0 = __ldcg(a0 + 32), s0 = __ldcg(a0 + 33), m1 = __ldcg(a1 + 32), s1 = __ldcg(a1 + 33);
        const float M = fmaxf(m0, m1);
        const float w0 = __expf(m0 - M), w1 = __expf(m1 - M);
        const float rs = 1.0f / (s0 * w0 + s1 * w1);
        x = (__ldcg(a0 + lane) * w0 + __ldcg(a1 + lane) * w1) * rs;
        const int bs = p.row0 + 16 * g + warp;
        if (hf == 0 && warp == jj) p.ctxraw[((size_t)t * B + bs) * G + lane] = x;   // softmax tape gradient needs ctx . d ctx
        x *= lin_mult;                                                     // LSTM input dropout, context part
        if (hf == 0 && warp == jj) p.xctx[((size_t)t * B + bs) * G + lane] = x;
        if (warp == jj) {                                                  // own sample: normalise alpha of the own half
          const float fac = (hf == 0 ? w0 : w1) * rs;
          sc[PD_MAXR / 2] = fac;
        }
      }
      xs[lane * 16 + warp] = x;
    }
    __syncthreads();
    if (has_sample && warp == 0) {
      const float fac = sc[PD_MAXR / 2];
#pragma unroll
      for (int k = 0; k < PD_MAXR / 64; ++k) {
        const int rl = lane + 32 * k;
        if (rl < nloc) p.alpha[((size_t)t * B + b) * R + r0 + rl] = ev[k] * fac;
      }
    }
    PD_T(7);
    gemm_rows(ks * (G / 8), (ks + 1) * (G / 8));
#pragma unroll
    for (int i = 0; i < 4; ++i)
      *reinterpret_cast<float4*>(red + ks * 1024 + (4 * rg + i) * 64 + 32 * ch + 4 * cg) =
          make_float4(acc[i][0].x, acc[i][0].y, acc[i][1].x, acc[i][1].y);
    __syncthreads();
    PD_T(8);
    // ---- cell update for (crow, cu)                                        (Keras LSTMCell.call, gate order i f c~ o)
    if (tid < 256) {
      float z[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float s = zx[q];
#pragma unroll
        for (int k2 = 0; k2 < 8; ++k2) s += red[k2 * 1024 + crow * 64 + q * 16 + cul];
        z[q] = s;
      }
      const float gi = pd_sigmoid(z[0]), gf = pd_sigmoid(z[1]), gc = pd_tanh(z[2]), go = pd_sigmoid(z[3]);
      creg = gf * creg + gi * gc;
      const float h = go * pd_tanh(creg);
      if (cvalid) {
        const size_t e = (size_t)cb * U + cu;
        if (p.gates) {
          float* gr = p.gates + ((size_t)t * B + cb) * 4 * U + cu;
          gr[0] = gi; gr[U] = gf; gr[2 * U] = gc; gr[3 * U] = go;
        }
        p.Call[(size_t)(t + 1) * BU + e] = creg;
        p.Hall[(size_t)(t + 1) * BU + e] = h;
        if (p.hout) p.hout[(size_t)t * BU + e] = h * drop_mult(dout, (uint64_t)e);
      }
      if (t + 1 < T) p.XT[(((size_t)(t + 1) * p.NGtot + gg) * K + G + cu) * 16 + crow] = cvalid ? h : 0.0f;
    }
    PD_T(9);
    pd_arrive(barH);
  }
}

// ------------------------------------------------------------------------------------------------
// backward
// ------------------------------------------------------------------------------------------------
template <int U>
__global__ void __launch_bounds__(PD_THREADS, 1) pd_bwd_kernel(PdBwd p) {
  constexpr int G = 32, A = 32, K = G + U, NCG = U / 16, KP = K + 4;
  extern __shared__ __align__(16) float pd_smem[];
  float* WsT = pd_smem;                // [64][KP]  resident weight slice, column-major: WsT[c][k] = [W_ctx;U][k][col c]
  float* dzs = WsT + 64 * KP;          // [64][16]  dz_t of the own columns, column-major over the 16 rows
  float* red = dzs + 64 * 16;          // [16][16*32] split partials of the d ctx tile
  float* W2o = red + 16 * 512;         // [16][33]  W2 rows of the own units
  float* dqs = W2o + 16 * 33;          // [16][32]
  float* tmp = dqs + 16 * 32;          // [256]
  float* part = tmp + 256;             // [3][16][33] + [16]
  __shared__ int s_ok;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = blockIdx.x / NCG, j = blockIdx.x % NCG;
  const int gg = p.row0 / 16 + g;
  const int nrg = min(16, p.nrows - 16 * g);
  // attention: sample jj of the group is shared by CTAs jj (regions [0, Rh)) and 16 + jj (regions [Rh, R))
  const int jj = j & 15, hf = j >> 4;
  const bool has_sample = jj < nrg;
  const int b = p.row0 + 16 * g + jj;
  const int B = p.B, T = p.T, R = p.R;
  const int Rh = (((R + 1) >> 1) + 3) & ~3;
  const int r0 = hf * Rh;
  const int nloc = max(0, min(Rh, R - r0));
  const int sub = lane >> 3, a4 = lane & 7;
  const int crow = (tid >> 4) & 15, cul = tid & 15;
  const bool cvalid = crow < nrg;
  const int cb = p.row0 + 16 * g + crow, cu = 16 * j + cul;
  unsigned* bar1 = p.bar + 4 * g + 2;
  unsigned* bar2 = p.bar + 4 * g + 3;
  const size_t BU = (size_t)B * U;
  float* Pg = p.P + (size_t)gg * NCG * 16 * K;            // partials of this group: [NCG][16][K]
  float* Pj = Pg + (size_t)j * 16 * K;

  for (int idx = tid; idx < K * 64; idx += PD_THREADS) {
    const int k = idx >> 6, c = idx & 63;
    const float* src = (k < G ? p.lW + (size_t)k * 4 * U : p.lU + (size_t)(k - G) * 4 * U) + (c >> 4) * U + 16 * j + (c & 15);
    WsT[c * KP + k] = __ldg(src);
  }
  for (int idx = tid; idx < 16 * A; idx += PD_THREADS) W2o[(idx >> 5) * 33 + (idx & 31)] = __ldg(p.W2 + (size_t)(16 * j + (idx >> 5)) * A + (idx & 31));
  float pf[PD_HQ][4], ft[PD_HQ][4], dpf[PD_HQ][4];
#pragma unroll
  for (int i = 0; i < PD_HQ; ++i) {
    const int rl = 4 * (warp + 16 * i) + sub;
#pragma unroll
    for (int c = 0; c < 4; ++c) { pf[i][c] = 0.0f; ft[i][c] = 0.0f; dpf[i][c] = 0.0f; }
    if (has_sample && rl < nloc) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(p.Pf + ((size_t)b * R + r0 + rl) * A) + a4);
      const float4 f = __ldg(reinterpret_cast<const float4*>(p.feats + ((size_t)b * R + r0 + rl) * G) + a4);
      pf[i][0] = a.x; pf[i][1] = a.y; pf[i][2] = a.z; pf[i][3] = a.w;
      ft[i][0] = f.x; ft[i][1] = f.y; ft[i][2] = f.z; ft[i][3] = f.w;
    }
  }
  float v4[4];
#pragma unroll
  for (int c = 0; c < 4; ++c) v4[c] = __ldg(p.Vw + 4 * a4 + c);
  float dc = 0.0f, dh_rec = 0.0f;
  // operands of the pointwise LSTM gradient of step t: gates i f c~ o, d hout, c_t, c_{t-1}
  float pre[7] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
  auto prefetch = [&](int t) {
    if (tid < 256 && cvalid) {
      const size_t e = (size_t)cb * U + cu;
      const float* gr = p.gates + ((size_t)t * B + cb) * 4 * U + cu;
      pre[0] = __ldg(gr); pre[1] = __ldg(gr + U); pre[2] = __ldg(gr + 2 * U); pre[3] = __ldg(gr + 3 * U);
      pre[4] = __ldg(p.dhout + (size_t)t * BU + e);
      pre[5] = __ldg(p.Call + (size_t)(t + 1) * BU + e);
      pre[6] = __ldg(p.Call + (size_t)t * BU + e);
    }
  };
  prefetch(T - 1);
  __syncthreads();

  for (int s = 0; s < T; ++s) {
    const int t = T - 1 - s;
    DropSite dattn = p.dattn, dlin = p.dlin, dout = p.dout;
    dattn.step = (uint32_t)t; dlin.step = (uint32_t)t; dout.step = (uint32_t)t;
    PD_T(0);
    // ---- dz_t for the own units (Keras LSTMCell tape gradient) ---------------------------------------------
    if (tid < 256) {
      float dz[4] = {0.0f, 0.0f, 0.0f, 0.0f};
      if (cvalid) {
        const float gi = pre[0], gf = pre[1], gc = pre[2], go = pre[3];
        const float dh = pre[4] * drop_mult(dout, (uint64_t)cb * U + cu) + dh_rec;
        const float tch = pd_tanh(pre[5]);
        const float cprev = pre[6];
        const float dcc = dc + dh * go * (1.0f - tch * tch);
        dz[0] = dcc * gc * gi * (1.0f - gi);
        dz[1] = dcc * cprev * gf * (1.0f - gf);
        dz[2] = dcc * gi * (1.0f - gc * gc);
        dz[3] = dh * tch * go * (1.0f - go);
        dc = dcc * gf;
        float* dzr = p.dZ + ((size_t)t * B + cb) * 4 * U + cu;
        dzr[0] = dz[0]; dzr[U] = dz[1]; dzr[2 * U] = dz[2]; dzr[3 * U] = dz[3];
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) dzs[(q * 16 + cul) * 16 + crow] = dz[q];
      if (t > 0) prefetch(t - 1);                                   // lands while the partial product runs
    }
    __syncthreads();
    PD_T(1);
    // ---- partial dx[16 rows, K] = dz_t[:, own 64 columns] W_slice^T.  The d ctx tile (k < G) first: it is all the
    //      attention gradient waits for; the 512 h columns (the bulk) are multiplied while barrier B1 settles and are
    //      consumed only after barrier B2.  P is double-buffered by step parity so late readers are never overtaken.
    float* Pgs = Pg + (size_t)(s & 1) * p.NGtot * NCG * 16 * K;
    float* Pjs = Pgs + (size_t)j * 16 * K;
    const int rg = lane >> 3, kg = lane & 7;
    {
      float actx[4][4];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int c = 0; c < 4; ++c) actx[i][c] = 0.0f;
      // every warp takes 4 of the 64 columns, summed through shared memory
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) {
        const int c = 4 * warp + cc;
        const float4 x = *reinterpret_cast<const float4*>(dzs + c * 16 + 4 * rg);
        const float4 w = *reinterpret_cast<const float4*>(WsT + c * KP + 4 * kg);
        const float xv[4] = {x.x, x.y, x.z, x.w}, wv[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int k = 0; k < 4; ++k) actx[i][k] = fmaf(xv[i], wv[k], actx[i][k]);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
        *reinterpret_cast<float4*>(red + warp * 512 + (4 * rg + i) * 32 + 4 * kg) =
            make_float4(actx[i][0], actx[i][1], actx[i][2], actx[i][3]);
    }
    __syncthreads();
    {
      float v = 0.0f;
#pragma unroll
      for (int w = 0; w < 16; ++w) v += red[w * 512 + tid];
      Pjs[(size_t)(tid >> 5) * K + (tid & 31)] = v;
    }
    PD_T(2);
    pd_arrive(bar1);
    if (32 * warp < U) {
      float2 acc[4][2];
#pragma unroll
      for (int i = 0; i < 4; ++i) { acc[i][0] = make_float2(0.0f, 0.0f); acc[i][1] = make_float2(0.0f, 0.0f); }
      const float* wp = WsT + G + 32 * warp + 4 * kg;
#pragma unroll 4
      for (int c = 0; c < 64; ++c) {
        const float4 x = *reinterpret_cast<const float4*>(dzs + c * 16 + 4 * rg);
        const float4 w = *reinterpret_cast<const float4*>(wp + c * KP);
        const float xv[4] = {x.x, x.y, x.z, x.w};
        const float2 w01 = make_float2(w.x, w.y), w23 = make_float2(w.z, w.w);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 xx = make_float2(xv[i], xv[i]);
          acc[i][0] = __ffma2_rn(xx, w01, acc[i][0]);
          acc[i][1] = __ffma2_rn(xx, w23, acc[i][1]);
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
        *reinterpret_cast<float4*>(Pjs + (size_t)(4 * rg + i) * K + G + 32 * warp + 4 * kg) =
            make_float4(acc[i][0].x, acc[i][0].y, acc[i][1].x, acc[i][1].y);
    }
    // operands of the attention backward that do not depend on the barrier
    float al[PD_HQ], ae[PD_HQ];
    float q = 0.0f, craw = 0.0f, se = 0.0f;
    if (has_sample) {
#pragma unroll
      for (int i = 0; i < PD_HQ; ++i) {
        const int rl = 4 * (warp + 16 * i) + sub;
        al[i] = rl < nloc ? __ldg(p.alpha + ((size_t)t * B + b) * R + r0 + rl) : 0.0f;
        ae[i] = (p.attn_e != nullptr && rl < nloc) ? __ldg(p.attn_e + ((size_t)t * B + b) * R + r0 + rl) : 0.0f;
      }
      if (p.attn_e != nullptr) se = __ldg(p.attn_Se + (size_t)t * B + b);
      q = __ldg(p.qact + ((size_t)t * B + b) * A + lane);
      craw = __ldg(p.ctxraw + ((size_t)t * B + b) * G + lane);
    }
    if (!pd_wait(bar1, (unsigned)(s + 1) * NCG, p.abort_flag, &s_ok)) return;
    PD_T(3);

    if (has_sample) {
      // ---- d ctx_t of sample b = sum over the group's partials, then the LSTM input-dropout mask
      float v = 0.0f;
      for (int jp = warp; jp < NCG; jp += 16) v += __ldcg(Pgs + ((size_t)jp * 16 + jj) * K + lane);
      part[warp * 33 + lane] = v;
      __syncthreads();
      float dctx = 0.0f;
#pragma unroll
      for (int w = 0; w < 16; ++w) dctx += part[w * 33 + lane];
      dctx *= drop_mult(dlin, (uint64_t)b * p.GE + lane);
      if (warp == 0 && hf == 0) p.dxctx[((size_t)t * B + b) * G + lane] = dctx;
      float d4[4], q4[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        d4[c] = __shfl_sync(0xffffffffu, dctx, 4 * a4 + c);
        q4[c] = __shfl_sync(0xffffffffu, q, 4 * a4 + c);
      }
      // ---- d alpha_r = f_r . d ctx ; S = sum_r alpha_r d alpha_r = ctx . d ctx over ALL regions (softmax tape
      //      gradient): the context itself was saved by the forward pass, so neither half needs the other
      float da[PD_HQ];
#pragma unroll
      for (int i = 0; i < PD_HQ; ++i) {
        float d = 0.0f;
#pragma unroll
        for (int c = 0; c < 4; ++c) d = fmaf(ft[i][c], d4[c], d);
        d += __shfl_xor_sync(0xffffffffu, d, 1);
        d += __shfl_xor_sync(0xffffffffu, d, 2);
        d += __shfl_xor_sync(0xffffffffu, d, 4);
        da[i] = d + ae[i];                       // + direct gradient of the attention loss on alpha_r, if enabled
      }
      const float S = warp_sum(craw * dctx) + se;
      // ---- d e_r = alpha_r (d alpha_r - S); through V, the dropout mask and tanh     (attention.py:30-38)
      float dq4[4] = {0.0f, 0.0f, 0.0f, 0.0f}, dV4[4] = {0.0f, 0.0f, 0.0f, 0.0f};
      float dbv = 0.0f;
#pragma unroll
      for (int i = 0; i < PD_HQ; ++i) {
        const int rl = 4 * (warp + 16 * i) + sub;
        const int r = r0 + rl;
        if (rl < nloc) {
          const float de = al[i] * (da[i] - S);
          if (a4 == 0) dbv += de;
          float m[4];
          drop_mult4(dattn, (((uint64_t)b * R + r) * A) / 4 + a4, m);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const float sv = pd_tanh(pf[i][c] + q4[c]);
            dV4[c] = fmaf(de, sv * m[c], dV4[c]);
            const float dpre = de * v4[c] * m[c] * (1.0f - sv * sv);
            dpf[i][c] += dpre;
            dq4[c] += dpre;
          }
        }
      }
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        dq4[c] += __shfl_xor_sync(0xffffffffu, dq4[c], 8);
        dq4[c] += __shfl_xor_sync(0xffffffffu, dq4[c], 16);
        dV4[c] += __shfl_xor_sync(0xffffffffu, dV4[c], 8);
        dV4[c] += __shfl_xor_sync(0xffffffffu, dV4[c], 16);
      }
      dbv = warp_sum(dbv);
      if (lane < 8) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          part[(16 + warp) * 33 + 4 * lane + c] = dq4[c];
          part[(32 + warp) * 33 + 4 * lane + c] = dV4[c];
        }
      }
      if (lane == 0) part[3 * 16 * 33 + 16 + warp] = dbv;
      __syncthreads();
      if (warp == 0) {
        float tq = 0.0f, tv = 0.0f, tb = 0.0f;
#pragma unroll
        for (int w = 0; w < 16; ++w) {
          tq += part[(16 + w) * 33 + lane];
          tv += part[(32 + w) * 33 + lane];
          tb += part[3 * 16 * 33 + 16 + w];
        }
        // the two halves' sums meet in their consumers: d q after barrier B2, d V / d bV in the column sums
        p.DQ[(((size_t)t * p.NGtot + gg) * 2 + hf) * 512 + jj * 32 + lane] = tq * lrelu_grad_s(q, p.att_slope);
        p.dVpart[(((size_t)hf * T + t) * B + b) * A + lane] = tv;
        if (lane == 0) p.dbVpart[((size_t)hf * T + t) * B + b] = tb;
      }
    }
    PD_T(4);
    pd_arrive(bar2);
    if (!pd_wait(bar2, (unsigned)(s + 1) * NCG, p.abort_flag, &s_ok)) return;
    PD_T(5);
    // ---- dh_{t-1}[rows, own units] from the recurrent kernel: sum of the group's partials -------------------
    {
      const int half = tid >> 8;
      float v = 0.0f;
      for (int jp = half * (NCG / 2); jp < (half + 1) * (NCG / 2); ++jp)
        v += __ldcg(Pgs + ((size_t)jp * 16 + crow) * K + G + cu);
      if (half == 1) tmp[tid & 255] = v;
      __syncthreads();
      if (half == 0) dh_rec = v + tmp[tid];
    }
    PD_T(6);
    // ---- d q_t of the 16 rows = sum of the two half-attentions; CTA 0 of the group publishes it for dW2 / db2
    {
      const int row = tid >> 5;
      const float* dq0 = p.DQ + ((size_t)t * p.NGtot + gg) * 2 * 512;
      const float v = row < nrg ? __ldcg(dq0 + tid) + __ldcg(dq0 + 512 + tid) : 0.0f;
      dqs[tid] = v;
      if (j == 0 && row < nrg) p.dqpre[((size_t)t * B + p.row0 + 16 * g + row) * A + lane] = v;
    }
    if (s == T - 1) break;
    __syncthreads();
    // ---- dh_{t-1} += dq_t W2[own units]^T                                           (attention.py:32)
    if (tid < 256) {
      float a = 0.0f;
#pragma unroll
      for (int k = 0; k < A; ++k) a = fmaf(dqs[crow * 32 + k], W2o[cul * 33 + k], a);
      dh_rec += a;
    }
  }
  // d Pf accumulated over the steps                                                   (attention.py:31)
  if (has_sample) {
#pragma unroll
    for (int i = 0; i < PD_HQ; ++i) {
      const int rl = 4 * (warp + 16 * i) + sub;
      if (rl < nloc)
        *(reinterpret_cast<float4*>(p.dPf + ((size_t)b * R + r0 + rl) * A) + a4) = make_float4(dpf[i][0], dpf[i][1], dpf[i][2], dpf[i][3]);
    }
  }
}

// d feats[b][r][:] = sum_t alpha_t[b][r] * d ctx_t[b][:]                              (attention.py:42)
__global__ void __launch_bounds__(256)
att_dfeats_kernel(int B, int T, int R, int G, const float* __restrict__ alpha, const float* __restrict__ dxctx,
                  float* __restrict__ dfeats) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= (int64_t)B * R * G) return;
  const int gq = (int)(e % G);
  const int r = (int)((e / G) % R);
  const int b = (int)(e / ((int64_t)G * R));
  float s = 0.0f;
  for (int t = 0; t < T; ++t)
    s = fmaf(alpha[((size_t)t * B + b) * R + r], dxctx[((size_t)t * B + b) * G + gq], s);
  dfeats[e] = s;
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static constexpr int PD_U = 512, PD_K = 32 + PD_U, PD_NCG = PD_U / 16;
static constexpr size_t PD_FWD_SMEM = sizeof(float) * (PD_K * 64 + PD_K * 16 + 8 * 1024 + PD_MAXR + 16 * 33);
static constexpr size_t PD_BWD_SMEM =
    sizeof(float) * (64 * (PD_K + 4) + 64 * 16 + 16 * 512 + 16 * 33 + 16 * 32 + 256 + 3 * 16 * 33 + 32);

bool pd_shape_ok(const tnt_config& f) { return f.U == PD_U && f.G == 32 && f.A == 32 && f.R <= PD_MAXR; }

// workspace sizes in bytes for tnt_create
int64_t pd_ws_bytes(const tnt_config& f, const char* what) {
  const int64_t ng = (f.B_max + 15) / 16;
  if (!strcmp(what, "pd_XT")) return (int64_t)f.T * ng * PD_K * 16 * sizeof(float);
  if (!strcmp(what, "pd_P")) return 2 * ng * PD_NCG * 16 * PD_K * sizeof(float);   // double-buffered by step parity
  if (!strcmp(what, "pd_DQ")) return (int64_t)f.T * ng * 2 * 512 * sizeof(float);
  if (!strcmp(what, "pd_AX")) return (int64_t)f.T * ng * 2 * 16 * PD_AXW * sizeof(float);
  if (!strcmp(what, "pd_bar")) return (4 * ng + 16) * sizeof(unsigned);
  if (!strcmp(what, "pd_trace")) return 4 * 64 * 16 * sizeof(long long);
  if (!strcmp(what, "pd_dxctx") || !strcmp(what, "pd_ctx")) return (int64_t)f.T * f.B_max * f.G * sizeof(float);
  return 0;
}

// The persistent kernels cover U = 512, G = A = 32, R <= 384 (every BASELINE configuration); other shapes (the tiny
// test configurations) run the per-step kernels of decoder.cu by design.  Nothing is selected at run time: a shape the
// persistent kernels cover either runs them or fails the call.
bool pd_enabled(tnt_ctx* c) {
  return pd_shape_ok(c->cfg) && ws(c, "pd_XT") != nullptr && c->sm_count >= PD_NCG;
}

static bool pd_prepare(tnt_ctx* c) {
  if (c->pd_ready) return true;
  cudaError_t e1 = cudaFuncSetAttribute(pd_fwd_kernel<PD_U>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PD_FWD_SMEM);
  cudaError_t e2 = cudaFuncSetAttribute(pd_bwd_kernel<PD_U>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PD_BWD_SMEM);
  int coop = 0;
  cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, c->device);
  if (e1 != cudaSuccess || e2 != cudaSuccess || !coop) {
    cudaGetLastError();
    ctx_fail(c, TNT_ERR_CUDA, std::string("persistent decoder unavailable: ") +
                                  (!coop ? "no cooperative launch" : cudaGetErrorString(e1 != cudaSuccess ? e1 : e2)));
    return false;
  }
  c->pd_ready = 1;
  return true;
}

// one sticky device flag per context: set by any kernel whose inter-CTA wait timed out (persistent recurrence, fused head,
// one-shot all-reduce); the optimizer refuses to apply gradients while it is set
unsigned* pd_abort_flag(tnt_ctx* c) { return wsT<unsigned>(c, "abort_flag"); }

// the T-step recurrence for all B rows; Zx, Hall[0], Call[0] must be ready.  false => the context is marked failed
bool pd_forward(tnt_ctx* c, int B, int T, int training, uint64_t seed, float* gates, float* hout, float* alpha,
                float* s_out, cudaStream_t st) {
  if (!pd_prepare(c)) return false;
  const tnt_config& f = c->cfg;
  const int ng_per_launch = c->sm_count / PD_NCG;
  const int ng_tot = (B + 15) / 16;
  PdFwd p;
  p.B = B; p.T = T; p.R = f.R; p.GE = f.G + f.E; p.NGtot = ng_tot;
  p.W2 = c->w + c->o_W2; p.b2 = c->w + c->o_b2; p.Vw = c->w + c->o_V; p.bV = c->w + c->o_bV;
  p.lW = c->w + c->o_lW; p.lU = c->w + c->o_lU;
  p.Zx = wsT<float>(c, "Zx"); p.Pf = wsT<float>(c, "Pf"); p.feats = wsT<float>(c, "feats");
  p.Hall = wsT<float>(c, "Hall"); p.Call = wsT<float>(c, "Call"); p.gates = gates; p.alpha = alpha;
  p.qact = wsT<float>(c, "qact"); p.xctx = wsT<float>(c, "xctx"); p.hout = hout; p.XT = wsT<float>(c, "pd_XT"); p.AX = wsT<float>(c, "pd_AX"); p.ctxraw = wsT<float>(c, "pd_ctx");
  p.s_out = s_out;
  p.bar = wsT<unsigned>(c, "pd_bar"); p.abort_flag = pd_abort_flag(c);
  p.trace = getenv("TNT_PD_TRACE") ? wsT<long long>(c, "pd_trace") : nullptr;
  p.dattn = make_drop(f.dropout_attn, training, seed, SITE_ATTN, 0);
  p.dlin = make_drop(f.dropout_lstm, training, seed, SITE_LSTM_IN, 0);
  p.dout = make_drop(f.dropout_lstm, training, seed, SITE_LSTM_OUT, 0);
  p.att_slope = att_slope_of(f);
  for (int g0 = 0; g0 < ng_tot; g0 += ng_per_launch) {
    const int ng = std::min(ng_per_launch, ng_tot - g0);
    p.row0 = 16 * g0; p.nrows = std::min(B - p.row0, 16 * ng);
    cudaMemsetAsync(p.bar, 0, sizeof(unsigned) * 4 * ng, st);
    void* args[] = {&p};
    cudaError_t e = cudaLaunchCooperativeKernel((const void*)pd_fwd_kernel<PD_U>, dim3(ng * PD_NCG), dim3(PD_THREADS),
                                                args, PD_FWD_SMEM, st);
    if (e != cudaSuccess) {
      cudaGetLastError();
      ctx_fail(c, TNT_ERR_CUDA, std::string("pd_forward: cooperative launch failed: ") + cudaGetErrorString(e));
      return false;
    }
    c->launches += 1;
  }
  return true;
}

bool pd_backward(tnt_ctx* c, int B, int T, uint64_t seed, cudaStream_t st) {
  if (!pd_prepare(c)) return false;
  const tnt_config& f = c->cfg;
  const int ng_per_launch = c->sm_count / PD_NCG;
  const int ng_tot = (B + 15) / 16;
  PdBwd p;
  p.B = B; p.T = T; p.R = f.R; p.GE = f.G + f.E; p.NGtot = ng_tot;
  p.W2 = c->w + c->o_W2; p.Vw = c->w + c->o_V; p.lW = c->w + c->o_lW; p.lU = c->w + c->o_lU;
  p.Pf = wsT<float>(c, "Pf"); p.feats = wsT<float>(c, "feats"); p.gates = wsT<float>(c, "gates");
  p.Call = wsT<float>(c, "Call"); p.dhout = wsT<float>(c, "dhout"); p.alpha = wsT<float>(c, "alpha");
  p.qact = wsT<float>(c, "qact");
  p.dZ = wsT<float>(c, "dZ"); p.dPf = wsT<float>(c, "dPf"); p.dqpre = wsT<float>(c, "dqpre");
  p.dVpart = wsT<float>(c, "dVpart"); p.dbVpart = wsT<float>(c, "dbVpart"); p.dxctx = wsT<float>(c, "pd_dxctx");
  p.P = wsT<float>(c, "pd_P"); p.DQ = wsT<float>(c, "pd_DQ"); p.ctxraw = wsT<float>(c, "pd_ctx");
  p.attn_e = c->cfg.attn_loss ? wsT<float>(c, "attn_e") : nullptr;
  p.attn_Se = c->cfg.attn_loss ? wsT<float>(c, "attn_Se") : nullptr;
  p.bar = wsT<unsigned>(c, "pd_bar"); p.abort_flag = pd_abort_flag(c);
  p.trace = getenv("TNT_PD_TRACE") ? wsT<long long>(c, "pd_trace") + 2 * 64 * 16 : nullptr;
  p.dattn = make_drop(f.dropout_attn, 1, seed, SITE_ATTN, 0);
  p.dlin = make_drop(f.dropout_lstm, 1, seed, SITE_LSTM_IN, 0);
  p.dout = make_drop(f.dropout_lstm, 1, seed, SITE_LSTM_OUT, 0);
  p.att_slope = att_slope_of(f);
  for (int g0 = 0; g0 < ng_tot; g0 += ng_per_launch) {
    const int ng = std::min(ng_per_launch, ng_tot - g0);
    p.row0 = 16 * g0; p.nrows = std::min(B - p.row0, 16 * ng);
    cudaMemsetAsync(p.bar, 0, sizeof(unsigned) * 4 * ng, st);
    void* args[] = {&p};
    cudaError_t e = cudaLaunchCooperativeKernel((const void*)pd_bwd_kernel<PD_U>, dim3(ng * PD_NCG), dim3(PD_THREADS),
                                                args, PD_BWD_SMEM, st);
    if (e != cudaSuccess) {
      cudaGetLastError();
      ctx_fail(c, TNT_ERR_CUDA, std::string("pd_backward: cooperative launch failed: ") + cudaGetErrorString(e));
      return false;
    }
    c->launches += 1;
  }
  return true;                                       // d feats = sum_t alpha_t d ctx_t is folded into att_w1_bwd_kernel
}

// sticky: set by a kernel whose barrier wait timed out
int pd_read_abort(tnt_ctx* c, cudaStream_t st) {
  if (!ws(c, "abort_flag")) return 0;
  unsigned h = 0;
  cudaMemcpyAsync(&h, pd_abort_flag(c), sizeof(h), cudaMemcpyDeviceToHost, st);
  cudaStreamSynchronize(st);
  return (int)h;
}
