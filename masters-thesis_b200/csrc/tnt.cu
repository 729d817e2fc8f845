// libtnt.so host side: context, flat parameter layout, workspaces and the C ABI of include/tnt.h.
#include "tnt_internal.h"
#include <cstdio>
#include <cstring>
#include <cmath>
#include <algorithm>
#include <cstdlib>

#define ENC_KSPLIT 512     // indices of one parcel handled by one encoder CTA
#define OPT_CHUNK 8192     // floats per optimizer CTA

static std::string g_create_err;

void* ws(tnt_ctx* c, const char* name) {
  auto it = c->bufs.find(name);
  return it == c->bufs.end() ? nullptr : it->second.first;
}

static int fail(tnt_ctx* c, int code, const std::string& msg) {
  if (c) c->err = msg; else g_create_err = msg;
  return code;
}

static int check_cuda(tnt_ctx* c, const char* where) {
  if (c && c->fail_code) { cudaGetLastError(); return c->fail_code; }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(c, TNT_ERR_CUDA, std::string(where) + ": " + cudaGetErrorString(e));
  return TNT_OK;
}

static bool alloc(tnt_ctx* c, const char* name, int64_t bytes) {
  void* p = nullptr;
  if (bytes < 16) bytes = 16;
  if (cudaMalloc(&p, (size_t)bytes) != cudaSuccess) {
    c->err = std::string("cudaMalloc failed for workspace '") + name + "' (" + std::to_string(bytes) + " bytes)";
    cudaGetLastError();
    return false;
  }
  cudaMemset(p, 0, (size_t)bytes);
  c->bufs[name] = std::make_pair(p, bytes);
  return true;
}

static int64_t align4(int64_t x) { return (x + 3) & ~(int64_t)3; }

static void add_var(tnt_ctx* c, const std::string& name, int64_t off, int64_t rows, int64_t cols, float l2,
                    int trainable, int is_emb = 0) {
  tnt_var_info v;
  memset(&v, 0, sizeof(v));
  snprintf(v.name, sizeof(v.name), "%s", name.c_str());
  v.offset = off; v.rows = rows; v.cols = cols; v.l2 = l2; v.trainable = trainable;
  c->vars.push_back(v);
  if (trainable) {
    Seg s; s.off = off; s.len = rows * (cols > 0 ? cols : 1); s.l2 = l2; s.is_emb = is_emb;
    s.cols = (int32_t)cols; s.unit0 = 0;
    c->segs.push_back(s);
  }
}

extern "C" int tnt_create(const tnt_config* cfg, const int32_t* gidx, const int64_t* goff, tnt_ctx** out) {
  if (!cfg || !gidx || !goff || !out) return fail(nullptr, TNT_ERR_ARG, "tnt_create: null argument");
  const tnt_config& f = *cfg;
  if (f.G < 1 || f.G > 32 || f.A < 1 || f.A > 32)
    return fail(nullptr, TNT_ERR_ARG, "tnt_create: group_size and attn_units must be in [1,32]");
  if (f.n_enc < 1 || f.n_enc > 2 || f.R < 1 || f.U < 1 || f.E < 1 || f.H < 1 || f.V < 2 || f.T < 1 || f.B_max < 1 ||
      f.N < 1)
    return fail(nullptr, TNT_ERR_ARG, "tnt_create: bad dimension");
  if (f.precision != TNT_FP32 && f.precision != TNT_BF16X3 && f.precision != TNT_BF16)
    return fail(nullptr, TNT_ERR_ARG, "tnt_create: precision must be TNT_FP32, TNT_BF16X3 or TNT_BF16");
  if (f.attn_loss < 0 || f.attn_loss > 2 || (f.attn_loss && f.n_enc != 1))
    return fail(nullptr, TNT_ERR_ARG, "tnt_create: attn_loss must be 0, 1 or 2 and needs the single-encoder model");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return fail(nullptr, TNT_ERR_NO_DEVICE, "tnt_create: no CUDA device (libtnt has no CPU fallback)");
  }
  tnt_ctx* c = new tnt_ctx();
  c->cfg = f;
  cudaGetDevice(&c->device);
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, c->device);
  if (prop.major < 10) {
    delete c;
    return fail(nullptr, TNT_ERR_NO_DEVICE, "tnt_create: device is not sm_100 (libtnt is built for sm_100a only)");
  }
  c->sm_count = prop.multiProcessorCount;
  {
    // low-priority side stream for everything off the critical path of the step (weight gradients, metrics, the
    // next step's decoder prologue, the decoder variables' update)
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    if (cudaStreamCreateWithPriority(&c->side, cudaStreamNonBlocking, lo) != cudaSuccess) { cudaGetLastError(); c->side = nullptr; }
    if (c->side && (cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
                    cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming) != cudaSuccess ||
                    cudaEventCreateWithFlags(&c->ev_pro, cudaEventDisableTiming) != cudaSuccess ||
                    cudaEventCreateWithFlags(&c->ev_head, cudaEventDisableTiming) != cudaSuccess ||
                    cudaEventCreateWithFlags(&c->ev_w1, cudaEventDisableTiming) != cudaSuccess ||
                    cudaEventCreateWithFlags(&c->ev_metrics, cudaEventDisableTiming) != cudaSuccess)) {
      cudaGetLastError(); cudaStreamDestroy(c->side); c->side = nullptr;
    }
    if (!c->side) { delete c; return fail(nullptr, TNT_ERR_CUDA, "tnt_create: cannot create the side stream"); }
  }

  // ---- encoder tables + flat layout ------------------------------------------------------------
  int64_t off = 0, soff = 0;
  const int R = f.R, G = f.G;
  int max_slots = 1;
  for (int e = 0; e < f.n_enc; ++e) {
    EncTables& t = c->enc[e];
    const int64_t* go = goff + (int64_t)e * R;
    const int64_t base = go[0];
    t.n_lab = go[R] - base;
    std::vector<int64_t> loc(R + 1);
    std::vector<EncWork> work;
    std::vector<int32_t> slot_first(R + 1);
    int slots = 0;
    for (int r = 0; r <= R; ++r) loc[r] = go[r] - base;
    for (int r = 0; r < R; ++r) {
      int64_t n = loc[r + 1] - loc[r];
      if (n < 1) { delete c; return fail(nullptr, TNT_ERR_ARG, "tnt_create: empty parcel"); }
      for (int64_t k = 0; k < n; ++k) {
        int32_t v = gidx[base + loc[r] + k];
        if (v < 0 || v >= f.N) { delete c; return fail(nullptr, TNT_ERR_ARG, "tnt_create: parcel index outside [0,N)"); }
      }
      slot_first[r] = slots;
      int ns = (int)((n + ENC_KSPLIT - 1) / ENC_KSPLIT);
      for (int s = 0; s < ns; ++s) {
        EncWork w;
        w.r = r; w.kbeg = s * ENC_KSPLIT; w.kend = (int32_t)std::min<int64_t>(n, (int64_t)(s + 1) * ENC_KSPLIT); w.slot = s;
        work.push_back(w);
      }
      slots += ns;
    }
    slot_first[R] = slots;
    // largest work first: better tail behaviour
    std::stable_sort(work.begin(), work.end(), [](const EncWork& a, const EncWork& b) {
      return (a.kend - a.kbeg) > (b.kend - b.kbeg);
    });
    t.n_work = (int)work.size(); t.n_slots = slots;
    max_slots = std::max(max_slots, slots);
    cudaMalloc(&t.idx, sizeof(int32_t) * t.n_lab);
    cudaMalloc(&t.off, sizeof(int64_t) * (R + 1));
    cudaMalloc(&t.work, sizeof(EncWork) * work.size());
    cudaMalloc(&t.slot_first, sizeof(int32_t) * (R + 1));
    {
      // inverse of the index list: where every input vertex goes in the parcel-ordered copy of the batch (a vertex may
      // belong to several parcels -- overlapping groups -- or to none)
      std::vector<int64_t> ioff((size_t)f.N + 1, 0);
      for (int64_t k = 0; k < t.n_lab; ++k) ioff[(size_t)gidx[base + k] + 1] += 1;
      for (int64_t v = 0; v < f.N; ++v) ioff[(size_t)v + 1] += ioff[(size_t)v];
      std::vector<int32_t> ipos((size_t)t.n_lab);
      std::vector<int64_t> fill(ioff.begin(), ioff.end() - 1);
      for (int64_t k = 0; k < t.n_lab; ++k) ipos[(size_t)fill[(size_t)gidx[base + k]]++] = (int32_t)k;
      cudaMalloc(&t.inv_off, sizeof(int64_t) * ((size_t)f.N + 1));
      cudaMalloc(&t.inv_pos, sizeof(int32_t) * (size_t)t.n_lab);
      cudaMemcpy(t.inv_off, ioff.data(), sizeof(int64_t) * ((size_t)f.N + 1), cudaMemcpyHostToDevice);
      cudaMemcpy(t.inv_pos, ipos.data(), sizeof(int32_t) * (size_t)t.n_lab, cudaMemcpyHostToDevice);
    }
    cudaMemcpy(t.idx, gidx + base, sizeof(int32_t) * t.n_lab, cudaMemcpyHostToDevice);
    cudaMemcpy(t.off, loc.data(), sizeof(int64_t) * (R + 1), cudaMemcpyHostToDevice);
    cudaMemcpy(t.work, work.data(), sizeof(EncWork) * work.size(), cudaMemcpyHostToDevice);
    cudaMemcpy(t.slot_first, slot_first.data(), sizeof(int32_t) * (R + 1), cudaMemcpyHostToDevice);
    t.w_off = off; off = align4(off + t.n_lab * G);
    t.b_off = off; off = align4(off + (int64_t)R * G);
    t.gamma_off = off; off = align4(off + G);
    t.beta_off = off; off = align4(off + G);
    t.mean_off = soff; soff = align4(soff + G);
    t.var_off = soff; soff = align4(soff + G);
    // Keras order inside LocallyDense: dense_0 kernel, bias, dense_1 ..., then input_bn (layers.py:33,40)
    std::string p = "enc" + std::to_string(e) + "/";
    for (int r = 0; r < R; ++r) {
      add_var(c, p + "dense_" + std::to_string(r) + "/kernel", t.w_off + loc[r] * G, loc[r + 1] - loc[r], G, f.input_reg, 1);
      add_var(c, p + "dense_" + std::to_string(r) + "/bias", t.b_off + (int64_t)r * G, G, 0, 0.0f, 1);
    }
    add_var(c, p + "input_bn/gamma", t.gamma_off, G, 0, 0.0f, 1);
    add_var(c, p + "input_bn/beta", t.beta_off, G, 0, 0.0f, 1);
    add_var(c, p + "input_bn/moving_mean", t.mean_off, G, 0, 0.0f, 0);
    add_var(c, p + "input_bn/moving_variance", t.var_off, G, 0, 0.0f, 0);
  }
  c->dec_off = off;
  const int U = f.U, E = f.E, A = f.A, H = f.H, V = f.V;
  auto place = [&](int64_t n) { int64_t o = off; off = align4(off + n); return o; };
  c->o_W1 = place((int64_t)G * A); c->o_b1 = place(A);
  c->o_W2 = place((int64_t)U * A); c->o_b2 = place(A);
  c->o_V = place(A); c->o_bV = place(1);
  c->o_emb = place((int64_t)V * E);
  c->o_lW = place((int64_t)(G + E) * 4 * U); c->o_lU = place((int64_t)U * 4 * U); c->o_lb = place(4 * U);
  c->o_iW = place((int64_t)U * H); c->o_ib = place(H);
  c->o_oW = place((int64_t)H * V); c->o_ob = place(V);
  add_var(c, "attention/W1/kernel", c->o_W1, G, A, f.attn_reg, 1);
  add_var(c, "attention/W1/bias", c->o_b1, A, 0, 0.0f, 1);
  add_var(c, "attention/W2/kernel", c->o_W2, U, A, f.attn_reg, 1);
  add_var(c, "attention/W2/bias", c->o_b2, A, 0, 0.0f, 1);
  add_var(c, "attention/V/kernel", c->o_V, A, 1, 0.0f, 1);
  add_var(c, "attention/V/bias", c->o_bV, 1, 0, 0.0f, 1);
  add_var(c, "emb_text/embeddings", c->o_emb, V, E, 0.0f, 1, 1);
  add_var(c, "lstm/kernel", c->o_lW, G + E, 4 * U, f.lstm_reg, 1);
  add_var(c, "lstm/recurrent_kernel", c->o_lU, U, 4 * U, 0.0f, 1);
  add_var(c, "lstm/bias", c->o_lb, 4 * U, 0, 0.0f, 1);
  add_var(c, "time_distributed_nonlinear/kernel", c->o_iW, U, H, f.output_reg, 1);
  add_var(c, "time_distributed_nonlinear/bias", c->o_ib, H, 0, 0.0f, 1);
  add_var(c, "time_distributed_softmax/kernel", c->o_oW, H, V, f.output_reg, 1);
  add_var(c, "time_distributed_softmax/bias", c->o_ob, V, 0, 0.0f, 1);
  c->n_param = off; c->n_state = soff;

  // ---- optimizer tables ------------------------------------------------------------------------
  std::vector<OptChunk> chunks;
  for (size_t s = 0; s < c->segs.size(); ++s)
    for (int64_t o = 0; o < c->segs[s].len; o += OPT_CHUNK) {
      OptChunk ch; ch.seg = (int32_t)s; ch.off = c->segs[s].off + o;
      ch.len = (int32_t)std::min<int64_t>(OPT_CHUNK, c->segs[s].len - o);
      chunks.push_back(ch);
    }
  c->n_chunks = (int)chunks.size(); c->n_segs = (int)c->segs.size();
  c->n_chunks_enc = 0;
  for (const OptChunk& ch : chunks) if (ch.off < c->dec_off) c->n_chunks_enc += 1;      // encoder variables come first
  c->n_segs_enc = 0;
  for (const Seg& sg : c->segs) if (sg.off < c->dec_off) c->n_segs_enc += 1;
  // adaptive-clipping units (agc.py unitwise_norm): one per column of a 2-D variable, one per 1-D variable
  std::vector<AgcUnit> units;
  for (size_t si = 0; si < c->segs.size(); ++si) {
    Seg& sg = c->segs[si];
    sg.unit0 = (int32_t)units.size();
    if (sg.cols > 0)
      for (int32_t col = 0; col < sg.cols; ++col) units.push_back(AgcUnit{(int32_t)si, 0, sg.off + col, sg.len / sg.cols, sg.cols});
    else units.push_back(AgcUnit{(int32_t)si, 0, sg.off, sg.len, 1});
  }
  c->n_units = (int)units.size();
  cudaMalloc(&c->d_units, sizeof(AgcUnit) * units.size());
  cudaMemcpy(c->d_units, units.data(), sizeof(AgcUnit) * units.size(), cudaMemcpyHostToDevice);
  cudaMalloc(&c->d_segs, sizeof(Seg) * c->segs.size());
  cudaMalloc(&c->d_chunks, sizeof(OptChunk) * chunks.size());
  cudaMalloc(&c->seg_gsq, sizeof(double) * c->segs.size());
  cudaMalloc(&c->seg_wsq, sizeof(double) * c->segs.size());
  cudaMemcpy(c->d_segs, c->segs.data(), sizeof(Seg) * c->segs.size(), cudaMemcpyHostToDevice);
  cudaMemcpy(c->d_chunks, chunks.data(), sizeof(OptChunk) * chunks.size(), cudaMemcpyHostToDevice);

  // ---- workspaces ------------------------------------------------------------------------------
  const int64_t Bm = f.B_max, T = f.T, F = sizeof(float);
  bool ok = true;
  ok = ok && alloc(c, "enc_part", (int64_t)max_slots * Bm * G * F);
  for (const char* n : {"Y", "feats", "dfeats", "dYpre"}) ok = ok && alloc(c, n, Bm * R * G * F);
  for (const char* n : {"Pf", "dPf"}) ok = ok && alloc(c, n, Bm * R * A * F);
  ok = ok && alloc(c, "bn_fwd_stats", 128 * sizeof(double)) && alloc(c, "bn_bwd_stats", 128 * sizeof(double));
  ok = ok && alloc(c, "bn_coef", 256 * F);
  for (const char* n : {"xtext", "dxtext"}) ok = ok && alloc(c, n, T * Bm * E * F);
  for (const char* n : {"Zx", "gates", "dZ"}) ok = ok && alloc(c, n, T * Bm * 4 * U * F);
  ok = ok && alloc(c, "xctx", T * Bm * G * F) && alloc(c, "alpha", T * Bm * R * F);
  for (const char* n : {"qact", "dqpre"}) ok = ok && alloc(c, n, T * Bm * A * F);
  // per (step, sample) partial sums of d V / d bV; the persistent kernel writes one record per half-attention
  ok = ok && alloc(c, "dVpart", 2 * T * Bm * A * F) && alloc(c, "dbVpart", 2 * T * Bm * F);
  for (const char* n : {"Hall", "Call"}) ok = ok && alloc(c, n, (T + 1) * Bm * U * F);
  for (const char* n : {"hout", "dhout"}) ok = ok && alloc(c, n, T * Bm * U * F);
  for (const char* n : {"inter_pre", "inter_d", "dinter"}) ok = ok && alloc(c, n, T * Bm * H * F);
  ok = ok && alloc(c, "logits", T * Bm * V * F);
  ok = ok && alloc(c, "rowloss", T * Bm * F) && alloc(c, "rowcorrect", T * Bm * F);
  {
    const int64_t mt = (T * Bm + 127) / 128, nch = ((V + 127) / 128 + 2) / 3;
    ok = ok && alloc(c, "hf_stats", mt * nch * 4 * 128 * 4 * F) && alloc(c, "hf_counters", (mt + 16) * sizeof(unsigned)) &&
         alloc(c, "dob_part", mt * V * F) && alloc(c, "abort_flag", 16);
  }
  ok = ok && alloc(c, "dh_rec", Bm * (U + G) * F) && alloc(c, "dc", Bm * U * F);
  ok = ok && alloc(c, "w1part", (int64_t)4 * c->sm_count * (32 * 32 + 32) * F);
  ok = ok && alloc(c, "attn_colsum", 2 * T * R * F) && alloc(c, "metrics_raw", 16 * sizeof(double));
  ok = ok && alloc(c, "attn_e", T * Bm * R * F) && alloc(c, "attn_Se", T * Bm * F);
  for (const char* n : {"agc_gsq", "agc_wsq", "agc_factor"}) ok = ok && alloc(c, n, (int64_t)c->n_units * F);
  ok = ok && alloc(c, "emb_colsq", (int64_t)E * F) && alloc(c, "agc_emb_sq", sizeof(double));
  ok = ok && alloc(c, "word", Bm * sizeof(int32_t));
  ok = ok && alloc(c, "emb_proj", (int64_t)V * 4 * U * F);       // greedy decode: emb @ W_text + b, once per call
  if (f.G == 32) {
    // parcel-ordered copy of the batch, xP[position in the index list][64 batch rows] (<= 64 rows per encoder): both
    // encoder passes stream it; + one tile of slack so that whole-tile bulk copies never leave the buffer
    int64_t rows = 0;
    for (int e = 0; e < f.n_enc; ++e) rows += c->enc[e].n_lab + 128;
    ok = ok && alloc(c, "xT", rows * 64 * F);
  }
  if (pd_shape_ok(f))
    for (const char* n : {"pd_XT", "pd_P", "pd_AX", "pd_ctx", "pd_DQ", "pd_bar", "pd_dxctx", "pd_trace"}) ok = ok && alloc(c, n, pd_ws_bytes(f, n));
  {
    // arena for the packed (bf16-split) GEMM operands of one phase: every time-parallel operand in both orientations
    const int64_t TB = T * Bm;
    const int64_t el = 2 * (TB * E + (int64_t)E * 4 * U + TB * U + (int64_t)U * H + 2 * TB * H + (int64_t)H * V + TB * V +
                            TB * 4 * U + (int64_t)U * 4 * U) + (T + 1) * Bm * U + 8 * TB * (G + A) + 8 * (int64_t)G * 4 * U;
    tc_arena_reserve(c, (size_t)(el * 6 * 1.3) + ((size_t)32 << 20));
  }
  if (!ok || cudaGetLastError() != cudaSuccess) {
    g_create_err = c->err.empty() ? "tnt_create: CUDA allocation failed" : c->err;
    tnt_destroy(c);
    return TNT_ERR_CUDA;
  }
  *out = c;
  return TNT_OK;
}

extern "C" int tnt_destroy(tnt_ctx* c) {
  if (!c) return TNT_OK;
  for (auto& r : c->trace) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
  for (auto& kv : c->bufs) cudaFree(kv.second.first);
  for (int e = 0; e < 2; ++e) {
    cudaFree(c->enc[e].idx); cudaFree(c->enc[e].off); cudaFree(c->enc[e].work); cudaFree(c->enc[e].slot_first);
    cudaFree(c->enc[e].inv_off); cudaFree(c->enc[e].inv_pos);
  }
  if (c->side) { cudaStreamSynchronize(c->side); cudaStreamDestroy(c->side); }
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  if (c->ev_join) cudaEventDestroy(c->ev_join);
  if (c->ev_pro) cudaEventDestroy(c->ev_pro);
  if (c->ev_head) cudaEventDestroy(c->ev_head);
  if (c->ev_w1) cudaEventDestroy(c->ev_w1);
  if (c->ev_metrics) cudaEventDestroy(c->ev_metrics);
  cudaFree(c->d_segs); cudaFree(c->d_chunks); cudaFree(c->seg_gsq); cudaFree(c->seg_wsq); cudaFree(c->d_units);
  dp_destroy(c);
  tc_destroy(c);
  delete c;
  return TNT_OK;
}

extern "C" const char* tnt_last_error(tnt_ctx* c) { return c ? c->err.c_str() : g_create_err.c_str(); }
extern "C" int64_t tnt_param_floats(tnt_ctx* c) { return c ? c->n_param : 0; }
extern "C" int64_t tnt_state_floats(tnt_ctx* c) { return c ? c->n_state : 0; }
extern "C" int tnt_num_vars(tnt_ctx* c) { return c ? (int)c->vars.size() : 0; }
extern "C" int64_t tnt_launch_count(tnt_ctx* c) { return c ? c->launches : 0; }

extern "C" int tnt_var_info_get(tnt_ctx* c, int i, tnt_var_info* out) {
  if (!c || !out || i < 0 || i >= (int)c->vars.size()) return fail(c, TNT_ERR_ARG, "tnt_var_info_get: bad index");
  *out = c->vars[i];
  return TNT_OK;
}

extern "C" int tnt_bind_params(tnt_ctx* c, float* w, float* m, float* v, float* g, float* state) {
  if (!c || !w || !state) return fail(c, TNT_ERR_ARG, "tnt_bind_params: w and state are required");
  c->w = w; c->m = m; c->v = v; c->g = g; c->state = state;
  return TNT_OK;
}

extern "C" int tnt_set_adam_step(tnt_ctx* c, int64_t t) { if (!c) return TNT_ERR_ARG; c->adam_t = t; return TNT_OK; }
extern "C" int64_t tnt_get_adam_step(tnt_ctx* c) { return c ? c->adam_t : 0; }
extern "C" int tnt_set_lr(tnt_ctx* c, float lr) { if (!c) return TNT_ERR_ARG; c->cfg.lr = lr; return TNT_OK; }

extern "C" int tnt_buffer(tnt_ctx* c, const char* name, void** ptr, int64_t* bytes) {
  if (!c || !name) return TNT_ERR_ARG;
  auto it = c->bufs.find(name);
  if (it == c->bufs.end()) return fail(c, TNT_ERR_ARG, std::string("tnt_buffer: unknown buffer ") + name);
  if (ptr) *ptr = it->second.first;
  if (bytes) *bytes = it->second.second;
  return TNT_OK;
}

extern "C" int tnt_grad_bucket(tnt_ctx* c, int which, int64_t* offset, int64_t* len) {
  if (!c || !offset || !len) return TNT_ERR_ARG;
  if (which == 0) { *offset = 0; *len = c->dec_off; }
  else if (which == 1) { *offset = c->dec_off; *len = c->n_param - c->dec_off; }
  else if (which == 2) { *offset = c->o_iW; *len = c->n_param - c->o_iW; }          // head: final mid-decode
  else if (which == 3) { *offset = c->dec_off; *len = c->o_iW - c->dec_off; }       // decoder without the head
  else return TNT_ERR_ARG;
  return TNT_OK;
}

// Make `stream` wait for an intermediate result of the step just enqueued with tnt_phase_decode:
//   TNT_EV_HEAD_GRADS: the head gradients (bucket 2) are final -- before the recurrence gradient starts, so their
//                      all-reduce hides completely behind it;
//   TNT_EV_METRICS:    "metrics_raw" and "attn_colsum" are final (right after the embedding gradient, the first
//                      thing the side stream does once the recurrence gradient is done).
extern "C" int tnt_stream_wait_event(tnt_ctx* c, int which, void* stream) {
  if (!c || (which != TNT_EV_HEAD_GRADS && which != TNT_EV_METRICS)) return TNT_ERR_ARG;
  cudaEvent_t ev = which == TNT_EV_HEAD_GRADS ? c->ev_head : c->ev_metrics;
  if (c->side && ev) cudaStreamWaitEvent((cudaStream_t)stream, ev, 0);
  return check_cuda(c, "tnt_stream_wait_event");
}

// ------------------------------------------------------------------------------------------------
// steps
// ------------------------------------------------------------------------------------------------
static int check_batch(tnt_ctx* c, int B, const char* who) {
  if (!c) return TNT_ERR_ARG;
  if (c->fail_code) return c->fail_code;
  if (!c->w || !c->state) return fail(c, TNT_ERR_STATE, std::string(who) + ": parameters not bound");
  if (B < 1 || B > c->cfg.B_max) return fail(c, TNT_ERR_ARG, std::string(who) + ": batch outside [1,B_max]");
  if (c->cfg.n_enc == 2 && (B & 1)) return fail(c, TNT_ERR_ARG, std::string(who) + ": ms2 needs an even batch");
  return TNT_OK;
}

// before dec_forward: pick up an announced prologue (wait for the side stream) or start a fresh operand arena
static void pro_sync(tnt_ctx* c, const int32_t* cap, int B, int training, uint64_t seed, cudaStream_t st) {
  const bool hit = c->pro_valid && c->pro_cap == cap && c->pro_B == B && c->pro_training == training && c->pro_seed == seed;
  if (hit) {
    if (c->pro_pending) cudaStreamWaitEvent(st, c->ev_pro, 0);
  } else {
    if (c->pro_pending) cudaStreamWaitEvent(st, c->ev_pro, 0);   // a stale announcement still owns buffers: let it finish
    c->pro_valid = 0;
    side_join(c, st);
    tc_arena_reset(c);
  }
  c->pro_pending = 0;
}

static void half_of(tnt_ctx* c, int e, int B, int* b_begin, int* Bh) {
  if (c->cfg.n_enc == 2) { *Bh = B / 2; *b_begin = e * (B / 2); }
  else { *Bh = B; *b_begin = 0; }
}

// Announce the step: the encoder-independent head of the decoder (embedding, Zx, weight packing) starts on the side
// stream and overlaps tnt_phase_encode.  Optional: tnt_phase_decode does the same work inline when not announced.
extern "C" int tnt_phase_prologue(tnt_ctx* c, const int32_t* cap, int B, int training, int backward, uint64_t seed,
                                  void* stream) {
  int rc = check_batch(c, B, "tnt_phase_prologue");
  if (rc) return rc;
  if (!cap) return fail(c, TNT_ERR_ARG, "tnt_phase_prologue: null captions");
  cudaStream_t st = (cudaStream_t)stream;
  // no side_join here: a data-parallel caller may still have the previous step's decoder-bucket all-reduce and
  // optimizer part in flight on the side stream; the prologue below is ordered behind them on that stream, and the
  // main stream joins before its first use of a decoder weight (enc_bn_finish)
  if (!c->side) side_join(c, st);
  tc_arena_reset(c);
  if (c->side) {
    cudaEventRecord(c->ev_fork, st);                 // encoder weights / buffers of the previous step are final here
    cudaStreamWaitEvent(c->side, c->ev_fork, 0);
    tc_set_lane(c, 1);
    { TNT_TRACE(c, "prologue: emb_fwd + packs + Zx gemm_pk (side stream)", c->side);
      dec_prologue(c, cap, B, training, backward, seed, c->side); }
    tc_set_lane(c, 0);
    cudaEventRecord(c->ev_pro, c->side);
    c->pro_pending = 1;
  } else {
    dec_prologue(c, cap, B, training, backward, seed, st);
  }
  c->pro_valid = 1; c->pro_cap = cap; c->pro_B = B; c->pro_training = training; c->pro_seed = seed;
  return check_cuda(c, "tnt_phase_prologue");
}

extern "C" int tnt_phase_encode(tnt_ctx* c, const float* betas, int B, int B_global, int training,
                                uint64_t seed, void* stream) {
  int rc = check_batch(c, B, "tnt_phase_encode");
  if (rc) return rc;
  if (!betas) return fail(c, TNT_ERR_ARG, "tnt_phase_encode: null betas");
  cudaStream_t st = (cudaStream_t)stream;
  c->last_B = B; c->last_training = training; c->last_seed = seed;
  for (int e = 0; e < c->cfg.n_enc; ++e) {
    int b0, Bh; half_of(c, e, B, &b0, &Bh);
    enc_forward(c, e, betas, b0, Bh, B, training, seed, st);
  }
  (void)B_global;
  return check_cuda(c, "tnt_phase_encode");
}

extern "C" int tnt_phase_decode(tnt_ctx* c, const int32_t* cap, const int32_t* target, const float* h0,
                                const float* c0, int B, int B_global, int training, int backward,
                                uint64_t seed, void* stream) {
  int rc = check_batch(c, B, "tnt_phase_decode");
  if (rc) return rc;
  if (!cap || !h0 || !c0 || !target) return fail(c, TNT_ERR_ARG, "tnt_phase_decode: null input");
  if (backward && !c->g) return fail(c, TNT_ERR_STATE, "tnt_phase_decode: backward needs g");
  if (B_global < B || B_global % B) return fail(c, TNT_ERR_ARG, "tnt_phase_decode: B_global must be a multiple of B");
  cudaStream_t st = (cudaStream_t)stream;
  const int world = B_global / B;
  for (int e = 0; e < c->cfg.n_enc; ++e) {
    int b0, Bh; half_of(c, e, B, &b0, &Bh);
    enc_bn_finish(c, e, b0, Bh, Bh * world, training, seed, st);
  }
  cudaMemsetAsync(wsT<double>(c, "metrics_raw"), 0, 16 * sizeof(double), st);
  pro_sync(c, cap, B, training, seed, st);
  dec_forward(c, cap, h0, c0, B, training, seed, st);
  head_forward(c, B, training, seed, st, 0);
  if (c->cfg.attn_loss == TNT_ATTN_LOSS_ELEM) cudaMemsetAsync(wsT<double>(c, "metrics_raw") + 8, 0, sizeof(double), st);
  { TNT_TRACE(c, "head_fused_kernel (dense_out + softmax + CE + d logits, logits in TMEM only)", st);
    if (!head_fused_loss(c, target, B, B_global, backward, st)) return c->fail_code; }
  head_metrics(c, 1, B, backward, st);
  if (backward) {
    if (c->cfg.attn_loss == TNT_ATTN_LOSS_BATCH && B_global != B)
      return fail(c, TNT_ERR_ARG, "tnt_phase_decode: the batch-sum attention loss is not built for data-parallel runs "
                                  "(its gradient needs the global column sums before the recurrence gradient starts)");
    head_backward(c, B, seed, st);
    dec_attn_loss_grad(c, B, B_global, st);         // no-op unless the attention loss is part of the differentiated loss
    dec_backward(c, cap, B, B_global, seed, st);
    for (int e = 0; e < c->cfg.n_enc; ++e) {
      int b0, Bh; half_of(c, e, B, &b0, &Bh);
      enc_bwd_stats(c, e, b0, Bh, seed, st);
    }
  }
  return check_cuda(c, "tnt_phase_decode");
}

extern "C" int tnt_phase_encoder_backward(tnt_ctx* c, const float* betas, int B, int B_global, void* stream) {
  int rc = check_batch(c, B, "tnt_phase_encoder_backward");
  if (rc) return rc;
  if (!betas || !c->g) return fail(c, TNT_ERR_ARG, "tnt_phase_encoder_backward: null betas / g");
  cudaStream_t st = (cudaStream_t)stream;
  const int world = B_global / B;
  for (int e = 0; e < c->cfg.n_enc; ++e) {
    int b0, Bh; half_of(c, e, B, &b0, &Bh);
    enc_bwd_finish(c, e, betas, b0, Bh, Bh * world, c->last_seed, st);
  }
  return check_cuda(c, "tnt_phase_encoder_backward");
}

// In-situ timing: enable (clears the records), run steps, then read {name, ms} per recorded kernel group.
extern "C" int tnt_trace(tnt_ctx* c, int enable) {
  if (!c) return TNT_ERR_ARG;
  c->trace_on = enable ? 1 : 0;
  if (enable) c->trace_n = 0;
  return TNT_OK;
}
extern "C" int tnt_trace_read(tnt_ctx* c, char* names, int names_len, float* ms, int max_records) {
  if (!c || !names || !ms) return -1;
  cudaDeviceSynchronize();
  int n = 0, pos = 0;
  for (size_t i = 0; i < c->trace_n && n < max_records; ++i) {
    float t = 0.f;
    if (cudaEventElapsedTime(&t, c->trace[i].e0, c->trace[i].e1) != cudaSuccess) { cudaGetLastError(); continue; }
    int len = (int)strlen(c->trace[i].name);
    if (pos + len + 2 > names_len) break;
    memcpy(names + pos, c->trace[i].name, len); pos += len; names[pos++] = '\n';
    ms[n++] = t;
  }
  names[pos] = 0;
  return n;
}

extern "C" void* tnt_side_stream(tnt_ctx* c) { return c ? (void*)c->side : nullptr; }
extern "C" int tnt_side_join(tnt_ctx* c, void* stream) {
  if (!c) return TNT_ERR_ARG;
  side_join(c, (cudaStream_t)stream);
  return TNT_OK;
}

extern "C" int tnt_phase_optimizer(tnt_ctx* c, void* stream) {
  if (!c || !c->w || !c->m || !c->v || !c->g) return fail(c, TNT_ERR_STATE, "tnt_phase_optimizer: w/m/v/g not bound");
  cudaStream_t st = (cudaStream_t)stream;
  if (c->cfg.agc_clip > 0.0f) {
    // adaptive gradient clipping (agc.py, lc_NIC.py:388) sits between tape.gradient and apply_gradients: unit-wise
    // factors first, then the per-variable clipnorm of the clipped gradients, then Adam -- one pass over all variables
    side_join(c, st);
    opt_norms(c, 1, st);
    opt_l2(c, st);
    opt_agc(c, st);
    opt_adam(c, st);
    return check_cuda(c, "tnt_phase_optimizer");
  }
  if (c->side && c->side_pending) {
    // the encoder variables' gradients are all made on the main stream: update them while the side stream finishes
    // the decoder weight gradients, then join and update the rest.  (Moving this part and the data-parallel exchange
    // to a third stream so that the next step's enc_permute_kernel overlaps them was measured: no gain at N = 8, 10 us
    // slower at N = 1 -- the passes contend for HBM.)
    { TNT_TRACE(c, "opt_norm_kernel (encoder part)", st); opt_norms(c, 1, st, 0, 1); }
    { TNT_TRACE(c, "opt_adam_kernel (encoder part)", st); opt_adam(c, st, 0, 1); }
    {
      // The decoder variables are updated on the side stream, right behind the weight gradients it is still
      // computing: the main stream does not wait for them and goes on with the NEXT step's encoder forward (which
      // reads encoder weights only).  The side stream's next prologue is ordered behind the update; the main stream
      // joins before its first use of a decoder weight (enc_bn_finish) and before metrics are read.
      cudaEventRecord(c->ev_fork, st);                 // the L2 metric sums every variable's norm: after the encoder part
      cudaStreamWaitEvent(c->side, c->ev_fork, 0);
      { TNT_TRACE(c, "opt_norm_kernel (decoder part)", c->side); opt_norms(c, 1, c->side, 1, 0); }
      opt_l2(c, c->side);
      { TNT_TRACE(c, "opt_adam_kernel (decoder part)", c->side); opt_adam(c, c->side, 1, 0); }
      cudaEventRecord(c->ev_join, c->side);
      c->side_pending = 1;
      return check_cuda(c, "tnt_phase_optimizer");
    }
  }
  side_join(c, st);
  { TNT_TRACE(c, "opt_norm_kernel", st); opt_norms(c, 1, st); }
  opt_l2(c, st);
  { TNT_TRACE(c, "opt_adam_kernel", st); opt_adam(c, st); }
  return check_cuda(c, "tnt_phase_optimizer");
}

// Data-parallel form of phase 4, one call per gradient bucket in the order the reduced gradients arrive:
// part 1 = decoder variables, part 0 = encoder variables.  `first` clears the per-variable norm sums and advances the
// Adam iteration; `last` finishes the L2 sum.
extern "C" int tnt_phase_optimizer_part(tnt_ctx* c, int part, int first, int last, void* stream) {
  if (!c || !c->w || !c->m || !c->v || !c->g) return fail(c, TNT_ERR_STATE, "tnt_phase_optimizer_part: w/m/v/g not bound");
  if (part != 0 && part != 1) return fail(c, TNT_ERR_ARG, "tnt_phase_optimizer_part: part must be 0 or 1");
  cudaStream_t st = (cudaStream_t)stream;
  const bool on_side = c->side && st == c->side;      // deferred: runs behind the bucket's all-reduce on the side stream
  if (c->cfg.agc_clip > 0.0f) {
    // adaptive clipping needs every variable's unit norms first: the whole update runs with the LAST part
    if (!last) return TNT_OK;
    if (!on_side) side_join(c, st);
    opt_norms(c, 1, st);
    opt_l2(c, st);
    opt_agc(c, st);
    opt_adam(c, st);
    if (on_side) { cudaEventRecord(c->ev_join, c->side); c->side_pending = 1; }
    return check_cuda(c, "tnt_phase_optimizer_part");
  }
  if (!on_side && part == 1) side_join(c, st);        // encoder gradients are all made on the main stream
  opt_norms(c, 1, st, part, first);
  opt_adam(c, st, part, first);
  if (last) opt_l2(c, st);
  if (on_side) {
    cudaEventRecord(c->ev_join, c->side);
    c->side_pending = 1;
  }
  return check_cuda(c, "tnt_phase_optimizer_part");
}

extern "C" int tnt_eval_l2(tnt_ctx* c, void* stream) {
  if (!c || !c->w) return fail(c, TNT_ERR_STATE, "tnt_eval_l2: parameters not bound");
  side_join(c, (cudaStream_t)stream);
  opt_norms(c, 0, (cudaStream_t)stream);
  opt_l2(c, (cudaStream_t)stream);
  return check_cuda(c, "tnt_eval_l2");
}

// The general step: forward (+ loss / metrics) (+ tape gradient) (+ clip and Adam).  After tnt_dp_init every exchange of
// the data-parallel protocol is made here and in the phases: the host makes one call per step.
extern "C" int tnt_step(tnt_ctx* c, const float* betas, const int32_t* cap, const int32_t* target, const float* h0,
                        const float* c0, int B, int training, int backward, int apply, uint64_t seed, void* stream) {
  if (!c) return TNT_ERR_ARG;
  if (!target) return fail(c, TNT_ERR_ARG, "tnt_step: null target");
  cudaStream_t st = (cudaStream_t)stream;
  const int world = dp_world(c), Bg = B * world;
  int rc = (cap && backward) ? tnt_phase_prologue(c, cap, B, training, backward, seed, stream) : TNT_OK;
  if (rc) return rc;
  rc = tnt_phase_encode(c, betas, B, Bg, training, seed, stream);
  if (rc) return rc;
  if (world > 1 && training) {
    TNT_TRACE(c, "dp: all-reduce of the BatchNorm forward sums", st);
    dp_allreduce_small(c, wsT<double>(c, "bn_fwd_stats"), 64 * c->cfg.n_enc, st);
  }
  rc = tnt_phase_decode(c, cap, target, h0, c0, B, Bg, training, backward, seed, stream);
  if (rc) return rc;
  if (backward) {
    if (world > 1) {
      TNT_TRACE(c, "dp: all-reduce of the BatchNorm backward sums", st);
      dp_allreduce_small(c, wsT<double>(c, "bn_bwd_stats"), 64 * c->cfg.n_enc, st);
    }
    c->dp_push_on = world > 1 && dp_push_ready(c);   // the gradient kernels store into the owning ranks' staging
    c->dp_n_pushed = 0;
    rc = tnt_phase_encoder_backward(c, betas, B, Bg, stream);
    c->dp_push_on = 0;
    if (rc) return rc;
    if (world > 1) {
      // straight into the encoder-part optimizer when that is what runs next (tnt_phase_optimizer's split branch)
      const int direct = apply && c->cfg.agc_clip <= 0.0f && c->side && c->side_pending;
      TNT_TRACE(c, "dp: encoder gradient exchange (main stream)", st);
      dp_grad_exchange(c, c->g, c->dec_off, c->dp_pushed, c->dp_n_pushed, direct, st);
    }
    if (apply) {
      rc = tnt_phase_optimizer(c, stream);
      c->g_enc_src = nullptr;
      return rc;
    }
  }
  return tnt_eval_l2(c, stream);       // joins the side stream; L2 term only
}

extern "C" int tnt_train_step(tnt_ctx* c, const float* betas, const int32_t* cap, const int32_t* target,
                              const float* h0, const float* c0, int B, uint64_t seed, void* stream) {
  return tnt_step(c, betas, cap, target, h0, c0, B, 1, 1, 1, seed, stream);
}

extern "C" int tnt_eval_step(tnt_ctx* c, const float* betas, const int32_t* cap, const int32_t* target,
                             const float* h0, const float* c0, int B, int training, uint64_t seed, void* stream) {
  return tnt_step(c, betas, cap, target, h0, c0, B, training, 0, 0, seed, stream);
}

extern "C" int tnt_forward(tnt_ctx* c, const float* betas, const int32_t* cap, const float* h0, const float* c0,
                           int B, int training, uint64_t seed, float* probs_out, float* attn_out, void* stream) {
  int rc = tnt_phase_encode(c, betas, B, B, training, seed, stream);
  if (rc) return rc;
  if (!cap || !h0 || !c0) return fail(c, TNT_ERR_ARG, "tnt_forward: null input");
  cudaStream_t st = (cudaStream_t)stream;
  for (int e = 0; e < c->cfg.n_enc; ++e) {
    int b0, Bh; half_of(c, e, B, &b0, &Bh);
    enc_bn_finish(c, e, b0, Bh, Bh, training, seed, st);
  }
  pro_sync(c, cap, B, training, seed, st);
  dec_forward(c, cap, h0, c0, B, training, seed, st);
  head_forward(c, B, training, seed, st);
  if (probs_out) head_loss(c, nullptr, B, B, 0, probs_out, st);
  if (attn_out)
    cudaMemcpyAsync(attn_out, wsT<float>(c, "alpha"), (size_t)c->cfg.T * B * c->cfg.R * sizeof(float),
                    cudaMemcpyDeviceToDevice, st);
  return check_cuda(c, "tnt_forward");
}

static int decode_common(tnt_ctx* c, const char* who, const float* betas, const float* h0, const float* c0,
                         const int32_t* start, int B, int max_len, int enc, const void* out, cudaStream_t st) {
  if (!c) return TNT_ERR_ARG;
  if (c->fail_code) return c->fail_code;
  if (!c->w || !c->state) return fail(c, TNT_ERR_STATE, std::string(who) + ": parameters not bound");
  if (B < 1 || B > c->cfg.B_max) return fail(c, TNT_ERR_ARG, std::string(who) + ": batch outside [1,B_max]");
  if (!betas || !h0 || !c0 || !start || !out || max_len < 1 || enc < 0 || enc >= c->cfg.n_enc)
    return fail(c, TNT_ERR_ARG, std::string(who) + ": bad argument");
  // features = dense_in(x, training=False): BN moving statistics, no dropout   (lc_NIC.py:593-594)
  enc_forward(c, enc, betas, 0, B, B, 0, 0, st);
  enc_bn_finish(c, enc, 0, B, B, 0, 0, st);
  tc_arena_reset(c);
  return TNT_OK;
}

extern "C" int tnt_greedy(tnt_ctx* c, const float* betas, const float* h0, const float* c0, const int32_t* start,
                          int B, int max_len, int enc, int64_t* out_ids, float* out_probs, float* out_attn,
                          float* out_attn_hidden, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  int rc = decode_common(c, "tnt_greedy", betas, h0, c0, start, B, max_len, enc, out_ids, st);
  if (rc) return rc;
  dec_greedy(c, h0, c0, start, B, max_len, out_ids, out_probs, out_attn, out_attn_hidden, st);
  return check_cuda(c, "tnt_greedy");
}

extern "C" int tnt_sample(tnt_ctx* c, const float* betas, const float* h0, const float* c0, const int32_t* start,
                          int B, int max_len, int enc, int top_k, float top_p, uint64_t seed, int64_t* out_ids,
                          float* out_probs, float* out_attn, float* out_attn_hidden, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (c && c->cfg.V > 8192) return fail(c, TNT_ERR_ARG, "tnt_sample: vocabulary larger than 8192");
  int rc = decode_common(c, "tnt_sample", betas, h0, c0, start, B, max_len, enc, out_ids, st);
  if (rc) return rc;
  dec_greedy(c, h0, c0, start, B, max_len, out_ids, out_probs, out_attn, out_attn_hidden, st, 1, top_k, top_p, seed);
  return check_cuda(c, "tnt_sample");
}

extern "C" int tnt_beam_search(tnt_ctx* c, const float* betas, const float* h0, const float* c0, const int32_t* start,
                               int B, int max_len, int enc, int beam, int end_id, int64_t* out_ids, float* out_scores,
                               void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (c && (beam < 1 || beam > 8 || (int64_t)B * beam > c->cfg.B_max || !out_scores))
    return fail(c, TNT_ERR_ARG, "tnt_beam_search: need 1 <= beam <= 8, B * beam <= B_max and an out_scores buffer");
  int rc = decode_common(c, "tnt_beam_search", betas, h0, c0, start, B, max_len, enc, out_ids, st);
  if (rc) return rc;
  rc = dec_beam_search(c, h0, c0, start, B, max_len, beam, end_id, out_ids, out_scores, st);
  if (rc) return fail(c, rc, "tnt_beam_search: workspace allocation failed");
  return check_cuda(c, "tnt_beam_search");
}

extern "C" int tnt_check(tnt_ctx* c, void* stream) {
  if (!c) return TNT_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  side_join(c, st);
  if (cudaStreamSynchronize(st) != cudaSuccess) { check_cuda(c, "tnt_check"); return TNT_ERR_CUDA; }
  if (pd_read_abort(c, st)) ctx_fail(c, TNT_ERR_CUDA, "persistent decoder: a grid barrier timed out (results invalid, weights not updated)");
  return check_cuda(c, "tnt_check");
}

extern "C" const char* tnt_kernel_variants(tnt_ctx* c) {
  if (!c) return "";
  c->variants_str = std::string("enc_fwd=") + c->variant_enc_fwd + ";enc_bwd=" + c->variant_enc_bwd +
                    ";recurrence=" + c->variant_rec + ";head=" + c->variant_head +
                    ";gemm=gemm_pk_kernel(tcgen05," + (c->cfg.precision == TNT_FP32 ? "bf16x6" : c->cfg.precision == TNT_BF16X3 ? "bf16x3" : "bf16") + ")" +
                    ";dp_small_allreduce=" + dp_small_kind(c) + ";dp_encoder_gradients=" + dp_grad_kind(c);
  return c->variants_str.c_str();
}

extern "C" int tnt_set_attn_loss(tnt_ctx* c, int mode) {
  if (!c || mode < 0 || mode > 2 || (mode && c->cfg.n_enc != 1)) return fail(c, TNT_ERR_ARG, "tnt_set_attn_loss: bad mode");
  c->cfg.attn_loss = mode;
  return TNT_OK;
}

extern "C" int tnt_sam_perturb(tnt_ctx* c, float rho, float eps, void* stream) {
  if (!c || !c->w || !c->g) return fail(c, TNT_ERR_STATE, "tnt_sam_perturb: w/g not bound");
  int rc = sam_perturb(c, rho, eps, (cudaStream_t)stream);
  if (rc) return fail(c, rc, "tnt_sam_perturb: cannot allocate the perturbation buffer");
  return check_cuda(c, "tnt_sam_perturb");
}
extern "C" int tnt_sam_restore(tnt_ctx* c, void* stream) {
  if (!c || !c->w) return fail(c, TNT_ERR_STATE, "tnt_sam_restore: w not bound");
  int rc = sam_restore(c, (cudaStream_t)stream);
  if (rc) return fail(c, rc, "tnt_sam_restore: no perturbation to undo");
  return check_cuda(c, "tnt_sam_restore");
}

extern "C" int tnt_read_metrics(tnt_ctx* c, tnt_metrics* out, void* stream) {
  if (!c || !out) return TNT_ERR_ARG;
  double h[16];
  cudaStream_t st = (cudaStream_t)stream;
  side_join(c, st);
  if (cudaMemcpyAsync(h, wsT<double>(c, "metrics_raw"), sizeof(h), cudaMemcpyDeviceToHost, st) != cudaSuccess ||
      cudaStreamSynchronize(st) != cudaSuccess)
    return check_cuda(c, "tnt_read_metrics") ? TNT_ERR_CUDA : fail(c, TNT_ERR_CUDA, "tnt_read_metrics failed");
  if (pd_read_abort(c, st)) {
    ctx_fail(c, TNT_ERR_CUDA, "persistent decoder: a grid barrier timed out (results invalid, weights not updated)");
    return TNT_ERR_CUDA;
  }
  // data parallel: the sums are global (all-reduced inside the step), so are the row counts
  int B = c->last_B * dp_world(c), halves = c->cfg.n_enc == 2 ? 2 : 1;
  for (int i = 0; i < 2; ++i) {
    out->ce_sum[i] = h[0 + i]; out->correct[i] = h[2 + i]; out->attn_mse[i] = h[4 + i];
    out->rows[i] = i < halves ? (double)(B / halves) * c->cfg.T : 0.0;
  }
  out->l2 = h[6]; out->emb_sqnorm = h[7];
  out->attn_elem_mse = c->cfg.attn_loss == TNT_ATTN_LOSS_ELEM ? h[8] / ((double)c->cfg.T * B * c->cfg.R) : 0.0;
  return TNT_OK;
}

extern "C" int tnt_onehot_to_ids(tnt_ctx* c, const float* onehot, int64_t rows, int32_t* ids, void* stream) {
  if (!c || !onehot || !ids || rows < 0) return fail(c, TNT_ERR_ARG, "tnt_onehot_to_ids: bad argument");
  launch_onehot_to_ids(onehot, rows, c->cfg.V, ids, (cudaStream_t)stream);
  c->launches += 1;
  return check_cuda(c, "tnt_onehot_to_ids");
}
