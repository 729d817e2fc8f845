// Internal declarations shared by the translation units of libtnt.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include <map>
#include "../../include/tnt.h"
#include "common.cuh"

// One CTA's share of the parcel encoder: a k-range of one parcel's index list.
struct EncWork {
  int32_t r;       // parcel
  int32_t kbeg;    // first position inside the parcel's index list
  int32_t kend;    // one past the last position
  int32_t slot;    // which partial-sum slot of the parcel this CTA writes (0 .. nslots-1)
};

struct EncTables {
  int32_t* idx = nullptr;        // device: concatenated vertex indices of this encoder
  int64_t* off = nullptr;        // device: R+1 offsets into idx (== row offsets into the packed kernel)
  EncWork* work = nullptr;       // device: CTA work list
  int32_t* slot_first = nullptr; // device: R+1 -- partial-sum slots of parcel r are [slot_first[r], slot_first[r+1])
  int64_t* inv_off = nullptr;    // device: N+1 -- positions of vertex v in the concatenated index list: inv_pos[inv_off[v] ..)
  int32_t* inv_pos = nullptr;    // device: n_lab
  int n_work = 0, n_slots = 0;
  int64_t n_lab = 0;             // sum of parcel sizes
  // offsets (floats) into the flat parameter buffer / state buffer
  int64_t w_off = 0, b_off = 0, gamma_off = 0, beta_off = 0, mean_off = 0, var_off = 0;
  // encoder_tc.cu: tensor maps (CUtensorMap, opaque here) of the parcel-ordered batch copy and of dYpre, and the bases they
  // were encoded for
  alignas(64) unsigned char tmap_x[128] = {0};
  alignas(64) unsigned char tmap_dy[128] = {0};
  const void* tmap_x_base = nullptr;
  const void* tmap_dy_base = nullptr;
};

struct Seg {            // one Keras variable for the optimizer (clip norm is per variable)
  int64_t off, len;
  float l2;
  int32_t is_emb;
  int32_t cols;         // 2-D variable (rows, cols): cols; 1-D: 0
  int32_t unit0;        // first adaptive-clipping unit of the variable (one per column, or one for a 1-D variable)
};
struct AgcUnit { int32_t seg; int32_t pad; int64_t base, rows, stride; };   // agc.py unitwise_norm: what one norm covers
struct OptChunk { int32_t seg; int32_t len; int64_t off; };

struct TraceRec { const char* name; cudaEvent_t e0, e1; };

// ---- tensor-core GEMM state (gemm_tc.cu): owned by the context, never shared between contexts ----
struct TcOperand {        // an fp32 matrix split into bf16 planes in the tensor core's K-major block layout
  const uint8_t* planes;
  int64_t plane_stride;   // bytes between planes
  int mn_blocks;          // blocks of 128 along MN
  int k_tiles;            // blocks of 64 along K
};
struct TcSplitWs { float* part; unsigned* count; };
// point-wise tail fused into the GEMM epilogue (gemm_tc.cu: gemm_tc_epi)
enum { TC_EPI_NONE = 0, TC_EPI_INTER_FWD = 1 };
struct TcEpi {
  int mode = TC_EPI_NONE;
  float* C2 = nullptr;            // second fp32 output (forward: activated + dropped-out tensor)
  DropSite drop;                  // dropout site of the layer's output; element index ((b*T + t)*N + n), row = t*B + b
  int B = 1, T = 1;
  uint8_t* pk = nullptr; int64_t pk_stride = 0; int pk_mnb = 0, pk_kt = 0;   // packed A operand of the result
};
struct TcCacheEntry { const float* X; int rows, cols, ld, transpose, ns; TcOperand op; };
struct TcState {
  uint8_t* arena = nullptr;            // packed operands of one phase: lane 0 allocates upwards, lane 1 downwards
  size_t cap = 0, off = 0, top = 0, want = (size_t)256 << 20;
  int lane = 0;                        // 0: main compute stream, 1: side stream (own arena end + split-K workspace)
  TcSplitWs ws[2] = {{nullptr, nullptr}, {nullptr, nullptr}};
  std::vector<TcCacheEntry> cache;     // operands packed since the last reset
  int attr_done = 0;
};

struct tnt_ctx {
  tnt_config cfg;
  std::string err, variants_str;
  int device = 0;
  int sm_count = 148;
  int64_t launches = 0;
  // parameters (caller-owned)
  float *w = nullptr, *m = nullptr, *v = nullptr, *g = nullptr, *state = nullptr;
  int64_t n_param = 0, n_state = 0;
  int64_t adam_t = 0;
  std::vector<tnt_var_info> vars;
  std::vector<Seg> segs;
  EncTables enc[2];
  // offsets of decoder variables in the flat buffer
  int64_t o_W1, o_b1, o_W2, o_b2, o_V, o_bV, o_emb, o_lW, o_lU, o_lb, o_iW, o_ib, o_oW, o_ob;
  int64_t dec_off = 0;           // first float of the decoder part (everything after the encoders)
  // optimizer tables (device)
  Seg* d_segs = nullptr; OptChunk* d_chunks = nullptr; int n_chunks = 0; int n_segs = 0; int n_chunks_enc = 0; int n_segs_enc = 0;
  double* seg_gsq = nullptr; double* seg_wsq = nullptr;
  AgcUnit* d_units = nullptr; int n_units = 0;
  // workspaces
  std::map<std::string, std::pair<void*, int64_t>> bufs;
  int last_B = 0, last_training = 0;
  cudaStream_t side = nullptr;          // low-priority stream for work off the critical path of the step
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  int side_pending = 0;

  cudaEvent_t ev_pro = nullptr, ev_head = nullptr, ev_w1 = nullptr, ev_metrics = nullptr;          // the announced step's prologue (dec_prologue on the side stream)
  int pro_valid = 0, pro_pending = 0, pro_B = 0, pro_training = 0;
  const int32_t* pro_cap = nullptr;
  uint64_t pro_seed = 0;
  // in-situ kernel timing (tnt_trace): CUDA events on the launching stream around named kernel groups
  int trace_on = 0;
  size_t trace_n = 0;
  std::vector<TraceRec> trace;
  const float* xT_src[2] = {nullptr, nullptr};   // parcel-ordered copy of the batch (encoder.cu): what it was made from
  int xT_b0[2] = {0, 0}, xT_Bh[2] = {0, 0};
  int et_attr = 0;                   // encoder_tc.cu: function attributes set
  long long* et_trace = nullptr;     // encoder_tc.cu: diagnostic time line (device buffer), off in production
  int hf_attr = 0;                   // head_fused.cu: function attributes set (bit per plane count)
  int pd_ready = 0;                  // persistent decoder (decoder_persist.cu): function attributes set
  int l2_frozen = 0;                 // the next opt_l2 keeps metrics_raw[6] (sharpness-aware step, variants.cu)
  struct DpState* dp = nullptr;      // data parallelism inside the library (dp.cu); nullptr: single GPU
  const float* g_enc_src = nullptr;  // encoder-part optimizer reads its gradients here instead of g (dp.cu: the exchange's
                                     // output buffer, saves a 2 x 37 MB copy); nullptr: g
  int dp_push_on = 0;                // tnt_step: the encoder gradient kernels push into the owners' staging (dp.cu)
  int64_t dp_pushed[8]; int dp_n_pushed = 0;   // ranges pushed by the producing kernels in this step
  int fail_code = 0;                 // sticky: a launch the call depends on failed; every later call returns it
  uint64_t last_seed = 0;
  TcState tc;
  // which kernel variant the last call used (tnt_kernel_variants; bench.py records it)
  const char *variant_enc_fwd = "", *variant_enc_bwd = "", *variant_rec = "", *variant_head = "";
};

// RAII: records an event pair on `st` around the enclosed launches when tracing is on
struct TraceScope {
  tnt_ctx* c; cudaStream_t st; long idx;
  TraceScope(tnt_ctx* c_, const char* name, cudaStream_t st_) : c(c_), st(st_), idx(-1) {
    if (!c->trace_on) return;
    if (c->trace_n == c->trace.size()) {
      TraceRec r; r.name = name;
      if (cudaEventCreate(&r.e0) != cudaSuccess || cudaEventCreate(&r.e1) != cudaSuccess) { cudaGetLastError(); return; }
      c->trace.push_back(r);
    }
    idx = (long)c->trace_n++;
    c->trace[idx].name = name;
    cudaEventRecord(c->trace[idx].e0, st);
  }
  ~TraceScope() { if (idx >= 0) cudaEventRecord(c->trace[idx].e1, st); }
};
#define TNT_CAT2(a, b) a##b
#define TNT_CAT(a, b) TNT_CAT2(a, b)
#define TNT_TRACE(c, name, st) TraceScope TNT_CAT(trace_scope_, __LINE__)(c, name, st)

static inline float att_slope_of(const tnt_config& f) {
  return (f.variant & TNT_VARIANT_PLAIN_ATTENTION) ? 1.0f : 0.2f;
}
// marks the context failed (sticky; there is no alternative path to fall back to) and records the message
inline void ctx_fail(tnt_ctx* c, int code, const std::string& msg) { if (!c->fail_code) { c->fail_code = code; c->err = msg; } }
void* ws(tnt_ctx* c, const char* name);   // named workspace (throws nothing; returns nullptr if absent)
template <typename T> inline T* wsT(tnt_ctx* c, const char* name) { return reinterpret_cast<T*>(ws(c, name)); }

// ---- encoder.cu -------------------------------------------------------------------------------
void enc_forward(tnt_ctx* c, int e, const float* betas, int b_begin, int Bh, int B_total, int training,
                 uint64_t seed, cudaStream_t st);                       // -> Y, bn_fwd_stats
void enc_bn_finish(tnt_ctx* c, int e, int b_begin, int Bh, int Bh_global, int training, uint64_t seed,
                   cudaStream_t st);                                    // -> bn coef, feats, Pf
void enc_bwd_stats(tnt_ctx* c, int e, int b_begin, int Bh, uint64_t seed, cudaStream_t st);
void enc_bwd_finish(tnt_ctx* c, int e, const float* betas, int b_begin, int Bh, int Bh_global, uint64_t seed,
                    cudaStream_t st);
// ---- encoder_tc.cu ----------------------------------------------------------------------------
bool enc_fwd_tc(tnt_ctx* c, int e, const float* xP, int Bh, cudaStream_t st);                    // -> enc_part
bool enc_bwd_dw_tc(tnt_ctx* c, int e, const float* xP, int b_begin, int Bh, cudaStream_t st);   // -> g (packed kernels)
// ---- decoder.cu -------------------------------------------------------------------------------
void dec_prologue(tnt_ctx* c, const int32_t* cap, int B, int training, int backward, uint64_t seed, cudaStream_t st);
void dec_attn_colsum(tnt_ctx* c, int B, cudaStream_t st);
void dec_forward(tnt_ctx* c, const int32_t* cap, const float* h0, const float* c0, int B, int training,
                 uint64_t seed, cudaStream_t st);
void dec_backward(tnt_ctx* c, const int32_t* cap, int B, int B_global, uint64_t seed, cudaStream_t st);
// sample == 0: argmax (greedy_predict_attention); sample == 1: random choice (tnt_sample in include/tnt.h)
void dec_greedy(tnt_ctx* c, const float* h0, const float* c0, const int32_t* start, int B, int max_len,
                int64_t* out_ids, float* out_probs, float* out_attn, float* out_attn_hidden, cudaStream_t st,
                int sample = 0, int top_k = 0, float top_p = 0.0f, uint64_t seed = 0);
// ---- decoder_persist.cu -----------------------------------------------------------------------
bool pd_shape_ok(const tnt_config& f);
int64_t pd_ws_bytes(const tnt_config& f, const char* what);
bool pd_enabled(tnt_ctx* c);
bool pd_forward(tnt_ctx* c, int B, int T, int training, uint64_t seed, float* gates, float* hout, float* alpha,
                float* s_out, cudaStream_t st);
bool pd_backward(tnt_ctx* c, int B, int T, uint64_t seed, cudaStream_t st);
int pd_read_abort(tnt_ctx* c, cudaStream_t st);
unsigned* pd_abort_flag(tnt_ctx* c);   // device flag set by a persistent kernel whose barrier timed out; nullptr if none
// ---- head.cu ----------------------------------------------------------------------------------
void head_forward(tnt_ctx* c, int B, int training, uint64_t seed, cudaStream_t st, int want_logits = 1);   // hout -> inter (-> logits)
// head_fused.cu: dense_out + softmax + loss (+ d logits in packed operand form) without the logits ever leaving the SM
bool head_fused_loss(tnt_ctx* c, const int32_t* target, int B, int B_global, int backward, cudaStream_t st);
void head_loss(tnt_ctx* c, const int32_t* target, int B, int B_global, int backward, float* probs_out,
               cudaStream_t st);                                                        // logits -> probabilities (tnt_forward)
void head_metrics(tnt_ctx* c, int have_target, int B, int backward, cudaStream_t st);   // row losses, alpha -> metrics_raw (side stream)
void head_backward(tnt_ctx* c, int B, uint64_t seed, cudaStream_t st);                  // dlogits -> grads, dhout
void side_join(tnt_ctx* c, cudaStream_t st);     // main stream waits for the side-stream work of this phase
void launch_onehot_to_ids(const float* onehot, int64_t rows, int V, int32_t* ids, cudaStream_t st);
// ---- optimizer.cu -----------------------------------------------------------------------------
void opt_norms(tnt_ctx* c, int with_grad, cudaStream_t st, int part = -1, int first = 1);   // per-variable ||g + 2 l w||^2 and l*||w||^2
void opt_l2(tnt_ctx* c, cudaStream_t st);
void opt_launch_norm(const OptChunk* chunks, int n_chunks, const Seg* segs, const float* w, const float* g,
                     double* seg_gsq, double* seg_wsq, cudaStream_t st);
void opt_launch_adam(const OptChunk* chunks, int n_chunks, const Seg* segs, float* w, float* m, float* v,
                     const float* g, const double* seg_gsq, float clipnorm, float alpha, float beta1, float beta2,
                     float eps, cudaStream_t st);
void opt_adam(tnt_ctx* c, cudaStream_t st, int part = -1, int first = 1);
// ---- variants.cu ------------------------------------------------------------------------------
void dec_attn_loss_grad(tnt_ctx* c, int B, int B_global, cudaStream_t st);   // alpha -> attn_e, attn_Se
void opt_agc(tnt_ctx* c, cudaStream_t st);            // unit-wise norms -> agc_factor, clipped per-variable norms
int sam_perturb(tnt_ctx* c, float rho, float eps, cudaStream_t st);
int sam_restore(tnt_ctx* c, cudaStream_t st);
// ---- dp.cu ------------------------------------------------------------------------------------
int dp_world(tnt_ctx* c);
int dp_rank(tnt_ctx* c);
const char* dp_small_kind(tnt_ctx* c);
const char* dp_grad_kind(tnt_ctx* c);
void dp_allreduce_small(tnt_ctx* c, double* buf, int n, cudaStream_t st);     // <= 128 doubles, critical path
void dp_allreduce_f32_small(tnt_ctx* c, float* buf, int n, cudaStream_t st); // <= 8192 floats, side stream only
void dp_allreduce(tnt_ctx* c, void* buf, int64_t n, int is_double, int lane, cudaStream_t st);   // NCCL; lane 1 = side stream
void dp_destroy(tnt_ctx* c);
// Encoder-gradient exchange over peer memory (dp.cu), fused with the producing kernel: reduce-scatter by PUSH -- the
// kernel that computes a gradient row stores it straight into the owning rank's staging buffer over NVLink -- then the
// owner sums its slice in rank order and broadcasts it (all-gather by push).  Flat partition of g[0, L) into `world`
// slices of SL floats (a multiple of 4: a float4 never straddles two owners).
#define TNT_DP_MAX_WORLD 8
struct DpPush {
  float* in[TNT_DP_MAX_WORLD];   // in[q]: rank q's staging [world][SL] (partials of q's slice, one row per source rank)
  int64_t SL;                    // floats per slice
  int32_t rank, world;
};
bool dp_push_ready(tnt_ctx* c);                       // peer-memory gradient exchange available (after tnt_dp_init)
DpPush dp_push_params(tnt_ctx* c);
// g[0, L) := sum over ranks.  `pushed` = ranges [beg, end) (floats, multiples of 4) the producing kernels have already
// pushed; everything else of [0, L) is pushed here from g.  Replaces dp_allreduce(c, c->g, L, 0, 0, st).
// direct != 0: the reduced gradient is NOT copied back into g; c->g_enc_src points at it for the encoder-part optimizer.
void dp_grad_exchange(tnt_ctx* c, float* g, int64_t L, const int64_t* pushed, int n_pushed, int direct, cudaStream_t st);
void dec_reduce_metrics(tnt_ctx* c, cudaStream_t st);   // head.cu: all-reduce metric sums / attention column sums (no-op on one GPU)
// ---- sampling.cu ------------------------------------------------------------------------------
void dec_sample_words(tnt_ctx* c, int B, int step, int top_k, float top_p, uint64_t seed, const float* logits,
                      int32_t* word, int64_t* out_ids, int out_stride, float* probs, int64_t probs_stride, cudaStream_t st);
int dec_beam_search(tnt_ctx* c, const float* h0, const float* c0, const int32_t* start, int B, int max_len, int beam,
                    int end_id, int64_t* out_ids, float* out_scores, cudaStream_t st);
void dec_step(tnt_ctx* c, int B, int t, int training, uint64_t seed, const float* hprev, const float* cprev,
              float* Zx_t, float* alpha_t, float* qact_t, float* xctx_t, float* gates_t, float* cnew,
              float* hnew, float* hout_t, float* s_out, cudaStream_t st);
void head_decode_step(tnt_ctx* c, int B, const float* h, float* inter, float* logits, cudaStream_t st);
// ---- tensor-core GEMM on packed operands (gemm_tc.cu) -------------------------------------------
void tc_arena_reserve(tnt_ctx* c, size_t bytes);
void tc_arena_reset(tnt_ctx* c);
void tc_set_lane(tnt_ctx* c, int lane);    // 1: following packs / GEMMs are enqueued on the side stream
void tc_destroy(tnt_ctx* c);
size_t tc_packed_bytes(int64_t mn, int64_t k, int ns);
TcOperand tc_pack_cached(tnt_ctx* c, const float* X, int rows, int cols, int ld, bool transpose, cudaStream_t st);
TcOperand tc_pack(tnt_ctx* c, const float* X, int rows, int cols, int ld, bool transpose, cudaStream_t st);
TcOperand tc_reserve_operand(tnt_ctx* c, const float* key, int rows, int cols, int ld, bool transpose);
int tc_planes(tnt_ctx* c);     // bf16 planes per operand under the context's precision (3 / 2 / 1)
struct TcPackReq { const float* X; int rows, cols, ld; bool transpose; };   // arguments of tc_pack_cached
void tc_prepack(tnt_ctx* c, const TcPackReq* reqs, int n, cudaStream_t st);   // several operands, one launch
void gemm_packed(tnt_ctx* c, const TcOperand& A, const TcOperand& B, int M, int N, float* C, int ldc, const float* bias,
                 int accumulate, cudaStream_t st, const TcEpi* epi = nullptr);
void gemm_tc_epi(tnt_ctx* c, int cacheA, bool tb, int M, int N, int K, const float* A, int lda, const float* W, int ldw,
                 float* C, int ldc, const float* bias, int mode, float* C2, const float* pre, const DropSite& drop, int Bsz,
                 int T, const float* pk_key, cudaStream_t st);
// C[M,N] = opA(A) opB(B) (+ bias) (+ C) on row-major fp32 operands (TA: A stored [K,M]; TB: B stored [N,K]): packs both
// operands (cacheA / cacheB: the operand is not rewritten before the next tc_arena_reset(), its packed form is kept)
// and runs gemm_packed.  accumulate: 0 store, 1 read-add-store, 2 atomicAdd onto C.
void gemm_tc(tnt_ctx* c, int cacheA, int cacheB, bool ta, bool tb, int M, int N, int K, const float* A, int lda,
             const float* B, int ldb, float* C, int ldc, const float* bias, int accumulate, cudaStream_t st);
// ---- column sums (gemm.cu) --------------------------------------------------------------------
void colsum(tnt_ctx* c, int M, int N, const float* X, int ldx, float* out, cudaStream_t st);
struct ColsumReq { const float* X; float* out; int M, N, ldx; };
void colsum_multi(tnt_ctx* c, const ColsumReq* reqs, int n, cudaStream_t st);   // <= 4 column sums, one launch
