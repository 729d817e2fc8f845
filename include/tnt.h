/*
 * tnt.h -- C ABI of libtnt.so, the sm_100a CUDA engine behind the Think-and-Tell hot path.
 *
 * The reference (seang123/Masters-Thesis) has no FFI/plugin interface for this path: the boundary
 * is the Python surface of `class NIC(tf.keras.Model)` (AttemptFour/Model/lc_NIC.py:36-638).  The
 * Python mirror of that class (masters-thesis_b200/model.py) binds the entry points below through
 * ctypes; each entry point names the reference method whose arithmetic it replaces.
 *
 * Conventions
 *   - return 0 on success, negative tnt_status on error; tnt_last_error(ctx) gives the message.
 *     Nothing aborts or throws across this boundary.
 *   - all pointers are DEVICE pointers unless the name ends in _host; the caller owns every buffer
 *     it passes; the library owns only the workspaces inside tnt_ctx.
 *   - every call is asynchronous on the cudaStream_t passed as `stream` (a `void*` here so the
 *     header needs no CUDA include) and not re-entrant per ctx (one ctx per GPU, one host thread).
 *   - storage and element-wise arithmetic are fp32; GEMM operand precision per tnt_precision; ids are int32.
 *   - recurrent buffers are time-major; caller-visible tensors keep the reference layouts.
 */
#ifndef TNT_H
#define TNT_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct tnt_ctx tnt_ctx;

enum tnt_status {
  TNT_OK = 0,
  TNT_ERR_ARG = -1,       /* bad argument / unsupported dimension */
  TNT_ERR_CUDA = -2,      /* a CUDA runtime call failed */
  TNT_ERR_STATE = -3,     /* call order violated (e.g. params not bound) */
  TNT_ERR_NO_DEVICE = -4  /* no sm_100 device: there is NO CPU fallback */
};

/* Operand precision of the tensor-core GEMMs (accumulation is fp32 in TMEM in every mode; the encoder, the recurrence
 * and all element-wise arithmetic stay fp32).  Stated parity bounds against the fp32 reference arithmetic:
 *   TNT_FP32    three-way bf16 split of both operands (six MMAs, 24 mantissa bits: fp32-equivalent products);
 *               loss and gradients within 1e-3 relative (measured ~1e-6).  The default; what bench.py times.
 *   TNT_BF16X3  two-way split (three MMAs, 16 mantissa bits): loss within 1e-3, gradients within 5e-3 relative.
 *   TNT_BF16    plain bf16 operands (one MMA, 8 mantissa bits): loss within 1e-2, gradients within 1e-1 relative. */
enum tnt_precision { TNT_FP32 = 0, TNT_BF16X3 = 1, TNT_BF16 = 2 };
/* tnt_config.attn_loss: is the "attention loss" part of the loss that is differentiated?
 *   TNT_ATTN_LOSS_OFF    reported only (lc_NIC.py:365-367 with :384 commented out, as committed)
 *   TNT_ATTN_LOSS_BATCH  total_loss += MSE(1, sum over the BATCH axis of alpha) (lc_NIC.py:365-367 with :384 enabled)
 *   TNT_ATTN_LOSS_ELEM   total_loss += MSE(1, alpha) element-wise (train_step_sam's first pass, lc_NIC.py:751-752,774) */
enum tnt_attn_loss { TNT_ATTN_LOSS_OFF = 0, TNT_ATTN_LOSS_BATCH = 1, TNT_ATTN_LOSS_ELEM = 2 };
enum tnt_encoder_kind { TNT_ENC_PARCEL = 0 /* layers.py:43-53 */ };

/* Constructor arguments of NIC (lc_NIC.py:42) + optimizer (main.py:97) + sizes. */
typedef struct tnt_config {
  int32_t N;        /* input width: 327684 vertices (config.yaml:46); image control: R*K */
  int32_t R;        /* parcels per encoder (len(groups[0])) */
  int32_t G;        /* group_size (<= 32) */
  int32_t U;        /* units */
  int32_t E;        /* embedding_text */
  int32_t A;        /* attn_units (<= 32) */
  int32_t H;        /* dense_inter width (256, lc_NIC.py:141) */
  int32_t V;        /* vocab_size */
  int32_t T;        /* max_length */
  int32_t B_max;    /* largest batch any call will pass */
  int32_t n_enc;    /* 1: lc_NIC;  2: ms2_NIC (first half of the batch -> encoder a, second -> b) */
  int32_t precision;
  float dropout_input, dropout_features, dropout_text, dropout_attn, dropout_lstm, dropout_out;
  float input_reg, attn_reg, lstm_reg, output_reg;
  float lr, beta1, beta2, adam_eps;
  float clipnorm;                 /* <= 0: no clipping (Keras < 2.4 behaviour) */
  int32_t emb_unaggregated_norm;  /* 1: clip norm of the embedding over IndexedSlices values */
  int32_t second_feature_dropout; /* 1: ms2_NIC.py:214,257 applies self.dropout twice */
  int32_t variant;                /* bit flags TNT_VARIANT_*; 0 = AttemptFour lc_NIC / ms2_NIC / img_NIC */
  int32_t attn_loss;              /* enum tnt_attn_loss (n_enc == 1 only); tnt_set_attn_loss changes it between steps */
  float agc_clip, agc_eps;        /* adaptive gradient clipping before apply_gradients (agc.py:20-38, call site
                                     lc_NIC.py:388: clip_factor 0.01, eps 1e-3); agc_clip <= 0: off */
} tnt_config;

/* ThinkAndTell/att_model.py variant of the same decoder:
 * MASKED_SPARSE_CE: SparseCategoricalCrossentropy(from_logits=True) x pad mask (target != 0), mean over the batch
 *   including the masked zeros, summed over the T steps and divided by T+1 (att_model.py:258-262,306-321);
 * PLAIN_ATTENTION: the two attention projections carry no LeakyReLU (att_model.py:11-29: fc1, fc2, f3). */
#define TNT_VARIANT_MASKED_SPARSE_CE 1
#define TNT_VARIANT_PLAIN_ATTENTION 2

/* One Keras variable inside the flat parameter buffer. */
typedef struct tnt_var_info {
  char name[64];      /* e.g. "enc0/dense_17/kernel", "lstm/recurrent_kernel" */
  int64_t offset;     /* in floats, into the flat w/m/v/g buffers (state buffer if !trainable) */
  int64_t rows, cols; /* logical shape (cols == 0: 1-D of length rows) */
  float l2;           /* regulariser lambda (0 if none) */
  int32_t trainable;  /* 0: BatchNormalization moving_mean / moving_variance */
} tnt_var_info;

/* Raw per-step sums (host doubles); the Python class turns them into the dict train_step returns
 * (lc_NIC.py:408; ms2_NIC.py:366-374).  Index h = half (0: whole batch or subject A, 1: subject B). */
typedef struct tnt_metrics {
  double ce_sum[2];       /* sum over rows (b,t) of -log clip(p_y) */
  double correct[2];      /* number of rows with argmax(pred) == target */
  double attn_mse[2];     /* lc_NIC.py:365-367 */
  double l2;              /* sum lambda_k ||W_k||^2 */
  double emb_sqnorm;      /* sum of squares of the embedding IndexedSlices values */
  double rows[2];         /* rows counted in each half (B_half * T) */
  double attn_elem_mse;   /* mean (1 - alpha)^2 over (T,B,R): the SAM step's attention loss (lc_NIC.py:751-752);
                             only computed when attn_loss == TNT_ATTN_LOSS_ELEM */
} tnt_metrics;

/* ---- lifetime ------------------------------------------------------------------------------ */
/* group_idx_host: all parcels' vertex indices concatenated (encoder 0's parcels, then encoder 1's);
 * group_off_host: n_enc*R + 1 offsets into it.  = the `groups` argument of NIC (lc_NIC.py:42,84). */
int tnt_create(const tnt_config* cfg, const int32_t* group_idx_host, const int64_t* group_off_host,
               tnt_ctx** out);
int tnt_destroy(tnt_ctx* ctx);
const char* tnt_last_error(tnt_ctx* ctx);   /* ctx may be NULL: message of the last failed tnt_create */

/* ---- parameters (model.get_weights / set_weights / get_layer: main.py:155-162,172-190) ------- */
int64_t tnt_param_floats(tnt_ctx* ctx);     /* length of each of the flat w, m, v, g buffers */
int64_t tnt_state_floats(tnt_ctx* ctx);     /* length of the non-trainable buffer */
int tnt_num_vars(tnt_ctx* ctx);
int tnt_var_info_get(tnt_ctx* ctx, int i, tnt_var_info* out);
int tnt_bind_params(tnt_ctx* ctx, float* w, float* m, float* v, float* g, float* state);
int tnt_set_adam_step(tnt_ctx* ctx, int64_t t);   /* optimizer.iterations */
int64_t tnt_get_adam_step(tnt_ctx* ctx);
int tnt_set_lr(tnt_ctx* ctx, float lr);           /* LearningRateScheduler hook (main.py:218-219) */

/* ---- whole-step entry points ------------------------------------------------------------------ */
/* NIC.train_step (lc_NIC.py:328-408 / ms2_NIC.py:294-374): forward, loss, backward, clip + Adam.
 * betas (B,N) f32; cap (B,T) i32; target_ids (B,T) i32 (= argmax of the reference's one-hot target);
 * h0,c0 (B,U).  B_global = B * world under data parallelism (loss/BN normalisers), else B. */
int tnt_train_step(tnt_ctx* ctx, const float* betas, const int32_t* cap, const int32_t* target_ids,
                   const float* h0, const float* c0, int B, uint64_t dropout_seed, void* stream);
/* The general step behind both: forward + loss/metrics (training: BN batch statistics + dropout), tape gradient when
 * `backward`, clip + Adam when `apply`.  backward && !apply leaves the tape gradient in g (loss_and_gradients, the
 * sharpness-aware step).  After tnt_dp_init this call also makes every data-parallel exchange. */
int tnt_step(tnt_ctx* ctx, const float* betas, const int32_t* cap, const int32_t* target_ids, const float* h0,
             const float* c0, int B, int training, int backward, int apply, uint64_t dropout_seed, void* stream);
/* NIC.test_step (lc_NIC.py:410-459): metrics only.  training != 0 keeps BN batch stats + dropout
 * (ms2_NIC.py:415-427 forces training=True). */
int tnt_eval_step(tnt_ctx* ctx, const float* betas, const int32_t* cap, const int32_t* target_ids,
                  const float* h0, const float* c0, int B, int training, uint64_t dropout_seed,
                  void* stream);
/* NIC.__call__ (lc_NIC.py:163,223-263): probs (B,T,V) and attention (T,B,R) (either may be NULL). */
int tnt_forward(tnt_ctx* ctx, const float* betas, const int32_t* cap, const float* h0, const float* c0,
                int B, int training, uint64_t dropout_seed, float* probs_out, float* attn_out,
                void* stream);
/* NIC.greedy_predict (lc_NIC.py:577-638).  out_ids (B,max_len) i64; optional out_probs (B,max_len,V),
 * out_attn (max_len,B,R), out_attn_hidden (max_len,B,R,A).  enc: which encoder (ms2) the rows use. */
int tnt_greedy(tnt_ctx* ctx, const float* betas, const float* h0, const float* c0, const int32_t* start,
               int B, int max_len, int enc, int64_t* out_ids, float* out_probs, float* out_attn,
               float* out_attn_hidden, void* stream);
/* Decoders with a random choice: lc_NIC.sample_choice (tf.random.categorical(log p), lc_NIC.py:571-575),
 * top-k and nucleus selection (img_NIC.py:509-533; the reference sorts ascending, which would keep the k LEAST likely
 * words -- the intended "k most likely / smallest set with mass > p" is built).  One rule covers all three: sort the
 * step's probabilities descending (ties: lower id first), keep the first n (n = V | top_k | shortest prefix whose sum
 * exceeds top_p), draw u from the shared Philox contract (counter = (row, site 9, step, 0), key = seed) and return the
 * first kept word whose running sum exceeds u * (sum of the kept).  top_k <= 0 and top_p <= 0 or >= 1: plain
 * categorical sampling.  Outputs as tnt_greedy.  Needs V <= 8192. */
int tnt_sample(tnt_ctx* ctx, const float* betas, const float* h0, const float* c0, const int32_t* start,
               int B, int max_len, int enc, int top_k, float top_p, uint64_t seed, int64_t* out_ids,
               float* out_probs, float* out_attn, float* out_attn_hidden, void* stream);
/* Beam search over the same decoder step (what lc_NIC.beam_search, lc_NIC.py:640-692, set out to do; the committed
 * function does not run).  Scores are sums of log-probabilities, no length normalisation; a beam that emitted end_id
 * is frozen (extended with id 0 at no cost).  B * beam <= B_max, beam <= 8.  out_ids (B, beam, max_len) i64 sorted by
 * score, out_scores (B, beam) f32. */
int tnt_beam_search(tnt_ctx* ctx, const float* betas, const float* h0, const float* c0, const int32_t* start,
                    int B, int max_len, int enc, int beam, int end_id, int64_t* out_ids, float* out_scores,
                    void* stream);
/* Synchronises `stream` and returns the context's status: TNT_ERR_CUDA if a persistent kernel's barrier timed out
 * (results of the last call are invalid; weights were NOT updated) or an earlier launch failed. */
int tnt_check(tnt_ctx* ctx, void* stream);
/* Names of the kernel variants the last calls ran ("enc_fwd=...;enc_bwd=...;recurrence=...;head=..."): every choice is a
 * function of the configuration's shape, never of the environment; bench.py records it. */
const char* tnt_kernel_variants(tnt_ctx* ctx);
int tnt_set_attn_loss(tnt_ctx* ctx, int mode);   /* enum tnt_attn_loss, effective from the next step */
/* Sharpness-aware step (train_step_sam, lc_NIC.py:713-838), driven by the caller as: step with backward but no update
 * (tnt_phase_*) -> tnt_sam_perturb: e = g * rho / (global_norm(g) + eps), w += e, e kept -> second backward pass at
 * w + e -> tnt_sam_restore: w -= e -> tnt_phase_optimizer with the second gradient.  g is the tape gradient (L2 terms
 * included; the embedding enters the global norm through its un-aggregated IndexedSlices values). */
int tnt_sam_perturb(tnt_ctx* ctx, float rho, float eps, void* stream);
int tnt_sam_restore(tnt_ctx* ctx, void* stream);
/* Copies the raw sums of the last train/eval step to host memory (synchronises `stream`). */
int tnt_read_metrics(tnt_ctx* ctx, tnt_metrics* out_host, void* stream);
/* One-hot (rows,V) f32 -> ids (rows) i32, first maximal index (lc_NIC.py:481 argmax on the target). */
int tnt_onehot_to_ids(tnt_ctx* ctx, const float* onehot, int64_t rows, int32_t* ids_out, void* stream);

/* ---- data parallelism inside the library ------------------------------------------------------
 * One process per GPU; rank k holds rows [k*B, (k+1)*B) of the global batch.  Rank 0 calls tnt_dp_unique_id and ships the
 * 256 bytes (two ncclUniqueId) to every rank by any means (the Python class broadcasts them with torch.distributed);
 * then EVERY rank calls tnt_dp_init (collective).  From then on tnt_step / tnt_train_step / tnt_eval_step exchange
 * BatchNorm sums (one-shot all-reduce over peer memory), gradient buckets, metric sums and attention column sums
 * (NCCL, enqueued on the producing stream) themselves; tnt_read_metrics returns global sums.  Dropout masks are the
 * caller's business: pass a rank-dependent dropout_seed.  NCCL is bound with dlopen at the first call. */
int tnt_dp_unique_id(void* out256_host);
int tnt_dp_init(tnt_ctx* ctx, int rank, int world, const void* unique_id256_host, void* stream);
int tnt_dp_world(tnt_ctx* ctx);
/* ---- phases: tnt_step == the phases back to back (kept for timing and for hosts that interleave their own work) -- */
/* Between phases the caller all-reduces (sum) the named buffers over its process group:
 *   after phase 1: "bn_fwd_stats" (f64);  after phase 2: "bn_bwd_stats" (f64), "metrics_raw" (f64),
 *   "attn_colsum" (f32) and the decoder slice of g;  after phase 3: the encoder slice of g. */
/* Optional phase 0: announce the step.  The encoder-independent head of the decoder (embedding gather, the
 * time-parallel x_t @ W product, weight packing) starts on the library's side stream and overlaps phase 1;
 * tnt_phase_decode picks it up when (cap, B, training, dropout_seed) match, else it does that work inline. */
int tnt_phase_prologue(tnt_ctx* ctx, const int32_t* cap, int B, int training, int backward,
                       uint64_t dropout_seed, void* stream);
int tnt_phase_encode(tnt_ctx* ctx, const float* betas, int B, int B_global, int training,
                     uint64_t dropout_seed, void* stream);
int tnt_phase_decode(tnt_ctx* ctx, const int32_t* cap, const int32_t* target_ids, const float* h0,
                     const float* c0, int B, int B_global, int training, int backward,
                     uint64_t dropout_seed, void* stream);
int tnt_phase_encoder_backward(tnt_ctx* ctx, const float* betas, int B, int B_global, void* stream);
int tnt_phase_optimizer(tnt_ctx* ctx, void* stream);
/* Phase 4 per gradient bucket (data parallel: update the decoder variables while the encoder gradients are still in
 * flight): part 1 = decoder, part 0 = encoder; `first` on the first call of the step, `last` on the last. */
int tnt_phase_optimizer_part(tnt_ctx* ctx, int part, int first, int last, void* stream);
/* Work off the critical path (reported metrics, decoder weight gradients) is finished by a library-owned side
 * stream.  A data-parallel caller either enqueues the all-reduce of "metrics_raw", "attn_colsum" and the decoder
 * slice of g in that stream's order (so it overlaps phase 3), or calls tnt_side_join first.  NULL: no side stream. */
void* tnt_side_stream(tnt_ctx* ctx);
int tnt_side_join(tnt_ctx* ctx, void* stream);
/* test_step's `tf.add_n(self.losses)` (lc_NIC.py:457): the L2 sum only, no gradient, no update. */
int tnt_eval_l2(tnt_ctx* ctx, void* stream);
/* Named device buffers: the DP exchange buffers above and internal activations (tests/debug). */
int tnt_buffer(tnt_ctx* ctx, const char* name, void** ptr_out, int64_t* bytes_out);
/* [offset,len) in floats of the decoder / encoder parts of the flat gradient (DP buckets). */
int tnt_grad_bucket(tnt_ctx* ctx, int which /*0 encoder, 1 decoder = 3 + 2, 2 head (dense_inter, dense_out), 3 decoder
                                              without the head*/, int64_t* offset, int64_t* len);
/* Data parallel: make `stream` wait for an intermediate result of the step just enqueued with tnt_phase_decode.
 * TNT_EV_HEAD_GRADS: bucket 2 is final (before the recurrence gradient runs, so its all-reduce hides behind it);
 * TNT_EV_METRICS: "metrics_raw" and "attn_colsum" are final.  Without a side stream the call is a no-op (enqueue
 * the all-reduce in the main stream's order instead). */
enum tnt_event { TNT_EV_HEAD_GRADS = 0, TNT_EV_METRICS = 1 };
int tnt_stream_wait_event(tnt_ctx* ctx, int which, void* stream);
/* How many kernels the library has launched since creation (bench.py's gpu_launches). */
int64_t tnt_launch_count(tnt_ctx* ctx);
/* In-situ kernel timing (CUDA event pairs around each kernel group, on the stream the group runs on; bench.py's
 * roofline).  tnt_trace(ctx, 1) clears the records and starts recording, tnt_trace(ctx, 0) stops.  tnt_trace_read
 * synchronises the device and writes up to max_records entries: names as '\n'-separated text into `names`
 * (names_len bytes incl. the terminating NUL) and the matching durations in milliseconds into `ms`; returns the
 * number of records written, or -1 on bad arguments. */
int tnt_trace(tnt_ctx* ctx, int enable);
int tnt_trace_read(tnt_ctx* ctx, char* names, int names_len, float* ms, int max_records);
/* Diagnostic: per-tile clock64() stamps of CTA 0 of the two tcgen05 encoder kernels (tools/enc_trace.py).  dev_buf = device
 * buffer of 2 x 512 x 8 int64 (forward | gradient; load issued, poll begin, tile ready, conversion done, MMA go, MMA
 * committed, accumulator seen, drained), NULL switches it off again.  No reference counterpart. */
void tnt_debug_encoder_trace(tnt_ctx* ctx, long long* dev_buf);

/* ---- dense voxel encoder of AttemptFour/Model/NIC.py (SURVEY 8a row A1'') ---------------------------------
 * features = BatchNormalization()(expand_dims(Dropout(LeakyReLU_0.2(dropout_input(x) @ W + b)), 1))
 * (layer NIC.py:64-69, call NIC.py:122-128).  Standalone object; parameters are one flat fp32 buffer
 * [W (K,E) row-major = Keras kernel | bias E | gamma E | beta E] with matching m, v, g; state = [moving_mean E |
 * moving_variance E].  Same conventions as above (device pointers, async on `stream`, negative status on error). */
typedef struct tnt_dense tnt_dense;
typedef struct tnt_dense_config {
  int32_t K;       /* input_size: 62756 (visual cortex) or 327684 */
  int32_t E;       /* embedding_dim (multiple of 128; 512 in the reference) */
  int32_t B_max;   /* <= 64 */
  float dropout_input, dropout;   /* NIC.py:57-58 */
  float input_reg;                /* L2 on the kernel, NIC.py:53,67 */
  float lr, beta1, beta2, adam_eps, clipnorm;
} tnt_dense_config;
int tnt_dense_create(const tnt_dense_config* cfg, tnt_dense** out);
int tnt_dense_destroy(tnt_dense* d);
const char* tnt_dense_last_error(tnt_dense* d);
int64_t tnt_dense_param_floats(tnt_dense* d);
int64_t tnt_dense_launch_count(tnt_dense* d);
int tnt_dense_bind_params(tnt_dense* d, float* w, float* m, float* v, float* g, float* state);
int tnt_dense_set_adam_step(tnt_dense* d, int64_t t);
int tnt_dense_set_lr(tnt_dense* d, float lr);
/* x (B,K) -> features (B,E) [= (B,1,E)].  training: dropout (Philox sites INPUT / FEAT) + batch statistics. */
int tnt_dense_forward(tnt_dense* d, const float* x, int B, int training, uint64_t dropout_seed, float* features,
                      void* stream);
/* Inference forward (training = 0: moving statistics, no dropout) that allocates NOTHING -- no cudaMalloc / cudaFree and
 * no allocating library routine -- for callers under CUDA graph capture or in serving loops that forbid allocation.
 * NO COUNTERPART IN THE REFERENCE.  Same kernels and summation order as
 * tnt_dense_forward(d, x, B, 0, 0, features, stream): bit-identical features.  Its scratch is one file-scope
 * static __device__ array (32.1 MiB, held by every GPU that loads the library until the process exits), so:
 *   limits   K <= 65536, E <= 512, B <= B_max <= 64 (and S * 64 * E <= 4 Mi split-K partial sums, true for any such
 *            K, E on <= 170 SMs); anything larger returns TNT_ERR_ARG with a message -- use tnt_dense_forward;
 *   sharing  all calls on one GPU share the array and it keeps the last call's contents (every region a call reads is
 *            rewritten by that call first, nothing needs clearing).  Calls that OVERLAP on the GPU -- two streams, two
 *            host threads, two objects -- race; serialise them on one stream or with events.
 * Does not disturb the training state of `d` (may be interleaved with forward/backward/apply on the same stream). */
int tnt_dense_forward_noalloc(tnt_dense* d, const float* x, int B, float* features, void* stream);
/* d(loss)/d(features) (B,E) of the last training forward -> tape gradient of W, b, gamma, beta in g. */
int tnt_dense_backward(tnt_dense* d, const float* dfeatures, void* stream);
/* Adam(clipnorm).apply_gradients: per-variable clip, L2 gradient of the kernel, Adam. */
int tnt_dense_apply_gradients(tnt_dense* d, void* stream);

/* ---- batch assembly from HBM-resident tables ------------------------------------------------
 * Replaces the per-batch host work of DataGenerator.__data_generation (AttemptFour/DataLoaders/
 * data_generator_guse.py:129-171; two-subject layout data_generator_multisub.py:81-120): one np.load of a
 * 327 684-float row per sample, pad_sequences, the shifted target and to_categorical.  The subject's beta matrix
 * (n_rows, width) and the tokenised + padded caption table (n_caps, T) stay in HBM; a batch is an index gather.
 * Stateless (no tnt_ctx); device pointers; asynchronous on `stream`; an index outside its table never reads out of
 * bounds: the row is zero-filled and *bad_flag (device int32, caller-owned, never cleared here) is set to 1. */
/* out[j, :] = table[rows[j], :] for j < B            (data_generator_guse.py:146-153) */
int tnt_assemble_betas(const float* table, int64_t n_rows, int64_t width, const int32_t* rows, int B, float* out,
                       int32_t* bad_flag, void* stream);
/* cap_out[j, :] = cap_table[idx[j], :];  target_ids_out[j, t] = cap_table[idx[j], t + 1], 0 at t = T - 1
 * (data_generator_guse.py:160-162); onehot_out (B,T,V) f32 = to_categorical(target, V) (:163) when non-NULL.
 * cap_out / target_ids_out / onehot_out may each be NULL. */
int tnt_assemble_captions(const int32_t* cap_table, int64_t n_caps, int T, const int32_t* idx, int B,
                          int32_t* cap_out, int32_t* target_ids_out, float* onehot_out, int V, int32_t* bad_flag,
                          void* stream);
const char* tnt_data_last_error(void);
int64_t tnt_data_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* TNT_H */
