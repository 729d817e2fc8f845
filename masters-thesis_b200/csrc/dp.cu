// Data parallelism inside the library (tnt_dp_init; 8e): batch-sharded replicas, one process per GPU.
// After tnt_dp_init the step entry points do their own exchanges, so the host makes ONE call per step:
//   * BatchNorm batch sums (forward and backward; 64 doubles per encoder) are on the critical path of the step and far
//     below NCCL's launch + protocol floor (27-41 us measured): they go through a one-shot all-reduce over peer memory
//     -- every rank stores its payload + a sequence flag straight into every peer's mailbox over NVLink (CUDA IPC
//     mappings of one cudaMalloc'ed buffer per rank), spins on its own mailbox and sums in rank order (identical bits on
//     every rank).  If the IPC mapping is unavailable the same call goes through NCCL (recorded in tnt_kernel_variants).
//   * gradient buckets (head, rest of the decoder, encoder), the metric sums and the attention column sums are NCCL
//     all-reduces enqueued on the stream that produces them (two communicators: main-stream and side-stream traffic
//     never queue behind each other), so each transfer starts the moment its last producer kernel retires and overlaps
//     the rest of the backward pass.
// NCCL is bound at run time (dlopen of the copy already loaded into the process by PyTorch, else the system one): the
// library has no link-time dependency on it and single-GPU use never touches it.
#include "tnt_internal.h"
#include <dlfcn.h>
#include <nccl.h>
#include <cstring>
#include <cstdlib>
#include <algorithm>

struct NcclApi {
  void* h = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
static NcclApi g_nccl;

static bool nccl_load(std::string* err) {
  if (g_nccl.h) return true;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);       // the copy PyTorch already mapped, if any
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) { *err = std::string("cannot load libnccl.so.2: ") + dlerror(); return false; }
#define SYM(field, name) *(void**)(&g_nccl.field) = dlsym(h, name); if (!g_nccl.field) { *err = std::string("libnccl lacks ") + name; return false; }
  SYM(GetUniqueId, "ncclGetUniqueId"); SYM(CommInitRank, "ncclCommInitRank"); SYM(CommDestroy, "ncclCommDestroy");
  SYM(AllReduce, "ncclAllReduce"); SYM(AllGather, "ncclAllGather"); SYM(GroupStart, "ncclGroupStart");
  SYM(GroupEnd, "ncclGroupEnd"); SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
  g_nccl.h = h;
  return true;
}

#define DP_MAX_WORLD 16
#define DP_SLOTS 4
#define DP_PAYLOAD 128                       // doubles per small all-reduce (2 encoders x 64)
#define DP_ROW (DP_PAYLOAD + 8)              // payload + flag word (+ padding to 64 bytes)
// The mailbox has two LANES, one per stream that reduces (0 = main stream: BatchNorm sums; 1 = side stream: metric sums),
// each with its own call counter: calls of a lane are stream ordered, and a peer is never more than one call ahead.
#define DP_LANE_DOUBLES ((size_t)DP_SLOTS * DP_MAX_WORLD * DP_ROW)
// behind the two double lanes: a float lane for medium reductions (side stream: attention column sums, <= 8192 floats)
#define DP_F_PAYLOAD 8192
#define DP_F_ROW (DP_F_PAYLOAD + 16)         // payload + flag word (+ padding to 64 bytes)
#define DP_MAILBOX_BYTES (2 * DP_LANE_DOUBLES * sizeof(double) + (size_t)DP_SLOTS * DP_MAX_WORLD * DP_F_ROW * sizeof(float))

struct DpState {
  int rank = 0, world = 1;
  ncclComm_t comm_main = nullptr, comm_side = nullptr;
  double* mailbox = nullptr;                 // [2 lanes][DP_SLOTS][world][DP_ROW] doubles | [DP_SLOTS][world][DP_F_ROW] floats
  double* peers[DP_MAX_WORLD] = {};          // every rank's mailbox as seen from here (peers[rank] == mailbox)
  double** d_peers = nullptr;                // device copy of the table
  unsigned long long seq[3] = {0, 0, 0};     // calls made so far per lane (identical on every rank: program order)
  int use_peer = 0;
  // encoder-gradient exchange (dp_grad_exchange): one IPC buffer per rank = [flags 4 KB | in: world x SL | out: L]
  float* gbuf = nullptr;                     // this rank's buffer
  float* gpeers[DP_MAX_WORLD] = {};          // every rank's buffer as seen from here
  int64_t gL = 0, gSL = 0;                   // floats exchanged / floats per slice
  unsigned long long gseq = 0;               // barrier generation (identical on every rank: program order)
  int use_push = 0;
};
#define DP_GFLAGS 1024                       // floats reserved for the flag words in front of the buffer

static bool nccl_ok(tnt_ctx* c, ncclResult_t r, const char* where) {
  if (r == ncclSuccess) return true;
  ctx_fail(c, TNT_ERR_CUDA, std::string(where) + ": NCCL: " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?"));
  return false;
}

// ---- one-shot all-reduce over peer memory -------------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(DP_PAYLOAD)
dp_oneshot_kernel(double* const* __restrict__ peers, int rank, int world, unsigned long long seq, double* __restrict__ buf,
                  int n, unsigned* __restrict__ abort_flag, int lane) {
  __shared__ int s_ok;
  const int tid = threadIdx.x;
  const size_t slot = (size_t)lane * DP_LANE_DOUBLES + (size_t)(seq % DP_SLOTS) * world * DP_ROW;
  const double v = tid < n ? buf[tid] : 0.0;
  for (int p = 0; p < world; ++p)
    if (tid < n) peers[p][slot + (size_t)rank * DP_ROW + tid] = v;
  __threadfence_system();
  __syncthreads();
  if (tid < world)
    st_release_sys(reinterpret_cast<unsigned long long*>(peers[tid] + slot + (size_t)rank * DP_ROW + DP_PAYLOAD), seq + 1);
  if (tid == 0) s_ok = 1;
  __syncthreads();
  if (tid < world) {
    const unsigned long long* f = reinterpret_cast<const unsigned long long*>(peers[rank] + slot + (size_t)tid * DP_ROW + DP_PAYLOAD);
    const long long t0 = clock64();
    while (ld_acquire_sys(f) != seq + 1)
      if (clock64() - t0 > (1ll << 33)) { s_ok = 0; break; }           // ~4 s: a peer died; never hang the GPU
  }
  __syncthreads();
  if (!s_ok) { if (tid == 0 && abort_flag) atomicExch(abort_flag, 1u); return; }
  if (tid < n) {
    double s = 0.0;
    for (int p = 0; p < world; ++p) s += peers[rank][slot + (size_t)p * DP_ROW + tid];    // rank order: same bits everywhere
    buf[tid] = s;
  }
}

// the same for <= DP_F_PAYLOAD floats (one CTA of 1024 threads, 128-bit stores): attention column sums on the side stream
__global__ void __launch_bounds__(1024)
dp_oneshot_f32_kernel(double* const* __restrict__ peers, int rank, int world, unsigned long long seq, float* __restrict__ buf,
                      int n, unsigned* __restrict__ abort_flag) {
  __shared__ int s_ok;
  const int tid = threadIdx.x;
  const size_t slot = (size_t)(seq % DP_SLOTS) * world * DP_F_ROW;
  const int n4 = (n + 3) >> 2;
  for (int i = tid; i < n4; i += blockDim.x) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (4 * i + 3 < n) v = *reinterpret_cast<const float4*>(buf + 4 * i);     // buf is a 256-byte aligned workspace
    else { v.x = buf[4 * i]; if (4 * i + 1 < n) v.y = buf[4 * i + 1]; if (4 * i + 2 < n) v.z = buf[4 * i + 2]; }
    for (int p = 0; p < world; ++p) {
      float* mb = reinterpret_cast<float*>(peers[p] + 2 * DP_LANE_DOUBLES);
      reinterpret_cast<float4*>(mb + slot + (size_t)rank * DP_F_ROW)[i] = v;
    }
  }
  __threadfence_system();
  __syncthreads();
  if (tid < world) {
    float* mb = reinterpret_cast<float*>(peers[tid] + 2 * DP_LANE_DOUBLES);
    st_release_sys(reinterpret_cast<unsigned long long*>(mb + slot + (size_t)rank * DP_F_ROW + DP_F_PAYLOAD), seq + 1);
  }
  if (tid == 0) s_ok = 1;
  __syncthreads();
  const float* mine = reinterpret_cast<const float*>(peers[rank] + 2 * DP_LANE_DOUBLES);
  if (tid < world) {
    const unsigned long long* f = reinterpret_cast<const unsigned long long*>(mine + slot + (size_t)tid * DP_F_ROW + DP_F_PAYLOAD);
    const long long t0 = clock64();
    while (ld_acquire_sys(f) != seq + 1)
      if (clock64() - t0 > (1ll << 33)) { s_ok = 0; break; }
  }
  __syncthreads();
  if (!s_ok) { if (tid == 0 && abort_flag) atomicExch(abort_flag, 1u); return; }
  for (int i = tid; i < n; i += blockDim.x) {
    float s = 0.0f;
    for (int p = 0; p < world; ++p) s += mine[slot + (size_t)p * DP_F_ROW + i];      // rank order: same bits everywhere
    buf[i] = s;
  }
}

// ---- encoder-gradient exchange over peer memory --------------------------------------------------------------------
// buffer of rank q (floats): [0, DP_GFLAGS) flag words (uint64 flags[2][world]: barrier b, source rank) |
//                            in  [world][SL]  partials of q's slice, row = source rank |
//                            out [L]          the reduced gradient, slice s written by rank s
struct DpGx {
  float* buf[TNT_DP_MAX_WORLD];
  int64_t L, SL;
  int32_t rank, world;
  unsigned long long gen;                    // barrier generation of this call
};
__device__ __forceinline__ float* gx_in(const DpGx& x, int q) { return x.buf[q] + DP_GFLAGS; }
__device__ __forceinline__ float* gx_out(const DpGx& x, int q) { return x.buf[q] + DP_GFLAGS + (size_t)x.world * x.SL; }

// Cross-GPU barrier at the head of a kernel: everything this rank's stream did before this kernel (stream order) is
// released to every peer by CTA 0; every CTA waits until all peers have said the same.  false: a peer never arrived.
__device__ __forceinline__ bool gx_barrier(const DpGx& x, int which, unsigned* abort_flag) {
  __shared__ int s_ok;
  const int tid = threadIdx.x;
  if (tid == 0) s_ok = 1;
  if (blockIdx.x == 0 && tid < x.world) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<unsigned long long*>(x.buf[tid]) + which * TNT_DP_MAX_WORLD + x.rank, x.gen);
  }
  __syncthreads();
  if (tid < x.world) {
    const unsigned long long* f = reinterpret_cast<const unsigned long long*>(x.buf[x.rank]) + which * TNT_DP_MAX_WORLD + tid;
    const long long t0 = clock64();
    while (ld_acquire_sys(f) < x.gen)
      if (clock64() - t0 > (1ll << 33)) { s_ok = 0; break; }           // ~4 s: a peer died; never hang the GPU
  }
  __syncthreads();
  if (!s_ok && tid == 0 && abort_flag) atomicExch(abort_flag, 1u);
  return s_ok != 0;
}

struct DpRanges { int64_t beg[6], end[6]; int32_t n; };
// the parts of g[0, L) no producing kernel has pushed (biases, BatchNorm gamma / beta; everything on the generic paths)
__global__ void __launch_bounds__(256)
dp_push_ranges_kernel(const DpGx x, const float* __restrict__ g, const DpRanges r) {
  for (int k = 0; k < r.n; ++k) {
    const int64_t n4 = (r.end[k] - r.beg[k]) >> 2;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
      const int64_t o = r.beg[k] + 4 * i;
      const int q = (int)(o / x.SL);
      *reinterpret_cast<float4*>(gx_in(x, q) + (size_t)x.rank * x.SL + (o - (int64_t)q * x.SL)) =
          *reinterpret_cast<const float4*>(g + o);
    }
  }
}

// owner: slice := sum over source ranks in rank order (identical bits on every rank), broadcast into every rank's `out`
// (The cross-GPU barriers are one-warp kernels of their own, dp_wait_kernel: a bandwidth kernel that spins with 148 fat
// CTAs keeps the side stream's GEMMs and NCCL kernels off the SMs -- on every rank at once -- measured at N = 8.)
__global__ void __launch_bounds__(256)
dp_sum_bcast_kernel(const DpGx x) {
  const int64_t base = (int64_t)x.rank * x.SL;
  const int64_t n4 = (min(x.SL, x.L - base) + 3) >> 2;                     // <= 0 for a rank beyond the end
  const float4* in = reinterpret_cast<const float4*>(gx_in(x, x.rank));
  const int64_t sl4 = x.SL >> 2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 v[TNT_DP_MAX_WORLD];
#pragma unroll
    for (int p = 0; p < TNT_DP_MAX_WORLD; ++p)
      if (p < x.world) v[p] = in[(size_t)p * sl4 + i];
    float4 s = v[0];
#pragma unroll
    for (int p = 1; p < TNT_DP_MAX_WORLD; ++p)
      if (p < x.world) { s.x += v[p].x; s.y += v[p].y; s.z += v[p].z; s.w += v[p].w; }
#pragma unroll
    for (int q = 0; q < TNT_DP_MAX_WORLD; ++q)
      if (q < x.world) reinterpret_cast<float4*>(gx_out(x, q) + base)[i] = s;
  }
}

__global__ void dp_wait_kernel(const DpGx x, int which, unsigned* __restrict__ abort_flag) { gx_barrier(x, which, abort_flag); }

// every slice has been broadcast: out -> g (the optimizer and the gradient read-back keep reading g)
__global__ void __launch_bounds__(256)
dp_collect_kernel(const DpGx x, float* __restrict__ g) {
  const float4* out = reinterpret_cast<const float4*>(gx_out(x, x.rank));
  const int64_t n4 = x.L >> 2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x)
    reinterpret_cast<float4*>(g)[i] = out[i];
}

// ---- public to the rest of the library ----------------------------------------------------------------------------
int dp_world(tnt_ctx* c) { return c->dp ? c->dp->world : 1; }
int dp_rank(tnt_ctx* c) { return c->dp ? c->dp->rank : 0; }
const char* dp_small_kind(tnt_ctx* c) { return !c->dp ? "none" : (c->dp->use_peer ? "peer-memory one-shot" : "nccl"); }
const char* dp_grad_kind(tnt_ctx* c) {
  return !c->dp ? "none" : (c->dp->use_push ? "peer-memory push reduce-scatter fused into enc_bwd_dw_tc_kernel + sum/broadcast" : "nccl all-reduce");
}

// critical-path reductions of <= 128 doubles (BatchNorm sums)
void dp_allreduce_small(tnt_ctx* c, double* buf, int n, cudaStream_t st) {
  DpState* d = c->dp;
  if (!d || d->world == 1) return;
  const int lane = (c->side && st == c->side) ? 1 : 0;
  if (d->use_peer && n <= DP_PAYLOAD) {
    dp_oneshot_kernel<<<1, DP_PAYLOAD, 0, st>>>(d->d_peers, d->rank, d->world, d->seq[lane], buf, n, pd_abort_flag(c), lane);
    d->seq[lane] += 1;
    c->launches += 1;
    return;
  }
  nccl_ok(c, g_nccl.AllReduce(buf, buf, (size_t)n, ncclFloat64, ncclSum, lane ? d->comm_side : d->comm_main, st),
          "dp_allreduce_small");
}

// <= 8192 floats, one stream only (the float lane has one call counter): the attention column sums on the side stream
void dp_allreduce_f32_small(tnt_ctx* c, float* buf, int n, cudaStream_t st) {
  DpState* d = c->dp;
  if (!d || d->world == 1 || n <= 0) return;
  if (d->use_peer && n <= DP_F_PAYLOAD && d->world <= DP_MAX_WORLD) {
    dp_oneshot_f32_kernel<<<1, 1024, 0, st>>>(d->d_peers, d->rank, d->world, d->seq[2], buf, n, pd_abort_flag(c));
    d->seq[2] += 1;
    c->launches += 1;
    return;
  }
  nccl_ok(c, g_nccl.AllReduce(buf, buf, (size_t)n, ncclFloat32, ncclSum, d->comm_side, st), "dp_allreduce_f32_small");
}

// lane 0: main-stream traffic (encoder bucket), lane 1: side-stream traffic (decoder buckets, metrics)
void dp_allreduce(tnt_ctx* c, void* buf, int64_t n, int is_double, int lane, cudaStream_t st) {
  DpState* d = c->dp;
  if (!d || d->world == 1 || n <= 0) return;
  nccl_ok(c, g_nccl.AllReduce(buf, buf, (size_t)n, is_double ? ncclFloat64 : ncclFloat32, ncclSum,
                              lane ? d->comm_side : d->comm_main, st), "dp_allreduce");
}

bool dp_push_ready(tnt_ctx* c) { return c->dp && c->dp->world > 1 && c->dp->use_push; }

DpPush dp_push_params(tnt_ctx* c) {
  DpPush p = {};
  DpState* d = c->dp;
  for (int q = 0; q < d->world; ++q) p.in[q] = d->gpeers[q] + DP_GFLAGS;
  p.SL = d->gSL; p.rank = d->rank; p.world = d->world;
  return p;
}

void dp_grad_exchange(tnt_ctx* c, float* g, int64_t L, const int64_t* pushed, int n_pushed, int direct, cudaStream_t st) {
  DpState* d = c->dp;
  if (!d || d->world == 1 || L <= 0) return;
  if (!d->use_push || L != d->gL) {                       // not set up for this length: plain NCCL
    nccl_ok(c, g_nccl.AllReduce(g, g, (size_t)L, ncclFloat32, ncclSum, d->comm_main, st), "dp_grad_exchange");
    return;
  }
  DpGx x = {};
  for (int q = 0; q < d->world; ++q) x.buf[q] = d->gpeers[q];
  x.L = L; x.SL = d->gSL; x.rank = d->rank; x.world = d->world; x.gen = ++d->gseq;
  // complement of the pushed ranges (sorted, disjoint) in [0, L)
  DpRanges r = {};
  int64_t at = 0;
  for (int k = 0; k <= n_pushed; ++k) {
    const int64_t b = k < n_pushed ? pushed[2 * k] : L;
    if (b > at && r.n < 6) { r.beg[r.n] = at; r.end[r.n] = b; r.n += 1; }
    if (k < n_pushed) at = pushed[2 * k + 1];
  }
  int64_t rest = 0;
  for (int k = 0; k < r.n; ++k) rest += r.end[k] - r.beg[k];
  if (rest > 0) {
    const int grid = (int)std::min<int64_t>((rest / 4 + 255) / 256, (int64_t)c->sm_count * 4);
    dp_push_ranges_kernel<<<grid > 0 ? grid : 1, 256, 0, st>>>(x, g, r);
    c->launches += 1;
  }
  dp_wait_kernel<<<1, 32, 0, st>>>(x, 0, pd_abort_flag(c));            // every rank's pushes have landed
  dp_sum_bcast_kernel<<<c->sm_count, 256, 0, st>>>(x);
  dp_wait_kernel<<<1, 32, 0, st>>>(x, 1, pd_abort_flag(c));            // every slice has been broadcast
  if (direct) c->g_enc_src = d->gbuf + DP_GFLAGS + (size_t)d->world * d->gSL;
  else dp_collect_kernel<<<c->sm_count, 256, 0, st>>>(x, g);
  c->launches += direct ? 3 : 4;
}

void dp_destroy(tnt_ctx* c) {
  DpState* d = c->dp;
  if (!d) return;
  cudaDeviceSynchronize();
  for (int p = 0; p < d->world; ++p)
    if (p != d->rank && d->peers[p]) cudaIpcCloseMemHandle(d->peers[p]);
  for (int p = 0; p < d->world && p < TNT_DP_MAX_WORLD; ++p)
    if (p != d->rank && d->gpeers[p]) cudaIpcCloseMemHandle(d->gpeers[p]);
  if (d->gbuf) cudaFree(d->gbuf);
  if (d->mailbox) cudaFree(d->mailbox);
  if (d->d_peers) cudaFree(d->d_peers);
  if (d->comm_main) g_nccl.CommDestroy(d->comm_main);
  if (d->comm_side) g_nccl.CommDestroy(d->comm_side);
  cudaGetLastError();
  delete d;
  c->dp = nullptr;
}

extern "C" int tnt_dp_unique_id(void* out256) {
  std::string err;
  if (!out256 || !nccl_load(&err)) return TNT_ERR_CUDA;
  ncclUniqueId ids[2];
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  if (g_nccl.GetUniqueId(&ids[0]) != ncclSuccess || g_nccl.GetUniqueId(&ids[1]) != ncclSuccess) return TNT_ERR_CUDA;
  memcpy(out256, ids, 256);
  return TNT_OK;
}

extern "C" int tnt_dp_init(tnt_ctx* c, int rank, int world, const void* unique_id256, void* stream) {
  if (!c || !unique_id256 || world < 1 || world > DP_MAX_WORLD || rank < 0 || rank >= world) return TNT_ERR_ARG;
  if (c->dp) dp_destroy(c);
  if (world == 1) return TNT_OK;
  std::string err;
  if (!nccl_load(&err)) { c->err = err; return TNT_ERR_CUDA; }
  cudaStream_t st = (cudaStream_t)stream;
  DpState* d = new DpState();
  d->rank = rank; d->world = world;
  c->dp = d;
  ncclUniqueId ids[2];
  memcpy(ids, unique_id256, 256);
  bool ok = nccl_ok(c, g_nccl.GroupStart(), "tnt_dp_init");
  ok = ok && nccl_ok(c, g_nccl.CommInitRank(&d->comm_main, world, ids[0], rank), "tnt_dp_init");
  ok = ok && nccl_ok(c, g_nccl.CommInitRank(&d->comm_side, world, ids[1], rank), "tnt_dp_init");
  ok = ok && nccl_ok(c, g_nccl.GroupEnd(), "tnt_dp_init");
  if (!ok) { int rc = c->fail_code; dp_destroy(c); return rc ? rc : TNT_ERR_CUDA; }
  // mailboxes: one cudaMalloc'ed buffer per rank, IPC handles exchanged with an all-gather, every peer mapped
  const size_t mb = DP_MAILBOX_BYTES;
  cudaIpcMemHandle_t mine, *all_d = nullptr;
  std::vector<cudaIpcMemHandle_t> all(world);
  bool peer_ok = cudaMalloc(&d->mailbox, mb) == cudaSuccess && cudaMemset(d->mailbox, 0, mb) == cudaSuccess &&
                 cudaIpcGetMemHandle(&mine, d->mailbox) == cudaSuccess &&
                 cudaMalloc(&all_d, sizeof(cudaIpcMemHandle_t) * world) == cudaSuccess;
  int flag_local = peer_ok ? 1 : 0;
  if (peer_ok) {
    cudaMemcpyAsync(all_d + rank, &mine, sizeof(mine), cudaMemcpyHostToDevice, st);
    ok = nccl_ok(c, g_nccl.AllGather(all_d + rank, all_d, sizeof(mine), ncclUint8, d->comm_main, st), "tnt_dp_init");
    cudaMemcpyAsync(all.data(), all_d, sizeof(mine) * world, cudaMemcpyDeviceToHost, st);
    if (!ok || cudaStreamSynchronize(st) != cudaSuccess) peer_ok = false;
  }
  if (peer_ok) {
    for (int p = 0; p < world && peer_ok; ++p) {
      if (p == rank) { d->peers[p] = d->mailbox; continue; }
      void* ptr = nullptr;
      if (cudaIpcOpenMemHandle(&ptr, all[p], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); peer_ok = false; }
      d->peers[p] = (double*)ptr;
    }
    flag_local = peer_ok ? 1 : 0;
  }
  // gradient-exchange buffers (dp_grad_exchange): same recipe, one more handle per rank
  if (peer_ok && world <= TNT_DP_MAX_WORLD && c->dec_off > 0 && c->dec_off % 4 == 0) {
    d->gL = c->dec_off;
    d->gSL = ((d->gL + world - 1) / world + 31) / 32 * 32;
    const size_t gb = sizeof(float) * ((size_t)DP_GFLAGS + (size_t)world * d->gSL + (size_t)d->gL);
    cudaIpcMemHandle_t gmine;
    bool push_ok = cudaMalloc(&d->gbuf, gb) == cudaSuccess && cudaMemset(d->gbuf, 0, gb) == cudaSuccess &&
                   cudaIpcGetMemHandle(&gmine, d->gbuf) == cudaSuccess;
    if (!push_ok) { cudaGetLastError(); memset(&gmine, 0, sizeof(gmine)); }
    // every rank takes part in the all-gather whether or not its allocation worked
    cudaMemcpyAsync(all_d + rank, &gmine, sizeof(gmine), cudaMemcpyHostToDevice, st);
    ok = nccl_ok(c, g_nccl.AllGather(all_d + rank, all_d, sizeof(gmine), ncclUint8, d->comm_main, st), "tnt_dp_init");
    cudaMemcpyAsync(all.data(), all_d, sizeof(gmine) * world, cudaMemcpyDeviceToHost, st);
    if (!ok || cudaStreamSynchronize(st) != cudaSuccess) push_ok = false;
    for (int p = 0; p < world && push_ok; ++p) {
      if (p == rank) { d->gpeers[p] = d->gbuf; continue; }
      void* ptr = nullptr;
      if (cudaIpcOpenMemHandle(&ptr, all[p], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); push_ok = false; }
      d->gpeers[p] = (float*)ptr;
    }
    if (push_ok) flag_local = 2;
  }
  // every rank must take the same path: all-reduce the capability flag (min) through NCCL
  {
    int* f_d = nullptr;
    cudaMalloc(&f_d, sizeof(int));
    cudaMemcpyAsync(f_d, &flag_local, sizeof(int), cudaMemcpyHostToDevice, st);
    nccl_ok(c, g_nccl.AllReduce(f_d, f_d, 1, ncclInt32, ncclMin, d->comm_main, st), "tnt_dp_init");
    cudaMemcpyAsync(&flag_local, f_d, sizeof(int), cudaMemcpyDeviceToHost, st);
    cudaStreamSynchronize(st);
    cudaFree(f_d);
  }
  if (all_d) cudaFree(all_d);
  d->use_peer = flag_local >= 1;
  d->use_push = flag_local >= 2 && !getenv("TNT_DP_NCCL_GRADS");          // A/B switch for measurements (tools/gpu_dp.sh)
  if (d->use_peer) {
    cudaMalloc(&d->d_peers, sizeof(double*) * DP_MAX_WORLD);
    cudaMemcpy(d->d_peers, d->peers, sizeof(double*) * DP_MAX_WORLD, cudaMemcpyHostToDevice);
  }
  cudaGetLastError();
  if (c->fail_code) { int rc = c->fail_code; dp_destroy(c); return rc; }
  return TNT_OK;
}

extern "C" int tnt_dp_world(tnt_ctx* c) { return c ? dp_world(c) : 0; }
