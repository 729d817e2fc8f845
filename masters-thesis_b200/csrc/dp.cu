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

struct DpState {
  int rank = 0, world = 1;
  ncclComm_t comm_main = nullptr, comm_side = nullptr;
  double* mailbox = nullptr;                 // [DP_SLOTS][world][DP_ROW] on this rank
  double* peers[DP_MAX_WORLD] = {};          // every rank's mailbox as seen from here (peers[rank] == mailbox)
  double** d_peers = nullptr;                // device copy of the table
  unsigned long long seq = 0;                // calls made so far (identical on every rank: program order)
  int use_peer = 0;
};

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
                  int n, unsigned* __restrict__ abort_flag) {
  __shared__ int s_ok;
  const int tid = threadIdx.x;
  const size_t slot = (size_t)(seq % DP_SLOTS) * world * DP_ROW;
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

// ---- public to the rest of the library ----------------------------------------------------------------------------
int dp_world(tnt_ctx* c) { return c->dp ? c->dp->world : 1; }
int dp_rank(tnt_ctx* c) { return c->dp ? c->dp->rank : 0; }
const char* dp_small_kind(tnt_ctx* c) { return !c->dp ? "none" : (c->dp->use_peer ? "peer-memory one-shot" : "nccl"); }

// critical-path reductions of <= 128 doubles (BatchNorm sums)
void dp_allreduce_small(tnt_ctx* c, double* buf, int n, cudaStream_t st) {
  DpState* d = c->dp;
  if (!d || d->world == 1) return;
  if (d->use_peer && n <= DP_PAYLOAD) {
    dp_oneshot_kernel<<<1, DP_PAYLOAD, 0, st>>>(d->d_peers, d->rank, d->world, d->seq, buf, n, pd_abort_flag(c));
    d->seq += 1;
    c->launches += 1;
    return;
  }
  nccl_ok(c, g_nccl.AllReduce(buf, buf, (size_t)n, ncclFloat64, ncclSum, d->comm_main, st), "dp_allreduce_small");
}

// lane 0: main-stream traffic (encoder bucket), lane 1: side-stream traffic (decoder buckets, metrics)
void dp_allreduce(tnt_ctx* c, void* buf, int64_t n, int is_double, int lane, cudaStream_t st) {
  DpState* d = c->dp;
  if (!d || d->world == 1 || n <= 0) return;
  nccl_ok(c, g_nccl.AllReduce(buf, buf, (size_t)n, is_double ? ncclFloat64 : ncclFloat32, ncclSum,
                              lane ? d->comm_side : d->comm_main, st), "dp_allreduce");
}

void dp_destroy(tnt_ctx* c) {
  DpState* d = c->dp;
  if (!d) return;
  cudaDeviceSynchronize();
  for (int p = 0; p < d->world; ++p)
    if (p != d->rank && d->peers[p]) cudaIpcCloseMemHandle(d->peers[p]);
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
  const size_t mb = sizeof(double) * DP_SLOTS * world * DP_ROW;
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
  d->use_peer = flag_local;
  if (d->use_peer) {
    cudaMalloc(&d->d_peers, sizeof(double*) * DP_MAX_WORLD);
    cudaMemcpy(d->d_peers, d->peers, sizeof(double*) * DP_MAX_WORLD, cudaMemcpyHostToDevice);
  }
  cudaGetLastError();
  if (c->fail_code) { int rc = c->fail_code; dp_destroy(c); return rc; }
  return TNT_OK;
}

extern "C" int tnt_dp_world(tnt_ctx* c) { return c ? dp_world(c) : 0; }
