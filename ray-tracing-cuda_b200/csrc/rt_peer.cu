// rt_peer.cu — the image reduce of DistributedMain as ONE kernel over NVLink peer memory.
//
// Replaces D2H + MPI_Reduce + the host post-processing loop of GatherImageData (utils.cu:115-130, 232-241).
// Every rank renders into a "peer frame" — one cudaMalloc'ed block [exact sums | float sums | resolved image | flags]
// that the other ranks map through CUDA IPC.  After the render, rank r sums slice r of the image over the frames
// of ALL ranks (loads over NVLink), resolves it (/spp, clamp, sqrt) and stores sums + float image of that slice
// straight into the root's frame: a reduce-scatter, the resolve and the gather in one pass, no staging buffer, no
// NCCL call.  Integer sums, so the result is the one-GPU image bit for bit (rtcore.h, summation contract).
//
// Ordering between the processes is two counters per frame, written with system-scope atomics:
//   ready: +1 from every rank once its render is complete   (a rank reads peers only after ready == world * epoch)
//   done : +1 from every rank once its slice is stored      (a rank returns / clears only after done == world * epoch)
// Every wait is bounded (RT_PEER_TIMEOUT_MS, default 5000): a rank that never arrives sets the frame's error word
// instead of hanging the GPU.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "../../include/rtcore.h"
#include "rt_internal.h"

namespace rtc {

enum { PF_READY = 0, PF_DONE = 1, PF_ERROR = 2, PF_BLOCKS = 3, PF_WORDS = 32 };
constexpr int PEER_MAX = 16;

struct PeerFrame {  // device addresses inside one rank's block
  long long *sums;
  float *accum, *rgb;
  unsigned *flags;
};
struct PeerSet {
  PeerFrame f[PEER_MAX];
  int world, rank, root;
};

struct PeerState {
  void *local = nullptr;
  void *mapped[PEER_MAX] = {};
  PeerSet set{};
  int height = 0, width = 0;
  bool opened = false;
  unsigned epoch = 0;
};

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }
static size_t frame_bytes(size_t n) { return align256(n * 8) + 2 * align256(n * 4) + PF_WORDS * 4; }
static PeerFrame frame_at(void *base, size_t n) {
  char *b = static_cast<char *>(base);
  PeerFrame f;
  f.sums = reinterpret_cast<long long *>(b);
  f.accum = reinterpret_cast<float *>(b + align256(n * 8));
  f.rgb = reinterpret_cast<float *>(b + align256(n * 8) + align256(n * 4));
  f.flags = reinterpret_cast<unsigned *>(b + align256(n * 8) + 2 * align256(n * 4));
  return f;
}

__device__ __forceinline__ unsigned long long now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned load_acquire_sys(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// Spin until *flag has reached `target` (wrap-safe); false after timeout_ns.
__device__ bool wait_for(const unsigned *flag, unsigned target, unsigned long long timeout_ns) {
  const unsigned long long t0 = now_ns();
  while ((int)(load_acquire_sys(flag) - target) < 0) {
    if (now_ns() - t0 > timeout_ns) return false;
    __nanosleep(200);
  }
  return true;
}

// "My sums are complete": one increment of `ready` in every rank's frame (own included).
__global__ void k_peer_ready(PeerSet ps) {
  int p = threadIdx.x;
  if (p >= ps.world) return;
  __threadfence_system();
  atomicAdd_system(ps.f[p].flags + PF_READY, 1u);
}

// Slice `rank` of the image: sum over all frames, resolve, store into the root's frame.
__global__ void __launch_bounds__(256) k_peer_reduce_resolve(PeerSet ps, long long n, int spp_total, unsigned target,
                                                             unsigned long long timeout_ns) {
  __shared__ int ok;
  const PeerFrame me = ps.f[ps.rank];
  if (threadIdx.x == 0) {
    ok = wait_for(me.flags + PF_READY, target, timeout_ns) ? 1 : 0;
    if (!ok) atomicExch(me.flags + PF_ERROR, 1u);
  }
  __syncthreads();
  const long long lo = n * ps.rank / ps.world, hi = n * (ps.rank + 1) / ps.world;
  const PeerFrame root = ps.f[ps.root];
  for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi;
       i += (long long)gridDim.x * blockDim.x) {
    long long s = 0;
#pragma unroll 4
    for (int p = 0; p < ps.world; p++) s += ps.f[p].sums[i];
    const float sum = from_fixed(s);
    float c = sum / float(spp_total);
    c = fminf(fmaxf(c, 0.f), 1.f);
    c = sqrtf(c);
    root.sums[i] = s;
    root.accum[i] = sum;
    root.rgb[i] = c;
  }
  // last block out tells every rank that this slice is in place
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned arrived = atomicAdd(me.flags + PF_BLOCKS, 1u);
    ok = arrived == gridDim.x - 1;
  }
  __syncthreads();
  if (ok) {
    if (threadIdx.x == 0) me.flags[PF_BLOCKS] = 0;
    if ((int)threadIdx.x < ps.world) {
      __threadfence_system();
      atomicAdd_system(ps.f[threadIdx.x].flags + PF_DONE, 1u);
    }
  }
}

// Returns once every rank's slice is stored (so the root may read its frame and anyone may clear its sums).
__global__ void k_peer_done(PeerSet ps, unsigned target, unsigned long long timeout_ns) {
  const PeerFrame me = ps.f[ps.rank];
  if (!wait_for(me.flags + PF_DONE, target, timeout_ns)) atomicExch(me.flags + PF_ERROR, 1u);
}

}  // namespace rtc

using namespace rtc;

struct rt_peer_state : PeerState {};

namespace {
PeerState *state_of(rt_ctx *ctx) { return static_cast<PeerState *>(ctx->peer); }
}  // namespace

extern "C" int rt_peer_frame_destroy(rt_ctx *ctx) {
  if (!ctx) {
    set_error("rt_peer_frame_destroy: null context");
    return RT_ERR_INVALID;
  }
  PeerState *st = state_of(ctx);
  if (!st) return RT_OK;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  for (int p = 0; p < PEER_MAX; p++)
    if (st->mapped[p]) cudaIpcCloseMemHandle(st->mapped[p]);
  if (st->local) cudaFree(st->local);
  delete static_cast<rt_peer_state *>(st);
  ctx->peer = nullptr;
  return RT_OK;
}

extern "C" int rt_peer_frame_create(rt_ctx *ctx, int height, int width, char out_handle[RT_PEER_HANDLE_BYTES]) {
  if (!ctx || height <= 0 || width <= 0 || !out_handle) {
    set_error("rt_peer_frame_create: bad argument");
    return RT_ERR_INVALID;
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == RT_PEER_HANDLE_BYTES, "handle size");
  int rc = rt_peer_frame_destroy(ctx);
  if (rc) return rc;
  RT_CUDA(cudaSetDevice(ctx->device));
  rt_peer_state *st = new rt_peer_state();
  const size_t n = (size_t)height * width * 3;
  cudaError_t e = cudaMalloc(&st->local, frame_bytes(n));
  if (e == cudaSuccess) e = cudaMemset(st->local, 0, frame_bytes(n));
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, st->local);
  if (e != cudaSuccess) {
    if (st->local) cudaFree(st->local);
    delete st;
    return cuda_fail(e, "rt_peer_frame_create");
  }
  memcpy(out_handle, &h, sizeof(h));
  st->height = height;
  st->width = width;
  st->set.world = 1;
  st->set.rank = 0;
  st->set.f[0] = frame_at(st->local, n);
  ctx->peer = st;
  return RT_OK;
}

extern "C" int rt_peer_frame_buffers(rt_ctx *ctx, int64_t **d_accum_fixed, float **d_accum_rgb, float **d_rgb) {
  PeerState *st = ctx ? state_of(ctx) : nullptr;
  if (!st) {
    set_error("rt_peer_frame_buffers: no peer frame (call rt_peer_frame_create)");
    return RT_ERR_INVALID;
  }
  PeerFrame f = frame_at(st->local, (size_t)st->height * st->width * 3);
  if (d_accum_fixed) *d_accum_fixed = reinterpret_cast<int64_t *>(f.sums);
  if (d_accum_rgb) *d_accum_rgb = f.accum;
  if (d_rgb) *d_rgb = f.rgb;
  return RT_OK;
}

extern "C" int rt_peer_frame_open(rt_ctx *ctx, int world_size, int rank,
                                  const char (*handles)[RT_PEER_HANDLE_BYTES]) {
  PeerState *st = ctx ? state_of(ctx) : nullptr;
  if (!st || world_size < 1 || world_size > PEER_MAX || rank < 0 || rank >= world_size ||
      (world_size > 1 && !handles)) {
    set_error("rt_peer_frame_open: bad argument (1 <= world_size <= 16, frame created first)");
    return RT_ERR_INVALID;
  }
  if (st->opened) {
    set_error("rt_peer_frame_open: frame already opened");
    return RT_ERR_INVALID;
  }
  RT_CUDA(cudaSetDevice(ctx->device));
  const size_t n = (size_t)st->height * st->width * 3;
  st->set.world = world_size;
  st->set.rank = rank;
  for (int p = 0; p < world_size; p++) {
    void *base = st->local;
    if (p != rank) {
      cudaIpcMemHandle_t h;
      memcpy(&h, handles[p], sizeof(h));
      RT_CUDA(cudaIpcOpenMemHandle(&st->mapped[p], h, cudaIpcMemLazyEnablePeerAccess));
      base = st->mapped[p];
    }
    st->set.f[p] = frame_at(base, n);
  }
  st->opened = true;
  return RT_OK;
}

extern "C" int rt_reduce_resolve_peers(rt_ctx *ctx, int spp_total, int root, void *stream) {
  PeerState *st = ctx ? state_of(ctx) : nullptr;
  if (!st || !st->opened || spp_total <= 0 || root < 0 || root >= st->set.world) {
    set_error("rt_reduce_resolve_peers: bad argument (frame created and opened first)");
    return RT_ERR_INVALID;
  }
  RT_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t s = (cudaStream_t)stream;
  const char *env = getenv("RT_PEER_TIMEOUT_MS");
  const unsigned long long timeout_ns = (unsigned long long)(env ? std::max(1, atoi(env)) : 5000) * 1000000ull;
  const long long n = (long long)st->height * st->width * 3;
  PeerSet ps = st->set;
  ps.root = root;
  st->epoch++;
  const unsigned target = st->epoch * (unsigned)ps.world;
  k_peer_ready<<<1, 32, 0, s>>>(ps);
  const long long slice = (n + ps.world - 1) / ps.world;
  int blocks = (int)std::min<long long>((slice + 255) / 256, (long long)ctx->n_sm * 8);
  k_peer_reduce_resolve<<<std::max(blocks, 1), 256, 0, s>>>(ps, n, spp_total, target, timeout_ns);
  k_peer_done<<<1, 1, 0, s>>>(ps, target, timeout_ns);
  RT_CUDA(cudaGetLastError());
  return RT_OK;
}

// Synchronises `stream` and reports whether any wait of this frame timed out since the last check.
extern "C" int rt_peer_frame_check(rt_ctx *ctx, void *stream) {
  PeerState *st = ctx ? state_of(ctx) : nullptr;
  if (!st) {
    set_error("rt_peer_frame_check: no peer frame");
    return RT_ERR_INVALID;
  }
  RT_CUDA(cudaSetDevice(ctx->device));
  unsigned err = 0;
  unsigned *flag = st->set.f[st->set.rank].flags + PF_ERROR;
  RT_CUDA(cudaMemcpyAsync(&err, flag, 4, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  RT_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  if (err) {
    RT_CUDA(cudaMemsetAsync(flag, 0, 4, (cudaStream_t)stream));
    set_error("rt_reduce_resolve_peers: a rank did not arrive within RT_PEER_TIMEOUT_MS");
    return RT_ERR_CUDA;
  }
  return RT_OK;
}
