// rt_internal.h — host-side context of librtcore.so (not part of the ABI).
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "rt_device.cuh"
#include "rt_flat.cuh"

namespace rtc {

void set_error(const std::string &msg);
int cuda_fail(cudaError_t e, const char *what);

#define RT_CUDA(call)                                    \
  do {                                                   \
    cudaError_t e_ = (call);                             \
    if (e_ != cudaSuccess) return rtc::cuda_fail(e_, #call); \
  } while (0)

// A device allocation that grows but never shrinks (buffers are reused across commits/renders).
struct DevBuf {
  void *p = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    cudaError_t e = cudaMalloc(&p, bytes ? bytes : 16);
    if (e == cudaSuccess) cap = bytes;
    return e;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <typename T>
  T *as() const {
    return reinterpret_cast<T *>(p);
  }
};

// derived trees (rt_ctx::trees): built on demand by ensure_trees
enum : unsigned { RT_TREE_W8 = 1u, RT_TREE_W4 = 2u, RT_TREE_C8 = 4u, RT_TREE_C4 = 8u };

struct RenderCounters {  // device-side counters of one rt_render call
  unsigned long long samples, rays, rays_ref, nodes, prims, seq_fallbacks, sched_iters, sched_lanes;
};

}  // namespace rtc

struct rt_ctx;
namespace rtc {
int trav_variant(const rt_ctx *ctx);                  // rt_render.cu: RT_TRAV or the scene's default
unsigned trees_for_trav(int trav);                    // which derived trees a render variant walks
unsigned trees_for_mode(int mode);                    // ... and a rt_trace_closest mode
int ensure_trees(rt_ctx *ctx, unsigned need, cudaStream_t st);  // rt_build.cu
}  // namespace rtc

struct rt_ctx {
  int device = 0;
  bool has_scene = false, has_camera = false;
  rtc::DScene scene{};
  rtc::DCamera camera{};
  unsigned trees = 0;  // RT_TREE_* present for the committed scene
  int n_wide[2] = {0, 0};  // 8-wide / 4-wide node counts
  rtc::FlatItems flat{};  // box list of scenes with <= FLAT_MAX primitives (flat.n = 0: none)
  rt_camera camera_host{};
  // scene storage
  rtc::DevBuf nodes, wnodes, wnodes4, cnodes8, cnodes4, recs, recs_seq, face_uv, face_uv_seq, mats, textures, tex_pixels;
  std::vector<void *> tex_allocs;
  // build scratch
  rtc::DevBuf scratch[12];
  // render scratch
  rtc::DevBuf counters, accum_tmp, rgb_tmp, rgb8_tmp, rays_tmp, hits_tmp;
  rtc::DevBuf wf[16];  // wavefront path state
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr;
  // RT_RENDER_SPLIT: side streams of the sample slices, fork / join events, per-slice sums (created on first use)
  static constexpr int kMaxSlices = 32;
  cudaStream_t slice_st[kMaxSlices - 1] = {};
  cudaEvent_t slice_ev[kMaxSlices] = {};
  rtc::DevBuf slice_sums;
  // RT_RENDER_ORDERED: per-tile probe cost and the tile list sorted by it
  rtc::DevBuf tile_cost, tile_order;
  std::vector<uint32_t> tile_cost_host, tile_order_host;
  rt_build_stats build_stats{};
  // nccl (loaded lazily)
  void *nccl_comm = nullptr;
};
