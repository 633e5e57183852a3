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

struct RenderCounters {  // device-side counters of one rt_render call
  unsigned long long samples, rays, rays_ref, nodes, prims, seq_fallbacks, max_nodes, reserved;
};

}  // namespace rtc

struct rt_ctx;
namespace rtc {
void static_scratch_setup(int device);               // rt_render.cu: address of the static scratch array (rt_render_static)
int default_trav(const rt_ctx *ctx, float sah_cost);  // rt_render.cu: the scene's traversal (RT_TRAV overrides), read at commit
}  // namespace rtc

struct rt_ctx {
  int device = 0;
  bool has_scene = false, has_camera = false;
  rtc::DScene scene{};
  rtc::DCamera camera{};
  int n_sm = 148;
  int trav = 0;               // traversal of the render kernel for the committed scene (default_trav)
  bool opt_wavefront = false; // RT_STRATEGY=wavefront at commit: the default strategy of rt_render
  rtc::FlatItems flat{};  // box list of scenes with <= FLAT_MAX primitives (flat.n = 0: none)
  rt_camera camera_host{};
  // scene storage
  rtc::DevBuf nodes, nodes_ploc, recs, recs_seq, face_uv, face_uv_seq, mats, textures, tex_pixels;
  std::vector<rtc::DevBuf> tex_allocs;  // texel storage, one buffer per texture slot, reused across commits
  // build scratch
  rtc::DevBuf scratch[12];
  // render scratch
  rtc::DevBuf counters, accum_tmp, fixed_tmp, rgb_tmp, rgb8_tmp, rays_tmp, hits_tmp;
  rtc::DevBuf wf[16];  // wavefront path state
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr;
  uint32_t last_iterations = 0, last_launches = 0, last_items = 0;
  int last_chunk = 0;
  rt_build_stats build_stats{};
  // nccl (loaded lazily)
  void *nccl_comm = nullptr;
  // peer frame of the NVLink reduce (rt_peer.cu)
  void *peer = nullptr;
};
