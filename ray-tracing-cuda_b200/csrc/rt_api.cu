// rt_api.cu — context, camera, workload split, NCCL reduce and error plumbing of librtcore.so.
#include <dlfcn.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "rt_internal.h"

namespace rtc {
static thread_local std::string g_err;
void set_error(const std::string &msg) { g_err = msg; }
int cuda_fail(cudaError_t e, const char *what) {
  g_err = std::string("CUDA error: ") + cudaGetErrorString(e) + " in " + what;
  cudaGetLastError();
  return RT_ERR_CUDA;
}
}  // namespace rtc
using namespace rtc;

extern "C" int rt_abi_version(void) { return RT_ABI_VERSION; }
extern "C" const char *rt_last_error_string(void) { return g_err.c_str(); }

extern "C" int rt_ctx_destroy(rt_ctx *c);
extern "C" int rt_nccl_destroy(rt_ctx *ctx);
extern "C" int rt_peer_frame_destroy(rt_ctx *ctx);

extern "C" int rt_ctx_create(int device, rt_ctx **out) {
  if (!out) {
    set_error("rt_ctx_create: null out pointer");
    return RT_ERR_INVALID;
  }
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    cudaGetLastError();
    set_error(std::string("rt_ctx_create: no usable CUDA device (") + cudaGetErrorString(e) +
              "); this library has no CPU fallback");
    return RT_ERR_CUDA;
  }
  if (device < 0 || device >= n) {
    set_error("rt_ctx_create: device index out of range");
    return RT_ERR_INVALID;
  }
  RT_CUDA(cudaSetDevice(device));
  static_scratch_setup(device);  // once per device, with it current (rt_render_static)
  rt_ctx *c = new rt_ctx();
  c->device = device;
  cudaError_t ee = cudaEventCreate(&c->ev0);
  if (ee == cudaSuccess) ee = cudaEventCreate(&c->ev1);
  if (ee == cudaSuccess) ee = cudaEventCreate(&c->ev2);
  if (ee == cudaSuccess) ee = cudaDeviceGetAttribute(&c->n_sm, cudaDevAttrMultiProcessorCount, device);
  if (ee != cudaSuccess) {  // nothing half-built survives a failed create
    rt_ctx_destroy(c);
    return cuda_fail(ee, "rt_ctx_create");
  }
  *out = c;
  return RT_OK;
}

extern "C" int rt_ctx_destroy(rt_ctx *c) {
  if (!c) return RT_OK;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  rt_nccl_destroy(c);  // a caller that forgot it does not leak the communicator
  rt_peer_frame_destroy(c);
  DevBuf *bufs[] = {&c->nodes, &c->nodes_ploc, &c->recs, &c->recs_seq, &c->face_uv, &c->face_uv_seq, &c->mats, &c->textures,
                    &c->tex_pixels, &c->counters, &c->accum_tmp, &c->fixed_tmp, &c->rgb_tmp, &c->rgb8_tmp, &c->rays_tmp,
                    &c->hits_tmp};
  for (DevBuf *b : bufs) b->release();
  for (DevBuf &b : c->scratch) b.release();
  for (DevBuf &b : c->wf) b.release();
  for (DevBuf &b : c->tex_allocs) b.release();
  if (c->ev0) cudaEventDestroy(c->ev0);
  if (c->ev1) cudaEventDestroy(c->ev1);
  if (c->ev2) cudaEventDestroy(c->ev2);
  delete c;
  return RT_OK;
}

// ---- camera (camera.cu:6-47) ------------------------------------------------------------------------
extern "C" int rt_camera_set(rt_ctx *ctx, const rt_camera *cam) {
  if (!ctx || !cam) {
    set_error("rt_camera_set: null argument");
    return RT_ERR_INVALID;
  }
  auto v = [](const float *p) { return mk(p[0], p[1], p[2]); };
  DCamera &c = ctx->camera;
  c.position = v(cam->position);
  c.llc = v(cam->lower_left_corner);
  c.horizontal = v(cam->horizontal);
  c.vertical = v(cam->vertical);
  c.u = v(cam->u);
  c.v = v(cam->v);
  c.is_defocus = cam->is_defocus;
  c.lens_radius = (float)cam->lens_radius;  // CudaRandomFloat(0, lens_radius_) takes floats (camera.cu:75)
  ctx->camera_host = *cam;
  ctx->has_camera = true;
  return RT_OK;
}

// camera.cu:6-38; aperture < 0 selects the pinhole constructor (:24-38).  Host arithmetic in the
// reference's expression order (this file is compiled without FMA contraction).
extern "C" int rt_camera_look_at(const float position[3], const float look_at[3], const float up[3],
                                 double field_of_view, double aspect, double aperture, double focus_distance,
                                 rt_camera *out) {
  if (!position || !look_at || !up || !out) {
    set_error("rt_camera_look_at: null argument");
    return RT_ERR_INVALID;
  }
  memset(out, 0, sizeof(*out));
  bool defocus = aperture >= 0;
  V3 pos = mk(position[0], position[1], position[2]);
  V3 w = normalize(pos - mk(look_at[0], look_at[1], look_at[2]));
  V3 u = normalize(cross(mk(up[0], up[1], up[2]), w));
  V3 v = normalize(cross(w, u));
  double half_height = defocus ? focus_distance * std::tan(field_of_view / 2) : std::tan(field_of_view / 2);
  double half_width = aspect * half_height;
  V3 horizontal = u * static_cast<float>(2 * half_width);
  V3 vertical = v * static_cast<float>(2 * half_height);
  V3 llc = pos - w - u * static_cast<float>(half_width) - v * static_cast<float>(half_height);
  auto st = [](float *d, V3 a) {
    d[0] = a.x;
    d[1] = a.y;
    d[2] = a.z;
  };
  st(out->position, pos);
  st(out->lower_left_corner, llc);
  st(out->horizontal, horizontal);
  st(out->vertical, vertical);
  st(out->u, u);
  st(out->v, v);
  st(out->w, w);
  out->is_defocus = defocus ? 1 : 0;
  out->lens_radius = defocus ? aperture / 2 : -1;
  return RT_OK;
}

// ---- GetWorkload (utils.cu:111-113) ---------------------------------------------------------------------
extern "C" int rt_get_workload(int rank, int world_size, int spp, int *spp_begin) {
  if (world_size <= 0 || rank < 0 || rank >= world_size || spp < 0) {
    if (spp_begin) *spp_begin = 0;
    return 0;
  }
  int base = spp / world_size, rem = spp % world_size;
  if (spp_begin) *spp_begin = rank * base + (rank < rem ? rank : rem);
  return base + (rank < rem ? 1 : 0);
}

// ---- NCCL: one reduce of the float accumulation buffer (replaces D2H + MPI_Reduce, utils.cu:115-125) ----
struct Id128 {
  char b[128];
};
namespace {
struct Nccl {
  void *lib = nullptr;
  int (*GetUniqueId)(void *) = nullptr;
  int (*CommInitRank)(void **, int, /* ncclUniqueId by value */ Id128, int) = nullptr;
  int (*CommDestroy)(void *) = nullptr;
  int (*Reduce)(const void *, void *, size_t, int, int, int, void *, cudaStream_t) = nullptr;
  int (*AllReduce)(const void *, void *, size_t, int, int, void *, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(int) = nullptr;
};
}  // namespace
static Nccl g_nccl;
static std::mutex g_nccl_mu;

static int nccl_load() {
  std::lock_guard<std::mutex> lk(g_nccl_mu);
  if (g_nccl.lib) return RT_OK;
  const char *names[] = {getenv("RT_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  void *h = nullptr;
  for (const char *n : names) {
    if (!n) continue;
    h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (h) break;
  }
  if (!h) {
    set_error(std::string("NCCL library not found: ") + dlerror());
    return RT_ERR_NCCL;
  }
  g_nccl.GetUniqueId = (decltype(g_nccl.GetUniqueId))dlsym(h, "ncclGetUniqueId");
  g_nccl.CommInitRank = (decltype(g_nccl.CommInitRank))dlsym(h, "ncclCommInitRank");
  g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy))dlsym(h, "ncclCommDestroy");
  g_nccl.Reduce = (decltype(g_nccl.Reduce))dlsym(h, "ncclReduce");
  g_nccl.AllReduce = (decltype(g_nccl.AllReduce))dlsym(h, "ncclAllReduce");
  g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))dlsym(h, "ncclGetErrorString");
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.CommDestroy || !g_nccl.Reduce || !g_nccl.GetErrorString) {
    set_error("NCCL library lacks required symbols");
    dlclose(h);
    return RT_ERR_NCCL;
  }
  g_nccl.lib = h;
  return RT_OK;
}
static int nccl_fail(int r, const char *what) {
  set_error(std::string("NCCL error: ") + g_nccl.GetErrorString(r) + " in " + what);
  return RT_ERR_NCCL;
}

extern "C" int rt_nccl_unique_id(char out_id[128]) {
  if (!out_id) {
    set_error("rt_nccl_unique_id: null argument");
    return RT_ERR_INVALID;
  }
  int rc = nccl_load();
  if (rc) return rc;
  int r = g_nccl.GetUniqueId(out_id);
  if (r) return nccl_fail(r, "ncclGetUniqueId");
  return RT_OK;
}

extern "C" int rt_nccl_init(rt_ctx *ctx, const char id[128], int world_size, int rank) {
  if (!ctx || !id || world_size <= 0 || rank < 0 || rank >= world_size) {
    set_error("rt_nccl_init: bad argument");
    return RT_ERR_INVALID;
  }
  int rc = nccl_load();
  if (rc) return rc;
  RT_CUDA(cudaSetDevice(ctx->device));
  if (ctx->nccl_comm) {
    g_nccl.CommDestroy(ctx->nccl_comm);
    ctx->nccl_comm = nullptr;
  }
  Id128 uid;
  memcpy(uid.b, id, 128);
  int r = g_nccl.CommInitRank(&ctx->nccl_comm, world_size, uid, rank);
  if (r) return nccl_fail(r, "ncclCommInitRank");
  return RT_OK;
}

extern "C" int rt_nccl_destroy(rt_ctx *ctx) {
  if (ctx && ctx->nccl_comm && g_nccl.lib) {
    g_nccl.CommDestroy(ctx->nccl_comm);
    ctx->nccl_comm = nullptr;
  }
  return RT_OK;
}

extern "C" int rt_reduce(rt_ctx *ctx, void *nccl_comm, float *d_accum_rgb, int64_t n_floats, int root, void *stream) {
  if (!ctx || !d_accum_rgb || n_floats < 0) {
    set_error("rt_reduce: bad argument");
    return RT_ERR_INVALID;
  }
  void *comm = nccl_comm ? nccl_comm : ctx->nccl_comm;
  if (!comm) {
    set_error("rt_reduce: no NCCL communicator (pass one or call rt_nccl_init)");
    return RT_ERR_INVALID;
  }
  int rc = nccl_load();
  if (rc) return rc;
  RT_CUDA(cudaSetDevice(ctx->device));
  int r = g_nccl.Reduce(d_accum_rgb, d_accum_rgb, (size_t)n_floats, /*ncclFloat*/ 7, /*ncclSum*/ 0, root, comm,
                        (cudaStream_t)stream);
  if (r) return nccl_fail(r, "ncclReduce");
  return RT_OK;
}

// The same reduce on the fixed-point sums (ncclInt64): integer addition is exact, so the image of N ranks is the
// image one rank would have rendered, bit for bit, whatever the partition.
extern "C" int rt_reduce_fixed(rt_ctx *ctx, void *nccl_comm, int64_t *d_accum_fixed, int64_t n_values, int root,
                               void *stream) {
  if (!ctx || !d_accum_fixed || n_values < 0) {
    set_error("rt_reduce_fixed: bad argument");
    return RT_ERR_INVALID;
  }
  void *comm = nccl_comm ? nccl_comm : ctx->nccl_comm;
  if (!comm) {
    set_error("rt_reduce_fixed: no NCCL communicator (pass one or call rt_nccl_init)");
    return RT_ERR_INVALID;
  }
  int rc = nccl_load();
  if (rc) return rc;
  RT_CUDA(cudaSetDevice(ctx->device));
  int r = g_nccl.Reduce(d_accum_fixed, d_accum_fixed, (size_t)n_values, /*ncclInt64*/ 4, /*ncclSum*/ 0, root, comm,
                        (cudaStream_t)stream);
  if (r) return nccl_fail(r, "ncclReduce");
  return RT_OK;
}
