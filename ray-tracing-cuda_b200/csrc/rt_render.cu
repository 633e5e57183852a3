// rt_render.cu — the render hot path: closest-hit queries, the fused persistent-path kernel, the
// wavefront pipeline (logic/shade -> extend) and the resolve kernel.
//
// Replaces, in the reference: PathTracing (ray_tracing.cu:56-85), Trace (ray_tracing.cu:12-54),
// HitableList::Hit (hitable_list.cu:7-25) and everything they call, plus the post-processing tail
// (ray_tracing.cu:78-83, utils.cu:126-129).
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <string>

#include "rt_internal.h"
#include "rt_walk.cuh"
#include "rt_cwide.cuh"

#ifndef RT_DEFAULT_TRAV
#define RT_DEFAULT_TRAV 0  // traversal loop of the parity hook and the wavefront extend kernel
#endif

namespace rtc {

struct RenderArgs {
  int32_t height, width;
  int32_t spp_begin, spp_end;
  int32_t max_depth;
  uint32_t k0, k1;
  int32_t tiles_x = 0;                   // RT_RENDER_ORDERED: tiles per image row ...
  const uint32_t *tile_order = nullptr;  // ... and the tile each block renders (heaviest first)
};

// rt_hit of a query result (what HitRecord carries, hitable.cuh:13-17)
__device__ __forceinline__ rt_hit make_rt_hit(const DScene &s, const RayQ &r, const Hit &h) {
  rt_hit o;
  if (h.rec == REC_NONE) {
    o.t = 0;
    o.obj = -1;
    o.sub = -1;
    o.u = o.v = 0.f;
    o.normal[0] = o.normal[1] = o.normal[2] = 0.f;
    o.material = -1;
  } else {
    Surface sf = surface_at(s, r, h, true);
    o.t = h.t;
    o.obj = sf.obj;
    o.sub = sf.sub;
    o.u = sf.u;
    o.v = sf.v;
    o.normal[0] = sf.normal.x;
    o.normal[1] = sf.normal.y;
    o.normal[2] = sf.normal.z;
    o.material = sf.material;
  }
  return o;
}

// Parity hook of the warp-synchronous traversals (every lane of a warp enters the query):
// mode 8 wide tree, per-lane record tests; 9 wide tree + cooperative record tests; 10 binary tree + cooperative;
// 11 4-wide tree, per-lane record tests.
__global__ void __launch_bounds__(128) k_trace_closest_warp(DScene s, const float *__restrict__ rays, int64_t n,
                                                            double t_from, double t_to, int mode, rt_hit *out) {
  __shared__ CoopWarp sm_all[4];
  CoopWarp &sm = sm_all[threadIdx.x >> 5];
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool active = i < n;
  RayQ r;
  r.o = r.d = mk(0.f, 0.f, 1.f);
  if (active) {
    r.o = mk(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]);
    r.d = mk(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]);
  }
  Hit h;
  if (mode == 8)
    h = world_hit_wide<false, false>(s, r, active, t_from, t_to, &sm, nullptr);
  else if (mode == 9)
    h = world_hit_wide<false, true>(s, r, active, t_from, t_to, &sm, nullptr);
  else if (mode == 11)
    h = world_hit_wide<false, false, 4>(s, r, active, t_from, t_to, &sm, nullptr);
  else
    h = world_hit_coop<false>(s, r, active, t_from, t_to, sm, nullptr);
  if (active) out[i] = make_rt_hit(s, r, h);
}

// Parity hook of the box-list traversal (mode 12; scenes with at most FLAT_MAX primitives).
__global__ void __launch_bounds__(128) k_trace_closest_flat(DScene s, FlatItems fl, const float *__restrict__ rays, int64_t n,
                                                            double t_from, double t_to, rt_hit *out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  RayQ r;
  r.o = mk(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]);
  r.d = mk(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]);
  Hit h = world_hit_flat<false>(s, fl, r, t_from, t_to, nullptr);
  out[i] = make_rt_hit(s, r, h);
}

// ---- parity hook: HitableList::Hit on explicit rays ------------------------------------------------
__global__ void __launch_bounds__(128) k_trace_closest(DScene s, const float *__restrict__ rays, int64_t n,
                                                       double t_from, double t_to, int mode, rt_hit *out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  RayQ r;
  r.o = mk(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]);
  r.d = mk(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]);
  Hit h;
  switch (mode) {
    case 1: h = world_hit_seq(s, r, t_from, t_to); break;
    case 2: h = world_hit_bvh<false, 0>(s, r, t_from, t_to, nullptr); break;
    case 3: h = world_hit_bvh<false, 1>(s, r, t_from, t_to, nullptr); break;
    case 4: h = world_hit_bvh<false, 2>(s, r, t_from, t_to, nullptr); break;
    case 5: h = world_hit_bvh<false, 3>(s, r, t_from, t_to, nullptr); break;
    case 6: h = world_hit_bvh<false, 4>(s, r, t_from, t_to, nullptr); break;
    case 7: h = world_hit_bvh<false, 5>(s, r, t_from, t_to, nullptr); break;
    case 13: h = world_hit_cwide<false, 4>(s, r, t_from, t_to, nullptr); break;
    case 14: h = world_hit_cwide<false, 8>(s, r, t_from, t_to, nullptr); break;
    default: h = world_hit_bvh<false, RT_DEFAULT_TRAV>(s, r, t_from, t_to, nullptr); break;
  }
  out[i] = make_rt_hit(s, r, h);
}

// Pixel-centre primary rays: ray_tracing.cu:68-74 with both jitters = 0.5.
__global__ void k_primary_rays(DCamera cam, int height, int width, float *rays) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= height * width) return;
  int i = idx / width, j = idx % width;
  double x = ((double)0.5f + double(j)) / double(width);
  double y = ((double)0.5f + double(height - i)) / double(height);
  x = 2 * x - 1;
  y = 2 * y - 1;
  RayQ r = camera_ray_at(cam, x, y, 0.5f * cam.lens_radius, 0.5f * cam.lens_radius);
  float *o = rays + 6 * (size_t)idx;
  o[0] = r.o.x;
  o[1] = r.o.y;
  o[2] = r.o.z;
  o[3] = r.d.x;
  o[4] = r.d.y;
  o[5] = r.d.z;
}

// One bounce of Trace (ray_tracing.cu:20-43) after the closest-hit query, in forward form:
// returns true when the path continues with `ray`/`T` updated, else adds the path's radiance.
__device__ __forceinline__ bool shade_hit(const DScene &s, RayQ &ray, const Hit &h, const PathKey &pk, int depth, V3 &T,
                                          V3 &color) {
  int mat_id;
  bool need_uv;
  {
    // material first: only textured materials need u, v
    if (h.rec == REC_SKY) {
      mat_id = s.sky_mat;
    } else {
      const float4 *recs = h.seq ? s.recs_seq : s.recs;
      mat_id = (int)(__float_as_uint(__ldg(recs + 3 * (size_t)h.rec + 2).w) >> KIND_BITS);
    }
  }
  DMaterial m;
  {
    const float4 *mp = reinterpret_cast<const float4 *>(s.mats + mat_id);
    float4 a = __ldg(mp), b = __ldg(mp + 1);
    m.color_fuzz = a;
    m.type = __float_as_int(b.x);
    m.texture = __float_as_int(b.y);
    m.eta_in = b.z;
    m.eta_out = b.w;
  }
  need_uv = material_needs_uv(m);
  float t = float(h.t);
  if (m.type == RT_MAT_SKY || m.type == RT_MAT_DIFFUSE_LIGHT) {
    // Scatter returns false (sky.cu:3-7, diffuse_light.cu:5-9): result = emitted
    float u = 0.f, v = 0.f;
    if (need_uv) {
      Surface sf = surface_at(s, ray, h, true);
      u = sf.u;
      v = sf.v;
    }
    V3 p = ray.o + t * ray.d;
    color = color + T * emit(s, m, u, v, p);
    return false;
  }
  Surface sf = surface_at(s, ray, h, need_uv);
  V3 att;
  RayQ out;
  if (!scatter(s, m, ray, t, sf, pk, depth, &att, &out)) return false;  // Emit() of these materials is 0
  T = T * att;
  ray = out;
  return true;
}

// shade_hit in two halves around a converged point (RT_TRAV=23: cooperative ball rejection, rt_device.cuh).
// shade_prepare: everything up to the Scatter call; returns true when a Scatter is pending (ShadePending filled,
// need_ball = Scatter will draw a ball sample), false when the path ended here (emission added).
struct ShadePending {
  DMaterial m;
  Surface sf;
  float t;
};
__device__ __forceinline__ bool shade_prepare(const DScene &s, const RayQ &ray, const Hit &h, const V3 &T, V3 &color,
                                              ShadePending &sp, bool &need_ball) {
  int mat_id;
  if (h.rec == REC_SKY) {
    mat_id = s.sky_mat;
  } else {
    const float4 *recs = h.seq ? s.recs_seq : s.recs;
    mat_id = (int)(__float_as_uint(__ldg(recs + 3 * (size_t)h.rec + 2).w) >> KIND_BITS);
  }
  DMaterial &m = sp.m;
  {
    const float4 *mp = reinterpret_cast<const float4 *>(s.mats + mat_id);
    float4 a = __ldg(mp), b = __ldg(mp + 1);
    m.color_fuzz = a;
    m.type = __float_as_int(b.x);
    m.texture = __float_as_int(b.y);
    m.eta_in = b.z;
    m.eta_out = b.w;
  }
  const bool need_uv = material_needs_uv(m);
  sp.t = float(h.t);
  if (m.type == RT_MAT_SKY || m.type == RT_MAT_DIFFUSE_LIGHT) {
    float u = 0.f, v = 0.f;
    if (need_uv) {
      Surface sf = surface_at(s, ray, h, true);
      u = sf.u;
      v = sf.v;
    }
    V3 p = ray.o + sp.t * ray.d;
    color = color + T * emit(s, m, u, v, p);
    return false;
  }
  sp.sf = surface_at(s, ray, h, need_uv);
  need_ball = scatter_needs_ball(m, ray, sp.sf);
  return true;
}
__device__ __forceinline__ bool shade_finish(const DScene &s, RayQ &ray, const ShadePending &sp, V3 ball, float ball_l,
                                             V3 &T) {
  V3 att;
  RayQ out;
  if (!scatter_with(s, sp.m, ray, sp.t, sp.sf, [&](float *l) { *l = ball_l; return ball; }, &att, &out)) return false;
  T = T * att;
  ray = out;
  return true;
}

__device__ __forceinline__ void flush_counters(RenderCounters *ctr, unsigned long long samples,
                                               unsigned long long rays, unsigned long long rays_ref,
                                               unsigned long long nodes, unsigned long long prims) {
  for (int o = 16; o > 0; o >>= 1) {
    samples += __shfl_xor_sync(0xffffffffu, samples, o);
    rays += __shfl_xor_sync(0xffffffffu, rays, o);
    rays_ref += __shfl_xor_sync(0xffffffffu, rays_ref, o);
    nodes += __shfl_xor_sync(0xffffffffu, nodes, o);
    prims += __shfl_xor_sync(0xffffffffu, prims, o);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&ctr->samples, samples);
    atomicAdd(&ctr->rays, rays);
    atomicAdd(&ctr->rays_ref, rays_ref);
    if (nodes) atomicAdd(&ctr->nodes, nodes);
    if (prims) atomicAdd(&ctr->prims, prims);
  }
}

// Out-of-line forms of the rare steps (streaming kernel; fused kernel at 64 registers): the node step is the hot loop of a long query
// and must not carry the registers (or the spills) of shading, camera sampling and the record tests.  Everything
// goes in and out BY VALUE: a reference parameter would pin the traversal state of the hot loop in local memory.
struct ShadeRet {
  RayQ ray;
  V3 T, color;
  bool alive;
};
static __device__ __noinline__ ShadeRet shade_hit_call(const DScene &s, RayQ ray, Hit h, PathKey pk, int depth, V3 T,
                                                       V3 color) {
  ShadeRet r;
  r.alive = shade_hit(s, ray, h, pk, depth, T, color);
  r.ray = ray;
  r.T = T;
  r.color = color;
  return r;
}
static __device__ __noinline__ RayQ camera_sample_call(const DCamera &cam, int i, int j, int height, int width,
                                                       PathKey pk) {
  return camera_sample(cam, i, j, height, width, pk);
}
struct LeafRet {
  Trav tv;
  int state;
};
static __device__ __noinline__ LeafRet trav_leaf_step_call(const DScene &s, Trav tv, Range rg, const uint2 *stack) {
  LeafRet r;
  r.state = trav_leaf_step(s, tv, rg, stack);
  r.tv = tv;
  return r;
}
static __device__ __noinline__ Hit trav_finish_call(const DScene &s, Trav tv, Range rg) { return trav_finish(s, tv, rg); }

// ---- fused kernel: one thread per pixel, paths regenerated in registers ---------------------------
// Each lane runs its pixel's samples in increasing order and starts the next sample the moment the
// current path ends, so a warp stays full until its pixels run out of samples.  No queue traffic.
// ORDERED (RT_RENDER_ORDERED): a 1-D grid whose block b renders tile a.tile_order[b] (heaviest tile first).
template <bool STATS, int TRAV, int MINB = 5, bool ORDERED = false>
__global__ void __launch_bounds__(128, MINB) k_render_fused(DScene s, DCamera cam, RenderArgs a, float *__restrict__ accum,
                                                      RenderCounters *ctr,
                                                      typename FlatParamT<TRAV == 20 || TRAV == 23>::type fl = {}) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int bx = blockIdx.x, by = blockIdx.y;
  if constexpr (ORDERED) {
    const uint32_t tile = a.tile_order[blockIdx.x];
    bx = (int)(tile % (uint32_t)a.tiles_x);
    by = (int)(tile / (uint32_t)a.tiles_x);
  }
  const int j = bx * 16 + (warp & 1) * 8 + (lane & 7);
  const int i = by * 8 + (warp >> 1) * 4 + (lane >> 3);
  const bool inside = i < a.height && j < a.width;
  const int idx = i * a.width + j;
  PathKey pk;
  pk.pixel = (uint32_t)idx;
  pk.k0 = a.k0;
  pk.k1 = a.k1;
  int k = inside ? a.spp_begin : a.spp_end;
  V3 color = mk(0.f, 0.f, 0.f), T = mk(1.f, 1.f, 1.f);
  RayQ ray;
  ray.o = ray.d = mk(0.f, 0.f, 0.f);
  int depth = 0;
  bool need_new = true;
  unsigned long long n_samples = 0, n_rays = 0, n_rays_ref = 0;
  TravStats ts;
  ts.nodes = ts.prims = 0;
  unsigned long long n_nodes = 0, n_prims = 0;
  for (;;) {
    if (need_new) {
      if (k >= a.spp_end) break;
      pk.sample = (uint32_t)k;
      if constexpr (MINB >= 8)  // 64 registers: keep camera sampling and shading out of the traversal loop's budget
        ray = camera_sample_call(cam, i, j, a.height, a.width, pk);
      else
        ray = camera_sample(cam, i, j, a.height, a.width, pk);
      T = mk(1.f, 1.f, 1.f);
      depth = 0;
      need_new = false;
      n_samples++;
    }
    n_rays_ref++;
    bool alive = depth < a.max_depth;  // ray_tracing.cu:23: the query at depth == limit yields 0
    ShadePending spend;
    bool pending = false, need_ball = false;
    if (alive) {
      Hit h;
      if constexpr (TRAV == 21 || TRAV == 22)  // if-if over the compact 4-/8-wide tree (rt_cwide.cuh)
        h = world_hit_cwide<STATS, TRAV == 21 ? 4 : 8>(s, ray, 1e-3, (double)CUDART_INF_F, &ts);
      else if constexpr (TRAV == 20 || TRAV == 23)  // box list (<= 32 primitives)
        h = world_hit_flat<STATS>(s, fl, ray, 1e-3, (double)CUDART_INF_F, &ts);
      else if constexpr (TRAV >= 18)  // while-while over the 4-/8-wide tree, per lane
        h = world_hit_wide<STATS, false, TRAV == 18 ? 4 : 8>(s, ray, true, 1e-3, (double)CUDART_INF_F, nullptr, &ts);
      else if constexpr (TRAV >= 16)
        h = world_hit_wide_ifif<STATS, TRAV == 16 ? 4 : 8>(s, ray, 1e-3, (double)CUDART_INF_F, &ts);
      else
        h = world_hit_bvh<STATS, TRAV>(s, ray, 1e-3, (double)CUDART_INF_F, &ts);
      n_rays++;
      if (STATS) {
        n_nodes += ts.nodes;
        n_prims += ts.prims;
        ts.nodes = ts.prims = 0;
      }
      if constexpr (TRAV == 23) {  // shading in two halves around the converged point below
        alive = false;
        if (h.rec != REC_NONE) pending = shade_prepare(s, ray, h, T, color, spend, need_ball);
      } else if constexpr (MINB >= 8) {
        alive = false;
        if (h.rec != REC_NONE) {
          ShadeRet sr = shade_hit_call(s, ray, h, pk, depth, T, color);
          alive = sr.alive;
          ray = sr.ray;
          T = sr.T;
          color = sr.color;
        }
      } else {
        alive = h.rec != REC_NONE && shade_hit(s, ray, h, pk, depth, T, color);
      }
    }
    if constexpr (TRAV == 23) {
      // every lane still rendering passes here converged: lanes without a pending ball sample try later
      // rejection iterations of their neighbours (ball_rejection_coop)
      float bl;
      const V3 br = ball_rejection_coop(need_ball, pk, depth, &bl);
      if (pending) alive = shade_finish(s, ray, spend, br, bl, T);
    }
    if (alive) {
      depth++;
    } else {
      k++;
      need_new = true;
    }
  }
  if (inside) {
    accum[3 * (size_t)idx + 0] += color.x;
    accum[3 * (size_t)idx + 1] += color.y;
    accum[3 * (size_t)idx + 2] += color.z;
  }
  flush_counters(ctr, n_samples, n_rays, n_rays_ref, n_nodes, n_prims);
}


// RT_RENDER_ORDERED probe: one primary ray per pixel (the first sample of the call) through the 8-wide tree; the
// cost of a 16x8 tile is the node steps + record tests of its 128 rays (a sky or ground tile: a few steps; a tile
// on the mesh: tens per ray, and its paths go on to bounce).  Only the ORDER of the blocks depends on it.
__global__ void __launch_bounds__(128) k_tile_cost(DScene s, DCamera cam, RenderArgs a, uint32_t *__restrict__ cost) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int j = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
  const int i = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
  uint32_t c = 0;
  if (i < a.height && j < a.width) {
    PathKey pk;
    pk.pixel = (uint32_t)(i * a.width + j);
    pk.k0 = a.k0;
    pk.k1 = a.k1;
    pk.sample = (uint32_t)a.spp_begin;
    const RayQ ray = camera_sample(cam, i, j, a.height, a.width, pk);
    TravStats ts;
    ts.nodes = ts.prims = 0;
    const Hit h = world_hit_wide_ifif<true, 8>(s, ray, 1e-3, (double)CUDART_INF_F, &ts);
    c = (uint32_t)ts.nodes + (uint32_t)ts.prims + (h.rec != REC_NONE && h.rec != REC_SKY ? 64u : 0u);
  }
  c = __reduce_add_sync(0xffffffffu, c);
  if (lane == 0 && c) atomicAdd(&cost[blockIdx.y * gridDim.x + blockIdx.x], c);
}

// ---- warp-synchronous kernel: the fused kernel over the wide tree / with cooperative record tests ----
// Same path bookkeeping as k_render_fused (one lane per pixel, samples in increasing order, a path is
// regenerated in registers when the previous one ends), but every lane of the warp enters every
// closest-hit query — lanes whose path is at the depth limit or whose pixel is finished pass
// active = false — so the query can pool the warp's record tests (rt_coop.cuh).
//   MODE 0: binary tree + cooperative record tests      (RT_TRAV=9)
//   MODE 1: 8-wide tree, per-lane record tests          (RT_TRAV=10)
//   MODE 2: 8-wide tree + cooperative record tests      (RT_TRAV=11)
//   MODE 3: 4-wide tree, per-lane record tests          (RT_TRAV=15)
// Image and counters are bit-identical to k_render_fused.
template <bool STATS, int MODE, int MINB>
__global__ void __launch_bounds__(128, MINB) k_render_warp(DScene s, DCamera cam, RenderArgs a,
                                                           float *__restrict__ accum, RenderCounters *ctr) {
  constexpr unsigned FULL = 0xffffffffu;
  __shared__ CoopWarp sm_all[(MODE == 1 || MODE == 3) ? 1 : 4];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  CoopWarp &sm = sm_all[(MODE == 1 || MODE == 3) ? 0 : warp];
  const int j = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
  const int i = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
  const bool inside = i < a.height && j < a.width;
  const int idx = i * a.width + j;
  PathKey pk;
  pk.pixel = (uint32_t)idx;
  pk.k0 = a.k0;
  pk.k1 = a.k1;
  int k = a.spp_begin;
  V3 color = mk(0.f, 0.f, 0.f), T = mk(1.f, 1.f, 1.f);
  RayQ ray;
  ray.o = mk(0.f, 0.f, 0.f);
  ray.d = mk(0.f, 0.f, 1.f);
  int depth = 0;
  bool need_new = true, finished = !inside;
  unsigned long long n_samples = 0, n_rays = 0, n_rays_ref = 0, n_nodes = 0, n_prims = 0;
  TravStats ts;
  ts.nodes = ts.prims = 0;
  for (;;) {
    if (need_new && !finished) {
      if (k >= a.spp_end) {
        finished = true;
      } else {
        pk.sample = (uint32_t)k;
        ray = camera_sample(cam, i, j, a.height, a.width, pk);
        T = mk(1.f, 1.f, 1.f);
        depth = 0;
        need_new = false;
        n_samples++;
      }
    }
    if (__all_sync(FULL, finished)) break;
    const bool query = !finished && depth < a.max_depth;  // ray_tracing.cu:23: the query at depth == limit yields 0
    if (!finished) n_rays_ref++;
    Hit h;
    if (MODE == 0)
      h = world_hit_coop<STATS>(s, ray, query, 1e-3, (double)CUDART_INF_F, sm, &ts);
    else if (MODE == 3)
      h = world_hit_wide<STATS, false, 4>(s, ray, query, 1e-3, (double)CUDART_INF_F, &sm, &ts);
    else
      h = world_hit_wide<STATS, MODE == 2>(s, ray, query, 1e-3, (double)CUDART_INF_F, &sm, &ts);
    bool alive = false;
    if (query) {
      n_rays++;
      alive = h.rec != REC_NONE && shade_hit(s, ray, h, pk, depth, T, color);
    }
    if (!finished) {
      if (alive) {
        depth++;
      } else {
        k++;
        need_new = true;
      }
    }
  }
  if (STATS) {
    n_nodes = ts.nodes;
    n_prims = ts.prims;
  }
  if (inside) {
    accum[3 * (size_t)idx + 0] += color.x;
    accum[3 * (size_t)idx + 1] += color.y;
    accum[3 * (size_t)idx + 2] += color.z;
  }
  flush_counters(ctr, n_samples, n_rays, n_rays_ref, n_nodes, n_prims);
}

// ---- persistent kernel: no reconvergence at the query boundary -----------------------------------------
// One loop iteration = one traversal ROUND of every lane (walk to a leaf, test the leaf).  A lane whose query
// ended in the previous round first shades its hit, scatters or regenerates, and starts its next query —
// then joins the round with the others.  Lanes never wait for the longest query of the warp; only the
// shade block runs with the few lanes that need it.  Samples of a pixel still run in increasing order in
// one lane: image and counters are bit-identical to k_render_fused.   (RT_TRAV=12 binary tree, 13 8-wide tree)
template <bool STATS, class W, int MINB>
__global__ void __launch_bounds__(128, MINB) k_render_persist(DScene s, DCamera cam, RenderArgs a,
                                                              float *__restrict__ accum, RenderCounters *ctr) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int j = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
  const int i = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
  const bool inside = i < a.height && j < a.width;
  const int idx = i * a.width + j;
  PathKey pk;
  pk.pixel = (uint32_t)idx;
  pk.sample = 0;
  pk.k0 = a.k0;
  pk.k1 = a.k1;
  int k = a.spp_begin;
  V3 color = mk(0.f, 0.f, 0.f), T = mk(1.f, 1.f, 1.f);
  RayQ ray;
  ray.o = mk(0.f, 0.f, 0.f);
  ray.d = mk(0.f, 0.f, 1.f);
  int depth = 0;
  bool need_new = true, finished = !inside, have_query = false, ambiguous = false;
  Range rg;
  rg.t_from = 1e-3;
  rg.t_to = (double)CUDART_INF_F;
  rg.tf_f = float_ge(1e-3);
  Best b;
  best_init(b, rg.t_to);
  W w;
  w.begin(s, ray, rg);
  w.done = true;
  uint2 stack[W::kStack];
  unsigned long long n_samples = 0, n_rays = 0, n_rays_ref = 0;
  TravStats ts;
  ts.nodes = ts.prims = 0;
  for (;;) {
    if (w.done) {  // between two queries of this lane
      if (have_query) {
        Hit h;
        if (ambiguous || w.overflow) {
          h = world_hit_seq(s, ray, rg.t_from, rg.t_to);
        } else {
          h.t = b.rec == REC_NONE ? 0.0 : b.t;
          h.rec = b.rec;
          h.seq = 0;
        }
        n_rays++;
        const bool alive = h.rec != REC_NONE && shade_hit(s, ray, h, pk, depth, T, color);
        if (alive) {
          depth++;
        } else {
          k++;
          need_new = true;
        }
        have_query = false;
      }
      while (!finished) {
        if (need_new) {
          if (k >= a.spp_end) {
            finished = true;
            break;
          }
          pk.sample = (uint32_t)k;
          ray = camera_sample(cam, i, j, a.height, a.width, pk);
          T = mk(1.f, 1.f, 1.f);
          depth = 0;
          need_new = false;
          n_samples++;
        }
        n_rays_ref++;
        if (depth < a.max_depth) {  // ray_tracing.cu:23: the query at depth == limit yields 0
          have_query = true;
          ambiguous = false;
          query_begin(s, rg, b);
          w.begin(s, ray, rg);
          break;
        }
        k++;
        need_new = true;
      }
    }
    if (finished) break;
    int first = 0;
    const int n = w.template walk<STATS>(s, ray, b, stack, &first, &ts);
    if (STATS) ts.prims += n;
    for (int q = first, end = first + n; q < end;) q += rec_test(s.recs, q, ray, rg, b, ambiguous);
  }
  if (inside) {
    accum[3 * (size_t)idx + 0] += color.x;
    accum[3 * (size_t)idx + 1] += color.y;
    accum[3 * (size_t)idx + 2] += color.z;
  }
  flush_counters(ctr, n_samples, n_rays, n_rays_ref, STATS ? ts.nodes : 0, STATS ? ts.prims : 0);
}

// ---- streaming kernel: single steps, deferred minority work -----------------------------------------------
// For scenes whose queries are long (hundreds of node steps: the 1M-10M triangle soups) the per-lane loops leave
// 5 of 32 lanes active (profiles/r01_j_soup10m_fused_ncu_summary.txt): the warp reconverges after every query and
// waits for its longest one.  Here one loop iteration = one NODE step for every lane that has one pending; lanes
// whose next step is a record test (LEAF) or the work between two queries (SHADE: finish the query, shade,
// scatter or regenerate, begin the next query) wait until `leaf_min` / `shade_min` lanes of the warp wait for
// the same kind — or nothing else is left to do — and then run it together.  Unlike the vote kernel (one kind
// per iteration, the minority idles) the majority kind never stops, so a waiting lane costs one idle lane for a
// few iterations instead of stalling the warp.  Same step functions as the vote kernel: identical hits, image
// and counters.   (RT_TRAV=14; RT_STREAM_LEAF / RT_STREAM_SHADE set the two thresholds)
template <bool STATS, int MINB>
__global__ void __launch_bounds__(128, MINB) k_render_stream(DScene s, DCamera cam, RenderArgs a, int leaf_min,
                                                             int shade_min, float *__restrict__ accum,
                                                             RenderCounters *ctr) {
  constexpr unsigned FULL = 0xffffffffu;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int j = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
  const int i = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
  const bool inside = i < a.height && j < a.width;
  const int idx = i * a.width + j;
  PathKey pk;
  pk.pixel = (uint32_t)idx;
  pk.sample = 0;
  pk.k0 = a.k0;
  pk.k1 = a.k1;
  int k = a.spp_begin;
  V3 color = mk(0.f, 0.f, 0.f), T = mk(1.f, 1.f, 1.f);
  int depth = 0;
  bool need_new = true, have_query = false;
  int state = inside ? ST_SHADE : ST_DONE;
  Range rg;
  rg.t_from = 1e-3;
  rg.t_to = (double)CUDART_INF_F;
  rg.tf_f = float_ge(1e-3);
  Trav tv;
  tv.r.o = mk(0.f, 0.f, 0.f);
  tv.r.d = mk(0.f, 0.f, 1.f);
  tv.sp = 0;
  tv.ambiguous = tv.overflow = false;
  uint2 stack[STACK_SIZE];
  unsigned long long n_samples = 0, n_rays = 0, n_rays_ref = 0, n_nodes = 0, n_prims = 0, n_iters = 0, n_lanes = 0;
  for (;;) {
    const unsigned mN = __ballot_sync(FULL, state == ST_NODE);
    const unsigned mL = __ballot_sync(FULL, state == ST_LEAF);
    const unsigned mS = __ballot_sync(FULL, state == ST_SHADE);
    if ((mN | mL | mS) == 0u) break;
    const bool do_leaf = __popc(mL) >= leaf_min || mN == 0u;
    const bool do_shade = __popc(mS) >= shade_min || mN == 0u;
    if (STATS && lane == 0) {
      n_iters++;
      n_lanes += __popc(mN) + (do_leaf ? __popc(mL) : 0) + (do_shade ? __popc(mS) : 0);
    }
    if (do_shade && state == ST_SHADE) {  // between two queries of this lane
      if (have_query) {
        Hit h = trav_finish_call(s, tv, rg);
        n_rays++;
        bool alive = false;
        if (h.rec != REC_NONE) {
          ShadeRet sr = shade_hit_call(s, tv.r, h, pk, depth, T, color);
          alive = sr.alive;
          tv.r = sr.ray;
          T = sr.T;
          color = sr.color;
        }
        if (alive) {
          depth++;
        } else {
          k++;
          need_new = true;
        }
        have_query = false;
      }
      for (;;) {
        if (need_new) {
          if (k >= a.spp_end) {
            accum[3 * (size_t)idx + 0] += color.x;
            accum[3 * (size_t)idx + 1] += color.y;
            accum[3 * (size_t)idx + 2] += color.z;
            state = ST_DONE;
            break;
          }
          pk.sample = (uint32_t)k;
          tv.r = camera_sample_call(cam, i, j, a.height, a.width, pk);
          T = mk(1.f, 1.f, 1.f);
          depth = 0;
          need_new = false;
          n_samples++;
        }
        n_rays_ref++;
        if (depth < a.max_depth) {  // ray_tracing.cu:23: the query at depth == limit yields 0
          have_query = true;
          state = trav_begin(s, tv, rg);
          break;
        }
        k++;
        need_new = true;
      }
    } else if (do_leaf && state == ST_LEAF) {
      if (STATS) n_prims++;
      LeafRet lr = trav_leaf_step_call(s, tv, rg, stack);
      tv = lr.tv;
      state = lr.state;
    } else if (state == ST_NODE) {
      if (STATS) n_nodes += 2;
      state = trav_node_step(s, tv, stack);
    }
  }
  flush_counters(ctr, n_samples, n_rays, n_rays_ref, n_nodes, n_prims);
  if (STATS && lane == 0) {
    atomicAdd(&ctr->sched_iters, n_iters);
    atomicAdd(&ctr->sched_lanes, n_lanes);
  }
}

// ---- vote kernel: the fused kernel as a warp-scheduled state machine ------------------------------
// Same work as k_render_fused (one lane per pixel, samples in increasing order, paths regenerated in
// registers) but every lane carries an explicit state — GEN (needs a camera sample), NODE (inner-node
// step pending), LEAF (record test pending), SHADE (query finished) — and each iteration the warp
// executes ONE kind of step: the kind most of its lanes are waiting for.  Lanes in other states keep
// their state and wait, so their number grows until they are the majority.  This sorts the work of
// a warp by kind in time instead of in memory (no queues): node steps, record tests and shading each
// run with most lanes active instead of the 8 of 32 measured for the lane-by-lane loop
// (profiles/r01_a_cornell_fused_ncu_summary.txt).  Image and counters are identical to the fused kernel.
template <bool STATS>
__global__ void __launch_bounds__(128) k_render_vote(DScene s, DCamera cam, RenderArgs a, float *__restrict__ accum,
                                                     RenderCounters *ctr) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int j = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
  const int i = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
  const bool inside = i < a.height && j < a.width;
  const int idx = i * a.width + j;
  PathKey pk;
  pk.pixel = (uint32_t)idx;
  pk.sample = 0;
  pk.k0 = a.k0;
  pk.k1 = a.k1;
  int k = a.spp_begin;
  V3 color = mk(0.f, 0.f, 0.f), T = mk(1.f, 1.f, 1.f);
  int depth = 0;
  int state = inside ? ST_GEN : ST_DONE;
  Range rg;
  rg.t_from = 1e-3;
  rg.t_to = (double)CUDART_INF_F;
  rg.tf_f = float_ge(1e-3);
  Trav tv;
  tv.r.o = tv.r.d = mk(0.f, 0.f, 0.f);
  tv.sp = 0;
  uint2 stack[STACK_SIZE];
  unsigned long long n_samples = 0, n_rays = 0, n_rays_ref = 0, n_nodes = 0, n_prims = 0, n_iters = 0, n_lanes = 0;
  // next closest-hit query of the current path, or the end of the path at the depth limit
  // (ray_tracing.cu:23: the query at depth == limit yields 0 and is skipped)
  auto start_query = [&]() {
    n_rays_ref++;
    if (depth < a.max_depth) {
      n_rays++;
      state = trav_begin(s, tv, rg);
    } else {
      k++;
      state = ST_GEN;
    }
  };
  for (;;) {
    const unsigned mN = __ballot_sync(0xffffffffu, state == ST_NODE);
    const unsigned mL = __ballot_sync(0xffffffffu, state == ST_LEAF);
    const unsigned mS = __ballot_sync(0xffffffffu, state == ST_SHADE);
    const unsigned mG = __ballot_sync(0xffffffffu, state == ST_GEN);
    if ((mN | mL | mS | mG) == 0u) break;
    int pick = ST_NODE, best = __popc(mN);
    int c = __popc(mL);
    if (c > best) {
      pick = ST_LEAF;
      best = c;
    }
    c = __popc(mS);
    if (c > best) {
      pick = ST_SHADE;
      best = c;
    }
    c = __popc(mG);
    if (c > best) {
      pick = ST_GEN;
      best = c;
    }
    if (STATS && lane == 0) {
      n_iters++;
      n_lanes += best;
    }
    if (state != pick) continue;
    if (pick == ST_NODE) {
      if (STATS) n_nodes += 2;
      state = trav_node_step(s, tv, stack);
    } else if (pick == ST_LEAF) {
      if (STATS) n_prims++;
      state = trav_leaf_step(s, tv, rg, stack);
    } else if (pick == ST_SHADE) {
      Hit h = trav_finish(s, tv, rg);
      bool alive = h.rec != REC_NONE && shade_hit(s, tv.r, h, pk, depth, T, color);
      if (alive) {
        depth++;
        start_query();
      } else {
        k++;
        state = ST_GEN;
      }
    } else {  // ST_GEN
      if (k >= a.spp_end) {
        accum[3 * (size_t)idx + 0] += color.x;
        accum[3 * (size_t)idx + 1] += color.y;
        accum[3 * (size_t)idx + 2] += color.z;
        state = ST_DONE;
      } else {
        pk.sample = (uint32_t)k;
        tv.r = camera_sample(cam, i, j, a.height, a.width, pk);
        T = mk(1.f, 1.f, 1.f);
        depth = 0;
        n_samples++;
        start_query();
      }
    }
  }
  flush_counters(ctr, n_samples, n_rays, n_rays_ref, n_nodes, n_prims);
  if (STATS && lane == 0) {
    atomicAdd(&ctr->sched_iters, n_iters);
    atomicAdd(&ctr->sched_lanes, n_lanes);
  }
}

// ---- pool kernel: K path slots per lane, warp-scheduled ------------------------------------------------
// The vote kernel loses because a lane whose only path waits for a minority step idles.  Here every lane owns K
// pixels (K path slots, state in per-thread local memory = L1), and each iteration the warp executes the step
// kind for which the MOST LANES have a ready slot; a lane runs it on its first slot in that state.  Waiting slots
// cost storage, not issue slots: a lane idles only when none of its K slots wants the chosen step.  Samples of a
// pixel still run in increasing order in one slot, so the image is bit-identical to the fused kernel's.
struct PathSlot {
  V3 T, color;
  int32_t depth, k;
};

template <bool STATS, int K>
__global__ void __launch_bounds__(128) k_render_pool(DScene s, DCamera cam, RenderArgs a, float *__restrict__ accum,
                                                     RenderCounters *ctr) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int j0 = blockIdx.x * (16 * K) + (warp & 1) * 8 + (lane & 7);
  const int i = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
  Range rg;
  rg.t_from = 1e-3;
  rg.t_to = (double)CUDART_INF_F;
  rg.tf_f = float_ge(1e-3);
  Trav tv[K];
  PathSlot ps[K];
  uint2 stack[K][STACK_SIZE];
  uint32_t mG = 0, mN = 0, mL = 0, mS = 0;  // bit q: slot q is in that state
#pragma unroll
  for (int q = 0; q < K; q++) {
    ps[q].color = mk(0.f, 0.f, 0.f);
    ps[q].T = mk(1.f, 1.f, 1.f);
    ps[q].depth = 0;
    ps[q].k = a.spp_begin;
    tv[q].sp = 0;
    if (i < a.height && j0 + 16 * q < a.width) mG |= 1u << q;
  }
  unsigned long long n_samples = 0, n_rays = 0, n_rays_ref = 0, n_nodes = 0, n_prims = 0, n_iters = 0, n_lanes = 0;
  auto set_state = [&](int q, int state) {
    const uint32_t bit = 1u << q;
    mG &= ~bit, mN &= ~bit, mL &= ~bit, mS &= ~bit;
    if (state == ST_GEN) mG |= bit;
    else if (state == ST_NODE) mN |= bit;
    else if (state == ST_LEAF) mL |= bit;
    else if (state == ST_SHADE) mS |= bit;
  };
  auto start_query = [&](int q) {
    n_rays_ref++;
    if (ps[q].depth < a.max_depth) {
      n_rays++;
      set_state(q, trav_begin(s, tv[q], rg));
    } else {
      ps[q].k++;
      set_state(q, ST_GEN);
    }
  };
  for (;;) {
    const unsigned bN = __ballot_sync(0xffffffffu, mN != 0), bL = __ballot_sync(0xffffffffu, mL != 0);
    const unsigned bS = __ballot_sync(0xffffffffu, mS != 0), bG = __ballot_sync(0xffffffffu, mG != 0);
    if ((bN | bL | bS | bG) == 0u) break;
    int pick = ST_NODE, best = __popc(bN);
    int c = __popc(bL);
    if (c > best) pick = ST_LEAF, best = c;
    c = __popc(bS);
    if (c > best) pick = ST_SHADE, best = c;
    c = __popc(bG);
    if (c > best) pick = ST_GEN, best = c;
    if (STATS && lane == 0) {
      n_iters++;
      n_lanes += best;
    }
    if (pick == ST_NODE) {
      if (mN) {
        const int q = __ffs(mN) - 1;
        if (STATS) n_nodes += 2;
        set_state(q, trav_node_step(s, tv[q], stack[q]));
      }
    } else if (pick == ST_LEAF) {
      if (mL) {
        const int q = __ffs(mL) - 1;
        if (STATS) n_prims++;
        set_state(q, trav_leaf_step(s, tv[q], rg, stack[q]));
      }
    } else if (pick == ST_SHADE) {
      if (mS) {
        const int q = __ffs(mS) - 1;
        PathKey pk;
        pk.pixel = (uint32_t)(i * a.width + j0 + 16 * q);
        pk.sample = (uint32_t)ps[q].k;
        pk.k0 = a.k0;
        pk.k1 = a.k1;
        Hit h = trav_finish(s, tv[q], rg);
        V3 T = ps[q].T, color = ps[q].color;
        bool alive = h.rec != REC_NONE && shade_hit(s, tv[q].r, h, pk, ps[q].depth, T, color);
        ps[q].T = T;
        ps[q].color = color;
        if (alive) {
          ps[q].depth++;
          start_query(q);
        } else {
          ps[q].k++;
          set_state(q, ST_GEN);
        }
      }
    } else if (mG) {
      const int q = __ffs(mG) - 1;
      const int j = j0 + 16 * q;
      const int idx = i * a.width + j;
      if (ps[q].k >= a.spp_end) {
        accum[3 * (size_t)idx + 0] += ps[q].color.x;
        accum[3 * (size_t)idx + 1] += ps[q].color.y;
        accum[3 * (size_t)idx + 2] += ps[q].color.z;
        set_state(q, ST_DONE);
      } else {
        PathKey pk;
        pk.pixel = (uint32_t)idx;
        pk.sample = (uint32_t)ps[q].k;
        pk.k0 = a.k0;
        pk.k1 = a.k1;
        tv[q].r = camera_sample(cam, i, j, a.height, a.width, pk);
        ps[q].T = mk(1.f, 1.f, 1.f);
        ps[q].depth = 0;
        n_samples++;
        start_query(q);
      }
    }
  }
  flush_counters(ctr, n_samples, n_rays, n_rays_ref, n_nodes, n_prims);
  if (STATS && lane == 0) {
    atomicAdd(&ctr->sched_iters, n_iters);
    atomicAdd(&ctr->sched_lanes, n_lanes);
  }
}

// ---- wavefront pipeline ------------------------------------------------------------------------------
// One path slot per pixel, paths regenerated in place (sample k+1 starts in the slot the moment
// sample k ends), so the pool stays full until pixels run out of samples and every pixel sums its
// samples in increasing order — the image is bit-identical to the fused kernel's.
//   logic  : shade the slot's hit (Material::Scatter/Emit), regenerate (Camera::RayAt), enqueue
//   extend : closest-hit query for every queued slot
// Path state is SoA, float4 per field: 16-byte coalesced accesses.
struct WfState {
  float4 *ray_o;   // o.xyz, w = hit t (float)
  float4 *ray_d;   // d.xyz, w = hit record (encoded)
  float4 *thr;     // T.xyz, w = depth | flags
  float4 *color;   // colour sum of the pixel, w = current sample index
  uint32_t *queue[2];
  uint32_t *count;   // [2] queue lengths, ping-pong
  uint32_t *mqueue;  // material-sorted shade queues: WF_CLASSES segments of n_slots entries (null: unsorted)
  uint32_t *mcount;  // [2][WF_CLASSES] lengths, ping-pong with `count`
  uint32_t n_slots;
};
// Shade classes: the five material types (RT_MAT_*) and "no hit".  A warp of the sorted logic pass runs one
// Material::Scatter / Emit body instead of a mixture (north_star (3): material-sorted shading queues).
constexpr int WF_CLASSES = 6;
__device__ __forceinline__ int shade_class(const DScene &s, const Hit &h) {
  if (h.rec == REC_NONE) return 5;
  int mat_id;
  if (h.rec == REC_SKY) {
    mat_id = s.sky_mat;
  } else {
    const float4 *recs = h.seq ? s.recs_seq : s.recs;
    mat_id = (int)(__float_as_uint(__ldg(recs + 3 * (size_t)h.rec + 2).w) >> KIND_BITS);
  }
  int t = __float_as_int(__ldg(reinterpret_cast<const float4 *>(s.mats + mat_id) + 1).x);
  return (t < 0 || t > 4) ? 5 : t;
}
constexpr int WF_DONE = 1 << 30, WF_NEW = 1 << 29;

__device__ __forceinline__ int hit_encode(const Hit &h) { return (h.rec < 0) ? h.rec : (h.seq ? (h.rec | 0x40000000) : h.rec); }
__device__ __forceinline__ Hit hit_decode(int enc, float t) {
  Hit h;
  h.t = (double)t;
  if (enc < 0) {
    h.rec = enc;
    h.seq = 0;
  } else {
    h.rec = enc & 0x3fffffff;
    h.seq = (enc >> 30) & 1;
  }
  return h;
}

// Warp-aggregated append of `slot` to a queue.
__device__ __forceinline__ void queue_push(bool pred, uint32_t slot, uint32_t *queue, uint32_t *count) {
  unsigned m = __ballot_sync(0xffffffffu, pred);
  if (m == 0) return;
  int lane = threadIdx.x & 31;
  int leader = __ffs(m) - 1;
  uint32_t base = 0;
  if (lane == leader) base = atomicAdd(count, (uint32_t)__popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (pred) queue[base + __popc(m & ((1u << lane) - 1))] = slot;
}

template <bool FIRST>
__global__ void __launch_bounds__(256) k_wf_logic(DScene s, DCamera cam, RenderArgs a, WfState w, int qi,
                                                  float *__restrict__ accum, RenderCounters *ctr) {
  const int n_slots = a.height * a.width;
  const uint32_t n_in = FIRST ? (uint32_t)n_slots : w.count[qi];
  unsigned long long n_samples = 0, n_rays_ref = 0;
  // sorted input: position q of the concatenated class queues (each class padded to a warp boundary, so that no
  // warp mixes two classes)
  uint32_t mend[WF_CLASSES], mtotal = 0;
  const bool sorted = !FIRST && w.mqueue != nullptr;
  const uint32_t *mc = w.mcount + qi * WF_CLASSES;
  if (sorted) {
#pragma unroll
    for (int c = 0; c < WF_CLASSES; c++) {
      mtotal += (mc[c] + 31u) & ~31u;
      mend[c] = mtotal;
    }
  }
  const uint32_t n_iter = sorted ? mtotal : ((n_in + 31u) & ~31u);
  for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < n_iter; q += gridDim.x * blockDim.x) {
    bool valid = q < n_in;
    uint32_t slot = 0;
    if (sorted) {
      int c = 0;
      uint32_t begin = 0;
#pragma unroll
      for (int k = 0; k < WF_CLASSES - 1; k++)
        if (q >= mend[k]) {
          c = k + 1;
          begin = mend[k];
        }
      valid = q - begin < mc[c];
      if (valid) slot = w.mqueue[(size_t)c * w.n_slots + (q - begin)];
    } else if (valid) {
      slot = FIRST ? q : w.queue[qi][q];
    }
    bool push = false;
    if (valid) {
      const int i = slot / a.width, j = slot % a.width;
      PathKey pk;
      pk.pixel = slot;
      pk.k0 = a.k0;
      pk.k1 = a.k1;
      V3 color, T;
      RayQ ray;
      int k, depth;
      bool need_new;
      if (FIRST) {
        color = mk(0.f, 0.f, 0.f);
        T = mk(1.f, 1.f, 1.f);
        k = a.spp_begin;
        depth = 0;
        need_new = true;
        ray.o = ray.d = mk(0.f, 0.f, 0.f);
      } else {
        float4 ro = w.ray_o[slot], rd = w.ray_d[slot], th = w.thr[slot], co = w.color[slot];
        ray.o = xyz(ro);
        ray.d = xyz(rd);
        T = xyz(th);
        color = xyz(co);
        depth = __float_as_int(th.w);
        k = __float_as_int(co.w);
        pk.sample = (uint32_t)k;
        Hit h = hit_decode(__float_as_int(rd.w), ro.w);
        bool alive = h.rec != REC_NONE && shade_hit(s, ray, h, pk, depth, T, color);
        if (alive) {
          depth++;
          need_new = false;
        } else {
          k++;
          need_new = true;
        }
      }
      // next query of this slot (a path at depth == limit contributes 0: ray_tracing.cu:23-26)
      for (;;) {
        if (need_new) {
          if (k >= a.spp_end) break;
          pk.sample = (uint32_t)k;
          ray = camera_sample(cam, i, j, a.height, a.width, pk);
          T = mk(1.f, 1.f, 1.f);
          depth = 0;
          need_new = false;
          n_samples++;
        }
        n_rays_ref++;
        if (depth < a.max_depth) {
          push = true;
          break;
        }
        k++;
        need_new = true;
      }
      if (push) {
        w.ray_o[slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, 0.f);
        w.ray_d[slot] = make_float4(ray.d.x, ray.d.y, ray.d.z, 0.f);
        w.thr[slot] = make_float4(T.x, T.y, T.z, __int_as_float(depth));
        w.color[slot] = make_float4(color.x, color.y, color.z, __int_as_float(k));
      } else {
        accum[3 * (size_t)slot + 0] += color.x;
        accum[3 * (size_t)slot + 1] += color.y;
        accum[3 * (size_t)slot + 2] += color.z;
      }
    }
    queue_push(push, slot, w.queue[qi ^ 1], w.count + (qi ^ 1));
  }
  flush_counters(ctr, n_samples, 0, n_rays_ref, 0, 0);
}

template <bool STATS, int TRAV, int MINB = 5>
__global__ void __launch_bounds__(128, MINB) k_wf_extend(DScene s, WfState w, int qi, RenderCounters *ctr) {
  const uint32_t n_in = w.count[qi];
  unsigned long long n_rays = 0, n_nodes = 0, n_prims = 0;
  TravStats ts;
  ts.nodes = ts.prims = 0;
  for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < n_in; q += gridDim.x * blockDim.x) {
    uint32_t slot = w.queue[qi][q];
    float4 ro = w.ray_o[slot], rd = w.ray_d[slot];
    RayQ ray;
    ray.o = xyz(ro);
    ray.d = xyz(rd);
    Hit h = world_hit_bvh<STATS, TRAV>(s, ray, 1e-3, (double)CUDART_INF_F, &ts);
    n_rays++;
    // only the two .w words change: t and the record
    reinterpret_cast<float *>(w.ray_o + slot)[3] = float(h.t);
    reinterpret_cast<int *>(w.ray_d + slot)[3] = hit_encode(h);
    if (w.mqueue) {  // one warp-aggregated push per class present in the warp
      const int c = shade_class(s, h);
      const unsigned peers = __match_any_sync(__activemask(), c);
      const int leader = __ffs(peers) - 1, lane = threadIdx.x & 31;
      uint32_t base = 0;
      if (lane == leader) base = atomicAdd(w.mcount + qi * WF_CLASSES + c, (uint32_t)__popc(peers));
      base = __shfl_sync(peers, base, leader);
      w.mqueue[(size_t)c * w.n_slots + base + __popc(peers & ((1u << lane) - 1))] = slot;
    }
  }
  if (STATS) {
    n_nodes = ts.nodes;
    n_prims = ts.prims;
  }
  flush_counters(ctr, 0, n_rays, 0, n_nodes, n_prims);
}

__global__ void k_wf_reset(uint32_t *count, int which, uint32_t *mcount) {
  if (threadIdx.x == 0) count[which] = 0;
  if (threadIdx.x < WF_CLASSES) mcount[which * WF_CLASSES + threadIdx.x] = 0;
}

// ---- resolve: ray_tracing.cu:78-83 + WriteImage's *255 truncation (utils.cu:87-96) ---------------
__global__ void k_resolve(const float *__restrict__ accum, int64_t n, int spp_total, float *rgb, uint8_t *rgb8) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float c = accum[i] / float(spp_total);
  c = fminf(fmaxf(c, 0.f), 1.f);  // glm::clamp = min(max(x, lo), hi)
  c = sqrtf(c);
  if (rgb) rgb[i] = c;
  if (rgb8) rgb8[i] = (uint8_t)(c * 255);
}

// RT_RENDER_SPLIT: adds the sums of slices 1..S-1 to the image in slice order, so that the result is the one of S
// consecutive rt_render calls over the sample sub-ranges (slice 0 has already been added by its own kernel).
__global__ void k_slice_sum(float *__restrict__ accum, const float *__restrict__ sums, int64_t n, int n_extra) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float c = accum[i];
  for (int z = 0; z < n_extra; z++) c += sums[(int64_t)z * n + i];
  accum[i] = c;
}

}  // namespace rtc

// =================================================================================================
// C-ABI
// =================================================================================================
using namespace rtc;

namespace {
inline int nblk(size_t n, int b) { return (int)((n + b - 1) / b); }

// Traversal variant of the render kernels: RT_TRAV=0..22 overrides the default (diagnostics).
// Default (measured, profiles/r01_h_variants_sweep.txt, r01_o_wide_ifif_sweep.txt, r01_v_box_list_sweep.txt):
//   <= 32 primitives (Cornell, birthday): no tree — box list, every lane tests every box (rt_flat.cuh);
//   <= 4096 records (spheres, birthday): binary tree, while-while loop — incoherent bounce rays in an
//      L1-resident tree;  <= 64 records without spheres (Cornell): the same loop over the 4-wide tree, +4 %
//      (with FP64 sphere tests in the loop the wide step spills at 80 registers and loses);
//   meshes up to 500 k records (bunny): 8-wide tree, per-lane if-if loop — 4-5 dependent steps instead of 12,
//      +13 % over the binary if-if loop;
//   larger (soups): binary tree, if-if loop at 64 registers — latency bound, 32 warps/SM beat fewer, fatter steps.

int check_ready(rt_ctx *ctx, bool need_camera, const char *who) {
  if (!ctx) {
    set_error(std::string(who) + ": null context");
    return RT_ERR_INVALID;
  }
  if (!ctx->has_scene) {
    set_error(std::string(who) + ": no scene committed");
    return RT_ERR_INVALID;
  }
  if (need_camera && !ctx->has_camera) {
    set_error(std::string(who) + ": no camera set");
    return RT_ERR_INVALID;
  }
  return RT_OK;
}
}  // namespace

int rtc::trav_variant(const rt_ctx *ctx) {
  const char *e = getenv("RT_TRAV");
  int dflt = ctx->scene.n_recs <= 4096 ? 1 : (ctx->scene.n_recs < 500000 ? 17 : 0);
  if (ctx->scene.n_recs <= 64 && !ctx->scene.has_spheres) dflt = 18;
  if (ctx->flat.n > 0) dflt = 20;  // <= 32 primitives: box list (rt_flat.cuh), +37 % on Cornell, +20 % on birthday
  int v = e ? atoi(e) : dflt;
  if (v == 20 && ctx->flat.n == 0) v = dflt;  // no box list for this scene
  if (v == 23 && ctx->flat.n == 0) v = dflt;
  return (v < 0 || v > 23) ? dflt : v;
}

unsigned rtc::trees_for_trav(int trav) {
  switch (trav) {
    case 10: case 11: case 13: case 17: case 19: return RT_TREE_W8;
    case 15: case 16: case 18: return RT_TREE_W4;
    case 21: return RT_TREE_C4;
    case 22: return RT_TREE_C8;
    default: return 0u;
  }
}
unsigned rtc::trees_for_mode(int mode) {
  switch (mode) {
    case 8: case 9: return RT_TREE_W8;
    case 11: return RT_TREE_W4;
    case 13: return RT_TREE_C4;
    case 14: return RT_TREE_C8;
    default: return 0u;
  }
}

extern "C" int rt_trace_closest_device(rt_ctx *ctx, const float *d_rays, int64_t n, double t_from, double t_to, int mode,
                                       rt_hit *d_hits, void *stream) {
  int rc = check_ready(ctx, false, "rt_trace_closest");
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!d_rays || !d_hits)) || (mode < 0 || mode > 14)) {
    set_error("rt_trace_closest: bad argument");
    return RT_ERR_INVALID;
  }
  if (n == 0) return RT_OK;
  RT_CUDA(cudaSetDevice(ctx->device));
  rc = ensure_trees(ctx, trees_for_mode(mode), (cudaStream_t)stream);
  if (rc) return rc;
  if (mode == 12 && ctx->flat.n == 0 && ctx->scene.n_recs > 0) {
    set_error("rt_trace_closest: mode 12 (box list) needs a scene of at most 32 primitives");
    return RT_ERR_INVALID;
  }
  if (mode == 12)
    k_trace_closest_flat<<<nblk((size_t)n, 128), 128, 0, (cudaStream_t)stream>>>(ctx->scene, ctx->flat, d_rays, n, t_from,
                                                                                 t_to, d_hits);
  else if (mode >= 8 && mode <= 11)
    k_trace_closest_warp<<<nblk((size_t)n, 128), 128, 0, (cudaStream_t)stream>>>(ctx->scene, d_rays, n, t_from, t_to,
                                                                                 mode, d_hits);
  else
    k_trace_closest<<<nblk((size_t)n, 128), 128, 0, (cudaStream_t)stream>>>(ctx->scene, d_rays, n, t_from, t_to, mode,
                                                                            d_hits);
  RT_CUDA(cudaGetLastError());
  return RT_OK;
}

extern "C" int rt_trace_closest_host(rt_ctx *ctx, const float *h_rays, int64_t n, double t_from, double t_to, int mode,
                                     rt_hit *h_hits) {
  int rc = check_ready(ctx, false, "rt_trace_closest");
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!h_rays || !h_hits))) {
    set_error("rt_trace_closest: bad argument");
    return RT_ERR_INVALID;
  }
  if (n == 0) return RT_OK;
  RT_CUDA(cudaSetDevice(ctx->device));
  RT_CUDA(ctx->rays_tmp.reserve((size_t)n * 24));
  RT_CUDA(ctx->hits_tmp.reserve((size_t)n * sizeof(rt_hit)));
  RT_CUDA(cudaMemcpy(ctx->rays_tmp.p, h_rays, (size_t)n * 24, cudaMemcpyHostToDevice));
  rc = rt_trace_closest_device(ctx, ctx->rays_tmp.as<float>(), n, t_from, t_to, mode, ctx->hits_tmp.as<rt_hit>(), nullptr);
  if (rc) return rc;
  RT_CUDA(cudaMemcpy(h_hits, ctx->hits_tmp.p, (size_t)n * sizeof(rt_hit), cudaMemcpyDeviceToHost));
  return RT_OK;
}

extern "C" int rt_primary_rays_host(rt_ctx *ctx, int height, int width, float *h_rays) {
  if (!ctx || !ctx->has_camera || height <= 0 || width <= 0 || !h_rays) {
    set_error("rt_primary_rays_host: bad argument or no camera");
    return RT_ERR_INVALID;
  }
  RT_CUDA(cudaSetDevice(ctx->device));
  size_t n = (size_t)height * width;
  RT_CUDA(ctx->rays_tmp.reserve(n * 24));
  k_primary_rays<<<nblk(n, 256), 256>>>(ctx->camera, height, width, ctx->rays_tmp.as<float>());
  RT_CUDA(cudaGetLastError());
  RT_CUDA(cudaMemcpy(h_rays, ctx->rays_tmp.p, n * 24, cudaMemcpyDeviceToHost));
  return RT_OK;
}

static int render_wavefront(rt_ctx *ctx, const RenderArgs &a, bool stats_on, bool sort, float *d_accum,
                            RenderCounters *ctr, cudaStream_t st, uint32_t *iterations, uint32_t *launches) {
  size_t n = (size_t)a.height * a.width;
  for (int i = 0; i < 4; i++) RT_CUDA(ctx->wf[i].reserve(n * 16));
  RT_CUDA(ctx->wf[4].reserve(n * 4));
  RT_CUDA(ctx->wf[5].reserve(n * 4));
  RT_CUDA(ctx->wf[6].reserve(64));
  WfState w;
  w.ray_o = ctx->wf[0].as<float4>();
  w.ray_d = ctx->wf[1].as<float4>();
  w.thr = ctx->wf[2].as<float4>();
  w.color = ctx->wf[3].as<float4>();
  w.queue[0] = ctx->wf[4].as<uint32_t>();
  w.queue[1] = ctx->wf[5].as<uint32_t>();
  w.count = ctx->wf[6].as<uint32_t>();
  w.n_slots = (uint32_t)n;
  w.mqueue = nullptr;
  w.mcount = w.count + 4;
  if (sort) {
    RT_CUDA(ctx->wf[7].reserve(n * 4 * WF_CLASSES));
    w.mqueue = ctx->wf[7].as<uint32_t>();
  }
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
  // resident blocks per SM the extend kernel is compiled for: long, latency-bound queries want warps, not registers
  const char *wmb = getenv("RT_WF_MINB");
  const int wf_minb = wmb ? atoi(wmb) : (ctx->scene.n_recs >= 500000 ? 10 : 5);
  RT_CUDA(cudaMemsetAsync(w.count, 0, 64, st));
  // first logic pass reads no queue and fills queue 1
  int qi = 0;
  k_wf_logic<true><<<std::min(nblk(n, 256), sms * 8), 256, 0, st>>>(ctx->scene, ctx->camera, a, w, qi, d_accum, ctr);
  *launches += 1;
  qi ^= 1;
  uint32_t live = 0;
  for (;;) {
    RT_CUDA(cudaMemcpyAsync(&live, w.count + qi, 4, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));
    if (live == 0) break;
    int gb = std::min(nblk(live, 128), sms * 32);
    if (stats_on)
      k_wf_extend<true, RT_DEFAULT_TRAV><<<gb, 128, 0, st>>>(ctx->scene, w, qi, ctr);
    else if (wf_minb >= 12)
      k_wf_extend<false, RT_DEFAULT_TRAV, 12><<<gb, 128, 0, st>>>(ctx->scene, w, qi, ctr);
    else if (wf_minb >= 10)
      k_wf_extend<false, RT_DEFAULT_TRAV, 10><<<gb, 128, 0, st>>>(ctx->scene, w, qi, ctr);
    else if (wf_minb >= 8)
      k_wf_extend<false, RT_DEFAULT_TRAV, 8><<<gb, 128, 0, st>>>(ctx->scene, w, qi, ctr);
    else
      k_wf_extend<false, RT_DEFAULT_TRAV><<<gb, 128, 0, st>>>(ctx->scene, w, qi, ctr);
    k_wf_reset<<<1, 32, 0, st>>>(w.count, qi ^ 1, w.mcount);
    // sorted: up to WF_CLASSES * 31 padding slots on top of `live`
    k_wf_logic<false><<<std::min(nblk(live + 32 * WF_CLASSES, 256), sms * 8), 256, 0, st>>>(ctx->scene, ctx->camera, a, w, qi,
                                                                                          d_accum, ctr);
    *launches += 3;
    *iterations += 1;
    qi ^= 1;
  }
  RT_CUDA(cudaGetLastError());
  return RT_OK;
}

extern "C" int rt_render(rt_ctx *ctx, const rt_render_params *p, float *d_accum_rgb, rt_render_stats *stats,
                         void *stream) {
  int rc = check_ready(ctx, true, "rt_render");
  if (rc) return rc;
  if (!p || !d_accum_rgb || p->height <= 0 || p->width <= 0 || p->spp_count < 0 || p->spp_begin < 0 ||
      p->max_depth < 0 || (int64_t)p->height * p->width > 0x3fffffff ||
      (int64_t)p->spp_begin + p->spp_count > 0x7fffffff) {
    set_error("rt_render: bad parameters");
    return RT_ERR_INVALID;
  }
  RT_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  RenderArgs a;
  a.height = p->height;
  a.width = p->width;
  a.spp_begin = p->spp_begin;
  a.spp_end = p->spp_begin + p->spp_count;
  a.max_depth = p->max_depth;
  a.k0 = (uint32_t)p->seed;
  a.k1 = (uint32_t)(p->seed >> 32);
  size_t n = (size_t)p->height * p->width;
  RT_CUDA(ctx->counters.reserve(sizeof(RenderCounters)));
  RenderCounters *ctr = ctx->counters.as<RenderCounters>();
  RT_CUDA(cudaMemsetAsync(ctr, 0, sizeof(RenderCounters), st));
  if (p->flags & RT_RENDER_CLEAR) RT_CUDA(cudaMemsetAsync(d_accum_rgb, 0, n * 12, st));
  bool stats_on = (p->flags & RT_RENDER_PROFILE) != 0;
  bool fused = (p->flags & RT_RENDER_FUSED) != 0;
  if (!(p->flags & (RT_RENDER_FUSED | RT_RENDER_WAVEFRONT))) {
    // default strategy: scenes whose nodes + records stay cache resident run fused (no queue traffic)
    const char *env = getenv("RT_STRATEGY");
    fused = env ? strcmp(env, "wavefront") != 0 : true;
  }
  uint32_t iterations = 0, launches = 0;
  if (fused) {  // a variant selected by RT_TRAV after the commit: derive its tree now (outside the timed region)
    int rc_t = ensure_trees(ctx, trees_for_trav(trav_variant(ctx)), st);
    if (rc_t) return rc_t;
  }
  RT_CUDA(cudaEventRecord(ctx->ev0, st));
  if (fused) {
    dim3 grid((p->width + 15) / 16, (p->height + 7) / 8);
    int trav = trav_variant(ctx);
    // resident blocks per SM the kernel is compiled for (register budget): measured best per scene class
    // (profiles/r01_d_occupancy_sweep.txt): 6 for small scenes, 8 for >= 500k records (latency bound), else 5
    const char *mb = getenv("RT_MINB");
    int minb = mb ? atoi(mb) : (ctx->scene.n_recs <= 4096 ? 6 : (ctx->scene.n_recs >= 500000 ? 8 : 5));
    // RT_RENDER_SPLIT: an image of fewer than 8 waves of blocks (spheres 400x225: 725 blocks on 888 slots, one wave whose
    // length is the slowest warp's 16 samples) is rendered as S sample slices on S streams, so that the block scheduler
    // has S times the blocks of 1/S the length to balance.  Slice sums are added in slice order: the image is the one
    // of S consecutive calls over [begin + z*chunk, begin + (z+1)*chunk).
    int S = 1, chunk = p->spp_count;
    if ((p->flags & RT_RENDER_SPLIT) && p->spp_count >= 2) {
      int n_sm = 148;
      cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, ctx->device);
      const long long blocks = (long long)grid.x * grid.y, slots = (long long)n_sm * minb;
      const char *es = getenv("RT_SPLIT");
      const char *ew = getenv("RT_SPLIT_WAVES");
      const long long waves = ew ? atoll(ew) : 16;  // measured on spheres 400x225x16: 8 slices 1338, 16 slices 1400 Msamples/s
      long long want = blocks >= 8 * slots ? 1 : (waves * slots + blocks - 1) / blocks;
      // Meshes behind a tree (4 k - 500 k records): block cost follows the image (bunny 1080p: 1.17 rays per sample on
      // average, an order of magnitude more on the mesh), so a frame ends with a long tail of heavy blocks whatever the
      // wave count - measured 77 % of the bunny frame (S = 1 / 2 / 4 / 8 / 16 / 32: 2 785 / 4 541 / 6 636 / 8 575 / 10 033 /
      // 10 784 Msamples/s).  As many slices as there are: the tail shrinks by S.  Not for the box-list scenes (Cornell
      // -2 % at 16, -5 % at 32 slices) nor the soups (uniform images: 423 -> 384 at 16), profiles/r01_ab_sample_slices.txt.
      if (ctx->flat.n == 0 && ctx->scene.n_recs > 4096 && ctx->scene.n_recs < 500000) want = rt_ctx::kMaxSlices;
      if (es) want = atoll(es);
      const char *em = getenv("RT_SPLIT_MINCHUNK");
      const int min_chunk = em ? std::max(1, atoi(em)) : 1;
      want = std::min<long long>(std::min<long long>(want, rt_ctx::kMaxSlices), p->spp_count / min_chunk);
      want = std::min<long long>(want, 1 + (long long)((2ull << 30) / (n * 12)));  // slice sums: at most 2 GiB
      if (want > 1) {
        chunk = (p->spp_count + (int)want - 1) / (int)want;
        S = (p->spp_count + chunk - 1) / chunk;
      }
    }
    // RT_RENDER_ORDERED: probe, sort the tiles by cost on the host (16 200 tiles at 1080p; one small round trip per
    // call, before the first slice is launched), hand the list to the render kernel.
    const bool ordered = (p->flags & RT_RENDER_ORDERED) && trav == 17 && !stats_on && (minb == 5 || minb == 6);
    if (ordered) {
      const uint32_t n_tiles = grid.x * grid.y;
      RT_CUDA(ctx->tile_cost.reserve((size_t)n_tiles * 4));
      RT_CUDA(ctx->tile_order.reserve((size_t)n_tiles * 4));
      RT_CUDA(cudaMemsetAsync(ctx->tile_cost.p, 0, (size_t)n_tiles * 4, st));
      k_tile_cost<<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, ctx->tile_cost.as<uint32_t>());
      ctx->tile_cost_host.resize(n_tiles);
      ctx->tile_order_host.resize(n_tiles);
      RT_CUDA(cudaMemcpyAsync(ctx->tile_cost_host.data(), ctx->tile_cost.p, (size_t)n_tiles * 4, cudaMemcpyDeviceToHost, st));
      RT_CUDA(cudaStreamSynchronize(st));
      for (uint32_t t = 0; t < n_tiles; t++) ctx->tile_order_host[t] = t;
      const uint32_t *cost = ctx->tile_cost_host.data();
      std::stable_sort(ctx->tile_order_host.begin(), ctx->tile_order_host.end(),
                       [cost](uint32_t x, uint32_t y) { return cost[x] > cost[y]; });
      RT_CUDA(cudaMemcpyAsync(ctx->tile_order.p, ctx->tile_order_host.data(), (size_t)n_tiles * 4, cudaMemcpyHostToDevice, st));
      RT_CUDA(cudaStreamSynchronize(st));  // the host list is reused by the next call
      a.tiles_x = (int32_t)grid.x;
      a.tile_order = ctx->tile_order.as<uint32_t>();
    }
    const RenderArgs a_all = a;
    float *const accum_all = d_accum_rgb;
    const cudaStream_t st_all = st;
    if (S > 1) {
      for (int z = 0; z < S; z++)
        if (!ctx->slice_ev[z]) RT_CUDA(cudaEventCreateWithFlags(&ctx->slice_ev[z], cudaEventDisableTiming));
      for (int z = 0; z + 1 < S; z++)
        if (!ctx->slice_st[z]) RT_CUDA(cudaStreamCreateWithFlags(&ctx->slice_st[z], cudaStreamNonBlocking));
      RT_CUDA(ctx->slice_sums.reserve((size_t)(S - 1) * n * 12));
      RT_CUDA(cudaMemsetAsync(ctx->slice_sums.p, 0, (size_t)(S - 1) * n * 12, st_all));
      RT_CUDA(cudaEventRecord(ctx->slice_ev[0], st_all));
    }
    for (int z = 0; z < S; z++) {
      RenderArgs a = a_all;
      a.spp_begin = a_all.spp_begin + z * chunk;
      a.spp_end = std::min(a_all.spp_end, a.spp_begin + chunk);
      float *d_accum_rgb = z ? ctx->slice_sums.as<float>() + (size_t)(z - 1) * n * 3 : accum_all;
      const cudaStream_t st = z ? ctx->slice_st[z - 1] : st_all;
      if (z) RT_CUDA(cudaStreamWaitEvent(st, ctx->slice_ev[0], 0));
#define RT_FUSED_CASE(T)                                                                               \
  case T:                                                                                              \
    if (stats_on)                                                                                      \
      k_render_fused<true, T><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);     \
    else if (minb == 6 && T < 2)                                                                       \
      k_render_fused<false, T < 2 ? T : 0, 6><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr); \
    else if (minb == 8 && T < 2)                                                                       \
      k_render_fused<false, T < 2 ? T : 0, 8><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr); \
    else                                                                                               \
      k_render_fused<false, T><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);    \
    break;
    if (trav == 21 || trav == 22) {  // if-if over the compact wide trees
#define RT_CWIDE_CASE(T)                                                                                         \
  if (stats_on)                                                                                                  \
    k_render_fused<true, T><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);                 \
  else if (minb == 6)                                                                                            \
    k_render_fused<false, T, 6><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);             \
  else if (minb == 8)                                                                                            \
    k_render_fused<false, T, 8><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);             \
  else                                                                                                           \
    k_render_fused<false, T><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);
      if (trav == 21) {
        RT_CWIDE_CASE(21)
      } else {
        RT_CWIDE_CASE(22)
      }
#undef RT_CWIDE_CASE
    } else if (trav == 23) {  // box list + cooperative ball rejection
      if (stats_on)
        k_render_fused<true, 23><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr, ctx->flat);
      else if (minb == 6)
        k_render_fused<false, 23, 6><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr, ctx->flat);
      else
        k_render_fused<false, 23><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr, ctx->flat);
    } else if (trav == 20) {  // box list
      if (stats_on)
        k_render_fused<true, 20><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr, ctx->flat);
      else if (minb == 6)
        k_render_fused<false, 20, 6><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr, ctx->flat);
      else if (minb == 8)
        k_render_fused<false, 20, 8><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr, ctx->flat);
      else
        k_render_fused<false, 20><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr, ctx->flat);
    } else if (trav >= 16) {  // fused kernel over the wide trees, per lane: if-if 4-wide (16) / 8-wide (17), while-while 4-wide (18) / 8-wide (19)
#define RT_WIDE_IFIF_CASE(T)                                                                                     \
  if (stats_on)                                                                                                  \
    k_render_fused<true, T><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);                 \
  else if (minb == 6)                                                                                            \
    k_render_fused<false, T, 6><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);             \
  else if (minb == 8)                                                                                            \
    k_render_fused<false, T, 8><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);             \
  else                                                                                                           \
    k_render_fused<false, T><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);
      if (ordered) {  // trav 17, heaviest tile first
        const dim3 ogrid(grid.x * grid.y);
        if (minb == 6)
          k_render_fused<false, 17, 6, true><<<ogrid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);
        else
          k_render_fused<false, 17, 5, true><<<ogrid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);
      } else if (trav == 16) {
        RT_WIDE_IFIF_CASE(16)
      } else if (trav == 17) {
        RT_WIDE_IFIF_CASE(17)
      } else if (trav == 18) {
        RT_WIDE_IFIF_CASE(18)
      } else {
        RT_WIDE_IFIF_CASE(19)
      }
#undef RT_WIDE_IFIF_CASE
    } else if (trav == 15) {  // warp-synchronous kernel over the 4-wide tree
#define RT_WARP4_CASE                                                                                            \
  if (stats_on)                                                                                                  \
    k_render_warp<true, 3, 5><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);               \
  else if (minb == 6)                                                                                            \
    k_render_warp<false, 3, 6><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);              \
  else if (minb == 8)                                                                                            \
    k_render_warp<false, 3, 8><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);              \
  else                                                                                                           \
    k_render_warp<false, 3, 5><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);
      RT_WARP4_CASE
#undef RT_WARP4_CASE
    } else if (trav == 14) {  // streaming kernel
      const char *e1 = getenv("RT_STREAM_LEAF"), *e2 = getenv("RT_STREAM_SHADE");
      const int leaf_min = e1 ? atoi(e1) : 8, shade_min = e2 ? atoi(e2) : 8;
      if (stats_on)
        k_render_stream<true, 5><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, leaf_min, shade_min, d_accum_rgb, ctr);
      else if (minb == 8)
        k_render_stream<false, 8><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, leaf_min, shade_min, d_accum_rgb, ctr);
      else if (minb == 6)
        k_render_stream<false, 6><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, leaf_min, shade_min, d_accum_rgb, ctr);
      else
        k_render_stream<false, 5><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, leaf_min, shade_min, d_accum_rgb, ctr);
    } else if (trav >= 12) {  // persistent kernel: 12 binary tree, 13 8-wide tree
#define RT_PERSIST_CASE(WK)                                                                                      \
  if (stats_on)                                                                                                  \
    k_render_persist<true, WK, 5><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);           \
  else if (minb == 6)                                                                                            \
    k_render_persist<false, WK, 6><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);          \
  else if (minb == 8)                                                                                            \
    k_render_persist<false, WK, 8><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);          \
  else if (minb == 4)                                                                                            \
    k_render_persist<false, WK, 4><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);          \
  else                                                                                                           \
    k_render_persist<false, WK, 5><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);
      if (trav == 12) {
        RT_PERSIST_CASE(BinWalker)
      } else {
        RT_PERSIST_CASE(WideWalker)
      }
#undef RT_PERSIST_CASE
    } else if (trav >= 9) {  // warp-synchronous kernel: 9 binary + coop, 10 wide, 11 wide + coop
#define RT_WARP_CASE(M)                                                                                          \
  if (stats_on)                                                                                                  \
    k_render_warp<true, M, 5><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);               \
  else if (minb == 6)                                                                                            \
    k_render_warp<false, M, 6><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);              \
  else if (minb == 8)                                                                                            \
    k_render_warp<false, M, 8><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);              \
  else                                                                                                           \
    k_render_warp<false, M, 5><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);
      if (trav == 9) {
        RT_WARP_CASE(0)
      } else if (trav == 10) {
        RT_WARP_CASE(1)
      } else {
        RT_WARP_CASE(2)
      }
#undef RT_WARP_CASE
    } else if (trav >= 6) {  // pool kernel, K = trav - 4 slots per lane (6 -> 2, 7 -> 3, 8 -> 4)
      const int K = trav - 4;
      dim3 pgrid((p->width + 16 * K - 1) / (16 * K), (p->height + 7) / 8);
      if (K == 2) {
        if (stats_on) k_render_pool<true, 2><<<pgrid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);
        else k_render_pool<false, 2><<<pgrid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);
      } else if (K == 3) {
        if (stats_on) k_render_pool<true, 3><<<pgrid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);
        else k_render_pool<false, 3><<<pgrid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);
      } else {
        if (stats_on) k_render_pool<true, 4><<<pgrid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);
        else k_render_pool<false, 4><<<pgrid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);
      }
    } else
    switch (trav) {
      case 5:
        if (stats_on)
          k_render_vote<true><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);
        else
          k_render_vote<false><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, d_accum_rgb, ctr);
        break;
      RT_FUSED_CASE(0)
      RT_FUSED_CASE(1)
      RT_FUSED_CASE(2)
      RT_FUSED_CASE(3)
      default:
      RT_FUSED_CASE(4)
    }
      if (z) {
        RT_CUDA(cudaEventRecord(ctx->slice_ev[z], st));
        RT_CUDA(cudaStreamWaitEvent(st_all, ctx->slice_ev[z], 0));
      }
    }
#undef RT_FUSED_CASE
    if (S > 1)
      k_slice_sum<<<nblk(n * 3, 256), 256, 0, st_all>>>(accum_all, ctx->slice_sums.as<float>(), (int64_t)n * 3, S - 1);
    RT_CUDA(cudaGetLastError());
    launches = (S > 1 ? S + 1 : 1) + (ordered ? 1 : 0);
    iterations = S;  // sample slices (1: no split)
  } else {
    rc = render_wavefront(ctx, a, stats_on, (p->flags & RT_RENDER_SORT) != 0, d_accum_rgb, ctr, st, &iterations, &launches);
    if (rc) return rc;
  }
  RT_CUDA(cudaEventRecord(ctx->ev1, st));
  if (stats) {
    RenderCounters hc;
    RT_CUDA(cudaMemcpyAsync(&hc, ctr, sizeof(hc), cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaEventSynchronize(ctx->ev1));
    RT_CUDA(cudaStreamSynchronize(st));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    memset(stats, 0, sizeof(*stats));
    stats->samples = hc.samples;
    stats->rays = hc.rays;
    stats->rays_ref_equiv = hc.rays_ref;
    stats->node_visits = hc.nodes;
    stats->prim_tests = hc.prims;
    stats->iterations = iterations;
    stats->kernel_launches = launches;
    stats->render_ms = ms;
    stats->sched_iters = hc.sched_iters;
    stats->sched_lanes = hc.sched_lanes;
  }
  return RT_OK;
}

extern "C" int rt_resolve(rt_ctx *ctx, const float *d_accum_rgb, int height, int width, int spp_total, float *d_rgb,
                          uint8_t *d_rgb8, void *stream) {
  if (!ctx || !d_accum_rgb || height <= 0 || width <= 0 || spp_total <= 0) {
    set_error("rt_resolve: bad argument");
    return RT_ERR_INVALID;
  }
  RT_CUDA(cudaSetDevice(ctx->device));
  int64_t n = (int64_t)height * width * 3;
  k_resolve<<<nblk((size_t)n, 256), 256, 0, (cudaStream_t)stream>>>(d_accum_rgb, n, spp_total, d_rgb, d_rgb8);
  RT_CUDA(cudaGetLastError());
  return RT_OK;
}

extern "C" int rt_render_host(rt_ctx *ctx, const rt_render_params *p, float *h_accum_rgb, float *h_rgb,
                              rt_render_stats *stats) {
  int rc = check_ready(ctx, true, "rt_render_host");
  if (rc) return rc;
  if (!p || p->height <= 0 || p->width <= 0) {
    set_error("rt_render_host: bad parameters");
    return RT_ERR_INVALID;
  }
  RT_CUDA(cudaSetDevice(ctx->device));
  size_t n = (size_t)p->height * p->width * 3;
  RT_CUDA(ctx->accum_tmp.reserve(n * 4));
  rt_render_params q = *p;
  q.flags |= RT_RENDER_CLEAR;
  rc = rt_render(ctx, &q, ctx->accum_tmp.as<float>(), stats, nullptr);
  if (rc) return rc;
  if (h_accum_rgb) RT_CUDA(cudaMemcpy(h_accum_rgb, ctx->accum_tmp.p, n * 4, cudaMemcpyDeviceToHost));
  if (h_rgb) {
    RT_CUDA(ctx->rgb_tmp.reserve(n * 4));
    rc = rt_resolve(ctx, ctx->accum_tmp.as<float>(), p->height, p->width, std::max(p->spp_count, 1),
                    ctx->rgb_tmp.as<float>(), nullptr, nullptr);
    if (rc) return rc;
    RT_CUDA(cudaMemcpy(h_rgb, ctx->rgb_tmp.p, n * 4, cudaMemcpyDeviceToHost));
  }
  RT_CUDA(cudaDeviceSynchronize());
  return RT_OK;
}
