// rt_render.cu — the render hot path: closest-hit queries, the persistent path kernel with its dynamic
// (tile, sample-chunk) work distribution, the wavefront pipeline (logic/shade -> extend) and the resolve kernels.
//
// Replaces, in the reference: PathTracing (ray_tracing.cu:56-85), Trace (ray_tracing.cu:12-54),
// HitableList::Hit (hitable_list.cu:7-25) and everything they call, plus the post-processing tail
// (ray_tracing.cu:78-83, utils.cu:126-129).
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <string>

#include "rt_internal.h"
#include "rt_flat.cuh"

namespace rtc {

// Work distribution of one rt_render call (k_render).  A SUB-UNIT is (pixel, chunk of samples): what one lane renders
// before it asks for more work.  An ITEM is (8x4-pixel tile, some consecutive chunks) = 32 sub-units per chunk: what a
// warp draws from the global counter with one atomic.  The sample range is cut into up to PLAN_MAX PHASES of
// decreasing item size (plan_work): big items first — a warp stays on its tile and both the per-lane hand-out and the
// draw of a new item are rare — small items last, so that the launch ends within a few samples' time on every SM.
// Items are numbered phase, then chunk group, then tile, so every pass sweeps the whole image and the expensive
// tiles are spread over the launch.
constexpr int PLAN_MAX = 4;
struct PlanPhase {
  uint32_t item_begin;  // first item of the phase
  int32_t k_begin;      // first sample of the phase (absolute index)
  int32_t ch, r;        // samples per chunk, chunks per item
};
struct RenderArgs {
  int32_t height, width;
  int32_t spp_begin, spp_end;
  int32_t max_depth;
  uint32_t k0, k1;
  int32_t n_phases;
  PlanPhase phase[PLAN_MAX + 1];  // phase[n_phases] = {n_items, spp_end, -, -}: the end of the last phase
  int32_t tiles_x, n_tiles;
  uint32_t tile_stride;  // tile of the i-th item of a sweep = i * tile_stride mod n_tiles (1: raster order)
  uint32_t n_items;
  unsigned int *work_counter;
  long long *fixed;  // [height * width * 3] radiance sums, 2^-30 units
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

// Traversals (RT_TRAV): 0 binary tree if-if, 1 binary tree while-while, 20 box list.
// EXACT_T = false: the caller only uses (float)Hit::t (render kernels; finalize_sphere_hit).
#ifndef RT_SM_STACK
#define RT_SM_STACK 16  // shared-memory levels of the traversal stack of the large-scene kernel (TRAV 0); 0: all local
#endif
template <bool STATS, int TRAV, bool SPH = true, bool EXACT_T = true, int SM_LEVELS = 0>
__device__ __forceinline__ Hit world_hit(const DScene &s, const FlatItems *fl, const RayQ &r, double t_from, double t_to,
                                         TravStats *st, int self_box = -1, uint2 *sm = nullptr) {
  if constexpr (TRAV == 20)
    return world_hit_flat<STATS, SPH, EXACT_T>(s, *fl, r, t_from, t_to, st, self_box);
  else
    return world_hit_bvh<STATS, TRAV, SPH, EXACT_T, SM_LEVELS>(s, r, t_from, t_to, st, sm);
}

// ---- parity hook: HitableList::Hit on explicit rays ------------------------------------------------
// mode 1 sequential scan, 2 if-if, 3 while-while, 4 box list (mode 0 is resolved by the host to the traversal the
// scene renders with).
__global__ void __launch_bounds__(128) k_trace_closest(DScene s, FlatItems fl, const float *__restrict__ rays, int64_t n,
                                                       double t_from, double t_to, int mode, rt_hit *out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  RayQ r;
  r.o = mk(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]);
  r.d = mk(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]);
  Hit h;
  switch (mode) {
    case 1: h = world_hit_seq(s, r, t_from, t_to); break;
    case 3: h = world_hit<false, 1>(s, &fl, r, t_from, t_to, nullptr); break;
    case 4: h = world_hit<false, 20>(s, &fl, r, t_from, t_to, nullptr); break;
    default: h = world_hit<false, 0>(s, &fl, r, t_from, t_to, nullptr); break;
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

// One bounce of Trace (ray_tracing.cu:20-43) after the closest-hit query, in forward form: returns true when the
// path continues with `ray` / `T` updated; else the path ends and `L` is its radiance (T * emitted, or 0).
// A material either emits or scatters (DiffuseLight / Sky never scatter, the others emit 0: material.cu:3-6), so a
// path contributes at most once, at its end — the back-to-front fold of ray_tracing.cu:50-52 collapses to T * emitted.
template <bool SPH = true>
__device__ __forceinline__ bool shade_hit(const DScene &s, const DCamera &cam, RayQ &ray, const Hit &h, const PathKey &pk,
                                          int depth, PhiloxCache *pc, V3 &T, V3 &L) {
  int mat_id;
  if (h.rec == REC_SKY) {
    mat_id = s.sky_mat;
  } else {
    const float4 *recs = h.seq ? s.recs_seq : s.recs;
    mat_id = (int)(__float_as_uint(__ldg(recs + 3 * (size_t)h.rec + 2).w) >> KIND_BITS);
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
  const bool need_uv = material_needs_uv(m);  // only textured materials read u, v
  float t = float(h.t);
  L = mk(0.f, 0.f, 0.f);
  if (m.type == RT_MAT_SKY || m.type == RT_MAT_DIFFUSE_LIGHT) {
    // Scatter returns false (sky.cu:3-7, diffuse_light.cu:5-9): result = emitted
    float u = 0.f, v = 0.f;
    if (need_uv) {
      Surface sf = surface_at<SPH>(s, ray, h, true);
      u = sf.u;
      v = sf.v;
    }
    V3 p = ray.o + t * ray.d;
    L = T * emit(s, m, u, v, p);
    return false;
  }
  Surface sf = surface_at<SPH>(s, ray, h, need_uv);
  V3 att;
  RayQ out;
  if (!scatter(s, m, ray, t, sf, pk, depth, scatter_word(cam, depth), pc, &att, &out)) return false;  // Emit() is 0
  T = T * att;
  ray = out;
  return true;
}

__device__ __forceinline__ void add_radiance(long long *fixed, int idx, V3 L) {
  const long long qx = to_fixed(L.x), qy = to_fixed(L.y), qz = to_fixed(L.z);
  unsigned long long *f = reinterpret_cast<unsigned long long *>(fixed) + 3 * (size_t)idx;
  if (qx) atomicAdd(f + 0, (unsigned long long)qx);
  if (qy) atomicAdd(f + 1, (unsigned long long)qy);
  if (qz) atomicAdd(f + 2, (unsigned long long)qz);
}

__device__ __forceinline__ void flush_counters(RenderCounters *ctr, unsigned long long samples,
                                               unsigned long long rays, unsigned long long rays_ref,
                                               unsigned long long nodes, unsigned long long prims,
                                               unsigned int seq = 0, unsigned int max_nodes = 0) {
  seq = __reduce_add_sync(0xffffffffu, seq);
  max_nodes = __reduce_max_sync(0xffffffffu, max_nodes);
  if ((threadIdx.x & 31) == 0) {
    if (seq) atomicAdd(&ctr->seq_fallbacks, (unsigned long long)seq);
    if (max_nodes) atomicMax(&ctr->max_nodes, (unsigned long long)max_nodes);
  }
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

// Out-of-line forms of the rare steps (64-register kernels): the node step is the hot loop of a long query and must
// not carry the registers (or the spills) of shading and camera sampling.  Everything goes in and out BY VALUE: a
// reference parameter would pin the traversal state of the hot loop in local memory.
#ifndef RT_WW_MINB  // resident blocks per SM of the while-while kernel (experiment builds: 8 = 64 registers, shading out of line)
#define RT_WW_MINB 6
#endif
struct ShadeRet {
  RayQ ray;
  V3 T, L;
  PhiloxCache pc;
  bool alive;
};
template <bool SPH>
static __device__ __noinline__ ShadeRet shade_hit_call(const DScene &s, const DCamera &cam, RayQ ray, Hit h, PathKey pk,
                                                       int depth, PhiloxCache pc, V3 T) {
  ShadeRet r;
  r.alive = shade_hit<SPH>(s, cam, ray, h, pk, depth, &pc, T, r.L);
  r.ray = ray;
  r.T = T;
  r.pc = pc;
  return r;
}
struct CameraRet {
  RayQ ray;
  PhiloxCache pc;
};
static __device__ __noinline__ CameraRet camera_sample_call(const DCamera &cam, int i, int j, int height, int width,
                                                            PathKey pk) {
  CameraRet r;
  r.ray = camera_sample(cam, i, j, height, width, pk, &r.pc);
  return r;
}

// The hand-out of sub-units to the lanes of a warp (see k_render).  Warp-synchronous: every lane of the warp calls it
// in every iteration; `want` = this lane has run out of samples.  The lanes that want work are ranked by ballot and
// take consecutive sub-units of the warp's current item; when the item is used up lane 0 draws the next one from the
// global counter.  A lane that gets nothing because the launch has run out of items becomes `finished`.
struct WorkPool {
  uint32_t next = 0, end = 0;
  int tile_px = 0, tile_py = 0, k0 = 0, ch = 1, kend = 0;
};
__device__ __forceinline__ void hand_out(const RenderArgs &a, WorkPool &p, bool want, bool &finished, int &pi, int &pj,
                                         int &k, int &k_end, uint32_t &pixel) {
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
  unsigned wm = __ballot_sync(FULL, want);
  while (wm) {
    if (p.next >= p.end) {  // warp-uniform: draw the next item
      uint32_t it = 0;
      if (lane == 0) it = atomicAdd(a.work_counter, 1u);
      it = __shfl_sync(FULL, it, 0);
      if (it >= a.n_items) {
        finished |= want;
        break;
      }
      int ph = 0;
#pragma unroll
      for (int q = 1; q < PLAN_MAX; q++) ph += (q < a.n_phases && it >= a.phase[q].item_begin) ? 1 : 0;
      const PlanPhase pp = a.phase[ph];
      const uint32_t it2 = it - pp.item_begin;
      const uint32_t group = it2 / (uint32_t)a.n_tiles;
      const uint32_t tile = (uint32_t)(((unsigned long long)(it2 - group * (uint32_t)a.n_tiles) * a.tile_stride) % (uint32_t)a.n_tiles);
      p.tile_px = (int)(tile % (uint32_t)a.tiles_x) * 8;
      p.tile_py = (int)(tile / (uint32_t)a.tiles_x) * 4;
      p.ch = pp.ch;
      p.k0 = pp.k_begin + (int)group * pp.r * pp.ch;
      p.kend = a.phase[ph + 1].k_begin;
      p.next = 0;
      p.end = 32u * (uint32_t)pp.r;
    }
    const uint32_t rank = __popc(wm & lt_mask), take = min((uint32_t)__popc(wm), p.end - p.next);
    if (want && rank < take) {
      const uint32_t su = p.next + rank;
      pj = p.tile_px + (int)(su & 7u);
      pi = p.tile_py + (int)((su >> 3) & 3u);
      const int kk = p.k0 + (int)(su >> 5) * p.ch;
      // edge tiles: sub-units outside the image are dropped; so are the chunks of a phase's last item beyond its end
      if (pi < a.height && pj < a.width && kk < p.kend) {
        k = kk;
        k_end = min(kk + p.ch, p.kend);
        pixel = (uint32_t)(pi * a.width + pj);
        want = false;
      }
    }
    p.next += take;
    wm = __ballot_sync(FULL, want);
  }
}

// ---- the render kernel -------------------------------------------------------------------------------------------
// A persistent grid (blocks = SMs x MINB) of independent warps.  Every lane renders sub-units — (pixel, chunk of
// a.ch samples), samples in increasing order, a path regenerated in registers the moment the previous one ends —
// and asks for the next sub-unit when it runs out, so a warp stays full until the whole launch runs out of work:
// no lane waits for the longest pixel of its tile, no block for the longest tile of the image (round 1: 77 % of the
// one-launch bunny frame was that tail, and heaviest-tile-first block order recovered 4 % of it).  The hand-out is
// warp-synchronous and lives in registers: the lanes that want work are ranked by ballot and take consecutive
// sub-units of the warp's current item (pixels of one 8x4 tile first, so a warp's primary rays stay coherent); when
// the item is used up lane 0 draws the next one from the global counter.  Radiance goes to the 64-bit fixed-point
// sums (rt_device.cuh), so the image does not depend on who rendered what.
template <bool STATS, int TRAV, int MINB, bool SPH>
__global__ void __launch_bounds__(128, MINB) k_render(DScene s, DCamera cam, RenderArgs a, RenderCounters *ctr,
                                                      typename FlatParamT<TRAV == 20>::type fl = {}) {
  constexpr unsigned FULL = 0xffffffffu;
  constexpr bool SLIM = MINB >= 8;  // 64 registers: shading and camera sampling out of line
  WorkPool pool;  // warp-uniform: the item being handed out
  // lane: the sub-unit being rendered
  int pi = 0, pj = 0, k = 0, k_end = 0;
  bool finished = false, need_new = true;
  PathKey pk;
  pk.pixel = 0;
  pk.sample = 0;
  pk.k0 = a.k0;
  pk.k1 = a.k1;
  PhiloxCache pc;
  pc.block = PHILOX_NO_BLOCK;
  pc.w2 = pc.w3 = 0;
  V3 T = mk(1.f, 1.f, 1.f);
  RayQ ray;
  ray.o = ray.d = mk(0.f, 0.f, 0.f);
  int depth = 0;
  int self_box = -1;  // box list: the box the current ray starts on
  unsigned long long n_samples = 0, n_rays = 0, n_rays_ref = 0, n_nodes = 0, n_prims = 0;
  unsigned int n_seq = 0, max_nodes = 0;
  TravStats ts;
  ts.nodes = ts.prims = 0;
  const FlatItems *flp = nullptr;
  if constexpr (TRAV == 20) flp = &fl;
  constexpr int SM_LEVELS = TRAV == 0 ? RT_SM_STACK : 0;
  __shared__ uint2 s_stack[SM_LEVELS > 0 ? SM_LEVELS : 1][SM_LEVELS > 0 ? 128 : 1];
  uint2 *const sm = SM_LEVELS > 0 ? &s_stack[0][threadIdx.x] : nullptr;
  for (;;) {
    hand_out(a, pool, !finished && need_new && k >= k_end, finished, pi, pj, k, k_end, pk.pixel);
    if (__all_sync(FULL, finished)) break;
    if (finished) continue;
    if (need_new) {
      pk.sample = (uint32_t)k;
      if constexpr (SLIM) {
        CameraRet cr = camera_sample_call(cam, pi, pj, a.height, a.width, pk);
        ray = cr.ray;
        pc = cr.pc;
      } else {
        ray = camera_sample(cam, pi, pj, a.height, a.width, pk, &pc);
      }
      T = mk(1.f, 1.f, 1.f);
      depth = 0;
      self_box = -1;
      need_new = false;
      n_samples++;
    }
    n_rays_ref++;
    bool alive = depth < a.max_depth;  // ray_tracing.cu:23: the query at depth == limit yields 0
    if (alive) {
      const Hit h = world_hit<STATS, TRAV, SPH, false, SM_LEVELS>(s, flp, ray, 1e-3, (double)CUDART_INF_F, &ts, self_box, sm);
      self_box = h.box;
      n_rays++;
      n_seq += (unsigned int)h.seq;  // the query was answered by the sequential scan (world_hit_seq)
      if (STATS) {
        n_nodes += ts.nodes;
        n_prims += ts.prims;
        max_nodes = max(max_nodes, ts.nodes);
        ts.nodes = ts.prims = 0;
      }
      alive = false;
      if (h.rec != REC_NONE) {
        V3 L;
        if constexpr (SLIM) {
          ShadeRet sr = shade_hit_call<SPH>(s, cam, ray, h, pk, depth, pc, T);
          alive = sr.alive;
          ray = sr.ray;
          T = sr.T;
          pc = sr.pc;
          L = sr.L;
        } else {
          alive = shade_hit<SPH>(s, cam, ray, h, pk, depth, &pc, T, L);
        }
        if (!alive) add_radiance(a.fixed, (int)pk.pixel, L);
      }
    }
    if (alive) {
      depth++;
    } else {
      k++;
      need_new = true;
    }
  }
  flush_counters(ctr, n_samples, n_rays, n_rays_ref, n_nodes, n_prims, n_seq, max_nodes);
}

// ---- wavefront pipeline ------------------------------------------------------------------------------
// One path slot per pixel, paths regenerated in place (sample k+1 starts in the slot the moment
// sample k ends), so the pool stays full until pixels run out of samples.  Same draws, same fixed-point sums:
// the image is bit-identical to k_render's.
//   logic  : shade the slot's hit (Material::Scatter/Emit), regenerate (Camera::RayAt), enqueue
//   extend : closest-hit query for every queued slot
// Path state is SoA, float4 per field: 16-byte coalesced accesses.
struct WfState {
  float4 *ray_o;   // o.xyz, w = hit t (float)
  float4 *ray_d;   // d.xyz, w = hit record (encoded)
  float4 *thr;     // T.xyz, w = depth
  uint32_t *sample;  // current sample index of the slot
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
                                                  RenderCounters *ctr) {
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
      V3 T;
      RayQ ray;
      int k, depth;
      bool need_new;
      if (FIRST) {
        T = mk(1.f, 1.f, 1.f);
        k = a.spp_begin;
        depth = 0;
        need_new = true;
        ray.o = ray.d = mk(0.f, 0.f, 0.f);
      } else {
        float4 ro = w.ray_o[slot], rd = w.ray_d[slot], th = w.thr[slot];
        ray.o = xyz(ro);
        ray.d = xyz(rd);
        T = xyz(th);
        depth = __float_as_int(th.w);
        k = (int)w.sample[slot];
        pk.sample = (uint32_t)k;
        Hit h = hit_decode(__float_as_int(rd.w), ro.w);
        bool alive = false;
        if (h.rec != REC_NONE) {
          V3 L;
          alive = shade_hit(s, cam, ray, h, pk, depth, nullptr, T, L);
          if (!alive) add_radiance(a.fixed, (int)slot, L);
        }
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
          ray = camera_sample(cam, i, j, a.height, a.width, pk, nullptr);
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
        w.sample[slot] = (uint32_t)k;
      }
    }
    queue_push(push, slot, w.queue[qi ^ 1], w.count + (qi ^ 1));
  }
  flush_counters(ctr, n_samples, 0, n_rays_ref, 0, 0);
}

// `queue`: the slots to extend (w.queue[qi], or its reordered copy)
template <bool STATS, int MINB>
__global__ void __launch_bounds__(128, MINB) k_wf_extend(DScene s, WfState w, int qi, const uint32_t *__restrict__ queue,
                                                         RenderCounters *ctr) {
  const uint32_t n_in = w.count[qi];
  unsigned long long n_rays = 0, n_nodes = 0, n_prims = 0;
  TravStats ts;
  ts.nodes = ts.prims = 0;
  for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < n_in; q += gridDim.x * blockDim.x) {
    uint32_t slot = queue[q];
    float4 ro = w.ray_o[slot], rd = w.ray_d[slot];
    RayQ ray;
    ray.o = xyz(ro);
    ray.d = xyz(rd);
    Hit h = world_hit_bvh<STATS, 0, true, false>(s, ray, 1e-3, (double)CUDART_INF_F, &ts);
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

// ---- ray reordering for the extend pass (RT_RENDER_RAYSORT) ---------------------------------------------------
// Bounce rays of a large scene leave a warp's lanes walking unrelated parts of the tree.  Counting sort of the
// extend queue by (direction octant, 32^3 cell of the origin in the scene box): rays that start close together and
// head the same way become neighbours.  Three passes over the queue, all sized on the device (no host round trip):
// histogram, exclusive scan of the 2^18 buckets, scatter.  The order inside a bucket is whatever the atomics give;
// the image does not depend on it (each slot's hit is its own, sums are exact).
constexpr int RAYSORT_BUCKETS = 1 << 18;
__device__ __forceinline__ uint32_t raysort_key(const DScene &s, const float4 ro, const float4 rd) {
  const float4 lo = __ldg(s.nodes), hi = __ldg(s.nodes + 1);  // node 0 = root: the scene box
  uint32_t cell[3];
  const float o[3] = {ro.x, ro.y, ro.z}, l[3] = {lo.x, lo.y, lo.z}, h[3] = {hi.x, hi.y, hi.z};
#pragma unroll
  for (int a = 0; a < 3; a++) {
    const float e = h[a] - l[a];
    float x = e > 0.f ? (o[a] - l[a]) / e : 0.f;
    x = fminf(fmaxf(x, 0.f), 0.999f);
    cell[a] = (uint32_t)(x * 32.f);
  }
  uint32_t m = 0;
#pragma unroll
  for (int b = 0; b < 5; b++)
    m |= ((cell[0] >> b & 1u) << (3 * b + 2)) | ((cell[1] >> b & 1u) << (3 * b + 1)) | ((cell[2] >> b & 1u) << (3 * b));
  const uint32_t oct = (rd.x < 0.f ? 4u : 0u) | (rd.y < 0.f ? 2u : 0u) | (rd.z < 0.f ? 1u : 0u);
  return oct << 15 | m;
}
__global__ void __launch_bounds__(256) k_wf_hist(DScene s, WfState w, int qi, uint32_t *hist) {
  const uint32_t n_in = w.count[qi];
  for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < n_in; q += gridDim.x * blockDim.x) {
    const uint32_t slot = w.queue[qi][q];
    atomicAdd(hist + raysort_key(s, w.ray_o[slot], w.ray_d[slot]), 1u);
  }
}
// one block: hist -> exclusive offsets (in place)
__global__ void __launch_bounds__(1024) k_wf_scan(uint32_t *hist) {
  __shared__ uint32_t part[1024];
  constexpr int PER = RAYSORT_BUCKETS / 1024;
  const int t = threadIdx.x;
  uint32_t sum = 0;
  for (int i = 0; i < PER; i++) sum += hist[t * PER + i];
  part[t] = sum;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {
    uint32_t v = t >= o ? part[t - o] : 0u;
    __syncthreads();
    part[t] += v;
    __syncthreads();
  }
  uint32_t run = part[t] - sum;
  for (int i = 0; i < PER; i++) {
    const uint32_t c = hist[t * PER + i];
    hist[t * PER + i] = run;
    run += c;
  }
}
__global__ void __launch_bounds__(256) k_wf_scatter(DScene s, WfState w, int qi, uint32_t *offs, uint32_t *sorted) {
  const uint32_t n_in = w.count[qi];
  for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < n_in; q += gridDim.x * blockDim.x) {
    const uint32_t slot = w.queue[qi][q];
    sorted[atomicAdd(offs + raysort_key(s, w.ray_o[slot], w.ray_d[slot]), 1u)] = slot;
  }
}

// ---- resolve: ray_tracing.cu:78-83 + WriteImage's *255 truncation (utils.cu:87-96) ---------------
// In-place use (rgb == accum) is allowed: each thread reads its element before it writes it.
__global__ void k_resolve(const float *accum, int64_t n, int spp_total, float *rgb, uint8_t *rgb8) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float c = accum[i] / float(spp_total);
  c = fminf(fmaxf(c, 0.f), 1.f);  // glm::clamp = min(max(x, lo), hi)
  c = sqrtf(c);
  if (rgb) rgb[i] = c;
  if (rgb8) rgb8[i] = (uint8_t)(c * 255);
}
// The same from the fixed-point sums: the float image is the correctly rounded exact sum.
__global__ void k_resolve_fixed(const long long *__restrict__ fixed, int64_t n, int spp_total, float *accum, float *rgb,
                                uint8_t *rgb8) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float sum = from_fixed(fixed[i]);
  if (accum) accum[i] = sum;
  float c = sum / float(spp_total);
  c = fminf(fmaxf(c, 0.f), 1.f);
  c = sqrtf(c);
  if (rgb) rgb[i] = c;
  if (rgb8) rgb8[i] = (uint8_t)(c * 255);
}
// rt_render's float interface: accum (+)= the sum of this call.
__global__ void k_fixed_to_accum(const long long *__restrict__ fixed, int64_t n, float *accum, int clear) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float sum = from_fixed(fixed[i]);
  accum[i] = clear ? sum : accum[i] + sum;
}

}  // namespace rtc

// =================================================================================================
// C-ABI
// =================================================================================================
using namespace rtc;

namespace {
inline int nblk(size_t n, int b) { return (int)((n + b - 1) / b); }

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

// Traversal of the render kernel, chosen at commit (RT_TRAV overrides it for diagnostics, read once per commit).
//   <= 32 primitives (Cornell, birthday): no tree — box list, every lane tests every box (rt_flat.cuh);
//   otherwise one binary tree, walked by
//     the while-while loop at 80 registers when a ray's walk is short — at most 4096 records, or a tree of SAH cost
//     below 40 (spheres 3.4, bunny 2.2, a 54-parallelogram Cornell box 18: 9 - 25 box tests per ray, L1-resident);
//     the if-if loop at 64 registers with the first stack levels in shared memory when it is long and incoherent
//     (random soups: SAH cost 61 at 6 k triangles, 95 at 20 k, 372 at 1 M; 38 - 275 box tests per ray, latency bound).
// Measured per scene class (profiles/r02_z_small_builder_overhead_inbetween.txt, r02_za_trav_rule.txt): meshes of
// 1 k - 69 k triangles prefer while-while by 16 - 25 % (bunny 10 677 -> 12 815 Msamples/s), soups of 6 k triangles
// and more prefer if-if by 8 - 45 %; the record count alone (round 1's rule: <= 4096) put the bunny on the wrong side.
// Round 1 also kept an 8-wide tree for 4 k - 500 k records; the binary loops are level with it
// (profiles/r02_opq_large_scene_experiments.txt) and the wide tree, its builder and its kernel are gone.
int rtc::default_trav(const rt_ctx *ctx, float sah_cost) {
  int dflt = (ctx->scene.n_recs <= 4096 || sah_cost < 40.0f) ? 1 : 0;
  if (ctx->flat.n > 0) dflt = 20;
  const char *e = getenv("RT_TRAV");
  int v = e ? atoi(e) : dflt;
  if (v != 0 && v != 1 && v != 20) v = dflt;
  if (v == 20 && ctx->flat.n == 0) v = dflt;
  return v;
}

static int mode_to_trav(const rt_ctx *ctx, int mode) {
  switch (mode) {
    case 0: return ctx->trav;
    case 2: return 0;
    case 3: return 1;
    case 4: return 20;
    default: return -1;  // 1: sequential scan
  }
}

extern "C" int rt_trace_closest_device(rt_ctx *ctx, const float *d_rays, int64_t n, double t_from, double t_to, int mode,
                                       rt_hit *d_hits, void *stream) {
  int rc = check_ready(ctx, false, "rt_trace_closest");
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!d_rays || !d_hits)) || (mode < 0 || mode > 4)) {
    set_error("rt_trace_closest: bad argument");
    return RT_ERR_INVALID;
  }
  if (n == 0) return RT_OK;
  RT_CUDA(cudaSetDevice(ctx->device));
  const int trav = mode_to_trav(ctx, mode);
  if (trav == 20 && ctx->flat.n == 0 && ctx->scene.n_recs > 0) {
    set_error("rt_trace_closest: mode 4 (box list) needs a scene of at most 32 primitives");
    return RT_ERR_INVALID;
  }
  const int kmode = trav < 0 ? 1 : (trav == 0 ? 2 : (trav == 1 ? 3 : 4));
  k_trace_closest<<<nblk((size_t)n, 128), 128, 0, (cudaStream_t)stream>>>(ctx->scene, ctx->flat, d_rays, n, t_from, t_to,
                                                                          kmode, d_hits);
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

// Wavefront driver.  The queue length lives on the device; the host reads it back once per BATCH of iterations
// (round 1 read it every iteration: a 4-byte D2H + stream sync between every extend and logic pass), the kernels
// of the iterations launched past the end find an empty queue and return.
static int render_wavefront(rt_ctx *ctx, const RenderArgs &a, bool stats_on, bool sort, bool raysort, RenderCounters *ctr,
                            cudaStream_t st, uint32_t *iterations, uint32_t *launches) {
  size_t n = (size_t)a.height * a.width;
  for (int i = 0; i < 3; i++) RT_CUDA(ctx->wf[i].reserve(n * 16));
  RT_CUDA(ctx->wf[3].reserve(n * 4));
  RT_CUDA(ctx->wf[4].reserve(n * 4));
  RT_CUDA(ctx->wf[5].reserve(n * 4));
  RT_CUDA(ctx->wf[6].reserve(64));
  WfState w;
  w.ray_o = ctx->wf[0].as<float4>();
  w.ray_d = ctx->wf[1].as<float4>();
  w.thr = ctx->wf[2].as<float4>();
  w.sample = ctx->wf[3].as<uint32_t>();
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
  uint32_t *rs_hist = nullptr, *rs_sorted = nullptr;
  if (raysort) {
    RT_CUDA(ctx->wf[8].reserve((size_t)RAYSORT_BUCKETS * 4));
    RT_CUDA(ctx->wf[9].reserve(n * 4));
    rs_hist = ctx->wf[8].as<uint32_t>();
    rs_sorted = ctx->wf[9].as<uint32_t>();
  }
  const int sms = ctx->n_sm;
  // resident blocks per SM the extend kernel is compiled for: long, latency-bound queries want warps, not registers
  const bool big = ctx->scene.n_recs >= 500000;
  RT_CUDA(cudaMemsetAsync(w.count, 0, 64, st));
  // first logic pass reads no queue and fills queue 1
  int qi = 0;
  k_wf_logic<true><<<std::min(nblk(n, 256), sms * 8), 256, 0, st>>>(ctx->scene, ctx->camera, a, w, qi, ctr);
  *launches += 1;
  qi ^= 1;
  const char *eb = getenv("RT_WF_BATCH");
  const int batch = eb ? std::max(1, atoi(eb)) : 8;
  const int gb_e = std::min(nblk(n, 128), sms * 32), gb_l = std::min(nblk(n + 32 * WF_CLASSES, 256), sms * 8);
  uint32_t live = 1;
  while (live) {
    for (int b = 0; b < batch; b++) {
      const uint32_t *queue = w.queue[qi];
      if (raysort) {
        RT_CUDA(cudaMemsetAsync(rs_hist, 0, (size_t)RAYSORT_BUCKETS * 4, st));
        k_wf_hist<<<gb_l, 256, 0, st>>>(ctx->scene, w, qi, rs_hist);
        k_wf_scan<<<1, 1024, 0, st>>>(rs_hist);
        k_wf_scatter<<<gb_l, 256, 0, st>>>(ctx->scene, w, qi, rs_hist, rs_sorted);
        queue = rs_sorted;
        *launches += 3;
      }
      if (stats_on)
        k_wf_extend<true, 5><<<gb_e, 128, 0, st>>>(ctx->scene, w, qi, queue, ctr);
      else if (big)
        k_wf_extend<false, 10><<<gb_e, 128, 0, st>>>(ctx->scene, w, qi, queue, ctr);
      else
        k_wf_extend<false, 5><<<gb_e, 128, 0, st>>>(ctx->scene, w, qi, queue, ctr);
      k_wf_reset<<<1, 32, 0, st>>>(w.count, qi ^ 1, w.mcount);
      k_wf_logic<false><<<gb_l, 256, 0, st>>>(ctx->scene, ctx->camera, a, w, qi, ctr);
      *launches += 3;
      *iterations += 1;
      qi ^= 1;
    }
    RT_CUDA(cudaMemcpyAsync(&live, w.count + qi, 4, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));
  }
  RT_CUDA(cudaGetLastError());
  return RT_OK;
}

// Work distribution of a k_render launch (RenderArgs), sized against a warp's share of the samples.  Measured on the
// B200 (profiles/r02_*; Cornell box 1024 x 1024, ms per sample-per-pixel with every item of the launch alike):
//     4-sample chunks, 1 chunk per item   0.2723  (fixed cost of the launch 0.26 ms)
//    16-sample chunks, 1 chunk per item   0.2672  (0.8 ms: the last items are 512 samples = 0.45 ms of a warp)
//    16-sample chunks, 4 chunks per item  0.2600  (1.7 ms)
// — the per-lane hand-out (~35 warp instructions) and the draw of a new item both cost, and big items leave a tail.
//  * scenes whose cost per ray is BOUNDED BY CONSTRUCTION (the box list: every ray tests the same <= 32 boxes) get
//    three phases: A = 16-sample chunks, rA chunks per item; B = 16-sample chunks, one per item, long enough to
//    absorb the last items of A (3 of them per warp); C = chunks of <= 4 samples, long enough to absorb the last
//    items of B.  Round 2's first plan was A + C with C = 1/4 of the samples;
//  * behind a tree the cost of a sample is unbounded in practice: the bunny frame has 8x4 tiles whose rays need 40x the
//    box tests of the average mesh tile (644 against 14.7 per ray), one such tile as an 80-samples-per-pixel item is
//    22 ms on one warp, and it held the launch for 39 % of its time (profiles/r02_i_bunny_two_phase_ncu_summary.txt:
//    one SM active for 130.7 M cycles, the others for 88-90 M).  Tree scenes are therefore all phase C: 4-sample
//    chunks, one chunk per item (bunny 7 423 -> 10 222; the 1 M soup pays 5 % for the lost tile locality, 494 -> 468);
//  * a launch with little work per lane (spheres 400x225x16: 13 samples) uses 1-sample chunks.
// Diagnostics / tests: RT_PLAN="16x4:64,16x1:48,4x1" (chunk x chunks-per-item : samples, the last phase takes the
// rest); RT_CHUNK / RT_ITEM_CHUNKS: one phase with that chunk size and that many chunks per item over all whole
// items, the remainder with the same chunk size, one chunk per item.
static void plan_work(int n_sm, RenderArgs &a, int minb, bool bounded_cost) {
  const long long lanes = (long long)n_sm * minb * 128, warps = lanes / 32;
  const long long spp = a.spp_end - a.spp_begin, px = (long long)a.height * a.width;
  const long long per_lane = std::max<long long>(1, px * spp / lanes);
  a.tiles_x = (a.width + 7) / 8;
  a.n_tiles = a.tiles_x * ((a.height + 3) / 4);
  {  // RT_TILE_ORDER (experiment): 0 raster (default), 1 golden-ratio stride (heavy image regions spread over a sweep)
    const char *eo = getenv("RT_TILE_ORDER");
    a.tile_stride = 1;
    if (eo && atoi(eo) == 1) {
      auto gcd = [](long long x, long long y) { while (y) { long long t = x % y; x = y; y = t; } return x; };
      long long st = (long long)(a.n_tiles * 0.6180339887) | 1;
      while (st > 1 && gcd(st, a.n_tiles) != 1) st -= 2;
      a.tile_stride = (uint32_t)std::max<long long>(st, 1);
    }
  }
  struct P { long long ch, r, n; };  // n: samples of the phase
  P ph[PLAN_MAX];
  int np = 0;
  auto clampll = [](long long v, long long lo, long long hi) { return std::min(std::max(v, lo), hi); };
  const long long ch_small = clampll(per_lane / 64, 1, 4);
  const char *ep = getenv("RT_PLAN"), *ec = getenv("RT_CHUNK"), *ei = getenv("RT_ITEM_CHUNKS");
  if (ep && *ep) {
    long long left = spp;
    const char *q = ep;
    while (*q && np < PLAN_MAX && left > 0) {
      char *e = nullptr;
      long long ch = strtoll(q, &e, 10), r = 1, n = left;
      if (e == q) break;  // not a number
      if (*e == 'x') r = strtoll(e + 1, &e, 10);
      if (*e == ':') n = clampll(strtoll(e + 1, &e, 10), 0, left);
      q = *e == ',' ? e + 1 : e;
      if (*q == 0 || np == PLAN_MAX - 1) n = left;  // the last phase takes the rest
      if (n > 0) {
        ph[np++] = {clampll(ch, 1, 64), clampll(r, 1, 64), n};
        left -= n;
      }
    }
    if (left > 0) {
      if (np == 0) ph[np++] = {ch_small, 1, left};
      else ph[np - 1].n += left;
    }
  } else if (ec) {
    const long long ch = std::min(clampll(atoll(ec), 1, 16), std::max<long long>(spp, 1)), r = clampll(ei ? atoll(ei) : 1, 1, 16);
    const long long n1 = r > 1 ? spp / (ch * r) * (ch * r) : 0;
    if (n1 > 0) ph[np++] = {ch, r, n1};
    if (spp - n1 > 0) ph[np++] = {ch, 1, spp - n1};
  } else if (bounded_cost && per_lane >= 512 && spp >= 64) {
    // Geometric cascade: 16-sample chunks in items of r, r/4, 1 chunks, then chunks of ch_small.  A phase takes as
    // many whole sweeps of the image as leave, after it, the samples that absorb its tail: 3 of its items per warp
    // (spp_for).  r starts at 1/12 of a warp's share of the launch (<= 32).  Cornell 1024 spp: 16x32:512, 16x8:384,
    // 16x1:112, 4x1:16 — the best of the plans measured (profiles/r02_zb_plan_sweep.txt), 262.2 ms against 266.8 ms
    // for round 2's first plan (16x9:720, 4x1:304); at 128 spp (the 8-GPU share) 16x6:96, 16x1:16, 4x1:16.
    auto spp_for = [&](long long s) { return (s * warps + px - 1) / px; };  // samples per pixel = s samples per warp
    const long long nC = std::max<long long>(2 * ch_small, (spp_for(3 * 512) + ch_small - 1) / ch_small * ch_small);
    long long left = spp, r = clampll(per_lane / 192, 1, 32);
    while (np < PLAN_MAX - 1 && left > nC) {
      const long long reserve = r == 1 ? nC : std::max(nC, spp_for(3 * 512 * r));
      const long long n = left > reserve ? (left - reserve) / (16 * r) * (16 * r) : 0;
      if (n > 0) {
        ph[np++] = {16, r, n};
        left -= n;
        if (r == 1) break;
        r = r >= 16 ? r / 4 : 1;
      } else {
        if (r == 1) break;
        r = r > 2 ? r / 2 : 1;  // the sweep does not fit: smaller items
      }
    }
    ph[np++] = {ch_small, 1, left};
  } else {
    ph[np++] = {std::min<long long>(ch_small, std::max<long long>(spp, 1)), 1, spp};
  }
  long long k = a.spp_begin, item = 0;
  a.n_phases = np;
  for (int i = 0; i < np; i++) {
    a.phase[i].item_begin = (uint32_t)item;
    a.phase[i].k_begin = (int32_t)k;
    a.phase[i].ch = (int32_t)ph[i].ch;
    a.phase[i].r = (int32_t)ph[i].r;
    const long long per_item = ph[i].ch * ph[i].r;
    item += (ph[i].n + per_item - 1) / per_item * a.n_tiles;
    k += ph[i].n;
  }
  a.phase[np].item_begin = (uint32_t)item;
  a.phase[np].k_begin = a.spp_end;
  a.phase[np].ch = a.phase[np].r = 1;
  a.n_items = (uint32_t)item;
}

// The plan as data (diagnostics and the CPU tests of the plan arithmetic; no device needed): what a k_render launch of
// this shape would be handed.  phases[i] = {first item, first sample, samples per chunk, chunks per item}; entry
// n_phases closes the last phase ({items of the launch, spp_begin + spp_count, -, -}).
extern "C" int rt_work_plan(int n_sm, int blocks_per_sm, int height, int width, int spp_begin, int spp_count,
                            int bounded_cost, int32_t *phases /* [5][4] */, int32_t *n_tiles) {
  if (n_sm <= 0 || blocks_per_sm <= 0 || height <= 0 || width <= 0 || spp_begin < 0 || spp_count <= 0 || !phases ||
      (int64_t)height * width > 0x3fffffff || (int64_t)spp_begin + spp_count > 0x7fffffff) {
    set_error("rt_work_plan: bad argument");
    return RT_ERR_INVALID;
  }
  RenderArgs a;
  memset(&a, 0, sizeof(a));
  a.height = height;
  a.width = width;
  a.spp_begin = spp_begin;
  a.spp_end = spp_begin + spp_count;
  plan_work(n_sm, a, blocks_per_sm, bounded_cost != 0);
  for (int i = 0; i <= PLAN_MAX; i++) {
    const bool used = i <= a.n_phases;
    phases[4 * i + 0] = used ? (int32_t)a.phase[i].item_begin : 0;
    phases[4 * i + 1] = used ? a.phase[i].k_begin : 0;
    phases[4 * i + 2] = used ? a.phase[i].ch : 0;
    phases[4 * i + 3] = used ? a.phase[i].r : 0;
  }
  if (n_tiles) *n_tiles = a.n_tiles;
  return a.n_phases;
}

// Renders the samples of `p` into the fixed-point sums `fixed` (added).
static int render_into_fixed(rt_ctx *ctx, const rt_render_params *p, long long *fixed, rt_render_stats *stats,
                             cudaStream_t st, bool record_begin) {
  RenderArgs a;
  memset(&a, 0, sizeof(a));
  a.height = p->height;
  a.width = p->width;
  a.spp_begin = p->spp_begin;
  a.spp_end = p->spp_begin + p->spp_count;
  a.max_depth = p->max_depth;
  a.k0 = (uint32_t)p->seed;
  a.k1 = (uint32_t)(p->seed >> 32);
  a.fixed = fixed;
  RT_CUDA(ctx->counters.reserve(sizeof(RenderCounters) + 64));
  RenderCounters *ctr = ctx->counters.as<RenderCounters>();
  a.work_counter = reinterpret_cast<unsigned int *>(ctr + 1);
  RT_CUDA(cudaMemsetAsync(ctr, 0, sizeof(RenderCounters) + 64, st));
  const bool stats_on = (p->flags & RT_RENDER_PROFILE) != 0;
  bool fused = (p->flags & RT_RENDER_FUSED) != 0;
  if (!(p->flags & (RT_RENDER_FUSED | RT_RENDER_WAVEFRONT))) fused = !ctx->opt_wavefront;
  uint32_t iterations = 0, launches = 0;
  if (record_begin) RT_CUDA(cudaEventRecord(ctx->ev0, st));
  if (p->spp_count > 0 && fused) {
    const int trav = ctx->trav;
    // resident blocks per SM the kernel is compiled for (register budget): 6 (80 registers) for the cache-resident
    // scenes (box list, while-while), 8 (64 registers) for the if-if loop of large scenes, which is latency bound
    const int minb = trav == 0 ? 8 : (trav == 1 ? RT_WW_MINB : 6);
    plan_work(ctx->n_sm, a, minb, trav == 20);
    const int grid = ctx->n_sm * minb;
#define RT_LAUNCH(T, M, ...)                                                                             \
  if (stats_on && sph)                                                                                   \
    k_render<true, T, M, true><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, ctr, ##__VA_ARGS__);    \
  else if (stats_on)                                                                                     \
    k_render<true, T, M, false><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, ctr, ##__VA_ARGS__);   \
  else if (sph)                                                                                          \
    k_render<false, T, M, true><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, ctr, ##__VA_ARGS__);   \
  else                                                                                                   \
    k_render<false, T, M, false><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, ctr, ##__VA_ARGS__);
    const bool sph = ctx->scene.has_spheres != 0;  // scenes without a Sphere run kernels without the FP64 test
    if (trav == 20) {
      RT_LAUNCH(20, 6, ctx->flat)
    } else if (trav == 1) {
      RT_LAUNCH(1, RT_WW_MINB)
    } else {
      RT_LAUNCH(0, 8)
    }
#undef RT_LAUNCH
    RT_CUDA(cudaGetLastError());
    launches = 1;
  } else if (p->spp_count > 0) {
    int rc = render_wavefront(ctx, a, stats_on, (p->flags & RT_RENDER_SORT) != 0, (p->flags & RT_RENDER_RAYSORT) != 0, ctr,
                              st, &iterations, &launches);
    if (rc) return rc;
  }
  ctx->last_iterations = iterations;
  ctx->last_launches = launches;
  ctx->last_chunk = a.phase[0].ch;
  ctx->last_items = a.n_items;
  (void)stats;
  return RT_OK;
}

static int finish_stats(rt_ctx *ctx, rt_render_stats *stats, cudaStream_t st, uint32_t extra_launches) {
  RT_CUDA(cudaEventRecord(ctx->ev1, st));
  if (!stats) return RT_OK;
  RenderCounters hc;
  RT_CUDA(cudaMemcpyAsync(&hc, ctx->counters.p, sizeof(hc), cudaMemcpyDeviceToHost, st));
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
  stats->iterations = ctx->last_iterations;
  stats->kernel_launches = ctx->last_launches + extra_launches;
  stats->render_ms = ms;
  stats->chunk_samples = (uint32_t)ctx->last_chunk;
  stats->work_items = ctx->last_items;
  stats->seq_fallbacks = (uint32_t)std::min<unsigned long long>(hc.seq_fallbacks, 0xffffffffull);
  stats->max_box_tests_per_ray = (uint32_t)hc.max_nodes;
  return RT_OK;
}

static int check_params(rt_ctx *ctx, const rt_render_params *p, const void *buf, const char *who) {
  int rc = check_ready(ctx, true, who);
  if (rc) return rc;
  if (!p || !buf || p->height <= 0 || p->width <= 0 || p->spp_count < 0 || p->spp_begin < 0 || p->max_depth < 0 ||
      (int64_t)p->height * p->width > 0x3fffffff || (int64_t)p->spp_begin + p->spp_count > 0x7fffffff) {
    set_error(std::string(who) + ": bad parameters");
    return RT_ERR_INVALID;
  }
  return RT_OK;
}

extern "C" int rt_render_fixed(rt_ctx *ctx, const rt_render_params *p, int64_t *d_accum_fixed, rt_render_stats *stats,
                               void *stream) {
  int rc = check_params(ctx, p, d_accum_fixed, "rt_render_fixed");
  if (rc) return rc;
  RT_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  const size_t n = (size_t)p->height * p->width;
  RT_CUDA(cudaEventRecord(ctx->ev0, st));
  if (p->flags & RT_RENDER_CLEAR) RT_CUDA(cudaMemsetAsync(d_accum_fixed, 0, n * 24, st));
  rc = render_into_fixed(ctx, p, reinterpret_cast<long long *>(d_accum_fixed), stats, st, false);
  if (rc) return rc;
  return finish_stats(ctx, stats, st, 0);
}

extern "C" int rt_render(rt_ctx *ctx, const rt_render_params *p, float *d_accum_rgb, rt_render_stats *stats,
                         void *stream) {
  int rc = check_params(ctx, p, d_accum_rgb, "rt_render");
  if (rc) return rc;
  RT_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;
  const size_t n = (size_t)p->height * p->width;
  RT_CUDA(ctx->fixed_tmp.reserve(n * 24));
  RT_CUDA(cudaEventRecord(ctx->ev0, st));
  RT_CUDA(cudaMemsetAsync(ctx->fixed_tmp.p, 0, n * 24, st));
  rc = render_into_fixed(ctx, p, ctx->fixed_tmp.as<long long>(), stats, st, false);
  if (rc) return rc;
  k_fixed_to_accum<<<nblk(n * 3, 256), 256, 0, st>>>(ctx->fixed_tmp.as<long long>(), (int64_t)n * 3, d_accum_rgb,
                                                     (p->flags & RT_RENDER_CLEAR) ? 1 : 0);
  RT_CUDA(cudaGetLastError());
  return finish_stats(ctx, stats, st, 1);
}

// ---- rt_render_static: rt_render without any allocation ------------------------------------------------------------
// The scratch of rt_render (the exact int64 sums of the call + the counters / work counter of the kernel) lives in
// one file-scope device array instead of the context's cudaMalloc'ed buffers.  Every GPU that loads the library
// holds the array (6 MiB) until exit.  In host code `g_static_scratch` is only a stand-in for the symbol: the device
// address comes from cudaGetSymbolAddress, resolved once per device at rt_ctx_create (static_scratch_setup).
constexpr size_t STATIC_SUMS_BYTES = (size_t)RT_STATIC_MAX_PIXELS * 3 * sizeof(long long);
constexpr size_t STATIC_CTR_BYTES = 256;  // RenderCounters (64 B) + the work counter, as render_into_fixed lays them out
static __device__ __align__(256) unsigned char g_static_scratch[STATIC_SUMS_BYTES + STATIC_CTR_BYTES];
static_assert(sizeof(RenderCounters) + 64 <= STATIC_CTR_BYTES, "counter block");

namespace {
constexpr int STATIC_MAX_DEVICES = 64;
void *g_static_ptr[STATIC_MAX_DEVICES] = {};
cudaError_t g_static_err[STATIC_MAX_DEVICES] = {};
bool g_static_done[STATIC_MAX_DEVICES] = {};
}  // namespace

namespace rtc {
// Called with `device` current (rt_ctx_create).  A failure is remembered and reported by rt_render_static; it does
// not fail the context.
void static_scratch_setup(int device) {
  if (device < 0 || device >= STATIC_MAX_DEVICES || g_static_done[device]) return;
  void *ptr = nullptr;
  cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);
  if (e != cudaSuccess) cudaGetLastError();
  g_static_ptr[device] = e == cudaSuccess ? ptr : nullptr;
  g_static_err[device] = e;
  g_static_done[device] = true;
}
}  // namespace rtc

extern "C" int rt_render_static(rt_ctx *ctx, const rt_render_params *p, float *d_accum_rgb, void *stream) {
  int rc = check_params(ctx, p, d_accum_rgb, "rt_render_static");
  if (rc) return rc;
  const size_t n = (size_t)p->height * p->width;
  if (n > (size_t)RT_STATIC_MAX_PIXELS) {
    set_error("rt_render_static: height * width = " + std::to_string(n) + " exceeds RT_STATIC_MAX_PIXELS = " +
              std::to_string((long long)RT_STATIC_MAX_PIXELS) + " (the size of the static scratch array); use rt_render");
    return RT_ERR_LIMIT;
  }
  if (p->flags & (RT_RENDER_WAVEFRONT | RT_RENDER_PROFILE | RT_RENDER_SORT | RT_RENDER_RAYSORT)) {
    set_error("rt_render_static: the wavefront pipeline and the profile counters need allocated buffers; use rt_render");
    return RT_ERR_INVALID;
  }
  const int dev = ctx->device;
  if (dev < 0 || dev >= STATIC_MAX_DEVICES || !g_static_done[dev] || !g_static_ptr[dev]) {
    set_error(std::string("rt_render_static: no address for the static scratch array on this device (") +
              (dev >= 0 && dev < STATIC_MAX_DEVICES && g_static_done[dev] ? cudaGetErrorString(g_static_err[dev])
                                                                           : "device not set up") + ")");
    return RT_ERR_CUDA;
  }
  RT_CUDA(cudaSetDevice(dev));
  cudaStream_t st = (cudaStream_t)stream;
  unsigned char *base = static_cast<unsigned char *>(g_static_ptr[dev]);
  long long *sums = reinterpret_cast<long long *>(base);
  RenderCounters *ctr = reinterpret_cast<RenderCounters *>(base + STATIC_SUMS_BYTES);
  // the array keeps the last call's contents: the sums of this call and the counters start at zero
  RT_CUDA(cudaMemsetAsync(sums, 0, n * 24, st));
  RT_CUDA(cudaMemsetAsync(ctr, 0, sizeof(RenderCounters) + 64, st));
  if (p->spp_count > 0) {
    RenderArgs a;
    memset(&a, 0, sizeof(a));
    a.height = p->height;
    a.width = p->width;
    a.spp_begin = p->spp_begin;
    a.spp_end = p->spp_begin + p->spp_count;
    a.max_depth = p->max_depth;
    a.k0 = (uint32_t)p->seed;
    a.k1 = (uint32_t)(p->seed >> 32);
    a.fixed = sums;
    a.work_counter = reinterpret_cast<unsigned int *>(ctr + 1);
    const int trav = ctx->trav;
    const int minb = trav == 0 ? 8 : (trav == 1 ? RT_WW_MINB : 6);
    plan_work(ctx->n_sm, a, minb, trav == 20);
    const int grid = ctx->n_sm * minb;
    const bool sph = ctx->scene.has_spheres != 0;
#define RT_LAUNCH_STATIC(T, M, ...)                                                                      \
  if (sph)                                                                                               \
    k_render<false, T, M, true><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, ctr, ##__VA_ARGS__);   \
  else                                                                                                   \
    k_render<false, T, M, false><<<grid, 128, 0, st>>>(ctx->scene, ctx->camera, a, ctr, ##__VA_ARGS__);
    if (trav == 20) {
      RT_LAUNCH_STATIC(20, 6, ctx->flat)
    } else if (trav == 1) {
      RT_LAUNCH_STATIC(1, RT_WW_MINB)
    } else {
      RT_LAUNCH_STATIC(0, 8)
    }
#undef RT_LAUNCH_STATIC
    RT_CUDA(cudaGetLastError());
  }
  k_fixed_to_accum<<<nblk(n * 3, 256), 256, 0, st>>>(sums, (int64_t)n * 3, d_accum_rgb,
                                                     (p->flags & RT_RENDER_CLEAR) ? 1 : 0);
  RT_CUDA(cudaGetLastError());
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

extern "C" int rt_resolve_fixed(rt_ctx *ctx, const int64_t *d_accum_fixed, int height, int width, int spp_total,
                                float *d_accum_rgb, float *d_rgb, uint8_t *d_rgb8, void *stream) {
  if (!ctx || !d_accum_fixed || height <= 0 || width <= 0 || spp_total <= 0) {
    set_error("rt_resolve_fixed: bad argument");
    return RT_ERR_INVALID;
  }
  RT_CUDA(cudaSetDevice(ctx->device));
  int64_t n = (int64_t)height * width * 3;
  k_resolve_fixed<<<nblk((size_t)n, 256), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const long long *>(d_accum_fixed), n, spp_total, d_accum_rgb, d_rgb, d_rgb8);
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
