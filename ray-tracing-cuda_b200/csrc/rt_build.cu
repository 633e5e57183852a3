// rt_build.cu — rt_scene_commit: flattening of the scene description into type-tagged records and
// the device-side LBVH build (Morton keys -> radix sort -> Karras hierarchy -> bottom-up refit with
// SAH leaf collapse) emitting 32-byte nodes.
//
// Replaces, in the reference: scene construction by device-side `new` + HitableList::Append
// (hitable_list.cu:27-29), the single-thread BVHNode constructor with its sequential thrust::sort
// (bvh.cuh:113-121), Face::GetBV (bvh.cuh:71-82) and the virtual object graph as a whole.
#include <cub/cub.cuh>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "rt_internal.h"

namespace rtc {

constexpr int kRefLeafMin = 2048;  // bvh.cuh:105: ranges of <= kMin faces are never sorted

__device__ __forceinline__ uint32_t float_orderable(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// ---- mesh faces -> records ---------------------------------------------------------------------
// Sort key of Face::operator< (bvh.cuh:96-98): positions_[0].x, with -0 == +0 as `<` sees them.
__global__ void k_face_keys(const float *faces, int stride, int n, uint32_t *keys, uint32_t *vals) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float x = faces[(size_t)i * stride];
  if (x == 0.0f) x = 0.0f;
  keys[i] = float_orderable(x);
  vals[i] = i;
}

__global__ void k_face_records(const float *faces, int stride, const uint32_t *perm, int n, int obj, uint32_t word,
                               float4 *recs, float *uv) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int src = perm ? (int)perm[i] : i;
  const float *f = faces + (size_t)src * stride;
  V3 p0 = mk(f[0], f[1], f[2]), p1 = mk(f[3], f[4], f[5]), p2 = mk(f[6], f[7], f[8]);
  V3 e1 = p1 - p0, e2 = p2 - p0;  // utils.cu:54-55
  recs[3 * (size_t)i + 0] = make_float4(p0.x, p0.y, p0.z, __int_as_float(obj));
  recs[3 * (size_t)i + 1] = make_float4(e1.x, e1.y, e1.z, __int_as_float(i));
  recs[3 * (size_t)i + 2] = make_float4(e2.x, e2.y, e2.z, __uint_as_float(word));
  if (uv) {
    for (int k = 0; k < 6; k++) uv[6 * (size_t)i + k] = stride == 15 ? f[9 + k] : 0.f;
  }
}

// ---- build primitives ----------------------------------------------------------------------------
// Every step of the build is a per-element body d_*(i, ...) with two callers: the grid-wide kernel k_* below it (one
// launch per step) and k_build_small, which runs ALL steps of a small scene in one block (one launch per commit).
__device__ __forceinline__ uint32_t d_head_flag(const float4 *recs, int i) {
  uint32_t kind = __float_as_uint(recs[3 * (size_t)i + 2].w) & KIND_MASK;
  return (kind == K_TRI || kind == K_QUAD0 || kind == K_SPHERE) ? 1u : 0u;
}
__global__ void k_head_flags(const float4 *recs, int n, uint32_t *flags) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  flags[i] = d_head_flag(recs, i);
}
__global__ void k_scatter_heads(const uint32_t *flags, const uint32_t *pos, int n, uint32_t *bp_first) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (flags[i]) bp_first[pos[i]] = i;
}

__device__ __forceinline__ void grow(V3 &lo, V3 &hi, V3 p) {
  lo = mk(fminf(lo.x, p.x), fminf(lo.y, p.y), fminf(lo.z, p.z));
  hi = mk(fmaxf(hi.x, p.x), fmaxf(hi.y, p.y), fmaxf(hi.z, p.z));
}
__device__ __forceinline__ void rec_bounds(const float4 *recs, size_t k, V3 &lo, V3 &hi) {
  float4 r0 = recs[3 * k], r1 = recs[3 * k + 1], r2 = recs[3 * k + 2];
  uint32_t kind = __float_as_uint(r2.w) & KIND_MASK;
  if (kind == K_SPHERE) {
    float rad = fabsf(__double2float_ru(fabs(rec_radius(r1))));
    rad = rad + rad * 1e-5f;
    grow(lo, hi, xyz(r0) - mk(rad, rad, rad));
    grow(lo, hi, xyz(r0) + mk(rad, rad, rad));
  } else {
    grow(lo, hi, xyz(r0));
    grow(lo, hi, xyz(r0) + xyz(r1));
    grow(lo, hi, xyz(r0) + xyz(r2));
  }
}

// Padded box per build primitive + centroid bounds of the scene (ordered-int atomics).
__device__ __forceinline__ void d_bp_bounds(int i, const float4 *recs, const uint32_t *bp_first, float4 *bp_lo,
                                            float4 *bp_hi, uint32_t *scene_bounds) {
  size_t k = bp_first[i];
  V3 lo = mk(CUDART_INF_F, CUDART_INF_F, CUDART_INF_F), hi = mk(-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F);
  rec_bounds(recs, k, lo, hi);
  uint32_t kind = __float_as_uint(recs[3 * k + 2].w) & KIND_MASK;
  int nrec = 1;
  if (kind == K_QUAD0) {
    rec_bounds(recs, k + 1, lo, hi);
    nrec = 2;
  }
  // guard band: MT accepts rays within rounding of the edges; culling must stay conservative
  float m = fmaxf(fmaxf(hi.x - lo.x, hi.y - lo.y), hi.z - lo.z);
  m = fmaxf(m, fmaxf(fmaxf(fabsf(lo.x), fabsf(lo.y)), fabsf(lo.z)));
  m = fmaxf(m, fmaxf(fmaxf(fabsf(hi.x), fabsf(hi.y)), fabsf(hi.z)));
  float pad = m * 2e-5f + 1e-30f;
  lo = lo - mk(pad, pad, pad);
  hi = hi + mk(pad, pad, pad);
  bp_lo[i] = make_float4(lo.x, lo.y, lo.z, __int_as_float(nrec));
  bp_hi[i] = make_float4(hi.x, hi.y, hi.z, 0.f);
  V3 c = (lo + hi) * 0.5f;
  atomicMin(scene_bounds + 0, float_orderable(c.x));
  atomicMin(scene_bounds + 1, float_orderable(c.y));
  atomicMin(scene_bounds + 2, float_orderable(c.z));
  atomicMax(scene_bounds + 3, float_orderable(c.x));
  atomicMax(scene_bounds + 4, float_orderable(c.y));
  atomicMax(scene_bounds + 5, float_orderable(c.z));
}
__global__ void k_bp_bounds(const float4 *recs, const uint32_t *bp_first, int n_bp, float4 *bp_lo, float4 *bp_hi,
                            uint32_t *scene_bounds) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_bp) return;
  d_bp_bounds(i, recs, bp_first, bp_lo, bp_hi, scene_bounds);
}

__device__ __forceinline__ float orderable_float(uint32_t u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}
__device__ __forceinline__ uint64_t expand21(uint64_t v) {
  v &= 0x1fffffull;
  v = (v | v << 32) & 0x1f00000000ffffull;
  v = (v | v << 16) & 0x1f0000ff0000ffull;
  v = (v | v << 8) & 0x100f00f00f00f00full;
  v = (v | v << 4) & 0x10c30c30c30c30c3ull;
  v = (v | v << 2) & 0x1249249249249249ull;
  return v;
}
__device__ __forceinline__ uint64_t d_morton(int i, const float4 *bp_lo, const float4 *bp_hi, const uint32_t *scene_bounds) {
  float4 lo = bp_lo[i], hi = bp_hi[i];
  float c[3] = {(lo.x + hi.x) * 0.5f, (lo.y + hi.y) * 0.5f, (lo.z + hi.z) * 0.5f};
  uint64_t q[3];
  for (int a = 0; a < 3; a++) {
    float mn = orderable_float(scene_bounds[a]), mx = orderable_float(scene_bounds[3 + a]);
    double ext = (double)mx - (double)mn;
    double x = ext > 0 ? ((double)c[a] - (double)mn) / ext : 0.0;
    x = fmin(fmax(x * 2097152.0, 0.0), 2097151.0);
    q[a] = (uint64_t)x;
  }
  return expand21(q[0]) << 2 | expand21(q[1]) << 1 | expand21(q[2]);
}
__global__ void k_morton(const float4 *bp_lo, const float4 *bp_hi, int n_bp, const uint32_t *scene_bounds,
                         uint64_t *keys, uint32_t *vals) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_bp) return;
  keys[i] = d_morton(i, bp_lo, bp_hi, scene_bounds);
  vals[i] = i;
}

__global__ void k_sorted_nrec(const float4 *bp_lo, const uint32_t *order, int n_bp, uint32_t *nrec) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_bp) return;
  nrec[j] = (uint32_t)__float_as_int(bp_lo[order[j]].w);
}

// Records copied into BVH (Morton) order; single-primitive leaf nodes written; per-leaf SAH data.
__device__ __forceinline__ void d_emit(int j, const float4 *recs_seq, const float *uv_seq, const uint32_t *bp_first,
                                       const float4 *bp_lo, const float4 *bp_hi, const uint32_t *order,
                                       const uint32_t *off, int n_bp, float4 *recs, float *uv, float4 *nodes,
                                       float *cost, int *depth) {
  uint32_t bp = order[j];
  size_t src = bp_first[bp];
  uint32_t dst = off[j], n = off[j + 1] - off[j];
  for (uint32_t r = 0; r < n; r++) {
    for (int q = 0; q < 3; q++) recs[3 * (size_t)(dst + r) + q] = recs_seq[3 * (src + r) + q];
    if (uv)
      for (int q = 0; q < 6; q++) uv[6 * (size_t)(dst + r) + q] = uv_seq[6 * (src + r) + q];
  }
  size_t node = (size_t)(n_bp - 1) + j;
  float4 lo = bp_lo[bp], hi = bp_hi[bp];
  nodes[2 * node] = make_float4(lo.x, lo.y, lo.z, __uint_as_float(dst));
  nodes[2 * node + 1] = make_float4(hi.x, hi.y, hi.z, __uint_as_float(LEAF_FLAG | n));
  cost[node] = (float)n;
  depth[node] = 1;
}
__global__ void k_emit(const float4 *recs_seq, const float *uv_seq, const uint32_t *bp_first, const float4 *bp_lo,
                       const float4 *bp_hi, const uint32_t *order, const uint32_t *off, int n_bp, float4 *recs,
                       float *uv, float4 *nodes, float *cost, int *depth) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_bp) return;
  d_emit(j, recs_seq, uv_seq, bp_first, bp_lo, bp_hi, order, off, n_bp, recs, uv, nodes, cost, depth);
}

// ---- Karras 2012: one thread per internal node ----------------------------------------------------
__device__ __forceinline__ int lbvh_delta(const uint64_t *keys, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  uint64_t a = keys[i], b = keys[j];
  if (a == b) return 64 + __clz(i ^ j);
  return __clzll(a ^ b);
}
__device__ __forceinline__ void d_hierarchy(int i, const uint64_t *keys, int n, uint32_t *left, uint32_t *right,
                                            uint32_t *parent, uint2 *range) {
  int d = lbvh_delta(keys, n, i, i + 1) - lbvh_delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
  int dmin = lbvh_delta(keys, n, i, i - d);
  int lmax = 2;
  while (lbvh_delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
  int l = 0;
  for (int t = lmax / 2; t >= 1; t /= 2)
    if (lbvh_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
  int j = i + l * d;
  int dnode = lbvh_delta(keys, n, i, j);
  int s = 0, t = l;
  do {
    t = (t + 1) / 2;
    if (lbvh_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
  } while (t > 1);
  int gamma = i + s * d + min(d, 0);
  int lo = min(i, j), hi = max(i, j);
  uint32_t lc = (lo == gamma) ? (uint32_t)(n - 1 + gamma) : (uint32_t)gamma;
  uint32_t rc = (hi == gamma + 1) ? (uint32_t)(n - 1 + gamma + 1) : (uint32_t)(gamma + 1);
  left[i] = lc;
  right[i] = rc;
  parent[lc] = i;
  parent[rc] = i;
  range[i] = make_uint2((uint32_t)lo, (uint32_t)hi);
  if (i == 0) parent[0] = 0xffffffffu;
}
__global__ void k_hierarchy(const uint64_t *keys, int n, uint32_t *left, uint32_t *right, uint32_t *parent,
                            uint2 *range) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  d_hierarchy(i, keys, n, left, right, parent, range);
}

__device__ __forceinline__ float half_area(float4 lo, float4 hi) {
  float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
  return dx * dy + dy * dz + dz * dx;
}

// Bottom-up refit: the second thread to reach a node owns it.  Writes the node's box and words;
// collapses the subtree into a leaf when the SAH says so (records of a subtree are contiguous).
__device__ __forceinline__ void d_refit(int j, const uint32_t *left, const uint32_t *right, const uint32_t *parent,
                                        const uint2 *range, const uint32_t *off, int n, int max_leaf, float c_trav,
                                        uint32_t *visits, float4 *nodes, float *cost, int *depth) {
  uint32_t cur = parent[(size_t)(n - 1) + j];
  while (cur != 0xffffffffu) {
    __threadfence();
    if (atomicAdd(visits + cur, 1u) == 0) return;
    __threadfence();
    uint32_t lc = left[cur], rc = right[cur];
    float4 l0 = __ldcg(nodes + 2 * (size_t)lc), l1 = __ldcg(nodes + 2 * (size_t)lc + 1);
    float4 r0 = __ldcg(nodes + 2 * (size_t)rc), r1 = __ldcg(nodes + 2 * (size_t)rc + 1);
    float4 lo = make_float4(fminf(l0.x, r0.x), fminf(l0.y, r0.y), fminf(l0.z, r0.z), 0.f);
    float4 hi = make_float4(fmaxf(l1.x, r1.x), fmaxf(l1.y, r1.y), fmaxf(l1.z, r1.z), 0.f);
    float a = half_area(lo, hi);
    float cl = __ldcg(cost + lc), cr = __ldcg(cost + rc);
    float c_split = a > 0.f ? c_trav + (half_area(l0, l1) * cl + half_area(r0, r1) * cr) / a : c_trav + cl + cr;
    uint2 rg = range[cur];
    uint32_t first = off[rg.x], nrec = off[rg.y + 1] - first;
    float c_leaf = (float)nrec;
    bool collapse = nrec <= (uint32_t)max_leaf && c_leaf <= c_split;
    if (collapse) {
      lo.w = __uint_as_float(first);
      hi.w = __uint_as_float(LEAF_FLAG | nrec);
      cost[cur] = c_leaf;
      depth[cur] = 1;
    } else {
      lo.w = __uint_as_float(lc);
      hi.w = __uint_as_float(rc);
      cost[cur] = c_split;
      depth[cur] = 1 + max(__ldcg(depth + lc), __ldcg(depth + rc));
    }
    nodes[2 * (size_t)cur] = lo;
    nodes[2 * (size_t)cur + 1] = hi;
    cur = parent[cur];
  }
}
__global__ void k_refit(const uint32_t *left, const uint32_t *right, const uint32_t *parent, const uint2 *range,
                        const uint32_t *off, int n, int max_leaf, float c_trav, uint32_t *visits, float4 *nodes,
                        float *cost, int *depth) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  d_refit(j, left, right, parent, range, off, n, max_leaf, c_trav, visits, nodes, cost, depth);
}

// ---- small scenes: the whole build in ONE launch ----------------------------------------------------------------------
// A scene of at most SMALL_MAX_SEQ records / SMALL_MAX_BP build primitives (the Cornell box, birthday, the 489 spheres of
// scenes/spheres.cu) is built by one block of 1024 threads that runs the steps above back to back with a block barrier
// in between — head flags + scan, boxes, Morton keys, a block-wide radix sort (stable, like the device-wide sort: equal
// keys keep Append order), record offsets, leaf emission, Karras hierarchy, refit — and leaves the root words, SAH cost
// and depth in `result` for ONE read-back.  Same bodies, same arithmetic: the tree is the tree of the step-per-launch
// path bit for bit (tests/test_gpu_parity.py::test_small_scene_builder_identical).  The step-per-launch path costs
// ~15 launches and three host round trips, which on the spheres config outlasted the 0.8 ms frame it prepares.
constexpr int SMALL_T = 1024, SMALL_MAX_BP = SMALL_T, SMALL_MAX_SEQ = 2 * SMALL_T;
struct SmallBuild {
  const float4 *recs_seq;
  const float *uv_seq;
  int n_seq, n_bp, max_leaf;
  float c_trav;
  uint32_t *bp_first, *order, *off, *left, *right, *parent, *visits;
  float4 *bp_lo, *bp_hi, *recs, *nodes;
  uint64_t *keys;
  uint2 *range;
  float *uv, *cost;
  int *depth;
  uint32_t *result;  // [0..7] root node, [8] SAH cost (float bits), [9] depth, [10] primitives found, [11] records emitted
};
__global__ void __launch_bounds__(SMALL_T) k_build_small(SmallBuild a) {
  using Scan = cub::BlockScan<uint32_t, SMALL_T>;
  using Sort = cub::BlockRadixSort<uint64_t, SMALL_T, 1, uint32_t>;
  __shared__ union {
    typename Scan::TempStorage scan;
    typename Sort::TempStorage sort;
  } tmp;
  __shared__ uint32_t s_bounds[6];
  const int t = threadIdx.x, n_bp = a.n_bp;
  // build primitives: the records that start one (two consecutive records per thread)
  uint32_t flag[2], pos[2], found = 0;
  for (int q = 0; q < 2; q++) flag[q] = 2 * t + q < a.n_seq ? d_head_flag(a.recs_seq, 2 * t + q) : 0u;
  Scan(tmp.scan).ExclusiveSum(flag, pos, found);
  if (found != (uint32_t)n_bp) {  // block-uniform; the host reports it (never seen: both counts come from one description)
    if (t == 0) a.result[10] = found, a.result[11] = 0u;
    return;
  }
  for (int q = 0; q < 2; q++)
    if (flag[q]) a.bp_first[pos[q]] = 2 * t + q;
  if (t < 6) s_bounds[t] = t < 3 ? 0xffffffffu : 0u;
  __syncthreads();
  if (t < n_bp) d_bp_bounds(t, a.recs_seq, a.bp_first, a.bp_lo, a.bp_hi, s_bounds);
  __syncthreads();
  // Morton order; the padding keys of the idle threads sort behind every real key (< 2^63)
  uint64_t key[1] = {~0ull};
  uint32_t val[1] = {0xffffffffu};
  if (t < n_bp) {
    key[0] = d_morton(t, a.bp_lo, a.bp_hi, s_bounds);
    val[0] = (uint32_t)t;
  }
  Sort(tmp.sort).Sort(key, val);
  __syncthreads();
  uint32_t nrec = 0, first = 0, total = 0;
  if (t < n_bp) {
    a.keys[t] = key[0];
    a.order[t] = val[0];
    nrec = (uint32_t)__float_as_int(a.bp_lo[val[0]].w);
  }
  Scan(tmp.scan).ExclusiveSum(nrec, first, total);
  if (t < n_bp) a.off[t] = first;
  if (t == 0) a.off[n_bp] = total;
  __syncthreads();
  if (t < n_bp)
    d_emit(t, a.recs_seq, a.uv_seq, a.bp_first, a.bp_lo, a.bp_hi, a.order, a.off, n_bp, a.recs, a.uv, a.nodes, a.cost,
           a.depth);
  if (n_bp > 1) {
    if (t < n_bp - 1) {
      d_hierarchy(t, a.keys, n_bp, a.left, a.right, a.parent, a.range);
      a.visits[t] = 0u;
    }
    __syncthreads();
    if (t < n_bp)
      d_refit(t, a.left, a.right, a.parent, a.range, a.off, n_bp, a.max_leaf, a.c_trav, a.visits, a.nodes, a.cost, a.depth);
  }
  __syncthreads();
  if (t == 0) {
    const float4 lo = __ldcg(a.nodes), hi = __ldcg(a.nodes + 1);
    a.result[0] = __float_as_uint(lo.x), a.result[1] = __float_as_uint(lo.y), a.result[2] = __float_as_uint(lo.z);
    a.result[3] = __float_as_uint(lo.w);
    a.result[4] = __float_as_uint(hi.x), a.result[5] = __float_as_uint(hi.y), a.result[6] = __float_as_uint(hi.z);
    a.result[7] = __float_as_uint(hi.w);
    a.result[8] = __float_as_uint(__ldcg(a.cost));
    a.result[9] = (uint32_t)__ldcg(a.depth);
    a.result[10] = found;
    a.result[11] = total;
  }
}

// ---- PLOC: the upper tree of large scenes rebuilt by parallel locally-ordered clustering ----------------------------
// (Meister & Bittner 2018.)  Replaces the reference's median split on the sorted v0.x (bvh.cuh:113-121) a second time:
// the LBVH above splits wherever the Morton code flips a bit, whatever the boxes look like.
// For scenes behind the if-if traversal the tree ABOVE the collapsed leaves is rebuilt bottom-up instead: the leaves
// (k_refit's SAH-collapsed subtrees, <= max_leaf records, contiguous by construction) are the initial clusters, in Morton
// order; every round each cluster looks r neighbours to either side for the partner with the smallest merged
// surface area, mutual choices merge (node ids handed out by a prefix sum: deterministic), the cluster array is
// compacted, until one cluster is left.  The result is a new, compact node array: inner nodes [0, m-1) with the root
// at 0, leaves [m-1, 2m-1); the dead nodes of the LBVH are gone.
constexpr int PLOC_R_MAX = 32;  // search radius: RT_PLOC_R (default 8)

// Top-most LEAF-flagged ancestor of every sorted primitive (or its own leaf node): the cluster it belongs to.
__global__ void k_ploc_top_leaf(const float4 *__restrict__ nodes, const uint32_t *__restrict__ parent, int n_bp,
                                uint32_t *cluster_node) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_bp) return;
  uint32_t cur = (uint32_t)(n_bp - 1 + j), top = cur;
  while (true) {
    const uint32_t p = parent[cur];
    if (p == 0xffffffffu) break;
    if (__float_as_uint(nodes[2 * (size_t)p + 1].w) & LEAF_FLAG) top = p;
    cur = p;
  }
  cluster_node[j] = top;
}
__global__ void k_ploc_heads(const uint32_t *cluster_node, int n_bp, uint32_t *flag) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_bp) return;
  flag[j] = (j == 0 || cluster_node[j] != cluster_node[j - 1]) ? 1u : 0u;
}
// cluster c = {lo.xyz, node id}, {hi.xyz, depth}; its leaf node goes to slot (m - 1) + c of the new array
__global__ void k_ploc_init(const float4 *__restrict__ nodes, const uint32_t *cluster_node, const uint32_t *flag,
                            const uint32_t *pos, int n_bp, int m, float4 *cl, float4 *new_nodes) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_bp || !flag[j]) return;
  const uint32_t c = pos[j], o = cluster_node[j], id = (uint32_t)(m - 1) + c;
  const float4 lo = nodes[2 * (size_t)o], hi = nodes[2 * (size_t)o + 1];
  new_nodes[2 * (size_t)id] = lo;
  new_nodes[2 * (size_t)id + 1] = hi;
  cl[2 * (size_t)c] = make_float4(lo.x, lo.y, lo.z, __uint_as_float(id));
  cl[2 * (size_t)c + 1] = make_float4(hi.x, hi.y, hi.z, __int_as_float(1));
}
__device__ __forceinline__ float union_area(float4 alo, float4 ahi, float4 blo, float4 bhi) {
  const float dx = fmaxf(ahi.x, bhi.x) - fminf(alo.x, blo.x), dy = fmaxf(ahi.y, bhi.y) - fminf(alo.y, blo.y),
              dz = fmaxf(ahi.z, bhi.z) - fminf(alo.z, blo.z);
  return dx * dy + dy * dz + dz * dx;
}
// nearest neighbour (smallest merged area) within PLOC_R to either side; ties: the lower index
// The rounds are driven from the device: state = {clusters left, next node id from the top, rounds done, stuck flag}.
// The host launches a batch of rounds with grids sized for the cluster count it last read back (threads beyond the
// real count leave at once) and reads the state once per batch, instead of one host round trip per round (44 - 62 on
// the bunny: 2.7 ms of its 5.4 ms commit were those round trips).
__global__ void __launch_bounds__(256) k_ploc_nn(const float4 *__restrict__ cl, const uint32_t *__restrict__ state, int PLOC_R,
                                                 uint32_t *nn) {
  __shared__ float4 s_lo[256 + 2 * PLOC_R_MAX], s_hi[256 + 2 * PLOC_R_MAX];
  const int m = (int)state[0];
  const int b0 = blockIdx.x * 256;
  if (b0 >= m) return;  // block-uniform
  for (int t = threadIdx.x; t < 256 + 2 * PLOC_R; t += 256) {
    const int g = b0 - PLOC_R + t;
    if (g >= 0 && g < m) {
      s_lo[t] = cl[2 * (size_t)g];
      s_hi[t] = cl[2 * (size_t)g + 1];
    }
  }
  __syncthreads();
  const int i = b0 + threadIdx.x;
  if (i >= m) return;
  const float4 lo = s_lo[threadIdx.x + PLOC_R], hi = s_hi[threadIdx.x + PLOC_R];
  float best = CUDART_INF_F;
  int best_j = -1;
  for (int d = -PLOC_R; d <= PLOC_R; d++) {
    const int j = i + d;
    if (d == 0 || j < 0 || j >= m) continue;
    const float a = union_area(lo, hi, s_lo[threadIdx.x + PLOC_R + d], s_hi[threadIdx.x + PLOC_R + d]);
    if (a < best) {
      best = a;
      best_j = j;
    }
  }
  nn[i] = (uint32_t)best_j;
}
// key = (keeps its slot) | (creates a node) << 32, for one 64-bit exclusive scan
__global__ void k_ploc_flags(const uint32_t *__restrict__ nn, const uint32_t *__restrict__ state, int m_ub,
                             unsigned long long *key) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m_ub) return;
  const int m = (int)state[0];
  if (i >= m) {  // beyond the live clusters: neutral for the scan over m_ub
    key[i] = 0ull;
    return;
  }
  const uint32_t j = nn[i];
  const bool mutual = j < (uint32_t)m && nn[j] == (uint32_t)i;
  const bool merges = mutual && (uint32_t)i < j, dies = mutual && (uint32_t)i > j;
  key[i] = (dies ? 0ull : 1ull) | ((merges ? 1ull : 0ull) << 32);
}
__global__ void k_ploc_apply(const float4 *__restrict__ cl, const uint32_t *__restrict__ nn,
                             const unsigned long long *__restrict__ key, const unsigned long long *__restrict__ scan,
                             const uint32_t *__restrict__ state, float4 *cl_out, float4 *new_nodes) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int m = (int)state[0];
  const uint32_t id_top = state[1];
  if (i >= m) return;
  const unsigned long long k = key[i], sc = scan[i];
  if (!(k & 1ull)) return;  // merged into its partner
  const uint32_t pos = (uint32_t)sc;
  float4 lo = cl[2 * (size_t)i], hi = cl[2 * (size_t)i + 1];
  if (k >> 32) {
    const uint32_t j = nn[i], id = id_top - (uint32_t)(sc >> 32);
    const float4 jlo = cl[2 * (size_t)j], jhi = cl[2 * (size_t)j + 1];
    const float4 ulo = make_float4(fminf(lo.x, jlo.x), fminf(lo.y, jlo.y), fminf(lo.z, jlo.z), lo.w);
    const float4 uhi = make_float4(fmaxf(hi.x, jhi.x), fmaxf(hi.y, jhi.y), fmaxf(hi.z, jhi.z), jlo.w);
    new_nodes[2 * (size_t)id] = ulo;      // {box lo, child a}
    new_nodes[2 * (size_t)id + 1] = uhi;  // {box hi, child b}
    lo = make_float4(ulo.x, ulo.y, ulo.z, __uint_as_float(id));
    hi = make_float4(uhi.x, uhi.y, uhi.z, __int_as_float(1 + max(__float_as_int(hi.w), __float_as_int(jhi.w))));
  }
  cl_out[2 * (size_t)pos] = lo;
  cl_out[2 * (size_t)pos + 1] = hi;
}
// End of a round: the totals of the scan become the next round's state.  (A single cluster is carried from buffer to
// buffer by the kernels above without change, so rounds launched past the end are harmless.)
__global__ void k_ploc_advance(const unsigned long long *__restrict__ key, const unsigned long long *__restrict__ scan, int m_ub,
                               uint32_t *state) {
  if (state[0] <= 1u) return;
  const unsigned long long total = scan[m_ub - 1] + key[m_ub - 1];
  const uint32_t merged = (uint32_t)(total >> 32);
  if (merged == 0) state[3] = 1u;
  state[1] -= merged;
  state[0] = (uint32_t)total;
  state[2] += 1u;
}
// SAH cost of the finished tree: sum over nodes of area / root area * (c_trav | records)
__global__ void k_sah_cost(const float4 *__restrict__ nodes, int n_nodes, float c_trav, double *acc) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  double v = 0.0;
  if (i < n_nodes) {
    const float4 lo = nodes[2 * (size_t)i], hi = nodes[2 * (size_t)i + 1];
    const uint32_t b = __float_as_uint(hi.w);
    v = (double)half_area(lo, hi) * ((b & LEAF_FLAG) ? (double)(b & ~LEAF_FLAG) : (double)c_trav);
  }
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0 && v != 0.0) atomicAdd(acc, v);
}

}  // namespace rtc

// =================================================================================================
// host side
// =================================================================================================
using namespace rtc;

namespace {

inline int nblk(size_t n, int b = 256) { return (int)((n + b - 1) / b); }

struct HostRec {
  float4 r[3];
};
inline float as_f(int32_t i) {
  float f;
  memcpy(&f, &i, 4);
  return f;
}
inline float as_fu(uint32_t i) {
  float f;
  memcpy(&f, &i, 4);
  return f;
}
inline uint32_t rec_word(uint32_t kind, uint32_t flags, int material) {
  return kind | flags | ((uint32_t)material << KIND_BITS);
}
inline HostRec tri_rec(const float p0[3], const float p1[3], const float p2[3], int obj, int sub, uint32_t word) {
  HostRec h;
  // utils.cu:54-55: v0v1 = p1 - p0, v0v2 = p2 - p0 (single float subtractions)
  h.r[0] = make_float4(p0[0], p0[1], p0[2], as_f(obj));
  h.r[1] = make_float4(p1[0] - p0[0], p1[1] - p0[1], p1[2] - p0[2], as_f(sub));
  h.r[2] = make_float4(p2[0] - p0[0], p2[1] - p0[1], p2[2] - p0[2], as_fu(word));
  return h;
}

}  // namespace

extern "C" int rt_scene_commit(rt_ctx *ctx, const rt_scene_desc *d, rt_build_stats *stats) {
  if (!ctx || !d) {
    set_error("rt_scene_commit: null argument");
    return RT_ERR_INVALID;
  }
  if (d->n_hitables < 0 || d->n_materials < 0 || (d->n_hitables > 0 && !d->hitables) ||
      (d->n_materials > 0 && !d->materials)) {
    set_error("rt_scene_commit: bad counts");
    return RT_ERR_INVALID;
  }
  RT_CUDA(cudaSetDevice(ctx->device));
  ctx->has_scene = false;
  ctx->build_stats.reserved_[0] = ctx->build_stats.reserved_[1] = 0;
  cudaStream_t st = 0;
  RT_CUDA(cudaEventRecord(ctx->ev0, st));

  // ---- materials -------------------------------------------------------------------------------
  std::vector<DMaterial> hm(std::max(d->n_materials, 1));
  for (int i = 0; i < d->n_materials; i++) {
    const rt_material &m = d->materials[i];
    if (m.texture >= d->n_textures) {
      set_error("rt_scene_commit: material texture index out of range");
      return RT_ERR_INVALID;
    }
    hm[i].color_fuzz = make_float4(m.color[0], m.color[1], m.color[2], m.fuzz);
    hm[i].type = m.type;
    hm[i].texture = m.texture;
    hm[i].eta_in = (float)m.ior / 1.0f;
    hm[i].eta_out = 1.0f / (float)m.ior;
  }
  RT_CUDA(ctx->mats.reserve(hm.size() * sizeof(DMaterial)));
  RT_CUDA(cudaMemcpyAsync(ctx->mats.p, hm.data(), hm.size() * sizeof(DMaterial), cudaMemcpyHostToDevice, st));

  // ---- textures --------------------------------------------------------------------------------
  // texel buffers are kept per texture slot and only grow: a re-commit of the same scene (every frame of a caller that
  // rebuilds per frame, every e2e step of bench.py) neither frees nor allocates
  if ((int)ctx->tex_allocs.size() < d->n_textures) ctx->tex_allocs.resize(d->n_textures);
  std::vector<DTexture> ht(std::max(d->n_textures, 1));
  for (int i = 0; i < d->n_textures; i++) {
    const rt_texture &t = d->textures[i];
    if (!t.h_rgba || t.height <= 0 || t.width <= 0 || t.pitch_bytes < (int64_t)t.width * 4) {
      set_error("rt_scene_commit: bad texture");
      return RT_ERR_INVALID;
    }
    size_t bytes = (size_t)t.pitch_bytes * t.height;
    RT_CUDA(ctx->tex_allocs[i].reserve(bytes));
    void *p = ctx->tex_allocs[i].p;
    RT_CUDA(cudaMemcpyAsync(p, t.h_rgba, bytes, cudaMemcpyHostToDevice, st));
    ht[i].rgba = (const uint8_t *)p;
    ht[i].height = t.height;
    ht[i].width = t.width;
    ht[i].pitch = t.pitch_bytes;
  }
  RT_CUDA(ctx->textures.reserve(ht.size() * sizeof(DTexture)));
  RT_CUDA(cudaMemcpyAsync(ctx->textures.p, ht.data(), ht.size() * sizeof(DTexture), cudaMemcpyHostToDevice, st));

  // ---- sequential record list (Append order) ----------------------------------------------------
  size_t n_seq = 0;
  bool any_uv = false;
  int sky_obj = -1, sky_mat = -1, n_spheres = 0, n_quads = 0, n_tris = 0;
  for (int i = 0; i < d->n_hitables; i++) {
    const rt_hitable &h = d->hitables[i];
    if (h.material < 0 || h.material >= d->n_materials) {
      set_error("rt_scene_commit: hitable material index out of range");
      return RT_ERR_INVALID;
    }
    switch (h.type) {
      case RT_HIT_SPHERE: n_seq += 1; n_spheres++; break;
      case RT_HIT_TRIANGLE: n_seq += 1; n_tris++; break;
      case RT_HIT_PARALLELOGRAM: n_seq += 2; n_quads++; break;
      case RT_HIT_SKY:
        n_seq += 1;
        if (sky_obj < 0) {
          sky_obj = i;
          sky_mat = h.material;
        }
        break;
      case RT_HIT_MESH:
        if (h.mesh < 0 || h.mesh >= d->n_meshes || d->meshes[h.mesh].n_faces < 0 ||
            (d->meshes[h.mesh].n_faces > 0 && !d->meshes[h.mesh].h_faces)) {
          set_error("rt_scene_commit: bad mesh reference");
          return RT_ERR_INVALID;
        }
        n_seq += (size_t)d->meshes[h.mesh].n_faces;
        n_tris += d->meshes[h.mesh].n_faces;
        any_uv |= d->meshes[h.mesh].has_texcoord != 0;
        break;
      default:
        set_error("rt_scene_commit: unknown hitable type");
        return RT_ERR_INVALID;
    }
  }
  if (n_seq >= 0x7fffffffull / 4) {
    set_error("rt_scene_commit: too many primitives");
    return RT_ERR_LIMIT;
  }
  RT_CUDA(ctx->recs_seq.reserve(std::max<size_t>(n_seq, 1) * 48));
  if (any_uv) {
    RT_CUDA(ctx->face_uv_seq.reserve(n_seq * 24));
    RT_CUDA(cudaMemsetAsync(ctx->face_uv_seq.p, 0, n_seq * 24, st));
  }
  float4 *recs_seq = ctx->recs_seq.as<float4>();
  float *uv_seq = any_uv ? ctx->face_uv_seq.as<float>() : nullptr;

  std::vector<HostRec> run;
  size_t run_start = 0, pos = 0;
  auto flush = [&]() -> cudaError_t {
    cudaError_t e = cudaSuccess;
    if (!run.empty())
      e = cudaMemcpy(recs_seq + 3 * run_start, run.data(), run.size() * sizeof(HostRec), cudaMemcpyHostToDevice);
    run.clear();
    return e;
  };
  for (int i = 0; i < d->n_hitables; i++) {
    const rt_hitable &h = d->hitables[i];
    if (run.empty()) run_start = pos;
    switch (h.type) {
      case RT_HIT_SPHERE: {
        HostRec r;
        uint64_t rb;
        memcpy(&rb, &h.radius, 8);
        r.r[0] = make_float4(h.p[0][0], h.p[0][1], h.p[0][2], as_f(i));
        r.r[1] = make_float4(as_fu((uint32_t)rb), as_fu((uint32_t)(rb >> 32)), 0.f, as_f(0));
        r.r[2] = make_float4(0.f, 0.f, 0.f, as_fu(rec_word(K_SPHERE, 0, h.material)));
        run.push_back(r);
        pos += 1;
        break;
      }
      case RT_HIT_TRIANGLE:
        run.push_back(tri_rec(h.p[0], h.p[1], h.p[2], i, 0, rec_word(K_TRI, FLAG_BARY, h.material)));
        pos += 1;
        break;
      case RT_HIT_PARALLELOGRAM:
        run.push_back(tri_rec(h.p[0], h.p[1], h.p[2], i, 0, rec_word(K_QUAD0, 0, h.material)));
        run.push_back(tri_rec(h.p[1], h.p[2], h.p[3], i, 1, rec_word(K_QUAD1, 0, h.material)));
        pos += 2;
        break;
      case RT_HIT_SKY: {
        HostRec r;
        r.r[0] = make_float4(0.f, 0.f, 0.f, as_f(i));
        r.r[1] = make_float4(0.f, 0.f, 0.f, as_f(0));
        r.r[2] = make_float4(0.f, 0.f, 0.f, as_fu(rec_word(K_SKY, 0, h.material)));
        run.push_back(r);
        pos += 1;
        break;
      }
      case RT_HIT_MESH: {
        RT_CUDA(flush());
        const rt_mesh &m = d->meshes[h.mesh];
        int n = m.n_faces, stride = m.has_texcoord ? 15 : 9;
        if (n > 0) {
          size_t bytes = (size_t)n * stride * 4;
          RT_CUDA(ctx->scratch[0].reserve(bytes));
          RT_CUDA(cudaMemcpyAsync(ctx->scratch[0].p, m.h_faces, bytes, cudaMemcpyDefault, st));
          const uint32_t *perm = nullptr;
          if (n > kRefLeafMin) {  // bvh.cuh:116-117: stable sort on positions_[0].x
            RT_CUDA(ctx->scratch[1].reserve((size_t)n * 4));
            RT_CUDA(ctx->scratch[2].reserve((size_t)n * 4));
            RT_CUDA(ctx->scratch[3].reserve((size_t)n * 4));
            RT_CUDA(ctx->scratch[4].reserve((size_t)n * 4));
            k_face_keys<<<nblk(n), 256, 0, st>>>(ctx->scratch[0].as<float>(), stride, n, ctx->scratch[1].as<uint32_t>(),
                                                 ctx->scratch[2].as<uint32_t>());
            size_t tmp = 0;
            cub::DeviceRadixSort::SortPairs(nullptr, tmp, ctx->scratch[1].as<uint32_t>(), ctx->scratch[3].as<uint32_t>(),
                                            ctx->scratch[2].as<uint32_t>(), ctx->scratch[4].as<uint32_t>(), n, 0, 32, st);
            RT_CUDA(ctx->scratch[5].reserve(tmp));
            RT_CUDA(cub::DeviceRadixSort::SortPairs(ctx->scratch[5].p, tmp, ctx->scratch[1].as<uint32_t>(),
                                                    ctx->scratch[3].as<uint32_t>(), ctx->scratch[2].as<uint32_t>(),
                                                    ctx->scratch[4].as<uint32_t>(), n, 0, 32, st));
            perm = ctx->scratch[4].as<uint32_t>();
          }
          k_face_records<<<nblk(n), 256, 0, st>>>(ctx->scratch[0].as<float>(), stride, perm, n, i,
                                                  rec_word(K_TRI, 0, h.material), recs_seq + 3 * pos,
                                                  uv_seq ? uv_seq + 6 * pos : nullptr);
          RT_CUDA(cudaGetLastError());
          RT_CUDA(cudaStreamSynchronize(st));  // scratch[0] is reused by the next mesh
        }
        pos += (size_t)n;
        break;
      }
    }
  }
  RT_CUDA(flush());

  // ---- build primitives -------------------------------------------------------------------------
  // Both counts follow from the description (a sphere / triangle / face is one primitive of one record, a
  // parallelogram one primitive of two, the sky none): nothing is read back from the device to size the build.
  const int n_bp = n_spheres + n_tris + n_quads;
  const size_t n_recs_expected = (size_t)n_spheres + (size_t)n_tris + 2 * (size_t)n_quads;
  DevBuf *S = ctx->scratch;
  size_t n_recs = 0;
  int n_nodes = 0, max_depth = 0;
  float sah = 0.f;
  uint32_t root_words[2] = {0, LEAF_FLAG};
  const char *env = getenv("RT_MAX_LEAF");
  const int max_leaf = env ? std::min(std::max(1, atoi(env)), 7) : 4;
  env = getenv("RT_SAH_CT");
  const float c_trav = env ? (float)atof(env) : 1.2f;
  env = getenv("RT_BUILD_SMALL");  // 0: small scenes through the step-per-launch path too (tests compare the two)
  const bool small = n_bp > 0 && n_bp <= SMALL_MAX_BP && n_seq <= (size_t)SMALL_MAX_SEQ && !(env && atoi(env) == 0);
  bool flat_done = false;
  if (small) {
    // ---- one launch (k_build_small), one read-back ----
    n_recs = n_recs_expected;
    n_nodes = 2 * n_bp - 1;
    RT_CUDA(ctx->recs.reserve(n_recs * 48));
    RT_CUDA(ctx->nodes.reserve((size_t)n_nodes * 32));
    if (any_uv) RT_CUDA(ctx->face_uv.reserve(n_recs * 24));
    RT_CUDA(S[0].reserve((size_t)n_nodes * 8));
    RT_CUDA(S[2].reserve(((size_t)n_bp + 1) * 4));
    RT_CUDA(S[3].reserve((size_t)n_bp * 4));
    RT_CUDA(S[4].reserve(((size_t)n_bp * 2 + (size_t)n_nodes) * 4));
    RT_CUDA(S[6].reserve((size_t)n_bp * 16));
    RT_CUDA(S[7].reserve((size_t)n_bp * 16));
    RT_CUDA(S[8].reserve((size_t)n_bp * 8));
    RT_CUDA(S[9].reserve((size_t)n_bp * 4));
    RT_CUDA(S[10].reserve(64));
    RT_CUDA(S[11].reserve((size_t)n_bp * 12));
    SmallBuild a;
    a.recs_seq = recs_seq;
    a.uv_seq = uv_seq;
    a.n_seq = (int)n_seq;
    a.n_bp = n_bp;
    a.max_leaf = max_leaf;
    a.c_trav = c_trav;
    a.bp_first = S[3].as<uint32_t>();
    a.order = S[9].as<uint32_t>();
    a.off = S[2].as<uint32_t>();
    a.left = S[4].as<uint32_t>();
    a.right = a.left + n_bp;
    a.parent = a.right + n_bp;
    a.range = S[11].as<uint2>();
    a.visits = (uint32_t *)(a.range + n_bp);
    a.bp_lo = S[6].as<float4>();
    a.bp_hi = S[7].as<float4>();
    a.recs = ctx->recs.as<float4>();
    a.nodes = ctx->nodes.as<float4>();
    a.keys = S[8].as<uint64_t>();
    a.uv = any_uv ? ctx->face_uv.as<float>() : nullptr;
    a.cost = S[0].as<float>();
    a.depth = (int *)(a.cost + n_nodes);
    a.result = S[10].as<uint32_t>();
    k_build_small<<<1, SMALL_T, 0, st>>>(a);
    RT_CUDA(cudaGetLastError());
    uint32_t res[12];
    RT_CUDA(cudaMemcpyAsync(res, a.result, sizeof(res), cudaMemcpyDeviceToHost, st));
    if (n_bp <= FLAT_MAX)  // the box list of rt_flat.cuh rides on the same round trip
      RT_CUDA(cudaMemcpyAsync(ctx->flat.box, ctx->nodes.as<float4>() + 2 * (size_t)(n_bp - 1), (size_t)n_bp * 32,
                              cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));
    flat_done = true;
    if (res[10] != (uint32_t)n_bp || res[11] != (uint32_t)n_recs) {
      set_error("rt_scene_commit: the device build found " + std::to_string(res[10]) + " primitives / " +
                std::to_string(res[11]) + " records, the description has " + std::to_string(n_bp) + " / " +
                std::to_string(n_recs));
      return RT_ERR_CUDA;
    }
    root_words[0] = res[3];
    root_words[1] = res[7];
    memcpy(&sah, &res[8], 4);
    max_depth = (int)res[9];
  } else if (n_bp > 0) {
    RT_CUDA(S[1].reserve((n_seq + 1) * 4));
    RT_CUDA(S[2].reserve((n_seq + 1) * 4));
    k_head_flags<<<nblk(n_seq), 256, 0, st>>>(recs_seq, (int)n_seq, S[1].as<uint32_t>());
    {
      size_t tmp = 0;
      cub::DeviceScan::ExclusiveSum(nullptr, tmp, S[1].as<uint32_t>(), S[2].as<uint32_t>(), (int)n_seq, st);
      RT_CUDA(S[5].reserve(tmp));
      RT_CUDA(cub::DeviceScan::ExclusiveSum(S[5].p, tmp, S[1].as<uint32_t>(), S[2].as<uint32_t>(), (int)n_seq, st));
    }
    // S[3] bp_first, S[6] bp_lo, S[7] bp_hi, S[8] scene bounds + keys, S[9] vals ...
    RT_CUDA(S[3].reserve((size_t)n_bp * 4));
    k_scatter_heads<<<nblk(n_seq), 256, 0, st>>>(S[1].as<uint32_t>(), S[2].as<uint32_t>(), (int)n_seq,
                                                 S[3].as<uint32_t>());
    RT_CUDA(S[6].reserve((size_t)n_bp * 16));
    RT_CUDA(S[7].reserve((size_t)n_bp * 16));
    RT_CUDA(S[10].reserve(32));
    uint32_t binit[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0, 0, 0};
    RT_CUDA(cudaMemcpyAsync(S[10].p, binit, sizeof(binit), cudaMemcpyHostToDevice, st));
    k_bp_bounds<<<nblk(n_bp), 256, 0, st>>>(recs_seq, S[3].as<uint32_t>(), n_bp, S[6].as<float4>(), S[7].as<float4>(),
                                            S[10].as<uint32_t>());
    // keys/vals double buffers: S[8] = keys in | keys out, S[9] = vals in | vals out
    RT_CUDA(S[8].reserve((size_t)n_bp * 16));
    RT_CUDA(S[9].reserve((size_t)n_bp * 8));
    uint64_t *keys_in = S[8].as<uint64_t>(), *keys = keys_in + n_bp;
    uint32_t *vals_in = S[9].as<uint32_t>(), *order = vals_in + n_bp;
    k_morton<<<nblk(n_bp), 256, 0, st>>>(S[6].as<float4>(), S[7].as<float4>(), n_bp, S[10].as<uint32_t>(), keys_in,
                                         vals_in);
    size_t tmp = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmp, keys_in, keys, vals_in, order, n_bp, 0, 63, st);
    RT_CUDA(S[5].reserve(tmp));
    RT_CUDA(cub::DeviceRadixSort::SortPairs(S[5].p, tmp, keys_in, keys, vals_in, order, n_bp, 0, 63, st));
    // record offsets of the sorted primitives: S[1] nrec, S[2] off (n_bp + 1)
    RT_CUDA(S[1].reserve(((size_t)n_bp + 1) * 4));
    RT_CUDA(S[2].reserve(((size_t)n_bp + 1) * 4));
    RT_CUDA(cudaMemsetAsync(S[1].p, 0, ((size_t)n_bp + 1) * 4, st));
    k_sorted_nrec<<<nblk(n_bp), 256, 0, st>>>(S[6].as<float4>(), order, n_bp, S[1].as<uint32_t>());
    cub::DeviceScan::ExclusiveSum(nullptr, tmp, S[1].as<uint32_t>(), S[2].as<uint32_t>(), n_bp + 1, st);
    RT_CUDA(S[5].reserve(tmp));
    RT_CUDA(cub::DeviceScan::ExclusiveSum(S[5].p, tmp, S[1].as<uint32_t>(), S[2].as<uint32_t>(), n_bp + 1, st));
    uint32_t total = 0;  // checked against the description with the root read-back below
    RT_CUDA(cudaMemcpyAsync(&total, S[2].as<uint32_t>() + n_bp, 4, cudaMemcpyDeviceToHost, st));
    n_recs = n_recs_expected;
    n_nodes = 2 * n_bp - 1;
    RT_CUDA(ctx->recs.reserve(n_recs * 48));
    RT_CUDA(ctx->nodes.reserve((size_t)n_nodes * 32));
    if (any_uv) RT_CUDA(ctx->face_uv.reserve(n_recs * 24));
    // S[0] cost | depth, S[4] left | right | parent, S[11] range | visits
    RT_CUDA(S[0].reserve((size_t)n_nodes * 8));
    float *cost = S[0].as<float>();
    int *depth = (int *)(cost + n_nodes);
    k_emit<<<nblk(n_bp), 256, 0, st>>>(recs_seq, uv_seq, S[3].as<uint32_t>(), S[6].as<float4>(), S[7].as<float4>(),
                                       order, S[2].as<uint32_t>(), n_bp, ctx->recs.as<float4>(),
                                       any_uv ? ctx->face_uv.as<float>() : nullptr, ctx->nodes.as<float4>(), cost,
                                       depth);
    if (n_bp > 1) {
      RT_CUDA(S[4].reserve(((size_t)n_bp * 2 + (size_t)n_nodes) * 4));
      uint32_t *left = S[4].as<uint32_t>(), *right = left + n_bp, *parent = right + n_bp;
      RT_CUDA(S[11].reserve((size_t)n_bp * 12));
      uint2 *range = S[11].as<uint2>();
      uint32_t *visits = (uint32_t *)(range + n_bp);
      RT_CUDA(cudaMemsetAsync(visits, 0, (size_t)n_bp * 4, st));
      k_hierarchy<<<nblk(n_bp - 1), 256, 0, st>>>(keys, n_bp, left, right, parent, range);
      k_refit<<<nblk(n_bp), 256, 0, st>>>(left, right, parent, range, S[2].as<uint32_t>(), n_bp, max_leaf, c_trav,
                                          visits, ctx->nodes.as<float4>(), cost, depth);
    }
    RT_CUDA(cudaGetLastError());
    float4 root[2];
    RT_CUDA(cudaMemcpyAsync(root, ctx->nodes.p, 32, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaMemcpyAsync(&sah, cost, 4, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaMemcpyAsync(&max_depth, depth, 4, cudaMemcpyDeviceToHost, st));
    RT_CUDA(cudaStreamSynchronize(st));
    if ((size_t)total != n_recs) {
      set_error("rt_scene_commit: the device build emitted " + std::to_string(total) + " records, the description has " +
                std::to_string(n_recs));
      return RT_ERR_CUDA;
    }
    // ---- large scenes: the tree above the leaves rebuilt by PLOC (RT_PLOC=0: keep the LBVH) ----
    {
      const char *ep = getenv("RT_PLOC");
      const bool ploc = n_recs > 4096 && n_bp > FLAT_MAX && !(ep && atoi(ep) == 0);
      ctx->build_stats.reserved_[0] = 0;
      if (ploc) {
        uint32_t *parent = S[4].as<uint32_t>() + 2 * (size_t)n_bp;
        // S[1] cluster_node | flag, S[2] pos (n_bp + 1)  (nrec / off are dead after k_refit)
        RT_CUDA(S[1].reserve((size_t)n_bp * 8 + 8));
        uint32_t *cluster_node = S[1].as<uint32_t>(), *flag = cluster_node + n_bp;
        uint32_t *pos = S[2].as<uint32_t>();
        k_ploc_top_leaf<<<nblk(n_bp), 256, 0, st>>>(ctx->nodes.as<float4>(), parent, n_bp, cluster_node);
        k_ploc_heads<<<nblk(n_bp), 256, 0, st>>>(cluster_node, n_bp, flag);
        size_t tmp = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, tmp, flag, pos, n_bp, st);
        RT_CUDA(S[5].reserve(tmp));
        RT_CUDA(cub::DeviceScan::ExclusiveSum(S[5].p, tmp, flag, pos, n_bp, st));
        uint32_t last_pos = 0, last_flag = 0;
        RT_CUDA(cudaMemcpyAsync(&last_pos, pos + n_bp - 1, 4, cudaMemcpyDeviceToHost, st));
        RT_CUDA(cudaMemcpyAsync(&last_flag, flag + n_bp - 1, 4, cudaMemcpyDeviceToHost, st));
        RT_CUDA(cudaStreamSynchronize(st));
        int m = (int)(last_pos + last_flag);
        if (m >= 2) {
          const int n_new = 2 * m - 1;
          RT_CUDA(ctx->nodes_ploc.reserve((size_t)n_new * 32));
          float4 *new_nodes = ctx->nodes_ploc.as<float4>();
          // cluster arrays (double buffered): S[6] / S[7] (bp_lo / bp_hi, dead after k_emit; 32 B per cluster)
          RT_CUDA(S[6].reserve((size_t)m * 32));
          RT_CUDA(S[7].reserve((size_t)m * 32));
          float4 *cl = S[6].as<float4>(), *cl_out = S[7].as<float4>();
          k_ploc_init<<<nblk(n_bp), 256, 0, st>>>(ctx->nodes.as<float4>(), cluster_node, flag, pos, n_bp, m, cl, new_nodes);
          // S[8]: nn (4 B) | key (8 B) | scan (8 B) per cluster
          RT_CUDA(S[8].reserve((size_t)m * 24 + 64));
          unsigned long long *key = S[8].as<unsigned long long>(), *scan = key + m;
          uint32_t *nn = (uint32_t *)(scan + m);
          cub::DeviceScan::ExclusiveSum(nullptr, tmp, key, scan, m, st);
          RT_CUDA(S[5].reserve(tmp));
          const char *er = getenv("RT_PLOC_R");
          const int ploc_r = er ? std::min(std::max(atoi(er), 1), PLOC_R_MAX) : 8;
          // device-side round state: {clusters, next id from the top, rounds, stuck}
          RT_CUDA(S[10].reserve(64));
          uint32_t *state = S[10].as<uint32_t>() + 8;
          uint32_t hstate[4] = {(uint32_t)m, (uint32_t)(m - 2), 0u, 0u};
          RT_CUDA(cudaMemcpyAsync(state, hstate, sizeof(hstate), cudaMemcpyHostToDevice, st));
          int cur = m, rounds = 0;
          while (cur > 1) {
            for (int b = 0; b < 16; b++) {  // a batch of rounds sized for `cur` clusters, no host round trip in between
              k_ploc_nn<<<nblk(cur), 256, 0, st>>>(cl, state, ploc_r, nn);
              k_ploc_flags<<<nblk(cur), 256, 0, st>>>(nn, state, cur, key);
              size_t t2 = tmp;
              RT_CUDA(cub::DeviceScan::ExclusiveSum(S[5].p, t2, key, scan, cur, st));
              k_ploc_apply<<<nblk(cur), 256, 0, st>>>(cl, nn, key, scan, state, cl_out, new_nodes);
              k_ploc_advance<<<1, 1, 0, st>>>(key, scan, cur, state);
              std::swap(cl, cl_out);
            }
            RT_CUDA(cudaMemcpyAsync(hstate, state, sizeof(hstate), cudaMemcpyDeviceToHost, st));
            RT_CUDA(cudaStreamSynchronize(st));
            if (hstate[3] || (int)hstate[0] >= cur) {
              set_error("PLOC: a round merged nothing");
              return RT_ERR_CUDA;
            }
            cur = (int)hstate[0];
            rounds = (int)hstate[2];
          }
          // statistics of the new tree
          RT_CUDA(S[10].reserve(32));
          double *acc = S[10].as<double>();
          RT_CUDA(cudaMemsetAsync(acc, 0, 8, st));
          k_sah_cost<<<nblk(n_new), 256, 0, st>>>(new_nodes, n_new, c_trav, acc);
          double total_cost = 0;
          float4 top[2];
          RT_CUDA(cudaMemcpyAsync(&total_cost, acc, 8, cudaMemcpyDeviceToHost, st));
          RT_CUDA(cudaMemcpyAsync(top, cl, 32, cudaMemcpyDeviceToHost, st));
          RT_CUDA(cudaMemcpyAsync(root, new_nodes, 32, cudaMemcpyDeviceToHost, st));
          RT_CUDA(cudaStreamSynchronize(st));
          const float ra = (top[1].x - top[0].x) * (top[1].y - top[0].y) + (top[1].y - top[0].y) * (top[1].z - top[0].z) +
                           (top[1].z - top[0].z) * (top[1].x - top[0].x);
          // keep whichever tree has the lower SAH cost: clustering wins where primitives differ in size or the scene
          // has structure (bunny + backdrop: 2.2 against 5.0, 25 % fewer box tests per ray), the Morton split wins on a
          // uniformly random soup (372 against 377; profiles/r02_uv_ploc.txt)
          const float sah_ploc = ra > 0 ? (float)(total_cost / ra) : 3.0e38f;
          if (sah_ploc < sah) {
            memcpy(&ctx->build_stats.reserved_[1], &sah, 4);
            sah = sah_ploc;
            memcpy(&max_depth, &top[1].w, 4);
            std::swap(ctx->nodes, ctx->nodes_ploc);
            n_nodes = n_new;
            ctx->build_stats.reserved_[0] = rounds;
          } else {
            memcpy(&ctx->build_stats.reserved_[1], &sah_ploc, 4);
            ctx->build_stats.reserved_[0] = -rounds;
            RT_CUDA(cudaMemcpyAsync(root, ctx->nodes.p, 32, cudaMemcpyDeviceToHost, st));
            RT_CUDA(cudaStreamSynchronize(st));
          }
        }
      }
    }
    memcpy(&root_words[0], &root[0].w, 4);
    memcpy(&root_words[1], &root[1].w, 4);

    RT_CUDA(cudaGetLastError());
  }
  DScene &s = ctx->scene;
  s.nodes = ctx->nodes.as<float4>();
  s.recs = ctx->recs.as<float4>();
  s.recs_seq = recs_seq;
  s.face_uv = any_uv ? ctx->face_uv.as<float>() : nullptr;
  s.face_uv_seq = uv_seq;
  s.mats = ctx->mats.as<DMaterial>();
  s.textures = ctx->textures.as<DTexture>();
  s.root_a = root_words[0];
  s.root_b = root_words[1];
  s.n_recs = (int)n_recs;
  s.n_recs_seq = (int)n_seq;
  s.n_nodes = n_nodes;
  s.sky_obj = sky_obj;
  s.sky_mat = sky_mat;
  s.has_spheres = n_spheres > 0;
  {  // RT_SPHERE_POW=ieee: Sphere::Hit in plain IEEE arithmetic (the CPU oracle's form); default: CUDA's pow, as the reference
    const char *env = getenv("RT_SPHERE_POW");
    s.sphere_pow = (env && strcmp(env, "ieee") == 0) ? 0 : 1;
  }
  // box list for the flat traversal (rt_flat.cuh): the single-primitive leaf nodes k_emit wrote
  ctx->flat.n = 0;
  if (n_bp > 0 && n_bp <= FLAT_MAX) {
    if (!flat_done)
      RT_CUDA(cudaMemcpy(ctx->flat.box, ctx->nodes.as<float4>() + 2 * (size_t)(n_bp - 1), (size_t)n_bp * 32,
                         cudaMemcpyDeviceToHost));
    ctx->flat.n = n_bp;
    // {lo, hi} -> {centre, half extent}: the half extent covers the rounding of the centre
    for (int i = 0; i < n_bp; i++) {
      float4 &lo = ctx->flat.box[2 * i], &hi = ctx->flat.box[2 * i + 1];
      float *l = &lo.x, *h = &hi.x;
      for (int a = 0; a < 3; a++) {
        const double c = 0.5 * ((double)l[a] + (double)h[a]);
        const float cf = (float)c;
        const double he = std::max((double)h[a] - (double)cf, (double)cf - (double)l[a]);
        float hf = (float)he;
        if ((double)hf < he) hf = std::nextafter(hf, INFINITY);
        l[a] = cf;
        h[a] = hf;
      }
    }
  }
  ctx->has_scene = true;

  rt_build_stats &bs = ctx->build_stats;
  // the traversal this scene renders with (RT_TRAV / RT_STRATEGY are read here, once per commit) and its tree;
  // part of the build time
  ctx->trav = default_trav(ctx, sah);
  {
    const char *env = getenv("RT_STRATEGY");
    ctx->opt_wavefront = env && strcmp(env, "wavefront") == 0;
  }
  RT_CUDA(cudaEventRecord(ctx->ev1, st));
  RT_CUDA(cudaEventSynchronize(ctx->ev1));
  float ms = 0.f;
  cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
  bs.n_prims = n_bp;
  bs.n_nodes = n_nodes;
  bs.n_spheres = n_spheres;
  bs.n_quads = n_quads;
  bs.n_tris = n_tris;
  bs.max_depth = max_depth;
  bs.build_ms = ms;
  bs.upload_ms = 0.f;
  bs.sah_cost = sah;
  if (stats) *stats = bs;
  return RT_OK;
}
