// rt_wide.cuh — closest hit over the 8-wide (or 4-wide, W = 4) tree (traversal variants RT_TRAV=10/11/13/15;
// parity modes 8/9/11).
//
// Why wide: ncu on the binary while-while loop (profiles/r01_g_cornell_ww_ncu_summary.txt) shows the inner-node
// steps — 48 % of all warp instructions — running with 9 of 32 lanes: a ray takes 7-8 dependent two-box steps
// and the lanes of a warp need different numbers of them.  With 8 children per node a Cornell-box ray takes
// 1-2 steps, and per child the box test shrinks to 6 FFMA + 2 three-input min/max because the ray selects its
// near / far planes by address (direction octant).
// Measured (profiles/r01_h_cornell_wide_ncu_summary.txt, r01_h_variants_sweep.txt): -7 % warp instructions and
// +2 % on Cornell, equal on bunny, slower on spheres and the soups (more box tests per ray, 14 LDG.128 per step):
// NOT the default; kept as a variant with its evidence.
//
// Box culling only has to be conservative (rt_device.cuh, slab2): planes are evaluated as
// fma(plane, 1/d, -(o/d)) with 1/d from slab_rcp (finite); the rounding of that form is covered by the builder's
// relative box padding plus the relative guard bands on t below.
//
// Visit order: nearest hit child first; the other hit children are pushed with their entry distance and
// dropped at pop time once the closest hit has moved in front of them.  Hits are decided by rec_test /
// coop_leaf_phase, identical to every other traversal.
#pragma once
#include "rt_coop.cuh"

namespace rtc {

constexpr int WSTACK = 96;  // up to 7 pushes per level

// Every lane of the warp must call it when COOP (lanes without a ray pass active = false); sm is the warp's
// CoopWarp.  !COOP: per-lane leaf tests, sm unused, any subset of lanes may call.
template <bool STATS, bool COOP, int W = 8>
__device__ __forceinline__ Hit world_hit_wide(const DScene &s, const RayQ &r, bool active, double t_from, double t_to,
                                              CoopWarp *sm, TravStats *st) {
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  Range rg;
  rg.t_from = t_from;
  rg.t_to = t_to;
  rg.tf_f = float_ge(t_from);
  Best b;
  best_init(b, t_to);
  if (s.sky_obj >= 0) {  // sky.cu:18-27: always a candidate at t = 1e9
    double t = 1e9;
    if (t_from <= t && t <= t_to) {
      b.t = t;
      b.t_f = 1e9f;
      b.rec = REC_SKY;
      b.obj = s.sky_obj;
      b.sub = 0;
    }
  }
  bool ambiguous = false, overflow = false;
  bool done = !active || s.n_recs <= 0;
  if (COOP) {
    sm->ray[0][lane] = r.o.x;
    sm->ray[1][lane] = r.o.y;
    sm->ray[2][lane] = r.o.z;
    sm->ray[3][lane] = r.d.x;
    sm->ray[4][lane] = r.d.y;
    sm->ray[5][lane] = r.d.z;
  }
  const float idx = slab_rcp(r.d.x), idy = slab_rcp(r.d.y), idz = slab_rcp(r.d.z);
  const float nox = -(r.o.x * idx), noy = -(r.o.y * idy), noz = -(r.o.z * idz);
  // float4 offsets of this ray's near / far plane arrays inside a node
  using L = WideLayout<W>;
  const int nx = idx < 0.f ? L::HI : 0, ny = L::G + (idy < 0.f ? L::HI : 0), nz = 2 * L::G + (idz < 0.f ? L::HI : 0);
  const int fx = L::HI - nx, fy = 2 * L::G + L::HI - ny, fz = 4 * L::G + L::HI - nz;
  const float t_lo = rg.tf_f - fabsf(rg.tf_f) * 1e-5f;
  uint2 stack[WSTACK];
  int sp = 0;
  uint32_t cur = 0;  // wide node 0 = root
  float cur_tn = 0.f;  // entry distance of `cur` (for the cull after a leaf phase)
  const float4 *__restrict__ wn = W == 8 ? s.wnodes : s.wnodes4;
  float t_hi = b.t_f + fabsf(b.t_f) * 1e-5f;
  auto pop = [&]() -> bool {
    while (sp > 0) {
      uint2 e = stack[--sp];
      if (__uint_as_float(e.y) > t_hi * 1.00001f) continue;  // the closest hit moved in front of it
      cur = e.x;
      cur_tn = __uint_as_float(e.y);
      return true;
    }
    return false;
  };
  for (;;) {
    if (COOP) {
      if (!__any_sync(FULL, !done)) break;
    } else if (done) {
      break;
    }
    // ---- walk: inner nodes until this lane holds a leaf ----
    uint32_t leaf = 0;
    if (!done) {
      t_hi = b.t_f + fabsf(b.t_f) * 1e-5f;
      // the node popped before the last leaf phase may now lie behind the closest hit
      if (cur_tn > t_hi * 1.00001f && !pop()) done = true;
      while (!done) {
        if (cur & WLEAF) {
          leaf = cur;
          if (!pop()) done = true;
          break;
        }
        const float4 *__restrict__ n = wn + (size_t)L::F4 * cur;
        const uint32_t valid = __float_as_uint(__ldg(n + L::META).x);
        if (STATS) st->nodes += __popc(valid);
        uint32_t best_ref = WREF_EMPTY;
        float best_tn = CUDART_INF_F;
#define RT_WCHILD(NX, NY, NZ, FX, FY, FZ, REF, BIT)                                                            \
  {                                                                                                            \
    float tn_ = fmaxf(fmaxf(__fmaf_rn(NX, idx, nox), __fmaf_rn(NY, idy, noy)), fmaxf(__fmaf_rn(NZ, idz, noz), t_lo)); \
    float tf_ = fminf(fminf(__fmaf_rn(FX, idx, nox), __fmaf_rn(FY, idy, noy)), fminf(__fmaf_rn(FZ, idz, noz), t_hi)); \
    bool hit_ = (tn_ <= tf_ * 1.00001f) && (valid & (BIT));                                                    \
    bool nearer_ = hit_ && tn_ < best_tn;                                                                      \
    uint32_t pr_ = nearer_ ? best_ref : __float_as_uint(REF);                                                  \
    float pt_ = nearer_ ? best_tn : tn_;                                                                       \
    if (hit_ && pr_ != WREF_EMPTY) {                                                                           \
      if (sp < WSTACK)                                                                                         \
        stack[sp++] = make_uint2(pr_, __float_as_uint(pt_));                                                   \
      else                                                                                                     \
        overflow = true;                                                                                       \
    }                                                                                                          \
    best_ref = nearer_ ? __float_as_uint(REF) : best_ref;                                                      \
    best_tn = nearer_ ? tn_ : best_tn;                                                                         \
  }
#pragma unroll
        for (int h = 0; h < L::G; h++) {
          const float4 ax = __ldg(n + nx + h), ay = __ldg(n + ny + h), az = __ldg(n + nz + h);
          const float4 bx = __ldg(n + fx + h), by = __ldg(n + fy + h), bz = __ldg(n + fz + h);
          const float4 rf = __ldg(n + L::REF + h);
          RT_WCHILD(ax.x, ay.x, az.x, bx.x, by.x, bz.x, rf.x, 1u << (4 * h + 0))
          RT_WCHILD(ax.y, ay.y, az.y, bx.y, by.y, bz.y, rf.y, 1u << (4 * h + 1))
          RT_WCHILD(ax.z, ay.z, az.z, bx.z, by.z, bz.z, rf.z, 1u << (4 * h + 2))
          RT_WCHILD(ax.w, ay.w, az.w, bx.w, by.w, bz.w, rf.w, 1u << (4 * h + 3))
        }
#undef RT_WCHILD
        if (best_ref != WREF_EMPTY) {
          cur = best_ref;
          cur_tn = best_tn;
        } else if (!pop()) {
          done = true;
          break;
        }
      }
    }
    // ---- leaf phase ----
    const int first = (int)(leaf & WLEAF_FIRST_MASK);
    const int n = leaf ? (int)((leaf >> WLEAF_COUNT_SHIFT) & 7u) : 0;
    if (STATS) st->prims += n;
    if (COOP) {
      coop_leaf_phase(s, *sm, first, n, r, rg, b, ambiguous);
    } else {
      int k = first, end = first + n;
      while (k < end) k += rec_test(s.recs, k, r, rg, b, ambiguous);
    }
  }
  Hit h;
  h.t = 0.0;
  h.rec = REC_NONE;
  h.seq = 0;
  if (!active) return h;
  if (ambiguous || overflow) return world_hit_seq(s, r, t_from, t_to);
  h.t = b.rec == REC_NONE ? 0.0 : b.t;
  h.rec = b.rec;
  return h;
}

// Per-lane if-if form over the wide tree (the loop shape that wins on large scenes: a leaf is tested the moment
// it is reached, no second loop level).  Any subset of lanes may call.  RT_TRAV=16 (W = 4) / 17 (W = 8).
template <bool STATS, int W>
__device__ __forceinline__ Hit world_hit_wide_ifif(const DScene &s, const RayQ &r, double t_from, double t_to,
                                                   TravStats *st) {
  using L = WideLayout<W>;
  Range rg;
  rg.t_from = t_from;
  rg.t_to = t_to;
  rg.tf_f = float_ge(t_from);
  Best b;
  best_init(b, t_to);
  if (s.sky_obj >= 0) {  // sky.cu:18-27: always a candidate at t = 1e9
    double t = 1e9;
    if (t_from <= t && t <= t_to) {
      b.t = t;
      b.t_f = 1e9f;
      b.rec = REC_SKY;
      b.obj = s.sky_obj;
      b.sub = 0;
    }
  }
  bool ambiguous = false, overflow = false;
  if (s.n_recs > 0) {
    const float idx = slab_rcp(r.d.x), idy = slab_rcp(r.d.y), idz = slab_rcp(r.d.z);
    const float nox = -(r.o.x * idx), noy = -(r.o.y * idy), noz = -(r.o.z * idz);
    const int nx = idx < 0.f ? L::HI : 0, ny = L::G + (idy < 0.f ? L::HI : 0), nz = 2 * L::G + (idz < 0.f ? L::HI : 0);
    const int fx = L::HI - nx, fy = 2 * L::G + L::HI - ny, fz = 4 * L::G + L::HI - nz;
    const float t_lo = rg.tf_f - fabsf(rg.tf_f) * 1e-5f;
    float t_hi = b.t_f + fabsf(b.t_f) * 1e-5f;
    uint2 stack[WSTACK];
    int sp = 0;
    uint32_t cur = 0;  // wide node 0 = root
    const float4 *__restrict__ wn = W == 8 ? s.wnodes : s.wnodes4;
    for (;;) {
      if (cur & WLEAF) {
        int k = (int)(cur & WLEAF_FIRST_MASK);
        const int end = k + (int)((cur >> WLEAF_COUNT_SHIFT) & 7u);
        if (STATS) st->prims += end - k;
        while (k < end) k += rec_test(s.recs, k, r, rg, b, ambiguous);
        t_hi = b.t_f + fabsf(b.t_f) * 1e-5f;
      } else {
        const float4 *__restrict__ n = wn + (size_t)L::F4 * cur;
        const uint32_t valid = __float_as_uint(__ldg(n + L::META).x);
        if (STATS) st->nodes += __popc(valid);
        uint32_t best_ref = WREF_EMPTY;
        float best_tn = CUDART_INF_F;
#define RT_WCHILD(NX, NY, NZ, FX, FY, FZ, REF, BIT)                                                            \
  {                                                                                                            \
    float tn_ = fmaxf(fmaxf(__fmaf_rn(NX, idx, nox), __fmaf_rn(NY, idy, noy)), fmaxf(__fmaf_rn(NZ, idz, noz), t_lo)); \
    float tf_ = fminf(fminf(__fmaf_rn(FX, idx, nox), __fmaf_rn(FY, idy, noy)), fminf(__fmaf_rn(FZ, idz, noz), t_hi)); \
    bool hit_ = (tn_ <= tf_ * 1.00001f) && (valid & (BIT));                                                    \
    bool nearer_ = hit_ && tn_ < best_tn;                                                                      \
    uint32_t pr_ = nearer_ ? best_ref : __float_as_uint(REF);                                                  \
    float pt_ = nearer_ ? best_tn : tn_;                                                                       \
    if (hit_ && pr_ != WREF_EMPTY) {                                                                           \
      if (sp < WSTACK)                                                                                         \
        stack[sp++] = make_uint2(pr_, __float_as_uint(pt_));                                                   \
      else                                                                                                     \
        overflow = true;                                                                                       \
    }                                                                                                          \
    best_ref = nearer_ ? __float_as_uint(REF) : best_ref;                                                      \
    best_tn = nearer_ ? tn_ : best_tn;                                                                         \
  }
#pragma unroll
        for (int h = 0; h < L::G; h++) {
          const float4 ax = __ldg(n + nx + h), ay = __ldg(n + ny + h), az = __ldg(n + nz + h);
          const float4 bx = __ldg(n + fx + h), by = __ldg(n + fy + h), bz = __ldg(n + fz + h);
          const float4 rf = __ldg(n + L::REF + h);
          RT_WCHILD(ax.x, ay.x, az.x, bx.x, by.x, bz.x, rf.x, 1u << (4 * h + 0))
          RT_WCHILD(ax.y, ay.y, az.y, bx.y, by.y, bz.y, rf.y, 1u << (4 * h + 1))
          RT_WCHILD(ax.z, ay.z, az.z, bx.z, by.z, bz.z, rf.z, 1u << (4 * h + 2))
          RT_WCHILD(ax.w, ay.w, az.w, bx.w, by.w, bz.w, rf.w, 1u << (4 * h + 3))
        }
#undef RT_WCHILD
        if (best_ref != WREF_EMPTY) {
          cur = best_ref;
          continue;
        }
      }
      // pop the next node that can still hold a closer hit
      bool got = false;
      while (sp > 0) {
        uint2 e = stack[--sp];
        if (__uint_as_float(e.y) > t_hi * 1.00001f) continue;
        cur = e.x;
        got = true;
        break;
      }
      if (!got) break;
    }
  }
  if (ambiguous || overflow) return world_hit_seq(s, r, t_from, t_to);
  Hit h;
  h.t = b.rec == REC_NONE ? 0.0 : b.t;
  h.rec = b.rec;
  h.seq = 0;
  return h;
}

}  // namespace rtc
