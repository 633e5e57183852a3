// rt_flat.cuh — box-list closest hit for scenes of at most FLAT_MAX primitives (Cornell box: 18 parallelograms).
//
// On such scenes a tree does not pay: profiles/r01_q_cornell_fused_ncu_summary.txt shows 47 % of the warp instructions
// of the 4-wide while-while loop in node stepping (stack, leaf detection, child ordering) at 10 of 32 lanes, for 13 box
// tests per ray.  Here every lane tests ALL primitive boxes in one uniform loop (the boxes are a kernel parameter, i.e.
// constant-bank operands: no loads, no divergence, 32 of 32 lanes), keeps the hits as a bit mask, runs the record
// tests of the NEAREST box first, and then walks the remaining mask bits, re-testing each box against the closest hit
// found so far before its records are tested.  Record tests, tie rules and the fallback are those of the tree loops
// (rec_test / best_offer are order independent): identical hits.
//
// The boxes are the single-primitive leaf nodes the LBVH builder writes (rt_build.cu k_emit: nodes[n_bp-1 .. 2n_bp-2],
// padded by k_bp_bounds), copied to the host at commit time.
#pragma once
#include "rt_device.cuh"

namespace rtc {

constexpr int FLAT_MAX = 32;
#ifndef RT_FLAT_UNROLL
#define RT_FLAT_UNROLL 1
#endif
constexpr int kFlatUnroll = RT_FLAT_UNROLL;

struct FlatItems {
  float4 box[2 * FLAT_MAX];  // {centre.xyz, first record}, {half extent.xyz (rounded up), LEAF_FLAG | record count}
  int32_t n;                 // 0: no box list for this scene
  int32_t pad[3];
};
struct FlatNone {};
template <bool ON>
struct FlatParamT {
  using type = FlatNone;
};
template <>
struct FlatParamT<true> {
  using type = FlatItems;
};

#define RT_FLAT_SLAB(LO, HI, T_HI, TN, TF)                                                       \
  {                                                                                              \
    const float ax_ = __fmaf_rn((LO).x, idx, nox), bx_ = __fmaf_rn((HI).x, idx, nox);            \
    const float ay_ = __fmaf_rn((LO).y, idy, noy), by_ = __fmaf_rn((HI).y, idy, noy);            \
    const float az_ = __fmaf_rn((LO).z, idz, noz), bz_ = __fmaf_rn((HI).z, idz, noz);            \
    TN = fmaxf(fmaxf(fminf(ax_, bx_), fminf(ay_, by_)), fmaxf(fminf(az_, bz_), t_lo));           \
    TF = fminf(fminf(fmaxf(ax_, bx_), fmaxf(ay_, by_)), fminf(fmaxf(az_, bz_), T_HI));           \
  }

// self_box: the box the ray starts on (the record its path hit last: Hit::box), or -1.  That box always passes the
// slab test at entry distance t_lo — the ray origin lies inside its padding — so "nearest box first" used to spend
// its first, best-culling slot on the one candidate that almost never yields a hit, and every other candidate was
// then tested in INDEX order against an unshrunk bound (profiles/r02_d_cornell_first_cut_ncu_summary.txt: record tests
// = 45 % of all warp instructions at 12 of 32 lanes, a warp running the 3 candidates of its worst lane for an average
// of 1.66).  Now the nearest box OTHER than self_box goes first (usually the hit, which culls what lies behind it),
// then self_box (its own test cannot be skipped: grazing rays do re-hit their own parallelogram beyond t = 1e-3,
// and the reference shows that acne), then the rest in index order, each re-tested against the closest hit.
template <bool STATS, bool SPH = true, bool EXACT_T = true>
__device__ __forceinline__ Hit world_hit_flat(const DScene &s, const FlatItems &fl, const RayQ &r, double t_from,
                                              double t_to, TravStats *st, int self_box = -1) {
  Range rg;
  rg.t_from = t_from;
  rg.t_to = t_to;
  rg.tf_f = float_ge(t_from);
  rg.lib_pow = s.sphere_pow != 0;
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
  bool ambiguous = false;
  int hit_box = -1;
  const int n = fl.n;
  if (n > 0) {
    const float idx = slab_rcp(r.d.x), idy = slab_rcp(r.d.y), idz = slab_rcp(r.d.z);
    const float nox = -(r.o.x * idx), noy = -(r.o.y * idy), noz = -(r.o.z * idz);
    const float t_lo = rg.tf_f - fabsf(rg.tf_f) * 1e-5f;
    const float t_hi0 = b.t_f + fabsf(b.t_f) * 1e-5f;
    // ---- all boxes, uniformly: hit mask + nearest entry.  Centre / half-extent form: 3 FFMA per axis
    //      (m = c/d - o/d, near = m - h/|d|, far = m + h/|d|) instead of 2 FFMA + min + max.
    const float adx = fabsf(idx), ady = fabsf(idy), adz = fabsf(idz);
#ifdef RT_FLAT_NO_SELF  // experiment: round-1 order (nearest box first, whichever it is)
    self_box = -1;
#endif
    const uint32_t self_bit = self_box >= 0 ? 1u << self_box : 0u;
    uint32_t mask = 0, bit = 1u, bit_min = 0u;
    float tn_min = CUDART_INF_F;
#pragma unroll kFlatUnroll
    for (int c = 0; c < n; c++) {
      const float4 ce = fl.box[2 * c], he = fl.box[2 * c + 1];
      const float mx = __fmaf_rn(ce.x, idx, nox), my = __fmaf_rn(ce.y, idy, noy), mz = __fmaf_rn(ce.z, idz, noz);
      const float tn = fmaxf(fmaxf(__fmaf_rn(-he.x, adx, mx), __fmaf_rn(-he.y, ady, my)),
                             fmaxf(__fmaf_rn(-he.z, adz, mz), t_lo));
      const float tf = fminf(fminf(__fmaf_rn(he.x, adx, mx), __fmaf_rn(he.y, ady, my)),
                             fminf(__fmaf_rn(he.z, adz, mz), t_hi0));
      const bool hit = tn <= tf * 1.00001f;
      mask |= hit ? bit : 0u;
      const bool nearer = hit && tn < tn_min && bit != self_bit;
      tn_min = nearer ? tn : tn_min;
      bit_min = nearer ? bit : bit_min;
      bit += bit;
    }
    if (STATS) st->nodes += n;
    bool self_pending = (mask & self_bit) != 0;
    mask &= ~self_bit;
    int j = __ffs(bit_min) - 1;
    if (j >= 0) {
      mask &= ~(1u << j);
    } else if (self_pending) {  // nothing but the ray's own box
      j = self_box;
      self_pending = false;
    }
    // ---- candidates
    const float4 *__restrict__ leaf = s.nodes + 2 * (size_t)(n - 1);
    uint32_t w_first = 0, w_count = 0;
    if (j >= 0) {
      w_first = __float_as_uint(__ldg(leaf + 2 * j).w);
      w_count = __float_as_uint(__ldg(leaf + 2 * j + 1).w) & ~LEAF_FLAG;
    }
    while (j >= 0) {
      int k = (int)w_first;
      const int end = k + (int)w_count;
      const int rec_before = b.rec;
      while (k < end) {
        int used = rec_test<SPH>(s.recs, k, r, rg, b, ambiguous);
        if (STATS) st->prims += used;
        k += used;
      }
      hit_box = b.rec != rec_before ? j : hit_box;
      const float t_hi = b.t_f + fabsf(b.t_f) * 1e-5f;
      j = -1;
      if (self_pending) {
        self_pending = false;
        j = self_box;
        w_first = __float_as_uint(__ldg(leaf + 2 * j).w);
        w_count = __float_as_uint(__ldg(leaf + 2 * j + 1).w) & ~LEAF_FLAG;
        continue;
      }
      while (mask) {
        const int c = __ffs(mask) - 1;
        mask &= mask - 1;
        const float4 lo = __ldg(leaf + 2 * c), hi = __ldg(leaf + 2 * c + 1);
        float tn, tf;
        RT_FLAT_SLAB(lo, hi, t_hi, tn, tf)
        if (STATS) st->nodes += 1;
        if (tn <= tf * 1.00001f) {
          j = c;
          w_first = __float_as_uint(lo.w);
          w_count = __float_as_uint(hi.w) & ~LEAF_FLAG;
          break;
        }
      }
    }
  }
  resolve_quad_order(s, b, ambiguous);
  finalize_sphere_hit<SPH, EXACT_T>(s, r, rg, b, ambiguous);
  if (ambiguous) return world_hit_seq(s, r, t_from, t_to);
  Hit h;
  h.t = b.rec == REC_NONE ? 0.0 : b.t;
  h.rec = b.rec;
  h.seq = 0;
  h.box = hit_box;
  return h;
}
#undef RT_FLAT_SLAB

}  // namespace rtc
