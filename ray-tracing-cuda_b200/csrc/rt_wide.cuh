// rt_wide.cuh — closest hit over the 8-wide tree (256-byte nodes, rt_device.cuh / rt_build.cu k_wide_expand): the
// default traversal of meshes between 4 k and 500 k records (bunny: +13 % over the binary if-if loop — 1-2 dependent
// node steps per ray instead of 6, at 25 % more box tests; profiles/r01_o_wide_ifif_sweep.txt).
//
// Per child the box test is 6 FFMA + 2 three-input min/max: the ray selects its near / far planes by address
// (direction octant).  Box culling only has to be conservative (rt_device.cuh, slab2): planes are evaluated as
// fma(plane, 1/d, -(o/d)) with 1/d from slab_rcp (finite); the rounding of that form is covered by the builder's
// relative box padding plus the relative guard bands on t below.
//
// Visit order: nearest hit child first; the other hit children are pushed with their entry distance and
// dropped at pop time once the closest hit has moved in front of them.  Hits are decided by rec_test, identical to
// every other traversal.  (The warp-synchronous, cooperative and quantised variants of round 1 were measured slower
// and removed: DESIGN.md section 6b.)
#pragma once
#include "rt_device.cuh"

namespace rtc {

constexpr int WSTACK = 96;  // up to 7 pushes per level

// Per-lane if-if form over the wide tree (the loop shape that wins on large scenes: a leaf is tested the moment
// it is reached, no second loop level).  Any subset of lanes may call.  RT_TRAV=17.
template <bool STATS, int W, bool SPH = true, bool EXACT_T = true>
__device__ __forceinline__ Hit world_hit_wide_ifif(const DScene &s, const RayQ &r, double t_from, double t_to,
                                                   TravStats *st) {
  using L = WideLayout<W>;
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
    const float4 *__restrict__ wn = s.wnodes;
    for (;;) {
      if (cur & WLEAF) {
        int k = (int)(cur & WLEAF_FIRST_MASK);
        const int end = k + (int)((cur >> WLEAF_COUNT_SHIFT) & 7u);
        if (STATS) st->prims += end - k;
        while (k < end) k += rec_test<SPH>(s.recs, k, r, rg, b, ambiguous);
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
  resolve_quad_order(s, b, ambiguous);
  finalize_sphere_hit<SPH, EXACT_T>(s, r, rg, b, ambiguous);
  if (ambiguous || overflow) return world_hit_seq(s, r, t_from, t_to);
  Hit h;
  h.t = b.rec == REC_NONE ? 0.0 : b.t;
  h.rec = b.rec;
  h.seq = 0;
  h.box = -1;
  return h;
}

}  // namespace rtc
