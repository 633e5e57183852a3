// rt_cwide.cuh — closest hit over the COMPACT wide tree (RT_TRAV=21: 4 children / 64-byte node, 22: 8 children /
// 96-byte node; parity modes 13 / 14).
//
// Why: the soups (1 M - 10 M triangles, incoherent bounce rays) are bound by the latency of DEPENDENT node loads
// (profiles/r01_j_soup10m_fused_ncu_summary.txt: 68 % long_scoreboard stalls, 21 % of the HBM peak).  The binary loop
// fetches 64 B per step and opens two children; the plain wide nodes (rt_wide.cuh) halve / third the number of
// dependent steps but fetch 128 / 256 B in 7 / 14 LDG.128 per step and lost on the soups for that reason.  Here a
// step costs 4 (6) LDG.128 = 64 (96) contiguous bytes — the bytes of ONE binary step — and opens 4 (8) children:
// boxes are 8-bit grid coordinates relative to the node's origin with a power-of-two step per axis (the encoding of
// Ylitie, Karras, Laine 2017, without their octant-ordered child permutation: the order is decided per ray as below).
//
// Plane distance: t = fma(q, (1/d) * step, fma(origin, 1/d, -(o/d))).  (1/d) * step is exact (power of two); the two
// FMA roundings are <= 6e-8 * (|origin - o| / |d| + t): the t part is inside the relative guard bands on t, the
// rest is a plane shift of 6e-8 * (node extent), and k_wide_compress leaves >= 1/1024 of a grid step (4e-6 * node
// extent) of slack on every plane it moved.  Culling only; hits are decided by rec_test, so the
// result is bit-identical to every other traversal (test_traversal_variants_identical, modes 13 / 14).
//
// Visit order as in world_hit_wide_ifif: nearest hit child next, the others pushed with their entry distance and
// dropped at pop time once the closest hit has moved in front of them.
#pragma once
#include "rt_wide.cuh"

namespace rtc {

template <bool STATS, int W>
__device__ __forceinline__ Hit world_hit_cwide(const DScene &s, const RayQ &r, double t_from, double t_to,
                                               TravStats *st) {
  using C = CWideLayout<W>;
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
    const bool negx = idx < 0.f, negy = idy < 0.f, negz = idz < 0.f;
    const float t_lo = rg.tf_f - fabsf(rg.tf_f) * 1e-5f;
    float t_hi = b.t_f + fabsf(b.t_f) * 1e-5f;
    uint2 stack[WSTACK];
    int sp = 0;
    uint32_t cur = 0;  // wide node 0 = root
    const uint4 *__restrict__ cn = W == 8 ? s.cnodes8 : s.cnodes4;
    for (;;) {
      if (cur & WLEAF) {
        int k = (int)(cur & WLEAF_FIRST_MASK);
        const int end = k + (int)((cur >> WLEAF_COUNT_SHIFT) & 7u);
        if (STATS) st->prims += end - k;
        while (k < end) k += rec_test(s.recs, k, r, rg, b, ambiguous);
        t_hi = b.t_f + fabsf(b.t_f) * 1e-5f;
      } else {
        const uint4 *__restrict__ n = cn + (size_t)C::U4 * cur;
        const uint4 h0 = __ldg(n);
        const uint32_t valid = h0.w >> 24;
        if (STATS) st->nodes += __popc(valid);
        // ray in the node's grid: t(q) = q * a + c per axis
        const float ax = idx * __uint_as_float((h0.w & 0xffu) << 23);
        const float ay = idy * __uint_as_float((h0.w >> 8 & 0xffu) << 23);
        const float az = idz * __uint_as_float((h0.w >> 16 & 0xffu) << 23);
        const float cx = __fmaf_rn(__uint_as_float(h0.x), idx, nox);
        const float cy = __fmaf_rn(__uint_as_float(h0.y), idy, noy);
        const float cz = __fmaf_rn(__uint_as_float(h0.z), idz, noz);
        uint32_t best_ref = WREF_EMPTY;
        float best_tn = CUDART_INF_F;
#define RT_CBYTE(WORD, I) __uint2float_rn(((WORD) >> (8 * (I))) & 0xffu)
#define RT_CCHILD(NX, NY, NZ, FX, FY, FZ, I, REF, BIT)                                                         \
  {                                                                                                            \
    float tn_ = fmaxf(fmaxf(__fmaf_rn(RT_CBYTE(NX, I), ax, cx), __fmaf_rn(RT_CBYTE(NY, I), ay, cy)),           \
                      fmaxf(__fmaf_rn(RT_CBYTE(NZ, I), az, cz), t_lo));                                        \
    float tf_ = fminf(fminf(__fmaf_rn(RT_CBYTE(FX, I), ax, cx), __fmaf_rn(RT_CBYTE(FY, I), ay, cy)),           \
                      fminf(__fmaf_rn(RT_CBYTE(FZ, I), az, cz), t_hi));                                        \
    bool hit_ = (tn_ <= tf_ * 1.00001f) && (valid & (BIT));                                                    \
    bool nearer_ = hit_ && tn_ < best_tn;                                                                      \
    uint32_t pr_ = nearer_ ? best_ref : (REF);                                                                 \
    float pt_ = nearer_ ? best_tn : tn_;                                                                       \
    if (hit_ && pr_ != WREF_EMPTY) {                                                                           \
      if (sp < WSTACK)                                                                                         \
        stack[sp++] = make_uint2(pr_, __float_as_uint(pt_));                                                   \
      else                                                                                                     \
        overflow = true;                                                                                       \
    }                                                                                                          \
    best_ref = nearer_ ? (REF) : best_ref;                                                                     \
    best_tn = nearer_ ? tn_ : best_tn;                                                                         \
  }
#define RT_CGROUP(LX, LY, LZ, HX, HY, HZ, RF, B0)                                                              \
  {                                                                                                            \
    const uint32_t nx_ = negx ? (HX) : (LX), fx_ = negx ? (LX) : (HX);                                         \
    const uint32_t ny_ = negy ? (HY) : (LY), fy_ = negy ? (LY) : (HY);                                         \
    const uint32_t nz_ = negz ? (HZ) : (LZ), fz_ = negz ? (LZ) : (HZ);                                         \
    RT_CCHILD(nx_, ny_, nz_, fx_, fy_, fz_, 0, RF.x, 1u << ((B0) + 0))                                         \
    RT_CCHILD(nx_, ny_, nz_, fx_, fy_, fz_, 1, RF.y, 1u << ((B0) + 1))                                         \
    RT_CCHILD(nx_, ny_, nz_, fx_, fy_, fz_, 2, RF.z, 1u << ((B0) + 2))                                         \
    RT_CCHILD(nx_, ny_, nz_, fx_, fy_, fz_, 3, RF.w, 1u << ((B0) + 3))                                         \
  }
        if constexpr (W == 4) {
          const uint4 h1 = __ldg(n + 1), h2 = __ldg(n + 2), rf = __ldg(n + 3);
          RT_CGROUP(h1.x, h1.y, h1.z, h1.w, h2.x, h2.y, rf, 0)
        } else {
          const uint4 h1 = __ldg(n + 1), h2 = __ldg(n + 2), h3 = __ldg(n + 3), rf0 = __ldg(n + 4), rf1 = __ldg(n + 5);
          RT_CGROUP(h1.x, h1.z, h2.x, h2.z, h3.x, h3.z, rf0, 0)
          RT_CGROUP(h1.y, h1.w, h2.y, h2.w, h3.y, h3.w, rf1, 4)
        }
#undef RT_CGROUP
#undef RT_CCHILD
#undef RT_CBYTE
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
