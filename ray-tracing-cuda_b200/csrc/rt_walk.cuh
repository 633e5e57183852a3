// rt_walk.cuh — resumable traversal rounds for the persistent render loop (k_render_persist).
//
// A closest-hit query is a sequence of ROUNDS: walk inner nodes until a leaf is held, then test that leaf's
// records.  The per-lane loops of rt_device.cuh run all rounds of one query inside a function, so the lanes of
// a warp reconverge after every query and lanes with short queries idle until the longest one ends
// (profiles/r01_g_*: 10 of 32 lanes per instruction).  A walker keeps the state of a query in registers across
// rounds, so the render loop can finish a lane's query, shade it and start the lane's next ray between two
// rounds while the other lanes keep traversing (Aila & Laine 2009, "persistent while-while" with the ray fetch
// replaced by shade + scatter).  Visit order, culling and record tests are those of world_hit_bvh_ww /
// world_hit_wide: identical hits.
#pragma once
#include "rt_wide.cuh"

namespace rtc {

// Best of a fresh query: sky.cu:18-27 is always a candidate at t = 1e9
__device__ __forceinline__ void query_begin(const DScene &s, const Range &rg, Best &b) {
  best_init(b, rg.t_to);
  if (s.sky_obj >= 0) {
    double t = 1e9;
    if (rg.t_from <= t && t <= rg.t_to) {
      b.t = t;
      b.t_f = 1e9f;
      b.rec = REC_SKY;
      b.obj = s.sky_obj;
      b.sub = 0;
    }
  }
}

// ---- binary tree (32-byte nodes), the walk of world_hit_bvh_ww ----
struct BinWalker {
  static constexpr int kStack = STACK_SIZE;
  float idx, idy, idz, t_lo;
  uint32_t ca, cb;
  int sp;
  bool done, overflow;

  __device__ __forceinline__ void begin(const DScene &s, const RayQ &r, const Range &rg) {
    idx = 1.0f / r.d.x, idy = 1.0f / r.d.y, idz = 1.0f / r.d.z;
    t_lo = rg.tf_f - fabsf(rg.tf_f) * 1e-5f;
    ca = s.root_a, cb = s.root_b;
    sp = 0;
    done = s.n_recs <= 0;
    overflow = false;
  }
  __device__ __forceinline__ bool pop(const uint2 *stack) {
    if (sp == 0) return false;
    uint2 e = stack[--sp];
    ca = e.x;
    cb = e.y;
    return true;
  }
  // One round's walk.  Returns the held leaf as (first record, count), count = 0: none (query finished).
  template <bool STATS>
  __device__ __forceinline__ int walk(const DScene &s, const RayQ &r, const Best &b, uint2 *stack, int *first,
                                      TravStats *st) {
    const float4 *__restrict__ nodes = s.nodes;
    const float t_hi = b.t_f + fabsf(b.t_f) * 1e-5f;
    while (!done) {
      if (cb & LEAF_FLAG) {
        *first = (int)ca;
        int n = (int)(cb & ~LEAF_FLAG);
        if (!pop(stack)) done = true;
        return n;
      }
      if (STATS) st->nodes += 2;
      float4 l0 = __ldg(nodes + 2 * (size_t)ca), l1 = __ldg(nodes + 2 * (size_t)ca + 1);
      float4 r0 = __ldg(nodes + 2 * (size_t)cb), r1 = __ldg(nodes + 2 * (size_t)cb + 1);
      float ax = (l0.x - r.o.x) * idx, bx = (l1.x - r.o.x) * idx;
      float ay = (l0.y - r.o.y) * idy, by = (l1.y - r.o.y) * idy;
      float az = (l0.z - r.o.z) * idz, bz = (l1.z - r.o.z) * idz;
      float ln = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), t_lo));
      float lf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), t_hi));
      ax = (r0.x - r.o.x) * idx, bx = (r1.x - r.o.x) * idx;
      ay = (r0.y - r.o.y) * idy, by = (r1.y - r.o.y) * idy;
      az = (r0.z - r.o.z) * idz, bz = (r1.z - r.o.z) * idz;
      float rn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), t_lo));
      float rf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), t_hi));
      bool hl = ln <= lf + fabsf(lf) * 1e-5f, hr = rn <= rf + fabsf(rf) * 1e-5f;
      uint32_t la = __float_as_uint(l0.w), lb = __float_as_uint(l1.w);
      uint32_t ra = __float_as_uint(r0.w), rb = __float_as_uint(r1.w);
      if (hl && hr) {
        bool left_first = ln <= rn;
        if (sp < kStack) {
          stack[sp++] = left_first ? make_uint2(ra, rb) : make_uint2(la, lb);
        } else {
          overflow = true;
        }
        ca = left_first ? la : ra;
        cb = left_first ? lb : rb;
      } else if (hl) {
        ca = la;
        cb = lb;
      } else if (hr) {
        ca = ra;
        cb = rb;
      } else if (!pop(stack)) {
        done = true;
      }
    }
    return 0;
  }
};

// ---- 8-wide tree (256-byte nodes), the walk of world_hit_wide ----
struct WideWalker {
  static constexpr int kStack = WSTACK;
  float idx, idy, idz, nox, noy, noz, t_lo, cur_tn;
  int nx, ny, nz;
  uint32_t cur;
  int sp;
  bool done, overflow;

  __device__ __forceinline__ void begin(const DScene &s, const RayQ &r, const Range &rg) {
    idx = 1.0f / r.d.x, idy = 1.0f / r.d.y, idz = 1.0f / r.d.z;
    nox = -(r.o.x * idx), noy = -(r.o.y * idy), noz = -(r.o.z * idz);
    nx = idx < 0.f ? 6 : 0, ny = idy < 0.f ? 8 : 2, nz = idz < 0.f ? 10 : 4;
    t_lo = rg.tf_f - fabsf(rg.tf_f) * 1e-5f;
    cur = 0;
    cur_tn = 0.f;
    sp = 0;
    done = s.n_recs <= 0;
    overflow = false;
  }
  __device__ __forceinline__ bool pop(const uint2 *stack, float t_hi) {
    while (sp > 0) {
      uint2 e = stack[--sp];
      if (__uint_as_float(e.y) > t_hi * 1.00001f) continue;  // the closest hit moved in front of it
      cur = e.x;
      cur_tn = __uint_as_float(e.y);
      return true;
    }
    return false;
  }
  template <bool STATS>
  __device__ __forceinline__ int walk(const DScene &s, const RayQ &r, const Best &b, uint2 *stack, int *first,
                                      TravStats *st) {
    const float4 *__restrict__ wn = s.wnodes;
    const float t_hi = b.t_f + fabsf(b.t_f) * 1e-5f;
    const int fx = 6 - nx, fy = 10 - ny, fz = 14 - nz;
    // the node popped before the last leaf test may now lie behind the closest hit
    if (!done && cur_tn > t_hi * 1.00001f && !pop(stack, t_hi)) done = true;
    while (!done) {
      if (cur & WLEAF) {
        *first = (int)(cur & WLEAF_FIRST_MASK);
        int n = (int)((cur >> WLEAF_COUNT_SHIFT) & 7u);
        if (!pop(stack, t_hi)) done = true;
        return n;
      }
      const float4 *__restrict__ n = wn + (size_t)WNODE_F4 * cur;
      const uint32_t valid = __float_as_uint(__ldg(n + 14).x);
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
      if (sp < kStack)                                                                                         \
        stack[sp++] = make_uint2(pr_, __float_as_uint(pt_));                                                   \
      else                                                                                                     \
        overflow = true;                                                                                       \
    }                                                                                                          \
    best_ref = nearer_ ? __float_as_uint(REF) : best_ref;                                                      \
    best_tn = nearer_ ? tn_ : best_tn;                                                                         \
  }
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const float4 ax = __ldg(n + nx + h), ay = __ldg(n + ny + h), az = __ldg(n + nz + h);
        const float4 bx = __ldg(n + fx + h), by = __ldg(n + fy + h), bz = __ldg(n + fz + h);
        const float4 rf = __ldg(n + 12 + h);
        RT_WCHILD(ax.x, ay.x, az.x, bx.x, by.x, bz.x, rf.x, 1u << (4 * h + 0))
        RT_WCHILD(ax.y, ay.y, az.y, bx.y, by.y, bz.y, rf.y, 1u << (4 * h + 1))
        RT_WCHILD(ax.z, ay.z, az.z, bx.z, by.z, bz.z, rf.z, 1u << (4 * h + 2))
        RT_WCHILD(ax.w, ay.w, az.w, bx.w, by.w, bz.w, rf.w, 1u << (4 * h + 3))
      }
#undef RT_WCHILD
      if (best_ref != WREF_EMPTY) {
        cur = best_ref;
        cur_tn = best_tn;
      } else if (!pop(stack, t_hi)) {
        done = true;
      }
    }
    return 0;
  }
};

}  // namespace rtc
