// rt_coop.cuh — warp-cooperative closest hit: the record tests of a warp's rays are pooled and
// spread over all 32 lanes.
//
// Why: ncu on the per-lane loops (profiles/r01_a_*, r01_g_*) shows the record tests — Möller–Trumbore as
// TriangleHit writes it (utils.cu:49-85), twice per Parallelogram (parallelogram.cu:17-44) — running with
// 4-5 of 32 lanes: the lanes of a warp reach their leaves at different times, a leaf holds 1-2 quads, and
// each lane tests its own records one after the other.  Here the traversal is warp-synchronous:
//   walk : every lane that still traverses walks inner nodes until it holds a leaf (as in while-while);
//   pool : the records of all held leaves become work items (owner lane, record) in shared memory;
//   test : item g is tested by lane g % 32 against the OWNER's ray (read from shared memory), whatever
//          lane owns it — a quad's two triangles run side by side on two lanes, idle lanes work too;
//   apply: each owner folds the results of its items in record order with rec_test's own rules.
// The arithmetic per (ray, record) pair and the order in which an owner sees its records are those of
// rec_test, so hits are bit-identical to the per-lane traversals (tests/test_gpu_parity.py).
// Measured (profiles/r01_h_variants_sweep.txt): with the while-while loop the record tests already run at 12 lanes
// and the pooling (scan, 3 __syncwarp, shared-memory round trip per round) costs more than it saves: 1559 vs
// 1909 Msamples/s on Cornell.  NOT the default; kept as a variant (RT_TRAV=9/11) with its evidence.
#pragma once
#include "rt_device.cuh"

namespace rtc {

constexpr int COOP_MAX_LEAF = 4;              // records per leaf the pool takes (the builder's default max_leaf)
constexpr int COOP_CAP = 32 * COOP_MAX_LEAF;  // work items per round

struct CoopWarp {  // shared memory of one warp: 3456 B
  float ray[6][32];              // o.xyz, d.xyz of every lane's current ray
  uint4 res[COOP_CAP];           // per item: {t bits | double lo, sub | double hi, obj, kind | hit << 8}
  uint32_t item_rec[COOP_CAP];   // record index
  uint8_t item_owner[COOP_CAP];  // owner lane
};

// One work item: record `rec` against ray `q`.  The decision arithmetic is tri_test / sphere_test.
__device__ __forceinline__ uint4 coop_test_item(const float4 *__restrict__ recs, uint32_t rec, const RayQ &q,
                                                const Range &rg) {
  float4 r0 = __ldg(recs + 3 * (size_t)rec), r1 = __ldg(recs + 3 * (size_t)rec + 1), r2 = __ldg(recs + 3 * (size_t)rec + 2);
  uint32_t kind = __float_as_uint(r2.w) & KIND_MASK;
  uint4 out;
  out.z = __float_as_uint(r0.w);
  if (kind == K_SPHERE) {
    double t = 0;
    bool hit = sphere_test(xyz(r0), rec_radius(r1), q, rg.t_from, &t);
    out.x = (uint32_t)__double2loint(t);
    out.y = (uint32_t)__double2hiint(t);
    out.w = kind | (hit ? 256u : 0u);
  } else {
    float t = 0, u, v;
    bool hit = tri_test(xyz(r0), xyz(r1), xyz(r2), q, rg.tf_f, &t, &u, &v);
    out.x = __float_as_uint(t);
    out.y = __float_as_uint(r1.w);
    out.w = kind | (hit ? 256u : 0u);
  }
  return out;
}

// Folds the results of one leaf (items [base, base + n) of the pool, records first .. first + n) into the
// owner's closest hit: rec_test's rules, record by record.
__device__ __forceinline__ void coop_apply(const CoopWarp &sm, int base, int n, int first, Best &b, bool &ambiguous) {
  int k = 0;
  while (k < n) {
    uint4 a = sm.res[base + k];
    uint32_t kind = a.w & 255u;
    bool hit = (a.w >> 8) != 0;
    if (kind == K_QUAD0) {
      uint4 c = sm.res[base + k + 1];  // its QUAD1 (leaves keep pairs together)
      bool h1 = (c.w >> 8) != 0;
      float t0 = __uint_as_float(a.x), t1 = __uint_as_float(c.x);
      if (hit) {
        if (h1 && t1 < t0) ambiguous = true;
        if (t0 <= b.t_f) {
          if (best_offer(b, (double)t0, first + k, (int32_t)a.z, 0)) b.t_f = t0;
        }
      } else if (h1 && t1 <= b.t_f) {
        if (best_offer(b, (double)t1, first + k + 1, (int32_t)c.z, 1)) b.t_f = t1;
      }
      k += 2;
    } else if (kind == K_SPHERE) {
      double t = __hiloint2double((int)a.y, (int)a.x);
      if (hit && t <= b.t) {
        if (best_offer(b, t, first + k, (int32_t)a.z, 0)) b.t_f = float_le(t);
      }
      k += 1;
    } else {  // K_TRI (a lone K_QUAD1 never starts a leaf)
      float t = __uint_as_float(a.x);
      if (kind == K_TRI && hit && t <= b.t_f) {
        if (best_offer(b, (double)t, first + k, (int32_t)a.z, (int32_t)a.y)) b.t_f = t;
      }
      k += 1;
    }
  }
}

// The leaf phase of one traversal round.  EVERY lane of the warp calls it; a lane holding a leaf passes
// its records [first, first + n), the others n = 0.  sm.ray must hold every lane's ray.
__device__ __forceinline__ void coop_leaf_phase(const DScene &s, CoopWarp &sm, int first, int n, const RayQ &r,
                                                const Range &rg, Best &b, bool &ambiguous) {
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  if (n > COOP_MAX_LEAF) {  // oversized leaf (RT_MAX_LEAF override): this lane tests it alone
    int k = first, end = first + n;
    while (k < end) k += rec_test(s.recs, k, r, rg, b, ambiguous);
  }
  // ---- pool: held leaves -> work items ----
  const int nc = n <= COOP_MAX_LEAF ? n : 0;
  int incl = nc;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int v = __shfl_up_sync(FULL, incl, o);
    if (lane >= o) incl += v;
  }
  const int base = incl - nc;
  const int total = __shfl_sync(FULL, incl, 31);
  if (total == 0) return;
  for (int k = 0; k < nc; k++) {
    sm.item_rec[base + k] = (uint32_t)(first + k);
    sm.item_owner[base + k] = (uint8_t)lane;
  }
  __syncwarp();
  // ---- test: item g on lane g % 32 ----
  for (int g = lane; g < total; g += 32) {
    const uint32_t rec = sm.item_rec[g];
    const int ow = sm.item_owner[g];
    RayQ q;
    q.o = mk(sm.ray[0][ow], sm.ray[1][ow], sm.ray[2][ow]);
    q.d = mk(sm.ray[3][ow], sm.ray[4][ow], sm.ray[5][ow]);
    sm.res[g] = coop_test_item(s.recs, rec, q, rg);
  }
  __syncwarp();
  // ---- apply ----
  if (nc) coop_apply(sm, base, nc, first, b, ambiguous);
  __syncwarp();
}

// HitableList::Hit for the rays of one warp.  EVERY lane of the warp must call it (lanes without a ray
// pass active = false and only work for the others).  sm is this warp's CoopWarp.
template <bool STATS>
__device__ __forceinline__ Hit world_hit_coop(const DScene &s, const RayQ &r, bool active, double t_from, double t_to,
                                              CoopWarp &sm, TravStats *st) {
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
  sm.ray[0][lane] = r.o.x;
  sm.ray[1][lane] = r.o.y;
  sm.ray[2][lane] = r.o.z;
  sm.ray[3][lane] = r.d.x;
  sm.ray[4][lane] = r.d.y;
  sm.ray[5][lane] = r.d.z;
  const float idx = 1.0f / r.d.x, idy = 1.0f / r.d.y, idz = 1.0f / r.d.z;
  const float t_lo = rg.tf_f - fabsf(rg.tf_f) * 1e-5f;
  uint2 stack[STACK_SIZE];
  int sp = 0;
  uint32_t ca = s.root_a, cb = s.root_b;
  const float4 *__restrict__ nodes = s.nodes;
  auto pop = [&]() -> bool {
    if (sp == 0) return false;
    uint2 e = stack[--sp];
    ca = e.x;
    cb = e.y;
    return true;
  };
  while (__any_sync(FULL, !done)) {
    // ---- walk: inner nodes until this lane holds a leaf (world_hit_bvh_ww, no speculation) ----
    uint32_t leaf_a = 0, leaf_b = 0;
    if (!done) {
      for (;;) {
        if (cb & LEAF_FLAG) {
          leaf_a = ca;
          leaf_b = cb;
          if (!pop()) done = true;
          break;
        }
        if (STATS) st->nodes += 2;
        float4 l0 = __ldg(nodes + 2 * (size_t)ca), l1 = __ldg(nodes + 2 * (size_t)ca + 1);
        float4 r0 = __ldg(nodes + 2 * (size_t)cb), r1 = __ldg(nodes + 2 * (size_t)cb + 1);
        float t_hi = b.t_f + fabsf(b.t_f) * 1e-5f;
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
          if (sp < STACK_SIZE) {
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
        } else if (!pop()) {
          done = true;
          break;
        }
      }
    }
    // ---- pool / test / apply ----
    const int n = leaf_b ? (int)(leaf_b & ~LEAF_FLAG) : 0;
    if (STATS) st->prims += n;
    coop_leaf_phase(s, sm, (int)leaf_a, n, r, rg, b, ambiguous);
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

}  // namespace rtc
