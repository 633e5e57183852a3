// rt_device.cuh — device-side data layout and arithmetic of the B200 path-tracing core.
//
// Everything that decides a hit or a colour is written in the reference's expression order with
// one IEEE operation per source operation: this translation unit set is compiled with
// --fmad=false, so `a*b+c` is FMUL+FADD exactly like the reference built with --fmad=false (and
// like oracle/rt_oracle.cpp built with -ffp-contract=off).  Box culling, which only has to be
// conservative, is free to use explicit __fmaf_rn.
//
// glm 0.9.9.8 definitions the reference relies on (third_party/glm/glm/detail/func_geometric.inl):
//   dot :47-55   cross :65-77   normalize :80-89 (v * inversesqrt, func_exponential.inl:134-139)
//   reflect :102-109   refract :111-123
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include "../../include/rtcore.h"

namespace rtc {

// ------------------------------------------------------------------------------------------------
// vec3 in glm's expression order
// ------------------------------------------------------------------------------------------------
struct V3 {
  float x, y, z;
};
__host__ __device__ __forceinline__ V3 mk(float x, float y, float z) { return V3{x, y, z}; }
__host__ __device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__host__ __device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__host__ __device__ __forceinline__ V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
__host__ __device__ __forceinline__ V3 operator*(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
__host__ __device__ __forceinline__ V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
__host__ __device__ __forceinline__ V3 operator*(float s, V3 a) { return mk(s * a.x, s * a.y, s * a.z); }
__host__ __device__ __forceinline__ float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__host__ __device__ __forceinline__ V3 cross(V3 x, V3 y) {
  return mk(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}
__host__ __device__ __forceinline__ float length(V3 v) { return sqrtf(dot(v, v)); }
__host__ __device__ __forceinline__ V3 normalize(V3 v) { return v * (1.0f / sqrtf(dot(v, v))); }
__host__ __device__ __forceinline__ V3 reflect(V3 I, V3 N) { return I - N * dot(N, I) * 2.0f; }
__host__ __device__ __forceinline__ V3 refract(V3 I, V3 N, float eta) {
  float const d = dot(N, I);
  float const k = 1.0f - eta * eta * (1.0f - d * d);
  return (k >= 0.0f) ? (eta * I - (eta * d + sqrtf(k)) * N) : mk(0.f, 0.f, 0.f);
}
__host__ __device__ __forceinline__ V3 xyz(float4 a) { return mk(a.x, a.y, a.z); }

// ------------------------------------------------------------------------------------------------
// Records (48 B = 3 x float4): the type-tagged SoA replacement of the virtual Hitable objects.
//   r0 = { a.xyz , obj  }   r1 = { b.xyz , sub }   r2 = { c.xyz , kind | flags | material << 4 }
// kind TRI   : a = v0, b = v1 - v0, c = v2 - v0                 (Triangle, Face<>: utils.cu:54-55)
// kind QUAD0 : first triangle (p0,p1,p2) of a Parallelogram; the next record is its QUAD1
// kind QUAD1 : second triangle (p1,p2,p3)                        (parallelogram.cu:25-41)
// kind SPHERE: a = centre, b.x/b.y = bit halves of the double radius (sphere.cuh:20)
// kind SKY   : no geometry; only present in the sequential list     (sky.cu:18-27)
// ------------------------------------------------------------------------------------------------
enum : uint32_t { K_TRI = 0, K_QUAD0 = 1, K_QUAD1 = 2, K_SPHERE = 3, K_SKY = 4 };
constexpr uint32_t KIND_BITS = 4, KIND_MASK = 7;
constexpr uint32_t FLAG_BARY = 8;  // top-level Triangle: HitRecord::u,v are the barycentrics (triangle.cu:11-19)
constexpr uint32_t LEAF_FLAG = 0x80000000u;
constexpr int REC_NONE = -1, REC_SKY = -2;
constexpr int STACK_SIZE = 64;

struct DMaterial {  // 32 B, two float4 loads
  float4 color_fuzz;  // albedo / attenuation / emission, Metal fuzz
  int32_t type;       // RT_MAT_*
  int32_t texture;    // -1: constant colour
  float eta_in;       // (float)ior / 1.0f   inner -> outer   dielectric.cu:27-28
  float eta_out;      // 1.0f / (float)ior   outer -> inner   dielectric.cu:36-37
};

struct DTexture {
  const uint8_t *rgba;
  int32_t height, width;
  int64_t pitch;
};

struct DScene {
  const float4 *nodes;     // 32-byte nodes: {lo.xyz, a}, {hi.xyz, b}
  const float4 *recs;      // records in BVH (Morton) order, 3 float4 each
  const float4 *recs_seq;  // records in list order (Append order; faces in post-sort order; incl. SKY)
  const float *face_uv;    // per BVH-order record: 6 floats (Face<true> texcoords), or null
  const float *face_uv_seq;
  const DMaterial *mats;
  const DTexture *textures;
  uint32_t root_a, root_b;  // the root's (a, b) words
  int32_t n_recs, n_recs_seq, n_nodes;
  int32_t sky_obj, sky_mat;  // first Sky in the list, or -1
  int32_t has_spheres;
  int32_t sphere_pow;  // 1: Sphere::Hit's pow() calls go to CUDA's pow like the reference's (sphere_test); 0: IEEE x * x / sqrt
};

struct DCamera {  // camera.cuh:12-14
  V3 position, llc, horizontal, vertical, u, v;
  int32_t is_defocus;
  float lens_radius;
};

// Closest-hit result of a query (superset of HitRecord, hitable.cuh:13-17)
struct Hit {
  double t;
  int32_t rec;   // record index (BVH order in mode 0, list order in mode 1), REC_SKY, REC_NONE
  int32_t seq;   // 1: rec indexes recs_seq
  int32_t box;   // box-list traversal: index of the box whose record was hit (the box a scattered ray starts on), else -1
};

// ------------------------------------------------------------------------------------------------
// Philox4x32-10 keyed (seed) with counter (pixel, sample, dimension, iteration)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                              uint32_t k1, uint32_t out[4]) {
#pragma unroll
  for (int r = 0; r < 10; r++) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0;
    c1 = lo1;
    c2 = n2;
    c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0;
  out[1] = c1;
  out[2] = c2;
  out[3] = c3;
}
// curand_uniform's mapping to (0, 1] (curand_kernel.h _curand_uniform)
__device__ __forceinline__ float u01(uint32_t x) {
  return __uint2float_rn(x) * 2.3283064365386963e-10f + 1.16415321826934814453125e-10f;
}

// ------------------------------------------------------------------------------------------------
// Ray with the float thresholds of the reference's double range test
// ------------------------------------------------------------------------------------------------
struct RayQ {
  V3 o, d;
};

struct Range {
  double t_from, t_to;  // as passed to Hit()
  float tf_f;           // smallest float >= t_from : (double)t >= t_from  <=>  t >= tf_f
  bool lib_pow;         // DScene::sphere_pow (sphere_test)
};
__host__ __device__ __forceinline__ float float_ge(double x) {  // smallest float >= x
#ifdef __CUDA_ARCH__
  return __double2float_ru(x);
#else
  float f = (float)x;
  if ((double)f < x) f = nextafterf(f, INFINITY);
  return f;
#endif
}
__device__ __forceinline__ float float_le(double x) { return __double2float_rd(x); }  // largest float <= x

constexpr float DET_EPS_F = 1.00000001168609742308035492897e-7f;  // smallest float >= 1e-7 (utils.cu:53,60)
static_assert((double)DET_EPS_F >= 1e-7, "DET_EPS_F must round up");

// utils.cu:49-85 TriangleHit, decision part.  Returns true when u, v pass and t >= tf_f; the
// caller applies the upper bound.  a = v0, e1 = v1 - v0, e2 = v2 - v0.
// Straight-line form: det, u, v, t are always computed (the same IEEE operations on the same operands as the
// reference's early-exit form, so the same bits) and the five rejections are ANDed.  Each rejection is written as
// the negation of the reference's `if (...) return false`, which keeps its NaN behaviour.  In the per-lane loops the
// early exits only pay when every lane of the warp leaves (profiles/r01_x_cornell_boxlist_ncu_summary.txt: 11.5 -> 6.7
// lanes across the function, i.e. the warp almost always runs to the end anyway) and cost four divergent branches.
__device__ __forceinline__ bool tri_test(V3 a, V3 e1, V3 e2, const RayQ &r, float tf_f, float *out_t, float *out_u,
                                         float *out_v) {
  V3 pvec = cross(r.d, e2);
  float det = dot(e1, pvec);
  float invDet = 1 / det;
  V3 tvec = r.o - a;
  float u = dot(tvec, pvec) * invDet;
  V3 qvec = cross(tvec, e1);
  float v = dot(r.d, qvec) * invDet;
  float t = dot(e2, qvec) * invDet;
  *out_t = t;
  *out_u = u;
  *out_v = v;
  return !(fabsf(det) < DET_EPS_F) & !(u < 0 || u > 1) & !(v < 0 || u + v > 1) & (t >= tf_f);
}

__device__ __forceinline__ double rec_radius(float4 r1) {
  return __hiloint2double(__float_as_int(r1.y), __float_as_int(r1.x));
}

// sphere.cu:11-44 Sphere::Hit, decision part: the root the reference would return for an
// unbounded t_to (near root if >= t_from, else far root), or false.
//
// The reference squares and takes roots with pow(x, 2) / pow(x, 0.5) (sphere.cu:13-21), which resolve to CUDA's
// double pow(double, double).  That function is NOT correctly rounded: on the B200 its result differs from x * x on
// 6 % and from sqrt(x) on 14 % of the operands of these scenes, by one ulp (bench.py hit_bit_exact.pow_kat_device; the
// KAT kernel lives in baseline/ref_cuda_harness.cu).  lib_pow = true (the default, DScene::sphere_pow) therefore calls
// the same pow() on the same operands: primary-ray t equals the reference's own CUDA code (built --fmad=false) bit for
// bit.  The five pow calls cost ~10x the rest of the test, so they only run for rays the plain IEEE form (x * x, sqrt)
// cannot rule out: a ray misses for certain when the IEEE discriminant is below -E, E = 1e-14 * (b^2 + 4a(|oc|^2 +
// r^2)) — a bound ~40x the largest difference one ulp in each pow can make.  lib_pow = false is the IEEE form alone: what the
// CPU oracle computes (oracle/rt_oracle.cpp, pinned to the reference's HOST build, whose glibc pow is correctly
// rounded on all but ~1e-5 of the operands), used by the GPU-vs-oracle tests (RT_SPHERE_POW=ieee at commit).
#ifndef RT_SPHERE_INLINE
#define RT_SPHERE_INLINE __forceinline__
#endif
// The reference's quadratic with CUDA's pow, out of line and by value: its ~1000 instructions must not weigh on the
// registers of the loops that call sphere_test.  Returns {a, sqrt of the discriminant, or -1 for a miss}.
struct SphereLibPow {
  double a, sq;
};
static __device__ __noinline__ SphereLibPow sphere_quadratic_libdevice(double len_d, double len_oc, double radius, double b) {
  SphereLibPow o;
  o.a = pow(len_d, 2.0);
  const double cc = pow(len_oc, 2.0) - pow(radius, 2.0);
  const double disc = pow(b, 2.0) - 4 * o.a * cc;
  o.sq = disc < 0 ? -1.0 : pow(disc, 0.5);
  return o;
}
static __device__ RT_SPHERE_INLINE bool sphere_test(V3 c, double radius, const RayQ &r, double t_from, bool lib_pow,
                                                    double *out_t) {
  double len_d = (double)length(r.d);
  double a = len_d * len_d;
  V3 oc = r.o - c;
  double b = (double)(2 * dot(r.d, oc));
  double len_oc = (double)length(oc);
  double cc = len_oc * len_oc - radius * radius;
  double disc = b * b - 4 * a * cc;
  double sq;
  if (lib_pow) {
    const double guard = 1e-14 * (b * b + 4 * a * (len_oc * len_oc + radius * radius));
    if (disc < -guard) return false;
    const SphereLibPow q = sphere_quadratic_libdevice(len_d, len_oc, radius, b);
    if (q.sq < 0) return false;  // pow(disc, 0.5) of a non-negative disc is never negative
    a = q.a;
    sq = q.sq;
  } else {
    if (disc < 0) return false;
    sq = sqrt(disc);
  }
  double t = (-b - sq) / (2 * a);
  if (t_from <= t) {
    *out_t = t;
    return true;
  }
  t = (-b + sq) / (2 * a);
  if (t_from <= t) {
    *out_t = t;
    return true;
  }
  return false;
}

// Sphere::Hit for the traversal loops when lib_pow is on.  Calling CUDA's pow five times for every sphere a ray MIGHT
// hit cost 45 % of the spheres frame and 44 % of birthday's (one sphere; profiles/r02_f_*).  A traversal only has to
// ORDER candidates, and the IEEE form of the test is within a known distance of the pow form: the root t with
//   dt = guard / (sqrt(disc) * 2a) + 1e-13 |t|,   guard = 1e-14 (b^2 + 4a(|oc|^2 + r^2))
// (disc differs by < guard / 40, a and the roots by 1 ulp each; measured difference <= 8e-13 |t|).  Every decision that
// holds over the whole interval [t - dt, t + dt] — in range, nearer or farther than the best so far — is taken on the
// IEEE value; anything closer than that is UNSURE and sends the query to the sequential scan, which calls pow for
// every sphere (world_hit_seq; as rare as two surfaces within 1e-12 of each other).  The winner's t is then computed
// ONCE with pow (finalize_sphere_hit): the same value the reference's CUDA code returns.
enum : int { SPH_MISS = 0, SPH_HIT = 1, SPH_UNSURE = 2 };
struct SphereApprox {
  int state;
  double t, dt;
};
static __device__ RT_SPHERE_INLINE SphereApprox sphere_test_interval(V3 c, double radius, const RayQ &r, double t_from) {
  SphereApprox o;
  o.t = o.dt = 0;
  const double len_d = (double)length(r.d);
  const double a = len_d * len_d;
  const V3 oc = r.o - c;
  const double b = (double)(2 * dot(r.d, oc));
  const double len_oc = (double)length(oc);
  const double l2 = len_oc * len_oc, r2 = radius * radius;
  const double disc = b * b - 4 * a * (l2 - r2);
  const double guard = 1e-14 * (b * b + 4 * a * (l2 + r2));
  if (disc < -guard) {
    o.state = SPH_MISS;
    return o;
  }
  if (!(disc > guard)) {  // tangent within the guard (or NaN)
    o.state = SPH_UNSURE;
    return o;
  }
  const double sq = sqrt(disc), inv = 1.0 / (2 * a);
  const double dsq = guard / sq * inv;
  double t = (-b - sq) * inv, dt = dsq + 1e-13 * fabs(t);
  if (!(t >= t_from + dt)) {
    if (t > t_from - dt) {
      o.state = SPH_UNSURE;
      return o;
    }
    t = (-b + sq) * inv;
    dt = dsq + 1e-13 * fabs(t);
    if (!(t >= t_from + dt)) {
      o.state = t > t_from - dt ? SPH_UNSURE : SPH_MISS;
      return o;
    }
  }
  o.state = SPH_HIT;
  o.t = t;
  o.dt = dt;
  return o;
}

// ------------------------------------------------------------------------------------------------
// Closest-hit bookkeeping with the reference's tie rules made order independent:
//   smaller t wins; equal t: the lower list index wins (HitableList::Hit replaces only on
//   hr.t < t_to, hitable_list.cu:14-21); inside one mesh the higher post-sort face index wins
//   (leaf loop t <= t_to, bvh.cuh:127-134; right child wins, :152-156).
// ------------------------------------------------------------------------------------------------
struct Best {
  double t;    // current closest t (HitRecord::t is a double)
  float t_f;   // largest float <= t: a float candidate passes `t <= t_to` iff t <= t_f
  int32_t rec;
  int32_t obj, sub;
  double t_err;  // > 0: t is the IEEE stand-in of a sphere hit (sphere_test_interval), exact value within t +- t_err
  // one parallelogram whose two triangles both hit with the second one closer (resolve_quad_order); amb_rec < 0: none
  int32_t amb_rec;  // its QUAD1 record
  float amb_t0, amb_t1;
};

__device__ __forceinline__ void best_init(Best &b, double t_to) {
  b.t = t_to;
  b.t_f = float_le(t_to);
  b.rec = REC_NONE;
  b.obj = 0x7fffffff;
  b.sub = -1;
  b.t_err = 0.0;
  b.amb_rec = -1;
  b.amb_t0 = b.amb_t1 = 0.f;
}
// `cond`: the candidate passed its own test and t <= the current bound.  Predicated selects, no branches.
__device__ __forceinline__ bool best_offer(Best &b, bool cond, double t, int32_t rec, int32_t obj, int32_t sub) {
  // precondition when cond: t <= b.t
  const bool tie = (t == b.t) & (b.rec != REC_NONE);
  const bool win = !tie | (obj < b.obj) | ((obj == b.obj) & (sub > b.sub));
  const bool take = cond & win;
  b.t = take ? t : b.t;
  b.rec = take ? rec : b.rec;
  b.obj = take ? obj : b.obj;
  b.sub = take ? sub : b.sub;
  return take;
}

// One record against the ray.  Returns the number of records consumed (2 for a quad).
// Parallelogram::Hit's "first triangle that hits wins" (parallelogram.cu:25-41) depends on the t_to it is called
// with when both triangles hit and the second is closer (a ray through the shared diagonal, within rounding): the
// traversal reports the first triangle, as the reference does when nothing listed EARLIER is closer, remembers the pair
// in Best::amb_*, and resolve_quad_order settles the rest once the closest hit of everything else is known.
// `ambiguous` (the caller re-runs the ray through the sequential scan) is left for what that cannot settle.
// SPH = false: the scene has no Sphere (DScene::has_spheres) — the FP64 test and its registers are compiled out.
template <bool SPH = true>
__device__ __forceinline__ int rec_test(const float4 *__restrict__ recs, int k, const RayQ &r, const Range &rg, Best &b,
                                        bool &ambiguous) {
  float4 r0 = __ldg(recs + 3 * (size_t)k), r1 = __ldg(recs + 3 * (size_t)k + 1), r2 = __ldg(recs + 3 * (size_t)k + 2);
  uint32_t kind = __float_as_uint(r2.w) & KIND_MASK;
  if (kind == K_QUAD0) {  // both triangles of the parallelogram in one straight-line block
    float t0, t1, u, v;
    float4 s0 = __ldg(recs + 3 * (size_t)k + 3), s1 = __ldg(recs + 3 * (size_t)k + 4), s2 = __ldg(recs + 3 * (size_t)k + 5);
    const bool h0 = tri_test(xyz(r0), xyz(r1), xyz(r2), r, rg.tf_f, &t0, &u, &v);
    const bool h1 = tri_test(xyz(s0), xyz(s1), xyz(s2), r, rg.tf_f, &t1, &u, &v);
    if (h0 & h1 & (t1 < t0)) {
      ambiguous |= b.amb_rec >= 0;  // a second such pair on one ray: the sequential scan decides
      b.amb_rec = k + 1;
      b.amb_t0 = t0;
      b.amb_t1 = t1;
    }
    // parallelogram.cu:25-41: the first triangle that hits wins; the second is only looked at when the first misses
    const bool first = h0;
    const float t = first ? t0 : t1;
    const bool cond = (h0 | h1) & (t <= b.t_f);
    if constexpr (SPH) {  // too close to an approximate sphere hit to call
      ambiguous |= cond & (b.t_err > 0) & ((double)t >= b.t - b.t_err);
      b.t_err = cond ? 0.0 : b.t_err;
    }
    if (best_offer(b, cond, (double)t, first ? k : k + 1, __float_as_int(r0.w), first ? 0 : 1)) b.t_f = t;
    return 2;
  }
  if (kind == K_TRI) {
    float t, u, v;
    const bool h = tri_test(xyz(r0), xyz(r1), xyz(r2), r, rg.tf_f, &t, &u, &v);
    const bool cond = h & (t <= b.t_f);
    if constexpr (SPH) {
      ambiguous |= cond & (b.t_err > 0) & ((double)t >= b.t - b.t_err);
      b.t_err = cond ? 0.0 : b.t_err;
    }
    if (best_offer(b, cond, (double)t, k, __float_as_int(r0.w), __float_as_int(r1.w))) b.t_f = t;
    return 1;
  }
  if constexpr (SPH) {
    if (kind == K_SPHERE) {
      if (rg.lib_pow) {  // decide on the IEEE interval, compute the winner with pow afterwards (finalize_sphere_hit)
        const SphereApprox q = sphere_test_interval(xyz(r0), rec_radius(r1), r, rg.t_from);
        const double e = q.dt + b.t_err;
        if (q.state == SPH_UNSURE) {
          ambiguous = true;
        } else if (q.state == SPH_HIT && q.t <= b.t + e) {
          if (q.t < b.t - e) {
            b.t = q.t;
            b.t_err = q.dt;
            b.t_f = float_ge(q.t + q.dt);  // a triangle at or below this may still win: it is checked against t_err
            b.rec = k;
            b.obj = __float_as_int(r0.w);
            b.sub = 0;
          } else {
            ambiguous = true;
          }
        }
        return 1;
      }
      double t = 0;
      const bool h = sphere_test(xyz(r0), rec_radius(r1), r, rg.t_from, false, &t);
      if (best_offer(b, h && t <= b.t, t, k, __float_as_int(r0.w), 0)) b.t_f = float_le(t);
      return 1;
    }
  }
  return 1;  // K_QUAD1 reached on its own never happens (leaves keep pairs together); K_SKY is not in the BVH
}

// The pending parallelogram of Best::amb_* (triangles at t1 < t0, list index q), given the closest hit b of a traversal
// that let it report t0.  HitableList::Hit calls it with T = the closest report of the objects listed before it
// (hitable_list.cu:11-21) and it answers t0 if t0 <= T, else t1 if t1 <= T (parallelogram.cu:25-41):
//   b.t < t1, or b is the pair itself, or b.t >= t0   -> nothing changes (a closer object wins either way / T >= t0);
//   t1 <= b.t < t0 and b is listed BEFORE q            -> T = b.t: the pair answers t1, which wins unless t1 == b.t
//                                                         (HitableList replaces on `<` only);
//   t1 <= b.t < t0 and b is listed AFTER q             -> depends on whether something listed before q lies in
//                                                         [b.t, t0), a one-ulp window: the sequential scan decides.
// The scan used to be the answer to EVERY such ray; on a scene with a mesh it tests every face (bunny: 69 451, 12 ms
// for one lane, its warp waiting) and the handful of rays per frame that cross the backdrop's diagonal left 39 % of the
// SM cycles of the launch idle (profiles/r02_i_bunny_two_phase_ncu_summary.txt).
__device__ __forceinline__ void resolve_quad_order(const DScene &s, Best &b, bool &ambiguous) {
  if (b.amb_rec < 0) return;
  const double t0 = (double)b.amb_t0, t1 = (double)b.amb_t1;
  if (b.t_err > 0 && (fabs(b.t - t0) <= b.t_err || fabs(b.t - t1) <= b.t_err)) {
    ambiguous = true;
    return;
  }
  if (b.t < t1 || b.t >= t0 || b.rec == b.amb_rec - 1) return;
  const int q = __float_as_int(__ldg(s.recs + 3 * (size_t)b.amb_rec).w);
  if (b.rec != REC_NONE && b.obj < q) {
    if (t1 < b.t) {
      b.t = t1;
      b.t_f = b.amb_t1;
      b.rec = b.amb_rec;
      b.obj = q;
      b.sub = 1;
      b.t_err = 0.0;
    }
  } else {
    ambiguous = true;
  }
}

// The winner of a lib_pow traversal, when it is a sphere: its t as the reference computes it (CUDA's pow).
// EXACT_T = false (render kernels): Trace only ever uses (float)t — hit_point = o + (float)t * d, ray_tracing.cu:32,
// sphere.cu:23 — so when the whole interval [t - t_err, t + t_err] rounds to ONE float, that float is (float) of the
// pow value too and pow is not called; it is when the interval straddles a float rounding boundary (~1e-5 of hits).
template <bool SPH, bool EXACT_T = true>
__device__ __forceinline__ void finalize_sphere_hit(const DScene &s, const RayQ &r, const Range &rg, Best &b, bool &ambiguous) {
  if constexpr (SPH) {
    if (b.t_err > 0 && b.rec >= 0) {
      if constexpr (!EXACT_T) {
        if (__double2float_rn(b.t - b.t_err) == __double2float_rn(b.t + b.t_err)) return;
      }
      const float4 r0 = __ldg(s.recs + 3 * (size_t)b.rec), r1 = __ldg(s.recs + 3 * (size_t)b.rec + 1);
      double t = 0;
      if (sphere_test(xyz(r0), rec_radius(r1), r, rg.t_from, true, &t) && t <= rg.t_to)
        b.t = t;
      else
        ambiguous = true;  // the interval said hit, pow says miss: let the sequential scan decide
    }
  }
}

// ------------------------------------------------------------------------------------------------
// HitableList::Hit restated as the reference executes it: a sequential scan in Append order
// (hitable_list.cu:7-25), meshes scanned face by face in post-sort order (the exact-visibility
// form of BVHNode::Hit).  Used by mode 1 of rt_trace_closest and as the resolver of the rare
// order-dependent cases of the BVH traversal.
// ------------------------------------------------------------------------------------------------
static __device__ __noinline__ Hit world_hit_seq(const DScene &s, const RayQ &r, double t_from, double t_to) {
  bool ok = false;
  double T = t_to;
  int best_rec = REC_NONE;
  int cur_obj = -1;
  bool obj_hit = false;
  double obj_t = 0, obj_tto = T;
  int obj_rec = REC_NONE;
  float tf_f = float_ge(t_from);
  const float4 *recs = s.recs_seq;
  auto commit = [&]() {
    if (obj_hit) {
      if (!ok) {
        best_rec = obj_rec;
        T = fmin(T, obj_t);
        ok = true;
      } else if (obj_t < T) {
        T = obj_t;
        best_rec = obj_rec;
      }
    }
  };
  for (int k = 0; k < s.n_recs_seq; k++) {
    float4 r0 = recs[3 * (size_t)k], r1 = recs[3 * (size_t)k + 1], r2 = recs[3 * (size_t)k + 2];
    int obj = __float_as_int(r0.w);
    uint32_t kind = __float_as_uint(r2.w) & KIND_MASK;
    if (obj != cur_obj) {
      commit();
      cur_obj = obj;
      obj_hit = false;
      obj_tto = T;
    }
    if (kind == K_TRI) {
      float t, u, v;
      if (tri_test(xyz(r0), xyz(r1), xyz(r2), r, tf_f, &t, &u, &v) && (double)t <= obj_tto) {
        obj_tto = t;
        obj_t = t;
        obj_rec = k;
        obj_hit = true;
      }
    } else if (kind == K_QUAD0) {
      float t, u, v;
      if (tri_test(xyz(r0), xyz(r1), xyz(r2), r, tf_f, &t, &u, &v) && (double)t <= obj_tto) {
        obj_t = t;
        obj_rec = k;
        obj_hit = true;
      } else {
        float4 s0 = recs[3 * (size_t)k + 3], s1 = recs[3 * (size_t)k + 4], s2 = recs[3 * (size_t)k + 5];
        if (tri_test(xyz(s0), xyz(s1), xyz(s2), r, tf_f, &t, &u, &v) && (double)t <= obj_tto) {
          obj_t = t;
          obj_rec = k + 1;
          obj_hit = true;
        }
      }
      k++;  // the QUAD1 record
    } else if (kind == K_SPHERE) {
      double t;
      if (sphere_test(xyz(r0), rec_radius(r1), r, t_from, s.sphere_pow != 0, &t) && t <= obj_tto) {
        obj_t = t;
        obj_rec = k;
        obj_hit = true;
      }
    } else if (kind == K_SKY) {
      double t = 1e9;
      if (t_from <= t && t <= obj_tto) {
        obj_t = t;
        obj_rec = k;
        obj_hit = true;
      }
    }
  }
  commit();
  Hit h;
  h.t = ok ? T : 0.0;
  h.rec = best_rec;
  h.seq = 1;
  h.box = -1;
  return h;
}

// ------------------------------------------------------------------------------------------------
// BVH traversal (product path).  32-byte nodes, children tested before descent, near child first,
// far child's (a, b) words pushed on a short per-thread stack.  Boxes were padded at build time
// and the cull bound is widened, so that culling never removes a record the scan would accept.
// ------------------------------------------------------------------------------------------------
// 1/d for the slab tests, clamped to +-1e18: with an infinite reciprocal (d = +-0) the FFMA form turns a
// plane on the far side into inf - inf = NaN, and fmaxf(near, NaN) would return the NEAR plane as the exit.
// A finite stand-in keeps every plane distance finite and ordered (|t| >= 1e18 * |plane - o|: still far
// outside any scene unless the origin lies within rounding of the plane, where either answer is accepted
// by the padded box).
// The reciprocal itself is the 1-instruction approximation (2 ulp, denormals flushed to an infinite result that the
// clamp catches): a relative error of 2e-7 on every plane distance of the ray, far inside the 1e-5 guard bands.
__device__ __forceinline__ float slab_rcp(float d) { return fmaxf(fminf(__fdividef(1.0f, d), 1e18f), -1e18f); }

// Slab test of the two children of a binary node: entry / exit distances of the left (ln, lf) and right
// (rn, rf) box, clamped to [t_lo, t_hi].  Culling only has to be conservative, so the plane distances are
// one FFMA each, fma(plane, 1/d, -(o/d)): the rounding of that form is bounded by
// 3e-8 * (t + |plane| / |d|), which the builder's relative box padding (2e-5 * |coordinate|, rt_build.cu
// k_bp_bounds) and the relative guard bands on t (1e-5) cover.  1/d comes from slab_rcp (finite).
__device__ __forceinline__ void slab2(const float4 &l0, const float4 &l1, const float4 &r0, const float4 &r1, float idx,
                                      float idy, float idz, float nox, float noy, float noz, float t_lo, float t_hi,
                                      float &ln, float &lf, float &rn, float &rf) {
  float ax = __fmaf_rn(l0.x, idx, nox), bx = __fmaf_rn(l1.x, idx, nox);
  float ay = __fmaf_rn(l0.y, idy, noy), by = __fmaf_rn(l1.y, idy, noy);
  float az = __fmaf_rn(l0.z, idz, noz), bz = __fmaf_rn(l1.z, idz, noz);
  ln = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), t_lo));
  lf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), t_hi));
  ax = __fmaf_rn(r0.x, idx, nox), bx = __fmaf_rn(r1.x, idx, nox);
  ay = __fmaf_rn(r0.y, idy, noy), by = __fmaf_rn(r1.y, idy, noy);
  az = __fmaf_rn(r0.z, idz, noz), bz = __fmaf_rn(r1.z, idz, noz);
  rn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), t_lo));
  rf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), t_hi));
}

struct TravStats {
  uint32_t nodes, prims;
};

// SM_LEVELS > 0: the first SM_LEVELS entries of the traversal stack live in shared memory (sm = this thread's column of
// a [SM_LEVELS][blockDim.x] array, consecutive lanes in consecutive banks), only deeper entries in local memory.  On
// the soups the per-thread local stack is what fills L1 — 1024 threads x ~20 live entries x 8 B = 160 KB of the SM's
// 256 KB — and its pops were the largest single source of long-scoreboard stalls
// (profiles/r01_j_soup10m_fused_ncu_summary.txt: `stack[--sp]` 325 096 samples against 54 629 for the node loads).
template <bool STATS, bool SPH = true, bool EXACT_T = true, int SM_LEVELS = 0>
__device__ __forceinline__ Hit world_hit_bvh_ifif(const DScene &s, const RayQ &r, double t_from, double t_to,
                                                  TravStats *st, uint2 *sm = nullptr) {
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
    const float t_lo = rg.tf_f - fabsf(rg.tf_f) * 1e-5f;
    uint2 stack[STACK_SIZE - SM_LEVELS];
    int sp = 0;
    uint32_t ca = s.root_a, cb = s.root_b;
    const float4 *__restrict__ nodes = s.nodes;
    for (;;) {
      if (cb & LEAF_FLAG) {
        int k = (int)ca, end = (int)(ca + (cb & ~LEAF_FLAG));
        while (k < end) {
          int used = rec_test<SPH>(s.recs, k, r, rg, b, ambiguous);
          if (STATS) st->prims += used;  // records tested (a parallelogram = 2 triangle records)
          k += used;
        }
      } else {
        if (STATS) st->nodes += 2;
        float4 l0 = __ldg(nodes + 2 * (size_t)ca), l1 = __ldg(nodes + 2 * (size_t)ca + 1);
        float4 r0 = __ldg(nodes + 2 * (size_t)cb), r1 = __ldg(nodes + 2 * (size_t)cb + 1);
        // widened upper bound: never cull a box whose entry is within rounding of the best t
        float t_hi = b.t_f + fabsf(b.t_f) * 1e-5f;
        float ln, lf, rn, rf;
        slab2(l0, l1, r0, r1, idx, idy, idz, nox, noy, noz, t_lo, t_hi, ln, lf, rn, rf);
        // entry <= exit with a relative guard band on the exit (an exit < 0 never passes: entry >= t_lo > 0)
        bool hl = ln <= lf * 1.00001f, hr = rn <= rf * 1.00001f;
        uint32_t la = __float_as_uint(l0.w), lb = __float_as_uint(l1.w);
        uint32_t ra = __float_as_uint(r0.w), rb = __float_as_uint(r1.w);
        if (hl && hr) {
          bool left_first = ln <= rn;
          if (sp < STACK_SIZE) {
            const uint2 e = left_first ? make_uint2(ra, rb) : make_uint2(la, lb);
            if (SM_LEVELS > 0 && sp < SM_LEVELS)
              sm[sp * blockDim.x] = e;
            else
              stack[sp - SM_LEVELS] = e;
            sp++;
          } else {
            overflow = true;
          }
          ca = left_first ? la : ra;
          cb = left_first ? lb : rb;
          continue;
        }
        if (hl) {
          ca = la;
          cb = lb;
          continue;
        }
        if (hr) {
          ca = ra;
          cb = rb;
          continue;
        }
      }
      if (sp == 0) break;
      --sp;
      const uint2 e = (SM_LEVELS > 0 && sp < SM_LEVELS) ? sm[sp * blockDim.x] : stack[sp - SM_LEVELS];
      ca = e.x;
      cb = e.y;
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

// "while-while" form of the same traversal (Aila & Laine 2009): every lane first walks inner nodes
// until it holds a leaf, then the warp tests leaf records together, so node steps and record tests
// are not interleaved across lanes.  (Round 1 also measured speculative leaf postponement and pop-time culling:
// both slower; removed.)
template <bool STATS, bool SPH = true, bool EXACT_T = true>
__device__ __forceinline__ Hit world_hit_bvh_ww(const DScene &s, const RayQ &r, double t_from, double t_to,
                                                TravStats *st) {
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
    const float t_lo = rg.tf_f - fabsf(rg.tf_f) * 1e-5f;
    uint2 stack[STACK_SIZE];
    int sp = 0;
    uint32_t ca = s.root_a, cb = s.root_b;
    const float4 *__restrict__ nodes = s.nodes;
    bool done = false;
    while (!done) {
      uint32_t leaf_a = 0, leaf_b = 0;  // leaf_b == 0: no leaf held
      for (;;) {
        if (cb & LEAF_FLAG) {
          leaf_a = ca;
          leaf_b = cb;
          if (sp == 0) {
            done = true;
          } else {
            uint2 e = stack[--sp];
            ca = e.x;
            cb = e.y;
          }
          break;
        }
        if (STATS) st->nodes += 2;
        float4 l0 = __ldg(nodes + 2 * (size_t)ca), l1 = __ldg(nodes + 2 * (size_t)ca + 1);
        float4 r0 = __ldg(nodes + 2 * (size_t)cb), r1 = __ldg(nodes + 2 * (size_t)cb + 1);
        float t_hi = b.t_f + fabsf(b.t_f) * 1e-5f;
        float ln, lf, rn, rf;
        slab2(l0, l1, r0, r1, idx, idy, idz, nox, noy, noz, t_lo, t_hi, ln, lf, rn, rf);
        bool hl = ln <= lf * 1.00001f, hr = rn <= rf * 1.00001f;
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
        } else if (sp == 0) {
          done = true;
          break;
        } else {
          uint2 e = stack[--sp];
          ca = e.x;
          cb = e.y;
        }
      }
      if (leaf_b) {
        int k = (int)leaf_a, end = (int)(leaf_a + (leaf_b & ~LEAF_FLAG));
        while (k < end) {
          int used = rec_test<SPH>(s.recs, k, r, rg, b, ambiguous);
          if (STATS) st->prims += used;  // records tested (a parallelogram = 2 triangle records)
          k += used;
        }
      }
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

// Binary-tree traversal loops of the render kernels (identical hits): 0 if-if, 1 while-while.
template <bool STATS, int TRAV, bool SPH = true, bool EXACT_T = true, int SM_LEVELS = 0>
__device__ __forceinline__ Hit world_hit_bvh(const DScene &s, const RayQ &r, double t_from, double t_to,
                                             TravStats *st, uint2 *sm = nullptr) {
  if (TRAV == 1) return world_hit_bvh_ww<STATS, SPH, EXACT_T>(s, r, t_from, t_to, st);
  return world_hit_bvh_ifif<STATS, SPH, EXACT_T, SM_LEVELS>(s, r, t_from, t_to, st, sm);
}

// ------------------------------------------------------------------------------------------------
// Surface data of an accepted hit (what HitRecord carries): normal, u, v, material
// ------------------------------------------------------------------------------------------------
struct Surface {
  V3 normal;
  float u, v;  // HitRecord::u,v narrowed to float
  int32_t material, obj, sub;
  uint32_t kind;
};

// sphere.cu:52-64 GetUV
__device__ __forceinline__ void sphere_uv(V3 p, float *u, float *v) {
  float pi = 3.14159265358979323846264338327950288f;
  float theta = acosf(-p.y);
  float phi = atan2f(-p.z, p.x) + pi;
  *u = (float)((double)(phi / (2 * pi)));
  *v = (float)((double)(theta / pi));
}

// need_uv: only textured materials read u, v; callers that do not need them skip the arithmetic.
template <bool SPH = true>
__device__ __forceinline__ Surface surface_at(const DScene &s, const RayQ &r, const Hit &h, bool need_uv) {
  Surface sf;
  sf.normal = mk(0.f, 0.f, 0.f);
  sf.u = sf.v = 0.f;
  sf.sub = 0;
  if (h.rec == REC_SKY) {
    sf.material = s.sky_mat;
    sf.obj = s.sky_obj;
    sf.kind = K_SKY;
    return sf;
  }
  const float4 *recs = h.seq ? s.recs_seq : s.recs;
  float4 r0 = __ldg(recs + 3 * (size_t)h.rec), r1 = __ldg(recs + 3 * (size_t)h.rec + 1),
         r2 = __ldg(recs + 3 * (size_t)h.rec + 2);
  uint32_t w = __float_as_uint(r2.w);
  sf.kind = w & KIND_MASK;
  sf.material = (int32_t)(w >> KIND_BITS);
  sf.obj = __float_as_int(r0.w);
  sf.sub = __float_as_int(r1.w);
  if (sf.kind == K_SKY) return sf;
  if constexpr (SPH) {
    if (sf.kind == K_SPHERE) {
      V3 c = xyz(r0);
      V3 p = r.o + float(h.t) * r.d;
      sf.normal = normalize(p - c);
      if (need_uv) sphere_uv(sf.normal, &sf.u, &sf.v);
      return sf;
    }
  }
  V3 e1 = xyz(r1), e2 = xyz(r2);
  V3 n = normalize(cross(e1, e2));
  sf.normal = dot(r.d, n) < 0 ? n : -n;
  if (need_uv) {
    // re-run the decision arithmetic for u, v (same operations -> same bits)
    float t, u, v;
    tri_test(xyz(r0), e1, e2, r, -CUDART_INF_F, &t, &u, &v);
    float a = (float)(1 - (double)u - (double)v), b = u, c = v;
    if (sf.kind == K_QUAD0) {  // parallelogram.cu:26-29
      sf.u = 0.f * a + 1.f * b + 0.f * c;
      sf.v = 1.f * a + 1.f * b + 0.f * c;
    } else if (sf.kind == K_QUAD1) {  // parallelogram.cu:35-38
      sf.u = 1.f * a + 0.f * b + 1.f * c;
      sf.v = 1.f * a + 0.f * b + 0.f * c;
    } else {
      const float *uvs = h.seq ? s.face_uv_seq : s.face_uv;
      if (w & FLAG_BARY) {  // Triangle::Hit keeps the barycentrics (triangle.cu:11-19)
        sf.u = u;
        sf.v = v;
      } else if (uvs) {  // Face<true>::Hit, bvh.cuh:41-45 (Face<false> leaves u,v unset: reported 0)
        const float *q = uvs + 6 * (size_t)h.rec;
        sf.u = q[0] * a + q[2] * b + q[4] * c;
        sf.v = q[1] * a + q[3] * b + q[5] * c;
      }
    }
  }
  return sf;
}

// ------------------------------------------------------------------------------------------------
// textures and materials
// ------------------------------------------------------------------------------------------------
// image_texture.cu:9-15 over the texture object of :17-38 (normalized coords, point filter, wrap,
// uchar4 -> x/255), restated on raw memory so that the result does not depend on texture hardware.
__device__ __forceinline__ V3 image_value(const DTexture &tx, float u_in, float v_in) {
  double v = 1.0 - (double)v_in;
  float fu = u_in, fv = (float)v;
  float x = fu - floorf(fu), y = fv - floorf(fv);
  int ix = (int)floorf(x * (float)tx.width), iy = (int)floorf(y * (float)tx.height);
  if (ix >= tx.width) ix = tx.width - 1;
  if (iy >= tx.height) iy = tx.height - 1;
  if (ix < 0) ix = 0;
  if (iy < 0) iy = 0;
  uchar4 px = *reinterpret_cast<const uchar4 *>(tx.rgba + (size_t)iy * tx.pitch + (size_t)ix * 4);
  return mk(px.x / 255.0f, px.y / 255.0f, px.z / 255.0f);
}
__device__ __forceinline__ V3 texture_value(const DScene &s, const DMaterial &m, float u, float v) {
  if (m.texture < 0) return xyz(m.color_fuzz);  // constant_texture.cu:11-14
  return image_value(s.textures[m.texture], u, v);
}

struct PathKey {
  uint32_t pixel, sample, k0, k1;
};

// ------------------------------------------------------------------------------------------------
// Sampling.  The random words of a path are a counter-based stream: word w of (pixel, sample) is word w & 3 of
// the Philox block (pixel, sample, w >> 2, 0).  Layout (rtcore.h, "sampling contract"):
//   words 0, 1        pixel jitter                           (ray_tracing.cu:68-71)
//   words 2, 3        lens sample, thin-lens cameras only    (camera.cu:62-65)
//   words B+2k, B+2k+1   the direction sample of the scatter at depth k, B = 4 with a lens, else 2
// so one block serves two bounces (and the camera block serves bounce 0 of a pinhole camera); Metal with fuzz
// takes its radius from word 0 of block (pixel, sample, 0x40000000 + k, 0).
//
// Lambertian::Scatter normalises a rejection-sampled ball point (lambertian.cu:19-31), i.e. it draws a uniform
// DIRECTION; Metal::Scatter uses the ball point itself (metal.cu:27-36).  Both are drawn here without a loop:
// direction = (sqrt(1 - z^2) cos phi, sqrt(1 - z^2) sin phi, z), z = 1 - 2 u0, phi = 2 pi u1 (Archimedes), ball
// point = direction * cbrt(u2).  Same distributions as the reference's rejection loops, one Philox block instead
// of 1.91 on average — and no warp waiting for the longest of its lanes' trip counts (round 1: philox + rejection
// = 17 % of all warp instructions at 8 of 32 lanes).  sin / cos / cbrt are written out in IEEE operations (explicit
// fmaf, correctly rounded division) so that the CPU oracle reproduces them bit for bit; libdevice's would not be.
// ------------------------------------------------------------------------------------------------
struct PhiloxCache {  // words 2, 3 of the block generated last (the second bounce of a block costs nothing)
  uint32_t block, w2, w3;
};
constexpr uint32_t PHILOX_NO_BLOCK = 0xffffffffu;

__device__ __forceinline__ void path_words(const PathKey &pk, uint32_t w, PhiloxCache *pc, uint32_t *a, uint32_t *b) {
  const uint32_t block = w >> 2;
  if ((w & 2u) && pc && pc->block == block) {
    *a = pc->w2;
    *b = pc->w3;
    return;
  }
  uint32_t o[4];
  philox4x32_10(pk.pixel, pk.sample, block, 0u, pk.k0, pk.k1, o);
  if (pc) {
    pc->block = block;
    pc->w2 = o[2];
    pc->w3 = o[3];
  }
  *a = (w & 2u) ? o[2] : o[0];
  *b = (w & 2u) ? o[3] : o[1];
}

// (cos, sin)(2 pi u) * sqrt(1/2), u in (0, 1].  Quadrant k = floor(4u), x = (pi/2)(4u - k - 1/2) in [-pi/4, pi/4):
// Taylor polynomials of degree 9 / 8 (|error| < 3e-8), then the rotation by (k + 1/2) pi/2, whose sine and cosine
// are +-sqrt(1/2) — left to the caller's radius.
__host__ __device__ __forceinline__ void sincos_2pi_scaled(float u, float *c, float *s) {
  const float q = u * 4.0f;
  const float kf = floorf(q);
  const float x = ((q - kf) - 0.5f) * 1.57079637050628662109375f;
  const float x2 = x * x;
  float ps = fmaf(x2, 2.75573192e-6f, -1.98412698e-4f);
  ps = fmaf(x2, ps, 8.33333333e-3f);
  ps = fmaf(x2, ps, -1.66666667e-1f);
  const float sn = fmaf(x * x2, ps, x);
  float pc = fmaf(x2, 2.48015873e-5f, -1.38888889e-3f);
  pc = fmaf(x2, pc, 4.16666667e-2f);
  pc = fmaf(x2, pc, -0.5f);
  const float cs = fmaf(x2, pc, 1.0f);
  const float p = cs - sn, qq = cs + sn;
  const int k = (int)kf;
  const float cv = (k & 1) ? qq : p, sv = (k & 1) ? p : qq;
  *c = ((k + 1) & 2) ? -cv : cv;
  *s = (k & 2) ? -sv : sv;
}

// Uniform direction on the unit sphere from two words.
__host__ __device__ __forceinline__ V3 sphere_direction(float u0, float u1) {
  const float z = 1.0f - 2.0f * u0;
  const float r = sqrtf(fmaxf(1.0f - z * z, 0.0f)) * 0.707106769084930419921875f;
  float c, s;
  sincos_2pi_scaled(u1, &c, &s);
  return mk(r * c, r * s, z);
}

// Cube root of x in (0, 1]: exponent / 3 by integer division, three Newton steps (relative error < 2e-7).
__host__ __device__ __forceinline__ float cbrt_unit(float x) {
  uint32_t i;
#ifdef __CUDA_ARCH__
  i = __float_as_uint(x);
#else
  memcpy(&i, &x, 4);
#endif
  i = i / 3u + 709921077u;
  float y;
#ifdef __CUDA_ARCH__
  y = __uint_as_float(i);
#else
  memcpy(&y, &i, 4);
#endif
  y = y - (y * y * y - x) / (3.0f * (y * y));
  y = y - (y * y * y - x) / (3.0f * (y * y));
  y = y - (y * y * y - x) / (3.0f * (y * y));
  return y;
}

__device__ __forceinline__ bool material_needs_uv(const DMaterial &m) { return m.texture >= 0; }

// Material::Scatter (lambertian.cu:33-43, metal.cu:12-25, dielectric.cu:16-44, diffuse_light.cu:5-9,
// sky.cu:3-7).  Returns false when the path ends here.  The scattered ray goes through Ray::Ray,
// which normalizes (ray.cu:8-11).  word0 = index of the first stream word of this bounce.
__device__ __forceinline__ bool scatter(const DScene &s, const DMaterial &m, const RayQ &ray, float t, const Surface &sf,
                                        const PathKey &pk, int depth, uint32_t word0, PhiloxCache *pc, V3 *att,
                                        RayQ *out) {
  switch (m.type) {
    case RT_MAT_LAMBERTIAN: {
      if (dot(ray.d, sf.normal) >= 0) return false;
      V3 p = ray.o + t * ray.d;
      V3 albedo = texture_value(s, m, sf.u, sf.v);
      uint32_t a, b;
      path_words(pk, word0, pc, &a, &b);
      V3 r = sphere_direction(u01(a), u01(b));
      V3 next_dir = normalize(r + sf.normal);
      *att = albedo;
      out->o = p;
      out->d = normalize(next_dir);
      return true;
    }
    case RT_MAT_METAL: {
      if (dot(ray.d, sf.normal) >= 0) return false;
      V3 p = ray.o + t * ray.d;
      *att = xyz(m.color_fuzz);
      V3 reflected = reflect(ray.d, sf.normal);
      float fuzz = m.color_fuzz.w;
      out->o = p;
      if (fuzz > 0) {
        uint32_t a, b, o[4];
        path_words(pk, word0, pc, &a, &b);
        philox4x32_10(pk.pixel, pk.sample, 0x40000000u + (uint32_t)depth, 0u, pk.k0, pk.k1, o);
        V3 r = sphere_direction(u01(a), u01(b)) * cbrt_unit(u01(o[0]));
        out->d = normalize(reflected + fuzz * r);
      } else {
        out->d = normalize(reflected);
      }
      return true;
    }
    case RT_MAT_DIELECTRIC: {
      V3 p = ray.o + t * ray.d;
      V3 d;
      if (dot(ray.d, sf.normal) >= 0)
        d = refract(ray.d, -sf.normal, m.eta_in);
      else
        d = refract(ray.d, sf.normal, m.eta_out);
      if (d.x == 0 && d.y == 0 && d.z == 0) return false;
      if (isnan(d.x) || isnan(d.y) || isnan(d.z)) return false;
      *att = xyz(m.color_fuzz);
      out->o = p;
      out->d = normalize(d);
      return true;
    }
    default:
      return false;
  }
}

// ------------------------------------------------------------------------------------------------
// Fixed-point radiance sums.  A sample's radiance is rounded once to a multiple of 2^-30 and the samples of a pixel
// are summed as 64-bit integers: exact, hence independent of the order in which warps, sample chunks, rt_render
// calls or GPUs contribute them — any schedule and any partition gives the same bits.  |component| is clamped
// to 2^32 (a sum holds 2^31 such samples), NaN counts as 0.
// ------------------------------------------------------------------------------------------------
constexpr float FIXED_SCALE = 1073741824.0f, FIXED_CLAMP = 4294967296.0f;
constexpr double FIXED_INV = 1.0 / 1073741824.0;
__device__ __forceinline__ long long to_fixed(float c) {
  if (c != c) c = 0.0f;
  c = c > FIXED_CLAMP ? FIXED_CLAMP : c;
  c = c < -FIXED_CLAMP ? -FIXED_CLAMP : c;
  return __float2ll_rn(c * FIXED_SCALE);
}
__device__ __forceinline__ float from_fixed(long long q) { return __double2float_rn(__ll2double_rn(q) * FIXED_INV); }

// Material::Emit (diffuse_light.cu:11-13, sky.cu:9-14, material.cu:3-6)
__device__ __forceinline__ V3 emit(const DScene &s, const DMaterial &m, float u, float v, V3 p) {
  switch (m.type) {
    case RT_MAT_DIFFUSE_LIGHT:
      return texture_value(s, m, u, v);
    case RT_MAT_SKY: {
      V3 dir = normalize(p);
      float t = (float)(0.5 * ((double)dir.y + 1.0));
      return (1.0f - t) * mk(1.0f, 1.0f, 1.0f) + t * mk(0.5f, 0.7f, 1.0f);
    }
    default:
      return mk(0.f, 0.f, 0.f);
  }
}

// camera.cu:57-77 RayAt; x, y in [-1, 1].  (DiskRand samples the square [0, lens_radius]^2.)
__device__ __forceinline__ RayQ camera_ray_at(const DCamera &c, double x, double y, float lens_u, float lens_v) {
  x = (x + 1) / 2;
  y = (y + 1) / 2;
  V3 target = c.llc + static_cast<float>(x) * c.horizontal + static_cast<float>(y) * c.vertical;
  V3 origin;
  if (c.is_defocus) {
    origin = c.position + c.u * lens_u + c.v * lens_v;
  } else {
    origin = c.position;
  }
  RayQ r;
  r.o = origin;
  r.d = normalize(normalize(target - origin));
  return r;
}

// ray_tracing.cu:68-74: camera sample `sample` of pixel (i, j) from stream words 0..3 (block 0)
__device__ __forceinline__ RayQ camera_sample(const DCamera &cam, int i, int j, int height, int width,
                                              const PathKey &pk, PhiloxCache *pc) {
  uint32_t o[4];
  philox4x32_10(pk.pixel, pk.sample, 0u, 0u, pk.k0, pk.k1, o);
  if (pc) {
    pc->block = 0u;
    pc->w2 = o[2];
    pc->w3 = o[3];
  }
  double x = ((double)u01(o[0]) + double(j)) / double(width);
  double y = ((double)u01(o[1]) + double(height - i)) / double(height);
  x = 2 * x - 1;
  y = 2 * y - 1;
  float lu = 0, lv = 0;
  if (cam.is_defocus == 2) {
    // RT_CAMERA_DISK: the lens as a uniform DISK of radius lens_radius (r = R sqrt(u), phi = 2 pi v).  The
    // reference's DiskRand (camera.cu:74-77) draws two uniforms in (0, lens_radius]: a square in the positive quadrant
    float c, s;
    sincos_2pi_scaled(u01(o[3]), &c, &s);
    const float rad = sqrtf(u01(o[2])) * cam.lens_radius * 1.41421354f;
    lu = rad * c;
    lv = rad * s;
  } else if (cam.is_defocus) {
    lu = u01(o[2]) * cam.lens_radius;
    lv = u01(o[3]) * cam.lens_radius;
  }
  return camera_ray_at(cam, x, y, lu, lv);
}
// first stream word of the scatter at `depth`
__device__ __forceinline__ uint32_t scatter_word(const DCamera &cam, int depth) {
  return (cam.is_defocus ? 4u : 2u) + 2u * (uint32_t)depth;
}

}  // namespace rtc
