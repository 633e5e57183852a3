// rt_scene_api.cuh — the reference's C++ scene-construction API, re-implemented as plain tagged records.
//
// Host-side mirror of the interface the reference's scenes/*.cu are written against (class names,
// constructor signatures, host/device qualifiers, Append/list_len, Face<> memory layout, Main /
// DistributedMain): those files compile UNCHANGED against this directory (every header name of
// ray-tracing-cuda/ exists here and forwards to this file).  What is different by design:
//   * no virtual functions and no device-side Hit/Scatter/Emit: an object is a small record whose
//     first word is a type tag.  After init_world returns, Main()/DistributedMain() walk the list
//     once on the device (rt_scene_api.cu: FlattenWorld), hand the flat description to the C-ABI
//     (rt_scene_commit, include/rtcore.h) and the render runs in librtcore.so;
//   * BVH<Face<B>, AABB>'s constructor only registers the mesh (the reference builds the tree in
//     one device thread with a sequential sort, bvh.cuh:113-121); the LBVH is built in parallel by
//     rt_scene_commit;
//   * Parallelepiped stores its six Parallelograms inline instead of an 8 KB HitableList.
// Reference interface replaced, file by file: hitable.cuh:13-23, hitable_list.cuh:8-22, ray.cuh:5-14,
// camera.cuh:10-35, sphere.cuh:8-25, triangle.cuh:7-17, parallelogram.cuh:7-17,
// parallelepiped.cuh:16-29, sky.cuh:8-26, bvh.cuh:10-183, material.cuh:13-20, lambertian.cuh:11-29,
// metal.cuh:5-19, dielectric.cuh:5-17, diffuse_light.cuh:6-18, textures/*.cuh, cuda_copyable.cuh,
// utils.cuh:15-57, ray_tracing.cuh:9-21.
#pragma once

#include <cuda_runtime.h>
#include <curand_kernel.h>
#include <glog/logging.h>
#include <mpi.h>
#include <stddef.h>
#include <stdint.h>

#include <functional>
#include <glm/glm.hpp>
#include <nvfunctional>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/rtworld.h"

// ---- type tags: the values of include/rtworld.h (the records rt_scene_from_world walks) ---------------
namespace rtapi {
enum HitableKind : int32_t {
  kHitNone = RT_WORLD_HIT_NONE,
  kHitList = RT_WORLD_HIT_LIST,
  kHitSphere = RT_WORLD_HIT_SPHERE,
  kHitTriangle = RT_WORLD_HIT_TRIANGLE,
  kHitParallelogram = RT_WORLD_HIT_PARALLELOGRAM,
  kHitParallelepiped = RT_WORLD_HIT_PARALLELEPIPED,
  kHitSky = RT_WORLD_HIT_SKY,
  kHitMesh = RT_WORLD_HIT_MESH,      // BVH<Face<false>, AABB>
  kHitMeshUV = RT_WORLD_HIT_MESH_UV, // BVH<Face<true>, AABB>
  kHitAABB = RT_WORLD_HIT_AABB
};
enum MaterialKind : int32_t {
  kMatNone = RT_WORLD_MAT_NONE, kMatLambertian = RT_WORLD_MAT_LAMBERTIAN, kMatMetal = RT_WORLD_MAT_METAL,
  kMatDielectric = RT_WORLD_MAT_DIELECTRIC, kMatDiffuseLight = RT_WORLD_MAT_DIFFUSE_LIGHT, kMatSky = RT_WORLD_MAT_SKY
};
enum TextureKind : int32_t { kTexNone = RT_WORLD_TEX_NONE, kTexConstant = RT_WORLD_TEX_CONSTANT, kTexImage = RT_WORLD_TEX_IMAGE };
// layout checks against the records of rtworld.h (member offsets through a friend of every class)
struct LayoutCheck;
}  // namespace rtapi

// ---- cuda_copyable.cuh:6-19 ----------------------------------------------------------------------
class CudaCopyable {
 public:
  template <typename T>
  void ToDevice(T *symbol) {
    cudaError_t err = cudaMemcpy(symbol, this, sizeof(T), cudaMemcpyHostToDevice);
    CHECK(err == cudaSuccess) << cudaGetErrorString(err);
  }
  template <typename T>
  void FromDevice(T *symbol) {
    cudaError_t err = cudaMemcpy(this, symbol, sizeof(T), cudaMemcpyDeviceToHost);
    CHECK(err == cudaSuccess) << cudaGetErrorString(err);
  }
};

// ---- ray.cuh:5-14 (the constructor normalizes, ray.cu:8-11) ---------------------------------------
class Ray {
 public:
  __device__ __host__ Ray() {}
  __device__ __host__ Ray(glm::vec3 position, glm::vec3 direction)
      : position_(position), direction_(glm::normalize(direction)) {}
  __device__ __host__ const glm::vec3 position() const { return position_; }
  __device__ __host__ const glm::vec3 direction() const { return direction_; }

 private:
  glm::vec3 position_, direction_;
};

class Material;

// ---- hitable.cuh:13-17 ---------------------------------------------------------------------------
struct HitRecord {
  double t, u, v;
  glm::vec3 normal;
  Material *material_ptr;
};

// ---- textures/texture.cuh, constant_texture.cuh, image_texture.cuh --------------------------------
class Texture : public CudaCopyable {
  friend struct rtapi::LayoutCheck;
 public:
  __host__ __device__ int32_t rt_kind() const { return rt_kind_; }

 protected:
  int32_t rt_kind_ = rtapi::kTexNone;
};

class ConstantTexture : public Texture {
  friend struct rtapi::LayoutCheck;
 public:
  __host__ __device__ ConstantTexture() : color_(0.f) { rt_kind_ = rtapi::kTexConstant; }
  __host__ __device__ ConstantTexture(glm::vec3 color) : color_(color) { rt_kind_ = rtapi::kTexConstant; }
  __host__ __device__ glm::vec3 Value(double, double, const glm::vec3 &) const { return color_; }
  __host__ __device__ glm::vec3 color() const { return color_; }

 private:
  glm::vec3 color_;
};

class ImageTexture : public Texture {
  friend struct rtapi::LayoutCheck;
 public:
  __host__ __device__ ImageTexture(cudaTextureObject_t image_texture) : image_texture_(image_texture) {
    rt_kind_ = rtapi::kTexImage;
  }
  __host__ __device__ cudaTextureObject_t image_texture() const { return image_texture_; }
  // image_texture.cu:17-38: pitch2D uchar4, normalized float reads, normalized coords, point filter, wrap
  static cudaTextureObject_t CreateCudaTextureObj(uint8_t *dev_buffer, int height, int width, uint64_t pitch_in_bytes);

 private:
  cudaTextureObject_t image_texture_;
};

// ---- material.cuh:13-20 and the five materials ----------------------------------------------------
class Material : public CudaCopyable {
  friend struct rtapi::LayoutCheck;
 public:
  __host__ __device__ int32_t rt_kind() const { return rt_kind_; }

 protected:
  int32_t rt_kind_ = rtapi::kMatNone;
};

class Lambertian : public Material {
  friend struct rtapi::LayoutCheck;
 public:
  __host__ __device__ Lambertian(glm::vec3 color) : texture_ptr_(nullptr), color_(color), use_constant_tex_(true) {
    rt_kind_ = rtapi::kMatLambertian;
  }
  __host__ __device__ Lambertian(Texture *texture_ptr) : texture_ptr_(texture_ptr), use_constant_tex_(false) {
    rt_kind_ = rtapi::kMatLambertian;
  }
  __device__ __host__ __inline__ Texture *texture_ptr() { return use_constant_tex_ ? &color_ : texture_ptr_; }

 private:
  Texture *texture_ptr_;
  ConstantTexture color_;
  bool use_constant_tex_;
};

class Metal : public Material {
  friend struct rtapi::LayoutCheck;  // metal.cuh:8-9: constructors are device-only; fuzz clamped to <= 1 (metal.cu:10)
 public:
  Metal() = delete;
  __device__ Metal(glm::vec3 albedo) : albedo_(albedo), fuzz_(0.f) { rt_kind_ = rtapi::kMatMetal; }
  __device__ Metal(glm::vec3 albedo, float fuzz) : albedo_(albedo), fuzz_(fuzz < 1.f ? fuzz : 1.f) {
    rt_kind_ = rtapi::kMatMetal;
  }
  __host__ __device__ glm::vec3 albedo() const { return albedo_; }
  __host__ __device__ float fuzz() const { return fuzz_; }

 private:
  glm::vec3 albedo_;
  float fuzz_;
};

class Dielectric : public Material {
  friend struct rtapi::LayoutCheck;
 public:
  Dielectric() = delete;
  __device__ __host__ Dielectric(glm::vec3 attenuation, double refractive_index)
      : attenuation_(attenuation), refractive_index_(refractive_index) {
    rt_kind_ = rtapi::kMatDielectric;
  }
  __host__ __device__ glm::vec3 attenuation() const { return attenuation_; }
  __host__ __device__ double refractive_index() const { return refractive_index_; }

 private:
  glm::vec3 attenuation_;
  double refractive_index_;
};

class DiffuseLight : public Material {
  friend struct rtapi::LayoutCheck;
 public:
  DiffuseLight() = delete;
  __host__ __device__ DiffuseLight(Texture *texture_ptr) : texture_ptr_(texture_ptr) {
    rt_kind_ = rtapi::kMatDiffuseLight;
  }
  __host__ __device__ Texture *texture_ptr() const { return texture_ptr_; }

 private:
  Texture *texture_ptr_;
};

class SkyMaterial : public Material {
 public:
  __host__ __device__ SkyMaterial() { rt_kind_ = rtapi::kMatSky; }
};

// ---- hitable.cuh:19-23, hitable_list.cuh:8-22 -----------------------------------------------------
class Hitable {
  friend struct rtapi::LayoutCheck;
 public:
  __host__ __device__ int32_t rt_kind() const { return rt_kind_; }

 protected:
  int32_t rt_kind_ = rtapi::kHitNone;
};

class HitableList : public Hitable {
  friend struct rtapi::LayoutCheck;
 public:
  constexpr static int kMaxHitables = 1024;
  __host__ __device__ HitableList() { rt_kind_ = rtapi::kHitList; }
  __host__ __device__ void Append(Hitable *obj) {  // hitable_list.cu:27-29 (no bounds check there either)
    list_[list_len_++] = obj;
  }
  __inline__ __host__ __device__ int list_len() const { return list_len_; }
  __host__ __device__ Hitable *at(int i) const { return list_[i]; }

 private:
  Hitable *list_[kMaxHitables];
  int list_len_ = 0;
};

// ---- sphere.cuh:8-25 -----------------------------------------------------------------------------
class Sphere : public Hitable {
  friend struct rtapi::LayoutCheck;
 public:
  Sphere() = delete;
  __host__ __device__ Sphere(glm::vec3 position, double radius, Material *material_ptr)
      : radius_(radius), position_(position), material_ptr_(material_ptr) {
    rt_kind_ = rtapi::kHitSphere;
  }
  __device__ __host__ double radius() const { return radius_; }
  __device__ __host__ glm::vec3 position() const { return position_; }
  __device__ __host__ Material *material_ptr() const { return material_ptr_; }

 private:
  double radius_;
  glm::vec3 position_;
  Material *material_ptr_;
};

// ---- triangle.cuh:7-17 ---------------------------------------------------------------------------
class Triangle : public Hitable {
  friend struct rtapi::LayoutCheck;
 public:
  Triangle() = delete;
  __host__ __device__ Triangle(glm::vec3 p[], Material *material_ptr) : material_ptr_(material_ptr) {
    rt_kind_ = rtapi::kHitTriangle;
    for (int i = 0; i < 3; i++) p_[i] = p[i];
  }
  __host__ __device__ glm::vec3 point(int i) const { return p_[i]; }
  __host__ __device__ Material *material_ptr() const { return material_ptr_; }

 private:
  glm::vec3 p_[3];
  Material *material_ptr_;
};

// ---- parallelogram.cuh:7-17; p[3] = p[1] + p[2] - p[0] (parallelogram.cu:10-15) ---------------------
class Parallelogram : public Hitable {
  friend struct rtapi::LayoutCheck;
 public:
  __host__ __device__ Parallelogram() : material_ptr_(nullptr) { rt_kind_ = rtapi::kHitParallelogram; }
  __host__ __device__ Parallelogram(glm::vec3 p[3], Material *material_ptr) : material_ptr_(material_ptr) {
    rt_kind_ = rtapi::kHitParallelogram;
    for (int i = 0; i < 3; i++) p_[i] = p[i];
    p_[3] = p[1] + p[2] - p[0];
  }
  __host__ __device__ glm::vec3 point(int i) const { return p_[i]; }
  __host__ __device__ Material *material_ptr() const { return material_ptr_; }

 private:
  glm::vec3 p_[4];
  Material *material_ptr_;
};

// ---- parallelepiped.cuh:16-29: six parallelograms in AddCorner order (parallelepiped.cu:25-32) ------
class Parallelepiped : public Hitable {
  friend struct rtapi::LayoutCheck;
 public:
  Parallelepiped() = delete;
  __host__ __device__ Parallelepiped(glm::vec3 p[4], Material *material_ptr) {  // parallelepiped.cu:8-18
    rt_kind_ = rtapi::kHitParallelepiped;
    glm::vec3 q[4];
    q[3] = p[2] + p[1] - p[0];
    q[2] = p[3] + p[1] - p[0];
    q[1] = p[3] + p[2] - p[0];
    q[0] = q[3] + q[2] - p[1];
    AddCorner(p, material_ptr);
    AddCorner(q, material_ptr);
  }
  __host__ __device__ Parallelepiped(glm::vec3 lengths, Material *material_ptr,
                                     nvstd::function<glm::vec3(glm::vec3)> transform) {  // parallelepiped.cu:34-55
    rt_kind_ = rtapi::kHitParallelepiped;
    glm::vec3 p[4], q[4];
    for (int i = 0; i < 4; i++) {
      p[i] = glm::vec3(0);
      q[i] = lengths;
      if (i > 0) {
        p[i][i - 1] = lengths[i - 1];
        q[i][i - 1] = 0;
      }
    }
    for (int i = 0; i < 4; i++) {
      p[i] = transform(p[i]);
      q[i] = transform(q[i]);
    }
    AddCorner(p, material_ptr);
    AddCorner(q, material_ptr);
  }
  __host__ __device__ int face_count() const { return n_; }
  __host__ __device__ const Parallelogram &face(int i) const { return faces_[i]; }

 private:
  Parallelogram faces_[6];
  int n_ = 0;
  __device__ __host__ void AddCorner(glm::vec3 p[4], Material *material_ptr) {
    for (int i = 1; i <= 3; i++) {
      int y = i == 3 ? 1 : i + 1;
      glm::vec3 arr[3] = {p[0], p[i], p[y]};
      faces_[n_++] = Parallelogram(arr, material_ptr);
    }
  }
};

// ---- sky.cuh:8-26 ----------------------------------------------------------------------------------
class Sky : public Hitable {
  friend struct rtapi::LayoutCheck;
 public:
  __host__ __device__ Sky() { rt_kind_ = rtapi::kHitSky; }
  __device__ __host__ Material *material_ptr() { return &material_; }

 private:
  SkyMaterial material_;
};

// ---- bvh.cuh:10-99: AABB, FaceBase<>, Face<> (the 36 / 60 byte layouts are API: hosts memcpy them) ---
struct AABB : public Hitable {
 public:
  __host__ __device__ AABB() { rt_kind_ = rtapi::kHitAABB; }
  glm::vec3 min, max;
};

template <bool HasTexCoord>
class FaceBase {
 protected:
  glm::vec3 positions_[3];
  glm::vec2 tex_coords_[3];

 public:
  __device__ __host__ inline glm::vec3 &position(int i) { return positions_[i]; }
  __device__ __host__ inline glm::vec2 &tex_coord(int i) { return tex_coords_[i]; }
};
template <>
class FaceBase<false> {
 public:
  glm::vec3 positions_[3];
  __device__ __host__ inline glm::vec3 &position(int i) { return positions_[i]; }
};
template <bool HasTexCoord>
class Face : public FaceBase<HasTexCoord> {
 public:
  __host__ __device__ int operator<(const Face<HasTexCoord> &a) const {  // bvh.cuh:96-98
    return this->positions_[0].x < a.positions_[0].x;
  }
};
static_assert(sizeof(Face<false>) == 36 && sizeof(Face<true>) == 60, "Face<> layout is part of the API (bunny.cu:67-71)");

// ---- bvh.cuh:161-183: BVH<T, BV> — registers the mesh; one material per mesh ------------------------
template <typename T, typename BV>
class BVH : public Hitable {
  friend struct rtapi::LayoutCheck;
 public:
  __host__ __device__ explicit BVH(T *objs, int n, Material *material_ptr)
      : objs_(objs), n_(n), material_ptr_(material_ptr) {
    rt_kind_ = sizeof(T) == sizeof(Face<true>) ? rtapi::kHitMeshUV : rtapi::kHitMesh;
  }
  __host__ __device__ T *objs() const { return objs_; }
  __host__ __device__ int size() const { return n_; }
  __host__ __device__ Material *material_ptr() const { return material_ptr_; }

 private:
  T *objs_;
  int n_;
  Material *material_ptr_;
};

// ---- camera.cuh:10-35; constructors in the reference's expression order (camera.cu:6-47) ------------
class Camera : public CudaCopyable {
  friend struct rtapi::LayoutCheck;
 private:
  glm::vec3 position_, lower_left_corner_, horizontal_, vertical_, u_, v_, w_;
  bool is_defocus_camera_;
  double lens_radius_ = -1;
  __host__ __device__ void LookAt(glm::vec3 position, glm::vec3 look_at, glm::vec3 up, double half_height,
                                  double width_height_aspect) {
    position_ = position;
    w_ = glm::normalize(position - look_at);
    u_ = glm::normalize(glm::cross(up, w_));
    v_ = glm::normalize(glm::cross(w_, u_));
    double half_width = width_height_aspect * half_height;
    horizontal_ = u_ * static_cast<float>(2 * half_width);
    vertical_ = v_ * static_cast<float>(2 * half_height);
    lower_left_corner_ = position - w_ - u_ * static_cast<float>(half_width) - v_ * static_cast<float>(half_height);
  }

 public:
  Camera() = delete;
  __host__ __device__ Camera(glm::vec3 position, glm::vec3 look_at, glm::vec3 up, double field_of_view,
                             double width_height_aspect) {
    is_defocus_camera_ = false;
    LookAt(position, look_at, up, tan(field_of_view / 2), width_height_aspect);
  }
  __host__ __device__ Camera(glm::vec3 position, glm::vec3 look_at, glm::vec3 up, double field_of_view,
                             double width_height_aspect, double aperture, double focus_distance) {
    is_defocus_camera_ = true;
    LookAt(position, look_at, up, focus_distance * tan(field_of_view / 2), width_height_aspect);
    lens_radius_ = aperture / 2;
  }
  __host__ __device__ Camera(glm::vec3 position, glm::vec3 lower_left_corner, glm::vec3 horizontal,
                             glm::vec3 vertical)
      : position_(position), lower_left_corner_(lower_left_corner), horizontal_(horizontal), vertical_(vertical),
        u_(0.f), v_(0.f), w_(0.f), is_defocus_camera_(false) {}
  __host__ __device__ glm::vec3 position() const { return position_; }
  __host__ __device__ glm::vec3 lower_left_corner() const { return lower_left_corner_; }
  __host__ __device__ glm::vec3 horizontal() const { return horizontal_; }
  __host__ __device__ glm::vec3 vertical() const { return vertical_; }
  __host__ __device__ glm::vec3 u() const { return u_; }
  __host__ __device__ glm::vec3 v() const { return v_; }
  __host__ __device__ glm::vec3 w() const { return w_; }
  __host__ __device__ bool is_defocus_camera() const { return is_defocus_camera_; }
  __host__ __device__ double lens_radius() const { return lens_radius_; }
};

// ---- the classes above ARE the records of include/rtworld.h -------------------------------------------
#pragma nv_diag_suppress 1427  // offsetof on classes with bases: conditionally supported, exactly what is being checked
namespace rtapi {
struct LayoutCheck {
#define RTAPI_SAME(CLS, MEMBER, REC, FIELD) \
  static_assert(offsetof(CLS, MEMBER) == offsetof(REC, FIELD), #CLS "::" #MEMBER " is not at " #REC "::" #FIELD)
#define RTAPI_SIZE(CLS, REC) static_assert(sizeof(CLS) == sizeof(REC), #CLS " and " #REC " differ in size")
  RTAPI_SIZE(ConstantTexture, rt_world_constant_texture);
  RTAPI_SAME(ConstantTexture, rt_kind_, rt_world_constant_texture, kind);
  RTAPI_SAME(ConstantTexture, color_, rt_world_constant_texture, color);
  RTAPI_SIZE(ImageTexture, rt_world_image_texture);
  RTAPI_SAME(ImageTexture, image_texture_, rt_world_image_texture, texture_object);
  RTAPI_SIZE(Lambertian, rt_world_lambertian);
  RTAPI_SAME(Lambertian, rt_kind_, rt_world_lambertian, kind);
  RTAPI_SAME(Lambertian, texture_ptr_, rt_world_lambertian, texture);
  RTAPI_SAME(Lambertian, color_, rt_world_lambertian, color);
  RTAPI_SAME(Lambertian, use_constant_tex_, rt_world_lambertian, use_constant_texture);
  RTAPI_SIZE(Metal, rt_world_metal);
  RTAPI_SAME(Metal, albedo_, rt_world_metal, albedo);
  RTAPI_SAME(Metal, fuzz_, rt_world_metal, fuzz);
  RTAPI_SIZE(Dielectric, rt_world_dielectric);
  RTAPI_SAME(Dielectric, attenuation_, rt_world_dielectric, attenuation);
  RTAPI_SAME(Dielectric, refractive_index_, rt_world_dielectric, refractive_index);
  RTAPI_SIZE(DiffuseLight, rt_world_diffuse_light);
  RTAPI_SAME(DiffuseLight, texture_ptr_, rt_world_diffuse_light, texture);
  RTAPI_SIZE(SkyMaterial, rt_world_sky_material);
  RTAPI_SIZE(HitableList, rt_world_list);
  RTAPI_SAME(HitableList, rt_kind_, rt_world_list, kind);
  RTAPI_SAME(HitableList, list_, rt_world_list, list);
  RTAPI_SAME(HitableList, list_len_, rt_world_list, len);
  static_assert(HitableList::kMaxHitables == RT_WORLD_MAX_HITABLES, "kMaxHitables");
  RTAPI_SIZE(Sphere, rt_world_sphere);
  RTAPI_SAME(Sphere, radius_, rt_world_sphere, radius);
  RTAPI_SAME(Sphere, position_, rt_world_sphere, position);
  RTAPI_SAME(Sphere, material_ptr_, rt_world_sphere, material);
  RTAPI_SIZE(Triangle, rt_world_triangle);
  RTAPI_SAME(Triangle, p_, rt_world_triangle, p);
  RTAPI_SAME(Triangle, material_ptr_, rt_world_triangle, material);
  RTAPI_SIZE(Parallelogram, rt_world_parallelogram);
  RTAPI_SAME(Parallelogram, p_, rt_world_parallelogram, p);
  RTAPI_SAME(Parallelogram, material_ptr_, rt_world_parallelogram, material);
  RTAPI_SIZE(Parallelepiped, rt_world_parallelepiped);
  RTAPI_SAME(Parallelepiped, faces_, rt_world_parallelepiped, faces);
  RTAPI_SAME(Parallelepiped, n_, rt_world_parallelepiped, n_faces);
  RTAPI_SIZE(Sky, rt_world_sky);
  RTAPI_SAME(Sky, material_, rt_world_sky, material);
  typedef BVH<Face<false>, AABB> MeshBVH;
  RTAPI_SIZE(MeshBVH, rt_world_mesh);
  RTAPI_SAME(MeshBVH, objs_, rt_world_mesh, faces);
  RTAPI_SAME(MeshBVH, n_, rt_world_mesh, n_faces);
  RTAPI_SAME(MeshBVH, material_ptr_, rt_world_mesh, material);
  RTAPI_SIZE(Camera, rt_world_camera);
  RTAPI_SAME(Camera, position_, rt_world_camera, position);
  RTAPI_SAME(Camera, w_, rt_world_camera, w);
  RTAPI_SAME(Camera, is_defocus_camera_, rt_world_camera, is_defocus);
  RTAPI_SAME(Camera, lens_radius_, rt_world_camera, lens_radius);
#undef RTAPI_SAME
#undef RTAPI_SIZE
};
}  // namespace rtapi

// ---- ray_tracing.cuh:9-15 ----------------------------------------------------------------------------
struct Layer {
  glm::vec3 pos;
  double t;
  glm::vec3 target;
  glm::vec3 emitted;
  glm::vec3 attenuation;
};

// ---- utils.cuh:15-57 ---------------------------------------------------------------------------------
std::string BaseName(const std::string &path);
std::string ParentPath(const std::string &path);

__inline__ __device__ float CudaRandomFloat(float min, float max, curandState *state) {  // (min, max]
  float t = curand_uniform(state);
  return t * (max - min) + min;
}
__global__ void CudaRandomInit(uint64_t seed, curandState *state, int n);

void WriteImage(const std::vector<glm::vec3> &pixels, int height, int width, const std::string &path);

__host__ __device__ inline int GetWorkload(int rank, int world_size, int spp) {  // utils.cu:111-113
  return spp / world_size + (rank < spp % world_size ? 1 : 0);
}

__host__ void GatherImageData(int height, int width, const std::vector<glm::vec3> &send_buf, int spp,
                              std::vector<glm::vec3> *out_image);

// Environment overrides read by both drivers (the reference hard-codes all of these):
//   RT_WIDTH RT_HEIGHT RT_SPP RT_MAX_DEPTH (default 10 = TRACE_DEPTH_LIMIT) RT_SEED RT_OUT RT_STATS
__host__ void Main(curandState **d_states, Camera **d_camera, HitableList **d_world, glm::vec3 **d_image,
                   nvstd::function<void(HitableList *world, Camera *camera)> init_world, int height, int width,
                   int spp);
__host__ void DistributedMain(curandState **d_states, Camera **d_camera, HitableList **d_world, glm::vec3 **d_image,
                              nvstd::function<void(HitableList *world, Camera *camera)> init_world, int height,
                              int width, int spp);
