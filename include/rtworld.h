/*
 * rtworld.h — device-object ABI of the scene-construction API, for rt_scene_from_world (rtcore.h).
 *
 * The reference builds its scene ON THE DEVICE: a user kernel placement-news a HitableList and a Camera into two
 * cudaMalloc'ed slots and `new`s Sphere / Parallelogram / Lambertian / ... objects that it Appends
 * (scenes/cornell_box.cu:35-69).  The header set of ray-tracing-cuda_b200/scene_api/ keeps those class names and
 * constructors but lays every object out as one of the plain records below (first word = a type tag; no vptr, no
 * virtual Hit/Scatter/Emit).  rt_scene_from_world walks such a list in device memory, flattens it and commits it.
 * The scene API static_asserts that its classes have exactly these sizes and member offsets (rt_scene_api.cuh), so
 * any other producer of the same records (a driver of one's own, another language binding) can use the entry point
 * without the C++ headers.
 *
 * Pointers are device pointers.  vec3 = three floats.  Replaces, file by file: hitable_list.cuh:8-22, sphere.cuh:8-25,
 * triangle.cuh:7-17, parallelogram.cuh:7-17, parallelepiped.cuh:16-29, sky.cuh:8-26, bvh.cuh:161-183,
 * material.cuh:13-20, lambertian.cuh:11-29, metal.cuh:5-19, dielectric.cuh:5-17, diffuse_light.cuh:6-18,
 * textures/*.cuh, camera.cuh:10-35.
 */
#ifndef RTWORLD_H_
#define RTWORLD_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { /* rt_world_*::kind of a Hitable */
  RT_WORLD_HIT_NONE = 0,
  RT_WORLD_HIT_LIST = 1,
  RT_WORLD_HIT_SPHERE = 2,
  RT_WORLD_HIT_TRIANGLE = 3,
  RT_WORLD_HIT_PARALLELOGRAM = 4,
  RT_WORLD_HIT_PARALLELEPIPED = 5,
  RT_WORLD_HIT_SKY = 6,
  RT_WORLD_HIT_MESH = 7,    /* BVH<Face<false>, AABB>: 36-byte faces */
  RT_WORLD_HIT_MESH_UV = 8, /* BVH<Face<true>, AABB>: 60-byte faces */
  RT_WORLD_HIT_AABB = 9     /* not renderable */
};
enum { RT_WORLD_MAT_NONE = 0, RT_WORLD_MAT_LAMBERTIAN = 1, RT_WORLD_MAT_METAL = 2, RT_WORLD_MAT_DIELECTRIC = 3,
       RT_WORLD_MAT_DIFFUSE_LIGHT = 4, RT_WORLD_MAT_SKY = 5 };
enum { RT_WORLD_TEX_NONE = 0, RT_WORLD_TEX_CONSTANT = 1, RT_WORLD_TEX_IMAGE = 2 };

#define RT_WORLD_MAX_HITABLES 1024 /* HitableList::kMaxHitables, hitable_list.cuh:10 */

/* ---- textures ---- */
typedef struct rt_world_constant_texture {
  int32_t kind; /* RT_WORLD_TEX_CONSTANT */
  float color[3];
} rt_world_constant_texture; /* 16 bytes */
typedef struct rt_world_image_texture {
  int32_t kind; /* RT_WORLD_TEX_IMAGE */
  int32_t pad_;
  uint64_t texture_object; /* cudaTextureObject_t over a pitch2D uchar4 resource (image_texture.cu:17-38) */
} rt_world_image_texture; /* 16 bytes */

/* ---- materials ---- */
typedef struct rt_world_lambertian {
  int32_t kind; /* RT_WORLD_MAT_LAMBERTIAN */
  int32_t pad_;
  const void *texture; /* used when use_constant_texture == 0 */
  rt_world_constant_texture color;
  uint8_t use_constant_texture;
  uint8_t pad2_[7];
} rt_world_lambertian; /* 40 bytes */
typedef struct rt_world_metal {
  int32_t kind; /* RT_WORLD_MAT_METAL */
  float albedo[3];
  float fuzz; /* already clamped to <= 1 (metal.cu:10) */
} rt_world_metal; /* 20 bytes */
typedef struct rt_world_dielectric {
  int32_t kind; /* RT_WORLD_MAT_DIELECTRIC */
  float attenuation[3];
  double refractive_index;
} rt_world_dielectric; /* 24 bytes */
typedef struct rt_world_diffuse_light {
  int32_t kind; /* RT_WORLD_MAT_DIFFUSE_LIGHT */
  int32_t pad_;
  const void *texture;
} rt_world_diffuse_light; /* 16 bytes */
typedef struct rt_world_sky_material {
  int32_t kind; /* RT_WORLD_MAT_SKY */
} rt_world_sky_material; /* 4 bytes */

/* ---- hitables ---- */
typedef struct rt_world_list {
  int32_t kind; /* RT_WORLD_HIT_LIST */
  int32_t pad_;
  const void *list[RT_WORLD_MAX_HITABLES];
  int32_t len;
  int32_t pad2_;
} rt_world_list; /* 8208 bytes */
typedef struct rt_world_sphere {
  int32_t kind; /* RT_WORLD_HIT_SPHERE */
  int32_t pad_;
  double radius;
  float position[3];
  int32_t pad2_;
  const void *material;
} rt_world_sphere; /* 40 bytes */
typedef struct rt_world_triangle {
  int32_t kind; /* RT_WORLD_HIT_TRIANGLE */
  float p[3][3];
  const void *material;
} rt_world_triangle; /* 48 bytes */
typedef struct rt_world_parallelogram {
  int32_t kind; /* RT_WORLD_HIT_PARALLELOGRAM */
  float p[4][3]; /* p[3] = p[1] + p[2] - p[0], computed by the constructor (parallelogram.cu:10-15) */
  int32_t pad_;
  const void *material;
} rt_world_parallelogram; /* 64 bytes */
typedef struct rt_world_parallelepiped {
  int32_t kind; /* RT_WORLD_HIT_PARALLELEPIPED */
  int32_t pad_;
  rt_world_parallelogram faces[6]; /* AddCorner order (parallelepiped.cu:25-32) */
  int32_t n_faces;
  int32_t pad2_;
} rt_world_parallelepiped; /* 400 bytes */
typedef struct rt_world_sky {
  int32_t kind; /* RT_WORLD_HIT_SKY */
  rt_world_sky_material material;
} rt_world_sky; /* 8 bytes */
typedef struct rt_world_mesh {
  int32_t kind; /* RT_WORLD_HIT_MESH / RT_WORLD_HIT_MESH_UV */
  int32_t pad_;
  const void *faces; /* Face<false> (9 floats) or Face<true> (15 floats) records, device memory */
  int32_t n_faces;
  int32_t pad2_;
  const void *material; /* one material per mesh (bvh.cuh:178) */
} rt_world_mesh; /* 32 bytes */

/* ---- camera (camera.cuh:12-14) ---- */
typedef struct rt_world_camera {
  float position[3], lower_left_corner[3], horizontal[3], vertical[3], u[3], v[3], w[3];
  uint8_t is_defocus;
  uint8_t pad_[3];
  double lens_radius;
} rt_world_camera; /* 96 bytes */

#ifdef __cplusplus
}
#endif
#endif /* RTWORLD_H_ */
