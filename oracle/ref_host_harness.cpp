// ref_host_harness.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Host driver around the reference's OWN sources (compiled unmodified from /root/reference with
// g++ by oracle/Makefile into oracle/_ref/libref_host.so).  It builds scenes through the
// reference's constructors / its scenes' InitWorld functions, runs the reference's
// HitableList::Hit and PathTracing on CPU cores, and flattens the reference's object graph into
// the rt_scene_desc records of include/rtcore.h.  It contains no restated reference arithmetic:
// every intersection / scatter / emit executed here is the reference's code.
//
// Used to (a) pin oracle/rt_oracle.cpp, (b) generate tests/golden/*, (c) serve as the
// `cpu_baseline.kind = "reference"` timing arm of bench.py.

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <thread>
#include <tuple>
#include <vector>

#include <glm/glm.hpp>
#include <glm/gtc/constants.hpp>
#define GLM_ENABLE_EXPERIMENTAL
#include <glm/gtx/rotate_vector.hpp>

// Expose the reference classes' private fields to the flattening code below (layout unchanged).
#define private public
#define protected public
#include "bvh.cuh"
#include "camera.cuh"
#include "dielectric.cuh"
#include "diffuse_light.cuh"
#include "hitable_list.cuh"
#include "lambertian.cuh"
#include "metal.cuh"
#include "parallelepiped.cuh"
#include "parallelogram.cuh"
#include "ray_tracing.cuh"
#include "sky.cuh"
#include "sphere.cuh"
#include "textures/constant_texture.cuh"
#include "textures/image_texture.cuh"
#include "triangle.cuh"
#include "utils.cuh"
#undef private
#undef protected

#include "../include/rtcore.h"

namespace rt_shim {
thread_local Idx3 t_threadIdx{0, 0, 0}, t_blockIdx{0, 0, 0}, t_blockDim{1, 1, 1}, t_gridDim{1, 1, 1};

float4 Tex2DPointWrap(cudaTextureObject_t tex, float u, float v) {
  const HostTexture *t = reinterpret_cast<const HostTexture *>(tex);
  float x = u - std::floor(u), y = v - std::floor(v);
  int ix = (int)std::floor(x * (float)t->width), iy = (int)std::floor(y * (float)t->height);
  ix = std::min(std::max(ix, 0), t->width - 1);
  iy = std::min(std::max(iy, 0), t->height - 1);
  const uint8_t *px = t->data + (size_t)iy * t->pitch + (size_t)ix * 4;
  return make_float4(px[0] / 255.0f, px[1] / 255.0f, px[2] / 255.0f, px[3] / 255.0f);
}
}  // namespace rt_shim

// Entry points of the reference scenes' InitWorld/InitModel (oracle/ref_scenes.cpp).
namespace scene_spheres { void InitWorld(HitableList *, Camera *, curandState *); }
namespace scene_cornell_box { void InitWorld(HitableList *, Camera *); }
namespace scene_bunny {
void InitWorld(HitableList *, Camera *);
void InitModel(HitableList *, Face<false> *, int);
}
namespace scene_birthday { void InitWorld(HitableList *, Camera *, cudaTextureObject_t); }

namespace {

struct RefScene {
  HitableList *world = nullptr;
  Camera *camera = nullptr;
  curandState state0;     // states[0] after scene construction (spheres draws from it)
  bool has_state0 = false;
  uint64_t seed = 0;
  std::vector<uint8_t> tex_data;
  rt_shim::HostTexture tex{};
  std::vector<Face<false>> faces;  // storage handed to the reference BVH (sorted in place by it)
  std::vector<std::vector<Face<true>>> faces_uv;
  std::vector<std::vector<Face<false>>> faces_nouv;
  std::vector<std::vector<uint8_t>> tex_store;
  std::vector<std::unique_ptr<rt_shim::HostTexture>> tex_objs;

  // flattened view
  std::vector<rt_hitable> hitables;
  std::vector<Hitable *> flat_objs;       // reference object behind each flattened hitable
  std::vector<rt_material> materials;
  std::map<Material *, int> mat_ids;
  std::vector<const void *> mesh_objs;    // BVH<...>* per mesh
  std::vector<int> mesh_has_uv;
  std::vector<cudaTextureObject_t> tex_ids;
};

void st3(float *d, glm::vec3 v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; }

int texture_index(RefScene &s, cudaTextureObject_t t) {
  for (size_t i = 0; i < s.tex_ids.size(); i++)
    if (s.tex_ids[i] == t) return (int)i;
  s.tex_ids.push_back(t);
  return (int)s.tex_ids.size() - 1;
}

void describe_texture(RefScene &s, Texture *tx, rt_material *m) {
  if (auto c = dynamic_cast<ConstantTexture *>(tx)) {
    m->texture = -1;
    st3(m->color, c->color_);
  } else if (auto im = dynamic_cast<ImageTexture *>(tx)) {
    m->texture = texture_index(s, im->image_texture_);
  }
}

int material_index(RefScene &s, Material *mp) {
  auto it = s.mat_ids.find(mp);
  if (it != s.mat_ids.end()) return it->second;
  rt_material m;
  std::memset(&m, 0, sizeof(m));
  m.texture = -1;
  if (auto l = dynamic_cast<Lambertian *>(mp)) {
    m.type = RT_MAT_LAMBERTIAN;
    describe_texture(s, l->texture_ptr(), &m);
  } else if (auto me = dynamic_cast<Metal *>(mp)) {
    m.type = RT_MAT_METAL;
    st3(m.color, me->albedo_);
    m.fuzz = me->fuzz_;
  } else if (auto d = dynamic_cast<Dielectric *>(mp)) {
    m.type = RT_MAT_DIELECTRIC;
    st3(m.color, d->attenuation_);
    m.ior = d->refractive_index_;
  } else if (auto dl = dynamic_cast<DiffuseLight *>(mp)) {
    m.type = RT_MAT_DIFFUSE_LIGHT;
    describe_texture(s, dl->texture_ptr_, &m);
  } else if (dynamic_cast<SkyMaterial *>(mp)) {
    m.type = RT_MAT_SKY;
  } else {
    m.type = -1;
  }
  int id = (int)s.materials.size();
  s.materials.push_back(m);
  s.mat_ids[mp] = id;
  return id;
}

void flatten_one(RefScene &s, Hitable *h, int group) {
  rt_hitable r;
  std::memset(&r, 0, sizeof(r));
  r.mesh = -1;
  r.group = group;
  if (auto sp = dynamic_cast<Sphere *>(h)) {
    r.type = RT_HIT_SPHERE;
    st3(r.p[0], sp->position());
    r.radius = sp->radius();
    r.material = material_index(s, sp->material_ptr());
  } else if (auto tr = dynamic_cast<Triangle *>(h)) {
    r.type = RT_HIT_TRIANGLE;
    for (int i = 0; i < 3; i++) st3(r.p[i], tr->p_[i]);
    r.material = material_index(s, tr->material_ptr_);
  } else if (auto pg = dynamic_cast<Parallelogram *>(h)) {
    r.type = RT_HIT_PARALLELOGRAM;
    for (int i = 0; i < 4; i++) st3(r.p[i], pg->p_[i]);
    r.material = material_index(s, pg->material_ptr_);
  } else if (auto sk = dynamic_cast<Sky *>(h)) {
    r.type = RT_HIT_SKY;
    r.material = material_index(s, sk->material_ptr());
  } else if (auto pp = dynamic_cast<Parallelepiped *>(h)) {
    static int next_group = 0;
    int g = next_group++;
    for (int i = 0; i < pp->list_.list_len(); i++) flatten_one(s, pp->list_.list_[i], g);
    return;
  } else if (auto b0 = dynamic_cast<BVH<Face<false>, AABB> *>(h)) {
    r.type = RT_HIT_MESH;
    r.mesh = (int)s.mesh_objs.size();
    s.mesh_objs.push_back(b0);
    s.mesh_has_uv.push_back(0);
    r.material = material_index(s, b0->material_ptr_);
  } else if (auto b1 = dynamic_cast<BVH<Face<true>, AABB> *>(h)) {
    r.type = RT_HIT_MESH;
    r.mesh = (int)s.mesh_objs.size();
    s.mesh_objs.push_back(b1);
    s.mesh_has_uv.push_back(1);
    r.material = material_index(s, b1->material_ptr_);
  } else {
    r.type = -1;
  }
  s.hitables.push_back(r);
  s.flat_objs.push_back(h);
}

void flatten(RefScene &s) {
  s.hitables.clear();
  s.flat_objs.clear();
  s.materials.clear();
  s.mat_ids.clear();
  s.mesh_objs.clear();
  s.mesh_has_uv.clear();
  for (int i = 0; i < s.world->list_len(); i++) flatten_one(s, s.world->list_[i], -1);
}

Ray raw_ray(const float *r) {
  Ray ray;  // bypass Ray::Ray's normalization: the direction is used exactly as stored
  ray.position_ = glm::vec3(r[0], r[1], r[2]);
  ray.direction_ = glm::vec3(r[3], r[4], r[5]);
  return ray;
}

template <typename F>
void parallel_for(int n, int nthreads, F fn) {
  if (nthreads <= 0) nthreads = (int)std::thread::hardware_concurrency();
  if (nthreads <= 0) nthreads = 1;
  std::atomic<int> next{0};
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; t++)
    th.emplace_back([&]() {
      for (;;) {
        int i = next.fetch_add(1);
        if (i >= n) break;
        fn(i);
      }
    });
  for (auto &x : th) x.join();
}

// Which face produced `rec` inside a reference BVH: the highest index whose own Hit returns the
// same (t, normal).  Uses the reference's Face::Hit only.
template <bool UV>
int find_face(BVH<Face<UV>, AABB> *b, const Ray &ray, double t_from, const HitRecord &rec) {
  int found = -1;
  for (int i = 0; i < b->n_; i++) {
    HitRecord hr;
    if (b->objs_[i].Hit(ray, t_from, rec.t, &hr) && hr.t == rec.t && hr.normal == rec.normal) found = i;
  }
  return found;
}

}  // namespace

extern "C" {

// Build a named reference scene by running the scene file's own InitWorld (and InitModel) on the
// host.  faces9: caller-order Face<false> array for "bunny"; rgba/tex_*: texture for "birthday".
void *ref_scene_create_named(const char *name, uint64_t seed, const float *faces9, int n_faces,
                             const uint8_t *rgba, int tex_h, int tex_w) {
  RefScene *s = new RefScene();
  s->world = (HitableList *)::operator new(sizeof(HitableList));
  s->camera = (Camera *)::operator new(sizeof(Camera));
  s->seed = seed;
  std::string n(name);
  if (n == "spheres") {
    curand_init(seed, 0, 0, &s->state0);  // utils.cu:202 + CudaRandomInit idx 0
    s->has_state0 = true;
    scene_spheres::InitWorld(s->world, s->camera, &s->state0);
  } else if (n == "cornell_box") {
    scene_cornell_box::InitWorld(s->world, s->camera);
  } else if (n == "bunny") {
    scene_bunny::InitWorld(s->world, s->camera);
    s->faces.resize(n_faces);
    std::memcpy((void *)s->faces.data(), faces9, sizeof(Face<false>) * (size_t)n_faces);
    scene_bunny::InitModel(s->world, s->faces.data(), n_faces);
  } else if (n == "birthday") {
    s->tex_data.assign(rgba, rgba + (size_t)tex_h * tex_w * 4);
    s->tex = rt_shim::HostTexture{s->tex_data.data(), tex_h, tex_w, (uint64_t)tex_w * 4};
    scene_birthday::InitWorld(s->world, s->camera, (cudaTextureObject_t)(uintptr_t)&s->tex);
  } else {
    delete s;
    return nullptr;
  }
  flatten(*s);
  return s;
}

// Build a scene from a description through the reference's constructors (random test scenes).
void *ref_scene_create(const rt_scene_desc *d, const rt_camera *cam) {
  RefScene *s = new RefScene();
  s->world = new HitableList();
  std::vector<Texture *> texs;
  for (int i = 0; i < d->n_textures; i++) {
    const rt_texture &t = d->textures[i];
    s->tex_store.emplace_back(t.h_rgba, t.h_rgba + (size_t)t.pitch_bytes * t.height);
    s->tex_objs.emplace_back(new rt_shim::HostTexture{s->tex_store.back().data(), t.height, t.width, (uint64_t)t.pitch_bytes});
    texs.push_back(new ImageTexture((cudaTextureObject_t)(uintptr_t)s->tex_objs.back().get()));
  }
  std::vector<Material *> mats;
  for (int i = 0; i < d->n_materials; i++) {
    const rt_material &m = d->materials[i];
    glm::vec3 c(m.color[0], m.color[1], m.color[2]);
    switch (m.type) {
      case RT_MAT_LAMBERTIAN:
        mats.push_back(m.texture >= 0 ? new Lambertian(texs[m.texture]) : new Lambertian(c));
        break;
      case RT_MAT_METAL: mats.push_back(new Metal(c, m.fuzz)); break;
      case RT_MAT_DIELECTRIC: mats.push_back(new Dielectric(c, m.ior)); break;
      case RT_MAT_DIFFUSE_LIGHT:
        mats.push_back(new DiffuseLight(m.texture >= 0 ? texs[m.texture] : (Texture *)new ConstantTexture(c)));
        break;
      default: mats.push_back(nullptr); break;  // sky material is owned by Sky
    }
  }
  for (int i = 0; i < d->n_hitables; i++) {
    const rt_hitable &h = d->hitables[i];
    glm::vec3 p[4];
    for (int k = 0; k < 4; k++) p[k] = glm::vec3(h.p[k][0], h.p[k][1], h.p[k][2]);
    switch (h.type) {
      case RT_HIT_SPHERE: s->world->Append(new Sphere(p[0], h.radius, mats[h.material])); break;
      case RT_HIT_TRIANGLE: s->world->Append(new Triangle(p, mats[h.material])); break;
      case RT_HIT_PARALLELOGRAM: s->world->Append(new Parallelogram(p, mats[h.material])); break;
      case RT_HIT_SKY: s->world->Append(new Sky()); break;
      case RT_HIT_MESH: {
        const rt_mesh &m = d->meshes[h.mesh];
        if (m.has_texcoord) {
          s->faces_uv.emplace_back(m.n_faces);
          std::memcpy((void *)s->faces_uv.back().data(), m.h_faces, sizeof(Face<true>) * (size_t)m.n_faces);
          s->world->Append(new BVH<Face<true>, AABB>(s->faces_uv.back().data(), m.n_faces, mats[h.material]));
        } else {
          s->faces_nouv.emplace_back(m.n_faces);
          std::memcpy((void *)s->faces_nouv.back().data(), m.h_faces, sizeof(Face<false>) * (size_t)m.n_faces);
          s->world->Append(new BVH<Face<false>, AABB>(s->faces_nouv.back().data(), m.n_faces, mats[h.material]));
        }
        break;
      }
    }
  }
  s->camera = (Camera *)::operator new(sizeof(Camera));
  if (cam) {
    auto v = [](const float *a) { return glm::vec3(a[0], a[1], a[2]); };
    new (s->camera) Camera(v(cam->position), v(cam->lower_left_corner), v(cam->horizontal), v(cam->vertical));
    s->camera->u_ = v(cam->u);
    s->camera->v_ = v(cam->v);
    s->camera->w_ = v(cam->w);
    s->camera->is_defocus_camera_ = cam->is_defocus != 0;
    s->camera->lens_radius_ = cam->lens_radius;
  }
  flatten(*s);
  return s;
}

// Camera constructors (camera.cu:6-38) run by the reference itself.
void ref_camera_look_at(const float pos[3], const float at[3], const float up[3], double fov, double aspect,
                        double aperture, double focus, rt_camera *out) {
  auto v = [](const float *a) { return glm::vec3(a[0], a[1], a[2]); };
  Camera *c = aperture >= 0 ? new Camera(v(pos), v(at), v(up), fov, aspect, aperture, focus)
                            : new Camera(v(pos), v(at), v(up), fov, aspect);
  std::memset(out, 0, sizeof(*out));
  st3(out->position, c->position_);
  st3(out->lower_left_corner, c->lower_left_corner_);
  st3(out->horizontal, c->horizontal_);
  st3(out->vertical, c->vertical_);
  st3(out->u, c->u_);
  st3(out->v, c->v_);
  st3(out->w, c->w_);
  out->is_defocus = c->is_defocus_camera_ ? 1 : 0;
  out->lens_radius = c->lens_radius_;
  delete c;
}

void ref_scene_camera(void *sp, rt_camera *out) {
  RefScene *s = (RefScene *)sp;
  Camera *c = s->camera;
  std::memset(out, 0, sizeof(*out));
  st3(out->position, c->position_);
  st3(out->lower_left_corner, c->lower_left_corner_);
  st3(out->horizontal, c->horizontal_);
  st3(out->vertical, c->vertical_);
  st3(out->u, c->u_);
  st3(out->v, c->v_);
  st3(out->w, c->w_);
  out->is_defocus = c->is_defocus_camera_ ? 1 : 0;
  out->lens_radius = c->lens_radius_;
}

void ref_scene_destroy(void *sp) { delete (RefScene *)sp; /* reference objects leak, as in the reference */ }

// Flattened description of the reference object graph.
void ref_scene_counts(void *sp, int *n_hitables, int *n_materials, int *n_meshes, int *n_textures) {
  RefScene *s = (RefScene *)sp;
  *n_hitables = (int)s->hitables.size();
  *n_materials = (int)s->materials.size();
  *n_meshes = (int)s->mesh_objs.size();
  *n_textures = (int)s->tex_ids.size();
}
void ref_scene_describe(void *sp, rt_hitable *hitables, rt_material *materials) {
  RefScene *s = (RefScene *)sp;
  std::memcpy(hitables, s->hitables.data(), sizeof(rt_hitable) * s->hitables.size());
  std::memcpy(materials, s->materials.data(), sizeof(rt_material) * s->materials.size());
}
int ref_scene_mesh_info(void *sp, int mesh, int *has_uv) {
  RefScene *s = (RefScene *)sp;
  *has_uv = s->mesh_has_uv[mesh];
  return s->mesh_has_uv[mesh] ? ((BVH<Face<true>, AABB> *)s->mesh_objs[mesh])->n_
                              : ((BVH<Face<false>, AABB> *)s->mesh_objs[mesh])->n_;
}
// Faces of mesh `mesh` as the reference holds them after its in-place sort (9 or 15 floats each).
void ref_scene_mesh_faces(void *sp, int mesh, float *out) {
  RefScene *s = (RefScene *)sp;
  if (s->mesh_has_uv[mesh]) {
    auto b = (BVH<Face<true>, AABB> *)s->mesh_objs[mesh];
    std::memcpy(out, (void *)b->objs_, sizeof(Face<true>) * (size_t)b->n_);
  } else {
    auto b = (BVH<Face<false>, AABB> *)s->mesh_objs[mesh];
    std::memcpy(out, (void *)b->objs_, sizeof(Face<false>) * (size_t)b->n_);
  }
}
int ref_state0_draws_hint(void) { return 0; }

// world->Hit on explicit rays (hitable_list.cu:7-25 executed by the reference), plus ids
// recovered by re-running the reference's own per-object Hit in list order.
// Returns the number of rays whose id-recovery scan disagreed with world->Hit (must be 0).
int64_t ref_trace_closest(void *sp, const float *rays, int64_t n, double t_from, double t_to, rt_hit *out,
                          int nthreads) {
  RefScene *s = (RefScene *)sp;
  std::atomic<int64_t> bad{0};
  const int64_t chunk = 2048;
  int nchunks = (int)((n + chunk - 1) / chunk);
  parallel_for(nchunks, nthreads, [&](int c) {
    for (int64_t i = (int64_t)c * chunk; i < std::min(n, (int64_t)(c + 1) * chunk); i++) {
      Ray ray = raw_ray(rays + 6 * i);
      HitRecord rec;
      std::memset(&rec, 0, sizeof(rec));
      bool hit = s->world->Hit(ray, t_from, t_to, &rec);
      rt_hit h;
      std::memset(&h, 0, sizeof(h));
      if (!hit) {
        h.obj = -1;
        h.sub = -1;
        h.material = -1;
        out[i] = h;
        continue;
      }
      // id recovery: the same scan as HitableList::Hit over the flattened objects
      int obj = -1;
      double cur_to = t_to;
      bool ok = false;
      HitRecord best, hr;
      std::memset(&hr, 0, sizeof(hr));
      double t_to_at_winner = t_to;
      for (int k = 0; k < (int)s->flat_objs.size(); k++) {
        if (s->flat_objs[k]->Hit(ray, t_from, cur_to, &hr)) {
          if (!ok) {
            best = hr; obj = k; t_to_at_winner = cur_to;
            cur_to = std::fmin(cur_to, hr.t);
            ok = true;
          } else if (hr.t < cur_to) {
            t_to_at_winner = cur_to;
            cur_to = hr.t; best = hr; obj = k;
          }
        }
      }
      if (!ok || best.t != rec.t || best.material_ptr != rec.material_ptr) bad++;
      int sub = 0;
      const rt_hitable &fh = s->hitables[obj];
      if (fh.type == RT_HIT_PARALLELOGRAM) {
        auto pg = (Parallelogram *)s->flat_objs[obj];
        double tt, uu, vv;
        glm::vec3 nn;
        sub = TriangleHit(pg->p_, ray, t_from, t_to_at_winner, &tt, &nn, &uu, &vv) ? 0 : 1;
      } else if (fh.type == RT_HIT_MESH) {
        sub = s->mesh_has_uv[fh.mesh] ? find_face<true>((BVH<Face<true>, AABB> *)s->flat_objs[obj], ray, t_from, rec)
                                      : find_face<false>((BVH<Face<false>, AABB> *)s->flat_objs[obj], ray, t_from, rec);
      }
      bool sky = fh.type == RT_HIT_SKY;
      bool nouv_mesh = fh.type == RT_HIT_MESH && !s->mesh_has_uv[fh.mesh];
      h.t = rec.t;
      h.obj = obj;
      h.sub = sub;
      h.u = (sky || nouv_mesh) ? 0.f : (float)rec.u;
      h.v = (sky || nouv_mesh) ? 0.f : (float)rec.v;
      h.normal[0] = sky ? 0.f : rec.normal.x;
      h.normal[1] = sky ? 0.f : rec.normal.y;
      h.normal[2] = sky ? 0.f : rec.normal.z;
      h.material = material_index(*s, rec.material_ptr);
      out[i] = h;
    }
  });
  return bad.load();
}

// The reference's PathTracing kernel body (ray_tracing.cu:56-85) run once per pixel on the host,
// with the reference's curand streams: curand_init(seed, idx, 0) (utils.cu:43-47).
void ref_render(void *sp, int height, int width, int spp, int post_processing, float *out_rgb, int nthreads) {
  RefScene *s = (RefScene *)sp;
  std::vector<curandState> states((size_t)height * width);
  parallel_for(height, nthreads, [&](int i) {
    for (int j = 0; j < width; j++) {
      int idx = i * width + j;
      if (idx == 0 && s->has_state0)
        states[0] = s->state0;
      else
        curand_init(s->seed, idx, 0, &states[idx]);
    }
  });
  parallel_for(height, nthreads, [&](int i) {
    for (int j = 0; j < width; j++) {
      rt_shim::t_blockDim = {1, 1, 1};
      rt_shim::t_threadIdx = {0, 0, 0};
      rt_shim::t_blockIdx = {(unsigned)i, (unsigned)j, 0};
      PathTracing(s->world, s->camera, height, width, spp, post_processing != 0, states.data(),
                  (glm::vec3 *)out_rgb);
    }
  });
}

// Raw access to the reference's TriangleHit / AABB::Hit for known-answer tests.
int ref_triangle_hit(const float p9[9], const float ray6[6], double t_from, double t_to, double *t,
                     float normal[3], double *u, double *v) {
  glm::vec3 p[3] = {{p9[0], p9[1], p9[2]}, {p9[3], p9[4], p9[5]}, {p9[6], p9[7], p9[8]}};
  Ray ray = raw_ray(ray6);
  glm::vec3 n(0);
  *t = *u = *v = 0;
  int r = TriangleHit(p, ray, t_from, t_to, t, &n, u, v) ? 1 : 0;
  st3(normal, n);
  return r;
}
int ref_aabb_hit(const float mn[3], const float mx[3], const float ray6[6], double t_from, double t_to) {
  AABB b;
  b.min = glm::vec3(mn[0], mn[1], mn[2]);
  b.max = glm::vec3(mx[0], mx[1], mx[2]);
  Ray ray = raw_ray(ray6);
  HitRecord rec;
  return b.Hit(ray, t_from, t_to, &rec) ? 1 : 0;
}
int ref_get_workload(int rank, int world_size, int spp) { return GetWorkload(rank, world_size, spp); }
int ref_hardware_threads(void) { return (int)std::thread::hardware_concurrency(); }

}  // extern "C"
