// rt_world.cu — rt_scene_from_world: the device-resident object list of the scene API (include/rtworld.h) walked on the
// device, flattened into the C-ABI's scene description and committed.
//
// Replaces what the reference does implicitly by rendering straight from the virtual object graph its InitWorld kernels
// build (scenes/cornell_box.cu:35-69; HitableList::Hit, hitable_list.cu:7-25): here the graph is read ONCE, after
// init_world has run, and never touched again.  Nested lists are expanded in place (parallelepiped.cu:20-23 is the same
// scan nested), a Parallelepiped becomes its six parallelograms, materials are deduplicated by object address, image
// textures by texture object.
#include <map>
#include <vector>

#include "../../include/rtworld.h"
#include "rt_internal.h"

namespace {

struct FlatRec {
  rt_hitable h;   // material = -1 here; resolved on the host from material_ptr
  rt_material m;  // description of *material_ptr (texture = -1, or -2 when tex_obj is an image)
  unsigned long long material_ptr, tex_obj, faces_ptr;
  int32_t n_faces, has_uv;
};

__device__ int kind_of(const void *p) { return p ? *static_cast<const int32_t *>(p) : 0; }

__device__ void describe_texture(const void *t, FlatRec *out) {
  if (kind_of(t) == RT_WORLD_TEX_IMAGE) {
    out->m.texture = -2;
    out->tex_obj = static_cast<const rt_world_image_texture *>(t)->texture_object;
  } else if (kind_of(t) == RT_WORLD_TEX_CONSTANT) {
    const rt_world_constant_texture *c = static_cast<const rt_world_constant_texture *>(t);
    for (int k = 0; k < 3; k++) out->m.color[k] = c->color[k];
  }
}

__device__ void describe_material(const void *mp, FlatRec *out) {
  out->material_ptr = (unsigned long long)mp;
  out->tex_obj = 0;
  rt_material &m = out->m;
  m.type = -1;
  m.texture = -1;
  m.color[0] = m.color[1] = m.color[2] = 0.f;
  m.fuzz = 0.f;
  m.ior = 0.0;
  switch (kind_of(mp)) {
    case RT_WORLD_MAT_LAMBERTIAN: {
      const rt_world_lambertian *l = static_cast<const rt_world_lambertian *>(mp);
      m.type = RT_MAT_LAMBERTIAN;
      describe_texture(l->use_constant_texture ? static_cast<const void *>(&l->color) : l->texture, out);
      break;
    }
    case RT_WORLD_MAT_METAL: {
      const rt_world_metal *t = static_cast<const rt_world_metal *>(mp);
      m.type = RT_MAT_METAL;
      for (int k = 0; k < 3; k++) m.color[k] = t->albedo[k];
      m.fuzz = t->fuzz;
      break;
    }
    case RT_WORLD_MAT_DIELECTRIC: {
      const rt_world_dielectric *d = static_cast<const rt_world_dielectric *>(mp);
      m.type = RT_MAT_DIELECTRIC;
      for (int k = 0; k < 3; k++) m.color[k] = d->attenuation[k];
      m.ior = d->refractive_index;
      break;
    }
    case RT_WORLD_MAT_DIFFUSE_LIGHT:
      m.type = RT_MAT_DIFFUSE_LIGHT;
      describe_texture(static_cast<const rt_world_diffuse_light *>(mp)->texture, out);
      break;
    case RT_WORLD_MAT_SKY:
      m.type = RT_MAT_SKY;
      break;
    default:
      break;
  }
}

__device__ void clear_rec(FlatRec *r, int type, int group) {
  memset(r, 0, sizeof(*r));
  r->h.type = type;
  r->h.material = -1;
  r->h.mesh = -1;
  r->h.group = group;
}

__device__ void put_parallelogram(const rt_world_parallelogram &pg, int group, FlatRec *out) {
  clear_rec(out, RT_HIT_PARALLELOGRAM, group);
  for (int k = 0; k < 4; k++)
    for (int c = 0; c < 3; c++) out->h.p[k][c] = pg.p[k][c];
  describe_material(pg.material, out);
}

// Number of flat records under one Hitable (nested lists included), and the records themselves.
__device__ int count_flat(const void *h) {
  switch (kind_of(h)) {
    case RT_WORLD_HIT_LIST: {
      const rt_world_list *l = static_cast<const rt_world_list *>(h);
      int n = 0;
      for (int i = 0; i < l->len && i < RT_WORLD_MAX_HITABLES; i++) n += count_flat(l->list[i]);
      return n;
    }
    case RT_WORLD_HIT_PARALLELEPIPED:
      return static_cast<const rt_world_parallelepiped *>(h)->n_faces;
    case RT_WORLD_HIT_SPHERE:
    case RT_WORLD_HIT_TRIANGLE:
    case RT_WORLD_HIT_PARALLELOGRAM:
    case RT_WORLD_HIT_SKY:
    case RT_WORLD_HIT_MESH:
    case RT_WORLD_HIT_MESH_UV:
      return 1;
    default:
      return 0;  // AABB and unknown objects are not renderable
  }
}

__device__ int emit_flat(const void *h, int group, FlatRec *out) {
  switch (kind_of(h)) {
    case RT_WORLD_HIT_LIST: {
      const rt_world_list *l = static_cast<const rt_world_list *>(h);
      int n = 0;
      for (int i = 0; i < l->len && i < RT_WORLD_MAX_HITABLES; i++) n += emit_flat(l->list[i], group, out + n);
      return n;
    }
    case RT_WORLD_HIT_PARALLELEPIPED: {
      const rt_world_parallelepiped *pp = static_cast<const rt_world_parallelepiped *>(h);
      for (int i = 0; i < pp->n_faces; i++) put_parallelogram(pp->faces[i], group, out + i);
      return pp->n_faces;
    }
    case RT_WORLD_HIT_SPHERE: {
      const rt_world_sphere *s = static_cast<const rt_world_sphere *>(h);
      clear_rec(out, RT_HIT_SPHERE, -1);
      for (int c = 0; c < 3; c++) out->h.p[0][c] = s->position[c];
      out->h.radius = s->radius;
      describe_material(s->material, out);
      return 1;
    }
    case RT_WORLD_HIT_TRIANGLE: {
      const rt_world_triangle *t = static_cast<const rt_world_triangle *>(h);
      clear_rec(out, RT_HIT_TRIANGLE, -1);
      for (int k = 0; k < 3; k++)
        for (int c = 0; c < 3; c++) out->h.p[k][c] = t->p[k][c];
      describe_material(t->material, out);
      return 1;
    }
    case RT_WORLD_HIT_PARALLELOGRAM:
      put_parallelogram(*static_cast<const rt_world_parallelogram *>(h), -1, out);
      return 1;
    case RT_WORLD_HIT_SKY:
      clear_rec(out, RT_HIT_SKY, -1);
      describe_material(&static_cast<const rt_world_sky *>(h)->material, out);
      return 1;
    case RT_WORLD_HIT_MESH:
    case RT_WORLD_HIT_MESH_UV: {
      const rt_world_mesh *b = static_cast<const rt_world_mesh *>(h);
      clear_rec(out, RT_HIT_MESH, -1);
      describe_material(b->material, out);
      out->faces_ptr = (unsigned long long)b->faces;
      out->n_faces = b->n_faces;
      out->has_uv = b->kind == RT_WORLD_HIT_MESH_UV ? 1 : 0;
      return 1;
    }
    default:
      return 0;
  }
}

// pass 1: one thread per top-level entry counts its flat records; pass 2 writes them at the scanned offsets
__global__ void k_world_count(const rt_world_list *world, int *counts, int *list_len) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) *list_len = world->kind == RT_WORLD_HIT_LIST ? world->len : -1;
  if (world->kind == RT_WORLD_HIT_LIST && i < world->len && i < RT_WORLD_MAX_HITABLES) counts[i] = count_flat(world->list[i]);
}
__global__ void k_world_emit(const rt_world_list *world, const int *offsets, FlatRec *out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= world->len || i >= RT_WORLD_MAX_HITABLES) return;
  // a top-level Parallelepiped gets its own group id (its index in the list)
  int group = kind_of(world->list[i]) == RT_WORLD_HIT_PARALLELEPIPED ? i : -1;
  emit_flat(world->list[i], group, out + offsets[i]);
}

}  // namespace

using namespace rtc;

extern "C" int rt_scene_from_world(rt_ctx *ctx, const void *d_world, const void *d_camera, int lens_mode,
                                   rt_build_stats *stats) {
  if (!ctx || !d_world) {
    set_error("rt_scene_from_world: null argument");
    return RT_ERR_INVALID;
  }
  RT_CUDA(cudaSetDevice(ctx->device));
  constexpr int N = RT_WORLD_MAX_HITABLES;
  RT_CUDA(ctx->scratch[0].reserve(sizeof(int) * (2 * N + 1)));
  int *d_counts = ctx->scratch[0].as<int>(), *d_offsets = d_counts + N + 1;
  RT_CUDA(cudaMemset(d_counts, 0, sizeof(int) * (N + 1)));
  // recursion over nested lists: a few frames of a few dozen bytes each
  k_world_count<<<N / 128, 128>>>(static_cast<const rt_world_list *>(d_world), d_counts, d_counts + N);
  RT_CUDA(cudaGetLastError());
  std::vector<int> counts(N + 1);
  RT_CUDA(cudaMemcpy(counts.data(), d_counts, sizeof(int) * counts.size(), cudaMemcpyDeviceToHost));
  const int list_len = counts[N];
  if (list_len < 0 || list_len > N) {
    set_error(list_len < 0 ? "rt_scene_from_world: d_world is not a HitableList record (rtworld.h)"
                           : "rt_scene_from_world: HitableList overflow (more than 1024 entries)");
    return list_len < 0 ? RT_ERR_INVALID : RT_ERR_LIMIT;
  }
  std::vector<int> offsets(N, 0);
  int total = 0;
  for (int i = 0; i < list_len; i++) {
    offsets[i] = total;
    total += counts[i];
  }
  std::vector<FlatRec> flat((size_t)std::max(total, 1));
  if (total > 0) {
    RT_CUDA(ctx->scratch[1].reserve(sizeof(FlatRec) * (size_t)total));
    FlatRec *d_flat = ctx->scratch[1].as<FlatRec>();
    RT_CUDA(cudaMemcpy(d_offsets, offsets.data(), sizeof(int) * offsets.size(), cudaMemcpyHostToDevice));
    k_world_emit<<<N / 128, 128>>>(static_cast<const rt_world_list *>(d_world), d_offsets, d_flat);
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaMemcpy(flat.data(), d_flat, sizeof(FlatRec) * (size_t)total, cudaMemcpyDeviceToHost));
  }

  // materials deduplicated by object address, image textures by texture object
  std::vector<rt_hitable> hitables;
  std::vector<rt_material> materials;
  std::vector<rt_mesh> meshes;
  std::vector<rt_texture> textures;
  std::vector<std::vector<uint8_t>> texels;
  std::map<unsigned long long, int> material_ids, texture_ids;
  for (int i = 0; i < total; i++) {
    FlatRec &r = flat[i];
    if (r.m.type < 0) {
      set_error("rt_scene_from_world: hitable " + std::to_string(i) + " has no (known) material");
      return RT_ERR_INVALID;
    }
    auto it = material_ids.find(r.material_ptr);
    if (it == material_ids.end()) {
      rt_material m = r.m;
      if (m.texture == -2) {
        auto tt = texture_ids.find(r.tex_obj);
        if (tt == texture_ids.end()) {
          cudaResourceDesc res;
          RT_CUDA(cudaGetTextureObjectResourceDesc(&res, (cudaTextureObject_t)r.tex_obj));
          if (res.resType != cudaResourceTypePitch2D) {
            set_error("rt_scene_from_world: an ImageTexture must wrap a pitch2D resource (image_texture.cu:17-38)");
            return RT_ERR_INVALID;
          }
          rt_texture t;
          t.height = (int)res.res.pitch2D.height;
          t.width = (int)res.res.pitch2D.width;
          t.pitch_bytes = (int64_t)t.width * 4;
          texels.emplace_back((size_t)t.pitch_bytes * t.height);
          RT_CUDA(cudaMemcpy2D(texels.back().data(), t.pitch_bytes, res.res.pitch2D.devPtr, res.res.pitch2D.pitchInBytes,
                               (size_t)t.width * 4, t.height, cudaMemcpyDeviceToHost));
          t.h_rgba = nullptr;  // fixed up below
          tt = texture_ids.emplace(r.tex_obj, (int)textures.size()).first;
          textures.push_back(t);
        }
        m.texture = tt->second;
      }
      it = material_ids.emplace(r.material_ptr, (int)materials.size()).first;
      materials.push_back(m);
    }
    r.h.material = it->second;
    if (r.h.type == RT_HIT_MESH) {
      rt_mesh m;
      m.h_faces = (const float *)r.faces_ptr;  // device memory: rt_scene_commit copies with cudaMemcpyDefault
      m.n_faces = r.n_faces;
      m.has_texcoord = r.has_uv;
      r.h.mesh = (int)meshes.size();
      meshes.push_back(m);
    }
    hitables.push_back(r.h);
  }
  for (size_t i = 0; i < textures.size(); i++) textures[i].h_rgba = texels[i].data();

  rt_scene_desc d;
  memset(&d, 0, sizeof(d));
  d.hitables = hitables.data();
  d.n_hitables = (int)hitables.size();
  d.materials = materials.data();
  d.n_materials = (int)materials.size();
  d.meshes = meshes.data();
  d.n_meshes = (int)meshes.size();
  d.textures = textures.data();
  d.n_textures = (int)textures.size();
  int rc = rt_scene_commit(ctx, &d, stats);
  if (rc) return rc;

  if (d_camera) {  // the Camera record the callback placement-new'ed on the device
    rt_world_camera c;
    RT_CUDA(cudaMemcpy(&c, d_camera, sizeof(c), cudaMemcpyDeviceToHost));
    rt_camera cam;
    memset(&cam, 0, sizeof(cam));
    memcpy(cam.position, c.position, 12);
    memcpy(cam.lower_left_corner, c.lower_left_corner, 12);
    memcpy(cam.horizontal, c.horizontal, 12);
    memcpy(cam.vertical, c.vertical, 12);
    memcpy(cam.u, c.u, 12);
    memcpy(cam.v, c.v, 12);
    memcpy(cam.w, c.w, 12);
    cam.is_defocus = c.is_defocus ? (lens_mode == 2 ? 2 : 1) : 0;
    cam.lens_radius = c.lens_radius;
    rc = rt_camera_set(ctx, &cam);
  }
  return rc;
}
