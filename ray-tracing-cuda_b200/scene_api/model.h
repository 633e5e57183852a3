// model.h — Model<HasTexCoord> / Mesh<> / Image of the reference (model.h:14-114) without assimp.
//
// The reference imports any assimp-supported file; this reader handles Wavefront OBJ (what bunny.cu:102
// loads): `v`, `vt`, `f` with v, v/vt, v//vn, v/vt/vn corners, negative indices, polygons fan-triangulated
// (assimp's aiProcess_Triangulate), `usemtl` starting a new mesh, and `mtllib` + `map_Kd` naming the diffuse
// texture of a material, loaded with stbi_load from <model dir>/../../textures/<basename> as model.h:80-88 does
// when the translation unit has stb_image available.  Vertices are transformed on the host by `transform`
// exactly as RecursivelyInitNodes does (model.h:101-103): vec3(M * vec4(p, 1)) / w.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <glm/glm.hpp>
#include <map>
#include <sstream>
#include <string>
#include <vector>

#if defined(__has_include)
#if __has_include(<stb_image.h>)
#include <stb_image.h>  // optional: texture loading as model.h:84-90 (the reference vendors stb)
#endif
#endif

#include "rt_scene_api.cuh"

template <bool HasTexCoord>
struct Mesh {
 public:
  std::vector<Face<HasTexCoord>> faces;
  int32_t texture_id = -1;
};

struct Image {
 public:
  int height = 0, width = 0;
  std::string data;
};

namespace rtapi {
template <bool HasTexCoord>
inline void SetFaceCorner(Face<HasTexCoord> *face, int i, glm::vec3 position, glm::vec2 tex_coord) {
  face->position(i) = position;
  face->tex_coord(i) = tex_coord;
}
template <>
inline void SetFaceCorner<false>(Face<false> *face, int i, glm::vec3 position, glm::vec2) {
  face->position(i) = position;
}
// one "v/vt/vn" corner -> zero-based (position, texcoord) indices; texcoord -1 when absent
inline bool ParseObjCorner(const std::string &tok, int n_pos, int n_tex, int *vi, int *ti) {
  const char *s = tok.c_str();
  char *end = nullptr;
  long v = std::strtol(s, &end, 10);
  if (end == s) return false;
  long t = 0;
  if (*end == '/' && end[1] != '/' && end[1] != '\0') t = std::strtol(end + 1, nullptr, 10);
  *vi = (int)(v > 0 ? v - 1 : n_pos + v);
  *ti = t == 0 ? -1 : (int)(t > 0 ? t - 1 : n_tex + t);
  return *vi >= 0 && *vi < n_pos && *ti < n_tex;
}
}  // namespace rtapi

template <bool HasTexCoord>
struct Model {
 public:
  std::vector<Mesh<HasTexCoord>> meshes;
  std::vector<Image> textures;

  Model(const std::string &path, glm::mat4 transform) {
    std::ifstream in(path);
    CHECK(in.good()) << "cannot open model \"" << path << "\"";
    std::vector<glm::vec3> pos;
    std::vector<glm::vec2> tex;
    std::map<std::string, int> material_ids;  // usemtl name -> texture slot (assimp: material index)
    std::map<std::string, std::string> diffuse_maps;
    int cur = -1;
    auto mesh_for = [&](const std::string &name) {
      auto it = material_ids.find(name);
      if (it == material_ids.end()) {
        it = material_ids.emplace(name, (int)meshes.size()).first;
        meshes.emplace_back();
        textures.emplace_back();
      }
      return it->second;
    };
    std::string line;
    while (std::getline(in, line)) {
      std::istringstream ls(line);
      std::string tag;
      if (!(ls >> tag) || tag[0] == '#') continue;
      if (tag == "v") {
        glm::vec3 p(0);
        ls >> p.x >> p.y >> p.z;
        glm::vec4 q = transform * glm::vec4(p.x, p.y, p.z, 1);
        pos.push_back(glm::vec3(q) / q.w);
      } else if (tag == "vt") {
        glm::vec2 t(0);
        ls >> t.x >> t.y;
        tex.push_back(t);
      } else if (tag == "usemtl") {
        std::string name;
        ls >> name;
        cur = mesh_for(name);
      } else if (tag == "mtllib") {
        std::string name;
        ls >> name;
        LoadMtl(ParentPath(path) + "/" + name, &diffuse_maps);
      } else if (tag == "f") {
        if (cur < 0) cur = mesh_for("");
        std::vector<int> vi, ti;
        std::string tok;
        while (ls >> tok) {
          int v, t;
          CHECK(rtapi::ParseObjCorner(tok, (int)pos.size(), (int)tex.size(), &v, &t)) << "bad face corner \"" << tok << "\"";
          vi.push_back(v);
          ti.push_back(t);
        }
        for (size_t k = 1; k + 1 < vi.size(); k++) {
          Face<HasTexCoord> face;
          const size_t c[3] = {0, k, k + 1};
          for (int j = 0; j < 3; j++) {
            glm::vec2 uv(0);
            if (HasTexCoord && ti[c[j]] >= 0) uv = tex[ti[c[j]]];
            rtapi::SetFaceCorner(&face, j, pos[vi[c[j]]], uv);
          }
          meshes[cur].faces.emplace_back(face);
        }
      }
    }
    for (auto &kv : material_ids) {
      auto it = diffuse_maps.find(kv.first);
      if (it == diffuse_maps.end()) continue;
      std::string image_path = ParentPath(ParentPath(path)) + "/textures/" + BaseName(it->second);
      if (LoadImage(image_path, &textures[kv.second])) meshes[kv.second].texture_id = kv.second;
    }
    LOG(INFO) << "model \"" << path << "\": " << meshes.size() << " mesh(es), " << pos.size() << " vertices";
  }

 private:
  static void LoadMtl(const std::string &path, std::map<std::string, std::string> *diffuse_maps) {
    std::ifstream in(path);
    std::string line, cur;
    while (std::getline(in, line)) {
      std::istringstream ls(line);
      std::string tag, val;
      if (!(ls >> tag >> val)) continue;
      if (tag == "newmtl") cur = val;
      if (tag == "map_Kd") (*diffuse_maps)[cur] = val;
    }
  }
  static bool LoadImage(const std::string &path, Image *out) {
#ifdef STBI_INCLUDE_STB_IMAGE_H
    int channels = 0;
    LOG(INFO) << "loading texture at: \"" << path << "\"";
    unsigned char *data = stbi_load(path.c_str(), &out->width, &out->height, &channels, 4);
    if (!data) return false;
    out->data = std::string(data, data + (size_t)out->width * out->height * 4);
    stbi_image_free(data);
    return true;
#else
    (void)path;
    (void)out;
    return false;
#endif
  }
};
