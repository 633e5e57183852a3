// Test tool (tests/test_scene_api.py::test_model_obj_import): loads an OBJ through scene_api/model.h exactly as a
// scene file does (scenes/bunny.cu:64-77 defines STB_IMAGE_IMPLEMENTATION and lets model.h include stb) and prints
// every face of Model<true> and Model<false> as hex floats.  Host code only: no CUDA call is made.
#define STB_IMAGE_IMPLEMENTATION
#include <cstdio>
#include <glm/gtc/matrix_transform.hpp>
#include "model.h"
int main(int argc, char **argv) {
  glm::mat4 m = glm::translate(glm::mat4(1.f), glm::vec3(1.f, 2.f, 3.f)) * glm::scale(glm::mat4(1.f), glm::vec3(2.f));
  Model<true> mt(argv[1], m);
  Model<false> mf(argv[1], m);
  printf("%zu %zu\n", mt.meshes.size(), mf.meshes.size());
  for (auto &mesh : mt.meshes) {
    printf("mesh %zu %d\n", mesh.faces.size(), mesh.texture_id);
    for (auto &f : mesh.faces)
      for (int i = 0; i < 3; i++)
        printf("%a %a %a %a %a\n", f.position(i).x, f.position(i).y, f.position(i).z, f.tex_coord(i).x, f.tex_coord(i).y);
  }
  for (auto &mesh : mf.meshes) {
    printf("mesh %zu %d\n", mesh.faces.size(), mesh.texture_id);
    for (auto &f : mesh.faces)
      for (int i = 0; i < 3; i++) printf("%a %a %a\n", f.position(i).x, f.position(i).y, f.position(i).z);
  }
  for (auto &t : mt.textures) { printf("tex %d %d %zu", t.height, t.width, t.data.size()); for (unsigned char c : t.data) printf(" %d", (int)c); printf("\n"); }
  printf("sizes %zu %zu\n", sizeof(Face<true>), sizeof(Face<false>));
  return 0;
}
