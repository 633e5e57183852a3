// ref_cuda_harness.cu — MEASUREMENT INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Driver around the reference's OWN CUDA sources (compiled unmodified from /root/reference by
// baseline/Makefile).  It does what the reference's Main() does (utils.cu:132-179): cudaMalloc,
// CudaRandomInit, the scene file's own InitWorld<<<1,1>>>, then times PathTracing<<<>>> with a
// cudaEvent pair exactly as the reference does (utils.cu:155-170) — but with height/width/spp taken
// from the command line, the device heap/stack limits raised for big meshes, the raw float image
// written to a file, and an extra kernel that runs the reference's world->Hit on explicit rays.
//   ref_cuda <spheres|cornell_box|bunny|birthday|mesh> H W spp reps [--faces f.bin] [--tex f.bin h w]
//            [--camera px py pz lx ly lz fov aspect] [--image out.bin] [--rays rays.bin hits.bin]
//            [--seed s] [--post 0|1]
//   ref_cuda powkat 0 0 0 0     (known-answer test of pow(x, 2) / pow(x, 0.5) against x * x / sqrt(x))
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <string>
#include <tuple>
#include <vector>

#include <glm/glm.hpp>
#include <glm/gtc/constants.hpp>
#ifndef GLM_ENABLE_EXPERIMENTAL
#define GLM_ENABLE_EXPERIMENTAL
#endif
#include <glm/gtx/rotate_vector.hpp>

#define private public
#define protected public
#include "ray.cuh"
#undef private
#undef protected
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
#include "utils.cuh"

namespace scene_spheres {
#include "gen/spheres_head.inc"
}
namespace scene_cornell_box {
#include "gen/cornell_box_head.inc"
}
namespace scene_bunny {
#include "gen/bunny_head.inc"
}
namespace scene_birthday {
#include "gen/birthday_head.inc"
}

#define CK(x)                                                                     \
  do {                                                                            \
    cudaError_t e_ = (x);                                                         \
    if (e_ != cudaSuccess) {                                                      \
      fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(2);                                                                    \
    }                                                                             \
  } while (0)

// Generic mesh scene (BASELINE config 5): Sky + one BVH<Face<false>,AABB> with Lambertian(0.7).
__global__ void InitMeshWorld(HitableList *world, Camera *camera, Face<false> *faces, int n, glm::vec3 pos,
                              glm::vec3 look, double fov, double aspect) {
  new (world) HitableList();
  new (camera) Camera(pos, look, glm::vec3(0, 1, 0), fov, aspect);
  auto mat = new Lambertian(glm::vec3(0.7, 0.7, 0.7));
  world->Append(new BVH<Face<false>, AABB>(faces, n, mat));
  world->Append(new Sky());
}

struct RefHit {
  double t, u, v;
  float normal[3];
  int hit;
};

// world->Hit (hitable_list.cu:7-25) on explicit rays; directions used as stored.
__global__ void TraceRays(HitableList *world, const float *rays, long n, double t_from, double t_to, RefHit *out) {
  long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  Ray ray;
  ray.position_ = glm::vec3(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]);
  ray.direction_ = glm::vec3(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]);
  HitRecord rec;
  rec.t = rec.u = rec.v = 0;
  rec.normal = glm::vec3(0);
  bool hit = world->Hit(ray, t_from, t_to, &rec);
  RefHit h;
  h.t = hit ? rec.t : 0;
  h.u = rec.u;
  h.v = rec.v;
  h.normal[0] = rec.normal.x;
  h.normal[1] = rec.normal.y;
  h.normal[2] = rec.normal.z;
  h.hit = hit ? 1 : 0;
  out[i] = h;
}

// Known-answer test of the two pow() calls that decide a sphere hit (sphere.cu:13-21: pow(len, 2), pow(disc, 0.5);
// SURVEY 7.3-1): is CUDA's pow(x, 2) bitwise x * x and pow(x, 0.5) bitwise sqrt(x) on the operand ranges of the
// scenes?  Operands: lengths in [1e-3, 4e3] as float -> double (what length() returns), discriminants in [0, 1e9]
// as doubles with full mantissas, log-uniform, from a 64-bit LCG.  counts = {n, pow2 != x*x, pow.5 != sqrt, max ulp
// distance of each}.
__global__ void PowKat(unsigned long long n, unsigned long long *counts) {
  unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  unsigned long long bad2 = 0, badh = 0, ulp2 = 0, ulph = 0;
  for (; i < n; i += (unsigned long long)gridDim.x * blockDim.x) {
    unsigned long long z = i * 6364136223846793005ULL + 1442695040888963407ULL;
    z ^= z >> 29;
    z *= 0xbf58476d1ce4e5b9ULL;
    z ^= z >> 32;
    double u = (double)(z >> 11) * (1.0 / 9007199254740992.0);  // [0, 1)
    double len = (double)(float)exp(log(1e-3) + u * (log(4e3) - log(1e-3)));
    double disc = exp(log(1e-12) + u * (log(1e9) - log(1e-12)));
    double a = pow(len, 2), b = len * len;
    double c = pow(disc, 0.5), d = sqrt(disc);
    long long da = __double_as_longlong(a) - __double_as_longlong(b), dc = __double_as_longlong(c) - __double_as_longlong(d);
    if (da) bad2++;
    if (dc) badh++;
    ulp2 = max(ulp2, (unsigned long long)(da < 0 ? -da : da));
    ulph = max(ulph, (unsigned long long)(dc < 0 ? -dc : dc));
  }
  if (bad2) atomicAdd(counts + 1, bad2);
  if (badh) atomicAdd(counts + 2, badh);
  atomicMax(counts + 3, ulp2);
  atomicMax(counts + 4, ulph);
}

static std::vector<char> read_file(const char *path) {
  std::ifstream f(path, std::ios::binary);
  if (!f) {
    fprintf(stderr, "cannot open %s\n", path);
    exit(2);
  }
  return std::vector<char>((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
}

int main(int argc, char **argv) {
  if (argc < 6) {
    fprintf(stderr, "usage: ref_cuda scene H W spp reps [options]\n");
    return 2;
  }
  std::string scene = argv[1];
  if (scene == "powkat") {  // ref_cuda powkat _ _ _ _ : prints the KAT counts
    unsigned long long n = 1ull << 26, h[5] = {0, 0, 0, 0, 0}, *d;
    CK(cudaMalloc(&d, sizeof(h)));
    CK(cudaMemset(d, 0, sizeof(h)));
    PowKat<<<148 * 8, 256>>>(n, d);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost));
    printf("{\"powkat_operands\": %llu, \"pow2_ne_mul\": %llu, \"powhalf_ne_sqrt\": %llu, \"pow2_max_ulp\": %llu, "
           "\"powhalf_max_ulp\": %llu}\n", n, h[1], h[2], h[3], h[4]);
    return 0;
  }
  int H = atoi(argv[2]), W = atoi(argv[3]), spp = atoi(argv[4]), reps = atoi(argv[5]);
  const char *faces_path = nullptr, *tex_path = nullptr, *image_path = nullptr, *rays_path = nullptr, *hits_path = nullptr;
  int tex_h = 0, tex_w = 0, post = 0;
  uint64_t seed = 10086;  // DistributedMain's seed (utils.cu:202); spheres' layout depends on it
  float cam[8] = {0, 0, -4, 0, 0, 0, 0.6981317f, 16.f / 9.f};
  for (int i = 6; i < argc; i++) {
    std::string a = argv[i];
    if (a == "--faces") faces_path = argv[++i];
    else if (a == "--tex") { tex_path = argv[++i]; tex_h = atoi(argv[++i]); tex_w = atoi(argv[++i]); }
    else if (a == "--image") image_path = argv[++i];
    else if (a == "--rays") { rays_path = argv[++i]; hits_path = argv[++i]; }
    else if (a == "--seed") seed = strtoull(argv[++i], nullptr, 10);
    else if (a == "--post") post = atoi(argv[++i]);
    else if (a == "--camera") for (int k = 0; k < 8; k++) cam[k] = (float)atof(argv[++i]);
  }
  CK(cudaSetDevice(0));
  // the reference never raises these; big meshes need them (device-side new + thrust temp + recursion)
  CK(cudaDeviceSetLimit(cudaLimitMallocHeapSize, (size_t)4 << 30));
  CK(cudaDeviceSetLimit(cudaLimitStackSize, 16 << 10));

  curandState *d_states;
  Camera *d_camera;
  HitableList *d_world;
  glm::vec3 *d_image;
  size_t n = (size_t)H * W;
  CK(cudaMalloc(&d_states, sizeof(curandState) * n));
  CK(cudaMalloc(&d_image, sizeof(glm::vec3) * n));
  CK(cudaMalloc(&d_world, sizeof(HitableList)));
  CK(cudaMalloc(&d_camera, sizeof(Camera)));
  CudaRandomInit<<<(unsigned)((n + 63) / 64), 64>>>(seed, d_states, (int)n);
  CK(cudaDeviceSynchronize());

  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  float build_ms = 0;
  CK(cudaEventRecord(e0));
  if (scene == "spheres") {
    scene_spheres::InitWorld<<<1, 1>>>(d_world, d_camera, d_states);
  } else if (scene == "cornell_box") {
    scene_cornell_box::InitWorld<<<1, 1>>>(d_world, d_camera);
  } else if (scene == "bunny" || scene == "mesh") {
    if (!faces_path) { fprintf(stderr, "--faces required\n"); return 2; }
    auto buf = read_file(faces_path);
    int nf = (int)(buf.size() / sizeof(Face<false>));
    Face<false> *d_faces;
    CK(cudaMalloc(&d_faces, buf.size()));
    CK(cudaMemcpy(d_faces, buf.data(), buf.size(), cudaMemcpyHostToDevice));
    if (scene == "bunny") {
      scene_bunny::InitWorld<<<1, 1>>>(d_world, d_camera);
      scene_bunny::InitModel<<<1, 1>>>(d_world, d_faces, nf);
    } else {
      InitMeshWorld<<<1, 1>>>(d_world, d_camera, d_faces, nf, glm::vec3(cam[0], cam[1], cam[2]),
                              glm::vec3(cam[3], cam[4], cam[5]), (double)cam[6], (double)cam[7]);
    }
  } else if (scene == "birthday") {
    if (!tex_path) { fprintf(stderr, "--tex required\n"); return 2; }
    auto buf = read_file(tex_path);
    uint8_t *d_buf;
    size_t pitch;
    CK(cudaMallocPitch(&d_buf, &pitch, (size_t)tex_w * 4, tex_h));
    CK(cudaMemcpy2D(d_buf, pitch, buf.data(), (size_t)tex_w * 4, (size_t)tex_w * 4, tex_h, cudaMemcpyHostToDevice));
    auto tex = ImageTexture::CreateCudaTextureObj(d_buf, tex_h, tex_w, pitch);
    scene_birthday::InitWorld<<<1, 1>>>(d_world, d_camera, tex);
  } else {
    fprintf(stderr, "unknown scene %s\n", scene.c_str());
    return 2;
  }
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  CK(cudaGetLastError());
  CK(cudaEventElapsedTime(&build_ms, e0, e1));

  if (rays_path) {
    auto buf = read_file(rays_path);
    long nr = (long)(buf.size() / 24);
    float *d_rays;
    RefHit *d_hits;
    CK(cudaMalloc(&d_rays, buf.size()));
    CK(cudaMalloc(&d_hits, sizeof(RefHit) * nr));
    CK(cudaMemcpy(d_rays, buf.data(), buf.size(), cudaMemcpyHostToDevice));
    TraceRays<<<(unsigned)((nr + 63) / 64), 64>>>(d_world, d_rays, nr, 1e-3, INFINITY, d_hits);
    CK(cudaDeviceSynchronize());
    std::vector<RefHit> hits(nr);
    CK(cudaMemcpy(hits.data(), d_hits, sizeof(RefHit) * nr, cudaMemcpyDeviceToHost));
    std::ofstream f(hits_path, std::ios::binary);
    f.write((const char *)hits.data(), sizeof(RefHit) * nr);
  }

  std::vector<float> times;
  for (int r = 0; r < reps; r++) {
    dim3 block(8, 8);
    dim3 grid((H + block.x - 1) / block.x, (W + block.y - 1) / block.y);
    CK(cudaEventRecord(e0));
    PathTracing<<<grid, block>>>(d_world, d_camera, H, W, spp, post != 0, d_states, d_image);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    CK(cudaGetLastError());
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    times.push_back(ms);
  }
  if (image_path && reps > 0) {
    std::vector<glm::vec3> img(n);
    CK(cudaMemcpy(img.data(), d_image, sizeof(glm::vec3) * n, cudaMemcpyDeviceToHost));
    std::ofstream f(image_path, std::ios::binary);
    f.write((const char *)img.data(), sizeof(glm::vec3) * n);
  }
  float best = 1e30f, sum = 0;
  for (float t : times) { best = t < best ? t : best; sum += t; }
  printf("{\"impl\": \"reference-cuda\", \"scene\": \"%s\", \"height\": %d, \"width\": %d, \"spp\": %d, \"reps\": %d, "
         "\"ms_best\": %.4f, \"ms_mean\": %.4f, \"msamples_per_s\": %.4f, \"init_world_ms\": %.3f}\n",
         scene.c_str(), H, W, spp, reps, reps ? best : 0.f, reps ? sum / reps : 0.f,
         reps ? (double)n * spp / best / 1e3 : 0.0, build_ms);
  return 0;
}
