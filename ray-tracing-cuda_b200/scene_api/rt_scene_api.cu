// rt_scene_api.cu — drivers of the scene API: Main / DistributedMain (utils.cu:132-242), the flattening of
// the device-resident object list into the C-ABI's scene description, and the small utilities of utils.cu
// (CudaRandomInit :43-47, WriteImage :87-96, GatherImageData :115-130, BaseName/ParentPath :26-41).
//
// All rendering goes through librtcore.so (include/rtcore.h); nothing here traces a ray.
#include "rt_scene_api.cuh"

#include <fcntl.h>
#include <unistd.h>

#include <algorithm>
#include <cctype>
#include <cerrno>
#include <sstream>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <thread>

#include "../../include/rtcore.h"

#define RTAPI_CUDA(call)                                                     \
  do {                                                                       \
    cudaError_t e_ = (call);                                                 \
    CHECK(e_ == cudaSuccess) << #call << ": " << cudaGetErrorString(e_);     \
  } while (0)
#define RTAPI_RT(call)                                                       \
  do {                                                                       \
    int rc_ = (call);                                                        \
    CHECK(rc_ == RT_OK) << #call << ": " << rt_last_error_string();          \
  } while (0)

// ---- small utilities ---------------------------------------------------------------------------------
std::string BaseName(const std::string &path) {
  size_t k = path.find_last_of("/\\");
  return k == std::string::npos ? path : path.substr(k + 1);
}

std::string ParentPath(const std::string &path) {
  // the reference ignores a trailing separator and returns "" when there is no parent (utils.cu:33-41)
  if (path.size() < 2) return "";
  size_t k = path.find_last_of("/\\", path.size() - 2);
  return k == std::string::npos ? "" : path.substr(0, k);
}

__global__ void CudaRandomInit(uint64_t seed, curandState *state, int n) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < n) curand_init(seed, idx, 0, &state[idx]);
}

cudaTextureObject_t ImageTexture::CreateCudaTextureObj(uint8_t *dev_buffer, int height, int width,
                                                       uint64_t pitch_in_bytes) {
  cudaResourceDesc res;
  memset(&res, 0, sizeof(res));
  res.resType = cudaResourceTypePitch2D;
  res.res.pitch2D.devPtr = dev_buffer;
  res.res.pitch2D.desc = cudaCreateChannelDesc<uchar4>();
  res.res.pitch2D.width = width;
  res.res.pitch2D.height = height;
  res.res.pitch2D.pitchInBytes = pitch_in_bytes;
  cudaTextureDesc tex;
  memset(&tex, 0, sizeof(tex));  // point filter, wrap addressing, as the zeroed descriptor of the reference
  tex.readMode = cudaReadModeNormalizedFloat;
  tex.normalizedCoords = 1;
  cudaTextureObject_t obj = 0;
  RTAPI_CUDA(cudaCreateTextureObject(&obj, &res, &tex, nullptr));
  return obj;
}

void WriteImage(const std::vector<glm::vec3> &pixels, int height, int width, const std::string &path) {
  std::vector<uint8_t> rgb8((size_t)height * width * 3);
  for (size_t i = 0; i < (size_t)height * width; i++)
    for (int c = 0; c < 3; c++) rgb8[3 * i + c] = (uint8_t)(pixels[i][c] * 255);  // utils.cu:90-93: truncation
  LOG(INFO) << "Writing image to \"" << path << "\"";
  bool ppm = path.size() > 4 && path.compare(path.size() - 4, 4, ".ppm") == 0;
  RTAPI_RT(ppm ? rt_write_ppm(path.c_str(), width, height, rgb8.data())
               : rt_write_jpeg(path.c_str(), width, height, rgb8.data(), 100));
}

// ---- host side: commit the flattened world to the core -------------------------------------------------
namespace {

struct Options {
  int height, width, spp, max_depth;
  uint64_t seed;
  std::string out, stats;
};

int EnvInt(const char *name, int dflt) {
  const char *v = getenv(name);
  return v && *v ? atoi(v) : dflt;
}

Options ReadOptions(int height, int width, int spp) {
  Options o;
  o.height = EnvInt("RT_HEIGHT", height);
  o.width = EnvInt("RT_WIDTH", width);
  o.spp = EnvInt("RT_SPP", spp);
  o.max_depth = EnvInt("RT_MAX_DEPTH", 10);  // TRACE_DEPTH_LIMIT, ray_tracing.cu:10
  const char *s = getenv("RT_SEED");
  o.seed = s && *s ? strtoull(s, nullptr, 10) : 1;
  const char *out = getenv("RT_OUT");
  o.out = out && *out ? out : "image.jpeg";  // utils.cu:178
  const char *st = getenv("RT_STATS");
  o.stats = st ? st : "";
  return o;
}

// The device-resident object list -> the core (rt_scene_from_world walks the tagged records of include/rtworld.h, whose
// layout the classes of rt_scene_api.cuh are static_asserted against).  Returns the build statistics.
rt_build_stats CommitWorld(rt_ctx *ctx, HitableList *d_world, Camera *d_camera) {
  rt_build_stats bs;
  RTAPI_RT(rt_scene_from_world(ctx, d_world, d_camera, EnvInt("RT_CAMERA_DISK", 0) ? 2 : 1, &bs));
  return bs;
}

void WriteStats(const Options &o, int rank, int world_size, const rt_build_stats &bs, const rt_render_stats &rs,
                int local_spp, double total_ms) {
  if (o.stats.empty()) return;
  std::string path = o.stats;
  if (world_size > 1) path += "." + std::to_string(rank);
  std::ofstream f(path);
  double samples = (double)o.height * o.width * local_spp;
  f << "{\"rank\": " << rank << ", \"world_size\": " << world_size << ", \"height\": " << o.height
    << ", \"width\": " << o.width << ", \"spp\": " << o.spp << ", \"local_spp\": " << local_spp
    << ", \"max_depth\": " << o.max_depth << ", \"render_ms\": " << rs.render_ms << ", \"total_ms\": " << total_ms
    << ", \"msamples_per_s\": " << (rs.render_ms > 0 ? samples / rs.render_ms / 1e3 : 0.0) << ", \"rays\": " << rs.rays
    << ", \"rays_ref_equiv\": " << rs.rays_ref_equiv << ", \"build_ms\": " << bs.build_ms << ", \"prims\": " << bs.n_prims
    << ", \"nodes\": " << bs.n_nodes << "}\n";
}

// NCCL rendezvous between the ranks of one box: rank 0 publishes the ncclUniqueId in /dev/shm.
// The file name carries a per-run nonce, so that a file left behind by a run that died can never be taken for this
// run's: RT_JOB_ID (rtrun: its pid + start time), else the launcher's pid (the ranks' common parent: torchrun,
// mpirun) + that process's start time from /proc (a reused pid has another start time) + MASTER_PORT if set.
// The record is {magic, world size, id}; readers check all of it.
struct RendezvousRecord {
  char magic[8];
  int32_t world_size, reserved;
  char id[128];
};
std::string RendezvousPath() {
  const char *job = getenv("RT_JOB_ID");
  std::string key;
  if (job && *job) {
    key = job;
  } else {
    const long ppid = (long)getppid();
    key = "p" + std::to_string(ppid);
    std::ifstream st("/proc/" + std::to_string(ppid) + "/stat");
    std::string line;
    if (std::getline(st, line)) {  // field 22 (starttime) counted after the ")" that ends the command name
      size_t k = line.rfind(')');
      std::istringstream is(k == std::string::npos ? std::string() : line.substr(k + 1));
      std::string f;
      for (int i = 0; i < 20 && (is >> f); i++) {}
      if (!f.empty()) key += "-" + f;
    }
    const char *port = getenv("MASTER_PORT");
    if (port && *port) key += "-" + std::string(port);
  }
  for (char &c : key)
    if (!(isalnum((unsigned char)c) || c == '-' || c == '_')) c = '_';
  return "/dev/shm/rtcore_nccl_" + std::to_string((long)getuid()) + "_" + key;
}

void BindNccl(rt_ctx *ctx, int rank, int world_size) {
  RendezvousRecord rec;
  memset(&rec, 0, sizeof(rec));
  const std::string path = RendezvousPath();
  if (rank == 0) {
    memcpy(rec.magic, "RTNCCL01", 8);
    rec.world_size = world_size;
    RTAPI_RT(rt_nccl_unique_id(rec.id));
    const std::string tmp = path + ".tmp";
    unlink(path.c_str());  // nothing of an earlier run survives under this name
    unlink(tmp.c_str());
    int fd = open(tmp.c_str(), O_CREAT | O_EXCL | O_WRONLY, 0600);
    CHECK(fd >= 0) << "cannot create " << tmp << ": " << strerror(errno);
    CHECK(write(fd, &rec, sizeof(rec)) == (ssize_t)sizeof(rec)) << "short write to " << tmp;
    close(fd);
    CHECK(rename(tmp.c_str(), path.c_str()) == 0) << "cannot publish " << path;
  } else {
    bool ok = false;
    for (int tries = 0; tries < 6000 && !ok; tries++) {
      int fd = open(path.c_str(), O_RDONLY);
      if (fd >= 0) {
        ok = read(fd, &rec, sizeof(rec)) == (ssize_t)sizeof(rec) && memcmp(rec.magic, "RTNCCL01", 8) == 0 &&
             rec.world_size == world_size;
        close(fd);
      }
      if (!ok) std::this_thread::sleep_for(std::chrono::milliseconds(10));
    }
    CHECK(ok) << "rank " << rank << ": no ncclUniqueId for " << world_size << " ranks at " << path << " after 60 s";
  }
  RTAPI_RT(rt_nccl_init(ctx, rec.id, world_size, rank));
  if (rank == 0) unlink(path.c_str());  // every rank has joined once ncclCommInitRank returns
}

// Peer frames (rtcore.h, rt_peer_frame_*): every rank publishes the 64-byte CUDA IPC handle of its frame in its own
// file next to the NCCL one and reads the others'.  Returns the files this rank wrote (removed once the first reduce
// has shown that every rank has opened every frame).
struct PeerRecord {
  char magic[8];
  int32_t world_size, rank;
  char handle[RT_PEER_HANDLE_BYTES];
};
std::string OpenPeerFrames(rt_ctx *ctx, int rank, int world_size, int height, int width) {
  std::vector<PeerRecord> recs(world_size);
  memset(recs.data(), 0, sizeof(PeerRecord) * world_size);
  PeerRecord &mine = recs[rank];
  memcpy(mine.magic, "RTPEER01", 8);
  mine.world_size = world_size;
  mine.rank = rank;
  RTAPI_RT(rt_peer_frame_create(ctx, height, width, mine.handle));
  const std::string base = RendezvousPath() + ".frame";
  const std::string path = base + std::to_string(rank), tmp = path + ".tmp";
  unlink(path.c_str());
  unlink(tmp.c_str());
  int fd = open(tmp.c_str(), O_CREAT | O_EXCL | O_WRONLY, 0600);
  CHECK(fd >= 0) << "cannot create " << tmp << ": " << strerror(errno);
  CHECK(write(fd, &mine, sizeof(mine)) == (ssize_t)sizeof(mine)) << "short write to " << tmp;
  close(fd);
  CHECK(rename(tmp.c_str(), path.c_str()) == 0) << "cannot publish " << path;
  for (int r = 0; r < world_size; r++) {
    if (r == rank) continue;
    const std::string other = base + std::to_string(r);
    bool ok = false;
    for (int tries = 0; tries < 6000 && !ok; tries++) {
      int f = open(other.c_str(), O_RDONLY);
      if (f >= 0) {
        ok = read(f, &recs[r], sizeof(PeerRecord)) == (ssize_t)sizeof(PeerRecord) &&
             memcmp(recs[r].magic, "RTPEER01", 8) == 0 && recs[r].world_size == world_size && recs[r].rank == r;
        close(f);
      }
      if (!ok) std::this_thread::sleep_for(std::chrono::milliseconds(10));
    }
    CHECK(ok) << "rank " << rank << ": no peer frame of rank " << r << " at " << other << " after 60 s";
  }
  std::vector<char> handles((size_t)world_size * RT_PEER_HANDLE_BYTES);
  for (int r = 0; r < world_size; r++) memcpy(&handles[(size_t)r * RT_PEER_HANDLE_BYTES], recs[r].handle, RT_PEER_HANDLE_BYTES);
  RTAPI_RT(rt_peer_frame_open(ctx, world_size, rank,
                              reinterpret_cast<const char(*)[RT_PEER_HANDLE_BYTES]>(handles.data())));
  return path;
}

// The part of Main / DistributedMain after init_world: flatten, build, render, (reduce), resolve, write.
// Both drivers honour the launcher (rtrun -n N, torchrun, mpirun): with N ranks each renders its GetWorkload share of
// the samples (utils.cu:111-113), the exact sums are added into rank 0's frame by one kernel per rank over NVLink peer
// memory (rt_reduce_resolve_peers; RT_REDUCE=nccl: one ncclReduce instead) and rank 0 writes the image — the same
// bits as one rank renders alone.  The reference's Main is single-GPU only (utils.cu:132-179); `mpi` tells
// whether the scene itself initialised MPI and bound its device (spheres.cu:83-98) or Main has to.
void RenderWorld(curandState **d_states, Camera **d_camera, HitableList **d_world, glm::vec3 **d_image,
                 nvstd::function<void(HitableList *, Camera *)> init_world, int height, int width, int spp,
                 uint64_t curand_seed, bool distributed) {
  auto t_start = std::chrono::steady_clock::now();
  Options o = ReadOptions(height, width, spp);
  int rank = 0, world_size = 1;
  MPI_Comm_size(MPI_COMM_WORLD, &world_size);
  MPI_Comm_rank(MPI_COMM_WORLD, &rank);
  if (!distributed && world_size > 1) {  // Main under a launcher: bind rank -> device as spheres.cu:90-93 does
    int count = 1;
    RTAPI_CUDA(cudaGetDeviceCount(&count));
    RTAPI_CUDA(cudaSetDevice(rank % count));
  }
  int device = 0;
  RTAPI_CUDA(cudaGetDevice(&device));
  const size_t n = (size_t)o.height * o.width;

  // utils.cu:139-146 / :194-202: the caller's four device pointers, states seeded before init_world runs
  RTAPI_CUDA(cudaMalloc(d_states, sizeof(curandState) * n));
  RTAPI_CUDA(cudaMalloc(d_image, sizeof(glm::vec3) * n));
  RTAPI_CUDA(cudaMalloc(d_world, sizeof(HitableList)));
  RTAPI_CUDA(cudaMalloc(d_camera, sizeof(Camera)));
  LOG(INFO) << "cuda memory allocated";
  CudaRandomInit<<<(unsigned)((n + 63) / 64), 64>>>(curand_seed, *d_states, (int)n);
  LOG(INFO) << "cuda random initialized";
  init_world(*d_world, *d_camera);
  RTAPI_CUDA(cudaDeviceSynchronize());
  RTAPI_CUDA(cudaGetLastError());
  LOG(INFO) << "world initialized";

  rt_ctx *ctx = nullptr;
  RTAPI_RT(rt_ctx_create(device, &ctx));
  rt_build_stats bs = CommitWorld(ctx, *d_world, *d_camera);
  LOG(INFO) << "scene committed: " << bs.n_prims << " primitives, " << bs.n_nodes << " BVH nodes, build " << bs.build_ms
            << "ms";

  int spp_begin = 0;
  int local_spp = rt_get_workload(rank, world_size, o.spp, &spp_begin);
  if (distributed || world_size > 1) LOG(INFO) << "[" << rank << " / " << world_size << "] workload: " << local_spp;
  const char *reduce_env = getenv("RT_REDUCE");
  const bool peers = world_size > 1 && !(reduce_env && strcmp(reduce_env, "nccl") == 0);
  std::string peer_file;
  if (world_size > 1 && !peers) BindNccl(ctx, rank, world_size);

  // the accumulation buffer: exact 64-bit sums (rtcore.h, summation contract); inside this rank's peer frame when
  // the ranks reduce over NVLink
  int64_t *d_sums = nullptr;
  if (peers) {
    peer_file = OpenPeerFrames(ctx, rank, world_size, o.height, o.width);
    RTAPI_RT(rt_peer_frame_buffers(ctx, &d_sums, nullptr, nullptr));
  } else {
    RTAPI_CUDA(cudaMalloc(&d_sums, n * 3 * sizeof(int64_t)));
  }
  rt_render_params p;
  p.height = o.height;
  p.width = o.width;
  p.spp_begin = spp_begin;
  p.spp_count = local_spp;
  p.max_depth = o.max_depth;
  p.flags = RT_RENDER_CLEAR;
  p.seed = o.seed;
  rt_render_stats rs;
  RTAPI_RT(rt_render_fixed(ctx, &p, d_sums, &rs, nullptr));
  if (distributed || world_size > 1)
    LOG(INFO) << "[" << rank << " / " << world_size << "] Ray tracing finished in " << rs.render_ms << "ms.";
  else
    LOG(INFO) << "Ray tracing finished in " << rs.render_ms << "ms.";

  // utils.cu:124 (MPI_Reduce of host copies) -> the ranks' sums added into rank 0's frame over NVLink
  if (peers) {
    RTAPI_RT(rt_reduce_resolve_peers(ctx, std::max(o.spp, 1), 0, nullptr));
    RTAPI_RT(rt_peer_frame_check(ctx, nullptr));
    unlink(peer_file.c_str());  // every rank has opened every frame by now
  } else if (world_size > 1) {
    RTAPI_RT(rt_reduce_fixed(ctx, nullptr, d_sums, (int64_t)n * 3, 0, nullptr));
  }
  float *image = reinterpret_cast<float *>(*d_image);
  if (rank == 0) {
    // ray_tracing.cu:78-83 / utils.cu:126-129 on the device, then WriteImage's *255 truncation.  *d_image is left as
    // the reference leaves it: post-processed by Main (post_processing = true, utils.cu:162), raw sums by
    // DistributedMain (utils.cu:220)
    uint8_t *d_rgb8 = nullptr;
    RTAPI_CUDA(cudaMalloc(&d_rgb8, n * 3));
    RTAPI_RT(rt_resolve_fixed(ctx, d_sums, o.height, o.width, std::max(o.spp, 1), distributed ? image : nullptr,
                              distributed ? nullptr : image, d_rgb8, nullptr));
    std::vector<uint8_t> rgb8(n * 3);
    RTAPI_CUDA(cudaMemcpy(rgb8.data(), d_rgb8, n * 3, cudaMemcpyDeviceToHost));
    RTAPI_CUDA(cudaFree(d_rgb8));
    LOG(INFO) << "Writing image to \"" << o.out << "\"";
    bool ppm = o.out.size() > 4 && o.out.compare(o.out.size() - 4, 4, ".ppm") == 0;
    RTAPI_RT(ppm ? rt_write_ppm(o.out.c_str(), o.width, o.height, rgb8.data())
                 : rt_write_jpeg(o.out.c_str(), o.width, o.height, rgb8.data(), 100));
  } else {
    RTAPI_RT(rt_resolve_fixed(ctx, d_sums, o.height, o.width, std::max(o.spp, 1), image, nullptr, nullptr, nullptr));
  }
  RTAPI_CUDA(cudaDeviceSynchronize());
  if (!peers) RTAPI_CUDA(cudaFree(d_sums));
  double total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_start).count();
  WriteStats(o, rank, world_size, bs, rs, local_spp, total_ms);
  if (world_size > 1 && !peers) rt_nccl_destroy(ctx);
  // (peer frames: rt_peer_frame_check returned only after every rank had stored its slice, so nobody reads a frame
  // another rank is about to free)
  RTAPI_RT(rt_ctx_destroy(ctx));
}

}  // namespace

// utils.cu:115-130: SUM of the ranks' images at rank 0, then /spp, clamp, sqrt.  Host-vector form kept for API
// compatibility (the drivers above reduce on the device instead); only the single-rank case is supported here.
__host__ void GatherImageData(int height, int width, const std::vector<glm::vec3> &send_buf, int spp,
                              std::vector<glm::vec3> *out_image) {
  int world_size = 1;
  MPI_Comm_size(MPI_COMM_WORLD, &world_size);
  CHECK(world_size == 1) << "GatherImageData on host vectors is single-rank; DistributedMain reduces on the device";
  out_image->resize((size_t)height * width);
  for (size_t i = 0; i < out_image->size(); i++) {
    glm::vec3 c = send_buf[i] / float(spp);
    c = glm::clamp(c, 0.f, 1.f);
    (*out_image)[i] = glm::sqrt(c);
  }
}

__host__ void Main(curandState **d_states, Camera **d_camera, HitableList **d_world, glm::vec3 **d_image,
                   nvstd::function<void(HitableList *world, Camera *camera)> init_world, int height, int width,
                   int spp) {
  RenderWorld(d_states, d_camera, d_world, d_image, init_world, height, width, spp, 1024 /* utils.cu:146 */, false);
}

__host__ void DistributedMain(curandState **d_states, Camera **d_camera, HitableList **d_world, glm::vec3 **d_image,
                              nvstd::function<void(HitableList *world, Camera *camera)> init_world, int height,
                              int width, int spp) {
  RenderWorld(d_states, d_camera, d_world, d_image, init_world, height, width, spp, 10086 /* utils.cu:202 */, true);
}
