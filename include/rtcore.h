/*
 * rtcore.h — C-ABI of the B200-native path-tracing core (librtcore.so).
 *
 * This is the drop-in boundary for the render hot path of tigert1998/ray-tracing-cuda.
 * Every entry point cites the reference interface it replaces (paths relative to the
 * reference tree).  Plain pointers and sizes only; no C++/torch types.
 *
 * Conventions
 *   - every function returns RT_OK (0) or a negative RT_ERR_* code; rt_last_error_string()
 *     gives the message of the last failure on the calling thread.  The library never
 *     aborts (the reference CHECK-aborts: utils.cu:143-152); the C++ `Main` wrapper CHECKs.
 *   - a context is bound to one CUDA device; one context per device; not thread-safe.
 *   - "h_" pointers are host memory, "d_" pointers are device memory of the context's device.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *   - there is NO CPU fallback: every compute entry point fails with RT_ERR_CUDA when no
 *     CUDA device is usable.
 */
#ifndef RTCORE_H_
#define RTCORE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_ABI_VERSION 4

enum {
  RT_OK = 0,
  RT_ERR_INVALID = -1, /* bad argument / call order */
  RT_ERR_CUDA = -2,    /* CUDA runtime failure (message has cudaGetErrorString) */
  RT_ERR_NOMEM = -3,
  RT_ERR_LIMIT = -4,   /* exceeds a documented limit (e.g. HitableList::kMaxHitables) */
  RT_ERR_NCCL = -5
};

/* ---- scene description (type-tagged records replacing the virtual object graph) ---------- */

/* Material tags.  Replaces the virtual Material hierarchy (material.cuh:13-20). */
enum {
  RT_MAT_LAMBERTIAN = 0,    /* lambertian.cuh:11-29   color = albedo, or texture>=0          */
  RT_MAT_METAL = 1,         /* metal.cuh:5-19         color = albedo, fuzz (clamped to <=1)   */
  RT_MAT_DIELECTRIC = 2,    /* dielectric.cuh:5-17    color = attenuation, ior                */
  RT_MAT_DIFFUSE_LIGHT = 3, /* diffuse_light.cuh:6-18 color = emission, or texture>=0         */
  RT_MAT_SKY = 4            /* sky.cuh:8-15           gradient miss shader                    */
};

typedef struct rt_material {
  int32_t type;
  int32_t texture; /* -1: ConstantTexture(color) (constant_texture.cuh); >=0: image texture index */
  float color[3];
  float fuzz;
  double ior;
} rt_material; /* 32 bytes */

/* Top-level hitable tags.  Replaces Hitable subclasses appended to HitableList
 * (hitable_list.cuh:8-22).  A Parallelepiped (parallelepiped.cuh:16-29) is described as its
 * six Parallelograms in AddCorner order (parallelepiped.cu:25-32) sharing one `group` id. */
enum {
  RT_HIT_SPHERE = 0,        /* sphere.cuh:8-25        p[0]=centre, radius                      */
  RT_HIT_TRIANGLE = 1,      /* triangle.cuh:7-17      p[0..2]                                  */
  RT_HIT_PARALLELOGRAM = 2, /* parallelogram.cuh:7-17 p[0..3], p[3]=p[1]+p[2]-p[0] (ctor value) */
  RT_HIT_SKY = 3,           /* sky.cuh:17-26          hit at t=1e9                             */
  RT_HIT_MESH = 4           /* bvh.cuh:161-183        BVH<Face<B>,AABB> over meshes[mesh]      */
};

typedef struct rt_hitable {
  int32_t type;
  int32_t material; /* index into materials; for RT_HIT_MESH the mesh-wide material (bvh.cuh:178) */
  float p[4][3];
  double radius;
  int32_t mesh;  /* RT_HIT_MESH: index into meshes, else -1 */
  int32_t group; /* >=0: id of the enclosing Parallelepiped; -1 otherwise (informational) */
} rt_hitable; /* 72 bytes */

/* A triangle mesh in the reference's Face<> layout (bvh.cuh:21-66): has_texcoord=0 -> 9 floats
 * per face (3 positions); has_texcoord=1 -> 15 floats (3 positions, then 3 uv pairs).
 * Faces are given in caller order; the library reproduces the reference build's in-place
 * stable sort by positions_[0].x for n_faces > 2048 (bvh.cuh:96-98,113-121), and face indices
 * reported by rt_trace_closest refer to that post-sort order. */
typedef struct rt_mesh {
  const float *h_faces;
  int32_t n_faces;
  int32_t has_texcoord;
} rt_mesh;

/* RGBA8 image sampled like the reference's cudaTextureObject_t (image_texture.cu:17-38):
 * normalized coords, point filter, wrap addressing, uchar4 -> float/255, v flipped by caller. */
typedef struct rt_texture {
  const uint8_t *h_rgba;
  int32_t height, width;
  int64_t pitch_bytes;
} rt_texture;

/* Camera frame: the seven vectors + lens fields the reference Camera holds (camera.cuh:12-14). */
typedef struct rt_camera {
  float position[3], lower_left_corner[3], horizontal[3], vertical[3], u[3], v[3], w[3];
  int32_t is_defocus; /* 0 pinhole; 1 thin lens sampled as the reference does (DiskRand, camera.cu:74-77: a SQUARE
                         (0, lens_radius]^2 in the positive quadrant); 2 thin lens sampled as a uniform disk of radius
                         lens_radius (what the name promises; scene API: RT_CAMERA_DISK=1) */
  double lens_radius;
} rt_camera; /* 96 bytes */

typedef struct rt_scene_desc {
  const rt_hitable *hitables;
  int32_t n_hitables; /* the reference's HitableList holds at most 1024 (hitable_list.cuh:10; the scene API's Append
                         enforces it); rt_scene_commit itself accepts any count whose records fit 2^29 */
  const rt_material *materials;
  int32_t n_materials;
  const rt_mesh *meshes;
  int32_t n_meshes;
  const rt_texture *textures;
  int32_t n_textures;
} rt_scene_desc;

/* ---- results ------------------------------------------------------------------------------ */

/* Closest-hit record of the parity hook.  `t` is the double the reference stores in
 * HitRecord::t (hitable.cuh:13-17).  obj = index in the top-level list (Append order, a
 * Parallelepiped counting as its 6 parallelograms); sub = 0/1 for the two triangles of a
 * parallelogram, the post-sort face index for a mesh, 0 otherwise.  obj = -1: no hit. */
typedef struct rt_hit {
  double t;
  int32_t obj;
  int32_t sub;
  float u, v; /* HitRecord::u,v narrowed to float (sky leaves them undefined: reported 0) */
  float normal[3];
  int32_t material;
} rt_hit; /* 40 bytes */

typedef struct rt_build_stats {
  int32_t n_prims;      /* leaf primitives in the unified BVH (spheres+triangles+parallelograms+faces) */
  int32_t n_nodes;      /* 32-byte nodes */
  int32_t n_spheres, n_quads, n_tris;
  int32_t max_depth;
  float build_ms;       /* device time of the BVH build (CUDA events) */
  float upload_ms;
  float sah_cost;       /* of the tree the traversal walks */
  int32_t reserved_[2]; /* large scenes build two trees and keep the one of lower SAH cost: [0] > 0: PLOC rounds, the
                           clustered tree was kept; < 0: it was built and rejected; 0: not tried.  [1] float bits of the
                           SAH cost of the tree that was NOT kept */
} rt_build_stats;

typedef struct rt_render_stats {
  uint64_t samples;       /* camera samples started */
  uint64_t rays;          /* closest-hit queries executed by this core */
  uint64_t rays_ref_equiv;/* queries the reference loop would issue for the same paths (ray_tracing.cu:22) */
  uint64_t node_visits;   /* box tests; only with RT_RENDER_PROFILE, else 0 */
  uint64_t prim_tests;    /* record tests; only with RT_RENDER_PROFILE, else 0 */
  uint32_t iterations;    /* wavefront iterations (0: the persistent path kernel) */
  uint32_t kernel_launches;
  float render_ms;        /* device time, CUDA events on `stream` */
  uint32_t chunk_samples; /* path kernel: samples per (pixel, chunk) sub-unit of the work distribution */
  uint32_t work_items;    /* path kernel: (tile, chunk group) items drawn from the global counter */
  uint32_t seq_fallbacks; /* path kernel: queries answered by the sequential scan (order-dependent ties the tree walk
                             could not settle); each costs one test per primitive of the scene */
  uint32_t max_box_tests_per_ray; /* only with RT_RENDER_PROFILE */
  uint32_t reserved_;
} rt_render_stats;

/* ---- entry points -------------------------------------------------------------------------- */

typedef struct rt_ctx rt_ctx;

int rt_abi_version(void);
const char *rt_last_error_string(void);

/* Replaces the cudaMalloc block at the top of Main / DistributedMain (utils.cu:139-142,194-197). */
int rt_ctx_create(int device, rt_ctx **out);
int rt_ctx_destroy(rt_ctx *ctx);

/* Replaces scene construction by device-side `new` + HitableList::Append (hitable_list.cu:27-29)
 * and the single-thread BVHNode build (bvh.cuh:113-121): copies the description to the device,
 * flattens it into SoA records and builds the unified LBVH on the device. */
int rt_scene_commit(rt_ctx *ctx, const rt_scene_desc *desc, rt_build_stats *stats /*nullable*/);

/* The same from the device-resident object list the scene API's constructors build (include/rtworld.h: tagged records
 * in device memory, a HitableList record at d_world, a Camera record at d_camera or NULL): walks the list on the
 * device, expands nested lists and parallelepipeds, deduplicates materials and image textures, commits the scene and
 * sets the camera.  This is what Main / DistributedMain run after init_world (utils.cu:148-150); a driver of one's own
 * calls it instead of describing the scene by hand.  lens_mode: rt_camera::is_defocus for a thin-lens camera (1 or 2). */
int rt_scene_from_world(rt_ctx *ctx, const void *d_world, const void *d_camera, int lens_mode,
                        rt_build_stats *stats /*nullable*/);

/* Replaces the Camera constructors (camera.cu:6-47). rt_camera_look_at computes the frame with
 * the reference's arithmetic on the host; aperture<0 selects the pinhole constructor. */
int rt_camera_set(rt_ctx *ctx, const rt_camera *cam);
int rt_camera_look_at(const float position[3], const float look_at[3], const float up[3],
                      double field_of_view, double width_height_aspect, double aperture,
                      double focus_distance, rt_camera *out);

/* Parity hook: replaces HitableList::Hit (hitable_list.cu:7-25) on explicit rays.
 * rays = n x 6 floats (position xyz, direction xyz); the direction is used as stored (callers
 * pass Ray::direction_, i.e. already normalized by Ray::Ray, ray.cu:8-11).
 * mode 0: the traversal the committed scene renders with (product path); mode 1: sequential scan over every
 * primitive in list order, as HitableList::Hit executes it; modes 2..5: one specific traversal of the render kernel
 * (2 binary tree if-if, 3 binary tree while-while, 4 box list — scenes of at most 32 primitives only).
 * All modes must return identical hits. */
int rt_trace_closest_host(rt_ctx *ctx, const float *h_rays, int64_t n, double t_from, double t_to,
                          int mode, rt_hit *h_hits);
int rt_trace_closest_device(rt_ctx *ctx, const float *d_rays, int64_t n, double t_from, double t_to,
                            int mode, rt_hit *d_hits, void *stream);

/* Pixel-centre primary rays, as PathTracing would form them with both jitters = 0.5
 * (ray_tracing.cu:68-74).  h_rays = height*width*6 floats, row-major like out_image. */
int rt_primary_rays_host(rt_ctx *ctx, int height, int width, float *h_rays);

typedef struct rt_render_params {
  int32_t height, width;
  int32_t spp_begin, spp_count; /* global sample indices [spp_begin, spp_begin+spp_count) per pixel */
  int32_t max_depth;            /* TRACE_DEPTH_LIMIT (ray_tracing.cu:10); reference value 10 */
  int32_t flags;                /* RT_RENDER_* */
  uint64_t seed;                /* Philox key */
} rt_render_params;

enum {
  RT_RENDER_DEFAULT = 0,
  RT_RENDER_CLEAR = 1,     /* zero the accumulation buffer first */
  RT_RENDER_PROFILE = 2,   /* count box / record tests (rt_render_stats.node_visits, prim_tests); slower */
  RT_RENDER_FUSED = 4,     /* force the persistent path kernel (paths regenerated in registers; the default) */
  RT_RENDER_SORT = 8,      /* wavefront: material-sorted shade queues (one Scatter/Emit body per warp) instead of one queue;
                              same sums, measured 6-9 % slower (profiles/r01_t_wavefront_sorted_queues.txt) */
  RT_RENDER_WAVEFRONT = 16, /* force the wavefront pipeline (logic/shade -> extend over compacted queues) */
  RT_RENDER_RAYSORT = 128   /* wavefront: reorder the extend queue by (direction octant, origin cell) before every extend
                               pass (counting sort on the device); same sums */
  /* 32, 64: RT_RENDER_SPLIT / RT_RENDER_ORDERED of ABI 2 (sample slices on side streams, heaviest tile first) are gone:
     the path kernel now draws (tile, sample chunk) work items from a global counter; the bits are ignored. */
};

/* Sampling contract (what makes an image reproducible): Philox4x32-10 keyed by `seed`; word w of the random stream
 * of (pixel, global sample index) is word (w & 3) of block counter (pixel, sample, w >> 2, 0).  Words 0,1: pixel
 * jitter (ray_tracing.cu:68-71); 2,3: lens sample (thin-lens cameras only, camera.cu:62-65); B+2k, B+2k+1 with B = 4
 * (lens) or 2 (pinhole): the direction sample of the scatter at depth k — a uniform direction drawn without a
 * rejection loop (same distribution as lambertian.cu:19-31 / metal.cu:27-36); Metal fuzz takes its radius from word 0
 * of block (pixel, sample, 0x40000000 + k, 0).  uint32 -> (0, 1] as curand_uniform.
 *
 * Summation contract: the radiance of a sample is rounded once to a multiple of 2^-30 (|component| clamped to 2^32,
 * NaN = 0) and the samples of a pixel are added as 64-bit integers.  Integer addition is exact, so a pixel's sum does
 * not depend on the order or grouping of its samples: any split of the sample range over rt_render_fixed calls,
 * work chunks or GPUs (rt_reduce_fixed) gives the same bits. */

/* Replaces the PathTracing<<<>>> launch with post_processing=false (ray_tracing.cu:56-85,
 * utils.cu:220-221): adds the radiance SUM of the requested samples (the fixed-point sum of this call, rounded to
 * float once) into d_accum_rgb (height*width*3 floats, the reference's glm::vec3 image layout). */
int rt_render(rt_ctx *ctx, const rt_render_params *params, float *d_accum_rgb,
              rt_render_stats *stats /*nullable*/, void *stream);

/* The same launch accumulating into caller-owned exact sums: d_accum_fixed = height*width*3 int64, units of 2^-30
 * (RT_RENDER_CLEAR zeroes it first).  This is the accumulation buffer Main / DistributedMain keep (utils.cu:140). */
int rt_render_fixed(rt_ctx *ctx, const rt_render_params *params, int64_t *d_accum_fixed,
                    rt_render_stats *stats /*nullable*/, void *stream);

/* rt_render for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments and result as rt_render (the same bits in d_accum_rgb), but the call's scratch — the exact int64
 * sums and the kernel's counters — is a fixed device array inside the library instead of context buffers, so the
 * call performs no cudaMalloc / cudaFree and synchronises nothing.  Limits: height * width <= RT_STATIC_MAX_PIXELS
 * (RT_ERR_LIMIT beyond it); path kernel only (RT_RENDER_WAVEFRONT / SORT / RAYSORT / PROFILE: RT_ERR_INVALID); no
 * stats.  The array (6 MiB) is held by every GPU that loads the library until exit and is shared by ALL calls and
 * contexts of a device: it is cleared at the start of every call, and two calls that overlap on the GPU (different
 * streams, or two contexts of one device) race — serialise them.  Scene and camera are committed beforehand (those
 * calls allocate). */
#define RT_STATIC_MAX_PIXELS 262144 /* e.g. 512 x 512, 640 x 400 */
int rt_render_static(rt_ctx *ctx, const rt_render_params *params, float *d_accum_rgb, void *stream);

/* Replaces the post-processing tail of PathTracing (ray_tracing.cu:78-83) and the host loops in
 * GatherImageData / WriteImage (utils.cu:126-129, 87-96): out = sqrt(clamp(accum/spp_total,0,1));
 * either output may be NULL.  d_rgb8 gets trunc(out*255) like WriteImage. */
int rt_resolve(rt_ctx *ctx, const float *d_accum_rgb, int height, int width, int spp_total,
               float *d_rgb /*nullable*/, uint8_t *d_rgb8 /*nullable*/, void *stream);

/* The same from the exact sums; additionally writes the float sums (the reference's pre-post-processing image)
 * when d_accum_rgb is not NULL. */
int rt_resolve_fixed(rt_ctx *ctx, const int64_t *d_accum_fixed, int height, int width, int spp_total,
                     float *d_accum_rgb /*nullable*/, float *d_rgb /*nullable*/, uint8_t *d_rgb8 /*nullable*/,
                     void *stream);

/* End-to-end convenience through host buffers: commit-independent render + resolve + D2H.
 * h_accum_rgb (nullable) receives the raw sums, h_rgb (nullable) the post-processed floats. */
int rt_render_host(rt_ctx *ctx, const rt_render_params *params, float *h_accum_rgb, float *h_rgb,
                   rt_render_stats *stats /*nullable*/);

/* Replaces GetWorkload (utils.cu:111-113): samples of rank `rank`, and its first global index. */
int rt_get_workload(int rank, int world_size, int spp, int *spp_begin /*nullable*/);

/* The work plan of a path-kernel launch as data (no device needed): how the samples
 * [spp_begin, spp_begin + spp_count) of a height x width frame are cut into phases of (chunk, chunks per item) for a grid
 * of n_sm x blocks_per_sm blocks.  Replaces, as far as anything does, the fixed pixel -> thread map of the PathTracing
 * launch (utils.cu:158-162, ray_tracing.cu:61-64); exists for diagnostics and the CPU tests of the plan arithmetic.
 * phases = 5 x 4 int32: {first item, first sample, samples per chunk, chunks per item} per phase, closed by
 * {items of the launch, spp_begin + spp_count, 1, 1}.  Honours RT_PLAN / RT_CHUNK / RT_ITEM_CHUNKS like rt_render.
 * Returns the number of phases (1..4) or a negative RT_ERR_*. */
int rt_work_plan(int n_sm, int blocks_per_sm, int height, int width, int spp_begin, int spp_count,
                 int bounded_cost, int32_t *phases, int32_t *n_tiles /*nullable*/);

/* Replaces D2H + MPI_Reduce in GatherImageData (utils.cu:115-125,232-238): in-place SUM of the
 * accumulation buffers of all ranks to `root` with one ncclReduce on `stream`.
 * `nccl_comm` is an ncclComm_t passed as void*. */
int rt_reduce(rt_ctx *ctx, void *nccl_comm, float *d_accum_rgb, int64_t n_floats, int root,
              void *stream);

/* The same reduce on the exact sums (ncclInt64, SUM): the image of N ranks equals the image of one, bit for bit. */
int rt_reduce_fixed(rt_ctx *ctx, void *nccl_comm, int64_t *d_accum_fixed, int64_t n_values, int root,
                    void *stream);

/* Communicator helpers for callers that have no ncclComm_t of their own (the reference ranks talk
 * through MPI_COMM_WORLD, utils.cu:118-125): rank 0 creates a 128-byte ncclUniqueId, the caller
 * ships it to the other ranks with whatever plumbing it has (torch.distributed, a file, MPI), and
 * every rank binds a communicator to its context.  rt_reduce(ctx, NULL, ...) then uses it. */
int rt_nccl_unique_id(char out_id[128]);
int rt_nccl_init(rt_ctx *ctx, const char id[128], int world_size, int rank);
int rt_nccl_destroy(rt_ctx *ctx);

/* ---- The reduce as one kernel over NVLink peer memory (no NCCL call) -------------------------------------------
 * Replaces D2H + MPI_Reduce + the host post-processing loop of GatherImageData (utils.cu:115-130, 232-241) for
 * ranks that are processes of ONE node.  Each rank owns a "peer frame" (exact sums | float sums | resolved image)
 * that the other ranks map through CUDA IPC; after its render, rank r sums slice r of the image over the frames of
 * all ranks, resolves it and stores it into the root's frame.  The root's image is the one-GPU image bit for bit.
 *
 *   rank r:  rt_peer_frame_create(ctx, H, W, handle_r)          -> 64-byte handle, shipped to the other ranks by the
 *            rt_peer_frame_open(ctx, world, r, handles)            caller's own plumbing (file, MPI, torch.distributed)
 *            rt_peer_frame_buffers(ctx, &sums, &accum, &rgb)
 *   frame:   rt_render_fixed(ctx, &params(RT_RENDER_CLEAR), sums, ...)      every rank, its own sample range
 *            rt_reduce_resolve_peers(ctx, spp_total, root, stream)           every rank, same stream as the render
 *            root: accum / rgb of ITS frame now hold the whole image (stream-ordered)
 * Every rank must call rt_reduce_resolve_peers the same number of times.  A rank that does not arrive within
 * RT_PEER_TIMEOUT_MS (environment, default 5000) makes rt_peer_frame_check return RT_ERR_CUDA on the ranks that
 * waited for it; nothing hangs.  world_size 1 is allowed (no handle is opened; the call is a plain resolve). */
#define RT_PEER_HANDLE_BYTES 64
int rt_peer_frame_create(rt_ctx *ctx, int height, int width, char out_handle[RT_PEER_HANDLE_BYTES]);
int rt_peer_frame_open(rt_ctx *ctx, int world_size, int rank, const char (*handles)[RT_PEER_HANDLE_BYTES]);
int rt_peer_frame_buffers(rt_ctx *ctx, int64_t **d_accum_fixed /*nullable*/, float **d_accum_rgb /*nullable*/,
                          float **d_rgb /*nullable*/);
int rt_reduce_resolve_peers(rt_ctx *ctx, int spp_total, int root, void *stream);
/* Synchronises `stream`; RT_ERR_CUDA if a wait of rt_reduce_resolve_peers timed out since the last check. */
int rt_peer_frame_check(rt_ctx *ctx, void *stream);
int rt_peer_frame_destroy(rt_ctx *ctx);

/* Image writers replacing stbi_write_jpg in WriteImage (utils.cu:87-96). */
int rt_write_jpeg(const char *path, int width, int height, const uint8_t *h_rgb8, int quality);
int rt_write_ppm(const char *path, int width, int height, const uint8_t *h_rgb8);
int rt_write_pfm(const char *path, int width, int height, const float *h_rgb);

#ifdef __cplusplus
}
#endif
#endif /* RTCORE_H_ */
