// Test-infrastructure shim, force-included (-include) when the reference's CUDA sources are
// compiled with g++ for the host (oracle/_ref).  It supplies the handful of device-only names
// the reference uses (thread indices, min/max, tex2D) so that its *unmodified* arithmetic runs
// on CPU cores.  Nothing here is product code and nothing here restates reference logic.
#pragma once
#include <cuda_runtime.h>
#include <curand_kernel.h>
#include <thrust/execution_policy.h>
#include <thrust/sort.h>
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>

namespace rt_shim {
struct Idx3 { unsigned x, y, z; };
extern thread_local Idx3 t_threadIdx, t_blockIdx, t_blockDim, t_gridDim;

// Host model of a cudaTextureObject_t created the way the reference creates it
// (pitch2D, uchar4, normalized float read, normalized coords, point filter, wrap mode 0 = wrap).
struct HostTexture { const uint8_t *data; int height, width; uint64_t pitch; };
float4 Tex2DPointWrap(cudaTextureObject_t tex, float u, float v);
}  // namespace rt_shim

#define threadIdx ::rt_shim::t_threadIdx
#define blockIdx ::rt_shim::t_blockIdx
#define blockDim ::rt_shim::t_blockDim
#define gridDim ::rt_shim::t_gridDim

// CUDA exposes ::min/::max overloads in device code; give the host the same spellings.
static inline double min(double a, double b) { return std::fmin(a, b); }
static inline double max(double a, double b) { return std::fmax(a, b); }
static inline float min(float a, float b) { return std::fmin(a, b); }
static inline float max(float a, float b) { return std::fmax(a, b); }
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }

template <typename T>
static inline T tex2D(cudaTextureObject_t tex, double u, double v);
template <>
inline float4 tex2D<float4>(cudaTextureObject_t tex, double u, double v) {
  return ::rt_shim::Tex2DPointWrap(tex, (float)u, (float)v);
}
