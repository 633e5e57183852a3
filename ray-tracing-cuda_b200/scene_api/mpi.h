// mpi.h — single-box stand-in for the four MPI calls the reference's scenes make (spheres.cu:83-98,108;
// bunny.cu:79-94,105) so that they build and run without an MPI installation.  Ranks are processes started
// by `rtrun -n N ./scene` (or torchrun / mpirun): the rank and world size come from the environment.  The
// image reduce itself does not go through MPI: DistributedMain uses one ncclReduce (rt_reduce, include/rtcore.h).
#pragma once
#include <cstdlib>
#include <cstring>
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
#define MPI_COMM_WORLD 0
#define MPI_SUCCESS 0
#define MPI_FLOAT 4
#define MPI_SUM 0
namespace rtapi {
inline int env_int(const char *const *names, int dflt) {
  for (; *names; ++names) {
    const char *v = std::getenv(*names);
    if (v && *v) return std::atoi(v);
  }
  return dflt;
}
inline int mpi_rank() {
  static const char *const n[] = {"RT_RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK", "RANK", nullptr};
  return env_int(n, 0);
}
inline int mpi_world_size() {
  static const char *const n[] = {"RT_WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE", "WORLD_SIZE", nullptr};
  int s = env_int(n, 1);
  return s > 0 ? s : 1;
}
}  // namespace rtapi
static inline int MPI_Init(int *, char ***) { return MPI_SUCCESS; }
static inline int MPI_Finalize() { return MPI_SUCCESS; }
static inline int MPI_Comm_size(MPI_Comm, int *size) { *size = rtapi::mpi_world_size(); return MPI_SUCCESS; }
static inline int MPI_Comm_rank(MPI_Comm, int *rank) { *rank = rtapi::mpi_rank(); return MPI_SUCCESS; }
