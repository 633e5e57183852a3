// Test-infrastructure shim: single-rank stand-in for the MPI calls the reference makes.
#pragma once
#include <cstring>
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
#define MPI_COMM_WORLD 0
#define MPI_FLOAT 4
#define MPI_SUM 0
static inline int MPI_Init(int *, char ***) { return 0; }
static inline int MPI_Finalize() { return 0; }
static inline int MPI_Comm_size(MPI_Comm, int *s) { *s = 1; return 0; }
static inline int MPI_Comm_rank(MPI_Comm, int *r) { *r = 0; return 0; }
static inline int MPI_Reduce(const void *s, void *r, int n, MPI_Datatype, MPI_Op, int, MPI_Comm) {
  std::memcpy(r, s, (size_t)n * 4); return 0;
}
