// nompi/mpi.h — the MPI calls of the reference's drivers (demo/euler/sourceless.cpp:60-66, 101, 187-196, 222-227) on
// mini::b200::Comm, for images without an MPI library.  Put this directory on the include path ONLY where there is no
// real <mpi.h>:   g++ -I include -I include/mini/b200/nompi driver.cpp
// Supported: MPI_COMM_WORLD; MPI_DOUBLE with MPI_MIN / MPI_MAX / MPI_SUM; count >= 1.
#ifndef MINI_B200_NOMPI_MPI_H_
#define MINI_B200_NOMPI_MPI_H_

#include "mini/b200/comm.hpp"

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
constexpr MPI_Comm MPI_COMM_WORLD = 0;
constexpr MPI_Datatype MPI_DOUBLE = 1;
constexpr MPI_Op MPI_MIN = 0, MPI_MAX = 1, MPI_SUM = 2;
constexpr int MPI_SUCCESS = 0;

inline int MPI_Init(int *, char ***) { mini::b200::Comm::World(); return MPI_SUCCESS; }
inline int MPI_Finalize() { mini::b200::Comm::World().Finalize(); return MPI_SUCCESS; }
inline int MPI_Comm_rank(MPI_Comm, int *rank) { *rank = mini::b200::Comm::World().rank(); return MPI_SUCCESS; }
inline int MPI_Comm_size(MPI_Comm, int *size) { *size = mini::b200::Comm::World().size(); return MPI_SUCCESS; }
inline int MPI_Barrier(MPI_Comm) { mini::b200::Comm::World().Barrier(); return MPI_SUCCESS; }
inline double MPI_Wtime() { return mini::b200::Comm::World().Wtime(); }
inline int MPI_Allreduce(const void *send, void *recv, int count, MPI_Datatype type, MPI_Op op, MPI_Comm) {
  if (type != MPI_DOUBLE) throw std::runtime_error("nompi: MPI_Allreduce supports MPI_DOUBLE only");
  auto &world = mini::b200::Comm::World();
  for (int i = 0; i < count; ++i) {
    static_cast<double *>(recv)[i] = world.Allreduce(static_cast<const double *>(send)[i],
        op == MPI_MIN ? mini::b200::Comm::kMin : op == MPI_MAX ? mini::b200::Comm::kMax : mini::b200::Comm::kSum);
  }
  return MPI_SUCCESS;
}
#endif  // MINI_B200_NOMPI_MPI_H_
