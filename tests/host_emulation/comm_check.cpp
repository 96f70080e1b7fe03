// The MPI subset of include/mini/b200/nompi/mpi.h between real processes (no GPU): tests/test_host_logic.py
#include <cstdio>
#include "mpi.h"
int main() {
  MPI_Init(NULL, NULL);
  int n, r;
  MPI_Comm_size(MPI_COMM_WORLD, &n);
  MPI_Comm_rank(MPI_COMM_WORLD, &r);
  double lo, hi, sum, mine = 1.5 + r;
  for (int round = 0; round < 200; ++round) {   // many rounds: the barrier's sense reversal
    mine = 1.5 + r + round;
    MPI_Allreduce(&mine, &lo, 1, MPI_DOUBLE, MPI_MIN, MPI_COMM_WORLD);
    MPI_Allreduce(&mine, &hi, 1, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD);
    MPI_Allreduce(&mine, &sum, 1, MPI_DOUBLE, MPI_SUM, MPI_COMM_WORLD);
    if (lo != 1.5 + round || hi != 1.5 + (n - 1) + round || sum != n * (1.5 + round) + n * (n - 1) / 2.0) {
      std::printf("rank %d round %d: %g %g %g\n", r, round, lo, hi, sum);
      return 1;
    }
  }
  MPI_Barrier(MPI_COMM_WORLD);
  std::printf("rank %d of %d ok\n", r, n);
  MPI_Finalize();
  return 0;
}
