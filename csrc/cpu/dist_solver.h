// dist_solver.h -- the distributed CPU solver with its host loop in C++: the reference's stages 2 (pure MPI,
// solve_mpi S2:356-460) and 3 (MPI + OpenMP, S3:364-470) over the shared-memory transport of runtime/shm_comm.h.
#pragma once
#include "../runtime/shm_comm.h"
#include "oracle.h"

namespace pb {

struct DistTimes {
  double assemble = 0.0, comm = 0.0;  // seconds inside the solve: assembly; halo exchanges + all-reduces (this rank)
};

// Collective over the ranks of `comm` (rank r owns block r of the pg.Px x pg.Py grid; comm.size() = Px * Py).
// w_block, if not null, receives this rank's block in the padded LocalBlock layout ((nx + 2) x pitch doubles).
// stats (same on every rank): iterations, stop norm, error norms over the whole domain, seconds = barrier-to-barrier
// wall time of assembly + iterations (what the reference brackets with MPI_Barrier / MPI_Wtime, S2:483-489).
int cpu_dist_solve(int M, int N, ProcGrid pg, ShmComm& comm, const CpuOptions& opt, double* w_block, SolveStats* stats,
                   DistTimes* times);

}  // namespace pb
