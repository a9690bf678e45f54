// ref_comm_bench.cu -- micro-benchmark of the reference's OWN communication paths (BASELINE.json config 5).
//
// The unmodified stage-4 source (baseline/_ref/src, a verbatim copy of /root/reference) is compiled into this
// program by #include with its `main` renamed, so the functions timed here are the reference's, not restatements:
//   halo      exchange_halos_2d_gpu (S4:331-500): 4 x cudaMemcpy / cudaMemcpy2D device -> pageable host,
//             4 x MPI_Sendrecv, 4 x host -> device or cudaMemset on the boundary of Omega
//   dot       the steps of its dot_gpu lambda (S4:771-786; a lambda inside gradient_solver_mpi, so its four statements
//             are repeated here around the reference's own dot_kernel) + MPI_Allreduce (S4:871)
// Ranks are processes of the MPI shim (baseline/mpishim), one GPU each, on the reference's own process grid
// (choose_process_grid, S4:82-86); every rank owns an edge x edge block.  Times are MPI_Wtime between barriers, max
// over ranks (MPI_Reduce MAX), like the reference's own timers (S4:962-966).
//   usage: ref_comm_bench [iters]      -> one JSON line per measurement on rank 0
#define main reference_main
#include "stage4-mpi+cuda/poisson_mpi_cuda2.cu"
#undef main

#include <cstdio>

int main(int argc, char** argv) {
  MPI_Init(&argc, &argv);
  int rank, size;
  MPI_Comm_rank(MPI_COMM_WORLD, &rank);
  MPI_Comm_size(MPI_COMM_WORLD, &size);
  const int iters = argc > 1 ? std::atoi(argv[1]) : 200;
  int Px, Py;
  choose_process_grid(size, Px, Py);
  const int px = rank % Px, py = rank / Px;
  const int nbrs = (px > 0) + (px < Px - 1) + (py > 0) + (py < Py - 1);

  const int edges[] = {64, 256, 1024, 4096, 16384};
  for (int edge : edges) {
    const int local_nx = edge, local_ny = edge;
    const size_t elems = (size_t)(local_nx + 2) * (size_t)(local_ny + 2);
    double *d_p = nullptr, *d_q = nullptr, *d_partial = nullptr;
    checkCuda(cudaMalloc(&d_p, elems * sizeof(double)), "cudaMalloc p");
    checkCuda(cudaMalloc(&d_q, elems * sizeof(double)), "cudaMalloc q");
    checkCuda(cudaMalloc(&d_partial, 128 * 256 * sizeof(double)), "cudaMalloc partial");
    checkCuda(cudaMemset(d_p, 0, elems * sizeof(double)), "memset p");
    checkCuda(cudaMemset(d_q, 0, elems * sizeof(double)), "memset q");
    std::vector<double> h_partial(128 * 256);

    // ---- halo exchange: the reference's function, as called at S4:850 ----
    double t_comm = 0.0, t_copy = 0.0;
    for (int it = 0; it < 5; ++it) exchange_halos_2d_gpu(d_p, local_nx, local_ny, Px, Py, rank, MPI_COMM_WORLD, t_comm, t_copy);
    t_comm = t_copy = 0.0;
    MPI_Barrier(MPI_COMM_WORLD);
    double t0 = MPI_Wtime();
    for (int it = 0; it < iters; ++it) exchange_halos_2d_gpu(d_p, local_nx, local_ny, Px, Py, rank, MPI_COMM_WORLD, t_comm, t_copy);
    double mine[3] = {MPI_Wtime() - t0, t_comm, t_copy}, mx[3];
    MPI_Reduce(mine, mx, 3, MPI_DOUBLE, MPI_MAX, 0, MPI_COMM_WORLD);
    if (rank == 0)
      std::printf("{\"what\": \"halo_exchange\", \"impl\": \"reference exchange_halos_2d_gpu (cudaMemcpy + MPI_Sendrecv)\", "
                  "\"edge_doubles\": %d, \"us\": %.3f, \"us_mpi\": %.3f, \"us_copies\": %.3f, \"n_gpus\": %d, \"grid\": \"%dx%d\", "
                  "\"neighbours_rank0\": %d}\n",
                  edge, 1e6 * mx[0] / iters, 1e6 * mx[1] / iters, 1e6 * mx[2] / iters, size, Px, Py, nbrs);

    // ---- dot product + all-reduce: the statements of dot_gpu (S4:771-786) and S4:871 ----
    auto dot_allreduce = [&]() -> double {
      dot_kernel<<<128, 256>>>(d_p, d_q, d_partial, local_nx, local_ny, 1.0);
      checkCuda(cudaGetLastError(), "dot_kernel launch");
      checkCuda(cudaDeviceSynchronize(), "dot_kernel sync");
      checkCuda(cudaMemcpy(h_partial.data(), d_partial, h_partial.size() * sizeof(double), cudaMemcpyDeviceToHost), "partial D2H");
      double sum = 0.0;
      for (size_t i = 0; i < h_partial.size(); ++i) sum += h_partial[i];
      double global = 0.0;
      MPI_Allreduce(&sum, &global, 1, MPI_DOUBLE, MPI_SUM, MPI_COMM_WORLD);
      return global;
    };
    double keep = 0.0;
    for (int it = 0; it < 5; ++it) keep += dot_allreduce();
    MPI_Barrier(MPI_COMM_WORLD);
    t0 = MPI_Wtime();
    for (int it = 0; it < iters; ++it) keep += dot_allreduce();
    double dt = MPI_Wtime() - t0, dmx = 0.0;
    MPI_Reduce(&dt, &dmx, 1, MPI_DOUBLE, MPI_MAX, 0, MPI_COMM_WORLD);
    // the all-reduce alone (what remains of the path when the dot product itself is free)
    MPI_Barrier(MPI_COMM_WORLD);
    t0 = MPI_Wtime();
    for (int it = 0; it < iters; ++it) {
      double s = 1.0, g = 0.0;
      MPI_Allreduce(&s, &g, 1, MPI_DOUBLE, MPI_SUM, MPI_COMM_WORLD);
      keep += g;
    }
    double da = MPI_Wtime() - t0, damx = 0.0;
    MPI_Reduce(&da, &damx, 1, MPI_DOUBLE, MPI_MAX, 0, MPI_COMM_WORLD);
    if (rank == 0)
      std::printf("{\"what\": \"dot_allreduce\", \"impl\": \"reference dot_kernel + sync + 256 KiB D2H + host sum + MPI_Allreduce\", "
                  "\"edge_doubles\": %d, \"us\": %.3f, \"us_mpi_allreduce_alone\": %.3f, \"n_gpus\": %d, \"grid\": \"%dx%d\", "
                  "\"checksum\": %.1f}\n",
                  edge, 1e6 * dmx / iters, 1e6 * damx / iters, size, Px, Py, keep);
    cudaFree(d_p);
    cudaFree(d_q);
    cudaFree(d_partial);
  }
  MPI_Finalize();
  return 0;
}
