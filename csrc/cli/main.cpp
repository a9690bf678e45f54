// main.cpp -- command-line driver: `poisson_b200 M N [options]`.
//
// Keeps the reference's interface (two positional integers, default 40 40 when fewer than two are given,
// S4:995-999; delta = 1e-6 and max_iter = (M-1)(N-1), S4:1005-1006) and prints the stage-4 report verbatim
// (S4:1001-1002, S4:900-902, S4:969-978, S4:1027-1033), followed by the lines the reference lacks (error
// against the analytic solution, device time per iteration).  `--stage 0|1` selects the CPU programs'
// behaviour and formats (S0:176-196, S1:204-229): stage 0 without M N sweeps the grids the reference hard-codes
// (10^2, 20^2, 40^2, S0:177; `--grids` overrides), stage 1 prints the reference's banner block and one line per
// entry of `--threads a,b,c` (S1:208-226).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <string>
#include <vector>

#include <sys/wait.h>
#include <unistd.h>

#include "../../include/poisson_b200.h"

namespace {

double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

void usage() {
  std::cout << "usage: poisson_b200 [M N] [--gpus G] [--grid PxxPy|reference|auto] [--virtual] [--norm weighted|unweighted]\n"
               "                    [--delta D] [--max-iter K] [--method persistent|step_fused|step_classic]\n"
               "                    [--fast-div] [--dump FILE] [--stage 0|1|2|3|4] [--ranks P] [--threads T[,T...]]\n"
               "                    [--grids MxN[,MxN...]]\n"
               "                    [--quiet-extras]\n";
}

void dump_solution(const std::string& path, const std::vector<double>& w, int M, int N) {
  const bool bin = path.size() > 4 && path.substr(path.size() - 4) == ".bin";
  if (bin) {
    std::ofstream f(path, std::ios::binary);
    f << "poisson_b200 " << M << " " << N << " float64 little-endian row=i(x) col=j(y)\n";
    f.write(reinterpret_cast<const char*>(w.data()), (std::streamsize)(w.size() * sizeof(double)));
    return;
  }
  std::ofstream f(path);
  f << std::setprecision(17);
  const double h1 = 2.0 / M, h2 = 1.2 / N;
  for (int i = 0; i <= M; ++i) {
    for (int j = 0; j <= N; ++j) f << (-1.0 + i * h1) << " " << (-0.6 + j * h2) << " " << w[(size_t)i * (N + 1) + j] << "\n";
    f << "\n";
  }
}

// "4,8,16" -> {4,8,16}; "10x10,20x20" -> {10,10,20,20}
std::vector<int> parse_int_list(const std::string& text) {
  std::vector<int> out;
  size_t pos = 0;
  while (pos < text.size()) {
    size_t end = text.find_first_of(",x", pos);
    if (end == std::string::npos) end = text.size();
    out.push_back(std::atoi(text.substr(pos, end - pos).c_str()));
    pos = end + 1;
  }
  return out;
}

void print_extras(const pb_stats& st) {
  std::printf("   L2 error vs analytic (in D)  ~ %.9e\n", st.l2_in_D);
  std::printf("   L2 error vs analytic (all)   ~ %.9e\n", st.l2_all);
  std::printf("   Max error vs analytic      ~ %.9e\n", st.max_abs);
  if (st.iters > 0 && st.t_iterations > 0) std::printf("   Device time per iteration    ~ %.3f us\n", 1e6 * st.t_iterations / st.iters);
}

}  // namespace

int main(int argc, char** argv) {
  const double t_launch = now_s();
  double t_program_start = t_launch;
  int M = 40, N = 40, gpus = 1, Px = 0, Py = 0, stage = 4, ranks = 1;
  std::vector<int> thread_list{1}, grid_list;
  bool virt = false, extras = true, ref_grid = false;
  std::string dump, method = "persistent";
  pb_solve_options opt{};
  opt.method = PB_METHOD_PERSISTENT;
  opt.norm = PB_NORM_WEIGHTED;
  opt.delta = 1e-6;
  opt.max_iter = -1;
  opt.exact_div = 2;  // Markstein sequence: the bits of IEEE division at a quarter of the instructions
  opt.error_report = 1;
  bool norm_set = false;

  std::vector<std::string> pos;
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    auto next = [&]() -> std::string {
      if (i + 1 >= argc) {
        usage();
        std::exit(2);
      }
      return argv[++i];
    };
    if (a == "--gpus") gpus = std::atoi(next().c_str());
    else if (a == "--grid") {
      const std::string g = next();
      if (g == "reference") ref_grid = true;
      else if (g != "auto" && std::sscanf(g.c_str(), "%dx%d", &Px, &Py) != 2) { usage(); return 2; }
    } else if (a == "--virtual") virt = true;
    else if (a == "--norm") { opt.norm = next() == "unweighted" ? PB_NORM_UNWEIGHTED : PB_NORM_WEIGHTED; norm_set = true; }
    else if (a == "--delta") opt.delta = std::atof(next().c_str());
    else if (a == "--max-iter") opt.max_iter = std::atoll(next().c_str());
    else if (a == "--method") method = next();
    else if (a == "--fast-div") opt.exact_div = 0;
    else if (a == "--dump") dump = next();
    else if (a == "--stage") stage = std::atoi(next().c_str());
    else if (a == "--threads") thread_list = parse_int_list(next());
    else if (a == "--ranks") ranks = std::atoi(next().c_str());
    else if (a == "--grids") grid_list = parse_int_list(next());
    else if (a == "--quiet-extras") extras = false;
    else if (a == "-h" || a == "--help") { usage(); return 0; }
    else pos.push_back(a);
  }
  if (pos.size() >= 2) {  // a single positional argument is ignored, like the reference (argc >= 3)
    M = std::atoi(pos[0].c_str());
    N = std::atoi(pos[1].c_str());
  }
  if (method == "step_fused") opt.method = PB_METHOD_STEP_FUSED;
  else if (method == "step_classic") opt.method = PB_METHOD_STEP_CLASSIC;

  // ---- distributed CPU stages (2: "pure MPI", 3: "MPI + OpenMP"): this program is its own launcher ----
  if (stage == 2 || stage == 3) {
    // `--ranks P` forks P - 1 further processes; all of them attach to one shared-memory segment named after this
    // process (csrc/runtime/shm_comm.h) and run the C++ host loop (csrc/cpu/dist_solver.cpp).  Rank 0 prints the
    // reference's report (S2:476-497, S3:486-507).
    if (ranks < 1) { usage(); return 2; }
    if (Px == 0) pb_choose_proc_grid(ranks, &Px, &Py);
    if (Px * Py != ranks) {
      std::cerr << "error: --grid " << Px << "x" << Py << " does not match " << ranks << " ranks\n";
      return 2;
    }
    const int threads = stage == 3 ? thread_list.front() : 1;
    const std::string job = std::to_string((long long)getpid());
    int my_rank = 0;
    std::vector<pid_t> children;
    for (int r = 1; r < ranks; ++r) {
      const pid_t pid = fork();
      if (pid < 0) { std::cerr << "error: fork failed\n"; return 1; }
      if (pid == 0) { my_rank = r; children.clear(); break; }
      children.push_back(pid);
    }
    if (my_rank == 0) {
      if (stage == 3) std::cout << "MPI/OpenMP run with " << ranks << " MPI processes; M=" << M << ", N=" << N << std::endl;
      else std::cout << "Pure MPI 2D run with " << ranks << " processes; M=" << M << ", N=" << N << std::endl;
    }
    pb_stats st{};
    const int rc = pb_cpu_dist_solve(M, N, Px, Py, my_rank, job.c_str(), opt.norm, opt.delta, opt.max_iter, threads, nullptr, &st);
    if (my_rank != 0) _exit(rc == 0 ? 0 : 1);
    int bad = rc != 0;
    for (pid_t c : children) {
      int status = 0;
      waitpid(c, &status, 0);
      bad |= !(WIFEXITED(status) && WEXITSTATUS(status) == 0);
    }
    if (bad) { std::cerr << "error: " << pb_last_error() << "\n"; return 1; }
    if (st.status == PB_STATUS_CONVERGED)
      std::cout << "Converged after " << st.iters << " iterations (||w(k+1)-w(k)|| < " << opt.delta << ").\n";
    std::cout << "M=" << M << ", N=" << N << " | Iter=" << st.iters << " | Time=" << std::fixed << std::setprecision(6)
              << st.seconds << " s\n";
    std::cout.flush();
    if (extras) {
      print_extras(st);
      std::printf("   Halo exchange + all-reduce time (max over ranks) ~ %.6f s\n", st.t_comm);
    }
    return 0;
  }

  // ---- CPU stages ----
  if (stage == 0 || stage == 1) {
    if (!norm_set) opt.norm = stage == 0 ? PB_NORM_UNWEIGHTED : PB_NORM_WEIGHTED;
    if (thread_list.empty() || grid_list.size() % 2 != 0) { usage(); return 2; }
    if (grid_list.empty()) {
      if (pos.size() >= 2 || stage == 1) grid_list = {M, N};
      else grid_list = {10, 10, 20, 20, 40, 40};  // the reference's stage-0 program takes no arguments (S0:177)
    }
    for (size_t gi = 0; gi + 1 < grid_list.size(); gi += 2) {
      M = grid_list[gi], N = grid_list[gi + 1];
      if (stage == 1) {
        std::cout << "--- (Variant 9: Ellipse x^2 + 4y^2 < 1, OpenMP Test) ---\n";
        std::cout << "Grid: M=" << M << ", N=" << N << "\n";
        std::cout << "--------------------------------------------------------\n";
      }
      std::vector<double> w((size_t)(M + 1) * (N + 1));
      pb_stats st{};
      for (int threads : (stage == 0 ? std::vector<int>{1} : thread_list)) {
        if (pb_cpu_solve(M, N, opt.norm, opt.delta, opt.max_iter, threads, 1, w.data(), &st) != 0) {
          std::cerr << "error: " << pb_last_error() << "\n";
          return 1;
        }
        // both CPU programs print the literal symbol, not the value of delta (S0:157-158, S1:192)
        if (st.status == PB_STATUS_CONVERGED)
          std::cout << "Converged after " << st.iters << " iterations (||w(k+1)-w(k)|| < δ)." << std::endl;
        if (stage == 0)
          std::cout << "M=" << M << ", N=" << N << " | Iter=" << st.iters << " | Time=" << std::fixed << std::setprecision(4)
                    << st.seconds << " s\n";
        else
          std::cout << "Threads = " << std::setw(2) << threads << " | Time = " << std::fixed << std::setprecision(3)
                    << st.seconds << " s\n";
      }
      if (stage == 1) std::cout << "--------------------------------------------------------\n";
      std::cout.flush();
      if (extras) print_extras(st);
      if (!dump.empty()) dump_solution(dump, w, M, N);
    }
    return 0;
  }

  // ---- GPU (stage 4) ----
  // Runtime bootstrap = what MPI_Init and the launcher's GPU binding are to the reference (S4:986-990; its clock
  // starts after MPI_Init, S4:993): only the GPUs of this run are made visible (the driver then initialises only
  // those), their contexts are created and peer access is enabled, one host thread per GPU.  The program clock
  // starts afterwards, like the reference's; the bootstrap time is printed with the extras.
  if (!virt && gpus >= 1 && !std::getenv("CUDA_VISIBLE_DEVICES")) {
    std::string vis;
    for (int d = 0; d < gpus; ++d) vis += (d ? "," : "") + std::to_string(d);
    setenv("CUDA_VISIBLE_DEVICES", vis.c_str(), 0);
  }
  const int ndev = pb_cuda_device_count();
  if (ndev < 1) {
    std::cerr << "error: no CUDA device; use --stage 0 or --stage 1 for the CPU solvers\n";
    return 1;
  }
  int nranks = virt ? (Px > 0 ? Px * Py : gpus) : gpus;
  if (!virt && gpus > ndev) {
    std::cerr << "error: " << gpus << " GPUs requested, " << ndev << " visible\n";
    return 1;
  }
  std::vector<int> devs(virt ? 1 : nranks);
  for (size_t r = 0; r < devs.size(); ++r) devs[r] = (int)r;
  if (pb_runtime_bootstrap(devs.data(), (int)devs.size()) != 0) {
    std::cerr << "error: " << pb_last_error() << "\n";
    return 1;
  }
  std::string grid_note = "given on the command line";
  if (Px == 0) {
    // default: derived from the peer-to-peer topology of the GPUs (uniform NVSwitch fabric -> slabs along x with
    // contiguous halo rows; NVLink islands -> the x-cuts between the islands; see csrc/core/decomp.h);
    // --grid reference restores the reference's near-square choice (S4:82-86)
    if (!ref_grid && !virt && nranks > 1) {
      std::vector<int> perf(nranks * nranks), order(nranks);
      if (pb_p2p_matrix(devs.data(), nranks, perf.data(), nullptr, nullptr) != 0) {
        std::cerr << "error: " << pb_last_error() << "\n";
        return 1;
      }
      const int kind = pb_choose_grid_topology(nranks, perf.data(), M, &Px, &Py, order.data());
      for (int r = 0; r < nranks; ++r) devs[r] = order[r];
      grid_note = kind == 0 ? "P2P topology: every pair equally ranked" : kind == 1 ? "P2P topology: NVLink islands"
                                                                                   : "P2P topology irregular: reference grid";
    } else {
      pb_choose_proc_grid(nranks, &Px, &Py);
      grid_note = "reference rule (S4:82-86)";
    }
  }
  const double t_bootstrap = now_s() - t_launch;
  t_program_start = now_s();
  if (Px * Py != nranks) {
    std::cerr << "error: --grid " << Px << "x" << Py << " does not match " << nranks << " ranks\n";
    return 2;
  }
  std::cout << "MPI + CUDA 2D run with " << nranks << " processes; M=" << M << ", N=" << N << std::endl;

  const double t_before_solver = now_s();
  pb_stats st{};
  std::vector<double> w;
  int rc = 0;
  if (virt || nranks == 1) {
    pb_solver* s = pb_solver_create(M, N, Px, Py, 0, nranks, 0);
    if (!s || pb_solver_finalize(s) != 0) { std::cerr << "error: " << pb_last_error() << "\n"; return 1; }
    rc = pb_solver_solve(s, &opt, nullptr, &st);
    if (rc == 0 && !dump.empty()) {
      w.assign((size_t)(M + 1) * (N + 1), 0.0);
      rc = pb_solver_copy_solution(s, w.data(), nullptr);
    }
    pb_solver_destroy(s);
  } else {
    pb_group* g = pb_group_create(M, N, Px, Py, devs.data(), nranks);
    if (!g) { std::cerr << "error: " << pb_last_error() << "\n"; return 1; }
    rc = pb_group_solve(g, &opt, &st);
    if (rc == 0 && !dump.empty()) {
      w.assign((size_t)(M + 1) * (N + 1), 0.0);
      rc = pb_group_copy_solution(g, w.data());
    }
    pb_group_destroy(g);
  }
  if (rc != 0) { std::cerr << "error: " << pb_last_error() << "\n"; return 1; }
  const double t_after_solver = now_s();

  if (st.status == PB_STATUS_CONVERGED)
    std::cout << "Converged after " << st.iters << " iterations (||w(k+1)-w(k)|| < " << opt.delta << ").\n";
  else if (st.status == PB_STATUS_WATCHDOG)
    std::cout << "Aborted: a device-side wait timed out.\n";
  std::cout << "   GPU compute time (Ap + D^{-1}r, max over ranks) ~ " << st.t_gpu << " s\n";
  std::cout << "   Host<->Device copy time (max over ranks)        ~ " << st.t_copy << " s\n";
  std::cout << "   MPI halo exchange time (max over ranks)         ~ " << st.t_comm << " s\n";
  std::cout << "   Preconditioner CPU part time (max over ranks)   ~ " << st.t_precond << " s\n";
  std::cout << "   Dot products time (max over ranks)              ~ " << st.t_dot << " s\n";
  const double time_init = t_before_solver - t_program_start, time_solver = t_after_solver - t_before_solver;
  std::cout << "M=" << M << ", N=" << N << " | Iter=" << st.iters << " | Total Time=" << std::fixed
            << std::setprecision(6) << (time_init + time_solver) << " s\n";
  std::cout << "   Init time (program)      ~ " << time_init << " s\n";
  std::cout << "   Solver time (MPI+CUDA)   ~ " << time_solver << " s\n";
  std::cout << "   Finalization time        ~ " << 0.0 << " s\n";
  std::cout.flush();
  if (extras) {
    std::printf("   Device solve time (assembly + PCG + error report) ~ %.6f s\n", st.seconds);
    std::printf("   Runtime bootstrap (contexts, peer access; before the program clock, like MPI_Init) ~ %.6f s\n", t_bootstrap);
    std::printf("   Process grid %dx%d (%s)\n", Px, Py, grid_note.c_str());
    print_extras(st);
  }
  if (!dump.empty()) dump_solution(dump, w, M, N);
  return st.status == PB_STATUS_WATCHDOG ? 3 : 0;
}
