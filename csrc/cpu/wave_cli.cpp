// wave_cli.cpp — stand-alone native executables of the CPU solvers.
//
// One source, four personalities, selected by the binary's name (argv[0]) or by
// --mode; command lines are the reference's (report.pdf p.20-26 §5):
//   wave        N tau K                  (sequential;     report.pdf p.20 §5.1.3)
//   openmpwave  N tau K threads          (OpenMP;         report.pdf p.21 §5.2.3)
//   mpi         -np P  N tau K           (rank-threaded;  report.pdf p.22 §5.3.3 `mpirun -np P ./mpi`)
//   mpiomp      -np P  N tau K threads   (ranks x threads; report.pdf p.25 §5.4)
// Extra options: --L L | --L3 Lx Ly Lz | --Lpi  --bc ddd|ddp|...  --dims P Q R  --every n  --quiet
#include "wave3d_cpu.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

static void usage(const char* me) {
  std::fprintf(stderr,
               "usage: %s [--mode seq|omp|mpi|mpiomp] [-np P | --dims P Q R] N tau K [threads]\n"
               "          [--L L | --L3 Lx Ly Lz | --Lpi] [--bc ddd] [--every n] [--quiet]\n", me);
}

// default rank grid: prime factors of P (largest first) go to z, then y, whichever
// is currently smaller; x (contiguous in memory) is split only by --dims.  Same
// rule as parallel/decomp.py::default_dims.
static void dims_create(int p, int dims[3]) {
  dims[0] = dims[1] = dims[2] = 1;
  std::vector<int> f;
  for (int q = 2; p > 1;) {
    if (p % q == 0) { f.push_back(q); p /= q; } else ++q;
  }
  for (size_t i = f.size(); i-- > 0;) {
    int t = dims[2] <= dims[1] ? 2 : 1;
    dims[t] *= f[i];
  }
}

int main(int argc, char** argv) {
  std::string me = argv[0];
  size_t slash = me.find_last_of('/');
  if (slash != std::string::npos) me = me.substr(slash + 1);
  std::string mode = "seq";
  if (me.find("openmp") != std::string::npos) mode = "omp";
  else if (me.find("mpiomp") != std::string::npos) mode = "mpiomp";
  else if (me.find("mpi") != std::string::npos) mode = "mpi";

  W3dCpuRun run;
  std::memset(&run, 0, sizeof(run));
  run.L[0] = run.L[1] = run.L[2] = 1.0;
  run.dims[0] = run.dims[1] = run.dims[2] = 1;
  run.every = 2;
  run.verbose = 1;
  run.threads = 1;
  int np = 1;
  bool have_dims = false;
  std::vector<std::string> pos;
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    auto need = [&](int n) {
      if (i + n >= argc) { usage(argv[0]); std::exit(2); }
    };
    if (a == "--mode") { need(1); mode = argv[++i]; }
    else if (a == "-np" || a == "-n") { need(1); np = std::atoi(argv[++i]); }
    else if (a == "--dims") { need(3); for (int d = 0; d < 3; ++d) run.dims[d] = std::atoi(argv[++i]); have_dims = true; }
    else if (a == "--L") { need(1); run.L[0] = run.L[1] = run.L[2] = std::atof(argv[++i]); }
    else if (a == "--L3") { need(3); for (int d = 0; d < 3; ++d) run.L[d] = std::atof(argv[++i]); }
    else if (a == "--Lpi") { run.L[0] = run.L[1] = run.L[2] = M_PI; }
    else if (a == "--bc") { need(1); const char* s = argv[++i]; for (int d = 0; d < 3 && s[d]; ++d) run.periodic[d] = (s[d] == 'p'); }
    else if (a == "--every") { need(1); run.every = std::atoi(argv[++i]); }
    else if (a == "--quiet") { run.verbose = 0; }
    else if (a == "-h" || a == "--help") { usage(argv[0]); return 0; }
    else pos.push_back(a);
  }
  if (pos.size() < 3) { usage(argv[0]); return 2; }
  const int N = std::atoi(pos[0].c_str());
  run.N[0] = run.N[1] = run.N[2] = N;
  run.tau = std::atof(pos[1].c_str());
  run.K = std::atoi(pos[2].c_str());
  if (pos.size() >= 4) run.threads = std::atoi(pos[3].c_str());
  if (mode == "seq") { run.threads = 1; np = 1; }
  if (mode == "omp") np = 1;
  if (mode == "mpi") run.threads = 1;
  if (const char* e = std::getenv("OMP_NUM_THREADS"))
    if (pos.size() < 4 && (mode == "omp" || mode == "mpiomp")) run.threads = std::atoi(e);
  if (run.threads < 1) run.threads = 1;
  if (!have_dims) dims_create(np, run.dims);
  if (N < 2 || run.tau <= 0 || run.K < 1 || run.every < 1) { usage(argv[0]); return 2; }

  double courant = 0;
  for (int d = 0; d < 3; ++d) courant += 1.0 / ((run.L[d] / N) * (run.L[d] / N));
  courant = run.tau * std::sqrt(courant);
  if (courant > 1.0)
    std::fprintf(stderr, "warning: Courant number %.3f > 1, the scheme is unstable\n", courant);

  const int P = run.dims[0] * run.dims[1] * run.dims[2];
  if (P > 1) std::printf("Running test with grid size %d^3 and %d processes\n", N, P);
  else std::printf("Running test with grid size %d^3\n", N);
  std::vector<double> trace(2 * (run.K + 1));
  double ph[5];
  int rc = w3d_cpu_solve(&run, trace.data(), ph);
  if (rc != 0) { std::fprintf(stderr, "solve failed: %d\n", rc); return 1; }
  double upd = (double)run.K;
  for (int d = 0; d < 3; ++d) upd *= run.periodic[d] ? N : N - 1;
  std::printf("Total time: %.6f s (init %.6f, compute %.6f, boundary %.6f, exchange %.6f)\n",
              ph[0], ph[1], ph[2], ph[3], ph[4]);
  std::printf("ranks %d (%dx%dx%d) x threads %d, %.4f GLUP/s\n", P, run.dims[0], run.dims[1],
              run.dims[2], run.threads, upd / ph[0] / 1e9);
  return 0;
}
