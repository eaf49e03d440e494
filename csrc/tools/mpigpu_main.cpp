// mpigpu_main.cpp — stand-alone native GPU executable (no Python, no torch).
//
// Native equivalent of the reference's `mpirun -np P ./mpigpu-1 N tau K 1`
// (report.pdf p.15 §4.2.4): same positional command line, same "Step ..." lines.
// One process drives G GPUs (`-np G`, slab decomposition along z): every GPU gets
// its own host thread, peer access is enabled between neighbours and the fused
// step kernel pushes halos over NVLink itself — there is no MPI in this program.
//
//   mpigpu [-np G] N tau K [threads] [--L L | --Lpi] [--bc ddd] [--variant v] [--zchunk z]
//          [--every n] [--quiet] [--kernel pair|fused|tma|box]
// --kernel pair (default): two time steps per pass + fix-up pass; fused: one pass per step.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "wave3d.h"

#define CK(x)                                                                                   \
  do {                                                                                          \
    cudaError_t e_ = (x);                                                                       \
    if (e_ != cudaSuccess) {                                                                    \
      std::fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_));   \
      std::exit(1);                                                                             \
    }                                                                                           \
  } while (0)
#define W3(x)                                                                                   \
  do {                                                                                          \
    int r_ = (x);                                                                               \
    if (r_ != 0) {                                                                              \
      std::fprintf(stderr, "%s:%d %s -> %d: %s\n", __FILE__, __LINE__, #x, r_, w3d_last_error()); \
      std::exit(1);                                                                             \
    }                                                                                           \
  } while (0)

struct Opts {
  int N = 512, K = 20, G = 1, variant = -1, zchunk = 0, every = 2, verbose = 1;
  double tau = 1e-3, L = 1.0;
  int periodic[3] = {0, 0, 0};
  std::string kernel = "pair";
};

struct HostBarrier {
  explicit HostBarrier(int n) : n_(n) {}
  void wait() {
    std::unique_lock<std::mutex> lk(m_);
    int g = gen_;
    if (++c_ == n_) { c_ = 0; ++gen_; cv_.notify_all(); }
    else cv_.wait(lk, [&] { return g != gen_; });
  }
  std::mutex m_; std::condition_variable cv_; int n_, c_ = 0, gen_ = 0;
};

struct Rank {
  int dev = 0, lo = 0, n[3] = {0, 0, 0};
  W3dBlock blk;
  double* field[2] = {nullptr, nullptr};
  double* aux[2] = {nullptr, nullptr};      // two more arrays for the two-steps-per-pass kernel
  bool faces = false;                       // any face with a neighbour (other GPU or periodic wrap)
  unsigned long long* flags = nullptr;
  double* tab[3] = {nullptr, nullptr, nullptr};
  uint32_t* status = nullptr;
  void* ws = nullptr;
  double* trace = nullptr;     // (K+1) x 2 on device
  W3dPlan* plan = nullptr;
  std::vector<double> htrace;
  float loop_ms = 0.f, init_ms = 0.f;
};

static void split(int n, int parts, int r, int* lo, int* cnt) {
  int q = n / parts, rem = n % parts;
  *cnt = q + (r < rem ? 1 : 0);
  *lo = r * q + std::min(r, rem);
}

int main(int argc, char** argv) {
  Opts o;
  std::vector<std::string> pos;
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    auto nxt = [&]() -> const char* { if (i + 1 >= argc) { std::fprintf(stderr, "missing value for %s\n", a.c_str()); std::exit(2); } return argv[++i]; };
    if (a == "-np" || a == "-n") o.G = std::atoi(nxt());
    else if (a == "--L") o.L = std::atof(nxt());
    else if (a == "--Lpi") o.L = M_PI;
    else if (a == "--bc") { const char* s = nxt(); for (int d = 0; d < 3 && s[d]; ++d) o.periodic[d] = s[d] == 'p'; }
    else if (a == "--variant") o.variant = std::atoi(nxt());
    else if (a == "--zchunk") o.zchunk = std::atoi(nxt());
    else if (a == "--every") o.every = std::atoi(nxt());
    else if (a == "--kernel") o.kernel = nxt();
    else if (a == "--quiet") o.verbose = 0;
    else pos.push_back(a);
  }
  if (pos.size() < 3) {
    std::fprintf(stderr, "usage: %s [-np G] N tau K [threads] [--L L|--Lpi] [--bc ddd] [--variant v] [--zchunk z] [--every n] [--kernel fused|tma|box] [--quiet]\n", argv[0]);
    return 2;
  }
  o.N = std::atoi(pos[0].c_str()); o.tau = std::atof(pos[1].c_str()); o.K = std::atoi(pos[2].c_str());
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  if (o.G > ndev) { std::fprintf(stderr, "asked for %d GPUs, %d visible\n", o.G, ndev); return 1; }
  const int G = o.G, K = o.K, N = o.N;
  int unk[3], first[3];
  for (int d = 0; d < 3; ++d) { unk[d] = o.periodic[d] ? N : N - 1; first[d] = o.periodic[d] ? 0 : 1; }
  const double h = o.L / N, coef = o.tau * o.tau / (h * h);
  double at2 = 0;
  for (int d = 0; d < 3; ++d) { double m = o.periodic[d] ? 2.0 : 1.0; at2 += m * m / (o.L * o.L); }
  const double a_t = M_PI * std::sqrt(at2);
  const double courant = o.tau * std::sqrt(3.0) / h;
  if (courant > 1.0) std::fprintf(stderr, "warning: Courant number %.3f > 1, the scheme is unstable\n", courant);
  if (G > 1) std::printf("Running test with grid size %d^3 and %d processes\n", N, G);
  else std::printf("Running test with grid size %d^3\n", N);

  std::vector<Rank> R(G);
  // ---- allocation (slab along z) ----
  for (int r = 0; r < G; ++r) {
    Rank& k = R[r];
    k.dev = r;
    CK(cudaSetDevice(r));
    k.n[0] = unk[0]; k.n[1] = unk[1];
    split(unk[2], G, r, &k.lo, &k.n[2]);
    k.blk.n[0] = k.n[0]; k.blk.n[1] = k.n[1]; k.blk.n[2] = k.n[2];
    k.blk.xoff = W3D_XOFF; k.blk.pitch = w3d_pitch_for(k.n[0]);
    const size_t elems = (size_t)(k.n[2] + 2) * (k.n[1] + 2) * k.blk.pitch;
    for (int f = 0; f < 2; ++f) CK(cudaMalloc(&k.field[f], elems * sizeof(double)));
    for (int f = 0; f < 2; ++f) { CK(cudaMalloc(&k.aux[f], elems * sizeof(double))); CK(cudaMemset(k.aux[f], 0, elems * sizeof(double))); }
    CK(cudaMalloc(&k.flags, 64)); CK(cudaMemset(k.flags, 0, 64));
    CK(cudaMalloc(&k.status, 16)); CK(cudaMemset(k.status, 0, 16));
    CK(cudaMalloc(&k.ws, w3d_workspace_bytes())); CK(cudaMemset(k.ws, 0, w3d_workspace_bytes()));
    CK(cudaMalloc(&k.trace, sizeof(double) * 4 * (K + 1))); CK(cudaMemset(k.trace, 0, sizeof(double) * 4 * (K + 1)));   // 2 slots x (max, sum)
    for (int d = 0; d < 3; ++d) {
      const int lo = d == 2 ? first[2] + k.lo : first[d];
      std::vector<double> t(k.n[d] + 2);
      const int m = o.periodic[d] ? 2 : 1;
      for (int l = 0; l < k.n[d] + 2; ++l) {
        long g = (long)lo - 1 + l;
        if (o.periodic[d]) g = ((g % N) + N) % N;
        double v = std::sin(m * M_PI * (double)g / (double)N);
        if (g == 0 || g == N || (m == 2 && 2 * g == N)) v = 0.0;
        t[l] = v;
      }
      CK(cudaMalloc(&k.tab[d], t.size() * sizeof(double)));
      CK(cudaMemcpy(k.tab[d], t.data(), t.size() * sizeof(double), cudaMemcpyHostToDevice));
    }
  }
  // ---- peer access between z neighbours ----
  for (int r = 0; r < G; ++r)
    for (int q : {r - 1, r + 1, (r + G - 1) % G, (r + 1) % G}) {
      if (q < 0 || q >= G || q == r) continue;
      CK(cudaSetDevice(r));
      int can = 0;
      CK(cudaDeviceCanAccessPeer(&can, r, q));
      if (!can) { std::fprintf(stderr, "GPU %d cannot access GPU %d\n", r, q); return 1; }
      cudaError_t e = cudaDeviceEnablePeerAccess(q, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CK(e);
      (void)cudaGetLastError();
    }
  // ---- plans ----
  for (int r = 0; r < G; ++r) {
    Rank& k = R[r];
    CK(cudaSetDevice(r));
    W3dPlanDesc d;
    std::memset(&d, 0, sizeof(d));
    d.blk = k.blk;
    d.field[0] = k.field[0]; d.field[1] = k.field[1];
    for (int a = 0; a < 3; ++a) { d.tab[a] = k.tab[a]; d.coef[a] = coef; }
    auto set_peer = [&](int face, const Rank& q, int ghost, int their_face) {
      W3dPeer& pr = d.peer[face];
      pr.active = 1; pr.ghost_index = ghost;
      pr.field[0] = q.field[0]; pr.field[1] = q.field[1];
      pr.pitch = q.blk.pitch; pr.plane = (int64_t)(q.n[1] + 2) * q.blk.pitch;
      pr.flag = q.flags + their_face;
    };
    // x, y: periodic axes wrap onto the rank itself
    for (int a = 0; a < 2; ++a)
      if (o.periodic[a]) { set_peer(a * 2 + 0, k, k.n[a] + 1, a * 2 + 1); set_peer(a * 2 + 1, k, 0, a * 2 + 0); }
    int lo_n = r - 1, hi_n = r + 1;
    if (o.periodic[2]) { lo_n = (r + G - 1) % G; hi_n = (r + 1) % G; }
    if (lo_n >= 0) set_peer(4, R[lo_n], R[lo_n].n[2] + 1, 5);
    if (hi_n < G) set_peer(5, R[hi_n], 0, 4);
    d.flags_in = k.flags; d.status = k.status; d.workspace = k.ws; d.workspace_bytes = w3d_workspace_bytes();
    d.variant = o.variant; d.zchunk = o.zchunk; d.num_sms = 0; d.spin_timeout_ms = 0;
    for (int f = 0; f < 6; ++f) k.faces = k.faces || d.peer[f].active;
    W3(w3d_plan_create(&d, &k.plan));
    if (o.kernel == "pair") {
      double* peer_aux[12] = {nullptr};
      for (int f = 0; f < 6; ++f)
        if (d.peer[f].active) {
          // find the rank that owns this peer's field[0]
          for (int q = 0; q < G; ++q)
            if (R[q].field[0] == d.peer[f].field[0]) { peer_aux[2 * f] = R[q].aux[0]; peer_aux[2 * f + 1] = R[q].aux[1]; }
        }
      W3(w3d_plan_enable_pairs(k.plan, k.aux[0], k.aux[1], peer_aux, o.variant, o.zchunk));
    }
    if (o.verbose && r == 0) std::printf("plan: %s\n", w3d_plan_describe(k.plan));
  }
  // ---- run: one host thread per GPU ----
  HostBarrier bar(G);
  std::vector<double> ct(K + 1);
  for (int n = 0; n <= K; ++n) ct[n] = std::cos(a_t * o.tau * n);
  auto worker = [&](int r) {
    Rank& k = R[r];
    CK(cudaSetDevice(r));
    cudaStream_t st;
    CK(cudaStreamCreate(&st));
    cudaEvent_t e0, e1, e2;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); CK(cudaEventCreate(&e2));
    for (int rep = 0; rep < 2; ++rep) {   // rep 0 warms up (module load), rep 1 is timed
      bar.wait();
      CK(cudaEventRecord(e0, st));
      W3(w3d_plan_init(k.plan, st));
      CK(cudaStreamSynchronize(st));
      bar.wait();
      CK(cudaEventRecord(e1, st));
      auto coef = [](int n, double* a, double* b, double* cs) { *a = n == 1 ? 1.0 : 2.0; *b = n == 1 ? 0.0 : -1.0; *cs = n == 1 ? 0.5 : 1.0; };
      auto slot = [&](int n, int q) -> double* { return (n % o.every == 0) ? k.trace + 4 * n + 2 * q : nullptr; };
      if (o.kernel == "pair") {
        // array roles rotate over the four arrays: (0,1)->(2,3), (3,2)->(0,1), (1,0)->(2,3), ...
        int n = 0, cur = 0, prev = 1;
        double a1, b1, c1, a2, b2, c2;
        if (K % 2 == 1) { coef(1, &a1, &b1, &c1); W3(w3d_step_fused(k.plan, 0, a1, b1, c1, ct[1], slot(1, 0), st)); n = 1; cur = 1; prev = 0; }
        const bool faces = k.faces;
        while (n < K) {
          const int o1 = cur <= 1 ? 2 : 0, o2 = cur <= 1 ? 3 : 1;
          coef(n + 1, &a1, &b1, &c1); coef(n + 2, &a2, &b2, &c2);
          W3(w3d_step_pair(k.plan, n, cur, prev, o1, o2, a1, b1, c1, ct[n + 1], a2, b2, c2, ct[n + 2], slot(n + 1, 0), slot(n + 2, 0), st));
          if (faces) W3(w3d_step_fixup(k.plan, n + 1, o1, cur, o2, a2, b2, c2, ct[n + 2], slot(n + 2, 1), st));
          n += 2; cur = o2; prev = o1;
        }
      }
      for (int n = 1; n <= K && o.kernel != "pair"; ++n) {
        const int s = n - 1;
        const double a = n == 1 ? 1.0 : 2.0, b = n == 1 ? 0.0 : -1.0, cs = n == 1 ? 0.5 : 1.0;
        double* err = slot(n, 0);
        if (o.kernel == "fused") W3(w3d_step_fused(k.plan, s, a, b, cs, ct[n], err, st));
        else if (o.kernel == "tma") W3(w3d_step_tma(k.plan, s, a, b, cs, ct[n], nullptr, nullptr, err, 0, st));
        else W3(w3d_step_box(k.plan, s, a, b, cs, ct[n], nullptr, nullptr, err, 0, st));
      }
      CK(cudaEventRecord(e2, st));
      CK(cudaStreamSynchronize(st));
      bar.wait();
    }
    CK(cudaEventElapsedTime(&k.init_ms, e0, e1));
    CK(cudaEventElapsedTime(&k.loop_ms, e1, e2));
    k.htrace.resize(4 * (K + 1));
    CK(cudaMemcpy(k.htrace.data(), k.trace, sizeof(double) * 4 * (K + 1), cudaMemcpyDeviceToHost));
    uint32_t stt = 0;
    CK(cudaMemcpy(&stt, k.status, 4, cudaMemcpyDeviceToHost));
    if (stt) std::fprintf(stderr, "GPU %d: watchdog status 0x%x\n", r, stt);
  };
  if ((o.kernel != "fused" && o.kernel != "pair") && (G > 1 || o.periodic[0] || o.periodic[1] || o.periodic[2])) {
    std::fprintf(stderr, "--kernel %s has no halo exchange: single GPU, Dirichlet only\n", o.kernel.c_str());
    return 2;
  }
  std::vector<std::thread> th;
  for (int r = 0; r < G; ++r) th.emplace_back(worker, r);
  for (auto& t : th) t.join();
  // ---- report ----
  double cnt = (double)unk[0] * unk[1] * unk[2];
  for (int n = 1; n <= K; ++n) {
    if (n % o.every) continue;
    double mx = 0, s2 = 0;
    for (int r = 0; r < G; ++r)
      for (int q = 0; q < 2; ++q) { mx = std::max(mx, R[r].htrace[4 * n + 2 * q]); s2 += R[r].htrace[4 * n + 2 * q + 1]; }
    if (o.verbose) std::printf("Step %d, t = %.6f, Max Error = %.6e, L2 Error = %.6e\n", n, n * o.tau, mx, std::sqrt(s2 / cnt));
  }
  float loop = 0, init = 0;
  for (auto& k : R) { loop = std::max(loop, k.loop_ms); init = std::max(init, k.init_ms); }
  std::printf("Total time: %.6f s (init %.6f, compute %.6f, H2D/D2H 0.000000, exchange 0.000000 [in-kernel])\n",
              (loop + init) / 1e3, init / 1e3, loop / 1e3);
  std::printf("GPUs %d, %.3f ms/step, %.2f GLUP/s (loop), %.2f GLUP/s (total)\n", G, loop / K,
              cnt * K / (loop / 1e3) / 1e9, cnt * K / ((loop + init) / 1e3) / 1e9);
  for (auto& k : R) { cudaSetDevice(k.dev); w3d_plan_destroy(k.plan); }
  return 0;
}
