// kernels.cuh — sm_100a device code of the wave solver.
//
// What the reference's mpigpu-1.cu must have contained (K1 init,
// K2 Taylor, K3 leapfrog, K4 boundary, K5 pack/unpack, K6 error) re-designed for
// B200.  No reference source exists (readme.md:68 names the file only), so none
// of this is derived from reference code.
//
//  * `tma_role`   — z-marching stencil: one producer lane streams (TX+4)x(TY+2)
//                   planes of u^n and TXxTY planes of u^{n-1} into a shared-memory
//                   ring with cp.async.bulk.tensor (TMA) + mbarriers; consumer
//                   warps keep the z-direction in registers, update in place and
//                   fold max|err| / sum err^2 into the epilogue.
//  * fused halo exchange — CTAs whose tile touches a face that has a neighbour
//                   ("face CTAs") are numbered first, so the hardware dispatches
//                   them first.  Their producer lane acquires the neighbour's flag
//                   for the current level before the first TMA load; their consumer
//                   warps store every freshly computed face value locally AND into
//                   the neighbour's ghost layer (plain st.global on peer-mapped
//                   addresses: NVLink 5 / NVSwitch P2P); the last face CTA to finish
//                   releases the flags for the next level (st.release.sys).  The
//                   remaining CTAs stream the interior while those stores fly.
//  * `shell_compute` — one-thread-per-node path over a box list; used by the
//                   plain `box_step_kernel` (NCCL baseline's thin boxes, numerics checks)
//                   and by `fixup_kernel`, the second launch of the two-steps-per-pass
//                   scheme (pair_kernel.cuh), which also ships halos and releases flags.
#pragma once
#include "ptx.cuh"
#include "stencil_core.h"
#include "wave3d.h"

namespace w3d {

constexpr int kMaxBoxes = 6;

struct Geometry {
  int n[3];
  int xoff;
  long long pitch, plane;
  int gs;   // ghost depth - 1: rows / planes that exist in front of local index 0 (TMA maps of the
            // two-steps-per-pass kernel are encoded over the whole allocation and shift y, z by gs)
};

// Device-side exchange statistics (u64 words in the workspace, see W3D_XSTAT_* in wave3d.h):
// every face CTA reports how long it spun on the neighbours' flags and how long its halo
// copy-out took; the last face CTA of a launch folds the per-launch maxima into running totals.
// cur_*[2]: per-launch maxima, two slots used alternately (launch sequence parity) so that a
// launch can fold its predecessor's slot while its own CTAs already fill the other one.
struct XStat {
  unsigned long long cur_wait[2], cur_copy[2], sum_wait, sum_copy, launches, max_wait, max_copy;
};

__device__ __forceinline__ void xstat_fold(XStat* x, int slot) {
  const unsigned long long w = atomicExch(&x->cur_wait[slot], 0ull), c = atomicExch(&x->cur_copy[slot], 0ull);
  x->sum_wait += w;
  x->sum_copy += c;
  if (w > x->max_wait) x->max_wait = w;
  if (c > x->max_copy) x->max_copy = c;
}

struct PeerDev {
  int active;
  int ghost_index;
  double* field;          // peer array that holds the level being produced
  long long pitch, plane;
  unsigned long long* flag;
};

// Wide halo on 3-D blocks: the two-step stencil also reaches the diagonal neighbours across an edge
// (level 1 on the ghost layer of face a next to face b reads u^n on the edge line between them).
// Edge e = 4 * pair + 2 * sa + sb, pair 0 = (x,y), 1 = (x,z), 2 = (y,z); sa, sb = my sides (0 low, 1 high).
// My edge line (edge layer of both axes, all owned nodes of the third) fills the diagonal
// neighbour's ghost line (ga, gb); its incoming flag word for that edge is `flag`.
constexpr int kEdges = 12;
struct EdgeDev {
  int active;
  int ga, gb;
  double* field;          // the diagonal neighbour's array that holds the level being shipped
  long long pitch, plane;
  unsigned long long* flag;
};
__host__ __device__ __forceinline__ void edge_axes(int e, int* a, int* b, int* sa, int* sb) {
  const int pair = e >> 2;
  *a = pair == 2 ? 1 : 0;
  *b = pair == 0 ? 1 : 2;
  *sa = (e >> 1) & 1;
  *sb = e & 1;
}

struct BoxTable {
  int nboxes;
  int lo[kMaxBoxes][3];
  int ext[kMaxBoxes][3];
  long long prefix[kMaxBoxes + 1];
};

struct TileGeom {
  int lo[3], hi[3];   // box [lo,hi) in local unknown indices
  int x0;             // x index of the first tile column (<= lo[0]; chosen so that TMA boxes start 16-B aligned)
  int ntx, nty, zc, nchunks;
};

struct ErrCtl {
  double* out;            // 2 doubles or nullptr
  double* part_max;
  double* part_sum;
  unsigned* ticket;
  // cross-GPU part of the reduction: the same (max, sum) pair is also stored into this
  // rank's row of every other rank's table (peer-mapped addresses, NVLink P2P), so after
  // the time loop every rank already holds all partials — no collective, no host.
  int n_bcast;
  double* bcast[7];
};

struct alignas(64) StepParams {
  CUtensorMap map_halo;   // (TX+4)x(TY+2)x1 boxes of the current level
  CUtensorMap map_core;   // TX x TY x 1 boxes of the previous level (array updated in place)
  Geometry g;
  const double* cur;
  double* nxt;
  const double* prev_sep;  // previous level when it is NOT stored in `nxt` (fix-up pass of the pair scheme)
  const double* sx;
  const double* sy;
  const double* sz;
  double a, b, cx, cy, cz, ct;
  const double* ct_ptr;   // if non-null the time factor is read from device memory (host-fed per step)
  ErrCtl err;
  // interior
  TileGeom tile;
  // shell
  BoxTable boxes;
  int n_face_ctas;        // CTAs whose tile touches a face with a neighbour (numbered first)
  const int* order;       // block index -> tile index (face CTAs first), nullptr = identity
  int use_flags;
  PeerDev peer[6];
  EdgeDev edge[kEdges];            // wide-halo plans on 3-D blocks only (all inactive otherwise)
  const unsigned long long* flags_in;   // 6 face words, then 12 edge words
  unsigned* shell_counter;
  unsigned* status;
  XStat* xstat;
  int seq;                         // launch sequence number of this plan (parity selects the XStat slot)
  int defer_release;               // wide-halo plans: publish level `level` at kernel ENTRY (see entry_release)
  unsigned long long* tl;          // timeline slot of this launch: [0] earliest CTA start, [1] latest CTA end (or null)
  unsigned long long release;      // level published to the neighbours when the last face CTA is done
  int halo_depth;                  // layers per face copied out (1; 2 for the wide-halo pair kernel)
#ifdef W3D_DIAG
  int debug_skip;                  // diagnostic builds only (W3D_DEBUG_SKIP): 1 no flag wait, 2 no copy-out, 4 no fence/release
#endif
  unsigned long long level;        // level of `cur`
  unsigned long long timeout_ns;
};

// ---------------------------------------------------------------------------
// error reduction: warp shuffle -> CTA -> per-CTA partial -> last CTA folds all
// partials in index order (deterministic for a given launch geometry).
// ---------------------------------------------------------------------------
__device__ __forceinline__ void warp_reduce(double& mx, double& s2) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    s2 += __shfl_xor_sync(0xffffffffu, s2, o);
  }
}

// All threads of the CTA must call this. red_smem: >= 2*32 doubles + 1 int.
__device__ void publish_error(const ErrCtl& e, double mx, double s2, double* red_smem) {
  if (e.out == nullptr) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
  int* is_last = reinterpret_cast<int*>(red_smem + 64);
  warp_reduce(mx, s2);
  if (lane == 0) {
    red_smem[warp] = mx;
    red_smem[32 + warp] = s2;
  }
  __syncthreads();
  if (warp == 0) {
    mx = lane < nwarps ? red_smem[lane] : 0.0;
    s2 = lane < nwarps ? red_smem[32 + lane] : 0.0;
    warp_reduce(mx, s2);
    if (lane == 0) {
      e.part_max[blockIdx.x] = mx;
      e.part_sum[blockIdx.x] = s2;
      __threadfence();
      const unsigned t = atomicAdd(e.ticket, 1u);
      *is_last = (t == gridDim.x - 1);
    }
  }
  __syncthreads();
  if (*is_last) {
    __threadfence();
    double m = 0.0, s = 0.0;
    for (unsigned i = threadIdx.x; i < gridDim.x; i += blockDim.x) {
      m = fmax(m, __ldcg(e.part_max + i));
      s += __ldcg(e.part_sum + i);
    }
    warp_reduce(m, s);
    __syncthreads();
    if (lane == 0) {
      red_smem[warp] = m;
      red_smem[32 + warp] = s;
    }
    __syncthreads();
    if (warp == 0) {
      m = lane < nwarps ? red_smem[lane] : 0.0;
      s = lane < nwarps ? red_smem[32 + lane] : 0.0;
      warp_reduce(m, s);
      if (lane == 0) {
        e.out[0] = m;
        e.out[1] = s;
        for (int q = 0; q < e.n_bcast; ++q) {
          e.bcast[q][0] = m;
          e.bcast[q][1] = s;
        }
        *e.ticket = 0u;   // ready for the next launch on this slot
      }
    }
  }
}

// Two reductions (the two levels of a pair pass) with ONE ticket round: either ErrCtl may be
// inactive (out == nullptr).  Uses e1's ticket word when e1 is active, e2's otherwise.
// red_smem: >= 4*32 doubles; flag: one int elsewhere in shared memory.
__device__ void publish_error2(const ErrCtl& e1, const ErrCtl& e2, double mx1, double s21, double mx2, double s22,
                               double* red_smem, int* is_last) {
  const bool w1 = e1.out != nullptr, w2 = e2.out != nullptr;
  if (!w1 && !w2) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
  unsigned* ticket = w1 ? e1.ticket : e2.ticket;
  warp_reduce(mx1, s21);
  warp_reduce(mx2, s22);
  if (lane == 0) {
    red_smem[warp] = mx1; red_smem[32 + warp] = s21;
    red_smem[64 + warp] = mx2; red_smem[96 + warp] = s22;
  }
  __syncthreads();
  if (warp == 0) {
    mx1 = lane < nwarps ? red_smem[lane] : 0.0;      s21 = lane < nwarps ? red_smem[32 + lane] : 0.0;
    mx2 = lane < nwarps ? red_smem[64 + lane] : 0.0; s22 = lane < nwarps ? red_smem[96 + lane] : 0.0;
    warp_reduce(mx1, s21);
    warp_reduce(mx2, s22);
    if (lane == 0) {
      if (w1) { e1.part_max[blockIdx.x] = mx1; e1.part_sum[blockIdx.x] = s21; }
      if (w2) { e2.part_max[blockIdx.x] = mx2; e2.part_sum[blockIdx.x] = s22; }
      __threadfence();
      const unsigned t = atomicAdd(ticket, 1u);
      *is_last = (t == gridDim.x - 1);
    }
  }
  __syncthreads();
  if (*is_last) {
    __threadfence();
    double m1 = 0.0, q1 = 0.0, m2 = 0.0, q2 = 0.0;
    for (unsigned i = threadIdx.x; i < gridDim.x; i += blockDim.x) {
      if (w1) { m1 = fmax(m1, __ldcg(e1.part_max + i)); q1 += __ldcg(e1.part_sum + i); }
      if (w2) { m2 = fmax(m2, __ldcg(e2.part_max + i)); q2 += __ldcg(e2.part_sum + i); }
    }
    warp_reduce(m1, q1);
    warp_reduce(m2, q2);
    __syncthreads();
    if (lane == 0) {
      red_smem[warp] = m1; red_smem[32 + warp] = q1;
      red_smem[64 + warp] = m2; red_smem[96 + warp] = q2;
    }
    __syncthreads();
    if (warp == 0) {
      m1 = lane < nwarps ? red_smem[lane] : 0.0;      q1 = lane < nwarps ? red_smem[32 + lane] : 0.0;
      m2 = lane < nwarps ? red_smem[64 + lane] : 0.0; q2 = lane < nwarps ? red_smem[96 + lane] : 0.0;
      warp_reduce(m1, q1);
      warp_reduce(m2, q2);
      if (lane == 0) {
        if (w1) {
          e1.out[0] = m1; e1.out[1] = q1;
          for (int q = 0; q < e1.n_bcast; ++q) { e1.bcast[q][0] = m1; e1.bcast[q][1] = q1; }
        }
        if (w2) {
          e2.out[0] = m2; e2.out[1] = q2;
          for (int q = 0; q < e2.n_bcast; ++q) { e2.bcast[q][0] = m2; e2.bcast[q][1] = q2; }
        }
        *ticket = 0u;   // ready for the next launch on this slot
      }
    }
  }
}

// ---------------------------------------------------------------------------
// shell role
// ---------------------------------------------------------------------------
__device__ __forceinline__ long long gidx(const Geometry& g, int i, int j, int k) {
  return (long long)k * g.plane + (long long)j * g.pitch + g.xoff + i;
}

// first/stride: this thread's first linear node id and the stride between its nodes.
// Nodes are handled in batches of kShellBatch with all loads issued before any use:
// the shell is latency-bound (scattered 8-byte accesses), so memory-level parallelism
// per thread is what shortens the critical path in front of the flag release.
template <int kShellBatch>
__device__ void shell_compute(const StepParams& P, long long first, long long stride, long long total,
                              bool push, double& mx, double& s2) {
  const Geometry& g = P.g;
  const BoxTable& B = P.boxes;
  const bool want_err = P.err.out != nullptr;
  const double ct = P.ct_ptr ? __ldg(P.ct_ptr) : P.ct;
  for (long long base = first; base < total; base += (long long)kShellBatch * stride) {
    int ci[kShellBatch], cj[kShellBatch], ck[kShellBatch];
    long long off[kShellBatch];
    bool live[kShellBatch];
    double c0[kShellBatch], pv[kShellBatch], xm[kShellBatch], xp[kShellBatch], ym[kShellBatch],
        yp[kShellBatch], zm[kShellBatch], zp[kShellBatch];
#pragma unroll
    for (int u = 0; u < kShellBatch; ++u) {
      const long long id = base + (long long)u * stride;
      live[u] = id < total;
      int b = 0;
      while (b + 1 < B.nboxes && id >= B.prefix[b + 1]) ++b;
      long long l = live[u] ? id - B.prefix[b] : 0;
      const int ex = B.ext[b][0], ey = B.ext[b][1];
      ci[u] = B.lo[b][0] + (int)(l % ex);
      l /= ex;
      cj[u] = B.lo[b][1] + (int)(l % ey);
      ck[u] = B.lo[b][2] + (int)(l / ey);
      off[u] = gidx(g, ci[u], cj[u], ck[u]);
    }
#pragma unroll
    for (int u = 0; u < kShellBatch; ++u) {
      if (live[u]) {
        const double* c = P.cur + off[u];
        c0[u] = c[0];
        xm[u] = c[-1];
        xp[u] = c[1];
        ym[u] = c[-g.pitch];
        yp[u] = c[g.pitch];
        zm[u] = c[-g.plane];
        zp[u] = c[g.plane];
        pv[u] = P.prev_sep ? P.prev_sep[off[u]] : P.nxt[off[u]];
      }
    }
#pragma unroll
    for (int u = 0; u < kShellBatch; ++u) {
      if (!live[u]) continue;
      const int i = ci[u], j = cj[u], k = ck[u];
      const double v = w3d_node(c0[u], pv[u], xm[u], xp[u], ym[u], yp[u], zm[u], zp[u], P.a, P.b, P.cx, P.cy, P.cz);
      P.nxt[off[u]] = v;
      if (push) {
        const int co[3] = {i, j, k};
#pragma unroll
        for (int d = 0; d < 3; ++d) {
#pragma unroll
          for (int sd = 0; sd < 2; ++sd) {
            const PeerDev& pr = P.peer[d * 2 + sd];
            if (pr.active && co[d] == (sd == 0 ? 1 : g.n[d])) {
              int q[3] = {i, j, k};
              q[d] = pr.ghost_index;
              pr.field[(long long)q[2] * pr.plane + (long long)q[1] * pr.pitch + g.xoff + q[0]] = v;
            }
          }
        }
      }
      if (want_err) {
        const double e = w3d_abs_err(v, P.sx[i] * P.sy[j], P.sz[k] * ct);
        mx = fmax(mx, e);
        s2 = fma(e, e, s2);
      }
    }
  }
}

constexpr int kPipeBarrier = 2;        // named barrier id: consumers + producer warp
constexpr int kConsBarrier = 3;        // named barrier id: consumer warps only

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

#ifdef W3D_DIAG
#define W3D_DIAG_SKIP(P, bit) ((P).debug_skip & (bit))
#else
#define W3D_DIAG_SKIP(P, bit) 0
#endif

// Spin (bounded) until the neighbours named in `mask` have delivered level P.level.
// The ghost layers were written by the peer through the generic proxy (st.global) and are
// about to be read by TMA (async proxy): the acquire alone does not order those two, so the
// wait ends with a cross-proxy fence (PTX memory model, "proxy fence"; one instruction per
// waiting CTA).
// Returns the nanoseconds spent here (%globaltimer).
template <class PT>
__device__ __noinline__ unsigned long long wait_peer_flags(const PT& P, int mask) {
  const unsigned long long t_in = globaltimer_ns();
  for (int f = 0; f < 6; ++f) {
    if (!(mask & (1 << f))) continue;
    const unsigned long long t0 = globaltimer_ns();
    unsigned spins = 0;
    while (ld_acquire_sys(P.flags_in + f) < P.level) {
      if ((++spins & 1023u) == 0 && globaltimer_ns() - t0 > P.timeout_ns) {
        atomicExch(P.status, 0x100u + (unsigned)f);   // watchdog: neighbour never arrived
        break;
      }
    }
  }
  // edge lines of the diagonal neighbours: needed by a tile that reads the ghost layers of both faces
  for (int e = 0; e < kEdges; ++e) {
    if (!P.edge[e].active) continue;
    int a, b, sa, sb;
    edge_axes(e, &a, &b, &sa, &sb);
    if (!(mask & (1 << (a * 2 + sa))) || !(mask & (1 << (b * 2 + sb)))) continue;
    const unsigned long long t0 = globaltimer_ns();
    unsigned spins = 0;
    while (ld_acquire_sys(P.flags_in + 6 + e) < P.level) {
      if ((++spins & 1023u) == 0 && globaltimer_ns() - t0 > P.timeout_ns) {
        atomicExch(P.status, 0x200u + (unsigned)e);
        break;
      }
    }
  }
  fence_proxy_async_global();
  return globaltimer_ns() - t_in;
}

// Called by ONE thread of a face CTA after all of the CTA's stores are ordered before
// it (CTA-scope barrier).  The last face CTA publishes P.release to every neighbour.
// wait_ns / t_copy0: this CTA's flag-wait time and the %globaltimer stamp taken before its
// copy-out (exchange statistics, only touched when P.xstat is set).
template <class PT>
__device__ __noinline__ void face_cta_done(const PT& P, unsigned long long wait_ns = 0ull,
                                           unsigned long long t_copy0 = 0ull) {
  __threadfence_system();
  const int slot = P.seq & 1;
  if (P.xstat) {
    atomicMax(&P.xstat->cur_wait[slot], wait_ns);
    if (t_copy0) atomicMax(&P.xstat->cur_copy[slot], globaltimer_ns() - t_copy0);
    __threadfence();
  }
  const unsigned done = atomicAdd(P.shell_counter, 1u);
  if (done == (unsigned)P.n_face_ctas - 1u) {
    *P.shell_counter = 0u;
    __threadfence_system();
    for (int f = 0; f < 6; ++f)
      if (P.peer[f].active) st_release_sys(P.peer[f].flag, P.release);
    for (int e = 0; e < kEdges; ++e)
      if (P.edge[e].active) st_release_sys(P.edge[e].flag, P.release);
    if (P.xstat) {
      xstat_fold(P.xstat, slot);
      if (!P.defer_release) P.xstat->launches += 1ull;     // (deferred plans count at kernel entry)
    }
  }
}

// Wide-halo plans: the halo layers of level `level` were pushed into the neighbours by the PREVIOUS
// launch of this stream, and a kernel boundary lies in between — its completion makes those peer
// stores visible system-wide.  So the neighbours are told at the ENTRY of the next launch, by one
// thread, instead of at the end of the producing one by a fence + counter + release in every face
// CTA (measured: 4.8 us at the end of every pass on a 511 x 511 x 64 block, all CTAs in lock step).
// Every launch of such a plan starts with this, so a kernel that waits for level s on one rank
// always has a partner on the neighbouring rank that publishes level s.
template <class PT>
__device__ __noinline__ void entry_release(const PT& P) {
  __threadfence_system();                        // ONE system-scope fence, then relaxed stores = release pattern
  for (int f = 0; f < 6; ++f)
    if (P.peer[f].active) st_relaxed_sys(P.peer[f].flag, P.level);
  for (int e = 0; e < kEdges; ++e)
    if (P.edge[e].active) st_relaxed_sys(P.edge[e].flag, P.level);
  if (P.xstat) {
    xstat_fold(P.xstat, (P.seq & 1) ^ 1);      // the previous launch's maxima
    P.xstat->launches += 1ull;
  }
}

// ---------------------------------------------------------------------------
// TMA z-marching role
// ---------------------------------------------------------------------------
template <int TX, int RPT, int RG, int STAGES>
struct TileCfg {
  static constexpr int kTX = TX, kRPT = RPT, kRG = RG, kStages = STAGES;
  static constexpr int kTY = RPT * RG;
  // The innermost TMA coordinate must be a multiple of 16 bytes (2 doubles) — an odd
  // start faults with 'illegal instruction' on sm_100 (measured) — so the
  // halo box carries two ghost columns on each side instead of one.
  static constexpr int kW = TX + 4;
  static constexpr int kHaloRows = kTY + 2;
  static constexpr int kConsumers = TX * RG;
  static constexpr int kThreads = kConsumers + 32;   // consumers + producer warp
  static constexpr int kHaloBytes = kW * kHaloRows * 8;
  static constexpr int kHaloBytesPad = (kHaloBytes + 127) / 128 * 128;
  static constexpr int kCoreBytes = TX * kTY * 8;
  static constexpr int kStageBytes = kHaloBytesPad + kCoreBytes;
  static constexpr int kBarBytes = 128 + 1024;   // barriers + reduction scratch
  static constexpr int kSmemBytes = kBarBytes + STAGES * kStageBytes + 128;
  static_assert(TX % 32 == 0 && TX <= 252, "TX must be a warp multiple and fit a TMA box");
  static_assert(kHaloRows <= 256, "TMA box limit");
  static_assert(2 * STAGES * 8 <= 128, "barrier area");
};

// Faces with a neighbour that the tile [lo,hi) comes within `reach` layers of: bit axis*2+side.
// reach 1: the tile owns nodes of the edge layer (1 or n) — it reads the ghost layer and is the
// one that copies the edge layer out.  reach 2 (two steps per pass): the tile's level-1 ring
// reaches the edge layer, so it reads the ghost layer although it may own no edge node.
// The host mirrors this in face_mask_host() (wave_api.cu).
template <class PT>
__device__ __forceinline__ int tile_face_mask(const PT& P, int x0, int y0, int z0, int z1, int tx_w,
                                              int ty_w, int reach = 1) {
  if (!P.use_flags) return 0;
  int m = 0;
  const int lo[3] = {x0, y0, z0}, hi[3] = {x0 + tx_w, y0 + ty_w, z1};
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    if (P.peer[d * 2].active && lo[d] <= reach && hi[d] > 1) m |= 1 << (d * 2);
    if (P.peer[d * 2 + 1].active && lo[d] <= P.g.n[d] && hi[d] > P.g.n[d] - reach + 1) m |= 1 << (d * 2 + 1);
  }
  return m;
}

// After a face CTA has written its tile, its consumer threads copy the tile's face
// layers out of the local array (just written: L2 hits) into the neighbours' ghost
// layers.  Keeping this out of the streaming loop leaves that loop identical for face
// and interior CTAs.  tid/nthreads: consumer thread index and count.
// P.halo_depth layers per face: layer 1+m / n-m of mine fills the neighbour's ghost layer
// ghost_index+m / ghost_index-m (m = 0 .. depth-1); a tile copies the layers it owns.
template <class PT>
__device__ __noinline__ void face_copy_out(const PT& P, int mask, int x0, int y0, int z0, int z1, int tx_w, int ty_w,
                              int tid, int nthreads) {
  const Geometry& g = P.g;
  const TileGeom& T = P.tile;
  // clip the tile to the box it updates
  const int xl = max(x0, T.lo[0]), xh = min(x0 + tx_w, T.hi[0]);
  const int yl = max(y0, T.lo[1]), yh = min(y0 + ty_w, T.hi[1]);
  const int ex = xh - xl, ey = yh - yl, ez = z1 - z0;
  if (ex <= 0 || ey <= 0 || ez <= 0) return;
  const int tlo[3] = {xl, yl, z0}, thi[3] = {xh, yh, z1};
#pragma unroll 1
  for (int fm = 0; fm < 6 * P.halo_depth; ++fm) {
    const int f = fm % 6, m = fm / 6;
    if (!(mask & (1 << f))) continue;
    const PeerDev& pr = P.peer[f];
    const int d = f >> 1;
    const int layer = (f & 1) ? g.n[d] - m : 1 + m;   // my m-th layer from that face along axis d
    if (layer < tlo[d] || layer >= thi[d]) continue;  // another tile owns it
    const int ghost = pr.ghost_index + ((f & 1) ? -m : m);
    // iterate the two other axes, fastest first
    const int e0 = d == 0 ? ey : ex;                    // fast extent
    const int e1 = d == 2 ? ey : ez;                    // slow extent
    const int total = e0 * e1;
    for (int base = tid; base < total; base += 4 * nthreads) {
      double v[4];
      long long dst[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int id = base + u * nthreads;
        if (id < total) {
          const int a = id % e0, b = id / e0;
          int q[3];
          if (d == 0) { q[0] = layer; q[1] = yl + a; q[2] = z0 + b; }
          else if (d == 1) { q[0] = xl + a; q[1] = layer; q[2] = z0 + b; }
          else { q[0] = xl + a; q[1] = yl + b; q[2] = layer; }
          v[u] = __ldcg(P.nxt + gidx(g, q[0], q[1], q[2]));
          q[d] = ghost;
          dst[u] = (long long)q[2] * pr.plane + (long long)q[1] * pr.pitch + g.xoff + q[0];
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (base + u * nthreads < total) pr.field[dst[u]] = v[u];   // peer-mapped: goes over NVLink
    }
  }
}

// Edge lines (wide halo on 3-D blocks): a tile that owns nodes of the line where the edge layers of
// two faces with neighbours meet copies them into the diagonal neighbour's ghost line.
template <class PT>
__device__ __noinline__ void edge_copy_out(const PT& P, int x0, int y0, int z0, int z1, int tx_w, int ty_w, int tid,
                                           int nthreads) {
  const Geometry& g = P.g;
  const TileGeom& T = P.tile;
  const int tlo[3] = {max(x0, T.lo[0]), max(y0, T.lo[1]), z0};
  const int thi[3] = {min(x0 + tx_w, T.hi[0]), min(y0 + ty_w, T.hi[1]), z1};
#pragma unroll 1
  for (int e = 0; e < kEdges; ++e) {
    const EdgeDev& ed = P.edge[e];
    if (!ed.active) continue;
    int a, b, sa, sb;
    edge_axes(e, &a, &b, &sa, &sb);
    const int c = 3 - a - b;
    const int la = sa ? g.n[a] : 1, lb = sb ? g.n[b] : 1;
    if (la < tlo[a] || la >= thi[a] || lb < tlo[b] || lb >= thi[b]) continue;
    for (int i = tlo[c] + tid; i < thi[c]; i += nthreads) {
      int q[3];
      q[a] = la; q[b] = lb; q[c] = i;
      const double v = __ldcg(P.nxt + gidx(g, q[0], q[1], q[2]));
      q[a] = ed.ga; q[b] = ed.gb;
      ed.field[(long long)q[2] * ed.plane + (long long)q[1] * ed.pitch + g.xoff + q[0]] = v;   // peer-mapped: NVLink
    }
  }
}

template <class Cfg>
__device__ __forceinline__ void tma_role(const StepParams& P, int x0, int y0, int z0, int z1, int mask, uint8_t* smem,
                         double& mx, double& s2) {
  constexpr int TX = Cfg::kTX, RPT = Cfg::kRPT, STAGES = Cfg::kStages, W = Cfg::kW;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);
  uint64_t* empty = full + STAGES;
  uint8_t* stage0 = smem + Cfg::kBarBytes;

  const TileGeom& T = P.tile;
  const int nzc = z1 - z0;
  const int iters = nzc + 2;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full + s, 1);
      mbar_init(empty + s, Cfg::kConsumers / 32);
    }
    fence_barrier_init();
  }
  __syncthreads();

  const int warp = threadIdx.x >> 5;
  if (warp == Cfg::kConsumers / 32) {
    // ===== producer warp: one lane drives TMA =====
    // The whole warp walks the loop (only lane 0 issues) so that it reaches the
    // CTA-wide barrier after this role converged: an aligned bar.sync executed by
    // a diverged warp faults on sm_100.
    const bool leader = (threadIdx.x & 31) == 0;
    if (leader) {
      tma_prefetch_desc(&P.map_halo);
      tma_prefetch_desc(&P.map_core);
      // ghosts of the current level must have arrived before they are loaded
      if (mask && !W3D_DIAG_SKIP(P, 1)) {
        const unsigned long long w = wait_peer_flags(P, mask);
        *reinterpret_cast<unsigned long long*>(smem + 128 + 1016) = w;   // handed to consumer thread 0 (read after the CTA barriers below)
      }
    }
    for (int t = 0; t < iters; ++t) {
      if (leader) {
        const int s = t % STAGES;
        const uint32_t ph = (uint32_t)(t / STAGES) & 1u;
        mbar_wait(empty + s, ph ^ 1u);
        const bool need_prev = (t >= 1 && t <= nzc);
        uint8_t* st = stage0 + (size_t)s * Cfg::kStageBytes;
        mbar_arrive_expect_tx(full + s, (uint32_t)(Cfg::kHaloBytes + (need_prev ? Cfg::kCoreBytes : 0)));
        const int zz = z0 - 1 + t;
        tma_load_3d(st, &P.map_halo, full + s, P.g.xoff + x0 - 2, y0 - 1, zz);
        if (need_prev) tma_load_3d(st + Cfg::kHaloBytesPad, &P.map_core, full + s, P.g.xoff + x0, y0, zz);
      }
      __syncwarp();
    }
    return;
  }

  // ===== consumers =====
  const int cxi = threadIdx.x % TX;           // column inside the tile
  const int r0 = (threadIdx.x / TX) * RPT;    // first row of this thread
  const int gx = x0 + cxi;                    // local x index
  const bool x_ok = gx >= T.lo[0] && gx < T.hi[0];
  const bool want_err = P.err.out != nullptr;
  const double ct = P.ct_ptr ? __ldg(P.ct_ptr) : P.ct;
  double acc[RPT], zm[RPT], sxy[RPT];
  bool ok[RPT];
  const double sxv = x_ok ? P.sx[gx] : 0.0;
#pragma unroll
  for (int r = 0; r < RPT; ++r) {
    const int gy = y0 + r0 + r;
    ok[r] = x_ok && gy < T.hi[1];
    sxy[r] = ok[r] ? sxv * P.sy[gy] : 0.0;
    acc[r] = 0.0;
    zm[r] = 0.0;
  }
  double* out_col = P.nxt + (long long)(y0 + r0) * P.g.pitch + P.g.xoff + gx;
  for (int t = 0; t < iters; ++t) {
    const int s = t % STAGES;
    const uint32_t ph = (uint32_t)(t / STAGES) & 1u;
    mbar_wait(full + s, ph);
    const double* cs = reinterpret_cast<const double*>(stage0 + (size_t)s * Cfg::kStageBytes);
    const double* ps = reinterpret_cast<const double*>(stage0 + (size_t)s * Cfg::kStageBytes + Cfg::kHaloBytesPad);
    double col[RPT + 2];
#pragma unroll
    for (int q = 0; q < RPT + 2; ++q) col[q] = cs[(r0 + q) * W + cxi + 2];

    if (t >= 2) {
      // plane zo = z0 + t - 2 is complete: add the z+1 neighbour, store, fold the error
      const int zo = z0 + t - 2;
      double* o = out_col + (long long)zo * P.g.plane;
      // (one global load per plane; staging the table in shared memory was measured 1 % slower here:
      // this kernel is DRAM-bound, the extra prologue is not hidden)
      const double szt = want_err ? P.sz[zo] * ct : 0.0;
#pragma unroll
      for (int r = 0; r < RPT; ++r) {
        const double v = w3d_finish(acc[r], col[r + 1], P.cz);
        if (ok[r]) {
          o[(long long)r * P.g.pitch] = v;
          if (want_err) {
            const double e = w3d_abs_err(v, sxy[r], szt);
            mx = fmax(mx, e);
            s2 = fma(e, e, s2);
          }
        }
      }
    }
    if (t >= 1 && t <= nzc) {
#pragma unroll
      for (int r = 0; r < RPT; ++r) {
        const double xm = cs[(r0 + r + 1) * W + cxi + 1];
        const double xp = cs[(r0 + r + 1) * W + cxi + 3];
        const double pv = ps[(r0 + r) * TX + cxi];
        acc[r] = w3d_partial(col[r + 1], pv, xm, xp, col[r], col[r + 2], zm[r], P.a, P.b, P.cx, P.cy, P.cz);
      }
    }
#pragma unroll
    for (int r = 0; r < RPT; ++r) zm[r] = col[r + 1];
    __syncwarp();
    if ((threadIdx.x & 31) == 0) mbar_arrive(empty + s);
  }
  if (mask) {
    named_bar_sync(kConsBarrier, Cfg::kConsumers);   // the tile is written (CTA-visible)
    const unsigned long long t_copy0 = P.xstat ? globaltimer_ns() : 0ull;
    if (!W3D_DIAG_SKIP(P, 2)) {
      face_copy_out(P, mask, x0, y0, z0, z1, TX, Cfg::kTY, threadIdx.x, Cfg::kConsumers);
      edge_copy_out(P, x0, y0, z0, z1, TX, Cfg::kTY, threadIdx.x, Cfg::kConsumers);
    }
    named_bar_sync(kConsBarrier, Cfg::kConsumers);   // every consumer's peer stores are issued
    if (threadIdx.x == 0 && !W3D_DIAG_SKIP(P, 4))
      face_cta_done(P, *reinterpret_cast<volatile unsigned long long*>(smem + 128 + 1016), t_copy0);
  }
}

// ---------------------------------------------------------------------------
// kernels
// ---------------------------------------------------------------------------
template <class Cfg, int MINB>
__global__ void __launch_bounds__(Cfg::kThreads, MINB)
fused_step_kernel(const __grid_constant__ StepParams P) {
  // Indexed directly (no integer round trip) so the compiler keeps the shared address space:
  // LDS/STS with 32-bit addresses instead of generic LD/ST with 64-bit address arithmetic.
  extern __shared__ __align__(1024) uint8_t smem[];
  double* red = reinterpret_cast<double*>(smem + 128);
  double mx = 0.0, s2 = 0.0;
  if (P.tl && threadIdx.x == 0) atomicMin(P.tl, globaltimer_ns());
  if (P.defer_release && P.use_flags && blockIdx.x == 0 && threadIdx.x == 0) entry_release(P);
  const TileGeom& T = P.tile;
  const int work = P.order ? P.order[blockIdx.x] : (int)blockIdx.x;
  const int tx = work % T.ntx;
  const int ty = (work / T.ntx) % T.nty;
  const int ch = work / (T.ntx * T.nty);
  const int x0 = T.x0 + tx * Cfg::kTX;
  const int y0 = T.lo[1] + ty * Cfg::kTY;
  const int z0 = T.lo[2] + ch * T.zc;
  const int z1 = min(z0 + T.zc, T.hi[2]);
  // tiles that own one of the halo_depth outermost layers of a face with a neighbour (1 layer; 2 in
  // plans that keep a wide halo for the two-steps-per-pass kernel): they wait, copy out, are counted
  const int mask = tile_face_mask(P, x0, y0, z0, z1, Cfg::kTX, Cfg::kTY, P.halo_depth);
  tma_role<Cfg>(P, x0, y0, z0, z1, mask, smem, mx, s2);
  if (P.err.out != nullptr) {
    __syncthreads();
    publish_error(P.err, mx, s2, red);
  }
  if (P.tl && threadIdx.x == 0) atomicMax(P.tl + 1, globaltimer_ns());
}

// plain one-thread-per-node kernel over the box table (no communication)
__global__ void __launch_bounds__(256) box_step_kernel(const __grid_constant__ StepParams P) {
  __shared__ double red[66];
  double mx = 0.0, s2 = 0.0;
  shell_compute<4>(P, (long long)blockIdx.x * blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x,
                   P.boxes.prefix[P.boxes.nboxes], false, mx, s2);
  publish_error(P.err, mx, s2, red);
}

// Fix-up pass of the two-step scheme: the layer next to every face with a neighbour gets
// its second level once the neighbour's first-level edge has arrived.  One thread per node
// over the shell boxes; acquires the flags for P.level, stores locally and into the
// neighbours' ghost layers, and the last CTA releases P.level + 1.
// One wave of 1024-thread CTAs (<= one per SM): a single flag wait, one batch of loads and one
// system fence per CTA; with many small CTAs the per-CTA fence latency serialised over waves.
__global__ void __launch_bounds__(1024, 1) fixup_kernel(const __grid_constant__ StepParams P) {
  __shared__ double red[66];
  double mx = 0.0, s2 = 0.0;
  unsigned long long wait_ns = 0ull, t0 = 0ull;
  if (P.tl && threadIdx.x == 0) atomicMin(P.tl, globaltimer_ns());
  if (P.use_flags) {
    if (threadIdx.x == 0) {
      int mask = 0;
      for (int f = 0; f < 6; ++f)
        if (P.peer[f].active) mask |= 1 << f;
      wait_ns = wait_peer_flags(P, mask);
      t0 = globaltimer_ns();
    }
    __syncthreads();
  }
  shell_compute<2>(P, (long long)blockIdx.x * blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x,
                   P.boxes.prefix[P.boxes.nboxes], P.use_flags != 0, mx, s2);
  if (P.use_flags) {
    __syncthreads();
    if (threadIdx.x == 0) face_cta_done(P, wait_ns, t0);    // (the whole shell pass counts as halo copy-out time)
  }
  publish_error(P.err, mx, s2, red);
}

// K1: analytic layer 0 on every node (ghost layers — g.gs + 1 of them — and row padding
// included) of `u`, zeros in `z` (the array that will receive level 1; may be null when nobody
// reads it before it is overwritten): one warp per row, coalesced stores.  With two ghost
// layers the tables are read at indices -1 and n+2 as well.
__global__ void __launch_bounds__(256) init_kernel(double* __restrict__ u, double* __restrict__ z, Geometry g,
                                                   const double* __restrict__ sx, const double* __restrict__ sy,
                                                   const double* __restrict__ sz) {
  const int gs = g.gs;
  const int rows_per_plane = g.n[1] + 2 + 2 * gs;
  const long long rows = (long long)rows_per_plane * (g.n[2] + 2 + 2 * gs);
  const int lane = threadIdx.x & 31;
  const long long warp0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  const int pitch = (int)g.pitch;
  for (long long row = warp0; row < rows; row += nwarps) {
    const int j = (int)(row % rows_per_plane) - gs;
    const int k = (int)(row / rows_per_plane) - gs;
    const double syz = sy[j] * sz[k];
    const long long off = (long long)k * g.plane + (long long)j * g.pitch;
    double* ur = u + off;
    double* zr = z + off;
    for (int xe = lane; xe < pitch; xe += 32) {
      const int i = xe - g.xoff;
      double v = 0.0;
      if (i >= -gs && i <= g.n[0] + 1 + gs) v = sx[i] * syz;
      ur[xe] = v;
      if (z != nullptr) zr[xe] = 0.0;
    }
  }
}

// K6 stand-alone: error norms of a field
__global__ void __launch_bounds__(256) error_kernel(const double* u, Geometry g, const double* sx,
                                                    const double* sy, const double* sz, double ct,
                                                    ErrCtl e) {
  __shared__ double red[66];
  double mx = 0.0, s2 = 0.0;
  const long long rows = (long long)g.n[1] * g.n[2];
  // one warp per row keeps loads coalesced
  const int lane = threadIdx.x & 31;
  const long long warp0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long r = warp0; r < rows; r += nwarps) {
    const int j = 1 + (int)(r % g.n[1]);
    const int k = 1 + (int)(r / g.n[1]);
    const double* row = u + gidx(g, 0, j, k);
    const double syj = sy[j], szt = sz[k] * ct;
    for (int i = 1 + lane; i <= g.n[0]; i += 32) {
      const double e1 = w3d_abs_err(row[i], sx[i] * syj, szt);
      mx = fmax(mx, e1);
      s2 = fma(e1, e1, s2);
    }
  }
  publish_error(e, mx, s2, red);
}

// K5 (baseline only): face pack / unpack.  layer = index along `axis` to read/write.
__global__ void __launch_bounds__(256) face_kernel(double* u, Geometry g, int axis, int layer,
                                                   double* buf, int unpack) {
  const int a = (axis + 1) % 3, c = (axis + 2) % 3;   // same enumeration as the CPU library
  // buffer order: fastest index first over the two other axes in (x,y,z) order
  int f0, f1;   // fast, slow axes
  if (axis == 0) { f0 = 1; f1 = 2; } else if (axis == 1) { f0 = 0; f1 = 2; } else { f0 = 0; f1 = 1; }
  (void)a; (void)c;
  const long long total = (long long)g.n[f0] * g.n[f1];
  for (long long id = (long long)blockIdx.x * blockDim.x + threadIdx.x; id < total;
       id += (long long)gridDim.x * blockDim.x) {
    int q[3];
    q[axis] = layer;
    q[f0] = 1 + (int)(id % g.n[f0]);
    q[f1] = 1 + (int)(id / g.n[f0]);
    const long long o = gidx(g, q[0], q[1], q[2]);
    if (unpack) u[o] = buf[id]; else buf[id] = u[o];
  }
}

// periodic self-wrap of one axis: ghost <- opposite edge, for both sides
__global__ void __launch_bounds__(256) wrap_kernel(double* u, Geometry g, int axis) {
  int f0, f1;
  if (axis == 0) { f0 = 1; f1 = 2; } else if (axis == 1) { f0 = 0; f1 = 2; } else { f0 = 0; f1 = 1; }
  const int e0 = g.n[f0] + 2, e1 = g.n[f1] + 2;   // ghosts of the other axes too
  const long long total = (long long)e0 * e1;
  for (long long id = (long long)blockIdx.x * blockDim.x + threadIdx.x; id < total;
       id += (long long)gridDim.x * blockDim.x) {
    int q[3];
    q[f0] = (int)(id % e0);
    q[f1] = (int)(id / e0);
    q[axis] = g.n[axis];
    const double hi = u[gidx(g, q[0], q[1], q[2])];
    q[axis] = 1;
    const double lo = u[gidx(g, q[0], q[1], q[2])];
    q[axis] = 0;
    u[gidx(g, q[0], q[1], q[2])] = hi;
    q[axis] = g.n[axis] + 1;
    u[gidx(g, q[0], q[1], q[2])] = lo;
  }
}

}  // namespace w3d
