// pair_kernel.cuh — temporal blocking: TWO time steps per pass over memory.
//
// The single-step kernel moves the algorithmic minimum of 24 B per node and step
// (read u^n, read u^{n-1}, write u^{n+1}) and runs at the HBM roofline, so the only
// way to go faster is to move fewer bytes.  This kernel advances two levels in one
// sweep: it reads u^n and u^{n-1} once and writes u^{n+1} and u^{n+2} once —
// 32 B per node for two steps = 16 B per node and step.
//
// The reference does one halo exchange and one pass per step (report.pdf p.16 §4.4:
// exchange is an additive per-step phase); temporal blocking is listed as absent in
// SURVEY.md §2.4, so this has no counterpart there.
//
// Structure per CTA (tile TX x TY, chunk of planes [z0,z1)), marching up in z:
//   * TMA streams, per plane p, a (TX+4)x(TY+4) box of u^n (A) and a (TX+4)x(TY+2)
//     box of u^{n-1} (B) into the shared-memory ring (same producer/mbarrier scheme
//     as the single-step kernel);
//   * level 1 (u^{n+1}) is computed on the tile PLUS a one-node ring around it
//     ((TX+2)x(TY+2); ring nodes are owned by the first 2(TX+2)+2TY threads), with
//     the z direction in registers (partial/finish), written to global C for tile
//     nodes and to a double-buffered shared-memory plane for everyone;
//   * one consumer barrier per plane, then level 2 (u^{n+2}) is computed on the tile
//     from that shared plane (x/y neighbours) and registers (z neighbours, and u^n
//     as the "previous" level), and written to global D;
//   * both levels can feed the error epilogue.
// Level-1 values on ring nodes outside the local block are taken as 0 (Dirichlet
// boundary).  Faces with a neighbour (another GPU, or this one across a periodic axis)
// come in two protocols, chosen by the host (wave_api.cu):
//   * wide halo (default, one launch and one flag handshake per two steps): the arrays
//     carry TWO ghost layers.  u^n arrives two layers deep, so level 1 is also formed on
//     the first ghost layer (redundantly with the neighbour, bit-identical) and kept in
//     out1's ghost layer as the next pass's u^{n-1} halo; level 2 is then complete on
//     every owned node and the face CTAs ship its two outermost layers per face.  Same
//     halo volume per step as one layer per step.
//     On 3-D blocks the two-step stencil also reaches the diagonal neighbours across the
//     block's edges; with their edge lines (w3d_plan_set_edges, edge_copy_out) blocks run
//     the same protocol.
//   * fix-up (one ghost layer, or blocks without edge peers): level 2 is left out next to faces
//     with a neighbour (skip_lo/hi) and finished by `fixup_kernel` after an exchange of level 1.
// The arithmetic is the same w3d_partial/w3d_finish sequence, so results are
// bit-identical to two single steps.
#pragma once
#include "kernels.cuh"


namespace w3d {

// CLY = 2: thread-block clusters of two CTAs whose tiles are neighbours in y.  Each CTA leaves out the
// level-1 ring row on its partner's side and reads those values — the partner's first / last tile
// row of level 1 — from the partner's shared memory (DSMEM) instead; the per-plane CTA barrier
// becomes a cluster barrier.  Measured against the plain variant in profiles/cluster_variant.md.
template <int TX, int RPT, int RG, int STAGES, int MINB = 1, int CLY = 1>
struct PairCfg {
  static constexpr int kClusterY = CLY;
  static constexpr int kMinBlocks = MINB;                  // CTAs per SM the launch bounds ask for
  static constexpr int kTX = TX, kRPT = RPT, kRG = RG, kStages = STAGES;
  static constexpr int kTY = RPT * RG;
  static constexpr int kWA = TX + 4, kRowsA = kTY + 4;     // u^n box (two ghost layers)
  static constexpr int kWB = TX + 4, kRowsB = kTY + 2;     // u^{n-1} box (one ghost layer; x padded for alignment)
  static constexpr int kW1 = TX + 2, kRows1 = kTY + 2;     // level-1 plane kept in shared memory
  static constexpr int kConsumers = TX * RG;
  // no separate producer warp: thread 0 issues the TMA loads from inside the consumer loop.
  // 256 threads = 8 warps lifts the per-thread register ceiling from 168 (288 threads are
  // allocated as 12 warps) to 255, which this kernel needs to stay spill-free.
  static constexpr int kThreads = kConsumers;
  static constexpr int kABytes = (kWA * kRowsA * 8 + 127) / 128 * 128;
  static constexpr int kBBytes = (kWB * kRowsB * 8 + 127) / 128 * 128;
  static constexpr int kATx = kWA * kRowsA * 8, kBTx = kWB * kRowsB * 8;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kL1Bytes = (kW1 * kRows1 * 8 + 127) / 128 * 128;
  static constexpr int kSzEntries = 640;                   // staged slice of the z table (chunk + 4 planes)
  static constexpr int kBarBytes = 128 + 1024 + kSzEntries * 8;
  static constexpr int kSmemBytes = kBarBytes + STAGES * kStageBytes + 2 * kL1Bytes + 128;
  static constexpr int kRing = 2 * kW1 + 2 * kTY;
  static_assert(kRing <= kConsumers, "one ring node per consumer thread at most");
  static_assert(TX % 32 == 0 && kWA <= 256 && kRowsA <= 256, "TMA box limits");
};

struct alignas(64) PairParams {
  CUtensorMap map_a;      // (TX+4)x(TY+4)x1 boxes of u^n
  CUtensorMap map_b;      // (TX+4)x(TY+2)x1 boxes of u^{n-1}
  Geometry g;
  double* out1;           // u^{n+1}
  double* out2;           // u^{n+2}
  const double* sx;
  const double* sy;
  const double* sz;
  double a1, b1, c1x, c1y, c1z, ct1;   // level n+1 coefficients / time factor
  double a2, b2, c2x, c2y, c2z, ct2;   // level n+2
  const double* ct_ptr;                // if non-null: ct1 = ct_ptr[0], ct2 = ct_ptr[1] (host-fed device table)
  int skip_b;                          // b1 == 0: u^{n-1} is not read at all (its smem boxes are zero-filled once)
  ErrCtl err1, err2;
  TileGeom tile;
  const int* order;
  // fused halo exchange of level n+1 (same protocol and member names as StepParams, so the
  // face helpers in kernels.cuh work on both): face CTAs acquire `level` before loading,
  // copy the level-(n+1) face layers of `nxt` (= out1) into the neighbours and the last
  // one releases level+1.
  double* nxt;            // array whose face layers are shipped: out1 (fix-up protocol) or out2 (wide halo)
  int n_face_ctas;
  int use_flags;
  PeerDev peer[6];
  EdgeDev edge[kEdges];   // wide halo on 3-D blocks: the diagonal neighbours across the edges (all inactive otherwise)
  const unsigned long long* flags_in;
  unsigned* shell_counter;
  unsigned* status;
  XStat* xstat;
  int seq;                  // launch sequence number of this plan (parity selects the XStat slot)
  int defer_release;        // wide halo: level `level` is published at kernel entry (entry_release, kernels.cuh)
  unsigned long long* tl;   // timeline slot of this launch: [0] earliest CTA start, [1] latest CTA end (or null)
  unsigned long long level;
  unsigned long long release;
  unsigned long long timeout_ns;
  int halo_depth;         // 1 (fix-up protocol) or 2 (wide halo)
  // nodes of level 2 that must NOT be produced here: the layer next to a face with a
  // neighbour (their level-1 ghost is not known yet).  skip_lo/hi[d] = 1 when active.
  // Fix-up protocol only; all zero with the wide halo.
  int skip_lo[3], skip_hi[3];
  // wide halo: level 1 is meaningful on local indices in_lo[d] .. in_hi[d] (0 / n+1 included
  // where that face has a neighbour; 1 .. n otherwise), and is stored on the ghost layer too
  int in_lo[3], in_hi[3];
  int wide;
  // wide halo, z faces: the shipped planes of level n+2 go to the neighbour straight from the
  // registers that hold them when they are stored locally (bit 0: low face, bit 1: high face), so
  // nothing but the fence and the flag release is left for the end of the pass.  x / y faces
  // (strided) are copied out after the loop by face_copy_out.
  int zpush;
};

// Consumer side of one CTA.  FULL = every node of the tile is inside the block, is
// updated by this launch and is not excluded by a skip flag: all per-node predicates
// fold away and the eight rows of a thread become straight-line code that the compiler
// can interleave (the dependent FMA chains of one row otherwise expose their latency).
// one thread: arm the stage's barrier and issue the TMA loads of iteration t (plane z0-2+t)
template <class Cfg>
__device__ __forceinline__ void pair_issue(const PairParams& P, uint64_t* full, uint8_t* stage0, int t, int x0,
                                           int y0, int z0, int nzc) {
  const int s = t % Cfg::kStages;
  const int p = z0 - 2 + t;
  // planes z0-1 .. z1 of u^{n-1}; none at all when its coefficient is zero (Taylor first step)
  const bool need_b = (t >= 1 && t <= nzc + 2) && !P.skip_b;
  uint8_t* st = stage0 + (size_t)s * Cfg::kStageBytes;
  mbar_arrive_expect_tx(full + s, (uint32_t)(Cfg::kATx + (need_b ? Cfg::kBTx : 0)));
  const int gs = P.g.gs;   // the maps span the whole allocation: gs ghost rows / planes sit in front of index 0
  tma_load_3d(st, &P.map_a, full + s, P.g.xoff + x0 - 2, y0 - 2 + gs, p + gs);
  if (need_b) tma_load_3d(st + Cfg::kABytes, &P.map_b, full + s, P.g.xoff + x0 - 2, y0 - 1 + gs, p + gs);
}

template <class Cfg, bool FULL, bool XFULL>
__device__ __forceinline__ void pair_consumer(const PairParams& P, uint64_t* full, uint8_t* stage0,
                                              double* l1buf, const double* s_sz, int x0, int y0, int yown, int cl_ov,
                                              int z0, int z1, double& mx1, double& s21, double& mx2, double& s22) {
  constexpr int TX = Cfg::kTX, RPT = Cfg::kRPT, STAGES = Cfg::kStages;
  constexpr int WA = Cfg::kWA, WB = Cfg::kWB, W1 = Cfg::kW1;
  const TileGeom& T = P.tile;
  const Geometry& g = P.g;
  const int nzc = z1 - z0;
  const int iters = nzc + 4;
  const int tid = threadIdx.x;
  const int cxi = tid % TX;
  const int r0 = (tid / TX) * RPT;
  const int gx = x0 + cxi;
  const bool x_in = gx >= P.in_lo[0] && gx <= P.in_hi[0];    // level 1 exists here (block, or first ghost layer of a wide halo)
  const bool x_ok = gx >= T.lo[0] && gx < T.hi[0];           // this launch updates it
  const bool want1 = P.err1.out != nullptr, want2 = P.err2.out != nullptr;
  const double ct1 = P.ct_ptr ? __ldg(P.ct_ptr) : P.ct1, ct2 = P.ct_ptr ? __ldg(P.ct_ptr + 1) : P.ct2;
  double acc1[RPT], zA[RPT], acc2[RPT], zL[RPT], sxy[RPT];
  double zA1[RPT], zL1[RPT];            // second register set of the z pipeline (see `iter`)
  bool in1[RPT], st1[RPT], st2[RPT], own[RPT], ec1[RPT], ec2[RPT];
  const double sxv = x_in ? P.sx[gx] : 0.0;
#pragma unroll
  for (int r = 0; r < RPT; ++r) {
    const int gy = y0 + r0 + r;
    const bool y_in = gy >= P.in_lo[1] && gy <= P.in_hi[1];
    const bool y_ok = gy >= T.lo[1] && gy < T.hi[1];
    // FULL: the tile is complete in y and free of skipped rows, so only the thread-constant
    // x conditions remain (a partial last tile in x just switches whole columns off)
    // XFULL additionally: complete in x and free of skipped columns -> no predicate at all
    in1[r] = (XFULL || x_in) && (FULL || y_in);                             // level 1 is meaningful here
    own[r] = (XFULL || x_ok) && (FULL || y_ok);                             // an owned node this launch updates
    // level 1 is stored on owned nodes and, with a wide halo, on first-ghost-layer nodes that fall
    // inside a partial tile (ghost along exactly one of x / y; edges are never needed)
    const bool xg = gx < 1 || gx > g.n[0], yg = gy < 1 || gy > g.n[1];
    st1[r] = own[r] || (!(FULL && XFULL) && P.wide && in1[r] && ((xg && y_ok) || (yg && x_ok)));
    st2[r] = own[r] && (XFULL || (!(P.skip_lo[0] && gx == 1) && !(P.skip_hi[0] && gx == g.n[0]))) &&
             (FULL || (!(P.skip_lo[1] && gy == 1) && !(P.skip_hi[1] && gy == g.n[1])));
    // rows below yown belong to the previous tile row as well (a last tile row that was moved back so that
    // it is complete): both tiles compute and store them, identically; only one may count their error
    ec1[r] = own[r] && gy >= yown;
    ec2[r] = st2[r] && gy >= yown;
    sxy[r] = in1[r] ? sxv * P.sy[min(max(gy, 0), g.n[1] + 1)] : 0.0;
    acc1[r] = zA[r] = acc2[r] = zL[r] = zA1[r] = zL1[r] = 0.0;
  }
  // ring node of this thread (level 1 only), in level-1-plane coordinates
  const bool has_ring = tid < Cfg::kRing;
  int rcx = 0, rry = 0;
  if (has_ring) {
    if (tid < W1) { rcx = tid; rry = 0; }
    else if (tid < 2 * W1) { rcx = tid - W1; rry = Cfg::kRows1 - 1; }
    else { const int q = tid - 2 * W1; rcx = (q & 1) ? W1 - 1 : 0; rry = 1 + (q >> 1); }
  }
  // cluster variant: rank 0 owns the upper tile (smaller y) of the pair, rank 1 the lower one; the ring row
  // facing the partner is not computed (its values are the partner's own tile row, read through DSMEM)
  const uint32_t crank = Cfg::kClusterY == 2 ? cluster_ctarank() : 0u;
  const bool ring_partner_row = Cfg::kClusterY == 2 && has_ring && tid < 2 * W1 &&
                                (crank == 0 ? rry == Cfg::kRows1 - 1 : rry == 0);
  const int rgx = x0 - 1 + rcx, rgy = y0 - 1 + rry;
  const bool ring_in = has_ring && !ring_partner_row && rgx >= P.in_lo[0] && rgx <= P.in_hi[0] && rgy >= P.in_lo[1] && rgy <= P.in_hi[1];
  // wide halo: a ring node on the first ghost layer of x (or y) whose other coordinate is an
  // owned row (column) of this tile keeps its level-1 value in out1's ghost layer
  const bool ring_xg = rgx < 1 || rgx > g.n[0], ring_yg = rgy < 1 || rgy > g.n[1];
  const bool ring_st = P.wide && ring_in && (ring_xg != ring_yg) &&
                       (ring_xg ? (rgy >= max(y0, T.lo[1]) && rgy < min(y0 + Cfg::kTY, T.hi[1]))
                                : (rgx >= max(x0, T.lo[0]) && rgx < min(x0 + TX, T.hi[0])));
  double* oring = P.out1 + (long long)rgy * g.pitch + g.xoff + rgx;
  double racc = 0.0, rzA = 0.0;
  const int pitch = (int)g.pitch;
  // z: level 1 exists on planes in_lo[2] .. in_hi[2]; the first ghost plane is stored by the chunk next to it
  const int zl1_lo = P.in_lo[2], zl1_hi = P.in_hi[2];
  const int zg_lo = (P.wide && zl1_lo == 0 && z0 == 1) ? 0 : -1000000;
  const int zg_hi = (P.wide && zl1_hi == g.n[2] + 1 && z1 == g.n[2] + 1) ? g.n[2] + 1 : -1000000;
  double* o1 = P.out1 + (long long)(y0 + r0) * g.pitch + g.xoff + gx;
  double* o2 = P.out2 + (long long)(y0 + r0) * g.pitch + g.xoff + gx;
  // thread-constant shared-memory offsets (elements)
  const int offA = (r0 + 1) * WA + cxi + 2;      // A box: tile node (cxi, r0-1)
  const int offB = (r0 + 1) * WB + cxi + 2;      // B box: tile node (cxi, r0)
  const int off1 = (r0 + 1) * W1 + cxi + 1;      // level-1 plane: tile node (cxi, r0)
  const int offRA = (rry + 1) * WA + rcx + 1, offRB = rry * WB + rcx + 1, offR1 = rry * W1 + rcx;

  // wide halo, z faces pushed from the loop: per-thread base pointers into the neighbours' arrays
  // (ghost plane index is added per plane), null when that face is not pushed
  const PeerDev& pzl = P.peer[4];
  const PeerDev& pzh = P.peer[5];
  double* const zpl = (P.zpush & 1) ? pzl.field + (long long)(y0 + r0) * pzl.pitch + g.xoff + gx : nullptr;
  double* const zph = (P.zpush & 2) ? pzh.field + (long long)(y0 + r0) * pzh.pitch + g.xoff + gx : nullptr;
  const int nzl = g.n[2];
  // z table entries for the error epilogues, loaded one plane ahead: szc = s_sz[t-1] (level-1 plane
  // finished in l1_phase(t)), szo = s_sz[t-2] (level-2 plane finished in l2_phase(t)).  A load at the
  // point of use exposed the shared-memory latency to every warp once per plane and level.
  double szc = 0.0, szo = 0.0, szn = 0.0;
  // cluster variant: DSMEM addresses (buffer 0; buffer 1 is kL1Bytes further) of the level-1 value that
  // sits above this thread's first row / below its last row when that row belongs to the partner
  uint32_t ds_up = 0u, ds_dn = 0u;
  if (Cfg::kClusterY == 2) {
    // (cl_ov: rows by which the lower tile was moved back onto the upper one, see y0 in the kernel)
    if (crank == 1 && r0 == 0) ds_up = dsmem_addr(l1buf + (Cfg::kTY - cl_ov) * W1 + cxi + 1, 0u);      // row y0-1 in the partner's plane
    if (crank == 0 && r0 + RPT == Cfg::kTY) ds_dn = dsmem_addr(l1buf + (1 + cl_ov) * W1 + cxi + 1, 1u);  // row y0+TY in the partner's plane
  }

  // The z pipeline of one plane is split in two phases.  zAp/zLp hold u^n / level 1 of the previous
  // plane (own column), zAc/zLc receive those of the current plane; the caller alternates the two
  // register sets, so the pipeline advances without any register-to-register copies.
  //
  // l1_phase(t): plane p = z0-2+t of u^n has arrived -> level 1 of plane p-1 is finished (shared
  // memory, global, error) and the partial sums of plane p are formed.
  //
  // S ("steady", always a literal at the call sites, so it folds after inlining) marks the planes in the middle of the chunk, t in [6, nzc+2):
  // every CTA-uniform condition of the general code is known there (the level-1 plane is owned and
  // stored, level 2 is owned, no ghost plane, no shipped plane, no skipped plane), so the body has
  // no uniform compare-and-branch chains (ncu: 35 of 332 instructions per warp and plane) and the
  // scheduler sees one large basic block per phase.
  auto l1_phase = [&](const bool S, const int t, double (&zAp)[RPT], double (&zAc)[RPT], double (&zLc)[RPT]) {
    const int s = t % STAGES;
    const int p = z0 - 2 + t;                 // plane of u^n that has just arrived
    const double* As = reinterpret_cast<const double*>(stage0 + (size_t)s * Cfg::kStageBytes);
    const double* Bs = reinterpret_cast<const double*>(stage0 + (size_t)s * Cfg::kStageBytes + Cfg::kABytes);
    double* L1w = l1buf + (size_t)((t + 1) & 1) * (Cfg::kL1Bytes / 8);     // plane p-1 of level 1
    szn = s_sz[t];                                        // used one plane later
    const double aTop = As[offA];                         // row above this thread's rows
    const double aBot = As[offA + (RPT + 1) * WA];        // row below
#pragma unroll
    for (int r = 0; r < RPT; ++r) zAc[r] = As[offA + (r + 1) * WA];
    const int p1 = p - 1;                                        // level-1 plane completed now
    const bool p1_live = S || (t >= 2 && p1 >= zl1_lo && p1 <= zl1_hi);  // otherwise level 1 is 0 there
    const bool p1_store = S || (p1 >= z0 && p1 < z1);
    const bool p_l1 = S || (p >= z0 - 1 && p <= z1);                   // plane p needs a level-1 partial
#pragma unroll
    for (int r = 0; r < RPT; ++r) zLc[r] = w3d_finish(acc1[r], zAc[r], P.c1z);
    if (!p1_live) {
#pragma unroll
      for (int r = 0; r < RPT; ++r) zLc[r] = 0.0;
    } else if (!(FULL && XFULL)) {
#pragma unroll
      for (int r = 0; r < RPT; ++r) zLc[r] = in1[r] ? zLc[r] : 0.0;
    }
#pragma unroll
    for (int r = 0; r < RPT; ++r) L1w[off1 + r * W1] = zLc[r];
    if (!S && (p1 == zg_lo || p1 == zg_hi)) {
      // wide halo: level 1 on the first ghost plane, kept as the next pass's u^{n-1} halo
      double* o = o1 + (long long)p1 * g.plane;
#pragma unroll
      for (int r = 0; r < RPT; ++r)
        if (own[r]) o[r * pitch] = zLc[r];
    }
    if (p1_store) {
      double* o = o1 + (long long)p1 * g.plane;
#pragma unroll
      for (int r = 0; r < RPT; ++r)
        if (st1[r]) o[r * pitch] = zLc[r];
      if (want1) {
        const double szt = szc * ct1;              // szc = s_sz[t - 1]; s_sz[i] = sz[z0 - 2 + i]
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
          if (ec1[r]) {
            const double e = w3d_abs_err(zLc[r], sxy[r], szt);
            mx1 = fmax(mx1, e);
            s21 = fma(e, e, s21);
          }
        }
      }
    }
    double ring_c = 0.0;
    if (has_ring) {
      ring_c = As[offRA];
      double v = w3d_finish(racc, ring_c, P.c1z);
      if (!(ring_in && p1_live)) v = 0.0;
      L1w[offR1] = v;
      if (ring_st && p1_store) oring[(long long)p1 * g.plane] = v;
    }
    if (p_l1) {
#pragma unroll
      for (int r = 0; r < RPT; ++r) {
        const double xm = As[offA + (r + 1) * WA - 1];
        const double xp = As[offA + (r + 1) * WA + 1];
        const double pv = Bs[offB + r * WB];
        const double ym = r == 0 ? aTop : zAc[r - 1];
        const double yp = r == RPT - 1 ? aBot : zAc[r + 1];
        acc1[r] = w3d_partial(zAc[r], pv, xm, xp, ym, yp, zAp[r], P.a1, P.b1, P.c1x, P.c1y, P.c1z);
      }
      if (has_ring) {
        const double* a = As + offRA;
        racc = w3d_partial(ring_c, Bs[offRB], a[-1], a[1], a[-WA], a[WA], rzA, P.a1, P.b1, P.c1x, P.c1y, P.c1z);
      }
    }
    rzA = ring_c;
  };
  // l2_phase(t): the level-1 plane p-1 is complete in shared memory (CTA barrier) -> level 2 of
  // plane p-2 is finished (global, pushed to a z neighbour when it is one of the two outermost
  // planes, error) and the partial sums of plane p-1 are formed.
  auto l2_phase = [&](const bool S, const int t, double (&zAp)[RPT], double (&zLp)[RPT], double (&zLc)[RPT]) {
    const int p = z0 - 2 + t;
    const double* L1w = l1buf + (size_t)((t + 1) & 1) * (Cfg::kL1Bytes / 8);
    const int p1 = p - 1, p2 = p - 2;                              // level-2 plane completed now: p2
    if (S || (p2 >= z0 && p2 < z1)) {
      const bool skipz = !S && ((P.skip_lo[2] && p2 == 1) || (P.skip_hi[2] && p2 == g.n[2]));
      double v2[RPT];
#pragma unroll
      for (int r = 0; r < RPT; ++r) v2[r] = w3d_finish(acc2[r], zLc[r], P.c2z);
      if (!skipz) {
        double* o = o2 + (long long)p2 * g.plane;
#pragma unroll
        for (int r = 0; r < RPT; ++r)
          if (st2[r]) o[r * pitch] = v2[r];
        if (!S && zpl != nullptr && p2 <= 2) {            // my planes 1, 2 -> the low neighbour's ghost planes n'+1, n'+2
          double* q = zpl + (long long)(pzl.ghost_index + (p2 - 1)) * pzl.plane;
          const int pp = (int)pzl.pitch;
#pragma unroll
          for (int r = 0; r < RPT; ++r)
            if (st2[r]) q[r * pp] = v2[r];
        }
        if (!S && zph != nullptr && p2 >= nzl - 1) {      // my planes n-1, n -> the high neighbour's ghost planes -1, 0
          double* q = zph + (long long)(pzh.ghost_index - (nzl - p2)) * pzh.plane;
          const int pp = (int)pzh.pitch;
#pragma unroll
          for (int r = 0; r < RPT; ++r)
            if (st2[r]) q[r * pp] = v2[r];
        }
        if (want2) {
          const double szt = szo * ct2;              // szo = s_sz[t - 2]
#pragma unroll
          for (int r = 0; r < RPT; ++r) {
            if (ec2[r]) {
              const double e = w3d_abs_err(v2[r], sxy[r], szt);
              mx2 = fmax(mx2, e);
              s22 = fma(e, e, s22);
            }
          }
        }
      }
    }
    if (S || (p1 >= z0 && p1 < z1)) {
      double ym0, ypN;                                   // rows above / below this thread's rows
      if (Cfg::kClusterY == 2) {
        const uint32_t par = (uint32_t)((t + 1) & 1) * (uint32_t)Cfg::kL1Bytes;
        ym0 = ds_up ? ld_dsmem_f64(ds_up + par) : L1w[off1 - W1];
        ypN = ds_dn ? ld_dsmem_f64(ds_dn + par) : L1w[off1 + RPT * W1];
      } else {
        ym0 = L1w[off1 - W1];
        ypN = L1w[off1 + RPT * W1];
      }
#pragma unroll
      for (int r = 0; r < RPT; ++r) {
        const double xm = L1w[off1 + r * W1 - 1];
        const double xp = L1w[off1 + r * W1 + 1];
        const double ym = r == 0 ? ym0 : zLc[r - 1];
        const double yp = r == RPT - 1 ? ypN : zLc[r + 1];
        acc2[r] = w3d_partial(zLc[r], zAp[r], xm, xp, ym, yp, zLp[r], P.a2, P.b2, P.c2x, P.c2y, P.c2z);
      }
    }
    szo = szc;
    szc = szn;
  };
  auto wait_stage = [&](const int t) { mbar_wait(full + t % STAGES, (uint32_t)(t / STAGES) & 1u); };
  auto sync_and_refill = [&](const int t) {
    // level-1 plane p-1 is complete in smem AND every thread is done with stage t % STAGES
    // (cluster variant: complete in BOTH CTAs of the pair, and the partner is done reading the plane before)
    if (Cfg::kClusterY == 2) cluster_sync_all(); else __syncthreads();
    if (tid == 0 && t + STAGES < iters) pair_issue<Cfg>(P, full, stage0, t + STAGES, x0, y0, z0, nzc);
  };
  // General plane (any special case; a few planes at both ends of the chunk): ONE copy of the code, the
  // z pipeline registers are moved by hand.  Steady planes: two per call with the register sets
  // alternating.  Keeping the general code to a single call site matters: with head / steady / tail
  // each inlined twice the kernel grew from 6.8 k to 14.4 k instructions and ran up to 50 % slower.
  auto one_plane = [&](const int t) {
    wait_stage(t);
    l1_phase(false, t, zA, zA1, zL1);
    sync_and_refill(t);
    l2_phase(false, t, zA, zL, zL1);
#pragma unroll
    for (int r = 0; r < RPT; ++r) { zA[r] = zA1[r]; zL[r] = zL1[r]; }
  };
  auto two_steady_planes = [&](const int t) {
    wait_stage(t);
    l1_phase(true, t, zA, zA1, zL1);
    sync_and_refill(t);
    l2_phase(true, t, zA, zL, zL1);
    wait_stage(t + 1);
    l1_phase(true, t + 1, zA1, zA, zL);
    sync_and_refill(t + 1);
    l2_phase(true, t + 1, zA1, zL1, zL);
  };
  constexpr int kHead = 6;                                          // planes with special cases at the chunk's start
  const int t_steady_end = kHead + (max(nzc + 2 - kHead, 0) & ~1);  // steady planes: [kHead, nzc+2), whole pairs
#pragma unroll 1
  for (int t = 0; t < iters;) {
    if (t >= kHead && t < t_steady_end) { two_steady_planes(t); t += 2; }
    else { one_plane(t); ++t; }
  }
}

template <class Cfg>
__global__ void __launch_bounds__(Cfg::kThreads, Cfg::kMinBlocks) pair_step_kernel(const __grid_constant__ PairParams P) {
  constexpr int TX = Cfg::kTX, STAGES = Cfg::kStages;
  // Indexed directly (no integer round trip) so the compiler keeps the shared address space:
  // LDS/STS with 32-bit addresses instead of generic LD/ST with 64-bit address arithmetic.
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);
  double* red = reinterpret_cast<double*>(smem + 128);
  double* s_sz = reinterpret_cast<double*>(smem + 128 + 1024);
  uint8_t* stage0 = smem + Cfg::kBarBytes;
  double* l1buf = reinterpret_cast<double*>(stage0 + (size_t)STAGES * Cfg::kStageBytes);

  const TileGeom& T = P.tile;
  const Geometry& g = P.g;
  const int work = P.order ? P.order[blockIdx.x] : (int)blockIdx.x;
  const int tx = work % T.ntx;
  const int ty = (work / T.ntx) % T.nty;
  const int ch = work / (T.ntx * T.nty);
  const int x0 = T.x0 + tx * TX;
  // A last tile row that would be partial is moved back so that it ends exactly at the block's last row: it
  // overlaps the previous tile row in rows [y0, yown), which both compute and store identically (inputs and
  // outputs are different arrays), and every tile runs the predicate-free code.  In a one-wave launch the
  // slowest CTA sets the pass time, and the partial tiles were that CTA (mirrored in pair_tile_y0(), wave_api.cu).
  const int yown = T.lo[1] + ty * Cfg::kTY;
  const int y0 = (yown + Cfg::kTY > T.hi[1] && T.hi[1] - Cfg::kTY >= T.lo[1]) ? T.hi[1] - Cfg::kTY : yown;
  int cl_ov = 0;      // cluster variant: overlap rows of the LOWER tile of the pair (known to both CTAs)
  if (Cfg::kClusterY == 2) {
    const int ylow = T.lo[1] + (ty | 1) * Cfg::kTY;     // nominal first row of the pair's lower tile (odd tile row)
    cl_ov = (ylow + Cfg::kTY > T.hi[1] && T.hi[1] - Cfg::kTY >= T.lo[1]) ? ylow - (T.hi[1] - Cfg::kTY) : 0;
  }
  const int z0 = T.lo[2] + ch * T.zc;
  const int z1 = min(z0 + T.zc, T.hi[2]);
  const int nzc = z1 - z0;
  const int iters = nzc + 4;          // planes z0-2 .. z1+1 of u^n

  double mx1 = 0.0, s21 = 0.0, mx2 = 0.0, s22 = 0.0;
  if (P.tl && threadIdx.x == 0) atomicMin(P.tl, globaltimer_ns());
  // copy-out mask: faces of which this tile owns one of the halo_depth outermost layers; wait
  // mask: faces whose ghost layers it reads — the level-1 ring reaches one node further than the
  // tile, so a tile that ends one layer short of a face (hi == n, lo == 2) reads that ghost too
  // and has to wait for it.  With the wide halo (two layers shipped) the two masks coincide.
  const int mask = tile_face_mask(P, x0, y0, z0, z1, TX, Cfg::kTY, P.halo_depth);
  const int wmask = tile_face_mask(P, x0, y0, z0, z1, TX, Cfg::kTY, 2);
  unsigned long long wait_ns = 0ull;
  // wide halo: tell the neighbours that the layers pushed by the previous launch have landed (first thing, it
  // is on their critical path)
  if (P.defer_release && P.use_flags && blockIdx.x == 0 && threadIdx.x == 0) entry_release(P);
  // z slice of the analytic table for the error epilogue (a global load per plane inside the
  // loop exposed its full latency to every warp: ncu, top long-scoreboard stall)
  for (int i = threadIdx.x; i < iters && i < Cfg::kSzEntries; i += Cfg::kThreads) {
    const int k = z0 - 2 + i;
    s_sz[i] = (k >= 0 && k <= g.n[2] + 1) ? P.sz[k] : 0.0;
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) mbar_init(full + s, 1);
    fence_barrier_init();
    tma_prefetch_desc(&P.map_a);
    tma_prefetch_desc(&P.map_b);
  }
  if (threadIdx.x == 0) {
    if (wmask) wait_ns = wait_peer_flags(P, wmask);    // ghosts of u^n must have arrived before they are loaded
    for (int t = 0; t < STAGES && t < iters; ++t) pair_issue<Cfg>(P, full, stage0, t, x0, y0, z0, nzc);
  }
  if (P.skip_b) {
    // the previous level has coefficient 0 in this pass: instead of streaming it (8 B per node),
    // the consumers read zeros — fma(0, 0, acc) is exactly what a zero-initialised array gives
    for (int s = 0; s < STAGES; ++s) {
      double* bs = reinterpret_cast<double*>(stage0 + (size_t)s * Cfg::kStageBytes + Cfg::kABytes);
      for (int i = threadIdx.x; i < Cfg::kBBytes / 8; i += Cfg::kThreads) bs[i] = 0.0;
    }
  }
  __syncthreads();
  {
    // Fast paths, chosen per tile (one code path per CTA).  FULL: every row of the tile is owned, updated and
    // free of skipped rows — true for every tile row now that a last partial row is moved back (y0 above),
    // except in blocks thinner than a tile and next to a skipped face of the fix-up protocol; XFULL: the tile is
    // complete in x and free of skipped columns.  All instantiations have the same barrier structure.
    // (A choice per thread group was measured slower: three code paths live in one CTA, profiles/pair_schedule.md.)
    const int gy0 = y0, gy1 = y0 + Cfg::kTY;       // whole tile: one code path per CTA
    const bool rows_full = gy0 >= T.lo[1] && gy1 <= T.hi[1] && !(P.skip_lo[1] && gy0 <= 1) &&
                           !(P.skip_hi[1] && gy1 > g.n[1]);
    const bool xfull = x0 >= T.lo[0] && x0 >= 1 && x0 + TX <= T.hi[0] && !(P.skip_lo[0] && x0 <= 1) &&
                       !(P.skip_hi[0] && x0 + TX > g.n[0]);
    if (rows_full && xfull)
      pair_consumer<Cfg, true, true>(P, full, stage0, l1buf, s_sz, x0, y0, yown, cl_ov, z0, z1, mx1, s21, mx2, s22);
    else if (rows_full)
      pair_consumer<Cfg, true, false>(P, full, stage0, l1buf, s_sz, x0, y0, yown, cl_ov, z0, z1, mx1, s21, mx2, s22);
    else
      pair_consumer<Cfg, false, false>(P, full, stage0, l1buf, s_sz, x0, y0, yown, cl_ov, z0, z1, mx1, s21, mx2, s22);
    if (Cfg::kClusterY == 2) cluster_sync_all();   // the partner may still be reading this CTA's level-1 planes
    if (mask) {
      __syncthreads();   // the tile's shipped level is written (CTA-visible), z-face pushes are issued
      const unsigned long long t_copy0 = P.xstat ? globaltimer_ns() : 0ull;
      const int cmask = mask & ~((P.zpush & 1 ? 16 : 0) | (P.zpush & 2 ? 32 : 0));   // faces still to copy: x, y
      if (cmask) {
        face_copy_out(P, cmask, x0, y0, z0, z1, TX, Cfg::kTY, threadIdx.x, Cfg::kThreads);
        edge_copy_out(P, x0, y0, z0, z1, TX, Cfg::kTY, threadIdx.x, Cfg::kThreads);   // (every edge has a face in cmask)
        __syncthreads();
      }
      if (threadIdx.x == 0) {
        if (!P.defer_release) face_cta_done(P, wait_ns, t_copy0);
        else if (P.xstat) {             // kernel completion delivers the halos; the next launch publishes them
          atomicMax(&P.xstat->cur_wait[P.seq & 1], wait_ns);
          atomicMax(&P.xstat->cur_copy[P.seq & 1], globaltimer_ns() - t_copy0);
        }
      }
    }
  }
  if (P.err1.out != nullptr || P.err2.out != nullptr) {
    __syncthreads();
    publish_error2(P.err1, P.err2, mx1, s21, mx2, s22, red, reinterpret_cast<int*>(smem + 96));
  }
  if (P.tl && threadIdx.x == 0) atomicMax(P.tl + 1, globaltimer_ns());
}

}  // namespace w3d
