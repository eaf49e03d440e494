/* wave3d.h — C ABI of the sm_100a CUDA library (libwave3d_b200.so).
 *
 * B200-native replacement for the reference's MPI+CUDA program mpigpu-1.cu
 * (named at readme.md:68, described in report.pdf p.14-17 §4; not shipped).
 * Functional kernel inventory of that program -> entry points here
 * (SURVEY.md §2.6):
 *   K1 init layer 0                      -> w3d_plan_init
 *   K2 Taylor step, K3 leapfrog, K4 BCs  -> w3d_step_fused / w3d_step_tma / w3d_step_box
 *   K5 halo pack/unpack + MPI_Sendrecv   -> in-kernel NVLink P2P stores + flags (w3d_step_fused);
 *                                           w3d_pack_face / w3d_unpack_face for the NCCL baseline
 *   K6 error kernel + MPI_Reduce         -> epilogue of every step kernel; w3d_error_norms stand-alone
 *
 * All pointers are device pointers unless stated.  Every function returns 0 on
 * success or a negative W3D_E* code; w3d_last_error() gives the text (except for
 * w3d_error_norms_noalloc, which writes it into its errbuf argument).  Streams are
 * passed as void* (cudaStream_t).  No function synchronises the device unless it says so.
 */
#ifndef WAVE3D_H
#define WAVE3D_H
#include <stddef.h>
#include <stdint.h>
#include "wave3d_cpu.h" /* W3dBlock: the block layout is shared with the CPU library */

#ifdef __cplusplus
extern "C" {
#endif

#define W3D_OK 0
#define W3D_EINVAL (-1)
#define W3D_ECUDA (-2)
#define W3D_ENOMEM (-3)
#define W3D_EDRIVER (-4)
#define W3D_ETOOBIG (-5)     /* w3d_error_norms_noalloc only: block exceeds the built-in scratch */

#define W3D_XOFF 15          /* x ghost sits at element 15, first unknown at 16 (128 B aligned) */
#define W3D_PITCH_ALIGN 16   /* row pitch is a multiple of 16 doubles (128 B) */
#define W3D_ERR_SLOTS 8      /* independent error-reduction slots per step */
#define W3D_MAX_PARTIALS 16384

/* Recommended layout for n unknowns along x. */
int64_t w3d_pitch_for(int32_t nx);

/* One face of the block whose edge layer must be delivered to a neighbour's ghost
 * layer (the neighbour may be this rank itself for a periodic axis on one rank). */
typedef struct W3dPeer {
  int32_t active;        /* 0: Dirichlet boundary, nothing to send */
  int32_t ghost_index;   /* index along the face axis of the peer's ghost layer that my edge fills */
  double* field[2];      /* peer's two field arrays, mapped into this process */
  int64_t pitch;         /* peer's row pitch (elements) */
  int64_t plane;         /* peer's plane stride (elements) */
  unsigned long long* flag;   /* peer's incoming flag word for this face (I write it) */
  unsigned long long* err_slot; /* unused, reserved */
} W3dPeer;

typedef struct W3dPlanDesc {
  W3dBlock blk;
  double* field[2];             /* my two field arrays (time levels alternate) */
  const double* tab[3];         /* analytic tables sx, sy, sz with n+2 entries (ghosts incl.) */
  double coef[3];               /* tau^2 / h_d^2 */
  W3dPeer peer[6];              /* index axis*2 + side */
  unsigned long long* flags_in; /* 6 words written by my neighbours: level of the ghost data present */
  uint32_t* status;             /* device word, set non-zero by the spin watchdog */
  void* workspace;              /* scratch for the error reductions and shell counters */
  size_t workspace_bytes;       /* >= w3d_workspace_bytes() */
  int32_t variant;              /* TMA tile variant, -1 = auto */
  int32_t zchunk;               /* planes per CTA in the z-marching kernel, 0 = auto */
  int32_t num_sms;              /* 0 = query the current device */
  int32_t spin_timeout_ms;      /* watchdog for flag waits, 0 = 10000 */
  /* Cross-GPU error reduction (optional).  Every rank owns a table of identical layout;
   * err_table is mine, err_table_peer[q] are the other ranks' tables mapped into this
   * process.  When w3d_step_fused is given err_out inside err_table, the kernel stores
   * the pair at the same offset of every peer table as well (MPI_Reduce replacement:
   * report.pdf p.15-16 shows rank 0 printing global errors each reporting step). */
  double* err_table;
  double* err_table_peer[7];
  int32_t n_err_peers;
  /* Ghost depth of the field arrays (0 or 1: one ghost layer, the classic layout; 2: two).
   * With 2, field[] (and the aux arrays of w3d_plan_enable_pairs, and the peers' arrays) point at
   * local index (0,0,0) of an allocation that also holds rows / planes -1 and n+2 (and columns
   * xoff-1, xoff+n+2 when an x face has a neighbour), plane_stride >= (ny+4)*pitch, and the
   * tables tab[d] are readable at indices -1 and n+2 (0 beyond a Dirichlet boundary).  This is
   * what lets two time steps share ONE launch and ONE halo handshake (see w3d_step_pair). */
  int32_t ghost_depth;
  int64_t plane_stride;         /* elements between planes; 0 = (ny+2)*pitch */
} W3dPlanDesc;

typedef struct W3dPlan W3dPlan;

size_t w3d_workspace_bytes(void);
const char* w3d_last_error(void);
const char* w3d_version(void);

int w3d_plan_create(const W3dPlanDesc* desc, W3dPlan** out);
void w3d_plan_destroy(W3dPlan* plan);
/* Human-readable description of the launch geometry chosen (static storage). */
const char* w3d_plan_describe(W3dPlan* plan);
/* Number of kernel launches issued through this plan so far. */
long long w3d_plan_launches(W3dPlan* plan);

/* Optional: read the time factor cos(a_t t_{s+1}) of step s from dev_table[s+1] (device
 * memory that the host refreshes every step) instead of the by-value `ct` argument.
 * NULL switches back.  Used by the streaming / end-to-end mode. */
int w3d_plan_set_ct_table(W3dPlan* plan, const double* dev_table);

/* field[0] <- analytic u0 on every node incl. ghosts, field[1] <- 0; clears flags,
 * status and workspace counters (K1; report.pdf p.5 §2.2 item 1). */
int w3d_plan_init(W3dPlan* plan, void* stream);
/* zero_second = 0 leaves field[1] untouched: valid when the first thing that happens to it is
 * being overwritten — a pair-mode run with an even number of steps (its first pass does not read
 * the previous level at all, the second pass writes field[0..1]).  Saves 8 B per node of init. */
int w3d_plan_init_ex(W3dPlan* plan, void* stream, int zero_second);

/* One time step producing level s+1 from levels s (cur = field[s&1]) and s-1
 * (stored in field[(s+1)&1], overwritten in place):
 *    next = a*cur + b*prev + cscale * sum_d coef[d] * (cur_{+1} - 2 cur + cur_{-1})
 * Taylor first step: a=1,b=0,cscale=0.5; leapfrog: a=2,b=-1,cscale=1.
 * If err_out != NULL the kernel also reduces max|next - exact| and sum (..)^2 over
 * the nodes it updates into err_out[0..1] (exact = sx*sy*sz*ct), using reduction
 * slot `slot` of the workspace.
 *
 * w3d_step_fused: ONE launch over the whole block.  CTAs whose tile touches a face
 *   with a neighbour are numbered (hence dispatched) first: they acquire the
 *   neighbours' flags for level s before loading, stream their tile with TMA like
 *   every other CTA, then copy the freshly written face layers straight into the
 *   neighbours' ghost layers (NVLink P2P stores); the last of them releases the
 *   flags for level s+1.  The remaining CTAs stream the interior meanwhile.  The
 *   error pair is also broadcast into the peers' tables (see W3dPlanDesc).
 * w3d_step_tma: the TMA z-marching kernel on the box [lo,hi) only, no communication.
 * w3d_step_box: the plain one-thread-per-node kernel on the box [lo,hi), no
 *   communication (used by the NCCL baseline for the shell and as a numerics check). */
int w3d_step_fused(W3dPlan* plan, int s, double a, double b, double cscale, double ct,
                   double* err_out, void* stream);
int w3d_step_tma(W3dPlan* plan, int s, double a, double b, double cscale, double ct,
                 const int32_t lo[3], const int32_t hi[3], double* err_out, int slot, void* stream);
int w3d_step_box(W3dPlan* plan, int s, double a, double b, double cscale, double ct,
                 const int32_t lo[3], const int32_t hi[3], double* err_out, int slot, void* stream);

/* Temporal blocking: two time steps in one pass over memory (16 B per node and step instead
 * of 24; no counterpart in the reference, which does one pass per step).  Needs two more
 * arrays of the same layout; the four arrays are addressed by index: 0,1 = desc.field[0..1],
 * 2,3 = aux0, aux1.  w3d_step_pair reads level n from `cur` and level n-1 from `prev` and
 * writes level n+1 to `out1` and level n+2 to `out2` (all four distinct).  (a,b,cscale,ct)
 * are given per produced level.  peer_aux: for every face f with a neighbour, peer_aux[2f]
 * and peer_aux[2f+1] are that neighbour's aux0/aux1 mapped here.
 *
 * Faces with a neighbour — w3d_plan_pair_mode() tells which protocol the plan uses:
 *   W3D_PAIR_LOCAL  no face has a neighbour: nothing to exchange.
 *   W3D_PAIR_WIDE   ghost_depth 2 and neighbours along one axis, or along several with the edge peers
 *       of w3d_plan_set_edges: ONE launch and ONE handshake per two steps.  `cur` must hold two valid ghost layers and `prev` one; the
 *       pass also forms level n+1 on the first ghost layer (kept in out1's ghosts for the next
 *       pass), completes level n+2 on every owned node, ships its two outermost layers per face
 *       into the neighbours' ghost layers of their `out2` and releases level s+2.
 *   W3D_PAIR_FIXUP  otherwise (one ghost layer, or neighbours along two or three axes without the
 *       edge peers): nodes next to a
 *       face with a neighbour are left out of `out2`; the pass ships level n+1 (one layer,
 *       releases s+1) and the caller finishes the layer with w3d_step_fixup. */
/* Wide halo on 3-D blocks (neighbours along two or three axes): the two-step stencil also reaches the
 * DIAGONAL neighbours across the block's edges, so besides the six faces up to twelve edge lines are
 * exchanged.  Edge e = 4 * pair + 2 * sa + sb with pair 0 = (x,y), 1 = (x,z), 2 = (y,z) and sa, sb = my
 * sides of the two axes (0 low, 1 high).  edges[e] describes the rank diagonally across that edge: its four
 * arrays (local index origin, index order as in w3d_plan_enable_pairs), its strides, the indices
 * (ghost_a, ghost_b) of ITS ghost line that my edge line fills, and its incoming flag word for this edge —
 * word 6 + e' of its flags_in, e' being the same edge seen from there (sides flipped).  flags_in of this
 * plan must then hold W3D_FLAG_WORDS words.  Call before w3d_plan_enable_pairs; an edge is needed wherever
 * two faces on different axes both have a neighbour, and if one is missing the plan falls back to
 * W3D_PAIR_FIXUP.  Pass NULL to clear. */
#define W3D_EDGES 12
#define W3D_FLAG_WORDS 18
typedef struct W3dEdgePeer {
  int32_t active;
  int32_t ghost_a, ghost_b;
  double* arr[4];
  int64_t pitch, plane;
  unsigned long long* flag;
} W3dEdgePeer;
int w3d_plan_set_edges(W3dPlan* plan, const W3dEdgePeer* edges /* [W3D_EDGES] or NULL */);

#define W3D_PAIR_LOCAL 0
#define W3D_PAIR_WIDE 1
#define W3D_PAIR_FIXUP 2
int w3d_plan_enable_pairs(W3dPlan* plan, double* aux0, double* aux1, double* const* peer_aux /* [12] or NULL */,
                          int pair_variant, int zchunk);
int w3d_plan_pair_mode(W3dPlan* plan);
/* s = level held by `cur`.  Face CTAs acquire the neighbours' flags for level s before their
 * first load; see the protocols above for what they ship afterwards. */
int w3d_step_pair(W3dPlan* plan, int s, int cur, int prev, int out1, int out2, double a1, double b1,
                  double cscale1, double ct1, double a2, double b2, double cscale2, double ct2, double* err1,
                  double* err2, void* stream);
/* Fix-up pass: out = a*cur + b*prev + cscale*Laplace(cur) on the layer next to every face with a
 * neighbour (the nodes w3d_step_pair left out), after acquiring the flags for `level` (= the
 * level held by `cur`, whose ghosts the neighbours deliver during their pair pass); stores go to
 * the local array and into the neighbours' ghost layers; releases level + 1. */
int w3d_step_fixup(W3dPlan* plan, int level, int cur, int prev, int out, double a, double b, double cscale,
                   double ct, double* err_out, void* stream);

/* Single step on any two of the four arrays (after w3d_plan_enable_pairs): same kernel and
 * protocol as w3d_step_fused, with level s in array `cur` and level s-1 in array `prev`, which
 * receives level s+1.  Used for the last step of an odd-length run in pair mode. */
int w3d_step_fused_on(W3dPlan* plan, int s, int cur, int prev, double a, double b, double cscale, double ct,
                      double* err_out, void* stream);

/* Exchange statistics measured on the device (%globaltimer, nanoseconds): how long face CTAs
 * spun on their neighbours' flags and how long their halo copy-out (P2P stores + system fence)
 * took.  Per launch the maximum over the face CTAs is taken; out[] receives
 *   [0] sum over launches of the flag-wait maxima   [1] sum of the copy-out maxima
 *   [2] launches counted   [3] largest single flag wait   [4] largest single copy-out.
 * Counterpart of the reference's "MPI" and "H2D/D2H" time columns (report.pdf p.16 §4.4).
 * Synchronises the stream it is given.  w3d_plan_init clears the counters. */
#define W3D_XSTAT_WORDS 5
int w3d_plan_exchange_stats(W3dPlan* plan, unsigned long long out[W3D_XSTAT_WORDS], void* stream);

/* Device timeline (tracing): dev_buf holds `capacity` pairs of 64-bit %globaltimer stamps.  A step
 * or pair launch whose input level is s < capacity records into pair s: [2s] = earliest start of
 * any of its CTAs (atomicMin), [2s+1] = latest end (atomicMax).  The caller fills the buffer with
 * (UINT64_MAX, 0) pairs beforehand; NULL switches recording off.  The difference between one
 * launch's end and the next one's start is the launch gap that no event timing can see. */
int w3d_plan_set_timeline(W3dPlan* plan, unsigned long long* dev_buf, int capacity);

/* ---- time stepper ------------------------------------------------------------------------
 * Which arrays hold the newest level and the one before, advanced by the library: the loop body of
 * the reference's main() (report.pdf p.15 §4.2.4) without its copies.  ct is a HOST table of
 * cos(a_t t_n), n = 0..K (ignored by the kernels while a device table is attached with
 * w3d_plan_set_ct_table).  err pointers are device addresses of (max, sum) pairs or NULL.
 * w3d_advance_pair: levels +1 and +2 in one pass (plus the fix-up launch when the plan needs it);
 * w3d_advance_single: one level with the one-pass kernel (any plan; the last step of an odd-length
 * run in pair plans). */
typedef struct W3dStepper { int32_t level, cur, prev, reserved; } W3dStepper;
void w3d_stepper_reset(W3dStepper* st);                       /* level 0 in array 0, nothing before it */
int w3d_plan_pairs_enabled(W3dPlan* plan);                    /* 1 after w3d_plan_enable_pairs */
int w3d_advance_pair(W3dPlan* plan, W3dStepper* st, const double* ct, double* err1, double* err2, double* err_fix,
                     void* stream);
int w3d_advance_single(W3dPlan* plan, W3dStepper* st, const double* ct, double* err, void* stream);

/* ---- streaming session ------------------------------------------------------------------
 * Step-by-step use with per-step host traffic, natively: w3d_session_step_async(n) copies step n's
 * time factor from pinned host memory to the device (8 B), launches what can be launched (in pair
 * plans the call for the second step of a pair launches the pass for both; the call for the first
 * one only uploads), copies the error pairs of the finished levels back to pinned host memory and
 * records an event; w3d_session_result(n) waits for that event only and folds the slots of step n
 * into out[0] = max, out[1] = sum of squares (this rank).  Feed n = 1, 2, ... K in order; any number
 * of steps may be in flight.  All buffers belong to the caller: ct_host (K+1 doubles) and err_host
 * (K+1 rows of `slots` pairs) pinned, ct_dev (>= K+1 doubles) and err_dev (same shape as err_host:
 * row n slot 0 = step kernels, slot 1 = fix-up pass) on the device.  The session attaches ct_dev to
 * the plan (w3d_plan_set_ct_table) until it is destroyed.  Counterpart of the reference's per-step
 * H2D/D2H traffic (report.pdf p.16 §4.4), reduced to 8 bytes in and 16*slots bytes out per step. */
typedef struct W3dSession W3dSession;
typedef struct W3dSessionDesc {
  W3dPlan* plan;
  int32_t K;
  int32_t slots;
  const double* ct_host;
  double* ct_dev;
  double* err_dev;
  double* err_host;
} W3dSessionDesc;
int w3d_session_create(const W3dSessionDesc* desc, W3dSession** out);
void w3d_session_destroy(W3dSession* s);
int w3d_session_begin(W3dSession* s, const W3dStepper* st /* NULL = level 0 */);   /* after (re-)initialising the fields */
int w3d_session_step_async(W3dSession* s, int n, void* stream);
int w3d_session_result(W3dSession* s, int n, double out[2]);
int w3d_session_stepper(W3dSession* s, W3dStepper* out);

/* NCCL-baseline helpers (K5): contiguous face buffers of the unknowns of the two
 * other axes, fastest index first.  ghost=0 packs/unpacks the edge layer of
 * unknowns, ghost=1 the ghost layer. */
int w3d_pack_face(W3dPlan* plan, int parity, int axis, int side, int ghost, double* buf, void* stream);
int w3d_unpack_face(W3dPlan* plan, int parity, int axis, int side, int ghost, const double* buf, void* stream);
/* periodic axis owned by one rank: copy edges into the opposite ghosts */
int w3d_wrap_axis(W3dPlan* plan, int parity, int axis, void* stream);

/* Stand-alone error norms of one field (K6): out[0] = max |u - sx*sy*sz*ct| over the
 * unknowns, out[1] = sum of squares.  `workspace` is scratch of at least
 * w3d_error_workspace_bytes() bytes; its first 4 bytes must be zero on first use
 * (the kernel leaves them zero again). */
size_t w3d_error_workspace_bytes(void);
int w3d_error_norms(const double* u, const W3dBlock* blk, const double* sx, const double* sy,
                    const double* sz, double ct, double* out, void* workspace,
                    size_t workspace_bytes, void* stream);

/* Same norms for callers that cannot allocate device memory (no counterpart in the
 * reference: it is an addition of this library).  Meant for code that lives inside a
 * host application with its own fixed memory pool, or inside someone else's time loop
 * where a cudaMalloc would stall the whole device: such callers can supply inputs and
 * the 2-double output, but no workspace.  The scratch is a static array inside the
 * library instead — W3D_NOALLOC_SCRATCH_BYTES per GPU, held until the process exits —
 * which limits the block to ny*nz <= W3D_NOALLOC_MAX_ROWS (nx is not limited); a larger
 * block returns W3D_ETOOBIG and leaves `out` untouched.  No cudaMalloc/cudaFree or other
 * allocating call is made.  The first call on a device resolves the array's address and
 * should happen outside stream capture.  All calls on a device share the array: calls
 * ordered on one stream are fine, calls that may overlap on the GPU race.  On failure
 * the reason is written to errbuf (may be NULL).  out[0] = max error, out[1] = sum of
 * squares, as for w3d_error_norms (the sum may differ in the last bits: other order). */
#define W3D_NOALLOC_MAX_ROWS 262144
#define W3D_NOALLOC_SCRATCH_BYTES 33024
int w3d_error_norms_noalloc(const double* u, const W3dBlock* blk, const double* sx, const double* sy,
                            const double* sz, double ct, double* out, void* stream, char* errbuf,
                            size_t errbuf_len);

#ifdef __cplusplus
}
#endif
#endif
