// wave_api.cu — host side of libwave3d_b200.so: plans, tensor maps, launch
// geometry and the extern "C" entry points declared in include/wave3d.h.
//
// Replaces the host half of the reference's mpigpu-1.cu (report.pdf p.14-17 §4):
// there, every step staged halos D2H -> MPI_Sendrecv -> H2D; here
// a step is one kernel launch and nothing crosses the host.
#include "kernels.cuh"
#include "pair_kernel.cuh"

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

using namespace w3d;

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define CUDA_TRY(x)                                                                      \
  do {                                                                                   \
    cudaError_t e_ = (x);                                                                \
    if (e_ != cudaSuccess) return fail(W3D_ECUDA, "%s: %s", #x, cudaGetErrorString(e_)); \
  } while (0)

// ---- TMA tile variants ------------------------------------------------------------
//            TX  RPT RG STAGES   (+32 threads: the producer warp)
using Cfg0 = TileCfg<128, 8, 1, 4>;   // 128 x 8 tile,  160 threads, ~76 KB  -> 2 CTAs / SM
using Cfg1 = TileCfg<128, 8, 2, 4>;   // 128 x 16 tile, 288 threads, ~143 KB -> 1 CTA / SM
using Cfg2 = TileCfg<128, 8, 4, 3>;   // 128 x 32 tile, 544 threads, ~207 KB -> 1 CTA / SM
using Cfg3 = TileCfg<64, 8, 4, 4>;    // 64 x 32 tile,  288 threads, ~141 KB -> 1 CTA / SM  (default)
using Cfg4 = TileCfg<64, 8, 4, 5>;    // 64 x 32 tile,  5-stage ring, ~176 KB -> 1 CTA / SM
using Cfg5 = TileCfg<64, 8, 2, 6>;    // 64 x 16 tile,  160 threads, ~109 KB -> 2 CTAs / SM

struct Variant {
  int tx, ty, threads, smem, ctas_per_sm;
  void (*kernel)(const StepParams);
  const char* name;
};

const Variant kVariants[] = {
    {Cfg0::kTX, Cfg0::kTY, Cfg0::kThreads, Cfg0::kSmemBytes, 2, fused_step_kernel<Cfg0, 2>, "128x8x4st"},
    {Cfg1::kTX, Cfg1::kTY, Cfg1::kThreads, Cfg1::kSmemBytes, 1, fused_step_kernel<Cfg1, 1>, "128x16x4st"},
    {Cfg2::kTX, Cfg2::kTY, Cfg2::kThreads, Cfg2::kSmemBytes, 1, fused_step_kernel<Cfg2, 1>, "128x32x3st"},
    {Cfg3::kTX, Cfg3::kTY, Cfg3::kThreads, Cfg3::kSmemBytes, 1, fused_step_kernel<Cfg3, 1>, "64x32x4st"},
    {Cfg4::kTX, Cfg4::kTY, Cfg4::kThreads, Cfg4::kSmemBytes, 1, fused_step_kernel<Cfg4, 1>, "64x32x5st"},
    {Cfg5::kTX, Cfg5::kTY, Cfg5::kThreads, Cfg5::kSmemBytes, 2, fused_step_kernel<Cfg5, 2>, "64x16x6st"},
};
constexpr int kNumVariants = sizeof(kVariants) / sizeof(kVariants[0]);

// ---- temporal-blocking (two steps per pass) variants ----
using PCfg0 = PairCfg<64, 8, 4, 4>;    // 64 x 32 tile, 256 threads, 4-stage ring (~194 KB)
using PCfg1 = PairCfg<64, 8, 4, 3>;    // 3-stage ring (~150 KB)
using PCfg2 = PairCfg<128, 8, 4, 2>;   // 128 x 32 tile, 544 threads, 2-stage ring (~215 KB)
using PCfg3 = PairCfg<64, 4, 8, 4>;    // 64 x 32 tile, 4 rows per thread: 544 threads (16 consumer warps)
using PCfg4 = PairCfg<64, 4, 8, 3>;
using PCfg5 = PairCfg<64, 2, 15, 4>;   // 64 x 30 tile, 2 rows per thread: 960 threads (30 warps)
using PCfg6 = PairCfg<64, 4, 4, 3, 2>; // 64 x 16 tile, 256 threads, ~88 KB: TWO CTAs per SM (small blocks)
using PCfg7 = PairCfg<64, 4, 4, 4, 2>; // same with a 4-stage ring (~108 KB)
// 64 x 30 tiles: 511 rows make 18 tiles instead of 16 -> 8 x 18 = 144 CTAs on 148 SMs where
// 64 x 32 leaves 20 SMs idle (one-wave blocks: 512^3 on 4 and 8 GPUs)
using PCfg8 = PairCfg<64, 5, 6, 3>;    // 384 threads, 5 rows per thread
using PCfg9 = PairCfg<64, 3, 10, 3>;   // 640 threads, 3 rows per thread
using PCfg10 = PairCfg<64, 6, 5, 4>;   // 320 threads, 6 rows per thread, 4-stage ring
// (tried and dropped: 128 x 14 / 128 x 15 tiles — 4 x 37 = 148 / 4 x 35 = 140 CTAs at 511 x 511 — with 2 / 3
// rows per thread: 0.085 / 0.072 ms per step on the 511 x 511 x 64 block against 0.057 for 64 x 32)
using PCfg11 = PairCfg<64, 4, 8, 3, 1, 2>;  // the default tile in 2-CTA clusters along y (DSMEM ring rows, cluster barrier)
struct PairVariant {
  int tx, ty, threads, smem, ctas_per_sm;
  void (*kernel)(const PairParams);
  const char* name;
  int cluster_y;
};
const PairVariant kPairVariants[] = {
    {PCfg0::kTX, PCfg0::kTY, PCfg0::kThreads, PCfg0::kSmemBytes, 1, pair_step_kernel<PCfg0>, "pair 64x32x4st"},
    {PCfg1::kTX, PCfg1::kTY, PCfg1::kThreads, PCfg1::kSmemBytes, 1, pair_step_kernel<PCfg1>, "pair 64x32x3st"},
    {PCfg2::kTX, PCfg2::kTY, PCfg2::kThreads, PCfg2::kSmemBytes, 1, pair_step_kernel<PCfg2>, "pair 128x32x2st"},
    {PCfg3::kTX, PCfg3::kTY, PCfg3::kThreads, PCfg3::kSmemBytes, 1, pair_step_kernel<PCfg3>, "pair 64x32x4st rpt4"},
    {PCfg4::kTX, PCfg4::kTY, PCfg4::kThreads, PCfg4::kSmemBytes, 1, pair_step_kernel<PCfg4>, "pair 64x32x3st rpt4"},
    {PCfg5::kTX, PCfg5::kTY, PCfg5::kThreads, PCfg5::kSmemBytes, 1, pair_step_kernel<PCfg5>, "pair 64x30x4st rpt2"},
    {PCfg6::kTX, PCfg6::kTY, PCfg6::kThreads, PCfg6::kSmemBytes, 2, pair_step_kernel<PCfg6>, "pair 64x16x3st rpt4 2cta"},
    {PCfg7::kTX, PCfg7::kTY, PCfg7::kThreads, PCfg7::kSmemBytes, 2, pair_step_kernel<PCfg7>, "pair 64x16x4st rpt4 2cta"},
    {PCfg8::kTX, PCfg8::kTY, PCfg8::kThreads, PCfg8::kSmemBytes, 1, pair_step_kernel<PCfg8>, "pair 64x30x3st rpt5"},
    {PCfg9::kTX, PCfg9::kTY, PCfg9::kThreads, PCfg9::kSmemBytes, 1, pair_step_kernel<PCfg9>, "pair 64x30x3st rpt3"},
    {PCfg10::kTX, PCfg10::kTY, PCfg10::kThreads, PCfg10::kSmemBytes, 1, pair_step_kernel<PCfg10>, "pair 64x30x4st rpt6"},
    {PCfg11::kTX, PCfg11::kTY, PCfg11::kThreads, PCfg11::kSmemBytes, 1, pair_step_kernel<PCfg11>, "pair 64x32x3st rpt4 cluster 1x2 (DSMEM)", 2},
};
constexpr int kNumPairVariants = sizeof(kPairVariants) / sizeof(kPairVariants[0]);
constexpr int kDefaultVariant = 3;   // best of the measured sweep at 512^3 (profiles/variant_sweep.md)

// ---- workspace layout ---------------------------------------------------------------
constexpr size_t kCounterBytes = 256;   // tickets [0..7] at 0, face-CTA counter at +64, exchange statistics (XStat) at +128
constexpr size_t kXStatOffset = 128;
static_assert(kXStatOffset + sizeof(XStat) <= kCounterBytes, "workspace header layout");
constexpr size_t kOrderBytes = 2 * (size_t)W3D_MAX_PARTIALS * sizeof(int);   // block -> tile tables (even / odd steps)
constexpr size_t kSlotBytes = 2ull * W3D_MAX_PARTIALS * sizeof(double);
constexpr size_t kSlotBase = kCounterBytes + kOrderBytes;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

}  // namespace

struct W3dPlan {
  W3dPlanDesc d;
  Geometry g;
  int variant;
  int num_sms;
  CUtensorMap halo[4], core[4];    // per array (2, 3 = aux arrays once w3d_plan_enable_pairs has run)
  bool maps_ok;
  // fused geometry: the TMA role covers the whole block; CTAs that touch a face
  // with a neighbour come first in `order`
  TileGeom tile;
  int n_ctas, n_face_ctas;
  int* order_dev;                  // lives in the workspace
  bool any_peer;
  const double* ct_table;          // optional device table of cos(a_t t_n), host-fed
  // temporal blocking: four arrays (0,1 = desc.field, 2,3 = aux), two TMA maps each
  double* arr[4];
  double* peer_arr[6][4];          // the neighbours' four arrays, per face
  CUtensorMap pmap_a[4], pmap_b[4];
  int pair_variant;
  TileGeom ptile;
  int pn_ctas, pn_face_ctas;
  BoxTable shell;                  // layer next to faces with a neighbour (fix-up pass)
  bool pair_ready;
  int pair_mode;                   // W3D_PAIR_LOCAL / WIDE / FIXUP
  int step_depth;                  // halo layers the single-step kernel ships: 1, or 2 in wide-halo plans
  W3dEdgePeer edges[W3D_EDGES];    // diagonal neighbours across the edges (wide halo on 3-D blocks)
  bool have_edges;
  bool defer;                      // wide halo, (nearly) all pair CTAs are face CTAs: flags are published at the next launch's entry
  int launch_seq;                  // launches issued through this plan (parity selects the XStat slot)
  unsigned long long* timeline;    // optional device buffer of (start, end) stamps per launch, indexed by level
  int timeline_cap;
  long long launches;
  char text[1024];
};

namespace {

ErrCtl err_ctl(const W3dPlan* p, int slot, double* out) {
  ErrCtl e;
  char* ws = static_cast<char*>(p->d.workspace);
  e.out = out;
  e.ticket = reinterpret_cast<unsigned*>(ws) + slot;
  e.part_max = reinterpret_cast<double*>(ws + kSlotBase + (size_t)slot * kSlotBytes);
  e.part_sum = e.part_max + W3D_MAX_PARTIALS;
  e.n_bcast = 0;
  return e;
}

// One 3-D tiled map over an array.  `ghost` layers of rows / planes in front of local index 0
// are part of the map (the array pointer is moved back accordingly), so a kernel addresses
// local (x, y, z) as (xoff + x, y + ghost - 1 ... see Geometry::gs).
int encode_one(const W3dPlan* p, CUtensorMap* m, double* arr, int box_x, int box_y, int gs, const char* what) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return fail(W3D_EDRIVER, "cuTensorMapEncodeTiled is not available from the driver");
  const cuuint64_t dims[3] = {(cuuint64_t)p->g.pitch, (cuuint64_t)(p->g.n[1] + 2 + 2 * gs),
                              (cuuint64_t)(p->g.n[2] + 2 + 2 * gs)};
  const cuuint64_t strides[2] = {(cuuint64_t)p->g.pitch * 8, (cuuint64_t)p->g.plane * 8};
  const cuuint32_t estr[3] = {1, 1, 1};
  const cuuint32_t box[3] = {(cuuint32_t)box_x, (cuuint32_t)box_y, 1};
  double* base = arr - (long long)gs * (p->g.plane + p->g.pitch);
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(W3D_EDRIVER, "cuTensorMapEncodeTiled(%s) failed: %d", what, (int)r);
  return W3D_OK;
}

// maps of the single-step kernel for array q (one ghost layer is all it addresses)
int encode_step_maps(W3dPlan* p, int q, double* arr) {
  const Variant& v = kVariants[p->variant];
  int rc = encode_one(p, &p->halo[q], arr, v.tx + 4, v.ty + 2, 0, "halo");
  if (rc == W3D_OK) rc = encode_one(p, &p->core[q], arr, v.tx, v.ty, 0, "core");
  return rc;
}

int encode_maps(W3dPlan* p) {
  for (int f = 0; f < 2; ++f) {
    const int rc = encode_step_maps(p, f, p->d.field[f]);
    if (rc != W3D_OK) return rc;
  }
  p->maps_ok = true;
  return W3D_OK;
}

// pick planes-per-CTA so that the CTA count fills whole waves of the resident slots
void plan_tiles(const W3dPlan* p, const int lo[3], const int hi[3], int zchunk, TileGeom* t) {
  const Variant& v = kVariants[p->variant];
  for (int d = 0; d < 3; ++d) { t->lo[d] = lo[d]; t->hi[d] = hi[d]; }
  const int ex = std::max(0, hi[0] - lo[0]), ey = std::max(0, hi[1] - lo[1]), ez = std::max(0, hi[2] - lo[2]);
  // TMA boxes must start on an even element: shift the tile origin left by one if needed
  t->x0 = lo[0] - ((p->g.xoff + lo[0]) & 1);
  t->ntx = (hi[0] - t->x0 + v.tx - 1) / v.tx;
  t->nty = (ey + v.ty - 1) / v.ty;
  if (ex == 0 || ey == 0 || ez == 0) { t->ntx = t->nty = 0; t->zc = 1; t->nchunks = 0; return; }
  const long long tiles = (long long)t->ntx * t->nty;
  if (zchunk > 0) {
    t->zc = std::min(zchunk, ez);
  } else {
    const long long slots = (long long)p->num_sms * v.ctas_per_sm;
    double best = 1e300;
    int best_c = 1;
    const int cmax = std::max(1, ez / 16);
    for (int c = 1; c <= cmax; ++c) {
      const int zc = (ez + c - 1) / c;
      const long long ctas = tiles * ((ez + zc - 1) / zc);
      if (ctas > W3D_MAX_PARTIALS - 1024) break;
      const long long waves = (ctas + slots - 1) / slots;
      // per-CTA cost: zc planes + 2 halo planes (read only, ~1/3 of the traffic) + launch/drain
      const double cost = (double)waves * (zc + 2.0 / 3.0 + 1.5);
      if (cost < best - 1e-9) { best = cost; best_c = c; }
    }
    t->zc = (ez + best_c - 1) / best_c;
  }
  t->nchunks = (ez + t->zc - 1) / t->zc;
}

// first row of tile row `ty` of the pair kernel: a last tile row that would be partial is moved back so that
// it is complete (mirror of the y0 computation in pair_step_kernel)
int pair_tile_y0(const TileGeom& t, int ty, int tile_y) {
  const int y = t.lo[1] + ty * tile_y;
  return (y + tile_y > t.hi[1] && t.hi[1] - tile_y >= t.lo[1]) ? t.hi[1] - tile_y : y;
}

// the diagonal neighbours' array `ship` as kernel parameters (wide-halo plans on 3-D blocks)
void fill_edges(const W3dPlan* p, EdgeDev* out, int ship) {
  for (int e = 0; e < kEdges; ++e) {
    const W3dEdgePeer& q = p->edges[e];
    out[e].active = (p->pair_mode == W3D_PAIR_WIDE && q.active) ? 1 : 0;
    out[e].ga = q.ghost_a;
    out[e].gb = q.ghost_b;
    out[e].field = q.arr[ship];
    out[e].pitch = q.pitch;
    out[e].plane = q.plane;
    out[e].flag = q.flag;
  }
}

// host mirror of tile_face_mask() (kernels.cuh)
int face_mask_host(const W3dPlanDesc& d, const int n[3], const int lo[3], const int hi[3], int reach) {
  int m = 0;
  for (int a = 0; a < 3; ++a) {
    if (d.peer[a * 2].active && lo[a] <= reach && hi[a] > 1) m |= 1 << (a * 2);
    if (d.peer[a * 2 + 1].active && lo[a] <= n[a] && hi[a] > n[a] - reach + 1) m |= 1 << (a * 2 + 1);
  }
  return m;
}

void add_box(BoxTable* B, const int lo[3], const int hi[3]) {
  for (int d = 0; d < 3; ++d)
    if (hi[d] <= lo[d]) return;
  const int b = B->nboxes++;
  long long cnt = 1;
  for (int d = 0; d < 3; ++d) {
    B->lo[b][d] = lo[d];
    B->ext[b][d] = hi[d] - lo[d];
    cnt *= B->ext[b][d];
  }
  B->prefix[b + 1] = B->prefix[b] + cnt;
}

// block index -> tile index of the single-step kernel: tiles that own one of the step_depth
// outermost layers of a face with a neighbour come first (the hardware dispatches blocks in
// index order), so the exchange starts at the top of the kernel
int order_step_tiles(W3dPlan* p) {
  const Variant& v = kVariants[p->variant];
  std::vector<int> face, rest;
  const TileGeom& t = p->tile;
  for (int w = 0; w < p->n_ctas; ++w) {
    const int tx = w % t.ntx, ty = (w / t.ntx) % t.nty, ch = w / (t.ntx * t.nty);
    const int lo[3] = {t.x0 + tx * v.tx, t.lo[1] + ty * v.ty, t.lo[2] + ch * t.zc};
    const int hi[3] = {lo[0] + v.tx, lo[1] + v.ty, std::min(lo[2] + t.zc, t.hi[2])};
    (face_mask_host(p->d, p->g.n, lo, hi, p->step_depth) ? face : rest).push_back(w);
  }
  p->n_face_ctas = (int)face.size();
  face.insert(face.end(), rest.begin(), rest.end());
  if (face.empty()) return W3D_OK;
  CUDA_TRY(cudaMemcpy(p->order_dev, face.data(), face.size() * sizeof(int), cudaMemcpyHostToDevice));
  return W3D_OK;
}

void fill_common(const W3dPlan* p, StepParams* P, int s, double a, double b, double cscale, double ct,
                 int cur = -1, int nxt = -1) {
  std::memset(P, 0, sizeof(*P));
  if (cur < 0) { cur = s & 1; nxt = cur ^ 1; }
  P->map_halo = p->halo[cur];
  P->map_core = p->core[nxt];
  P->g = p->g;
  P->cur = cur < 2 ? p->d.field[cur] : p->arr[cur];
  P->nxt = nxt < 2 ? p->d.field[nxt] : p->arr[nxt];
  P->sx = p->d.tab[0];
  P->sy = p->d.tab[1];
  P->sz = p->d.tab[2];
  P->a = a;
  P->b = b;
  P->cx = cscale * p->d.coef[0];
  P->cy = cscale * p->d.coef[1];
  P->cz = cscale * p->d.coef[2];
  P->ct = ct;
  P->ct_ptr = p->ct_table ? p->ct_table + (s + 1) : nullptr;
  P->status = p->d.status;
  P->level = (unsigned long long)s;
  P->release = (unsigned long long)s + 1ull;
  P->halo_depth = p->step_depth;
  P->seq = const_cast<W3dPlan*>(p)->launch_seq++;
  P->defer_release = p->defer ? 1 : 0;
  P->tl = (p->timeline && s >= 0 && s < p->timeline_cap) ? p->timeline + 2 * (size_t)s : nullptr;
  P->xstat = reinterpret_cast<XStat*>(static_cast<char*>(p->d.workspace) + kXStatOffset);
  P->timeout_ns = (unsigned long long)(p->d.spin_timeout_ms > 0 ? p->d.spin_timeout_ms : 10000) * 1000000ull;
}

int launch_fused(W3dPlan* p, const StepParams& P, int nblocks, cudaStream_t st) {
  const Variant& v = kVariants[p->variant];
  if (nblocks <= 0) return W3D_OK;
  if (nblocks > W3D_MAX_PARTIALS) return fail(W3D_EINVAL, "launch of %d CTAs exceeds W3D_MAX_PARTIALS", nblocks);
  v.kernel<<<nblocks, v.threads, v.smem, st>>>(P);
  CUDA_TRY(cudaGetLastError());
  ++p->launches;
  return W3D_OK;
}

}  // namespace

extern "C" {

const char* w3d_last_error(void) { return g_err; }
const char* w3d_version(void) { return "wave3d-b200 1.0 (sm_100a)"; }
size_t w3d_workspace_bytes(void) { return kSlotBase + (size_t)W3D_ERR_SLOTS * kSlotBytes; }
size_t w3d_error_workspace_bytes(void) { return kCounterBytes + kSlotBytes; }

int64_t w3d_pitch_for(int32_t nx) {
  const int64_t need = (int64_t)W3D_XOFF + nx + 2;
  return (need + W3D_PITCH_ALIGN - 1) / W3D_PITCH_ALIGN * W3D_PITCH_ALIGN;
}

int w3d_plan_create(const W3dPlanDesc* desc, W3dPlan** out) {
  if (!desc || !out) return fail(W3D_EINVAL, "null argument");
  const W3dBlock& b = desc->blk;
  if (b.n[0] < 1 || b.n[1] < 1 || b.n[2] < 1) return fail(W3D_EINVAL, "empty block");
  if (b.pitch % 2 != 0 || b.pitch < b.xoff + b.n[0] + 2 || b.xoff < 1)
    return fail(W3D_EINVAL, "bad layout: pitch=%lld xoff=%d nx=%d (pitch must be even and >= xoff+nx+2, xoff >= 1)",
                (long long)b.pitch, b.xoff, b.n[0]);
  if (!desc->field[0] || !desc->field[1] || !desc->tab[0] || !desc->tab[1] || !desc->tab[2])
    return fail(W3D_EINVAL, "null field or table pointer");
  if ((reinterpret_cast<uintptr_t>(desc->field[0]) | reinterpret_cast<uintptr_t>(desc->field[1])) & 15)
    return fail(W3D_EINVAL, "field arrays must be 16-byte aligned");
  if (!desc->workspace || desc->workspace_bytes < w3d_workspace_bytes())
    return fail(W3D_EINVAL, "workspace too small: need %zu bytes", w3d_workspace_bytes());
  if (!desc->status || !desc->flags_in) return fail(W3D_EINVAL, "status / flags_in must be provided");
  W3dPlan* p = new (std::nothrow) W3dPlan();
  if (!p) return fail(W3D_ENOMEM, "out of host memory");
  p->d = *desc;
  p->g.n[0] = b.n[0]; p->g.n[1] = b.n[1]; p->g.n[2] = b.n[2];
  p->g.xoff = b.xoff;
  p->g.pitch = b.pitch;
  p->g.gs = desc->ghost_depth >= 2 ? 1 : 0;
  p->g.plane = desc->plane_stride > 0 ? desc->plane_stride : (long long)(b.n[1] + 2) * b.pitch;
  if (p->g.plane < (long long)(b.n[1] + 2 + 2 * p->g.gs) * b.pitch || p->g.plane % 2 != 0) {
    delete p;
    return fail(W3D_EINVAL, "plane_stride %lld too small for %d rows with ghost depth %d (or odd)",
                (long long)desc->plane_stride, b.n[1], p->g.gs + 1);
  }
  if (p->g.gs && (desc->peer[0].active || desc->peer[1].active) && b.pitch < b.xoff + b.n[0] + 3) {
    delete p;
    return fail(W3D_EINVAL, "ghost depth 2 with an x neighbour needs pitch >= xoff + nx + 3 and xoff >= 2");
  }
  p->pair_mode = W3D_PAIR_LOCAL;
  p->step_depth = 1;
  p->timeline = nullptr;
  p->timeline_cap = 0;
  p->launch_seq = 0;
  p->defer = false;
  p->have_edges = false;
  std::memset(p->edges, 0, sizeof(p->edges));
  p->variant = desc->variant >= 0 && desc->variant < kNumVariants ? desc->variant : kDefaultVariant;
  p->num_sms = desc->num_sms;
  if (p->num_sms <= 0) {
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    p->num_sms = sms;
  }
  p->maps_ok = false;
  p->pair_ready = false;
  p->launches = 0;
  p->ct_table = nullptr;
  p->any_peer = false;
  for (int f = 0; f < 6; ++f) p->any_peer = p->any_peer || desc->peer[f].active;
  const Variant& v = kVariants[p->variant];
  {
    const int lo[3] = {1, 1, 1}, hi[3] = {b.n[0] + 1, b.n[1] + 1, b.n[2] + 1};
    plan_tiles(p, lo, hi, desc->zchunk, &p->tile);
  }
  // shell = the layer next to every face with a neighbour, as disjoint boxes (z slabs, then y, then x)
  std::memset(&p->shell, 0, sizeof(p->shell));
  {
    int lo[3] = {1, 1, 1}, hi[3] = {b.n[0] + 1, b.n[1] + 1, b.n[2] + 1};
    for (int d = 2; d >= 0; --d) {
      const int ilo = 1 + (desc->peer[d * 2].active ? 1 : 0);
      const int ihi = std::max(ilo, b.n[d] + 1 - (desc->peer[d * 2 + 1].active ? 1 : 0));
      int l2[3] = {lo[0], lo[1], lo[2]}, h2[3] = {hi[0], hi[1], hi[2]};
      if (ilo > lo[d]) { h2[d] = ilo; add_box(&p->shell, l2, h2); h2[d] = hi[d]; }
      if (ihi < hi[d]) { l2[d] = ihi; add_box(&p->shell, l2, h2); }
      lo[d] = ilo;
      hi[d] = ihi;
    }
  }
  p->n_ctas = p->tile.ntx * p->tile.nty * p->tile.nchunks;
  if (p->n_ctas > W3D_MAX_PARTIALS) {
    const int need = p->n_ctas;
    delete p;
    return fail(W3D_EINVAL, "block needs %d CTAs, limit %d: raise zchunk", need, W3D_MAX_PARTIALS);
  }
  p->order_dev = reinterpret_cast<int*>(static_cast<char*>(desc->workspace) + kCounterBytes);
  {
    const int rc = order_step_tiles(p);
    if (rc != W3D_OK) { delete p; return rc; }
  }
  cudaError_t e = cudaFuncSetAttribute(v.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, v.smem);
  if (e != cudaSuccess) {
    delete p;
    return fail(W3D_ECUDA, "cudaFuncSetAttribute(smem=%d): %s", v.smem, cudaGetErrorString(e));
  }
  int rc = encode_maps(p);
  if (rc != W3D_OK) { delete p; return rc; }
  snprintf(p->text, sizeof(p->text),
           "block %dx%dx%d pitch %lld | variant %s (%d thr, %d B smem, %d CTA/SM) | tiles %dx%d zc %d chunks %d -> "
           "%d CTAs, %d of them face CTAs (numbered first, push halos over P2P) | SMs %d",
           b.n[0], b.n[1], b.n[2], (long long)b.pitch, v.name, v.threads, v.smem, v.ctas_per_sm, p->tile.ntx,
           p->tile.nty, p->tile.zc, p->tile.nchunks, p->n_ctas, p->n_face_ctas, p->num_sms);
  *out = p;
  return W3D_OK;
}

void w3d_plan_destroy(W3dPlan* plan) { delete plan; }
int w3d_plan_set_ct_table(W3dPlan* plan, const double* dev_table) {
  if (!plan) return fail(W3D_EINVAL, "null plan");
  plan->ct_table = dev_table;
  return W3D_OK;
}
const char* w3d_plan_describe(W3dPlan* plan) { return plan ? plan->text : ""; }
long long w3d_plan_launches(W3dPlan* plan) { return plan ? plan->launches : 0; }

int w3d_plan_init(W3dPlan* p, void* stream) { return w3d_plan_init_ex(p, stream, 1); }

int w3d_plan_init_ex(W3dPlan* p, void* stream, int zero_second) {
  if (!p) return fail(W3D_EINVAL, "null plan");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int gs = p->g.gs;
  const long long rows = (long long)(p->g.n[2] + 2 + 2 * gs) * (p->g.n[1] + 2 + 2 * gs);
  const int blocks = (int)std::max<long long>(1, std::min<long long>((rows + 7) / 8, (long long)p->num_sms * 32));
  init_kernel<<<blocks, 256, 0, st>>>(p->d.field[0], zero_second ? p->d.field[1] : nullptr, p->g, p->d.tab[0],
                                      p->d.tab[1], p->d.tab[2]);
  CUDA_TRY(cudaGetLastError());
  ++p->launches;
  CUDA_TRY(cudaMemsetAsync(p->d.flags_in, 0, (p->have_edges ? W3D_FLAG_WORDS : 6) * sizeof(unsigned long long), st));
  CUDA_TRY(cudaMemsetAsync(p->d.status, 0, sizeof(uint32_t), st));
  CUDA_TRY(cudaMemsetAsync(p->d.workspace, 0, kCounterBytes, st));
  return W3D_OK;
}

static int step_fused_impl(W3dPlan* p, int s, int cur, int nxt, double a, double b, double cscale, double ct,
                           double* err_out, void* stream) {
  StepParams P;
  fill_common(p, &P, s, a, b, cscale, ct, cur, nxt);
  P.err = err_ctl(p, 0, err_out);
  if (err_out && p->d.err_table && p->d.n_err_peers > 0) {
    const ptrdiff_t off = err_out - p->d.err_table;   // same offset in every rank's table
    P.err.n_bcast = std::min(p->d.n_err_peers, 7);
    for (int q = 0; q < P.err.n_bcast; ++q) P.err.bcast[q] = p->d.err_table_peer[q] + off;
  }
  P.tile = p->tile;
  P.order = p->n_face_ctas > 0 ? p->order_dev : nullptr;   // identity when nothing to reorder
  P.n_face_ctas = p->n_face_ctas;
  P.use_flags = p->any_peer ? 1 : 0;
  for (int f = 0; f < 6; ++f) {
    const W3dPeer& q = p->d.peer[f];
    P.peer[f].active = q.active;
    P.peer[f].ghost_index = q.ghost_index;
    P.peer[f].field = nxt < 2 ? q.field[nxt] : p->peer_arr[f][nxt];
    P.peer[f].pitch = q.pitch;
    P.peer[f].plane = q.plane;
    P.peer[f].flag = q.flag;
  }
  fill_edges(p, P.edge, nxt);
#ifdef W3D_DIAG
  if (const char* dbg = std::getenv("W3D_DEBUG_SKIP")) P.debug_skip = std::atoi(dbg);   // timing diagnostics, wrong results
#endif
  P.flags_in = p->d.flags_in;
  P.shell_counter = reinterpret_cast<unsigned*>(static_cast<char*>(p->d.workspace) + 64);
  return launch_fused(p, P, p->n_ctas, static_cast<cudaStream_t>(stream));
}

int w3d_step_fused(W3dPlan* p, int s, double a, double b, double cscale, double ct, double* err_out,
                   void* stream) {
  if (!p) return fail(W3D_EINVAL, "null plan");
  return step_fused_impl(p, s, s & 1, (s & 1) ^ 1, a, b, cscale, ct, err_out, stream);
}

int w3d_step_fused_on(W3dPlan* p, int s, int cur, int prev, double a, double b, double cscale, double ct,
                      double* err_out, void* stream) {
  if (!p) return fail(W3D_EINVAL, "null plan");
  if (cur < 0 || cur > 3 || prev < 0 || prev > 3 || cur == prev) return fail(W3D_EINVAL, "bad array indices");
  if ((cur > 1 || prev > 1) && !p->pair_ready) return fail(W3D_EINVAL, "arrays 2, 3 exist after w3d_plan_enable_pairs only");
  return step_fused_impl(p, s, cur, prev, a, b, cscale, ct, err_out, stream);
}

int w3d_plan_exchange_stats(W3dPlan* p, unsigned long long out[W3D_XSTAT_WORDS], void* stream) {
  if (!p || !out) return fail(W3D_EINVAL, "null argument");
  XStat h;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CUDA_TRY(cudaMemcpyAsync(&h, static_cast<char*>(p->d.workspace) + kXStatOffset, sizeof(h), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  // (per-launch maxima that no later launch has folded yet: the last launch of a deferred-release plan)
  const unsigned long long pw = std::max(h.cur_wait[0], h.cur_wait[1]), pc = std::max(h.cur_copy[0], h.cur_copy[1]);
  out[0] = h.sum_wait + pw; out[1] = h.sum_copy + pc; out[2] = h.launches;
  out[3] = std::max(h.max_wait, pw); out[4] = std::max(h.max_copy, pc);
  return W3D_OK;
}

int w3d_plan_set_timeline(W3dPlan* p, unsigned long long* dev_buf, int capacity) {
  if (!p) return fail(W3D_EINVAL, "null plan");
  p->timeline = dev_buf;
  p->timeline_cap = dev_buf ? capacity : 0;
  return W3D_OK;
}

int w3d_plan_set_edges(W3dPlan* p, const W3dEdgePeer* edges) {
  if (!p) return fail(W3D_EINVAL, "null plan");
  if (p->pair_ready) return fail(W3D_EINVAL, "w3d_plan_set_edges must come before w3d_plan_enable_pairs");
  std::memset(p->edges, 0, sizeof(p->edges));
  p->have_edges = false;
  if (!edges) return W3D_OK;
  for (int e = 0; e < W3D_EDGES; ++e) {
    if (!edges[e].active) continue;
    if (!edges[e].flag || !edges[e].arr[0] || !edges[e].arr[1] || !edges[e].arr[2] || !edges[e].arr[3])
      return fail(W3D_EINVAL, "edge %d is active but has a null array or flag pointer", e);
    p->edges[e] = edges[e];
    p->have_edges = true;
  }
  return W3D_OK;
}

int w3d_plan_pairs_enabled(W3dPlan* p) { return p && p->pair_ready ? 1 : 0; }

int w3d_plan_pair_mode(W3dPlan* p) { return p ? p->pair_mode : W3D_PAIR_LOCAL; }

int w3d_step_tma(W3dPlan* p, int s, double a, double b, double cscale, double ct, const int32_t lo[3],
                 const int32_t hi[3], double* err_out, int slot, void* stream) {
  if (!p) return fail(W3D_EINVAL, "null plan");
  if (slot < 0 || slot >= W3D_ERR_SLOTS) return fail(W3D_EINVAL, "bad slot %d", slot);
  StepParams P;
  fill_common(p, &P, s, a, b, cscale, ct);
  P.err = err_ctl(p, slot, err_out);
  int l[3], h[3];
  for (int d = 0; d < 3; ++d) {
    l[d] = lo ? lo[d] : 1;
    h[d] = hi ? hi[d] : p->g.n[d] + 1;
    if (l[d] < 1 || h[d] > p->g.n[d] + 1) return fail(W3D_EINVAL, "box outside the block");
  }
  plan_tiles(p, l, h, p->d.zchunk, &P.tile);
  P.n_face_ctas = 0;
  P.order = nullptr;
  P.use_flags = 0;
  const int nblocks = P.tile.ntx * P.tile.nty * P.tile.nchunks;
  if (nblocks == 0 && err_out) {
    CUDA_TRY(cudaMemsetAsync(err_out, 0, 2 * sizeof(double), static_cast<cudaStream_t>(stream)));
    return W3D_OK;
  }
  return launch_fused(p, P, nblocks, static_cast<cudaStream_t>(stream));
}

int w3d_step_box(W3dPlan* p, int s, double a, double b, double cscale, double ct, const int32_t lo[3],
                 const int32_t hi[3], double* err_out, int slot, void* stream) {
  if (!p) return fail(W3D_EINVAL, "null plan");
  if (slot < 0 || slot >= W3D_ERR_SLOTS) return fail(W3D_EINVAL, "bad slot %d", slot);
  StepParams P;
  fill_common(p, &P, s, a, b, cscale, ct);
  P.err = err_ctl(p, slot, err_out);
  int l[3], h[3];
  for (int d = 0; d < 3; ++d) {
    l[d] = lo ? lo[d] : 1;
    h[d] = hi ? hi[d] : p->g.n[d] + 1;
    if (l[d] < 1 || h[d] > p->g.n[d] + 1) return fail(W3D_EINVAL, "box outside the block");
  }
  add_box(&P.boxes, l, h);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long nodes = P.boxes.prefix[P.boxes.nboxes];
  if (nodes == 0) {
    if (err_out) CUDA_TRY(cudaMemsetAsync(err_out, 0, 2 * sizeof(double), st));
    return W3D_OK;
  }
  const int blocks = (int)std::min<long long>((nodes + 255) / 256, (long long)W3D_MAX_PARTIALS);
  box_step_kernel<<<blocks, 256, 0, st>>>(P);
  CUDA_TRY(cudaGetLastError());
  ++p->launches;
  return W3D_OK;
}

// per-CTA cost of a pair pass in "plane units" for a tile of ty rows and zc planes: zc planes of
// work plus 4 halo planes that are loaded and partly computed, plus a fixed fill / drain term
static double pair_cta_cost(int ty, int zc) { return (double)ty / 32.0 * (zc + 4.0 * 0.45 + 2.0); }

// tile geometry of pair variant `vi` over the whole block; returns the modelled pass cost
static double plan_pair_tiles(const W3dPlan* p, int vi, int zchunk, TileGeom* out) {
  const PairVariant& v = kPairVariants[vi];
  TileGeom t;
  for (int d = 0; d < 3; ++d) { t.lo[d] = 1; t.hi[d] = p->g.n[d] + 1; }
  t.x0 = 1 - ((p->g.xoff + 1) & 1);
  t.ntx = (t.hi[0] - t.x0 + v.tx - 1) / v.tx;
  t.nty = (p->g.n[1] + v.ty - 1) / v.ty;
  const int ez = p->g.n[2];
  const long long tiles = (long long)t.ntx * t.nty;
  const long long slots = (long long)p->num_sms * v.ctas_per_sm;
  double best = 1e300;
  if (zchunk > 0) {
    t.zc = std::min(std::min(zchunk, ez), 636);
    const long long ctas = tiles * ((ez + t.zc - 1) / t.zc);
    best = (double)((ctas + slots - 1) / slots) * pair_cta_cost(v.ty, t.zc);
  } else {
    int best_c = std::max(1, (ez + 635) / 636);
    for (int c = 1; c <= std::max(1, ez / 16); ++c) {
      const int zc = (ez + c - 1) / c;
      if (zc + 4 > 640) continue;                 // the kernel stages zc+4 table entries in shared memory
      const long long ctas = tiles * ((ez + zc - 1) / zc);
      if (ctas > W3D_MAX_PARTIALS) break;
      const long long waves = (ctas + slots - 1) / slots;
      const double cost = (double)waves * pair_cta_cost(v.ty, zc);
      if (cost < best - 1e-9) { best = cost; best_c = c; }
    }
    t.zc = (ez + best_c - 1) / best_c;
  }
  t.nchunks = (ez + t.zc - 1) / t.zc;
  *out = t;
  return best;
}

int w3d_plan_enable_pairs(W3dPlan* p, double* aux0, double* aux1, double* const* peer_aux, int pair_variant,
                          int zchunk) {
  if (!p || !aux0 || !aux1) return fail(W3D_EINVAL, "null argument");
  if ((reinterpret_cast<uintptr_t>(aux0) | reinterpret_cast<uintptr_t>(aux1)) & 15)
    return fail(W3D_EINVAL, "aux arrays must be 16-byte aligned");
  p->arr[0] = p->d.field[0]; p->arr[1] = p->d.field[1]; p->arr[2] = aux0; p->arr[3] = aux1;
  // protocol for faces with a neighbour (see wave3d.h)
  int active_axes = 0;
  for (int d = 0; d < 3; ++d) active_axes += (p->d.peer[d * 2].active || p->d.peer[d * 2 + 1].active) ? 1 : 0;
  // wide halo needs two ghost layers and, wherever two faces on different axes both have a neighbour, the rank
  // diagonally across that edge (w3d_plan_set_edges)
  bool edges_ok = true;
  for (int e = 0; e < kEdges; ++e) {
    int a, b, sa, sb;
    edge_axes(e, &a, &b, &sa, &sb);
    if (p->d.peer[a * 2 + sa].active && p->d.peer[b * 2 + sb].active && !p->edges[e].active) edges_ok = false;
  }
  p->pair_mode = active_axes == 0 ? W3D_PAIR_LOCAL : (p->g.gs == 1 && (active_axes == 1 || edges_ok)) ? W3D_PAIR_WIDE : W3D_PAIR_FIXUP;
  if (p->pair_mode == W3D_PAIR_WIDE)
    for (int d = 0; d < 3; ++d)
      if ((p->d.peer[d * 2].active || p->d.peer[d * 2 + 1].active) && p->g.n[d] < 2)
        return fail(W3D_EINVAL, "wide halo needs at least 2 unknowns along the axis with neighbours");
  // variant: given, or the best of the measured sweep.  (The 64x30 tilings put 144 instead of 128 CTAs on
  // the 148 SMs for 511 rows, but with 12 instead of 16 warps per SM each CTA is as slow as a 64x32 one:
  // measured 0.0626 vs 0.0613 ms per step on the 511 x 511 x 64 block — profiles/pair_schedule.md.)
  p->pair_variant = (pair_variant >= 0 && pair_variant < kNumPairVariants) ? pair_variant : 4;
  plan_pair_tiles(p, p->pair_variant, zchunk, &p->ptile);
  if (p->pair_mode == W3D_PAIR_WIDE && p->step_depth != 2) {
    // the single-step kernel keeps the wide halo alive too (it ships two layers), so single steps
    // and pair passes can follow each other in any order
    p->step_depth = 2;
    const int rc = order_step_tiles(p);
    if (rc != W3D_OK) return rc;
  }
  const PairVariant& v = kPairVariants[p->pair_variant];
  for (int f = 0; f < 4; ++f) {
    int rc = encode_one(p, &p->pmap_a[f], p->arr[f], v.tx + 4, v.ty + 4, p->g.gs, "pair A");
    if (rc == W3D_OK) rc = encode_one(p, &p->pmap_b[f], p->arr[f], v.tx + 4, v.ty + 2, p->g.gs, "pair B");
    if (rc == W3D_OK && f >= 2) rc = encode_step_maps(p, f, p->arr[f]);
    if (rc != W3D_OK) return rc;
  }
  CUDA_TRY(cudaFuncSetAttribute(v.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, v.smem));
  TileGeom& t = p->ptile;
  p->pn_ctas = t.ntx * t.nty * t.nchunks;
  if (p->pn_ctas > W3D_MAX_PARTIALS) return fail(W3D_EINVAL, "pair kernel needs %d CTAs, limit %d", p->pn_ctas, W3D_MAX_PARTIALS);
  // neighbours' arrays by index; face CTAs first in the block -> tile table of the pair kernel
  for (int f = 0; f < 6; ++f) {
    p->peer_arr[f][0] = p->d.peer[f].field[0];
    p->peer_arr[f][1] = p->d.peer[f].field[1];
    p->peer_arr[f][2] = peer_aux ? peer_aux[f * 2] : nullptr;
    p->peer_arr[f][3] = peer_aux ? peer_aux[f * 2 + 1] : nullptr;
    if (p->d.peer[f].active && (!p->peer_arr[f][2] || !p->peer_arr[f][3]))
      return fail(W3D_EINVAL, "face %d has a neighbour but no peer aux arrays were given", f);
  }
  {
    // order: tiles that copy halo layers out (they wait, copy and are counted for the release),
    // then tiles that only read a ghost layer through their level-1 ring (they wait), then the rest.
    // Wide halo: two layers are shipped, so both sets are the tiles within 2 layers of a face.
    const int depth = p->pair_mode == W3D_PAIR_WIDE ? 2 : 1;
    std::vector<int> face, waiters, rest;
    auto tile_class = [&](int w) {      // 2: copies halo layers out, 1: only waits, 0: neither
      const int tx = w % t.ntx, ty = (w / t.ntx) % t.nty, ch = w / (t.ntx * t.nty);
      const int lo[3] = {t.x0 + tx * v.tx, pair_tile_y0(t, ty, v.ty), t.lo[2] + ch * t.zc};
      const int hi[3] = {lo[0] + v.tx, lo[1] + v.ty, std::min(lo[2] + t.zc, t.hi[2])};
      return face_mask_host(p->d, p->g.n, lo, hi, depth) ? 2 : face_mask_host(p->d, p->g.n, lo, hi, 2) ? 1 : 0;
    };
    int n_face = 0;
    if (v.cluster_y == 2) {
      // clusters are consecutive block indices: the two tiles of a y pair stay together (upper tile = rank 0)
      if (t.nty % 2 != 0) return fail(W3D_EINVAL, "the cluster variant needs an even number of tile rows (have %d)", t.nty);
      for (int ch = 0; ch < t.nchunks; ++ch)
        for (int ty = 0; ty < t.nty; ty += 2)
          for (int tx = 0; tx < t.ntx; ++tx) {
            const int w0 = (ch * t.nty + ty) * t.ntx + tx, w1 = w0 + t.ntx;
            const int c0 = tile_class(w0), c1 = tile_class(w1);
            n_face += (c0 == 2) + (c1 == 2);
            std::vector<int>& dst = std::max(c0, c1) == 2 ? face : std::max(c0, c1) == 1 ? waiters : rest;
            dst.push_back(w0);
            dst.push_back(w1);
          }
    } else {
      for (int w = 0; w < p->pn_ctas; ++w) {
        const int c = tile_class(w);
        (c == 2 ? face : c == 1 ? waiters : rest).push_back(w);
      }
      n_face = (int)face.size();
    }
    p->pn_face_ctas = n_face;
    // A system fence + counter in every face CTA at the end of the pass costs 4.8 us when every CTA is a face
    // CTA (one-wave launches on thin slabs): the flags are published by the NEXT launch at its entry instead
    // (entry_release, kernels.cuh).
    // Measured (2 GPUs, 512^3 and 1024^3): publishing at the next launch's entry is as good or better even when
    // the face CTAs are a minority, so it is the rule for every wide-halo plan; W3D_DEFER=0 switches back to the
    // in-kernel release for experiments.
    const char* de = std::getenv("W3D_DEFER");
    p->defer = p->pair_mode == W3D_PAIR_WIDE && !(de && de[0] == '0');
    face.insert(face.end(), waiters.begin(), waiters.end());
    face.insert(face.end(), rest.begin(), rest.end());
    CUDA_TRY(cudaMemcpy(p->order_dev + W3D_MAX_PARTIALS, face.data(), face.size() * sizeof(int),
                        cudaMemcpyHostToDevice));
  }
  p->pair_ready = true;
  const size_t used = std::strlen(p->text);
  static const char* kModeName[3] = {"no neighbours", "wide halo: 1 launch + 1 handshake per 2 steps", "fix-up pass: 2 launches per 2 steps"};
  snprintf(p->text + used, sizeof(p->text) - used,
           " || %s (%d thr, %d B smem): tiles %dx%d zc %d chunks %d -> %d CTAs, %d face CTAs [%s%s]", v.name, v.threads,
           v.smem, t.ntx, t.nty, t.zc, t.nchunks, p->pn_ctas, p->pn_face_ctas, kModeName[p->pair_mode],
           p->pair_mode != W3D_PAIR_WIDE ? "" : p->defer ? ", flags published at the next launch's entry" : ", flags released by the last face CTA");
  return W3D_OK;
}

namespace {
void set_bcast(const W3dPlan* p, ErrCtl* e, double* err_out) {
  if (err_out && p->d.err_table && p->d.n_err_peers > 0) {
    const ptrdiff_t off = err_out - p->d.err_table;   // same offset in every rank's table
    e->n_bcast = std::min(p->d.n_err_peers, 7);
    for (int q = 0; q < e->n_bcast; ++q) e->bcast[q] = p->d.err_table_peer[q] + off;
  }
}
}  // namespace

int w3d_step_pair(W3dPlan* p, int s, int cur, int prev, int out1, int out2, double a1, double b1, double cscale1,
                  double ct1, double a2, double b2, double cscale2, double ct2, double* err1, double* err2,
                  void* stream) {
  if (!p || !p->pair_ready) return fail(W3D_EINVAL, "w3d_plan_enable_pairs has not been called");
  const int idx[4] = {cur, prev, out1, out2};
  for (int i = 0; i < 4; ++i) {
    if (idx[i] < 0 || idx[i] > 3) return fail(W3D_EINVAL, "array index out of range");
    for (int j = 0; j < i; ++j)
      if (idx[i] == idx[j]) return fail(W3D_EINVAL, "the four arrays of a pair step must be distinct");
  }
  const PairVariant& v = kPairVariants[p->pair_variant];
  const bool wide = p->pair_mode == W3D_PAIR_WIDE;
  PairParams P;
  std::memset(&P, 0, sizeof(P));
  P.map_a = p->pmap_a[cur];
  P.map_b = p->pmap_b[prev];
  P.g = p->g;
  P.out1 = p->arr[out1];
  P.out2 = p->arr[out2];
  P.sx = p->d.tab[0]; P.sy = p->d.tab[1]; P.sz = p->d.tab[2];
  P.a1 = a1; P.b1 = b1; P.c1x = cscale1 * p->d.coef[0]; P.c1y = cscale1 * p->d.coef[1]; P.c1z = cscale1 * p->d.coef[2];
  P.a2 = a2; P.b2 = b2; P.c2x = cscale2 * p->d.coef[0]; P.c2y = cscale2 * p->d.coef[1]; P.c2z = cscale2 * p->d.coef[2];
  P.ct1 = ct1; P.ct2 = ct2;
  P.ct_ptr = p->ct_table ? p->ct_table + (s + 1) : nullptr;
  P.skip_b = (b1 == 0.0) ? 1 : 0;
  P.err1 = err_ctl(p, 0, err1);
  P.err2 = err_ctl(p, 1, err2);
  set_bcast(p, &P.err1, err1);
  set_bcast(p, &P.err2, err2);
  P.tile = p->ptile;
  P.order = (p->pn_face_ctas > 0 || v.cluster_y == 2) ? p->order_dev + W3D_MAX_PARTIALS : nullptr;   // identity when nothing to reorder
  P.wide = wide ? 1 : 0;
  for (int d = 0; d < 3; ++d) {
    const int lo_act = p->d.peer[d * 2].active ? 1 : 0, hi_act = p->d.peer[d * 2 + 1].active ? 1 : 0;
    P.skip_lo[d] = wide ? 0 : lo_act;
    P.skip_hi[d] = wide ? 0 : hi_act;
    P.in_lo[d] = 1 - (wide ? lo_act : 0);
    P.in_hi[d] = p->g.n[d] + (wide ? hi_act : 0);
  }
  // fused exchange: level s+1 (array out1, one layer) with the fix-up protocol,
  // level s+2 (array out2, two layers) with the wide halo
  const int ship = wide ? out2 : out1;
  P.nxt = p->arr[ship];
  P.halo_depth = wide ? 2 : 1;
  P.zpush = wide ? ((p->d.peer[4].active ? 1 : 0) | (p->d.peer[5].active ? 2 : 0)) : 0;
  P.n_face_ctas = p->pn_face_ctas;
  P.use_flags = p->any_peer ? 1 : 0;
  for (int f = 0; f < 6; ++f) {
    const W3dPeer& q = p->d.peer[f];
    P.peer[f].active = q.active;
    P.peer[f].ghost_index = q.ghost_index;
    P.peer[f].field = p->peer_arr[f][ship];
    P.peer[f].pitch = q.pitch;
    P.peer[f].plane = q.plane;
    P.peer[f].flag = q.flag;
  }
  fill_edges(p, P.edge, ship);
  P.flags_in = p->d.flags_in;
  P.shell_counter = reinterpret_cast<unsigned*>(static_cast<char*>(p->d.workspace) + 64);
  P.status = p->d.status;
  P.xstat = reinterpret_cast<XStat*>(static_cast<char*>(p->d.workspace) + kXStatOffset);
  P.tl = (p->timeline && s >= 0 && s < p->timeline_cap) ? p->timeline + 2 * (size_t)s : nullptr;
  P.seq = p->launch_seq++;
  P.defer_release = p->defer ? 1 : 0;
  P.level = (unsigned long long)s;
  P.release = (unsigned long long)s + (wide ? 2ull : 1ull);
  P.timeout_ns = (unsigned long long)(p->d.spin_timeout_ms > 0 ? p->d.spin_timeout_ms : 10000) * 1000000ull;
  // (programmatic dependent launch was tried here: the graph's launch gap is 1.5 us and with one CTA per SM the
  // successor has nowhere to start early — no gain, profiles/pair_schedule.md)
  if (v.cluster_y == 2) {
    cudaLaunchConfig_t cfg;
    std::memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)p->pn_ctas);
    cfg.blockDim = dim3((unsigned)v.threads);
    cfg.dynamicSmemBytes = (size_t)v.smem;
    cfg.stream = static_cast<cudaStream_t>(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    CUDA_TRY(cudaLaunchKernelEx(&cfg, v.kernel, P));
  } else {
    v.kernel<<<p->pn_ctas, v.threads, v.smem, static_cast<cudaStream_t>(stream)>>>(P);
    CUDA_TRY(cudaGetLastError());
  }
  ++p->launches;
  return W3D_OK;
}

int w3d_step_fixup(W3dPlan* p, int level, int cur, int prev, int out, double a, double b, double cscale,
                   double ct, double* err_out, void* stream) {
  if (!p || !p->pair_ready) return fail(W3D_EINVAL, "w3d_plan_enable_pairs has not been called");
  if (cur < 0 || cur > 3 || prev < 0 || prev > 3 || out < 0 || out > 3 || cur == out || prev == out)
    return fail(W3D_EINVAL, "bad array indices");
  if (p->pair_mode != W3D_PAIR_FIXUP) return fail(W3D_EINVAL, "this plan's pair passes need no fix-up (w3d_plan_pair_mode)");
  const long long nodes = p->shell.prefix[p->shell.nboxes];
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (nodes == 0) {
    if (err_out) CUDA_TRY(cudaMemsetAsync(err_out, 0, 2 * sizeof(double), st));
    return W3D_OK;
  }
  StepParams P;
  fill_common(p, &P, level, a, b, cscale, ct, 0, 1);   // (sets ct_ptr = ct_table + level + 1 when a table is attached)
  P.cur = p->arr[cur];
  P.nxt = p->arr[out];
  P.prev_sep = p->arr[prev];
  P.err = err_ctl(p, 2, err_out);
  set_bcast(p, &P.err, err_out);
  P.boxes = p->shell;
  P.use_flags = p->any_peer ? 1 : 0;
  for (int f = 0; f < 6; ++f) {
    const W3dPeer& q = p->d.peer[f];
    P.peer[f].active = q.active;
    P.peer[f].ghost_index = q.ghost_index;
    P.peer[f].field = p->peer_arr[f][out];
    P.peer[f].pitch = q.pitch;
    P.peer[f].plane = q.plane;
    P.peer[f].flag = q.flag;
  }
  P.flags_in = p->d.flags_in;
  P.shell_counter = reinterpret_cast<unsigned*>(static_cast<char*>(p->d.workspace) + 64);
  P.level = (unsigned long long)level;     // waits for `level`, releases level + 1
  const int blocks = (int)std::max<long long>(1, std::min<long long>((nodes + 2047) / 2048, (long long)p->num_sms));
  P.n_face_ctas = blocks;
  fixup_kernel<<<blocks, 1024, 0, st>>>(P);
  CUDA_TRY(cudaGetLastError());
  ++p->launches;
  return W3D_OK;
}

static int face_layer(const W3dPlan* p, int axis, int side, int ghost) {
  const int n = p->g.n[axis];
  return side == 0 ? (ghost ? 0 : 1) : (ghost ? n + 1 : n);
}

int w3d_pack_face(W3dPlan* p, int parity, int axis, int side, int ghost, double* buf, void* stream) {
  if (!p || !buf || axis < 0 || axis > 2) return fail(W3D_EINVAL, "bad argument");
  const int f0 = axis == 0 ? 1 : 0, f1 = axis == 2 ? 1 : 2;
  const long long total = (long long)p->g.n[f0] * p->g.n[f1];
  const int blocks = (int)std::min<long long>((total + 255) / 256, 4096);
  face_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(p->d.field[parity & 1], p->g, axis,
                                                                   face_layer(p, axis, side, ghost), buf, 0);
  CUDA_TRY(cudaGetLastError());
  ++p->launches;
  return W3D_OK;
}

int w3d_unpack_face(W3dPlan* p, int parity, int axis, int side, int ghost, const double* buf, void* stream) {
  if (!p || !buf || axis < 0 || axis > 2) return fail(W3D_EINVAL, "bad argument");
  const int f0 = axis == 0 ? 1 : 0, f1 = axis == 2 ? 1 : 2;
  const long long total = (long long)p->g.n[f0] * p->g.n[f1];
  const int blocks = (int)std::min<long long>((total + 255) / 256, 4096);
  face_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      p->d.field[parity & 1], p->g, axis, face_layer(p, axis, side, ghost), const_cast<double*>(buf), 1);
  CUDA_TRY(cudaGetLastError());
  ++p->launches;
  return W3D_OK;
}

int w3d_wrap_axis(W3dPlan* p, int parity, int axis, void* stream) {
  if (!p || axis < 0 || axis > 2) return fail(W3D_EINVAL, "bad argument");
  const int f0 = axis == 0 ? 1 : 0, f1 = axis == 2 ? 1 : 2;
  const long long total = (long long)(p->g.n[f0] + 2) * (p->g.n[f1] + 2);
  const int blocks = (int)std::min<long long>((total + 255) / 256, 4096);
  wrap_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(p->d.field[parity & 1], p->g, axis);
  CUDA_TRY(cudaGetLastError());
  ++p->launches;
  return W3D_OK;
}

int w3d_error_norms(const double* u, const W3dBlock* blk, const double* sx, const double* sy,
                    const double* sz, double ct, double* out, void* workspace, size_t workspace_bytes,
                    void* stream) {
  if (!u || !blk || !sx || !sy || !sz || !out || !workspace) return fail(W3D_EINVAL, "null argument");
  if (workspace_bytes < w3d_error_workspace_bytes())
    return fail(W3D_EINVAL, "workspace too small: need %zu bytes", w3d_error_workspace_bytes());
  Geometry g;
  g.n[0] = blk->n[0]; g.n[1] = blk->n[1]; g.n[2] = blk->n[2];
  g.xoff = blk->xoff;
  g.pitch = blk->pitch;
  g.plane = (long long)(blk->n[1] + 2) * blk->pitch;
  ErrCtl e;
  char* ws = static_cast<char*>(workspace);
  e.out = out;
  e.ticket = reinterpret_cast<unsigned*>(ws);
  e.part_max = reinterpret_cast<double*>(ws + kCounterBytes);
  e.part_sum = e.part_max + W3D_MAX_PARTIALS;
  e.n_bcast = 0;
  const long long rows = (long long)g.n[1] * g.n[2];
  const int blocks = (int)std::max<long long>(1, std::min<long long>((rows + 7) / 8, 148 * 8));
  error_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(u, g, sx, sy, sz, ct, e);
  CUDA_TRY(cudaGetLastError());
  return W3D_OK;
}

}  // extern "C"
