// error_noalloc.cu — w3d_error_norms_noalloc(): error norms of one field for
// callers that are not allowed to allocate device memory.
//
// This entry point has no counterpart in the reference (AICCer1/MPI-CUDA); it is
// an addition for library users.
//
// Who needs it: code that runs where cudaMalloc is off limits — a solver embedded
// in a host application whose device memory is handed out once by its own pool,
// a callback inside somebody else's time loop where an allocation would force a
// device-wide synchronisation, or a plug-in that must not disturb a memory budget
// it does not own.  Those callers can pass inputs and an output, but they have
// nowhere to get the scratch that w3d_error_norms() asks for.
//
// How it works: the scratch (one ticket word plus one (max, sum) pair per CTA) is
// a file-scope `static __device__` array in this translation unit.  Every GPU
// that loads libwave3d_b200.so carries W3D_NOALLOC_SCRATCH_BYTES (33 024 bytes)
// of it until the process exits.  Because the array is fixed, the block that can
// be reduced is limited: ny * nz <= W3D_NOALLOC_MAX_ROWS (262 144, e.g. a
// 512 x 512 cross-section; nx is unlimited).  Larger blocks are rejected with
// W3D_ETOOBIG and a message; use w3d_error_norms() with your own workspace then.
//
// Things a caller must know:
//  * The first call on each device looks up the device address of the array
//    (cudaGetSymbolAddress).  Make that first call outside any stream capture.
//  * All calls on one device share the one array.  Each call clears the ticket
//    word on the given stream before its kernel runs and otherwise overwrites what
//    the previous call left behind, so calls issued one after another on the same
//    stream are fine.  Two calls that can overlap on the GPU (different streams,
//    or different host threads without ordering) race on the array and give
//    wrong results — order them yourself.
//  * Nothing here calls cudaMalloc, cudaFree or any allocating library routine.
//  * Errors are reported through the errbuf argument only; w3d_last_error() is not touched.
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "stencil_core.h"
#include "wave3d.h"

namespace {

constexpr int kRowsPerCta = 128;                      // 8 warps x 16 rows
constexpr int kMaxCtas = W3D_NOALLOC_MAX_ROWS / kRowsPerCta;   // 2048
constexpr int kTicketBytes = 256;
static_assert(kTicketBytes + 2 * kMaxCtas * (int)sizeof(double) == W3D_NOALLOC_SCRATCH_BYTES,
              "header constant out of date");

// The scratch itself.  In host code this name is only a handle for
// cudaGetSymbolAddress(); it is not a device address.
static __device__ __align__(256) unsigned char w3d_noalloc_scratch[W3D_NOALLOC_SCRATCH_BYTES];

struct Geo {
  int nx, ny, nz, xoff;
  long long pitch, plane;
};

__device__ __forceinline__ void warp_fold(double& m, double& s) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    s += __shfl_xor_sync(0xffffffffu, s, o);
  }
}

// CTA b reduces rows [b*128, b*128+128) of the (j,k) row list; the last CTA to
// finish (ticket) folds the per-CTA pairs in index order and writes out[0..1].
__global__ void __launch_bounds__(256) error_rows_kernel(const double* __restrict__ u, Geo g,
                                                         const double* __restrict__ sx,
                                                         const double* __restrict__ sy,
                                                         const double* __restrict__ sz, double ct,
                                                         double* out, unsigned char* scratch) {
  __shared__ double sm_max[8], sm_sum[8];
  __shared__ int last;
  unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
  double* part_max = reinterpret_cast<double*>(scratch + kTicketBytes);
  double* part_sum = part_max + kMaxCtas;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long rows = (long long)g.ny * g.nz;
  double m = 0.0, s = 0.0;
  for (int q = 0; q < 16; ++q) {
    const long long r = (long long)blockIdx.x * kRowsPerCta + warp * 16 + q;
    if (r >= rows) break;
    const int j = 1 + (int)(r % g.ny), k = 1 + (int)(r / g.ny);
    const double* row = u + (long long)k * g.plane + (long long)j * g.pitch + g.xoff;
    const double syj = sy[j], szt = sz[k] * ct;
    for (int i = 1 + lane; i <= g.nx; i += 32) {
      const double e = w3d_abs_err(row[i], sx[i] * syj, szt);
      m = fmax(m, e);
      s = fma(e, e, s);
    }
  }
  warp_fold(m, s);
  if (lane == 0) { sm_max[warp] = m; sm_sum[warp] = s; }
  __syncthreads();
  if (warp == 0) {
    m = lane < 8 ? sm_max[lane] : 0.0;
    s = lane < 8 ? sm_sum[lane] : 0.0;
    warp_fold(m, s);
    if (lane == 0) {
      part_max[blockIdx.x] = m;
      part_sum[blockIdx.x] = s;
      __threadfence();
      last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    }
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  m = 0.0; s = 0.0;
  for (unsigned i = threadIdx.x; i < gridDim.x; i += blockDim.x) {
    m = fmax(m, __ldcg(part_max + i));
    s += __ldcg(part_sum + i);
  }
  warp_fold(m, s);
  __syncthreads();
  if (lane == 0) { sm_max[warp] = m; sm_sum[warp] = s; }
  __syncthreads();
  if (warp == 0) {
    m = lane < 8 ? sm_max[lane] : 0.0;
    s = lane < 8 ? sm_sum[lane] : 0.0;
    warp_fold(m, s);
    if (lane == 0) { out[0] = m; out[1] = s; }
  }
}

int report(char* buf, size_t len, int code, const char* fmt, ...) {
  if (buf && len) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, len, fmt, ap);
    va_end(ap);
  }
  return code;
}

constexpr int kMaxDevices = 64;
std::mutex g_setup_mutex;                                   // taken on a device's FIRST call only
std::atomic<unsigned char*> g_scratch_ptr[kMaxDevices];     // device address of the array, per device (null = not resolved yet)

}  // namespace

extern "C" int w3d_error_norms_noalloc(const double* u, const W3dBlock* blk, const double* sx,
                                       const double* sy, const double* sz, double ct, double* out,
                                       void* stream, char* errbuf, size_t errbuf_len) {
  if (errbuf && errbuf_len) errbuf[0] = '\0';
  if (!u || !blk || !sx || !sy || !sz || !out)
    return report(errbuf, errbuf_len, W3D_EINVAL, "w3d_error_norms_noalloc: null argument");
  if (blk->n[0] < 1 || blk->n[1] < 1 || blk->n[2] < 1 || blk->xoff < 0 || blk->pitch < blk->xoff + blk->n[0] + 2)
    return report(errbuf, errbuf_len, W3D_EINVAL,
                  "w3d_error_norms_noalloc: bad block layout (n = %d %d %d, xoff = %d, pitch = %lld; need n >= 1, "
                  "xoff >= 0, pitch >= xoff + nx + 2)", blk->n[0], blk->n[1], blk->n[2], blk->xoff, (long long)blk->pitch);
  const long long rows = (long long)blk->n[1] * blk->n[2];
  if (rows > W3D_NOALLOC_MAX_ROWS)
    return report(errbuf, errbuf_len, W3D_ETOOBIG,
                  "w3d_error_norms_noalloc: block cross-section ny*nz = %lld exceeds the limit of %d rows "
                  "that the built-in %d-byte scratch can hold; use w3d_error_norms() with a workspace of your own",
                  rows, W3D_NOALLOC_MAX_ROWS, W3D_NOALLOC_SCRATCH_BYTES);
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess)
    return report(errbuf, errbuf_len, W3D_ECUDA, "w3d_error_norms_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices)
    return report(errbuf, errbuf_len, W3D_EINVAL,
                  "w3d_error_norms_noalloc: current device has index %d, this entry point keeps the scratch address "
                  "of devices 0..%d only", dev, kMaxDevices - 1);
  // Every call but the first on a device is one atomic load here.  The first one, with that device
  // current, turns the symbol into a device address (under a mutex, so two first calls do not both do it).
  unsigned char* scratch = g_scratch_ptr[dev].load(std::memory_order_acquire);
  if (!scratch) {
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    scratch = g_scratch_ptr[dev].load(std::memory_order_relaxed);
    if (!scratch) {
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, w3d_noalloc_scratch);
      if (e != cudaSuccess || !p)
        return report(errbuf, errbuf_len, W3D_ECUDA, "w3d_error_norms_noalloc: cudaGetSymbolAddress: %s",
                      cudaGetErrorString(e));
      scratch = static_cast<unsigned char*>(p);
      g_scratch_ptr[dev].store(scratch, std::memory_order_release);
    }
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // the ticket must start at zero whatever an earlier (possibly failed) call left there
  e = cudaMemsetAsync(scratch, 0, kTicketBytes, st);
  if (e != cudaSuccess)
    return report(errbuf, errbuf_len, W3D_ECUDA, "w3d_error_norms_noalloc: cudaMemsetAsync: %s",
                  cudaGetErrorString(e));
  Geo g;
  g.nx = blk->n[0]; g.ny = blk->n[1]; g.nz = blk->n[2];
  g.xoff = blk->xoff;
  g.pitch = blk->pitch;
  g.plane = (long long)(blk->n[1] + 2) * blk->pitch;
  const int ctas = (int)((rows + kRowsPerCta - 1) / kRowsPerCta);
  error_rows_kernel<<<ctas, 256, 0, st>>>(u, g, sx, sy, sz, ct, out, scratch);
  e = cudaGetLastError();
  if (e != cudaSuccess)
    return report(errbuf, errbuf_len, W3D_ECUDA, "w3d_error_norms_noalloc: launch: %s", cudaGetErrorString(e));
  return W3D_OK;
}
