// tma_probe2.cu — second bring-up probe: official libcu++ TMA wrappers vs raw PTX.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda/barrier>
#include <cstdio>
#include <cstdlib>
#include <vector>
namespace cde = cuda::device::experimental;
using barrier = cuda::barrier<cuda::thread_scope_block>;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("FAIL %s: %s\n", #x, cudaGetErrorString(e_)); exit(1);} } while (0)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

constexpr int W = 32, H = 8;

__global__ void k_bulk(const int* src, int* out) {
  __shared__ alignas(128) int buf[W * H];
#pragma nv_diag_suppress static_var_with_dynamic_init
  __shared__ barrier bar;
  if (threadIdx.x == 0) { init(&bar, blockDim.x); cde::fence_proxy_async_shared_cta(); }
  __syncthreads();
  barrier::arrival_token tok;
  if (threadIdx.x == 0) {
    cuda::memcpy_async(buf, src, cuda::aligned_size_t<16>(sizeof(buf)), bar);
    tok = bar.arrive();
  } else tok = bar.arrive();
  bar.wait(std::move(tok));
  for (int i = threadIdx.x; i < W * H; i += blockDim.x) out[i] = buf[i];
}

__global__ void k_tma2d(const __grid_constant__ CUtensorMap m, int* out, int x, int y) {
  __shared__ alignas(128) int buf[H][W];
#pragma nv_diag_suppress static_var_with_dynamic_init
  __shared__ barrier bar;
  if (threadIdx.x == 0) { init(&bar, blockDim.x); cde::fence_proxy_async_shared_cta(); }
  __syncthreads();
  barrier::arrival_token tok;
  if (threadIdx.x == 0) {
    cde::cp_async_bulk_tensor_2d_global_to_shared(&buf, &m, x, y, bar);
    tok = cuda::device::barrier_arrive_tx(bar, 1, sizeof(buf));
  } else tok = bar.arrive();
  bar.wait(std::move(tok));
  for (int i = threadIdx.x; i < W * H; i += blockDim.x) out[i] = buf[i / W][i % W];
}

int main() {
  const int GW = 256, GH = 64;
  std::vector<int> h(GW * GH);
  for (int i = 0; i < GW * GH; ++i) h[i] = i;
  int *d, *out;
  CK(cudaMalloc(&d, h.size() * 4)); CK(cudaMalloc(&out, W * H * 4));
  CK(cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
  k_bulk<<<1, 128>>>(d, out);
  cudaError_t e = cudaDeviceSynchronize();
  printf("bulk copy (UBLKCP via libcu++): %s\n", e == cudaSuccess ? "ok" : cudaGetErrorString(e));
  if (e != cudaSuccess) return 1;
  void* p = nullptr; cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
  printf("entry point query: %d ptr %p\n", (int)q, p);
  EncodeTiledFn enc = reinterpret_cast<EncodeTiledFn>(p);
  CUtensorMap m;
  const cuuint64_t dims[2] = {GW, GH}; const cuuint64_t strides[1] = {GW * 4};
  const cuuint32_t box[2] = {W, H}, es[2] = {1, 1};
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_INT32, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode: %d\n", (int)r);
  const unsigned long long* w = reinterpret_cast<const unsigned long long*>(&m);
  for (int i = 0; i < 16; ++i) printf("  desc[%2d] = %016llx\n", i, w[i]);
  k_tma2d<<<1, 128>>>(m, out, 8, 4);
  e = cudaDeviceSynchronize();
  printf("tma 2d (libcu++ wrappers): %s\n", e == cudaSuccess ? "ok" : cudaGetErrorString(e));
  if (e != cudaSuccess) return 1;
  std::vector<int> o(W * H);
  CK(cudaMemcpy(o.data(), out, W * H * 4, cudaMemcpyDeviceToHost));
  int bad = 0;
  for (int j = 0; j < H; ++j) for (int i = 0; i < W; ++i) if (o[j * W + i] != (4 + j) * GW + 8 + i) ++bad;
  printf("mismatches %d\n", bad);
  return 0;
}
