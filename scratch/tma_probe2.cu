#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda/barrier>
#include <cstdio>
#include <cstdint>
#include <vector>
using barrier = cuda::barrier<cuda::thread_scope_block>;
namespace cde = cuda::device::experimental;
__global__ void probe(const __grid_constant__ CUtensorMap tmap, int c0, int c1, int tw, int th, uint8_t *out) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ barrier bar;
  if (threadIdx.x == 0) { init(&bar, blockDim.x); cde::fence_proxy_async_shared_cta(); }
  __syncthreads();
  barrier::arrival_token token;
  if (threadIdx.x == 0) {
    cde::cp_async_bulk_tensor_2d_global_to_shared(smem, &tmap, c0, c1, bar);
    token = cuda::device::barrier_arrive_tx(bar, 1, tw * th);
  } else token = bar.arrive();
  bar.wait(std::move(token));
  for (int i = threadIdx.x; i < tw * th; i += blockDim.x) out[i] = smem[i];
}
typedef CUresult (*Enc)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char **argv) {
  int W = 1000, H = 1000, pitch = 1008, tw = argc > 1 ? atoi(argv[1]) : 96, th = argc > 2 ? atoi(argv[2]) : 81;
  int c0 = argc > 3 ? atoi(argv[3]) : 100, c1 = argc > 4 ? atoi(argv[4]) : 100;
  std::vector<uint8_t> h((size_t)pitch * H);
  for (int y = 0; y < H; ++y) for (int x = 0; x < pitch; ++x) h[(size_t)y * pitch + x] = (uint8_t)((x * 7 + y * 13) % 251 + 1);
  uint8_t *d, *o; cudaMalloc(&d, h.size()); cudaMalloc(&o, 65536); cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
  void *p = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  CUtensorMap tm; cuuint64_t dims[2] = {(cuuint64_t)W, (cuuint64_t)H}; cuuint64_t str[1] = {(cuuint64_t)pitch}; cuuint32_t box[2] = {(cuuint32_t)tw, (cuuint32_t)th}; cuuint32_t es[2] = {1, 1};
  CUresult r = ((Enc)p)(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode %d\n", (int)r);
  probe<<<1, 128, 32768>>>(tm, c0, c1, tw, th, o);
  cudaError_t e = cudaDeviceSynchronize(); printf("kernel: %s\n", cudaGetErrorString(e));
  std::vector<uint8_t> out(tw * th); cudaMemcpy(out.data(), o, out.size(), cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int y = 0; y < th; ++y) for (int x = 0; x < tw; ++x) {
    int gx = c0 + x, gy = c1 + y; uint8_t want = (gx < 0 || gy < 0 || gx >= W || gy >= H) ? 0 : h[(size_t)gy * pitch + gx];
    if (out[y * tw + x] != want) ++bad;
  }
  printf("tw %d th %d c0 %d c1 %d bad %d\n", tw, th, c0, c1, bad);
  return 0;
}
