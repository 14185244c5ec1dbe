#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void probe(const uint8_t *src, int pitch, int x0, int y0, int tw, int th, uint8_t *out) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t *bar = (uint64_t *)(smem + 16384);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(1) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0)
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(tw * th) : "memory");
  __syncthreads();
  for (int r = threadIdx.x; r < th; r += blockDim.x) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem + r * tw)),
                 "l"(src + (size_t)(y0 + r) * pitch + x0), "r"(tw), "r"(smem_u32(bar)) : "memory");
  }
  uint32_t done = 0;
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(bar)), "r"(0) : "memory");
  }
  for (int i = threadIdx.x; i < tw * th; i += blockDim.x) out[i] = smem[i];
}
int main() {
  int W = 1000, H = 1000, pitch = 1008, tw = 112, th = 81, x0 = 96, y0 = 100;
  std::vector<uint8_t> h((size_t)pitch * H);
  for (int y = 0; y < H; ++y) for (int x = 0; x < pitch; ++x) h[(size_t)y * pitch + x] = (uint8_t)((x * 7 + y * 13) % 251 + 1);
  uint8_t *d, *o; cudaMalloc(&d, h.size()); cudaMalloc(&o, 65536); cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
  probe<<<1, 128, 32768>>>(d, pitch, x0, y0, tw, th, o);
  cudaError_t e = cudaDeviceSynchronize(); printf("kernel: %s\n", cudaGetErrorString(e));
  std::vector<uint8_t> out(tw * th); cudaMemcpy(out.data(), o, out.size(), cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int y = 0; y < th; ++y) for (int x = 0; x < tw; ++x) if (out[y * tw + x] != h[(size_t)(y0 + y) * pitch + x0 + x]) ++bad;
  printf("bulk rows bad %d\n", bad);
  return 0;
}
