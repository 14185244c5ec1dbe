#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void probe(const __grid_constant__ CUtensorMap tmap_p, const CUtensorMap *tmap_g, int c0, int c1, int tw, int th, uint8_t *out) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t *bar = (uint64_t *)(smem + 16384);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(1) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(tw * th) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(smem_u32(smem)),
                 "l"(tmap_g ? reinterpret_cast<uint64_t>(tmap_g) : reinterpret_cast<uint64_t>(&tmap_p)), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
  }
  uint32_t done = 0;
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(bar)), "r"(0) : "memory");
  }
  for (int i = threadIdx.x; i < tw * th; i += blockDim.x) out[i] = smem[i];
}
typedef CUresult (*Enc)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char **argv) {
  int W = 1000, H = 1000, pitch = 1008, tw = argc > 1 ? atoi(argv[1]) : 96, th = argc > 2 ? atoi(argv[2]) : 81;
  int variant = argc > 5 ? atoi(argv[5]) : 0;
  int c0 = argc > 3 ? atoi(argv[3]) : 100, c1 = argc > 4 ? atoi(argv[4]) : 100;
  std::vector<uint8_t> h((size_t)pitch * H);
  for (int y = 0; y < H; ++y) for (int x = 0; x < pitch; ++x) h[(size_t)y * pitch + x] = (uint8_t)((x * 7 + y * 13) % 251 + 1);
  uint8_t *d, *o; cudaMalloc(&d, h.size()); cudaMalloc(&o, 65536); cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
  void *p = nullptr; cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  printf("entry %d %d %p\n", (int)e, (int)q, p);
  CUtensorMap tm; cuuint64_t dims[2] = {(cuuint64_t)W, (cuuint64_t)H}; cuuint64_t str[1] = {(cuuint64_t)pitch}; cuuint32_t box[2] = {(cuuint32_t)tw, (cuuint32_t)th}; cuuint32_t es[2] = {1, 1};
  CUresult r;
  if (variant == 1) r = ((Enc)p)(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  else if (variant == 2) { dims[0] = W / 4; box[0] = tw / 4; r = ((Enc)p)(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE); }
  else r = ((Enc)p)(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode %d\n", (int)r);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
  CUtensorMap *dtm = nullptr; if (variant == 3) { cudaMalloc(&dtm, sizeof(tm)); cudaMemcpy(dtm, &tm, sizeof(tm), cudaMemcpyHostToDevice); }
  probe<<<1, 128, 32768>>>(tm, dtm, variant == 2 ? c0 / 4 : c0, c1, tw, th, o);
  e = cudaDeviceSynchronize(); printf("kernel: %s\n", cudaGetErrorString(e));
  std::vector<uint8_t> out(tw * th); cudaMemcpy(out.data(), o, out.size(), cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int y = 0; y < th; ++y) for (int x = 0; x < tw; ++x) {
    int gx = c0 + x, gy = c1 + y; uint8_t want = (gx < 0 || gy < 0 || gx >= W || gy >= H) ? 0 : h[(size_t)gy * pitch + gx];
    if (out[y * tw + x] != want) ++bad;
  }
  printf("tw %d th %d c0 %d c1 %d bad %d\n", tw, th, c0, c1, bad);
  return 0;
}
