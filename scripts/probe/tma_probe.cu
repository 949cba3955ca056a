// Stand-alone probe: one fp32, unswizzled 3-D TMA box with negative start coordinates.  nvcc -arch=sm_100a tma_probe.cu -o tma_probe
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
struct Maps { CUtensorMap m[4]; };
__global__ void probe(const __grid_constant__ Maps maps, float* out, int bw, int x0, int y0, int z0, int which) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  const uint32_t b = (uint32_t)__cvta_generic_to_shared(&bar), d = (uint32_t)__cvta_generic_to_shared(smem);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const uint32_t bytes = bw * 4 * 3 * 4;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(d),
                 "l"(reinterpret_cast<uint64_t>(&maps.m[which])), "r"(b), "r"(x0), "r"(y0), "r"(z0)
                 : "memory");
  }
  uint32_t done = 0;
  while (!done)
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(b) : "memory");
  for (int i = threadIdx.x; i < bw * 12; i += blockDim.x) out[i] = reinterpret_cast<float*>(smem)[i];
}
__global__ void probe_plain(const __grid_constant__ CUtensorMap tm, float* out, int bw, int x0, int y0, int z0, int rank) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  const uint32_t b = (uint32_t)__cvta_generic_to_shared(&bar), d = (uint32_t)__cvta_generic_to_shared(smem);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const uint32_t bytes = rank == 3 ? bw * 4 * 3 * 4 : bw * 4 * 4;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
    if (rank == 3)
      asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(d),
                   "l"(reinterpret_cast<uint64_t>(&tm)), "r"(b), "r"(x0), "r"(y0), "r"(z0)
                   : "memory");
    else
      asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(d),
                   "l"(reinterpret_cast<uint64_t>(&tm)), "r"(b), "r"(x0), "r"(y0)
                   : "memory");
  }
  uint32_t done = 0;
  while (!done)
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(b) : "memory");
  for (int i = threadIdx.x; i < (int)bytes / 4; i += blockDim.x) out[i] = reinterpret_cast<float*>(smem)[i];
}
int main(int argc, char** argv) {
  const int bw = argc > 1 ? atoi(argv[1]) : 132, which = argc > 2 ? atoi(argv[2]) : 0;
  const int mode = argc > 3 ? atoi(argv[3]) : 0;  // 0: struct param, 3-D   1: plain param, 3-D   2: plain param, 2-D   3: plain 2-D, non-negative coords
  const int W = 256, H = 256, N = 2;
  std::vector<float> h((size_t)N * 3 * H * W);
  for (size_t i = 0; i < h.size(); ++i) h[i] = (float)(i % 100003);
  float *dx, *dout;
  cudaMalloc(&dx, h.size() * 4);
  cudaMalloc(&dout, bw * 12 * 4);
  cudaMemcpy(dx, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  EncodeTiledFn enc = (EncodeTiledFn)p;
  Maps maps;
  cuuint64_t gd[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N * 3}, gs[2] = {(cuuint64_t)W * 4, (cuuint64_t)W * H * 4};
  cuuint32_t bx[3] = {(cuuint32_t)bw, 4, 3}, es[3] = {1, 1, 1};
  for (int i = 0; i < 4; ++i) {
    CUresult r = enc(&maps.m[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, dx, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
  }
  int x0 = -1, y0 = -1, z0 = 3;
  if (mode == 3) x0 = 0, y0 = 4;
  if (mode >= 2) {
    cuuint64_t gd2[2] = {(cuuint64_t)W, (cuuint64_t)H * N * 3}, gs2[1] = {(cuuint64_t)W * 4};
    cuuint32_t bx2[2] = {(cuuint32_t)bw, 4}, es2[2] = {1, 1};
    CUresult r = enc(&maps.m[0], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, dx, gd2, gs2, bx2, es2, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode2 failed %d\n", (int)r); return 1; }
  }
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(probe_plain, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  if (mode == 0) probe<<<1, 128, 65536>>>(maps, dout, bw, x0, y0, z0, which);
  else probe_plain<<<1, 128, 65536>>>(maps.m[0], dout, bw, x0, y0, z0, mode == 1 ? 3 : 2);
  cudaError_t e = cudaDeviceSynchronize();
  printf("bw=%d which=%d mode=%d: %s\n", bw, which, mode, cudaGetErrorString(e));
  if (mode >= 2) return 0;
  if (e != cudaSuccess) return 2;
  std::vector<float> o(bw * 12);
  cudaMemcpy(o.data(), dout, o.size() * 4, cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 4; ++r)
      for (int j = 0; j < bw; ++j) {
        const int x = x0 + j, y = y0 + r, z = z0 + c;
        const float want = (x < 0 || x >= W || y < 0 || y >= H) ? 0.f : h[((size_t)z * H + y) * W + x];
        if (o[(c * 4 + r) * bw + j] != want) ++bad;
      }
  printf("mismatches: %d\n", bad);
  return bad != 0;
}
