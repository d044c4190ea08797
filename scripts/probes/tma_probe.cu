// Scratch probe (not part of the library): which origins a tiled TMA box may have -- an origin whose innermost coordinate is
// not a multiple of 16 bytes is an illegal instruction; built by hand with nvcc -gencode arch=compute_100a,code=sm_100a.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include "../../woft_b200/csrc/ptx.cuh"
using namespace woft;
struct P { CUtensorMap tm; float* out; int bx, by, bq, boxw, boxh; };
__global__ void k(const __grid_constant__ P p) {
  extern __shared__ __align__(128) float sm[];
  __shared__ uint64_t bar;
  if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); }
  __syncthreads();
  if (threadIdx.x == 0) {
    ptx::mbar_arrive_expect_tx(&bar, p.boxw * p.boxh * 4);
    ptx::tma_load_3d(sm, &p.tm, &bar, p.bx, p.by, p.bq);
  }
  while (!ptx::mbar_try_wait(&bar, 0)) {}
  for (int i = threadIdx.x; i < p.boxw * p.boxh; i += blockDim.x) p.out[i] = sm[i];
}
typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main() {
  void* fn; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  Enc enc = (Enc)fn;
  const int W = 160, H = 90, NQ = 64;
  float* d; cudaMalloc(&d, (size_t)W * H * NQ * 4);
  float* h = (float*)malloc((size_t)W * H * NQ * 4);
  for (size_t i = 0; i < (size_t)W * H * NQ; ++i) h[i] = float(i % 1000003);
  cudaMemcpy(d, h, (size_t)W * H * NQ * 4, cudaMemcpyHostToDevice);
  float* out; cudaMalloc(&out, 4096 * 4);
  for (int boxw : {16}) {
    for (int bx : {0, 4, -4, 152, -400, 4000}) {
      P p{};
      cuuint64_t dims[3] = {W, H, NQ};
      cuuint64_t strides[2] = {W * 4, (cuuint64_t)W * H * 4};
      cuuint32_t box[3] = {(cuuint32_t)boxw, 10, 1};
      cuuint32_t es[3] = {1, 1, 1};
      CUresult r = enc(&p.tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      p.out = out; p.bx = bx; p.by = 5; p.bq = 7; p.boxw = boxw; p.boxh = 10;
      k<<<1, 64, 16384>>>(p);
      cudaError_t e = cudaDeviceSynchronize();
      float ho[512]; cudaMemcpy(ho, out, boxw * 10 * 4, cudaMemcpyDeviceToHost);
      float want = bx >= 0 ? h[((size_t)7 * H + 5) * W + bx] : 0.f;
      printf("boxw %d bx %d: encode %d, run %s, out[0]=%g want %g, out[boxw]=%g want %g\n", boxw, bx, (int)r, cudaGetErrorString(e), ho[0], want, ho[boxw], h[((size_t)7 * H + 6) * W + (bx >= 0 ? bx : 0)]);
      if (e != cudaSuccess) return 1;
    }
  }
  return 0;
}
