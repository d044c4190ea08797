// K3 with the footprints fetched by TMA: radius-4 bilinear lookup into up to 4 pyramid levels, channels-last output.
//
// Same operator as lookup.cu (CorrBlock.__call__ + bilinear_sampler, pytracking/external/RAFT/raft_core/corr.py:29-59,
// raft_core/utils/utils.py:59-73) for the case the fused schedule runs 13 times per frame.  lookup.cu fetches a (2r+2)^2
// footprint as 10 rows of 3-4 float4 loads per row; every such 16-byte access of a scattered row is its own L1 wavefront
// and the kernel sat at 60 % L1/TEX utilisation and 0.33 of the HBM roofline.  Here ONE cp.async.bulk.tensor box per
// (query, level) brings the whole 10-row footprint (12 x 10 floats, 48-byte rows) into shared memory:
//   - level l is a 3-D tensor (W_l, H_l, queries); the box starts at (floor(x) - 4, floor(y) - 4, q) -- any element
//     alignment -- and TMA zero-fills what lies outside the map, which is the operator's zero padding;
//   - the load/store units only see the coalesced output stores; the 57 600 boxes of a 1280x720 lookup are address
//     generation work for the TMA units (10 rows each);
//   - persistent CTAs with two 16-query stages: the boxes of batch i+1 are in flight while batch i is interpolated.
// Interpolation and channel order are lookup.cu's channels-last path (a warp per footprint, lane = three consecutive
// channels of one column), fp32 or fp16 output.
// Algorithmic bytes per query pixel: SURVEY.md 8(d)'s 2904 B (4*10*10*4 read + 8 coords + 324*4 written).
#include <cuda.h>
#include <cuda_fp16.h>

#include "common.cuh"
#include "ptx.cuh"

namespace woft {

constexpr int kLtThreads = 128, kLtQueries = 16, kLtLevels = 4;
constexpr int kLtR = 4, kLtP = 9, kLtWin = 10, kLtPitch = 12;       // 12 floats = 48-byte box rows (16-byte multiple)
constexpr int kLtBox = kLtPitch * kLtWin;                            // floats per footprint
constexpr int kLtStage = kLtLevels * kLtQueries * kLtBox;            // floats per stage: 30 KB
constexpr int kLtPairs = kLtLevels * kLtQueries;

struct LookupTmaParams {
  CUtensorMap tm[kLtLevels];  // (W_l, H_l, queries), box (12, 10, 1)
  int num_levels;
  const float* coords;  // (B, 2, H1, W1)
  void* out;            // (queries, out_stride) fp32 or fp16
  int nq, hw1, out_f16, out_stride, batches;
  int* abort_flag;
};

__device__ int g_lookup_abort = 0;

__device__ __forceinline__ bool lt_wait(uint64_t* bar, uint32_t parity) {
  if (ptx::mbar_try_wait(bar, parity)) return true;
  const uint64_t t0 = ptx::globaltimer_ns();
  uint32_t spins = 0;
  while (!ptx::mbar_try_wait(bar, parity)) {
    if ((++spins & 63u) == 0) {
      if (*reinterpret_cast<volatile int*>(&g_lookup_abort) != 0) return false;
      if (ptx::globaltimer_ns() - t0 > 2000000000ull) {
        atomicExch(&g_lookup_abort, 1);
        return false;
      }
    }
  }
  return true;
}

__global__ void __launch_bounds__(kLtThreads, 3) corr_lookup_tma_kernel(const __grid_constant__ LookupTmaParams p) {
  extern __shared__ __align__(128) float lt_smem[];
  float* win = lt_smem;                                                    // [2 stages][level][query][10][12]
  float* s_ax = win + 2 * kLtStage;                                        // [2][64]
  float* s_ay = s_ax + 2 * kLtPairs;
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_ay + 2 * kLtPairs);       // [2]
  const int tid = threadIdx.x;
  if (tid == 0) {
    ptx::mbar_init(bars, 1);
    ptx::mbar_init(bars + 1, 1);
    ptx::fence_mbar_init();
  }
  __syncthreads();
  pdl_sync();   // the coordinates come from the previous kernel of the iteration

  // the footprints of one batch of 16 queries: thread t < 64 handles (level slot, query) = (t >> 4, t & 15)
  auto issue = [&](int batch, int stage) {
    if (tid == 0) ptx::mbar_arrive_expect_tx(bars + stage, uint32_t(p.num_levels * kLtQueries * kLtBox * 4));
    __syncthreads();   // the expectation is posted before any box can complete; also orders the reuse of the stage
    ptx::fence_proxy_async_smem();   // the stage was last read through the generic proxy, TMA writes it through the async one
    if (tid < kLtPairs) {
      const int slot = tid >> 4, ql = tid & 15;
      const int q = batch * kLtQueries + ql;
      if (slot < p.num_levels) {
        float cx = 0.f, cy = 0.f;
        const int qc = q < p.nq ? q : p.nq - 1;   // the tail batch fetches (and discards) the last query again
        {
          const int b = qc / p.hw1, pix = qc - b * p.hw1;
          cx = __ldg(p.coords + (size_t)(2 * b) * p.hw1 + pix);
          cy = __ldg(p.coords + (size_t)(2 * b + 1) * p.hw1 + pix);
        }
        const float inv = 1.0f / float(1 << slot);
        const float x = cx * inv, y = cy * inv;  // centroid_lvl = coords / 2**i (corr.py:41)
        float fx = floorf(x), fy = floorf(y);
        s_ax[stage * kLtPairs + tid] = x - fx;
        s_ay[stage * kLtPairs + tid] = y - fy;
        fx = fminf(fmaxf(fx, -1048576.f), 1048576.f);
        fy = fminf(fmaxf(fy, -1048576.f), 1048576.f);
        float* dst = win + stage * kLtStage + tid * kLtBox;
        const CUtensorMap* tm = &p.tm[slot];
        ptx::tma_load_3d(dst, tm, bars + stage, int(fx) - kLtR, int(fy) - kLtR, qc);
      }
    }
  };

  const int warp = tid >> 5, lane = tid & 31;
  const int xo = lane / 3, y3 = (lane - xo * 3) * 3;   // lane < 27: column xo, rows y3 .. y3 + 2 of the 9 x 9 window
  const bool active = lane < 27;
  uint32_t phase[2] = {0u, 0u};
  int it = 0;
  int batch = blockIdx.x;
  if (batch < p.batches) issue(batch, 0);
  for (; batch < p.batches; batch += gridDim.x, ++it) {
    const int stage = it & 1;
    const int next = batch + gridDim.x;
    if (next < p.batches) issue(next, stage ^ 1);   // its __syncthreads also closes the previous use of that stage
    if (!lt_wait(bars + stage, phase[stage])) break;
    phase[stage] ^= 1u;
    const float* wst = win + stage * kLtStage;
    for (int pair = warp; pair < p.num_levels * kLtQueries; pair += kLtThreads / 32) {
      const int slot = pair >> 4, ql = pair & 15;
      const int q = batch * kLtQueries + ql;
      if (q >= p.nq || !active) continue;
      const float* c = wst + pair * kLtBox + y3 * kLtPitch + xo;
      const float ax = s_ax[stage * kLtPairs + pair], ay = s_ay[stage * kLtPairs + pair];
      const float bx = 1.f - ax, by = 1.f - ay;
      float h[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) h[r] = fmaf(ax, c[r * kLtPitch + 1], bx * c[r * kLtPitch]);
      const size_t oi = (size_t)q * p.out_stride + (size_t)slot * kLtP * kLtP + xo * kLtP + y3;
      if (p.out_f16) {
        __half* o = reinterpret_cast<__half*>(p.out) + oi;
#pragma unroll
        for (int j = 0; j < 3; ++j) o[j] = __float2half_rn(fmaf(ay, h[j + 1], by * h[j]));
      } else {
        float* o = reinterpret_cast<float*>(p.out) + oi;
#pragma unroll
        for (int j = 0; j < 3; ++j) o[j] = fmaf(ay, h[j + 1], by * h[j]);
      }
    }
  }
}

typedef CUresult (*LtEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// Returns WOFT_OK after launching, or WOFT_ERR_UNSUPPORTED when this shape must take lookup.cu's path (levels whose row
// pitch is not a multiple of 16 bytes, radius other than 4, more than 4 levels, channel-first output).
int corr_lookup_tma(const float* const* levels, int num_levels, const float* coords, int batch, int h1, int w1, int h2, int w2,
                    void* out, int out_f16, int out_stride, cudaStream_t stream) {
  if (num_levels > kLtLevels) return WOFT_ERR_UNSUPPORTED;
  static LtEncodeFn enc = [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      fn = nullptr;
    return reinterpret_cast<LtEncodeFn>(fn);
  }();
  if (!enc) return WOFT_ERR_UNSUPPORTED;
  LookupTmaParams p{};
  const long long nq = (long long)batch * h1 * w1;
  int h = h2, w = w2;
  for (int l = 0; l < num_levels; ++l) {
    if ((w * 4) % 16 != 0 || (reinterpret_cast<uintptr_t>(levels[l]) & 15) != 0) return WOFT_ERR_UNSUPPORTED;
    cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)nq};
    cuuint64_t strides[2] = {(cuuint64_t)w * 4, (cuuint64_t)w * h * 4};
    cuuint32_t box[3] = {kLtPitch, kLtWin, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    if (enc(&p.tm[l], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(levels[l]), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return WOFT_ERR_UNSUPPORTED;
    h /= 2;
    w /= 2;
  }
  p.num_levels = num_levels;
  p.coords = coords;
  p.out = out;
  p.nq = (int)nq;
  p.hw1 = h1 * w1;
  p.out_f16 = out_f16;
  p.out_stride = out_stride;
  p.batches = (int)((nq + kLtQueries - 1) / kLtQueries);
  const int smem = (2 * kLtStage + 4 * kLtPairs) * 4 + 64;
  WOFT_CUDA_CHECK(cudaFuncSetAttribute(corr_lookup_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  int dev = 0, sms = 0;
  WOFT_CUDA_CHECK(cudaGetDevice(&dev));
  WOFT_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = p.batches < 3 * sms ? p.batches : 3 * sms;
  WOFT_CUDA_CHECK(launch_pdl(corr_lookup_tma_kernel, dim3(grid), dim3(kLtThreads), (size_t)smem, stream, p));
  return launch_status("corr_lookup_tma_kernel");
}

}  // namespace woft

extern "C" int woft_corr_lookup_check(void* stream) {
  using namespace woft;
  int flag = 0;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  WOFT_CUDA_CHECK(cudaMemcpyFromSymbolAsync(&flag, g_lookup_abort, sizeof(int), 0, cudaMemcpyDeviceToHost, s));
  WOFT_CUDA_CHECK(cudaStreamSynchronize(s));
  if (flag != 0) return fail(WOFT_ERR_CUDA, "lookup kernel aborted: a footprint load timed out (flag=%d)", flag);
  return WOFT_OK;
}
