// K3 with the footprints fetched by TMA: radius-4 bilinear lookup into up to 4 pyramid levels, channels-last output.
//
// Same operator as lookup.cu (CorrBlock.__call__ + bilinear_sampler, pytracking/external/RAFT/raft_core/corr.py:29-59,
// raft_core/utils/utils.py:59-73) for the case the fused schedule runs 13 times per frame.  lookup.cu fetches a (2r+2)^2
// footprint as 10 rows of 3-4 float4 loads per row; every such 16-byte access of a scattered row is its own L1 wavefront
// and the kernel sat at 60 % L1/TEX utilisation and 0.33 of the HBM roofline.  Here ONE cp.async.bulk.tensor box per
// (query, level) brings the whole 10-row footprint (16 x 10 floats, 64-byte rows) into shared memory:
//   - level l is a 3-D tensor (W_l, H_l, queries); the box starts at ((floor(x) - 4) & ~3, floor(y) - 4, q) -- the origin of
//     a box must be 16-byte aligned within its row -- and TMA zero-fills what lies outside the map, which is the
//     operator's zero padding;
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
constexpr int kLtR = 4, kLtP = 9, kLtWin = 10, kLtPitch = 16;       // box rows of 16 floats: the box must START on a 16-byte
                                                                     // boundary of its row (a box origin at 12 bytes is an illegal
                                                                     // instruction), so the 10 floats wanted lie at offset 0..3 in it
constexpr int kLtBox = kLtPitch * kLtWin;                            // floats per footprint box (640 bytes = 5 x 128)
constexpr int kLtSlot = kLtBox;                                      // floats between footprints: TMA writes to 128-byte aligned addresses
constexpr int kLtStage = kLtLevels * kLtQueries * kLtSlot;           // floats per stage: 40 KB
constexpr int kLtPairs = kLtLevels * kLtQueries;

struct LookupTmaParams {
  CUtensorMap tm0, tm1, tm2, tm3;  // level l: (W_l, H_l, queries), box (16, 10, 1).  Separate members: a run-time index into
                                   // an array of kernel parameters would make the compiler copy them to local memory,
                                   // where TMA cannot read a descriptor
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

__global__ void __launch_bounds__(kLtThreads, 2) corr_lookup_tma_kernel(const __grid_constant__ LookupTmaParams p) {
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

  // the footprints of one batch of 16 queries: thread t < 64 works out the box origin of (level slot, query) = (t >> 4,
  // t & 15); one elected lane of warp 0 then issues the 64 box loads (the instruction takes its operands from uniform
  // registers, so it is issued from converged code)
  int* s_bx = reinterpret_cast<int*>(bars + 2);                            // [64] box origins of the batch being issued
  int* s_by = s_bx + kLtPairs;
  int* s_bq = s_by + kLtPairs;
  int* s_sh = s_bq + kLtPairs;                                             // [2][64]
  float* s_out = reinterpret_cast<float*>(s_sh + 2 * kLtPairs);            // [16 queries][out_stride] fp32 or fp16
  auto issue = [&](int batch, int stage) {
    __syncthreads();   // every warp is done with this stage (two batches ago) and with the origin scratch
    if (tid < kLtPairs) {
      const int slot = tid >> 4, ql = tid & 15;
      const int q = batch * kLtQueries + ql;
      const int qc = q < p.nq ? q : p.nq - 1;   // the tail batch fetches (and discards) the last query again
      const int b = qc / p.hw1, pix = qc - b * p.hw1;
      const float cx = __ldg(p.coords + (size_t)(2 * b) * p.hw1 + pix);
      const float cy = __ldg(p.coords + (size_t)(2 * b + 1) * p.hw1 + pix);
      const float inv = 1.0f / float(1 << slot);
      const float x = cx * inv, y = cy * inv;  // centroid_lvl = coords / 2**i (corr.py:41)
      float fx = floorf(x), fy = floorf(y);
      s_ax[stage * kLtPairs + tid] = x - fx;
      s_ay[stage * kLtPairs + tid] = y - fy;
      fx = fminf(fmaxf(fx, -1048576.f), 1048576.f);
      fy = fminf(fmaxf(fy, -1048576.f), 1048576.f);
      s_bx[tid] = (int(fx) - kLtR) & ~3;                       // aligned box origin ...
      s_sh[stage * kLtPairs + tid] = (int(fx) - kLtR) & 3;     // ... and where the footprint starts inside the box
      s_by[tid] = int(fy) - kLtR;
      s_bq[tid] = qc;
    }
    __syncthreads();
    // warp w fetches level w: one elected lane issues its 16 boxes (four issuers per CTA)
    if (tid == 0) {
      ptx::fence_proxy_async_smem();   // the stage was last read through the generic proxy, TMA writes it through the async one
      ptx::mbar_arrive_expect_tx(bars + stage, uint32_t(p.num_levels * kLtQueries * kLtBox * 4));
    }
    __syncthreads();   // the expectation is posted before any box can complete
    const int lvl_w = tid >> 5;
    if (lvl_w < p.num_levels && ptx::elect_one()) {
      float* dst = win + stage * kLtStage + lvl_w * kLtQueries * kLtSlot;
      const int o = lvl_w * kLtQueries;
      if (lvl_w == 0) {
        for (int i = 0; i < kLtQueries; ++i) ptx::tma_load_3d(dst + i * kLtSlot, &p.tm0, bars + stage, s_bx[o + i], s_by[o + i], s_bq[o + i]);
      } else if (lvl_w == 1) {
        for (int i = 0; i < kLtQueries; ++i) ptx::tma_load_3d(dst + i * kLtSlot, &p.tm1, bars + stage, s_bx[o + i], s_by[o + i], s_bq[o + i]);
      } else if (lvl_w == 2) {
        for (int i = 0; i < kLtQueries; ++i) ptx::tma_load_3d(dst + i * kLtSlot, &p.tm2, bars + stage, s_bx[o + i], s_by[o + i], s_bq[o + i]);
      } else {
        for (int i = 0; i < kLtQueries; ++i) ptx::tma_load_3d(dst + i * kLtSlot, &p.tm3, bars + stage, s_bx[o + i], s_by[o + i], s_bq[o + i]);
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
      const float* c = wst + pair * kLtSlot + y3 * kLtPitch + xo + s_sh[stage * kLtPairs + pair];
      const float ax = s_ax[stage * kLtPairs + pair], ay = s_ay[stage * kLtPairs + pair];
      const float bx = 1.f - ax, by = 1.f - ay;
      float h[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) h[r] = fmaf(ax, c[r * kLtPitch + 1], bx * c[r * kLtPitch]);
      // the batch's output rows are assembled in shared memory and leave as 16-byte pieces of whole rows: written from
      // here, three values per lane at an 18- or 36-byte stride, every store instruction touched four to five lines
      const int oi = ql * p.out_stride + slot * kLtP * kLtP + xo * kLtP + y3;
      if (p.out_f16) {
        __half* o = reinterpret_cast<__half*>(s_out) + oi;
#pragma unroll
        for (int j = 0; j < 3; ++j) o[j] = __float2half_rn(fmaf(ay, h[j + 1], by * h[j]));
      } else {
        float* o = s_out + oi;
#pragma unroll
        for (int j = 0; j < 3; ++j) o[j] = fmaf(ay, h[j + 1], by * h[j]);
      }
    }
    __syncthreads();
    {
      const int row_bytes = p.out_stride * (p.out_f16 ? 2 : 4);           // a multiple of 16 (host check)
      const int pieces = row_bytes >> 4;
      const int nrows = min(kLtQueries, p.nq - batch * kLtQueries);
      const uint4* src = reinterpret_cast<const uint4*>(s_out);
      uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(p.out) + (size_t)batch * kLtQueries * row_bytes);
      for (int i = tid; i < nrows * pieces; i += kLtThreads) dst[i] = src[i];   // consecutive rows are contiguous
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
    cuuint32_t box[3] = {kLtPitch, kLtWin, 1};   // 16 x 10 floats
    cuuint32_t estr[3] = {1, 1, 1};
    CUtensorMap* tml = l == 0 ? &p.tm0 : l == 1 ? &p.tm1 : l == 2 ? &p.tm2 : &p.tm3;
    if (enc(tml, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(levels[l]), dims, strides, box, estr,
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
  if ((out_stride * (out_f16 ? 2 : 4)) % 16 != 0 || (reinterpret_cast<uintptr_t>(out) & 15) != 0) return WOFT_ERR_UNSUPPORTED;
  const int smem = (2 * kLtStage + 4 * kLtPairs) * 4 + 16 + 5 * kLtPairs * 4 + kLtQueries * out_stride * 4 + 64;
  WOFT_CUDA_CHECK(cudaFuncSetAttribute(corr_lookup_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  int dev = 0, sms = 0;
  WOFT_CUDA_CHECK(cudaGetDevice(&dev));
  WOFT_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = p.batches < 2 * sms ? p.batches : 2 * sms;
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
