// K1/K2, fp32 CUDA-core path (WOFT_CORR_FP32_SIMT) + the stages every mode shares.
//
// Replaces CorrBlock.corr + the three avg_pool2d passes of CorrBlock.__init__
// (pytracking/external/RAFT/raft_core/corr.py:13-27, 61-69).  This path keeps the reference's arithmetic
// class (fp32 multiply-add, no operand rounding); it exists as the exact-precision mode and as the
// cross-check of the tensor-core path in corr_tc.cu, not as the fast one.
//
//   corr_simt_kernel   128x128 output tile per CTA, 8x8 per thread, operands staged through shared memory.
//                      Both operands are (C, HW) with HW contiguous, so tiles load coalesced as they are.
//   pool2x2_kernel     one level from the previous one (used by this mode only; the tensor-core path pools in
//                      its epilogue).
//   row_mean kernels   mean over (H2, W2) of level 0 for every query, computed as
//                      fmap1[:, q] . mean_j fmap2[:, j] / sqrt(C)  (linearity; no re-read of the volume, which
//                      is what WeightHead.forward does, weighted_raft.py:358-361).
#include <cuda_fp16.h>

#include "common.cuh"

namespace woft {

constexpr int TM = 128, TN = 128, TK = 16;

__global__ void __launch_bounds__(256) corr_simt_kernel(const float* __restrict__ f1, const float* __restrict__ f2,
                                                        float* __restrict__ out, int C, int HW, float denom) {
  __shared__ __align__(16) float As[TK][TM];
  __shared__ __align__(16) float Bs[TK][TN];
  const int b = blockIdx.z;
  const float* A = f1 + (size_t)b * C * HW;
  const float* Bm = f2 + (size_t)b * C * HW;
  float* O = out + (size_t)b * HW * HW;
  const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < C; k0 += TK) {
    // 16 x 128 floats per operand, 256 threads -> 8 elements each, coalesced along HW
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int idx = e * 256 + threadIdx.x;
      const int kk = idx >> 7, mm = idx & 127;
      const int k = k0 + kk;
      As[kk][mm] = (k < C && m0 + mm < HW) ? __ldg(A + (size_t)k * HW + m0 + mm) : 0.f;
      Bs[kk][mm] = (k < C && n0 + mm < HW) ? __ldg(Bm + (size_t)k * HW + n0 + mm) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < TK; ++kk) {
      float a[8], bb[8];
      // rows ty*4..+3 and 64+ty*4..+3; columns tx*4..+3 and 64+tx*4..+3 (conflict-free float4 reads)
      *reinterpret_cast<float4*>(a) = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      *reinterpret_cast<float4*>(a + 4) = *reinterpret_cast<const float4*>(&As[kk][64 + ty * 4]);
      *reinterpret_cast<float4*>(bb) = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      *reinterpret_cast<float4*>(bb + 4) = *reinterpret_cast<const float4*>(&Bs[kk][64 + tx * 4]);
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (m >= HW) continue;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int n = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
      if (n < HW) O[(size_t)m * HW + n] = acc[i][j] / denom;  // corr / sqrt(C), corr.py:69
    }
  }
}

// level l from level l-1: (Q, Hi, Wi) -> (Q, Hi/2, Wi/2), mean of each 2x2 block, trailing row/column dropped
__global__ void pool2x2_kernel(const float* __restrict__ in, float* __restrict__ out, long long total, int Hi,
                               int Wi) {
  const int Ho = Hi / 2, Wo = Wi / 2;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int j = int(idx % Wo);
    const long long t = idx / Wo;
    const int i = int(t % Ho);
    const long long q = t / Ho;
    const float* s = in + (q * Hi + 2 * i) * (long long)Wi + 2 * j;
    out[idx] = (s[0] + s[1] + s[Wi] + s[Wi + 1]) * 0.25f;
  }
}

// mean over HW of every channel of fmap2: one CTA per (b, c)
__global__ void __launch_bounds__(256) channel_mean_kernel(const float* __restrict__ f2, float* __restrict__ mean,
                                                           int HW) {
  __shared__ float part[8];
  const float* src = f2 + (size_t)blockIdx.x * HW;
  float s = 0.f;
  for (int i = threadIdx.x; i < HW; i += 256) s += __ldg(src + i);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += part[w];
    mean[blockIdx.x] = t / float(HW);
  }
}

// row_mean[b, q] = sum_c fmap1[b, c, q] * mean2[b, c] / sqrt(C): 32 queries x 8 channel groups per CTA
__global__ void __launch_bounds__(256) row_mean_kernel(const float* __restrict__ f1, const float* __restrict__ mean2,
                                                       float* __restrict__ row_mean, int C, int HW, float denom) {
  __shared__ float part[8][33];
  const int b = blockIdx.y;
  const int qx = threadIdx.x & 31, cg = threadIdx.x >> 5;
  const int q = blockIdx.x * 32 + qx;
  const float* a = f1 + (size_t)b * C * HW + q;
  const float* m = mean2 + (size_t)b * C;
  float s = 0.f;
  if (q < HW) {
#pragma unroll 4
    for (int c = cg; c < C; c += 8) s = fmaf(__ldg(a + (size_t)c * HW), __ldg(m + c), s);
  }
  part[cg][qx] = s;
  __syncthreads();
  if (cg == 0 && q < HW) {
    float t = 0.f;
#pragma unroll
    for (int g = 0; g < 8; ++g) t += part[g][qx];
    row_mean[(size_t)b * HW + q] = t / denom;
  }
}

// The kernels for channels-last feature maps (B, HW, C).
// mean2[b, c] in two steps: per-strip partial sums (a CTA walks its strip of pixels, threads over channels: every pixel
// is one coalesced row), then one CTA per image adds the strips in a fixed order.
__device__ __forceinline__ float rm_ld(const float* p) { return __ldg(p); }
__device__ __forceinline__ float rm_ld(const __half* p) { return __half2float(*p); }

// T: float, or __half for maps handed over as fp16 (WOFT_CORR_FMAPS_F16); the sums are fp32 either way
template <typename T>
__global__ void __launch_bounds__(256) channel_partial_nhwc_kernel(const T* __restrict__ f2, float* __restrict__ partial,
                                                                   int C, int HW, int strips) {
  const int b = blockIdx.y, strip = blockIdx.x;
  const int len = (HW + strips - 1) / strips;
  const int p0 = strip * len, p1 = min(p0 + len, HW);
  for (int c = threadIdx.x; c < C; c += 256) {
    const T* src = f2 + ((size_t)b * HW) * C + c;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int p = p0;
    for (; p + 3 < p1; p += 4) {
      s0 += rm_ld(src + (size_t)p * C);
      s1 += rm_ld(src + (size_t)(p + 1) * C);
      s2 += rm_ld(src + (size_t)(p + 2) * C);
      s3 += rm_ld(src + (size_t)(p + 3) * C);
    }
    for (; p < p1; ++p) s0 += rm_ld(src + (size_t)p * C);
    partial[((size_t)b * strips + strip) * C + c] = (s0 + s1) + (s2 + s3);
  }
}

__global__ void __launch_bounds__(256) channel_finalize_nhwc_kernel(const float* __restrict__ partial, float* __restrict__ mean,
                                                                    int C, int HW, int strips) {
  const int b = blockIdx.x;
  for (int c = threadIdx.x; c < C; c += 256) {
    float t = 0.f;
    int s = 0;
    for (; s + 7 < strips; s += 8) {   // eight loads in flight, added in strip order (the sum is the same bit for bit)
      float v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = __ldg(partial + ((size_t)b * strips + s + k) * C + c);
#pragma unroll
      for (int k = 0; k < 8; ++k) t += v[k];
    }
    for (; s < strips; ++s) t += partial[((size_t)b * strips + s) * C + c];
    mean[(size_t)b * C + c] = t / float(HW);
  }
}

// row_mean[b, q]: one warp per query, lanes over channels
template <typename T>
__global__ void __launch_bounds__(256) row_mean_nhwc_kernel(const T* __restrict__ f1, const float* __restrict__ mean2,
                                                            float* __restrict__ row_mean, int C, int HW, int B,
                                                            float denom) {
  const long long q = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (q >= (long long)B * HW) return;
  const int b = int(q / HW), lane = threadIdx.x & 31;
  const T* a = f1 + (size_t)q * C;
  const float* m = mean2 + (size_t)b * C;
  float s = 0.f;
  for (int c = lane; c < C; c += 32) s = fmaf(rm_ld(a + c), __ldg(m + c), s);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) row_mean[q] = s / denom;
}

int launch_row_mean(const float* fmap1, const float* fmap2, int B, int C, int HW, float* row_mean, float* mean2,
                    float* nhwc_scratch, bool f16, cudaStream_t stream) {
  if (nhwc_scratch) {   // channels-last maps; scratch holds strips * B * C floats (strips = min(128, HW / 2))
    const int strips = HW / 2 < 128 ? (HW / 2 > 0 ? HW / 2 : 1) : 128;
    const unsigned rows = (unsigned)(((long long)B * HW + 7) / 8);
    if (f16) {
      const __half *h1 = reinterpret_cast<const __half*>(fmap1), *h2 = reinterpret_cast<const __half*>(fmap2);
      channel_partial_nhwc_kernel<__half><<<dim3(strips, B), 256, 0, stream>>>(h2, nhwc_scratch, C, HW, strips);
      channel_finalize_nhwc_kernel<<<B, 256, 0, stream>>>(nhwc_scratch, mean2, C, HW, strips);
      row_mean_nhwc_kernel<__half><<<rows, 256, 0, stream>>>(h1, mean2, row_mean, C, HW, B, sqrtf(float(C)));
    } else {
      channel_partial_nhwc_kernel<float><<<dim3(strips, B), 256, 0, stream>>>(fmap2, nhwc_scratch, C, HW, strips);
      channel_finalize_nhwc_kernel<<<B, 256, 0, stream>>>(nhwc_scratch, mean2, C, HW, strips);
      row_mean_nhwc_kernel<float><<<rows, 256, 0, stream>>>(fmap1, mean2, row_mean, C, HW, B, sqrtf(float(C)));
    }
    return launch_status("row_mean kernels (channels-last)");
  }
  channel_mean_kernel<<<B * C, 256, 0, stream>>>(fmap2, mean2, HW);
  dim3 grid(ceil_div(HW, 32), B);
  row_mean_kernel<<<grid, 256, 0, stream>>>(fmap1, mean2, row_mean, C, HW, sqrtf(float(C)));
  return launch_status("row_mean kernels");
}

int launch_pool_levels(float* const* levels, int num_levels, int first, long long Q, int H, int W,
                       cudaStream_t stream) {
  int h = H, w = W;
  for (int l = 1; l < num_levels; ++l) {
    const int ho = h / 2, wo = w / 2;
    if (l >= first) {
      const long long total = Q * ho * wo;
      const int grid = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
      pool2x2_kernel<<<grid > 0 ? grid : 1, 256, 0, stream>>>(levels[l - 1], levels[l], total, h, w);
    }
    h = ho;
    w = wo;
  }
  return launch_status("pool2x2_kernel");
}

int corr_pyramid_simt(const float* fmap1, const float* fmap2, int B, int C, int H, int W, int num_levels,
                      float* const* levels, cudaStream_t stream) {
  const int HW = H * W;
  dim3 grid(ceil_div(HW, TN), ceil_div(HW, TM), B);
  corr_simt_kernel<<<grid, 256, 0, stream>>>(fmap1, fmap2, levels[0], C, HW, sqrtf(float(C)));
  int rc = launch_status("corr_simt_kernel");
  if (rc != WOFT_OK) return rc;
  return launch_pool_levels(levels, num_levels, 1, (long long)B * HW, H, W, stream);
}

}  // namespace woft
