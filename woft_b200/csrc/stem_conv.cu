// First layer of both encoders: 7x7, stride 2, pad 3, 3 -> 64 channels (raft_core/extractor.py:131, 171-178) on the
// full-resolution frames.  cuDNN 9 has no sm_100 kernel for three input channels and falls back to an sm_80
// indexed implicit GEMM plus a layout conversion (189 + 20 us for the frame pair, 98 + 12 us for the context encoder at
// 1280x720, profiles/r1_conv_times.txt).  Here it is an implicit GEMM on the tensor cores through mma.sync m16n8k8 with
// TF32 operands (rounded to nearest, fp32 accumulation -- the precision class cuDNN's default TF32 path has):
//   M = output pixels, N = 64, K = 7*7*3 = 147 padded to 152.
// A CTA owns an 8 x 16 tile of output pixels per step and loops over tiles; the weights are staged once per CTA as
// [k][64] (row pitch 72 floats: conflict-free B fragments), the 21 x 37 x 3 input patch of a tile is staged per step
// and the A fragments are gathered from it through a k -> patch-offset table.  4 warps, each 2 m16 tiles x 8 n8 tiles.
// Input (n, H, W, 3) fp32 channels-last (woft_images_to_input), output (n, H/2, W/2, 64) channels-last, optional
// bias + ReLU (the context encoder's folded batch norm).
#include <cuda_fp16.h>

#include "common.cuh"

namespace woft {

constexpr int kStTH = 8, kStTW = 16;                     // output tile
constexpr int kStPH = kStTH * 2 + 5, kStPW = kStTW * 2 + 5;
constexpr int kStK = 147, kStKP = 152, kStN = 64, kStWPitch = 72;
constexpr int kStPatch = kStPH * kStPW * 3;              // 2331 floats

__device__ __forceinline__ float to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

__device__ __forceinline__ void mma_tf32_16x8x8(float (&d)[4], const float (&a)[4], float b0, float b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(__float_as_uint(a[0])), "r"(__float_as_uint(a[1])), "r"(__float_as_uint(a[2])), "r"(__float_as_uint(a[3])),
        "r"(__float_as_uint(b0)), "r"(__float_as_uint(b1)));
}

template <bool F16OUT>   // F16OUT: `out_any` holds __half (input of the fp16 tensor-core convolutions)
__global__ void __launch_bounds__(128) stem_conv7x7_kernel(const float* __restrict__ in,      // (n, H, W, 3)
                                                           const float* __restrict__ weight,  // (64, 3, 7, 7)
                                                           const float* __restrict__ bias,    // (64) or null
                                                           void* __restrict__ out_any,        // (n, Ho, Wo, 64)
                                                           int H, int W, int Ho, int Wo, int tiles_x, int tiles_y,
                                                           int total_tiles, int relu) {
  extern __shared__ __align__(16) float sm[];
  float* sw = sm;                                   // [152][72] tf32 weights, rows 147..151 zero
  float* patch = sw + kStKP * kStWPitch;            // [21][37][3] tf32 input values of the current tile
  int* koff = reinterpret_cast<int*>(patch + ((kStPatch + 3) & ~3));   // k -> offset inside the patch
  const int tid = threadIdx.x;

  for (int i = tid; i < (kStKP - kStK) * kStWPitch; i += 128) sw[kStK * kStWPitch + i] = 0.f;
  for (int i0 = tid; i0 < kStN * kStK; i0 += 8 * 128) {   // 8 loads in flight, then the transposing stores
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = i0 + u * 128;
      v[u] = i < kStN * kStK ? __ldg(weight + i) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = i0 + u * 128;
      if (i < kStN * kStK) {
        const int co = i / kStK, rem = i - co * kStK;        // weight order (co, ci, ky, kx)
        const int ci = rem / 49, kk = rem - ci * 49;         // kk = ky * 7 + kx
        sw[(kk * 3 + ci) * kStWPitch + co] = to_tf32(v[u]);  // k = (ky * 7 + kx) * 3 + ci, the patch's channel order
      }
    }
  }
  for (int k = tid; k < kStKP; k += 128) {
    const int kk = k / 3, ci = k - kk * 3, ky = kk / 7, kx = kk - ky * 7;
    koff[k] = k < kStK ? (ky * kStPW + kx) * 3 + ci : 0;
  }

  const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  // this thread's column of the patch while staging: element j of a 111-float row = (pixel j / 3, channel j % 3)
  const int j = tid, jpx = j / 3;
  for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
    const int img = tile / (tiles_x * tiles_y), r = tile - img * tiles_x * tiles_y;
    const int ty = r / tiles_x, tx = r - ty * tiles_x;
    const int oy0 = ty * kStTH, ox0 = tx * kStTW;
    const int iy0 = oy0 * 2 - 3, ix0 = ox0 * 2 - 3;
    __syncthreads();   // weights staged / previous tile's fragments read before the patch is overwritten
    if (j < kStPW * 3) {
      const int ix = ix0 + jpx;
      const bool col_ok = ix >= 0 && ix < W;
      const float* src = in + ((size_t)img * H * W) * 3 + (ptrdiff_t)ix0 * 3 + j;
      float v[kStPH];
#pragma unroll
      for (int py = 0; py < kStPH; ++py) {
        const int iy = iy0 + py;
        v[py] = (col_ok && iy >= 0 && iy < H) ? __ldg(src + (ptrdiff_t)iy * W * 3) : 0.f;
      }
#pragma unroll
      for (int py = 0; py < kStPH; ++py) patch[py * (kStPW * 3) + j] = to_tf32(v[py]);
    }
    __syncthreads();

    int base[2][2];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int h = 0; h < 2; ++h) base[mt][h] = ((2 * warp + mt) * 2 * kStPW + (g + 8 * h) * 2) * 3;
    float acc[2][8][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 8; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[mt][nt][e] = 0.f;
#pragma unroll 1
    for (int k0 = 0; k0 < kStKP; k0 += 8) {
      const int o0 = koff[k0 + t], o1 = koff[k0 + t + 4];
      float a[2][4];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        a[mt][0] = patch[base[mt][0] + o0];
        a[mt][1] = patch[base[mt][1] + o0];
        a[mt][2] = patch[base[mt][0] + o1];
        a[mt][3] = patch[base[mt][1] + o1];
      }
      const float* w0 = sw + (k0 + t) * kStWPitch + g;
      const float* w1 = sw + (k0 + t + 4) * kStWPitch + g;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        const float b0 = w0[nt * 8], b1 = w1[nt * 8];
        mma_tf32_16x8x8(acc[0][nt], a[0], b0, b1);
        mma_tf32_16x8x8(acc[1][nt], a[1], b0, b1);
      }
    }
    // accumulator (row g | g + 8, columns 2t, 2t + 1) of n-tile nt -> pixel (oy, ox0 + g | g + 8), channels nt*8 + 2t, +1
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
      const int oy = oy0 + 2 * warp + mt;
      if (oy >= Ho) continue;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int ox = ox0 + g + 8 * h;
        if (ox >= Wo) continue;
        const size_t obase = (((size_t)img * Ho + oy) * Wo + ox) * kStN + 2 * t;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          float x = acc[mt][nt][2 * h], y = acc[mt][nt][2 * h + 1];
          if (bias) {
            x += __ldg(bias + nt * 8 + 2 * t);
            y += __ldg(bias + nt * 8 + 2 * t + 1);
          }
          if (relu) {
            x = fmaxf(x, 0.f);
            y = fmaxf(y, 0.f);
          }
          if (F16OUT)
            *reinterpret_cast<__half2*>(reinterpret_cast<__half*>(out_any) + obase + nt * 8) = __floats2half2_rn(x, y);
          else
            *reinterpret_cast<float2*>(reinterpret_cast<float*>(out_any) + obase + nt * 8) = make_float2(x, y);
        }
      }
    }
  }
}

}  // namespace woft

static int stem_conv7x7_impl(const float* in_nhwc, const float* weight, const float* bias, int batch, int height, int width,
                             int relu, void* out_nhwc, bool f16, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(in_nhwc && weight && out_nhwc, "woft_stem_conv7x7: null pointer");
  WOFT_REQUIRE(batch > 0 && height > 0 && width > 0, "woft_stem_conv7x7: bad extent");
  WOFT_REQUIRE((long long)batch * height * width * 3 < (1LL << 31), "woft_stem_conv7x7: input too large");
  const int Ho = (height + 2 * 3 - 7) / 2 + 1, Wo = (width + 2 * 3 - 7) / 2 + 1;
  const int tiles_x = ceil_div(Wo, kStTW), tiles_y = ceil_div(Ho, kStTH);
  const long long total = (long long)tiles_x * tiles_y * batch;
  WOFT_REQUIRE(total < (1LL << 31), "woft_stem_conv7x7: too many tiles");
  const size_t smem = ((size_t)kStKP * kStWPitch + ((kStPatch + 3) & ~3)) * sizeof(float) + kStKP * sizeof(int);
  const int grid = (int)(total < 148 * 4 ? total : 148 * 4);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (f16) {
    WOFT_CUDA_CHECK(cudaFuncSetAttribute(stem_conv7x7_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    stem_conv7x7_kernel<true><<<grid, 128, smem, s>>>(in_nhwc, weight, bias, out_nhwc, height, width, Ho, Wo, tiles_x, tiles_y,
                                                      (int)total, relu);
  } else {
    WOFT_CUDA_CHECK(cudaFuncSetAttribute(stem_conv7x7_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    stem_conv7x7_kernel<false><<<grid, 128, smem, s>>>(in_nhwc, weight, bias, out_nhwc, height, width, Ho, Wo, tiles_x, tiles_y,
                                                       (int)total, relu);
  }
  return launch_status("stem_conv7x7_kernel");
}

extern "C" int woft_stem_conv7x7(const float* in_nhwc, const float* weight, const float* bias, int batch, int height,
                                 int width, int relu, float* out_nhwc, void* stream) {
  return stem_conv7x7_impl(in_nhwc, weight, bias, batch, height, width, relu, out_nhwc, false, stream);
}

extern "C" int woft_stem_conv7x7_f16(const float* in_nhwc, const float* weight, const float* bias, int batch, int height,
                                     int width, int relu, void* out_nhwc_f16, void* stream) {
  return stem_conv7x7_impl(in_nhwc, weight, bias, batch, height, width, relu, out_nhwc_f16, true, stream);
}
