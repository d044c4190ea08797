// Two convolutions of the update block whose channel counts are too thin for cuDNN's tensor-core kernels
// (SURVEY.md 8f row 2).  Both run 12 times per flow pass at 1/8 resolution.
//
//   flow_conv7x7_relu_kernel   BasicMotionEncoder.convf1 + ReLU (raft_core/update.py:85, 92): 7x7, 2 -> 128 channels.
//                              cuDNN picks an sm_80 implicit-GEMM kernel that takes ~77 us for 0.36 GFLOP; the
//                              work is 98 multiply-adds per output value, done here in fp32 on CUDA cores with the
//                              50 KB weight tensor in shared memory.  Input NCHW (B,2,H,W) (coords1 - coords0 as it
//                              is), output channels-last (B,H,W,128).
//   flow_out3x3_kernel         FlowHead.conv2 (update.py:10, 14): 3x3, 256 -> 2 channels + bias.  One warp per
//                              pixel, lanes over input channels, warp-shuffle reduction.  Input channels-last
//                              (B,H,W,C), output NCHW (B,2,H,W) so it can be added to the coordinates directly.
#include "common.cuh"

namespace woft {

constexpr int kF7Tile = 8;                       // 8 x 8 output pixels per CTA
constexpr int kF7In = kF7Tile + 6;               // + 3-pixel halo on each side
constexpr int kF7Cout = 128;

__global__ void __launch_bounds__(256) flow_conv7x7_relu_kernel(const float* __restrict__ flow,   // (B,2,H,W)
                                                                const float* __restrict__ weight, // (128,2,7,7)
                                                                const float* __restrict__ bias,   // (128)
                                                                float* __restrict__ out,          // (B,H,W,128)
                                                                int H, int W) {
  extern __shared__ float sm[];
  float* sw = sm;                                // [98 taps][128 cout]
  float* sin = sm + 98 * kF7Cout;                // [2][14][14]
  const int b = blockIdx.z, y0 = blockIdx.y * kF7Tile, x0 = blockIdx.x * kF7Tile;
  for (int i = threadIdx.x; i < 98 * kF7Cout; i += 256) {
    const int co = i & (kF7Cout - 1), tap = i >> 7;   // tap = ci*49 + ky*7 + kx, matching the (ci,ky,kx) weight order
    sw[i] = __ldg(weight + co * 98 + tap);
  }
  for (int i = threadIdx.x; i < 2 * kF7In * kF7In; i += 256) {
    const int ci = i / (kF7In * kF7In), r = i - ci * kF7In * kF7In;
    const int yy = y0 + r / kF7In - 3, xx = x0 + r % kF7In - 3;
    sin[i] = (yy >= 0 && yy < H && xx >= 0 && xx < W) ? __ldg(flow + ((size_t)(b * 2 + ci) * H + yy) * W + xx) : 0.f;
  }
  __syncthreads();
  const int pix = threadIdx.x & 63, grp = threadIdx.x >> 6;   // 64 pixels x 4 groups of 32 output channels
  const int py = pix >> 3, px = pix & 7;
  float acc[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) acc[j] = __ldg(bias + grp * 32 + j);
  for (int ci = 0; ci < 2; ++ci)
    for (int ky = 0; ky < 7; ++ky)
#pragma unroll
      for (int kx = 0; kx < 7; ++kx) {
        const float v = sin[(ci * kF7In + py + ky) * kF7In + px + kx];
        const float4* wv = reinterpret_cast<const float4*>(sw + (ci * 49 + ky * 7 + kx) * kF7Cout + grp * 32);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 w4 = wv[j];
          acc[4 * j] = fmaf(v, w4.x, acc[4 * j]);
          acc[4 * j + 1] = fmaf(v, w4.y, acc[4 * j + 1]);
          acc[4 * j + 2] = fmaf(v, w4.z, acc[4 * j + 2]);
          acc[4 * j + 3] = fmaf(v, w4.w, acc[4 * j + 3]);
        }
      }
  const int y = y0 + py, x = x0 + px;
  if (y < H && x < W) {
    float4* o = reinterpret_cast<float4*>(out + (((size_t)b * H + y) * W + x) * kF7Cout + grp * 32);
#pragma unroll
    for (int j = 0; j < 8; ++j)
      o[j] = make_float4(fmaxf(acc[4 * j], 0.f), fmaxf(acc[4 * j + 1], 0.f), fmaxf(acc[4 * j + 2], 0.f),
                         fmaxf(acc[4 * j + 3], 0.f));
  }
}

// in: (B,H,W,C) channels-last; weight (2,C,3,3); out (B,2,H,W).  One warp per run of 4 pixels along x: lanes split the
// input channels, every input pixel and every weight vector loaded once serves up to 3 of the 4 outputs.
constexpr int kOutRun = 4;

__global__ void __launch_bounds__(256) flow_out3x3_kernel(const float* __restrict__ in, const float* __restrict__ weight,
                                                          const float* __restrict__ bias, float* __restrict__ out,
                                                          int B, int H, int W, int C) {
  extern __shared__ float sw[];                  // [9 taps][2 out][C]
  for (int i = threadIdx.x; i < 18 * C; i += 256) {
    const int c = i % C, o = (i / C) & 1, tap = i / (2 * C);
    sw[i] = __ldg(weight + ((size_t)o * C + c) * 9 + tap);
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int runs_per_row = (W + kOutRun - 1) / kOutRun;
  const long long run = (long long)blockIdx.x * 8 + warp;
  if (run >= (long long)B * H * runs_per_row) return;
  const int xr = int(run % runs_per_row) * kOutRun, y = int((run / runs_per_row) % H);
  const int b = int(run / ((long long)runs_per_row * H));
  float a0[kOutRun], a1[kOutRun];
#pragma unroll
  for (int j = 0; j < kOutRun; ++j) a0[j] = a1[j] = 0.f;
  for (int c4 = lane; c4 < C / 4; c4 += 32) {
    for (int ky = 0; ky < 3; ++ky) {
      const int yy = y + ky - 1;
      if (yy < 0 || yy >= H) continue;
      float4 wa[3], wb[3];
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        wa[kx] = reinterpret_cast<const float4*>(sw + ((ky * 3 + kx) * 2) * C)[c4];
        wb[kx] = reinterpret_cast<const float4*>(sw + ((ky * 3 + kx) * 2 + 1) * C)[c4];
      }
#pragma unroll
      for (int i = 0; i < kOutRun + 2; ++i) {   // input columns xr-1 .. xr+kOutRun
        const int xx = xr + i - 1;
        if (xx < 0 || xx >= W) continue;
        const float4 v = __ldg(reinterpret_cast<const float4*>(in + (((size_t)b * H + yy) * W + xx) * C) + c4);
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          const int j = i - kx;                 // output pixel this (input column, tap) pair feeds
          if (j < 0 || j >= kOutRun) continue;
          a0[j] = fmaf(v.x, wa[kx].x, fmaf(v.y, wa[kx].y, fmaf(v.z, wa[kx].z, fmaf(v.w, wa[kx].w, a0[j]))));
          a1[j] = fmaf(v.x, wb[kx].x, fmaf(v.y, wb[kx].y, fmaf(v.z, wb[kx].z, fmaf(v.w, wb[kx].w, a1[j]))));
        }
      }
    }
  }
#pragma unroll
  for (int j = 0; j < kOutRun; ++j) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      a0[j] += __shfl_xor_sync(0xffffffffu, a0[j], o);
      a1[j] += __shfl_xor_sync(0xffffffffu, a1[j], o);
    }
  }
  if (lane < kOutRun && xr + lane < W) {
    float v0 = a0[0], v1 = a1[0];
#pragma unroll
    for (int j = 1; j < kOutRun; ++j)
      if (lane == j) {
        v0 = a0[j];
        v1 = a1[j];
      }
    out[((size_t)(b * 2) * H + y) * W + xr + lane] = v0 + __ldg(bias);
    out[((size_t)(b * 2 + 1) * H + y) * W + xr + lane] = v1 + __ldg(bias + 1);
  }
}

}  // namespace woft

extern "C" int woft_flow_conv7x7_relu(const float* flow, const float* weight, const float* bias, int batch, int height,
                                      int width, float* out_nhwc, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(flow && weight && bias && out_nhwc, "woft_flow_conv7x7_relu: null pointer");
  WOFT_REQUIRE(batch > 0 && height > 0 && width > 0 && batch < 65536, "woft_flow_conv7x7_relu: bad extent");
  const int smem = (98 * kF7Cout + 2 * kF7In * kF7In) * sizeof(float);
  WOFT_CUDA_CHECK(cudaFuncSetAttribute(flow_conv7x7_relu_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  dim3 grid(ceil_div(width, kF7Tile), ceil_div(height, kF7Tile), batch);
  flow_conv7x7_relu_kernel<<<grid, 256, smem, static_cast<cudaStream_t>(stream)>>>(flow, weight, bias, out_nhwc, height,
                                                                                  width);
  return launch_status("flow_conv7x7_relu_kernel");
}

extern "C" int woft_flow_out3x3(const float* in_nhwc, const float* weight, const float* bias, int batch, int height,
                                int width, int channels, float* out_nchw, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(in_nhwc && weight && bias && out_nchw, "woft_flow_out3x3: null pointer");
  WOFT_REQUIRE(batch > 0 && height > 0 && width > 0, "woft_flow_out3x3: bad extent");
  WOFT_REQUIRE(channels % 4 == 0 && channels <= 1024, "woft_flow_out3x3: channels must be a multiple of 4, <= 1024");
  const long long pixels = (long long)batch * height * ((width + kOutRun - 1) / kOutRun);  // warps = runs of 4 pixels
  const int smem = 18 * channels * sizeof(float);
  if (smem > 48 * 1024)
    WOFT_CUDA_CHECK(cudaFuncSetAttribute(flow_out3x3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  flow_out3x3_kernel<<<(unsigned)((pixels + 7) / 8), 256, smem, static_cast<cudaStream_t>(stream)>>>(
      in_nhwc, weight, bias, out_nchw, batch, height, width, channels);
  return launch_status("flow_out3x3_kernel");
}
