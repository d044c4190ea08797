// Two convolutions of the update block whose channel counts are too thin for cuDNN's tensor-core kernels
// (SURVEY.md 8f row 2).  Both run 12 times per flow pass at 1/8 resolution.
//
//   flow_conv7x7_relu_kernel   BasicMotionEncoder.convf1 + ReLU (raft_core/update.py:85, 92): 7x7, 2 -> 128 channels.
//                              cuDNN picks an sm_80 implicit-GEMM kernel that takes ~77 us for 0.36 GFLOP; the
//                              work is 98 multiply-adds per output value, done here in fp32 on CUDA cores with the
//                              50 KB weight tensor in shared memory and 56 accumulators per thread.  Input NCHW
//                              (B,2,H,W) (coords1 - coords0 as it is), output channels-last (B,H,W,128).
//   flow_out3x3_kernel         FlowHead.conv2 (update.py:10, 14): 3x3, 256 -> 2 channels + bias.  One warp per
//                              pixel, lanes over input channels, warp-shuffle reduction.  Input channels-last
//                              (B,H,W,C), output NCHW (B,2,H,W) so it can be added to the coordinates directly.
//   flow_head_taps_kernel +    The same convolution in two steps that read every input value once: per input pixel the
//   flow_head_gather_kernel    18 products t[pixel][tap][o] = <in[pixel,:], w[o,:,tap]> (a 14400 x 256 x 18 product, one
//                              lane per pixel, four warps splitting the channels), then per output pixel the sum of
//                              the 9 shifted taps + bias, fused with the coordinate update that follows the flow
//                              head in the refinement loop (weighted_raft.py:244-246: coords1 = coords1 + delta_flow).
#include "common.cuh"

namespace woft {

// One CTA per run of <= 112 consecutive pixels of one image (the runs are sized so that a 160x90 map gives 147 CTAs for
// 148 SMs: the work is 181 M multiply-adds, so balance matters more than tile shape).  256 threads = 16 pixel slots x
// 16 groups of 8 output channels; a thread owns 7 pixels (slot, slot+16, ...) x 8 channels = 56 accumulators, so one
// weight vector pair read from shared memory feeds 56 FMAs.  The weights are staged [tap][cout] with a row stride of
// 132 floats: global reads coalesced, transposing writes 4-way conflicted instead of 32-way.
constexpr int kF7Slots = 16, kF7PerThread = 7, kF7MaxRun = kF7Slots * kF7PerThread;
constexpr int kF7Cout = 128, kF7WStride = 132;

__global__ void __launch_bounds__(256) flow_conv7x7_relu_kernel(const float* __restrict__ flow,   // (B,2,H,W)
                                                                const float* __restrict__ weight, // (128,2,7,7)
                                                                const float* __restrict__ bias,   // (128)
                                                                float* __restrict__ out,          // (B,H,W,128)
                                                                int H, int W, int run, int rows_max) {
  extern __shared__ __align__(16) float sm[];
  float* sw = sm;                                // [98 taps][132]
  float* sin = sm + 98 * kF7WStride;             // [2][rows][W + 6], 3 zero columns / rows of halo
  const int b = blockIdx.y;
  const int p0 = blockIdx.x * run, p1 = min(p0 + run, H * W);
  const int y_lo = p0 / W, y_hi = (p1 - 1) / W;
  const int rows = y_hi - y_lo + 7, pitch = W + 6, plane = rows_max * pitch;
  for (int i = threadIdx.x; i < 98 * kF7Cout; i += 256) {
    const int co = i / 98, tap = i - co * 98;    // tap = ci*49 + ky*7 + kx, the (ci,ky,kx) order of the weight tensor
    sw[tap * kF7WStride + co] = __ldg(weight + i);
  }
  for (int i = threadIdx.x; i < 2 * rows * pitch; i += 256) {
    const int ci = i / (rows * pitch), r = i - ci * rows * pitch;
    const int ry = r / pitch, rx = r - ry * pitch;
    const int yy = y_lo + ry - 3, xx = rx - 3;
    sin[ci * plane + r] = (yy >= 0 && yy < H && xx >= 0 && xx < W) ? __ldg(flow + ((size_t)(b * 2 + ci) * H + yy) * W + xx)
                                                                  : 0.f;
  }
  __syncthreads();
  const int grp = threadIdx.x & 15, slot = threadIdx.x >> 4;   // a warp = 2 slots x 16 groups: 512-byte output stores
  int base[kF7PerThread];
#pragma unroll
  for (int j = 0; j < kF7PerThread; ++j) {
    const int p = p0 + slot + kF7Slots * j;
    const int pc = p < p1 ? p : p0;              // out-of-run slots compute a valid pixel again and store nothing
    const int y = pc / W, x = pc - y * W;
    base[j] = (y - y_lo) * pitch + x;            // staged row of input (y + ky - 3) is (y - y_lo + ky), column x + kx
  }
  float acc[kF7PerThread][8];
  {
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(bias) + grp * 2);
    const float4 b1 = __ldg(reinterpret_cast<const float4*>(bias) + grp * 2 + 1);
#pragma unroll
    for (int j = 0; j < kF7PerThread; ++j) {
      acc[j][0] = b0.x; acc[j][1] = b0.y; acc[j][2] = b0.z; acc[j][3] = b0.w;
      acc[j][4] = b1.x; acc[j][5] = b1.y; acc[j][6] = b1.z; acc[j][7] = b1.w;
    }
  }
  for (int ci = 0; ci < 2; ++ci)
    for (int ky = 0; ky < 7; ++ky) {
      const float* srow = sin + ci * plane + ky * pitch;
      const float* wrow = sw + (ci * 49 + ky * 7) * kF7WStride + grp * 8;
#pragma unroll
      for (int kx = 0; kx < 7; ++kx) {
        const float4 w0 = *reinterpret_cast<const float4*>(wrow + kx * kF7WStride);
        const float4 w1 = *reinterpret_cast<const float4*>(wrow + kx * kF7WStride + 4);
#pragma unroll
        for (int j = 0; j < kF7PerThread; ++j) {
          const float v = srow[base[j] + kx];
          acc[j][0] = fmaf(v, w0.x, acc[j][0]); acc[j][1] = fmaf(v, w0.y, acc[j][1]);
          acc[j][2] = fmaf(v, w0.z, acc[j][2]); acc[j][3] = fmaf(v, w0.w, acc[j][3]);
          acc[j][4] = fmaf(v, w1.x, acc[j][4]); acc[j][5] = fmaf(v, w1.y, acc[j][5]);
          acc[j][6] = fmaf(v, w1.z, acc[j][6]); acc[j][7] = fmaf(v, w1.w, acc[j][7]);
        }
      }
    }
#pragma unroll
  for (int j = 0; j < kF7PerThread; ++j) {
    const int p = p0 + slot + kF7Slots * j;
    if (p < p1) {
      float4* o = reinterpret_cast<float4*>(out + ((size_t)b * H * W + p) * kF7Cout + grp * 8);
      o[0] = make_float4(fmaxf(acc[j][0], 0.f), fmaxf(acc[j][1], 0.f), fmaxf(acc[j][2], 0.f), fmaxf(acc[j][3], 0.f));
      o[1] = make_float4(fmaxf(acc[j][4], 0.f), fmaxf(acc[j][5], 0.f), fmaxf(acc[j][6], 0.f), fmaxf(acc[j][7], 0.f));
    }
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

// ---- flow head output in two steps -------------------------------------------------------------------------------
constexpr int kTapPix = 128;     // pixels per CTA: 4 groups of 32 (one lane per pixel) x 4 channel quarters = 16 warps
constexpr int kTapW = 20;        // floats per channel in the staged weights: 18 (tap, o) products padded to 5 float4

__global__ void __launch_bounds__(512) flow_head_taps_kernel(const float* __restrict__ in,      // (pixels, C)
                                                             const float* __restrict__ weight,  // (2, C, 3, 3)
                                                             float* __restrict__ taps,          // (pixels, 9, 2)
                                                             int pixels, int C) {
  extern __shared__ __align__(16) float smt[];
  float* sw = smt;                       // [C][20]
  float* red = smt + (size_t)C * kTapW;  // [3][128][19] partial sums of channel quarters 1..3
  for (int i = threadIdx.x; i < C * 18; i += 512) {
    const int c = i / 18, n = i - c * 18;
    sw[c * kTapW + n] = __ldg(weight + ((size_t)(n & 1) * C + c) * 9 + (n >> 1));
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int quarter = warp >> 2, pl = (warp & 3) * 32 + lane;
  const int p = blockIdx.x * kTapPix + pl;
  const int pc = p < pixels ? p : pixels - 1;
  const int cq = C >> 2;
  const float4* src = reinterpret_cast<const float4*>(in + (size_t)pc * C + quarter * cq);
  const float* wq = sw + (size_t)quarter * cq * kTapW;
  float acc[kTapW];
#pragma unroll
  for (int n = 0; n < kTapW; ++n) acc[n] = 0.f;
  for (int c4 = 0; c4 < (cq >> 2); ++c4) {
    const float4 v4 = __ldg(src + c4);
    const float v[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float4* wv = reinterpret_cast<const float4*>(wq + (c4 * 4 + k) * kTapW);
#pragma unroll
      for (int j = 0; j < 5; ++j) {
        const float4 w = wv[j];
        acc[4 * j] = fmaf(v[k], w.x, acc[4 * j]);
        acc[4 * j + 1] = fmaf(v[k], w.y, acc[4 * j + 1]);
        if (j < 4) {   // the fifth vector carries products 16 and 17 only
          acc[4 * j + 2] = fmaf(v[k], w.z, acc[4 * j + 2]);
          acc[4 * j + 3] = fmaf(v[k], w.w, acc[4 * j + 3]);
        }
      }
    }
  }
  if (quarter > 0) {
    float* r = red + ((size_t)(quarter - 1) * kTapPix + pl) * 19;
#pragma unroll
    for (int n = 0; n < 18; ++n) r[n] = acc[n];
  }
  __syncthreads();
  if (quarter == 0 && p < pixels) {
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      const float* r = red + ((size_t)q * kTapPix + pl) * 19;
#pragma unroll
      for (int n = 0; n < 18; ++n) acc[n] += r[n];
    }
    float2* o = reinterpret_cast<float2*>(taps + (size_t)p * 18);
#pragma unroll
    for (int t = 0; t < 9; ++t) o[t] = make_float2(acc[2 * t], acc[2 * t + 1]);
  }
}

// delta[o](y,x) = bias[o] + sum over taps of t[(y+ky-1, x+kx-1)][ky*3+kx][o] (zero padding).  Optional fused outputs:
// coords (B,2,H,W) updated in place, flow = coords - coords0 as NCHW and as (pixels, 2).
__global__ void flow_head_gather_kernel(const float* __restrict__ taps, const float* __restrict__ bias, int B, int H,
                                        int W, float* __restrict__ delta, float* coords, const float* __restrict__ coords0,
                                        float* __restrict__ flow_nchw, float* __restrict__ flow_nhwc) {
  const int hw = H * W;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * hw) return;
  const int b = i / hw, r = i - b * hw;
  const int y = r / W, x = r - y * W;
  float dx = __ldg(bias), dy = __ldg(bias + 1);
#pragma unroll
  for (int ky = 0; ky < 3; ++ky) {
    const int yy = y + ky - 1;
    if (yy < 0 || yy >= H) continue;
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      const int xx = x + kx - 1;
      if (xx < 0 || xx >= W) continue;
      const float2 t = __ldg(reinterpret_cast<const float2*>(taps + ((size_t)b * hw + yy * W + xx) * 18) + ky * 3 + kx);
      dx += t.x;
      dy += t.y;
    }
  }
  const size_t ix = (size_t)(b * 2) * hw + r, iy = ix + hw;
  if (delta) {
    delta[ix] = dx;
    delta[iy] = dy;
  }
  if (coords) {
    const float cx = coords[ix] + dx, cy = coords[iy] + dy;
    coords[ix] = cx;
    coords[iy] = cy;
    const float fx = cx - __ldg(coords0 + ix), fy = cy - __ldg(coords0 + iy);
    if (flow_nchw) {
      flow_nchw[ix] = fx;
      flow_nchw[iy] = fy;
    }
    if (flow_nhwc) reinterpret_cast<float2*>(flow_nhwc)[i] = make_float2(fx, fy);
  }
}

}  // namespace woft

extern "C" int woft_flow_conv7x7_relu(const float* flow, const float* weight, const float* bias, int batch, int height,
                                      int width, float* out_nhwc, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(flow && weight && bias && out_nhwc, "woft_flow_conv7x7_relu: null pointer");
  WOFT_REQUIRE(batch > 0 && height > 0 && width > 0 && batch < 65536, "woft_flow_conv7x7_relu: bad extent");
  // runs of consecutive pixels: a multiple of 148 CTAs per image batch when the map is large enough, <= 112 pixels each
  const int hw = height * width;
  const int waves = ceil_div(hw * batch, kF7MaxRun * 148);
  int run = ceil_div(hw, ceil_div(148 * waves, batch));
  run = run < kF7Slots ? kF7Slots : (run > kF7MaxRun ? kF7MaxRun : run);
  const int rows_max = (run - 1) / width + 2 + 6;   // rows a run can touch, plus the halo
  const size_t smem = (size_t)(98 * kF7WStride + 2 * rows_max * (width + 6)) * sizeof(float);
  WOFT_REQUIRE(smem <= 200 * 1024, "woft_flow_conv7x7_relu: width %d needs more shared memory than a CTA has", width);
  WOFT_CUDA_CHECK(cudaFuncSetAttribute(flow_conv7x7_relu_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid(ceil_div(hw, run), batch);
  flow_conv7x7_relu_kernel<<<grid, 256, smem, static_cast<cudaStream_t>(stream)>>>(flow, weight, bias, out_nhwc, height,
                                                                                  width, run, rows_max);
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

extern "C" size_t woft_flow_head_scratch_bytes(int batch, int height, int width) {
  if (batch <= 0 || height <= 0 || width <= 0) return 0;
  return (size_t)batch * height * width * 18 * sizeof(float);
}

extern "C" int woft_flow_head_step(const float* in_nhwc, const float* weight, const float* bias, int batch, int height,
                                   int width, int channels, void* scratch, size_t scratch_bytes, float* delta_nchw,
                                   float* coords, const float* coords0, float* flow_nchw, float* flow_nhwc,
                                   void* stream) {
  using namespace woft;
  WOFT_REQUIRE(in_nhwc && weight && bias && scratch, "woft_flow_head_step: null pointer");
  WOFT_REQUIRE(batch > 0 && height > 0 && width > 0, "woft_flow_head_step: bad extent");
  WOFT_REQUIRE(channels % 16 == 0 && channels <= 1024, "woft_flow_head_step: channels must be a multiple of 16, <= 1024");
  WOFT_REQUIRE(delta_nchw || coords, "woft_flow_head_step: no output requested");
  WOFT_REQUIRE(!coords || coords0, "woft_flow_head_step: coords needs coords0");
  const long long pixels = (long long)batch * height * width;
  WOFT_REQUIRE(pixels < (1LL << 30), "woft_flow_head_step: too many pixels");
  if (scratch_bytes < woft_flow_head_scratch_bytes(batch, height, width))
    return fail(WOFT_ERR_WORKSPACE, "woft_flow_head_step: scratch too small");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  float* taps = static_cast<float*>(scratch);
  const size_t smem = ((size_t)channels * kTapW + 3 * kTapPix * 19) * sizeof(float);
  WOFT_CUDA_CHECK(cudaFuncSetAttribute(flow_head_taps_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  flow_head_taps_kernel<<<(unsigned)ceil_div((int)pixels, kTapPix), 512, smem, s>>>(in_nhwc, weight, taps, (int)pixels,
                                                                                   channels);
  flow_head_gather_kernel<<<(unsigned)ceil_div((int)pixels, 128), 128, 0, s>>>(taps, bias, batch, height, width, delta_nchw,
                                                                               coords, coords0, flow_nchw, flow_nhwc);
  return launch_status("flow head kernels");
}
