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
//   flow_head_gather_kernel    18 products t[pixel][tap][o] = <in[pixel,:], w[o,:,tap]> (a 14400 x 256 x 18 product, four
//                              pixels per lane, eight warps splitting the channels), then per output pixel the sum of
//                              the 9 shifted taps + bias, fused with the coordinate update that follows the flow
//                              head in the refinement loop (weighted_raft.py:244-246: coords1 = coords1 + delta_flow).
#include <cuda_fp16.h>

#include "common.cuh"
#include "ptx.cuh"

namespace woft {

// The map is cut into row segments of 7 pixels; a CTA takes up to 16 consecutive segments (sized so that a 160x90 map
// gives 148 CTAs: the work is 181 M multiply-adds, so balance matters more than tile shape).  256 threads = 16 segment
// slots x 16 channel groups; a thread owns one segment x 8 output channels = 56 accumulators.  Per filter row it loads
// the 13 input values under its segment once and slides the 7 taps over them in registers, so one weight vector pair
// from shared memory feeds 56 FMAs and one input value up to 7 x 8.  Weights are staged [tap][cout] with a row stride of
// 132 floats (coalesced global reads, 4-way instead of 32-way conflicts on the transposing writes); a thread's
// channels are {4g..4g+3} and {64+4g..64+4g+3}, which keeps both the weight reads and the output stores contiguous
// across the 16 groups.
constexpr int kF7Seg = 7, kF7Slots = 16;
constexpr int kF7Cout = 128, kF7WStride = 132;

template <bool F16OUT>   // F16OUT: the output tensor holds __half (input of an fp16 tensor-core convolution)
__global__ void __launch_bounds__(256) flow_conv7x7_relu_kernel(const float* __restrict__ flow,   // (B,2,H,W)
                                                                const float* __restrict__ weight, // (128,2,7,7)
                                                                const float* __restrict__ bias,   // (128)
                                                                void* __restrict__ out_any,       // (B,H,W,128)
                                                                int H, int W, int segs_per_cta, int rows_max) {
  extern __shared__ __align__(16) float sm[];
  float* sw = sm;                                // [98 taps][132]
  float* sin = sm + 98 * kF7WStride;             // [2][rows][pitch], 3 zero columns / rows of halo
  const int b = blockIdx.y;
  const int segs_per_row = (W + kF7Seg - 1) / kF7Seg, total = segs_per_row * H;
  const int s0 = blockIdx.x * segs_per_cta, s1 = min(s0 + segs_per_cta, total);
  const int y_lo = s0 / segs_per_row, y_hi = (s1 - 1) / segs_per_row;
  const int rows = y_hi - y_lo + 7, pitch = segs_per_row * kF7Seg + 6, plane = rows_max * pitch;
  {
    // 98 * 128 = 49 * 256 weights: every thread issues its 49 loads before the first (transposing) store, otherwise
    // the staging is 49 L2 round trips long and costs more than the convolution
    float wv[49];
#pragma unroll
    for (int k = 0; k < 49; ++k) wv[k] = __ldg(weight + threadIdx.x + 256 * k);
#pragma unroll
    for (int k = 0; k < 49; ++k) {
      const int i = threadIdx.x + 256 * k;
      const int co = i / 98, tap = i - co * 98;  // tap = ci*49 + ky*7 + kx, the (ci,ky,kx) order of the weight tensor
      sw[tap * kF7WStride + co] = wv[k];
    }
  }
  const int n_in = 2 * rows * pitch;
  pdl_sync();   // the filters above are constants; the flow is the previous kernel's output
  for (int i0 = threadIdx.x; i0 < n_in; i0 += 4 * 256) {
    float v[4];
    int dst[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * 256;
      const int ic = i < n_in ? i : 0;
      const int ci = ic / (rows * pitch), r = ic - ci * rows * pitch;
      const int ry = r / pitch, rx = r - ry * pitch;
      const int yy = y_lo + ry - 3, xx = rx - 3;
      dst[u] = i < n_in ? ci * plane + r : -1;
      v[u] = (yy >= 0 && yy < H && xx >= 0 && xx < W) ? __ldg(flow + ((size_t)(b * 2 + ci) * H + yy) * W + xx) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (dst[u] >= 0) sin[dst[u]] = v[u];
  }
  __syncthreads();
  const int grp = threadIdx.x & 15, slot = threadIdx.x >> 4;   // a warp = 2 segments x 16 groups
  const bool active = slot < segs_per_cta && s0 + slot < s1;
  const int seg = active ? s0 + slot : s0;       // idle slots recompute the first segment and store nothing
  const int y = seg / segs_per_row, x0 = (seg - y * segs_per_row) * kF7Seg;
  unsigned long long acc[kF7Seg][4];   // 8 channels as 4 packed pairs: FFMA2 halves the issue slots of the 5488 FMAs
  {
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(bias) + grp);
    const float4 b1 = __ldg(reinterpret_cast<const float4*>(bias) + 16 + grp);
#pragma unroll
    for (int j = 0; j < kF7Seg; ++j) {
      acc[j][0] = ptx::pack_f32x2(b0.x, b0.y); acc[j][1] = ptx::pack_f32x2(b0.z, b0.w);
      acc[j][2] = ptx::pack_f32x2(b1.x, b1.y); acc[j][3] = ptx::pack_f32x2(b1.z, b1.w);
    }
  }
  for (int ci = 0; ci < 2; ++ci)
#pragma unroll 1
    for (int ky = 0; ky < 7; ++ky) {
      // staged row of input row (y + ky - 3) is (y - y_lo + ky); staged column of input column (x - 3) is x
      const float* srow = sin + ci * plane + (y - y_lo + ky) * pitch + x0;
      float v[kF7Seg + 6];
#pragma unroll
      for (int i = 0; i < kF7Seg + 6; ++i) v[i] = srow[i];
      const float4* wrow = reinterpret_cast<const float4*>(sw + (ci * 49 + ky * 7) * kF7WStride) + grp;
#pragma unroll
      for (int kx = 0; kx < 7; ++kx) {
        const float4 w0 = wrow[kx * (kF7WStride / 4)];
        const float4 w1 = wrow[kx * (kF7WStride / 4) + 16];
        const unsigned long long p0 = ptx::pack_f32x2(w0.x, w0.y), p1 = ptx::pack_f32x2(w0.z, w0.w);
        const unsigned long long p2 = ptx::pack_f32x2(w1.x, w1.y), p3 = ptx::pack_f32x2(w1.z, w1.w);
#pragma unroll
        for (int j = 0; j < kF7Seg; ++j) {
          const float u = v[j + kx];
          ptx::ffma2_bcast(acc[j][0], u, p0);
          ptx::ffma2_bcast(acc[j][1], u, p1);
          ptx::ffma2_bcast(acc[j][2], u, p2);
          ptx::ffma2_bcast(acc[j][3], u, p3);
        }
      }
    }
  if (!active) return;
#pragma unroll
  for (int j = 0; j < kF7Seg; ++j) {
    if (x0 + j < W) {
      float a[8];
#pragma unroll
      for (int k = 0; k < 4; ++k) ptx::unpack_f32x2(acc[j][k], a[2 * k], a[2 * k + 1]);
      const size_t pix = ((size_t)b * H + y) * W + x0 + j;
      if (F16OUT) {
        __half2* o = reinterpret_cast<__half2*>(reinterpret_cast<__half*>(out_any) + pix * kF7Cout);
        uint2 lo, hi;
        __half2 t0 = __floats2half2_rn(fmaxf(a[0], 0.f), fmaxf(a[1], 0.f)), t1 = __floats2half2_rn(fmaxf(a[2], 0.f), fmaxf(a[3], 0.f));
        __half2 t2 = __floats2half2_rn(fmaxf(a[4], 0.f), fmaxf(a[5], 0.f)), t3 = __floats2half2_rn(fmaxf(a[6], 0.f), fmaxf(a[7], 0.f));
        lo.x = *reinterpret_cast<uint32_t*>(&t0); lo.y = *reinterpret_cast<uint32_t*>(&t1);
        hi.x = *reinterpret_cast<uint32_t*>(&t2); hi.y = *reinterpret_cast<uint32_t*>(&t3);
        reinterpret_cast<uint2*>(o)[grp] = lo;        // channels 4g..4g+3
        reinterpret_cast<uint2*>(o)[16 + grp] = hi;   // channels 64+4g..64+4g+3
      } else {
        float4* o = reinterpret_cast<float4*>(reinterpret_cast<float*>(out_any) + pix * kF7Cout);
        o[grp] = make_float4(fmaxf(a[0], 0.f), fmaxf(a[1], 0.f), fmaxf(a[2], 0.f), fmaxf(a[3], 0.f));
        o[16 + grp] = make_float4(fmaxf(a[4], 0.f), fmaxf(a[5], 0.f), fmaxf(a[6], 0.f), fmaxf(a[7], 0.f));
      }
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
constexpr int kTapPix = 128;     // pixels per CTA: a lane owns 4 of them (l, l+32, l+64, l+96)
constexpr int kTapSlices = 8;    // warps per CTA, each taking 1/8 of the input channels
constexpr int kTapW = 20;        // floats per channel in the staged weights: 18 (tap, o) products padded to 5 float4
constexpr int kTapMaxC = 320;    // weights + a 128-pixel tile of C + 4 floats per row must fit in shared memory

// The CTA's 128 input pixels are one contiguous block of global memory (channels-last), brought in by 128 bulk copies
// of one pixel row each into rows of C + 4 floats: with that pitch a lane-per-pixel LDS.128 is conflict free.  Weights
// are read as warp-wide broadcasts; an LDS.128 costs the LSU 4 cycles however few distinct addresses it has, so each
// weight vector has to feed many FMAs: 4 pixels per lane make that 5 loads per 72 FMAs.
__global__ void __launch_bounds__(256) flow_head_taps_kernel(const float* __restrict__ in,      // (pixels, C)
                                                             const float* __restrict__ weight,  // (2, C, 3, 3)
                                                             float* __restrict__ taps,          // (pixels, 9, 2)
                                                             int pixels, int C) {
  extern __shared__ __align__(128) float smt[];
  __shared__ __align__(8) uint64_t bar;
  float* sw = smt;                        // [C][20]
  float* tile = smt + (size_t)C * kTapW;  // [128][C + 4]; reused as [8 slices][128 * 18] partial sums afterwards
  const int pitch = C + 4;
  const int pbase = blockIdx.x * kTapPix;
  const int n_pix = min(kTapPix, pixels - pbase);
  if (threadIdx.x == 0) {
    ptx::mbar_init(&bar, 1);
    ptx::fence_mbar_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) ptx::mbar_arrive_expect_tx(&bar, uint32_t(n_pix) * uint32_t(C) * 4u);
  pdl_sync();        // `in` is the previous kernel's output
  __syncthreads();   // the expectation is posted before any copy can complete
  if (threadIdx.x < n_pix)
    ptx::bulk_load_1d(tile + (size_t)threadIdx.x * pitch, in + (size_t)(pbase + threadIdx.x) * C, uint32_t(C) * 4u, &bar);
  for (int i0 = threadIdx.x; i0 < C * 18; i0 += 4 * 256) {
    float v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * 256;
      const int ic = i < C * 18 ? i : 0;
      const int c = ic / 18, n = ic - c * 18;
      v[u] = __ldg(weight + ((size_t)(n & 1) * C + c) * 9 + (n >> 1));
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * 256;
      if (i < C * 18) sw[(i / 18) * kTapW + i % 18] = v[u];
    }
  }
  __syncthreads();
  for (unsigned spins = 0; !ptx::mbar_try_wait(&bar, 0); ++spins)
    if (spins > (1u << 24)) __trap();   // a copy that never lands must end the kernel, not hang the device
  const int slice = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cs = C / kTapSlices;         // channels per slice, a multiple of 4
  const float* wq = sw + (size_t)slice * cs * kTapW;
  const float* tq = tile + (size_t)lane * pitch + slice * cs;   // rows of pixels beyond n_pix hold stale data: never stored
  unsigned long long acc[4][9];   // (tap, o = 0 / 1) pairs, packed for FFMA2
#pragma unroll
  for (int q = 0; q < 4; ++q)
#pragma unroll
    for (int n = 0; n < 9; ++n) acc[q][n] = 0ull;
  for (int c4 = 0; c4 < (cs >> 2); ++c4) {
    float4 x[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) x[q] = *reinterpret_cast<const float4*>(tq + (size_t)(32 * q) * pitch + c4 * 4);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float4* wv = reinterpret_cast<const float4*>(wq + (c4 * 4 + k) * kTapW);
      unsigned long long w[10];
#pragma unroll
      for (int j = 0; j < 5; ++j) {
        const float4 t = wv[j];
        w[2 * j] = ptx::pack_f32x2(t.x, t.y);
        w[2 * j + 1] = ptx::pack_f32x2(t.z, t.w);
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float v = k == 0 ? x[q].x : (k == 1 ? x[q].y : (k == 2 ? x[q].z : x[q].w));
#pragma unroll
        for (int n = 0; n < 9; ++n) ptx::ffma2_bcast(acc[q][n], v, w[n]);
      }
    }
  }
  __syncthreads();   // every warp is done with the tile: its memory now takes the partial sums
  float* red = tile;
  float* r = red + (size_t)slice * (kTapPix * 18);
#pragma unroll
  for (int q = 0; q < 4; ++q)
#pragma unroll
    for (int n = 0; n < 9; ++n) {
      float lo, hi;
      ptx::unpack_f32x2(acc[q][n], lo, hi);
      *reinterpret_cast<float2*>(r + (lane + 32 * q) * 18 + 2 * n) = make_float2(lo, hi);
    }
  __syncthreads();
  for (int o = threadIdx.x; o < n_pix * 18; o += 256) {
    float t = 0.f;
#pragma unroll
    for (int sl = 0; sl < kTapSlices; ++sl) t += red[sl * (kTapPix * 18) + o];
    taps[(size_t)pbase * 18 + o] = t;
  }
}

// delta[o](y,x) = bias[o] + sum over taps of t[(y+ky-1, x+kx-1)][ky*3+kx][o] (zero padding).  Optional fused outputs:
// coords (B,2,H,W) updated in place, flow = coords - coords0 as NCHW and as (pixels, 2).
// flow_h1 / flow_h2 (optional): fp16 channels-last tensors (pixels, stride) that receive the new flow at channel offset
// `h_off` -- the two operand buffers of the GRU convolutions, where the flow sits behind the motion features
__global__ void flow_head_gather_kernel(const float* __restrict__ taps, const float* __restrict__ bias, int B, int H,
                                        int W, float* __restrict__ delta, float* coords, const float* __restrict__ coords0,
                                        float* __restrict__ flow_nchw, float* __restrict__ flow_nhwc,
                                        __half* __restrict__ flow_h1, __half* __restrict__ flow_h2, int h_stride, int h_off) {
  const int hw = H * W;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  pdl_sync();
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
    const __half2 fh = __floats2half2_rn(fx, fy);
    if (flow_h1) *reinterpret_cast<__half2*>(flow_h1 + (size_t)i * h_stride + h_off) = fh;
    if (flow_h2) *reinterpret_cast<__half2*>(flow_h2 + (size_t)i * h_stride + h_off) = fh;
  }
}

}  // namespace woft

static int flow_conv7x7_impl(const float* flow, const float* weight, const float* bias, int batch, int height, int width,
                             void* out_nhwc, bool f16, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(flow && weight && bias && out_nhwc, "woft_flow_conv7x7_relu: null pointer");
  WOFT_REQUIRE(batch > 0 && height > 0 && width > 0 && batch < 65536, "woft_flow_conv7x7_relu: bad extent");
  // segments of 7 pixels, up to 16 per CTA, a multiple of 148 CTAs per batch when the map is large enough
  const int segs_per_row = ceil_div(width, kF7Seg), total = segs_per_row * height;
  const int waves = ceil_div(total * batch, kF7Slots * 148);
  int spc = ceil_div(total, ceil_div(148 * waves, batch));
  spc = spc < 1 ? 1 : (spc > kF7Slots ? kF7Slots : spc);
  const int rows_max = (spc - 1) / segs_per_row + 2 + 6;   // rows the segments of a CTA can touch, plus the halo
  const size_t smem = (size_t)(98 * kF7WStride + 2 * rows_max * (segs_per_row * kF7Seg + 6)) * sizeof(float);
  WOFT_REQUIRE(smem <= 200 * 1024, "woft_flow_conv7x7_relu: width %d needs more shared memory than a CTA has", width);
  dim3 grid(ceil_div(total, spc), batch);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (f16) {
    WOFT_CUDA_CHECK(cudaFuncSetAttribute(flow_conv7x7_relu_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    WOFT_CUDA_CHECK(launch_pdl(flow_conv7x7_relu_kernel<true>, grid, dim3(256), smem, s, flow, weight, bias, out_nhwc, height,
                               width, spc, rows_max));
  } else {
    WOFT_CUDA_CHECK(cudaFuncSetAttribute(flow_conv7x7_relu_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    WOFT_CUDA_CHECK(launch_pdl(flow_conv7x7_relu_kernel<false>, grid, dim3(256), smem, s, flow, weight, bias, out_nhwc, height,
                               width, spc, rows_max));
  }
  return launch_status("flow_conv7x7_relu_kernel");
}

extern "C" int woft_flow_conv7x7_relu(const float* flow, const float* weight, const float* bias, int batch, int height,
                                      int width, float* out_nhwc, void* stream) {
  return flow_conv7x7_impl(flow, weight, bias, batch, height, width, out_nhwc, false, stream);
}

extern "C" int woft_flow_conv7x7_relu_f16(const float* flow, const float* weight, const float* bias, int batch, int height,
                                          int width, void* out_nhwc_f16, void* stream) {
  return flow_conv7x7_impl(flow, weight, bias, batch, height, width, out_nhwc_f16, true, stream);
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
  WOFT_REQUIRE(channels % 32 == 0 && channels <= kTapMaxC, "woft_flow_head_step: channels must be a multiple of 32, <= 320");
  WOFT_REQUIRE(reinterpret_cast<uintptr_t>(in_nhwc) % 16 == 0, "woft_flow_head_step: input must be 16-byte aligned");
  WOFT_REQUIRE(delta_nchw || coords, "woft_flow_head_step: no output requested");
  WOFT_REQUIRE(!coords || coords0, "woft_flow_head_step: coords needs coords0");
  const long long pixels = (long long)batch * height * width;
  WOFT_REQUIRE(pixels < (1LL << 30), "woft_flow_head_step: too many pixels");
  if (scratch_bytes < woft_flow_head_scratch_bytes(batch, height, width))
    return fail(WOFT_ERR_WORKSPACE, "woft_flow_head_step: scratch too small");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  float* taps = static_cast<float*>(scratch);
  const size_t tile_floats = (size_t)kTapPix * (channels + 4), red_floats = (size_t)kTapSlices * kTapPix * 18;
  const size_t smem = ((size_t)channels * kTapW + (tile_floats > red_floats ? tile_floats : red_floats)) * sizeof(float);
  WOFT_CUDA_CHECK(cudaFuncSetAttribute(flow_head_taps_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  WOFT_CUDA_CHECK(launch_pdl(flow_head_taps_kernel, dim3((unsigned)ceil_div((int)pixels, kTapPix)), dim3(256), smem, s, in_nhwc,
                             weight, taps, (int)pixels, channels));
  WOFT_CUDA_CHECK(launch_pdl(flow_head_gather_kernel, dim3((unsigned)ceil_div((int)pixels, 128)), dim3(128), 0, s,
                             (const float*)taps, bias, batch, height, width, delta_nchw, coords, coords0, flow_nchw, flow_nhwc,
                             (__half*)nullptr, (__half*)nullptr, 0, 0));
  return launch_status("flow head kernels");
}

extern "C" int woft_flow_head_gather(const float* taps, const float* bias, int batch, int height, int width, float* delta_nchw,
                                     float* coords, const float* coords0, float* flow_nchw, float* flow_nhwc,
                                     void* flow_f16_a, void* flow_f16_b, int f16_stride, int f16_offset, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(taps && bias, "woft_flow_head_gather: null pointer");
  WOFT_REQUIRE(batch > 0 && height > 0 && width > 0, "woft_flow_head_gather: bad extent");
  WOFT_REQUIRE(delta_nchw || coords, "woft_flow_head_gather: no output requested");
  WOFT_REQUIRE(!coords || coords0, "woft_flow_head_gather: coords needs coords0");
  WOFT_REQUIRE((!flow_f16_a && !flow_f16_b) || (coords && f16_stride % 2 == 0 && f16_offset % 2 == 0 &&
                                                  ((reinterpret_cast<uintptr_t>(flow_f16_a) | reinterpret_cast<uintptr_t>(flow_f16_b)) & 3) == 0),
               "woft_flow_head_gather: the fp16 flow copies need coords and even strides / offsets");
  const long long pixels = (long long)batch * height * width;
  WOFT_REQUIRE(pixels < (1LL << 30), "woft_flow_head_gather: too many pixels");
  WOFT_CUDA_CHECK(launch_pdl(flow_head_gather_kernel, dim3((unsigned)ceil_div((int)pixels, 128)), dim3(128), 0,
                             static_cast<cudaStream_t>(stream), taps, bias, batch, height, width, delta_nchw, coords, coords0,
                             flow_nchw, flow_nhwc, static_cast<__half*>(flow_f16_a), static_cast<__half*>(flow_f16_b),
                             f16_stride, f16_offset));
  return launch_status("flow_head_gather_kernel");
}
