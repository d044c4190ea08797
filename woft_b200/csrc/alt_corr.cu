// A'': on-the-fly local correlation, the operator behind the reference's one native extension
// (alt_cuda_corr.forward, pytracking/external/RAFT/alt_cuda_corr/correlation_kernel.cu:18-119, 260-286,
// called from AlternateCorrBlock.__call__, raft_core/corr.py:83-100).  No volume is materialised: for every
// query pixel the (2r+2)^2 dot products around floor(coords) are formed from channels-last feature maps and
// combined bilinearly into the (2r+1)^2 outputs.
//
// The reference kernel gives one warp a 4x8 pixel tile and walks 32 channels at a time through shared memory.
// Here one warp owns one query: fmap1's C-vector stays in registers, every fmap2 position is one coalesced
// C*4-byte row (fmap2 is L2-resident: 14.7 MB at 1280x720), dot products are warp-reduced, then the lanes
// form the outputs from the (2r+2)^2 table in shared memory.  Output channel = (yoff+r) + (2r+1)*(xoff+r)
// (correlation_kernel.cu:92-95), values unscaled (the caller divides by sqrt(C), corr.py:100).
#include "common.cuh"

namespace woft {

constexpr int kAltWarps = 4;
constexpr int kAltMaxVec = 4;  // C <= 4 * 128

__global__ void __launch_bounds__(kAltWarps * 32) alt_corr_kernel(const float* __restrict__ f1,
                                                                   const float* __restrict__ f2,
                                                                   const float* __restrict__ coords,
                                                                   float* __restrict__ out, int B, int H1, int W1,
                                                                   int H2, int W2, int C, int N, int r) {
  __shared__ float table[kAltWarps][10 * 10];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long total = (long long)B * N * H1 * W1;
  const long long item = (long long)blockIdx.x * kAltWarps + warp;
  if (item >= total) return;
  const int pix = int(item % (H1 * W1));
  const int n = int((item / (H1 * W1)) % N);
  const int b = int(item / ((long long)H1 * W1 * N));
  const int P = 2 * r + 1, WIN = P + 1;

  float4 a[kAltMaxVec];
  const float4* a_src = reinterpret_cast<const float4*>(f1 + ((size_t)b * H1 * W1 + pix) * C);
  const int nvec = C / 4;  // float4 per position
#pragma unroll
  for (int k = 0; k < kAltMaxVec; ++k) {
    const int v = lane + 32 * k;
    a[k] = v < nvec ? __ldg(a_src + v) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const float* cq = coords + (((size_t)b * N + n) * H1 * W1 + pix) * 2;
  const float x = __ldg(cq), y = __ldg(cq + 1);
  float fx = floorf(x), fy = floorf(y);
  const float dx = x - fx, dy = y - fy;
  fx = fminf(fmaxf(fx, -1048576.f), 1048576.f);
  fy = fminf(fmaxf(fy, -1048576.f), 1048576.f);
  const int x0 = int(fx) - r, y0 = int(fy) - r;

  for (int pos = 0; pos < WIN * WIN; ++pos) {
    const int iy = pos / WIN, ix = pos - iy * WIN;
    const int h2 = y0 + iy, w2 = x0 + ix;
    float s = 0.f;
    if (h2 >= 0 && h2 < H2 && w2 >= 0 && w2 < W2) {  // warp-uniform
      const float4* b_src = reinterpret_cast<const float4*>(f2 + (((size_t)b * H2 + h2) * W2 + w2) * C);
#pragma unroll
      for (int k = 0; k < kAltMaxVec; ++k) {
        const int v = lane + 32 * k;
        if (v < nvec) {
          const float4 bv = __ldg(b_src + v);
          s = fmaf(a[k].x, bv.x, s);
          s = fmaf(a[k].y, bv.y, s);
          s = fmaf(a[k].z, bv.z, s);
          s = fmaf(a[k].w, bv.w, s);
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    }
    if (lane == 0) table[warp][pos] = s;
  }
  __syncwarp();
  float* o_base = out + ((size_t)b * N + n) * P * P * H1 * W1 + pix;
  for (int o = lane; o < P * P; o += 32) {
    const int ixo = o / P, iyo = o - ixo * P;  // channel = iyo + P * ixo
    const float se = table[warp][iyo * WIN + ixo], sw = table[warp][iyo * WIN + ixo + 1];
    const float ne = table[warp][(iyo + 1) * WIN + ixo], nw = table[warp][(iyo + 1) * WIN + ixo + 1];
    // same four-corner sum as the reference scatter: nw*dy*dx + ne*dy*(1-dx) + sw*(1-dy)*dx + se*(1-dy)*(1-dx)
    const float val = nw * dy * dx + ne * dy * (1.f - dx) + sw * (1.f - dy) * dx + se * (1.f - dy) * (1.f - dx);
    o_base[(size_t)o * H1 * W1] = val;
  }
}

}  // namespace woft

extern "C" int woft_alt_corr_forward(const float* fmap1, const float* fmap2, const float* coords, int batch, int h1,
                                     int w1, int h2, int w2, int channels, int n_coords, int radius, float* out,
                                     void* stream) {
  using namespace woft;
  WOFT_REQUIRE(fmap1 && fmap2 && coords && out, "woft_alt_corr_forward: null pointer");
  WOFT_REQUIRE(batch > 0 && h1 > 0 && w1 > 0 && h2 > 0 && w2 > 0 && n_coords > 0, "woft_alt_corr_forward: bad extent");
  WOFT_REQUIRE(radius >= 1 && radius <= 4, "woft_alt_corr_forward: radius %d not in [1,4]", radius);
  WOFT_REQUIRE(channels % 4 == 0 && channels <= 128 * kAltMaxVec,
               "woft_alt_corr_forward: channels %d must be a multiple of 4 and <= %d", channels, 128 * kAltMaxVec);
  WOFT_REQUIRE((reinterpret_cast<uintptr_t>(fmap1) & 15) == 0 && (reinterpret_cast<uintptr_t>(fmap2) & 15) == 0,
               "woft_alt_corr_forward: feature maps must be 16-byte aligned");
  const long long total = (long long)batch * n_coords * h1 * w1;
  const long long blocks = (total + kAltWarps - 1) / kAltWarps;
  WOFT_REQUIRE(blocks < (1ll << 31), "woft_alt_corr_forward: too many query pixels");
  alt_corr_kernel<<<(unsigned)blocks, kAltWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
      fmap1, fmap2, coords, out, batch, h1, w1, h2, w2, channels, n_coords, radius);
  return launch_status("alt_corr_kernel");
}
