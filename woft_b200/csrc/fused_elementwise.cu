// Fused element-wise stages of the flow network that sit between cuDNN convolutions.
//
// SepConvGRU.forward (pytracking/external/RAFT/raft_core/update.py:45-60) per half step:
//     z = sigmoid(convz(hx)); r = sigmoid(convr(hx)); q = tanh(convq(cat[r*h, x])); h = (1-z)*h + z*q
// In PyTorch that is 3 bias adds + 2 sigmoids + 1 tanh + 1 mul + 4 blend kernels per half step, 24 half steps
// per flow pass.  Here the z and r convolutions are one bias-free convolution with stacked weights, and
//   gru_gates_kernel   z = sigmoid(zr[:, :C] + bz),  rh = sigmoid(zr[:, C:] + br) * h
//   gru_update_kernel  h' = (1 - z) * h + z * tanh(q + bq)
// BasicEncoder with norm_fn='instance' (raft_core/extractor.py:118-192) does conv -> InstanceNorm2d -> ReLU;
//   instance_norm_relu_kernel  one CTA per (n, c) plane: mean / biased variance, then (x - mean) * rstd, ReLU
//                              (eps 1e-5, no affine: nn.InstanceNorm2d defaults), second read served by L2.
// All tensors NCHW fp32 contiguous.
#include <cuda_fp16.h>

#include "common.cuh"

namespace woft {

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

// zr: (B, 2C, HW); h: (B, C, HW) -> z, rh: (B, C, HW).  One thread per 4 consecutive pixels.
__global__ void gru_gates_kernel(const float* __restrict__ zr, const float* __restrict__ bias, const float* __restrict__ h,
                                 float* __restrict__ z, float* __restrict__ rh, int C, int HW4, long long total4) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4;
       i += (long long)gridDim.x * blockDim.x) {
    const int p4 = int(i % HW4);
    const long long bc = i / HW4;
    const int c = int(bc % C);
    const long long b = bc / C;
    const float4 zp = __ldg(reinterpret_cast<const float4*>(zr) + ((b * 2 * C + c) * HW4 + p4));
    const float4 rp = __ldg(reinterpret_cast<const float4*>(zr) + ((b * 2 * C + C + c) * HW4 + p4));
    const float4 hv = __ldg(reinterpret_cast<const float4*>(h) + i);
    const float bz = __ldg(bias + c), br = __ldg(bias + C + c);
    float4 zo, ro;
    zo.x = sigmoidf_(zp.x + bz); zo.y = sigmoidf_(zp.y + bz); zo.z = sigmoidf_(zp.z + bz); zo.w = sigmoidf_(zp.w + bz);
    ro.x = sigmoidf_(rp.x + br) * hv.x; ro.y = sigmoidf_(rp.y + br) * hv.y;
    ro.z = sigmoidf_(rp.z + br) * hv.z; ro.w = sigmoidf_(rp.w + br) * hv.w;
    reinterpret_cast<float4*>(z)[i] = zo;
    reinterpret_cast<float4*>(rh)[i] = ro;
  }
}

__global__ void gru_update_kernel(const float* __restrict__ q, const float* __restrict__ bias, const float* __restrict__ z,
                                  const float* __restrict__ h, float* __restrict__ h_new, int C, int HW4,
                                  long long total4) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = int((i / HW4) % C);
    const float bq = __ldg(bias + c);
    const float4 qv = __ldg(reinterpret_cast<const float4*>(q) + i);
    const float4 zv = __ldg(reinterpret_cast<const float4*>(z) + i);
    const float4 hv = __ldg(reinterpret_cast<const float4*>(h) + i);
    float4 o;
    o.x = (1.f - zv.x) * hv.x + zv.x * tanhf(qv.x + bq);
    o.y = (1.f - zv.y) * hv.y + zv.y * tanhf(qv.y + bq);
    o.z = (1.f - zv.z) * hv.z + zv.z * tanhf(qv.z + bq);
    o.w = (1.f - zv.w) * hv.w + zv.w * tanhf(qv.w + bq);
    reinterpret_cast<float4*>(h_new)[i] = o;
  }
}

__global__ void __launch_bounds__(512) instance_norm_relu_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                                 int HW, float eps, int relu) {
  __shared__ float s_sum[16], s_sq[16];
  __shared__ float s_mean, s_rstd;
  const float* src = x + (size_t)blockIdx.x * HW;
  float* dst = y + (size_t)blockIdx.x * HW;
  const int n4 = HW >> 2;
  // pass 1: sum and sum of squares around a pivot (the first element) to keep the variance well conditioned
  const float pivot = __ldg(src);
  float s = 0.f, q = 0.f;
  for (int i = threadIdx.x; i < n4; i += blockDim.x) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(src) + i);
    const float a = v.x - pivot, b = v.y - pivot, c = v.z - pivot, d = v.w - pivot;
    s += (a + b) + (c + d);
    q += (a * a + b * b) + (c * c + d * d);
  }
  for (int i = (n4 << 2) + threadIdx.x; i < HW; i += blockDim.x) {
    const float a = __ldg(src + i) - pivot;
    s += a;
    q += a * a;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if ((threadIdx.x & 31) == 0) {
    s_sum[threadIdx.x >> 5] = s;
    s_sq[threadIdx.x >> 5] = q;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double ts = 0.0, tq = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) {
      ts += s_sum[w];
      tq += s_sq[w];
    }
    const double m = ts / HW;
    const double var = tq / HW - m * m;  // biased variance, as nn.InstanceNorm2d uses for normalisation
    s_mean = float(m) + pivot;
    s_rstd = rsqrtf(fmaxf(float(var), 0.f) + eps);
  }
  __syncthreads();
  const float mean = s_mean, rstd = s_rstd;
  const float floor_ = relu ? 0.f : -3.4e38f;
  for (int i = threadIdx.x; i < n4; i += blockDim.x) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(src) + i);
    float4 o;
    o.x = fmaxf((v.x - mean) * rstd, floor_);
    o.y = fmaxf((v.y - mean) * rstd, floor_);
    o.z = fmaxf((v.z - mean) * rstd, floor_);
    o.w = fmaxf((v.w - mean) * rstd, floor_);
    reinterpret_cast<float4*>(dst)[i] = o;
  }
  for (int i = (n4 << 2) + threadIdx.x; i < HW; i += blockDim.x) dst[i] = fmaxf((__ldg(src + i) - mean) * rstd, floor_);
}

// ---- channels-last (NHWC) variants: a pixel's channels are contiguous; operands may be channel slices of wider
// buffers (pixel stride > channel count), which is how the GRU avoids torch.cat: [h | x] and [r*h | x] live in two
// persistent (pixels, C + Cx) buffers whose x part is written once per iteration.
// Index math is 32-bit (the entry points bound the element count): a 64-bit division per element costs more than the
// loads it addresses.
__global__ void gru_gates_nhwc_kernel(const float* __restrict__ z_pre, const float* __restrict__ r_pre, int pre_stride,
                                      const float* __restrict__ bias_z, const float* __restrict__ bias_r,
                                      const float* __restrict__ h, int h_stride, float* __restrict__ z,
                                      float* __restrict__ rh, int rh_stride, unsigned C4, unsigned total4) {
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += gridDim.x * blockDim.x) {
    const unsigned pix = i / C4;
    const unsigned c4 = i - pix * C4;
    const float4 zp = __ldg(reinterpret_cast<const float4*>(z_pre + (size_t)pix * pre_stride) + c4);
    const float4 rp = __ldg(reinterpret_cast<const float4*>(r_pre + (size_t)pix * pre_stride) + c4);
    const float4 hv = __ldg(reinterpret_cast<const float4*>(h + (size_t)pix * h_stride) + c4);
    const float4 bz = __ldg(reinterpret_cast<const float4*>(bias_z) + c4);
    const float4 br = __ldg(reinterpret_cast<const float4*>(bias_r) + c4);
    float4 zo, ro;
    zo.x = sigmoidf_(zp.x + bz.x); zo.y = sigmoidf_(zp.y + bz.y); zo.z = sigmoidf_(zp.z + bz.z); zo.w = sigmoidf_(zp.w + bz.w);
    ro.x = sigmoidf_(rp.x + br.x) * hv.x; ro.y = sigmoidf_(rp.y + br.y) * hv.y;
    ro.z = sigmoidf_(rp.z + br.z) * hv.z; ro.w = sigmoidf_(rp.w + br.w) * hv.w;
    reinterpret_cast<float4*>(z)[i] = zo;
    reinterpret_cast<float4*>(rh + (size_t)pix * rh_stride)[c4] = ro;
  }
}

__global__ void gru_update_nhwc_kernel(const float* __restrict__ q, const float* __restrict__ bias,
                                       const float* __restrict__ z, const float* h, int h_stride, float* h_out,
                                       int h_out_stride, float* __restrict__ h_dense, unsigned C4, unsigned total4) {
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += gridDim.x * blockDim.x) {
    const unsigned pix = i / C4;
    const unsigned c4 = i - pix * C4;
    const float4 qv = __ldg(reinterpret_cast<const float4*>(q) + i);
    const float4 zv = __ldg(reinterpret_cast<const float4*>(z) + i);
    const float4 hv = *(reinterpret_cast<const float4*>(h + (size_t)pix * h_stride) + c4);  // may alias h_out
    const float4 bq = __ldg(reinterpret_cast<const float4*>(bias) + c4);
    float4 o;
    o.x = (1.f - zv.x) * hv.x + zv.x * tanhf(qv.x + bq.x);
    o.y = (1.f - zv.y) * hv.y + zv.y * tanhf(qv.y + bq.y);
    o.z = (1.f - zv.z) * hv.z + zv.z * tanhf(qv.z + bq.z);
    o.w = (1.f - zv.w) * hv.w + zv.w * tanhf(qv.w + bq.w);
    reinterpret_cast<float4*>(h_out + (size_t)pix * h_out_stride)[c4] = o;
    if (h_dense) reinterpret_cast<float4*>(h_dense)[i] = o;
  }
}

struct ConcatParams {
  const float* src[4];
  int ch[4];          // channels taken from each source
  int src_stride[4];  // floats between consecutive pixels of each source (>= ch: a source may be a channel prefix)
  int off[4];         // channel offset of each source inside the concatenation
  int n_src;
  int total_ch;
  float* dst[2];
  int dst_stride[2];
  int dst_off[2];
  int n_dst;
};

// dst[d][pix * stride_d + off_d + off_i + c] = src_i[pix * src_stride_i + c]; VW floats per thread and step.  The source
// is picked with unrolled selects: indexing the parameter struct with a run-time value would move it to local memory.
template <int VW>
__global__ void concat_nhwc_kernel(const __grid_constant__ ConcatParams p, unsigned total) {
  const unsigned per_pixel = unsigned(p.total_ch) / VW;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned pix = i / per_pixel;
    const int c = int(i - pix * per_pixel) * VW;  // channel within the concatenation
    const float* base = p.src[0];
    int stride = p.src_stride[0], off = 0;
#pragma unroll
    for (int s = 1; s < 4; ++s)
      if (s < p.n_src && c >= p.off[s]) {
        base = p.src[s];
        stride = p.src_stride[s];
        off = p.off[s];
      }
    const float* src = base + (size_t)pix * stride + (c - off);
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (VW == 4) {
      const float4 t = __ldg(reinterpret_cast<const float4*>(src));
      v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else if (VW == 2) {
      const float2 t = __ldg(reinterpret_cast<const float2*>(src));
      v[0] = t.x; v[1] = t.y;
    } else {
      v[0] = __ldg(src);
    }
#pragma unroll
    for (int d = 0; d < 2; ++d) {
      if (d >= p.n_dst) break;
      float* o = p.dst[d] + (size_t)pix * p.dst_stride[d] + p.dst_off[d] + c;
      if (VW == 4)
        *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
      else if (VW == 2)
        *reinterpret_cast<float2*>(o) = make_float2(v[0], v[1]);
      else
        o[0] = v[0];
    }
  }
}

// ---- instance norm + ReLU on channels-last tensors (N, HW, C): per-strip partial sums, a tiny finalise, then apply
// element type of the channels-last instance-norm kernels: fp32, or fp16 for the encoders that feed the fp16 tensor-core
// convolutions (statistics and arithmetic stay fp32)
template <typename T> __device__ __forceinline__ float4 ld4(const T* p, size_t i4);
template <> __device__ __forceinline__ float4 ld4<float>(const float* p, size_t i4) {
  return __ldg(reinterpret_cast<const float4*>(p) + i4);
}
template <> __device__ __forceinline__ float4 ld4<__half>(const __half* p, size_t i4) {
  const uint2 u = __ldg(reinterpret_cast<const uint2*>(p) + i4);
  const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&u.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&u.y));
  return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ void st4(float* p, size_t i4, float4 v) { reinterpret_cast<float4*>(p)[i4] = v; }
__device__ __forceinline__ void st4(__half* p, size_t i4, float4 v) {
  const __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
  uint2 u;
  u.x = *reinterpret_cast<const uint32_t*>(&a);
  u.y = *reinterpret_cast<const uint32_t*>(&b);
  reinterpret_cast<uint2*>(p)[i4] = u;
}
__device__ __forceinline__ float ld1(const float* p, size_t i) { return __ldg(p + i); }
__device__ __forceinline__ float ld1(const __half* p, size_t i) { return __half2float(p[i]); }

constexpr int kInStrips = 256;  // (strips x batch) CTAs in the partial-sum pass: enough to fill 148 SMs at batch 1-2

template <typename T>
__global__ void __launch_bounds__(256) in_partial_nhwc_kernel(const T* __restrict__ x, float* __restrict__ partial,
                                                              int HW, int C) {
  __shared__ float4 s_sum[256], s_sq[256];
  const int n = blockIdx.y, strip = blockIdx.x;
  const int c4n = C >> 2, ppb = 256 / c4n;
  const int cq = threadIdx.x % c4n, pl = threadIdx.x / c4n;
  const int len = (HW + kInStrips - 1) / kInStrips;
  const int p0 = strip * len, p1 = min(p0 + len, HW);
  const T* base = x + (size_t)n * HW * C;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f), q = s;
  if (pl < ppb) {
    const float4 pv = ld4<T>(base, cq);  // pivot: the plane's first pixel, keeps the variance well conditioned
    auto add = [&](const float4 v) {
      const float a = v.x - pv.x, b = v.y - pv.y, c = v.z - pv.z, d = v.w - pv.w;
      s.x += a; s.y += b; s.z += c; s.w += d;
      q.x += a * a; q.y += b * b; q.z += c * c; q.w += d * d;
    };
    int p = p0 + pl;
    for (; p + 3 * ppb < p1; p += 4 * ppb) {  // four loads in flight per thread: the pass is latency bound otherwise
      const float4 v0 = ld4<T>(base, (size_t)p * c4n + cq);
      const float4 v1 = ld4<T>(base, (size_t)(p + ppb) * c4n + cq);
      const float4 v2 = ld4<T>(base, (size_t)(p + 2 * ppb) * c4n + cq);
      const float4 v3 = ld4<T>(base, (size_t)(p + 3 * ppb) * c4n + cq);
      add(v0); add(v1); add(v2); add(v3);
    }
    for (; p < p1; p += ppb) add(ld4<T>(base, (size_t)p * c4n + cq));
  }
  s_sum[threadIdx.x] = s;
  s_sq[threadIdx.x] = q;
  __syncthreads();
  if (pl == 0) {
    for (int k = 1; k < ppb; ++k) {
      const float4 a = s_sum[k * c4n + cq], b = s_sq[k * c4n + cq];
      s.x += a.x; s.y += a.y; s.z += a.z; s.w += a.w;
      q.x += b.x; q.y += b.y; q.z += b.z; q.w += b.w;
    }
    float4* dst = reinterpret_cast<float4*>(partial + ((size_t)(n * kInStrips + strip) * 2) * C);
    dst[cq] = s;
    dst[c4n + cq] = q;
  }
}

// one warp per channel: lanes add up the strips, 8 channels per CTA
template <typename T>
__global__ void __launch_bounds__(256) in_finalize_nhwc_kernel(const T* __restrict__ x,
                                                               const float* __restrict__ partial,
                                                               float* __restrict__ stats, int HW, int C, float eps) {
  const int n = blockIdx.y;
  const int c = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (c >= C) return;
  double s = 0.0, q = 0.0;
  for (int k = lane; k < kInStrips; k += 32) {
    const float* pp = partial + ((size_t)(n * kInStrips + k) * 2) * C;
    s += pp[c];
    q += pp[C + c];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if (lane == 0) {
    const double m = s / HW, var = q / HW - m * m;
    stats[(size_t)n * 2 * C + c] = float(m) + ld1(x, (size_t)n * HW * C + c);
    stats[(size_t)n * 2 * C + C + c] = rsqrtf(fmaxf(float(var), 0.f) + eps);
  }
}

// The same from sums a producer gathered while it wrote x (woft_conv_tc `stats`): partial is (rows, batch, 2, C), row k the
// sums and sums of squares of the values CTA k stored.  No pivot here: the rows hold a few thousand values each in fp32,
// the combination over rows is done in double.
__global__ void __launch_bounds__(256) in_finalize_sums_kernel(const float* __restrict__ partial, int rows, int batch,
                                                               float* __restrict__ stats, int HW, int C, float eps) {
  const int n = blockIdx.y;
  const int c = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (c >= C) return;
  double s = 0.0, q = 0.0;
  for (int k = lane; k < rows; k += 32) {
    const float* pp = partial + ((size_t)(k * batch + n) * 2) * C;
    s += pp[c];
    q += pp[C + c];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  if (lane == 0) {
    const double m = s / HW, var = q / HW - m * m;
    stats[(size_t)n * 2 * C + c] = float(m);
    stats[(size_t)n * 2 * C + C + c] = rsqrtf(fmaxf(float(var), 0.f) + eps);
  }
}

// `residual` (optional): y = relu(residual + act(norm(x))), the tail of a residual block (extractor.py:56) in one pass
template <typename T>
__global__ void in_apply_nhwc_kernel(const T* __restrict__ x, const float* __restrict__ stats,
                                     const T* __restrict__ residual, T* __restrict__ y, unsigned HW, unsigned C,
                                     int relu, unsigned total4) {
  const unsigned c4n = C >> 2;
  const unsigned per_image = HW * c4n;
  const float floor_ = relu ? 0.f : -3.4e38f;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += gridDim.x * blockDim.x) {
    const unsigned cq = i % c4n;
    const unsigned n = i / per_image;
    const float4 v = ld4<T>(x, i);
    const float4 m = __ldg(reinterpret_cast<const float4*>(stats + (size_t)n * 2 * C) + cq);
    const float4 r = __ldg(reinterpret_cast<const float4*>(stats + (size_t)n * 2 * C + C) + cq);
    float4 o;
    o.x = fmaxf((v.x - m.x) * r.x, floor_);
    o.y = fmaxf((v.y - m.y) * r.y, floor_);
    o.z = fmaxf((v.z - m.z) * r.z, floor_);
    o.w = fmaxf((v.w - m.w) * r.w, floor_);
    if (residual) {
      const float4 a = ld4<T>(residual, i);
      o.x = fmaxf(a.x + o.x, 0.f); o.y = fmaxf(a.y + o.y, 0.f); o.z = fmaxf(a.z + o.z, 0.f); o.w = fmaxf(a.w + o.w, 0.f);
    }
    st4(y, i, o);
  }
}

// fp16, channels a multiple of 8: 16 bytes per thread and access.  x and y may be the same tensor (the normalisation of
// the encoders runs in place, so the tensor crosses L2 once instead of twice); neither is declared restrict.
__global__ void __launch_bounds__(256) in_apply_nhwc_h8_kernel(const __half* x, const float* __restrict__ stats,
                                                               const __half* __restrict__ residual, __half* y, unsigned HW,
                                                               unsigned C, int relu, unsigned total8) {
  const unsigned c8n = C >> 3;
  const unsigned per_image = HW * c8n;
  const float floor_ = relu ? 0.f : -3.4e38f;
  const unsigned stride = gridDim.x * blockDim.x;
  // the launch makes the grid stride a multiple of the channel octets, so a thread keeps one octet for the whole loop and
  // its mean / inverse deviation stay in registers (reloading 64 bytes of statistics per 16 bytes of data made the L1 the
  // busiest unit: 73 %); they are reloaded only when the loop crosses into the next image
  const bool fixed = stride % c8n == 0;
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned cq = i % c8n, n_cur = 0xffffffffu;
  float m[8], r[8];
  constexpr int U = 4;   // vectors in flight per thread: one 16-byte load per thread and round trip left DRAM at 30 %
  for (; i < total8; i += U * stride) {
    uint4 u[U], a[U];
#pragma unroll
    for (int k = 0; k < U; ++k) {
      const unsigned ik = i + k * stride;
      if (ik < total8) {
        u[k] = reinterpret_cast<const uint4*>(x)[ik];
        if (residual) a[k] = __ldg(reinterpret_cast<const uint4*>(residual) + ik);
      }
    }
#pragma unroll
    for (int k = 0; k < U; ++k) {
      const unsigned ik = i + k * stride;
      if (ik >= total8) break;
      if (!fixed) cq = ik % c8n;
      const unsigned n = ik / per_image;
      if (n != n_cur || !fixed) {
        const float4* mp = reinterpret_cast<const float4*>(stats + (size_t)n * 2 * C) + 2 * cq;
        const float4* rp = reinterpret_cast<const float4*>(stats + (size_t)n * 2 * C + C) + 2 * cq;
        const float4 m0 = __ldg(mp), m1 = __ldg(mp + 1), r0 = __ldg(rp), r1 = __ldg(rp + 1);
        m[0] = m0.x; m[1] = m0.y; m[2] = m0.z; m[3] = m0.w; m[4] = m1.x; m[5] = m1.y; m[6] = m1.z; m[7] = m1.w;
        r[0] = r0.x; r[1] = r0.y; r[2] = r0.z; r[3] = r0.w; r[4] = r1.x; r[5] = r1.y; r[6] = r1.z; r[7] = r1.w;
        n_cur = n;
      }
      const __half2* hp = reinterpret_cast<const __half2*>(&u[k]);
      float o[8];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = __half22float2(hp[e]);
        o[2 * e] = fmaxf((f.x - m[2 * e]) * r[2 * e], floor_);
        o[2 * e + 1] = fmaxf((f.y - m[2 * e + 1]) * r[2 * e + 1], floor_);
      }
      if (residual) {
        const __half2* ap = reinterpret_cast<const __half2*>(&a[k]);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = __half22float2(ap[e]);
          o[2 * e] = fmaxf(f.x + o[2 * e], 0.f);
          o[2 * e + 1] = fmaxf(f.y + o[2 * e + 1], 0.f);
        }
      }
      uint4 w;
      __half2* wp = reinterpret_cast<__half2*>(&w);
#pragma unroll
      for (int e = 0; e < 4; ++e) wp[e] = __floats2half2_rn(o[2 * e], o[2 * e + 1]);
      reinterpret_cast<uint4*>(y)[ik] = w;
    }
  }
}

// x[pix][c] = scale * (x[pix][c] + bias[c]) in place, optional ReLU: the bias of a convolution whose cuDNN call has no
// fused epilogue (ATen adds a broadcast bias to a channels-last tensor with its slow, non-vectorised kernel)
__global__ void bias_act_nhwc_kernel(float* __restrict__ x, const float* __restrict__ bias, unsigned C4, float scale,
                                     int relu, unsigned total4) {
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += gridDim.x * blockDim.x) {
    const float4 b = __ldg(reinterpret_cast<const float4*>(bias) + i % C4);
    float4 v = reinterpret_cast<float4*>(x)[i];
    v.x = scale * (v.x + b.x); v.y = scale * (v.y + b.y); v.z = scale * (v.z + b.z); v.w = scale * (v.w + b.w);
    if (relu) {
      v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
    }
    reinterpret_cast<float4*>(x)[i] = v;
  }
}

// Tail of the context encoder (raft_core/weighted_raft.py:215-219): cnet's output + bias split into the GRU's initial
// hidden state tanh(.) and the context features relu(.), written where the update block wants them: h dense and at the
// head of the [h | x] buffer, the context at channel offset C of both GRU buffers.
__global__ void context_split_nhwc_kernel(const float* __restrict__ y, const float* __restrict__ bias, unsigned C4,
                                          float* __restrict__ h_dense, float* __restrict__ buf1, float* __restrict__ buf2,
                                          int ld, unsigned total4) {
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += gridDim.x * blockDim.x) {
    const unsigned pix = i / (2 * C4), c4 = i - pix * 2 * C4;
    const float4 b = __ldg(reinterpret_cast<const float4*>(bias) + c4);
    float4 v = __ldg(reinterpret_cast<const float4*>(y) + i);
    v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;
    if (c4 < C4) {
      v = make_float4(tanhf(v.x), tanhf(v.y), tanhf(v.z), tanhf(v.w));
      reinterpret_cast<float4*>(h_dense)[pix * C4 + c4] = v;
      reinterpret_cast<float4*>(buf1 + (size_t)pix * ld)[c4] = v;
    } else {
      v = make_float4(fmaxf(v.x, 0.f), fmaxf(v.y, 0.f), fmaxf(v.z, 0.f), fmaxf(v.w, 0.f));
      reinterpret_cast<float4*>(buf1 + (size_t)pix * ld)[c4] = v;
      reinterpret_cast<float4*>(buf2 + (size_t)pix * ld)[c4] = v;
    }
  }
}

__global__ void add_relu_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ y,
                                long long total4) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4;
       i += (long long)gridDim.x * blockDim.x) {
    const float4 u = __ldg(reinterpret_cast<const float4*>(a) + i);
    const float4 v = __ldg(reinterpret_cast<const float4*>(b) + i);
    float4 o;
    o.x = fmaxf(u.x + v.x, 0.f); o.y = fmaxf(u.y + v.y, 0.f); o.z = fmaxf(u.z + v.z, 0.f); o.w = fmaxf(u.w + v.w, 0.f);
    reinterpret_cast<float4*>(y)[i] = o;
  }
}

constexpr long long kMaxElems32 = 0x7fffffffLL;  // the NHWC kernels index with 32-bit integers

static int ew_grid(long long total4) {
  const long long blocks = (total4 + 255) / 256;
  return int(blocks < 148 * 16 ? (blocks > 0 ? blocks : 1) : 148 * 16);
}

// fp32 channels (pixels, src_stride) -> fp16 channel slices of one or two (pixels, dst_stride) tensors: feeds the fp16
// operand buffers of the tensor-core convolutions (context features, hidden state, the 2-channel flow) from fp32 producers.
__global__ void cast_nhwc_f16_kernel(const float* __restrict__ src, int src_stride, int channels, unsigned total,
                                     __half* __restrict__ d1, int d1_stride, int d1_off, __half* __restrict__ d2,
                                     int d2_stride, int d2_off) {
  const unsigned per_pixel = unsigned(channels) / 2;
  pdl_sync();
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned pix = i / per_pixel;
    const int c = int(i - pix * per_pixel) * 2;
    const float2 v = __ldg(reinterpret_cast<const float2*>(src + (size_t)pix * src_stride + c));
    const __half2 hv = __floats2half2_rn(v.x, v.y);
    *reinterpret_cast<__half2*>(d1 + (size_t)pix * d1_stride + d1_off + c) = hv;
    if (d2) *reinterpret_cast<__half2*>(d2 + (size_t)pix * d2_stride + d2_off + c) = hv;
  }
}

}  // namespace woft

extern "C" int woft_gru_gates(const float* zr, const float* bias_zr, const float* h, int batch, int channels, int hw,
                              float* z, float* rh, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(zr && bias_zr && h && z && rh, "woft_gru_gates: null pointer");
  WOFT_REQUIRE(batch > 0 && channels > 0 && hw > 0 && hw % 4 == 0, "woft_gru_gates: H*W must be a positive multiple of 4");
  const long long total4 = (long long)batch * channels * (hw / 4);
  gru_gates_kernel<<<ew_grid(total4), 256, 0, static_cast<cudaStream_t>(stream)>>>(zr, bias_zr, h, z, rh, channels,
                                                                                   hw / 4, total4);
  return launch_status("gru_gates_kernel");
}

extern "C" int woft_gru_update(const float* q, const float* bias_q, const float* z, const float* h, int batch,
                               int channels, int hw, float* h_new, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(q && bias_q && z && h && h_new, "woft_gru_update: null pointer");
  WOFT_REQUIRE(batch > 0 && channels > 0 && hw > 0 && hw % 4 == 0, "woft_gru_update: H*W must be a positive multiple of 4");
  const long long total4 = (long long)batch * channels * (hw / 4);
  gru_update_kernel<<<ew_grid(total4), 256, 0, static_cast<cudaStream_t>(stream)>>>(q, bias_q, z, h, h_new, channels,
                                                                                    hw / 4, total4);
  return launch_status("gru_update_kernel");
}

extern "C" int woft_instance_norm_relu(const float* x, int planes, int hw, float eps, int relu, float* y, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(x && y, "woft_instance_norm_relu: null pointer");
  WOFT_REQUIRE(planes > 0 && hw > 0, "woft_instance_norm_relu: bad extent");
  WOFT_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(y) & 15) == 0 &&
                   (hw % 4 == 0),
               "woft_instance_norm_relu: planes must be 16-byte aligned (H*W a multiple of 4)");
  instance_norm_relu_kernel<<<planes, 512, 0, static_cast<cudaStream_t>(stream)>>>(x, y, hw, eps, relu);
  return launch_status("instance_norm_relu_kernel");
}

extern "C" int woft_gru_gates_nhwc(const float* zr, const float* bias_zr, const float* h, int h_stride,
                                   long long pixels, int channels, float* z, float* rh, int rh_stride, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(zr && bias_zr && h && z && rh, "woft_gru_gates_nhwc: null pointer");
  WOFT_REQUIRE(pixels > 0 && channels > 0 && channels % 4 == 0 && h_stride % 4 == 0 && rh_stride % 4 == 0,
               "woft_gru_gates_nhwc: channels and strides must be positive multiples of 4");
  const long long total4 = pixels * (channels / 4);
  WOFT_REQUIRE(total4 <= kMaxElems32, "woft_gru_gates_nhwc: more than 2^31 - 1 vectors");
  gru_gates_nhwc_kernel<<<ew_grid(total4), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      zr, zr + channels, 2 * channels, bias_zr, bias_zr + channels, h, h_stride, z, rh, rh_stride, channels / 4,
      unsigned(total4));
  return launch_status("gru_gates_nhwc_kernel");
}

extern "C" int woft_gru_gates_nhwc_split(const float* z_pre, const float* r_pre, int pre_stride, const float* bias_z,
                                         const float* bias_r, const float* h, int h_stride, long long pixels,
                                         int channels, float* z, float* rh, int rh_stride, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(z_pre && r_pre && bias_z && bias_r && h && z && rh, "woft_gru_gates_nhwc_split: null pointer");
  WOFT_REQUIRE(pixels > 0 && channels > 0 && channels % 4 == 0 && h_stride % 4 == 0 && rh_stride % 4 == 0 &&
                   pre_stride >= channels && pre_stride % 4 == 0,
               "woft_gru_gates_nhwc_split: channels and strides must be positive multiples of 4, pre_stride >= channels");
  const long long total4 = pixels * (channels / 4);
  WOFT_REQUIRE(total4 <= kMaxElems32, "woft_gru_gates_nhwc_split: more than 2^31 - 1 vectors");
  gru_gates_nhwc_kernel<<<ew_grid(total4), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      z_pre, r_pre, pre_stride, bias_z, bias_r, h, h_stride, z, rh, rh_stride, channels / 4, unsigned(total4));
  return launch_status("gru_gates_nhwc_kernel");
}

extern "C" int woft_gru_update_nhwc(const float* q, const float* bias_q, const float* z, const float* h, int h_stride,
                                    long long pixels, int channels, float* h_out, int h_out_stride, float* h_dense,
                                    void* stream) {
  using namespace woft;
  WOFT_REQUIRE(q && bias_q && z && h && h_out, "woft_gru_update_nhwc: null pointer");
  WOFT_REQUIRE(pixels > 0 && channels > 0 && channels % 4 == 0 && h_stride % 4 == 0 && h_out_stride % 4 == 0,
               "woft_gru_update_nhwc: channels and strides must be positive multiples of 4");
  const long long total4 = pixels * (channels / 4);
  WOFT_REQUIRE(total4 <= kMaxElems32, "woft_gru_update_nhwc: more than 2^31 - 1 vectors");
  gru_update_nhwc_kernel<<<ew_grid(total4), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      q, bias_q, z, h, h_stride, h_out, h_out_stride, h_dense, channels / 4, unsigned(total4));
  return launch_status("gru_update_nhwc_kernel");
}

static int concat_nhwc_impl(const float* const* srcs, const int* channels, const int* src_strides, int n_src,
                            long long pixels, float* dst, int dst_stride, int dst_offset, float* dst2, int dst2_stride,
                            int dst2_offset, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(srcs && channels && dst, "woft_concat_nhwc: null pointer");
  WOFT_REQUIRE(n_src >= 1 && n_src <= 4 && pixels > 0, "woft_concat_nhwc: 1..4 sources and a positive pixel count");
  ConcatParams p{};
  int off = 0, vw = 4;
  for (int i = 0; i < n_src; ++i) {
    WOFT_REQUIRE(srcs[i] && channels[i] > 0, "woft_concat_nhwc: source %d is empty", i);
    p.src[i] = srcs[i];
    p.ch[i] = channels[i];
    p.src_stride[i] = src_strides ? src_strides[i] : channels[i];
    WOFT_REQUIRE(p.src_stride[i] >= channels[i], "woft_concat_nhwc: source %d has a stride below its channel count", i);
    p.off[i] = off;
    off += channels[i];
    while (vw > 1 && (channels[i] % vw || p.off[i] % vw || p.src_stride[i] % vw ||
                      (reinterpret_cast<uintptr_t>(srcs[i]) % (4 * vw))))
      vw >>= 1;
  }
  p.n_src = n_src;
  p.total_ch = off;
  p.dst[0] = dst;
  p.dst_stride[0] = dst_stride;
  p.dst_off[0] = dst_offset;
  p.n_dst = 1;
  if (dst2) {
    p.dst[1] = dst2;
    p.dst_stride[1] = dst2_stride;
    p.dst_off[1] = dst2_offset;
    p.n_dst = 2;
  }
  for (int d = 0; d < p.n_dst; ++d)
    while (vw > 1 && (p.dst_stride[d] % vw || p.dst_off[d] % vw || (reinterpret_cast<uintptr_t>(p.dst[d]) % (4 * vw))))
      vw >>= 1;
  const long long total = pixels * (off / vw);
  WOFT_REQUIRE(total <= kMaxElems32, "woft_concat_nhwc: more than 2^31 - 1 vectors");
  const int grid = ew_grid(total);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (vw == 4)
    concat_nhwc_kernel<4><<<grid, 256, 0, s>>>(p, unsigned(total));
  else if (vw == 2)
    concat_nhwc_kernel<2><<<grid, 256, 0, s>>>(p, unsigned(total));
  else
    concat_nhwc_kernel<1><<<grid, 256, 0, s>>>(p, unsigned(total));
  return launch_status("concat_nhwc_kernel");
}

extern "C" int woft_concat_nhwc(const float* const* srcs, const int* channels, int n_src, long long pixels, float* dst,
                                int dst_stride, int dst_offset, float* dst2, int dst2_stride, int dst2_offset,
                                void* stream) {
  return concat_nhwc_impl(srcs, channels, nullptr, n_src, pixels, dst, dst_stride, dst_offset, dst2, dst2_stride,
                          dst2_offset, stream);
}

extern "C" int woft_concat_nhwc_strided(const float* const* srcs, const int* channels, const int* src_strides, int n_src,
                                        long long pixels, float* dst, int dst_stride, int dst_offset, float* dst2,
                                        int dst2_stride, int dst2_offset, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(src_strides, "woft_concat_nhwc_strided: null stride array");
  return concat_nhwc_impl(srcs, channels, src_strides, n_src, pixels, dst, dst_stride, dst_offset, dst2, dst2_stride,
                          dst2_offset, stream);
}

extern "C" int woft_cast_nhwc_f16(const float* src, int src_stride, int channels, long long pixels, void* dst,
                                  int dst_stride, int dst_offset, void* dst2, int dst2_stride, int dst2_offset, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(src && dst && pixels > 0 && channels > 0, "woft_cast_nhwc_f16: null pointer or empty tensor");
  WOFT_REQUIRE(channels % 2 == 0 && src_stride % 2 == 0 && dst_stride % 2 == 0 && dst_offset % 2 == 0 &&
               dst2_stride % 2 == 0 && dst2_offset % 2 == 0 && (reinterpret_cast<uintptr_t>(src) & 7) == 0 &&
               (reinterpret_cast<uintptr_t>(dst) & 3) == 0 && (reinterpret_cast<uintptr_t>(dst2) & 3) == 0,
               "woft_cast_nhwc_f16: channel counts, strides and offsets must be even and the pointers aligned to a pair");
  const long long total = pixels * (channels / 2);
  WOFT_REQUIRE(total <= kMaxElems32, "woft_cast_nhwc_f16: more than 2^31 - 1 pairs");
  WOFT_CUDA_CHECK(launch_pdl(cast_nhwc_f16_kernel, dim3(ew_grid(total)), dim3(256), 0, static_cast<cudaStream_t>(stream), src,
                             src_stride, channels, unsigned(total), static_cast<__half*>(dst), dst_stride, dst_offset,
                             static_cast<__half*>(dst2), dst2_stride, dst2_offset));
  return launch_status("cast_nhwc_f16_kernel");
}

extern "C" size_t woft_instance_norm_nhwc_scratch_bytes(int batch, int channels) {
  if (batch <= 0 || channels <= 0) return 0;
  return (size_t)batch * channels * 2 * (woft::kInStrips + 1) * sizeof(float);
}

template <typename T>
static int instance_norm_nhwc_impl(const char* who, const T* x, int batch, int hw, int channels, float eps, int relu,
                                   const T* residual, T* y, void* scratch, size_t scratch_bytes, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(x && y && scratch, "instance norm (NHWC): null pointer");
  WOFT_REQUIRE(batch > 0 && hw > 0 && channels > 0 && channels % 4 == 0 && channels <= 1024 && batch < 65536,
               "instance norm (NHWC): channels must be a multiple of 4 (<= 1024)");
  if (scratch_bytes < woft_instance_norm_nhwc_scratch_bytes(batch, channels))
    return fail(WOFT_ERR_WORKSPACE, "%s: scratch too small", who);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  float* partial = static_cast<float*>(scratch);
  float* stats = partial + (size_t)batch * kInStrips * 2 * channels;
  in_partial_nhwc_kernel<T><<<dim3(kInStrips, batch), 256, 0, s>>>(x, partial, hw, channels);
  in_finalize_nhwc_kernel<T><<<dim3(ceil_div(channels, 8), batch), 256, 0, s>>>(x, partial, stats, hw, channels, eps);
  const long long total4 = (long long)batch * hw * (channels / 4);
  WOFT_REQUIRE(total4 <= kMaxElems32, "instance norm (NHWC): more than 2^31 - 1 vectors");
  in_apply_nhwc_kernel<T><<<ew_grid(total4), 256, 0, s>>>(x, stats, residual, y, hw, channels, relu, unsigned(total4));
  return launch_status("instance norm (NHWC) kernels");
}

extern "C" int woft_instance_norm_relu_nhwc(const float* x, int batch, int hw, int channels, float eps, int relu,
                                            float* y, void* scratch, size_t scratch_bytes, void* stream) {
  return instance_norm_nhwc_impl<float>("woft_instance_norm_relu_nhwc", x, batch, hw, channels, eps, relu, nullptr, y,
                                        scratch, scratch_bytes, stream);
}

extern "C" int woft_instance_norm_relu_nhwc_f16(const void* x, int batch, int hw, int channels, float eps, int relu,
                                                const void* residual, void* y, void* scratch, size_t scratch_bytes,
                                                void* stream) {
  return instance_norm_nhwc_impl<__half>("woft_instance_norm_relu_nhwc_f16", static_cast<const __half*>(x), batch, hw,
                                         channels, eps, relu, static_cast<const __half*>(residual), static_cast<__half*>(y),
                                         scratch, scratch_bytes, stream);
}

extern "C" int woft_instance_norm_from_sums_f16(const void* x, const float* partial, int rows, int batch, int hw, int channels,
                                                float eps, int relu, const void* residual, void* y, void* scratch,
                                                size_t scratch_bytes, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(x && y && partial && scratch, "woft_instance_norm_from_sums_f16: null pointer");
  WOFT_REQUIRE(batch > 0 && hw > 0 && rows > 0 && channels > 0 && channels % 4 == 0 && channels <= 1024 && batch < 65536,
               "woft_instance_norm_from_sums_f16: channels must be a multiple of 4 (<= 1024)");
  WOFT_REQUIRE(!residual || relu, "woft_instance_norm_from_sums_f16: the residual tail applies ReLU");
  if (scratch_bytes < (size_t)batch * 2 * channels * sizeof(float))
    return fail(WOFT_ERR_WORKSPACE, "woft_instance_norm_from_sums_f16: scratch too small");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  float* stats = static_cast<float*>(scratch);
  in_finalize_sums_kernel<<<dim3(ceil_div(channels, 8), batch), 256, 0, s>>>(partial, rows, batch, stats, hw, channels, eps);
  const long long total4 = (long long)batch * hw * (channels / 4);
  WOFT_REQUIRE(total4 <= kMaxElems32, "woft_instance_norm_from_sums_f16: more than 2^31 - 1 vectors");
  const bool wide = channels % 8 == 0 && ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) |
                                           reinterpret_cast<uintptr_t>(residual)) & 15) == 0;
  if (wide) {   // x == y allowed
    // 8 CTAs per SM's worth of threads; the grid is a multiple of 3 so that 256 * grid is a multiple of 12 and 16 octets too
    const long long total8 = total4 / 2;
    int grid = 148 * 8 + 2;   // 1186 = 2 * 593; made a multiple of 3 below
    grid = (grid + 2) / 3 * 3;
    if ((long long)grid * 256 > total8) grid = int((total8 + 255) / 256);
    in_apply_nhwc_h8_kernel<<<grid, 256, 0, s>>>(static_cast<const __half*>(x), stats,
                                                                 static_cast<const __half*>(residual), static_cast<__half*>(y),
                                                                 hw, channels, relu, unsigned(total4 / 2));
  } else {
    in_apply_nhwc_kernel<__half><<<ew_grid(total4), 256, 0, s>>>(static_cast<const __half*>(x), stats,
                                                                  static_cast<const __half*>(residual), static_cast<__half*>(y),
                                                                  hw, channels, relu, unsigned(total4));
  }
  return launch_status("instance norm (from sums) kernels");
}

extern "C" int woft_instance_norm_relu_add_nhwc(const float* x, int batch, int hw, int channels, float eps,
                                                const float* residual, float* y, void* scratch, size_t scratch_bytes,
                                                void* stream) {
  using namespace woft;
  WOFT_REQUIRE(residual, "woft_instance_norm_relu_add_nhwc: null residual");
  return instance_norm_nhwc_impl<float>("woft_instance_norm_relu_add_nhwc", x, batch, hw, channels, eps, 1, residual, y,
                                        scratch, scratch_bytes, stream);
}

extern "C" int woft_add_relu(const float* a, const float* b, long long n, float* y, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(a && b && y, "woft_add_relu: null pointer");
  WOFT_REQUIRE(n > 0 && n % 4 == 0, "woft_add_relu: element count must be a positive multiple of 4");
  add_relu_kernel<<<ew_grid(n / 4), 256, 0, static_cast<cudaStream_t>(stream)>>>(a, b, y, n / 4);
  return launch_status("add_relu_kernel");
}

extern "C" int woft_bias_act_nhwc(float* x, const float* bias, long long pixels, int channels, float scale, int relu,
                                  void* stream) {
  using namespace woft;
  WOFT_REQUIRE(x && bias, "woft_bias_act_nhwc: null pointer");
  WOFT_REQUIRE(pixels > 0 && channels > 0 && channels % 4 == 0, "woft_bias_act_nhwc: channels must be a positive multiple of 4");
  const long long total4 = pixels * (channels / 4);
  WOFT_REQUIRE(total4 <= kMaxElems32, "woft_bias_act_nhwc: more than 2^31 - 1 vectors");
  bias_act_nhwc_kernel<<<ew_grid(total4), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, bias, channels / 4, scale, relu,
                                                                                      unsigned(total4));
  return launch_status("bias_act_nhwc_kernel");
}

extern "C" int woft_context_split_nhwc(const float* y, const float* bias, long long pixels, int channels, float* h_dense,
                                       float* buf1, float* buf2, int buf_stride, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(y && bias && h_dense && buf1 && buf2, "woft_context_split_nhwc: null pointer");
  WOFT_REQUIRE(pixels > 0 && channels > 0 && channels % 4 == 0 && buf_stride % 4 == 0 && buf_stride >= 2 * channels,
               "woft_context_split_nhwc: channels and stride must be multiples of 4, stride >= 2 * channels");
  const long long total4 = pixels * (channels / 2);
  WOFT_REQUIRE(total4 <= kMaxElems32, "woft_context_split_nhwc: more than 2^31 - 1 vectors");
  context_split_nhwc_kernel<<<ew_grid(total4), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      y, bias, channels / 4, h_dense, buf1, buf2, buf_stride, unsigned(total4));
  return launch_status("context_split_nhwc_kernel");
}
