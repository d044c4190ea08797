// K3: radius-r bilinear lookup into the 4-level correlation pyramid.
//
// Replaces CorrBlock.__call__ + bilinear_sampler of the reference
// (pytracking/external/RAFT/raft_core/corr.py:29-59, raft_core/utils/utils.py:59-73), which runs
// linspace/meshgrid + one grid_sample per level + cat + permute, 13 times per flow pass.
//
// Memory-bound gather.  Per query pixel and level the (2r+1)^2 bilinear taps share one fractional offset,
// so the kernel loads the (2r+2)^2 integer footprint once and forms all taps from it.  A CTA of 128 threads
// owns 16 consecutive query pixels and up to 4 levels at a time:
//   load phase     one thread per footprint ROW (4 levels x 16 queries x 10 rows = 640 rows, 5 per thread).
//                  A row is 40 contiguous bytes at an arbitrary 4-byte offset; when every level's width is a
//                  multiple of 4 the thread fetches it as the 3-4 aligned float4 chunks that cover it (a chunk
//                  is then wholly inside or wholly outside the map, so zero padding is a per-chunk predicate)
//                  and drops the values into shared memory with the sub-chunk shift applied; otherwise it
//                  falls back to 10 predicated scalar loads.  All of a thread's loads are issued before any
//                  is consumed, so a CTA keeps tens of kilobytes in flight.
//   compute phase  warp = level, lane = (query, upper/lower half of the window): horizontal then vertical
//                  lerp in registers, row by row, footprints read from shared memory at an odd per-query
//                  stride (conflict free).  Every tap is stored with consecutive lanes = consecutive queries,
//                  i.e. two full 64-byte segments per store instruction in the reference's channel-first layout.
// Algorithmic bytes per query pixel (r=4, 4 levels): 4*10*10*4 B read + 8 B coords + 324*4 B written = 2904 B.
// DRAM moves more than that: a 40-byte row at a random offset costs one or two 64-byte sectors pairs.
#include <cuda_fp16.h>

#include <type_traits>

#include "common.cuh"

namespace woft {

struct LookupParams {
  const float* level[kMaxLevels];   // const __half* in the instantiations for fp16 storage (flag 32)
  int lh[kMaxLevels];
  int lw[kMaxLevels];
  int wide[kMaxLevels];   // fp16 levels only: rows of this level may be read in 16-byte pieces (width % 8 == 0, 16-byte base)
  int num_levels;
  const float* coords;  // (B, 2, H1, W1)
  float* out;           // (B, L*P*P, H1, W1)
  int nq;               // B*H1*W1
  int hw1;              // H1*W1
  int out_f16;          // channels-last output holds __half (flag 16): the A operand of an fp16 tensor-core convolution
  int out_stride;       // channels-last output: elements between consecutive queries (channels, rounded up to 8 for fp16)
  // Optional (woft_corr_lookup_gather): the flow head's 9-tap gather and coordinate update (flow_head_gather_kernel,
  // weighted_raft.py:244-246) done by this launch for its own queries before it looks anything up.  `g_coords` is advanced
  // in place; the new flow goes out as NCHW, as (pixels, 2) and as fp16 into the two GRU operand buffers.
  const float* g_taps;     // (pixels, 18) or null
  const float* g_bias;     // (2)
  float* g_coords;         // (B, 2, H1, W1), read and written
  const float* g_coords0;
  float* g_flow_nchw;
  float* g_flow_nhwc;
  __half* g_h1;
  __half* g_h2;
  int g_stride, g_off, g_H, g_W;
};

constexpr int kLookupThreads = 128;
constexpr int kQueriesPerCta = 16;
constexpr int kLevelsPerPass = 4;  // one warp per level in the compute phase

template <int R>
struct LookupGeom {
  static constexpr int P = 2 * R + 1;
  static constexpr int WIN = P + 1;
  static constexpr int STRIDE = WIN * WIN + 1;  // odd => lanes (queries) hit distinct banks
  static constexpr int NCHUNK = (WIN + 3 + 3) / 4;  // aligned float4 chunks that cover any WIN-float run
  static constexpr int NWIDE = (WIN + 7 + 7) / 8;   // aligned 8-value (16-byte fp16) pieces that cover any WIN-value run
  static constexpr int RAW = 4 * NWIDE > 2 * NCHUNK ? 4 * NWIDE : 2 * NCHUNK;   // 32-bit words of a row kept packed
  static constexpr int ROWS = kLevelsPerPass * kQueriesPerCta * WIN;
  static constexpr int ROWS_PER_THREAD = (ROWS + kLookupThreads - 1) / kLookupThreads;
};

// four consecutive level values at an index that is a multiple of 4 (16-byte / 8-byte aligned), and one value
__device__ __forceinline__ float4 lk_ld4(const float* row, int x) { return __ldg(reinterpret_cast<const float4*>(row + x)); }
__device__ __forceinline__ float4 lk_ld4(const __half* row, int x) {
  const uint2 u = __ldg(reinterpret_cast<const uint2*>(row + x));
  const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&u.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&u.y));
  return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ float lk_ld1(const float* row, int x) { return __ldg(row + x); }
__device__ __forceinline__ float lk_ld1(const __half* row, int x) { return __half2float(row[x]); }

// T: element type of the pyramid levels (float, or __half as written with WOFT_CORR_STORE_F16)
template <int R, bool VEC, bool NHWC, typename T = float>
__global__ void __launch_bounds__(kLookupThreads, 6) corr_lookup_kernel(const __grid_constant__ LookupParams p) {
  using G = LookupGeom<R>;
  constexpr int P = G::P, WIN = G::WIN, STRIDE = G::STRIDE;
  // fp16 levels keep their footprints as fp16 in shared memory: half the carve-out leaves the L1 twice the room, and the
  // pieces of a footprint row depend on the line their first piece brought in (7 CTAs/SM or no-allocate loads, both of which
  // shrink that room, measured 6 % and 35 % slower)
  using WT = typename std::conditional<sizeof(T) == 2, __half, float>::type;
  __shared__ WT win[kLevelsPerPass * kQueriesPerCta * STRIDE];
  __shared__ int s_x0[kLevelsPerPass * kQueriesPerCta], s_y0[kLevelsPerPass * kQueriesPerCta];
  __shared__ float s_ax[kLevelsPerPass * kQueriesPerCta], s_ay[kLevelsPerPass * kQueriesPerCta];
  __shared__ float s_cx[kQueriesPerCta], s_cy[kQueriesPerCta];
  pdl_sync();   // the coordinates come from the previous kernel of the iteration

  const int tid = threadIdx.x;
  const int q0 = blockIdx.x * kQueriesPerCta;
  const int channels = p.num_levels * P * P;
  const bool gather = p.g_taps != nullptr;
  if (gather) {
    // the arithmetic of flow_head_gather_kernel, operation for operation (the two must agree bit for bit)
    if (tid < kQueriesPerCta && q0 + tid < p.nq) {
      const int i = q0 + tid, hw = p.hw1, H = p.g_H, W = p.g_W;
      const int b = i / hw, r = i - b * hw;
      const int y = r / W, x = r - y * W;
      float dx = __ldg(p.g_bias), dy = __ldg(p.g_bias + 1);
#pragma unroll
      for (int ky = 0; ky < 3; ++ky) {
        const int yy = y + ky - 1;
        if (yy < 0 || yy >= H) continue;
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          const int xx = x + kx - 1;
          if (xx < 0 || xx >= W) continue;
          const float2 t = __ldg(reinterpret_cast<const float2*>(p.g_taps + ((size_t)b * hw + yy * W + xx) * 18) + ky * 3 + kx);
          dx += t.x;
          dy += t.y;
        }
      }
      const size_t ix = (size_t)(b * 2) * hw + r, iy = ix + hw;
      const float cx = p.g_coords[ix] + dx, cy = p.g_coords[iy] + dy;
      p.g_coords[ix] = cx;
      p.g_coords[iy] = cy;
      s_cx[tid] = cx;
      s_cy[tid] = cy;
      const float fx = cx - __ldg(p.g_coords0 + ix), fy = cy - __ldg(p.g_coords0 + iy);
      if (p.g_flow_nchw) {
        p.g_flow_nchw[ix] = fx;
        p.g_flow_nchw[iy] = fy;
      }
      if (p.g_flow_nhwc) reinterpret_cast<float2*>(p.g_flow_nhwc)[i] = make_float2(fx, fy);
      const __half2 fh = __floats2half2_rn(fx, fy);
      if (p.g_h1) *reinterpret_cast<__half2*>(p.g_h1 + (size_t)i * p.g_stride + p.g_off) = fh;
      if (p.g_h2) *reinterpret_cast<__half2*>(p.g_h2 + (size_t)i * p.g_stride + p.g_off) = fh;
    }
    __syncthreads();
  }

  for (int lv0 = 0; lv0 < p.num_levels; lv0 += kLevelsPerPass) {
    // ---- window origins and fractions for (level slot, query): 64 threads
    if (tid < kLevelsPerPass * kQueriesPerCta) {
      const int slot = tid >> 4, ql = tid & 15;
      const int q = q0 + ql, lvl = lv0 + slot;
      float cx = 0.f, cy = 0.f;
      if (q < p.nq && lvl < p.num_levels) {
        if (gather) {
          cx = s_cx[ql];
          cy = s_cy[ql];
        } else {
          const int b = q / p.hw1, pix = q - b * p.hw1;
          cx = __ldg(p.coords + (size_t)(2 * b) * p.hw1 + pix);
          cy = __ldg(p.coords + (size_t)(2 * b + 1) * p.hw1 + pix);
        }
      }
      const float inv = 1.0f / float(1 << (lvl & 31));
      const float x = cx * inv, y = cy * inv;  // centroid_lvl = coords / 2**i (corr.py:41)
      float fx = floorf(x), fy = floorf(y);
      s_ax[tid] = x - fx;
      s_ay[tid] = y - fy;
      fx = fminf(fmaxf(fx, -1048576.f), 1048576.f);
      fy = fminf(fmaxf(fy, -1048576.f), 1048576.f);
      s_x0[tid] = int(fx) - R;
      s_y0[tid] = int(fy) - R;
    }
    __syncthreads();

    // ---- load phase: one footprint row per thread and step, all loads issued before any store
    // fp16 levels, vector path: a row stays packed in registers (12 words instead of 16 floats) and is fetched in 16-byte
    // pieces where the level's width allows -- 2-3 loads per 20-byte row instead of 3-4: every load instruction of a
    // gather costs the L1 one wavefront per lane
    constexpr bool HALFV = VEC && sizeof(T) == 2;
    float vals[HALFV ? 1 : G::ROWS_PER_THREAD][VEC ? 4 * G::NCHUNK : WIN];
    uint32_t raw[HALFV ? G::ROWS_PER_THREAD : 1][G::RAW];
#pragma unroll
    for (int j = 0; j < G::ROWS_PER_THREAD; ++j) {
      const int s = tid + j * kLookupThreads;
      const int sq = s / WIN, r = s - sq * WIN;      // sq = slot*16 + ql
      const int slot = sq >> 4, ql = sq & 15;
      const int lvl = lv0 + slot;
      const bool live = s < G::ROWS && lvl < p.num_levels && (q0 + ql) < p.nq;
      const int lvc = live ? lvl : 0;
      const int H = p.lh[lvc], W = p.lw[lvc];
      const int x0 = s_x0[sq & (kLevelsPerPass * kQueriesPerCta - 1)];
      const int y = s_y0[sq & (kLevelsPerPass * kQueriesPerCta - 1)] + r;
      const bool row_in = live && y >= 0 && y < H;
      const T* row = reinterpret_cast<const T*>(p.level[lvc]) + ((size_t)(q0 + ql) * H + (row_in ? y : 0)) * W;
      if constexpr (HALFV) {
        const uint16_t* hrow = reinterpret_cast<const uint16_t*>(row);
#pragma unroll
        for (int k = 0; k < G::RAW; ++k) raw[j][k] = 0u;
        if (p.wide[lvc]) {
          const int xa = x0 & ~7, sh = x0 - xa;
#pragma unroll
          for (int k = 0; k < G::NWIDE; ++k) {
            const int xc = xa + 8 * k;
            if (row_in && 8 * k < sh + WIN && xc >= 0 && xc < W) {
              const uint4 c = __ldg(reinterpret_cast<const uint4*>(hrow + xc));
              raw[j][4 * k] = c.x; raw[j][4 * k + 1] = c.y; raw[j][4 * k + 2] = c.z; raw[j][4 * k + 3] = c.w;
            }
          }
        } else {
          const int xa = x0 & ~3, sh = x0 - xa;
#pragma unroll
          for (int k = 0; k < G::NCHUNK; ++k) {
            const int xc = xa + 4 * k;
            if (row_in && 4 * k < sh + WIN && xc >= 0 && xc < W) {
              const uint2 c = __ldg(reinterpret_cast<const uint2*>(hrow + xc));
              raw[j][2 * k] = c.x; raw[j][2 * k + 1] = c.y;
            }
          }
        }
      } else if (VEC) {
        const int xa = x0 & ~3;
        const int sh = x0 - xa;
#pragma unroll
        for (int k = 0; k < G::NCHUNK; ++k) {
          const int xc = xa + 4 * k;
          const bool need = (4 * k < sh + WIN);  // the last chunk is only touched for the largest shifts
          float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
          if (row_in && need && xc >= 0 && xc < W) c = lk_ld4(row, xc);
          vals[j][4 * k] = c.x;
          vals[j][4 * k + 1] = c.y;
          vals[j][4 * k + 2] = c.z;
          vals[j][4 * k + 3] = c.w;
        }
      } else {
#pragma unroll
        for (int c = 0; c < WIN; ++c) {
          const int xx = x0 + c;
          vals[j][c] = (row_in && xx >= 0 && xx < W) ? lk_ld1(row, xx) : 0.f;
        }
      }
    }
#pragma unroll
    for (int j = 0; j < G::ROWS_PER_THREAD; ++j) {
      const int s = tid + j * kLookupThreads;
      if (s >= G::ROWS) continue;
      const int sq = s / WIN, r = s - sq * WIN;
      WT* d = win + sq * STRIDE + r * WIN;
      if constexpr (HALFV) {
        const int lvl = lv0 + (sq >> 4);
        const int sh = s_x0[sq] & (p.wide[lvl < p.num_levels ? lvl : 0] ? 7 : 3);
#pragma unroll
        for (int k = 0; k < 2 * G::RAW; ++k) {
          const int c = k - sh;
          if (c >= 0 && c < WIN) {
            const __half_raw hr{(unsigned short)((raw[j][k >> 1] >> (16 * (k & 1))) & 0xffffu)};
            d[c] = WT(__half(hr));
          }
        }
      } else if (VEC) {
        const int sh = s_x0[sq] & 3;  // x0 - (x0 & ~3)
#pragma unroll
        for (int k = 0; k < 4 * G::NCHUNK; ++k) {
          const int c = k - sh;
          if (c >= 0 && c < WIN) d[c] = WT(vals[j][k]);
        }
      } else {
#pragma unroll
        for (int c = 0; c < WIN; ++c) d[c] = WT(vals[j][c]);
      }
    }
    __syncthreads();

    // ---- compute phase
    if (NHWC && P % 3 == 0) {
      // channels-last output (B, H1, W1, L*P*P), P divisible by 3 (r = 4, r = 1): a warp takes one (level, query)
      // footprint at a time; lane t < P*P/3 forms the three consecutive channels (xo, y3..y3+2) from a 4 x 2 patch of
      // the footprint (8 loads, separable lerp), so a warp writes one contiguous run of P*P floats per footprint
      const int warp = tid >> 5, lane = tid & 31;
      constexpr int GROUPS = P * P / 3;
      const int xo = lane / (P / 3), y3 = (lane - xo * (P / 3)) * 3;
      const bool active = lane < GROUPS;
      for (int pair = warp; pair < kLevelsPerPass * kQueriesPerCta; pair += kLookupThreads / 32) {
        const int slot = pair >> 4, ql = pair & 15;
        const int lvl = lv0 + slot, q = q0 + ql;
        if (lvl >= p.num_levels || q >= p.nq || !active) continue;
        const WT* c = win + pair * STRIDE + y3 * WIN + xo;
        const float ax = s_ax[pair], ay = s_ay[pair];
        const float bx = 1.f - ax, by = 1.f - ay;
        float h[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) h[r] = fmaf(ax, float(c[r * WIN + 1]), bx * float(c[r * WIN]));
        const size_t oi = (size_t)q * p.out_stride + (size_t)lvl * P * P + xo * P + y3;
        if (p.out_f16) {
          __half* o = reinterpret_cast<__half*>(p.out) + oi;
#pragma unroll
          for (int j = 0; j < 3; ++j) o[j] = __float2half_rn(fmaf(ay, h[j + 1], by * h[j]));
        } else {
          float* o = p.out + oi;
#pragma unroll
          for (int j = 0; j < 3; ++j) o[j] = fmaf(ay, h[j + 1], by * h[j]);
        }
      }
    } else if (NHWC) {
      // channels-last output, any radius: lanes take consecutive taps, 4-tap bilinear weights
      const int warp = tid >> 5, lane = tid & 31;
      constexpr int ROUNDS = (P * P + 31) / 32;
      int tap_off[ROUNDS];  // footprint offset of the tap this lane forms in each round (fixed for the whole kernel)
#pragma unroll
      for (int k = 0; k < ROUNDS; ++k) {
        const int t = lane + 32 * k;
        const int xo = t / P, yo = t - xo * P;  // channel = (xoff+r)*P + (yoff+r)
        tap_off[k] = t < P * P ? yo * WIN + xo : -1;
      }
      for (int pair = warp; pair < kLevelsPerPass * kQueriesPerCta; pair += kLookupThreads / 32) {
        const int slot = pair >> 4, ql = pair & 15;
        const int lvl = lv0 + slot, q = q0 + ql;
        if (lvl >= p.num_levels || q >= p.nq) continue;  // warp-uniform
        const WT* my = win + pair * STRIDE;
        const float ax = s_ax[pair], ay = s_ay[pair];
        const float w00 = (1.f - ax) * (1.f - ay), w01 = ax * (1.f - ay), w10 = (1.f - ax) * ay, w11 = ax * ay;
        const size_t oi = (size_t)q * p.out_stride + (size_t)lvl * P * P + lane;
#pragma unroll
        for (int k = 0; k < ROUNDS; ++k) {
          if (tap_off[k] >= 0) {
            const WT* c = my + tap_off[k];
            const float r = float(c[0]) * w00 + float(c[1]) * w01 + float(c[WIN]) * w10 + float(c[WIN + 1]) * w11;
            if (p.out_f16)
              reinterpret_cast<__half*>(p.out)[oi + 32 * k] = __float2half_rn(r);
            else
              p.out[oi + 32 * k] = r;
          }
        }
      }
    } else {
      // warp = level slot, lane = (half, query); taps stored coalesced across queries (channel-first output)
      const int slot = tid >> 5, lane = tid & 31;
      const int ql = lane & 15, half = lane >> 4;
      const int lvl = lv0 + slot;
      const int q = q0 + ql;
      if (lvl < p.num_levels && q < p.nq) {
        const int b = q / p.hw1, pix = q - b * p.hw1;
        const WT* my = win + (slot * kQueriesPerCta + ql) * STRIDE;
        const float ax = s_ax[slot * kQueriesPerCta + ql], ay = s_ay[slot * kQueriesPerCta + ql];
        const float bx = 1.f - ax, by = 1.f - ay;
        float* o = p.out + ((size_t)b * channels + (size_t)lvl * P * P) * p.hw1 + pix;
        constexpr int HALF = (P + 1) / 2;              // output rows [0, HALF) and [HALF, P)
        const int r_begin = half ? HALF : 0;
        const int r_end = half ? P : HALF;             // window rows r_begin .. r_end (inclusive)
        float prev[P], cur[P];
#pragma unroll
        for (int rr = 0; rr <= HALF; ++rr) {
          const int r = r_begin + rr;
          if (r <= r_end) {
            float v[WIN];
#pragma unroll
            for (int c = 0; c < WIN; ++c) v[c] = float(my[r * WIN + c]);
#pragma unroll
            for (int xo = 0; xo < P; ++xo) cur[xo] = fmaf(ax, v[xo + 1], bx * v[xo]);
            if (rr > 0) {
#pragma unroll
              for (int xo = 0; xo < P; ++xo)  // channel = (xoff+r)*P + (yoff+r): x walks slowest (corr.py:37-43)
                o[(size_t)(xo * P + (r - 1)) * p.hw1] = fmaf(ay, cur[xo], by * prev[xo]);
            }
#pragma unroll
            for (int xo = 0; xo < P; ++xo) prev[xo] = cur[xo];
          }
        }
      }
    }
    __syncthreads();  // the next pass (more than 4 levels) reuses the shared buffers
  }
}

template <int R>
static int launch_lookup(const LookupParams& p, bool vec, bool nhwc, bool f16_levels, cudaStream_t stream) {
  const dim3 grid(ceil_div(p.nq, kQueriesPerCta)), block(kLookupThreads);
  cudaError_t e;
  if (f16_levels && vec && nhwc)
    e = launch_pdl(corr_lookup_kernel<R, true, true, __half>, grid, block, 0, stream, p);
  else if (f16_levels && vec)
    e = launch_pdl(corr_lookup_kernel<R, true, false, __half>, grid, block, 0, stream, p);
  else if (f16_levels && nhwc)
    e = launch_pdl(corr_lookup_kernel<R, false, true, __half>, grid, block, 0, stream, p);
  else if (f16_levels)
    e = launch_pdl(corr_lookup_kernel<R, false, false, __half>, grid, block, 0, stream, p);
  else if (vec && nhwc)
    e = launch_pdl(corr_lookup_kernel<R, true, true>, grid, block, 0, stream, p);
  else if (vec)
    e = launch_pdl(corr_lookup_kernel<R, true, false>, grid, block, 0, stream, p);
  else if (nhwc)
    e = launch_pdl(corr_lookup_kernel<R, false, true>, grid, block, 0, stream, p);
  else
    e = launch_pdl(corr_lookup_kernel<R, false, false>, grid, block, 0, stream, p);
  if (e != cudaSuccess) return fail(WOFT_ERR_CUDA, "corr_lookup_kernel launch failed: %s", cudaGetErrorString(e));
  return launch_status("corr_lookup_kernel");
}

int corr_lookup_tma(const float* const*, int, const float*, int, int, int, int, int, void*, int, int, cudaStream_t);

}  // namespace woft

namespace woft {
struct LookupGather {   // the optional fused gather of woft_corr_lookup_gather
  const float *taps, *bias, *coords0;
  float *flow_nchw, *flow_nhwc;
  void *f16_a, *f16_b;
  int f16_stride, f16_offset;
};
static int corr_lookup_impl(const float* const* levels, int num_levels, const float* coords, int batch, int h1, int w1,
                            int h2, int w2, int radius, float* out, int flags, const LookupGather* g, void* stream);
}  // namespace woft

extern "C" int woft_corr_lookup(const float* const* levels, int num_levels, const float* coords, int batch, int h1,
                                int w1, int h2, int w2, int radius, float* out, int flags, void* stream) {
  return woft::corr_lookup_impl(levels, num_levels, coords, batch, h1, w1, h2, w2, radius, out, flags, nullptr, stream);
}

extern "C" int woft_corr_lookup_gather(const float* const* levels, int num_levels, float* coords, int batch, int h1, int w1,
                                       int h2, int w2, int radius, float* out, int flags, const float* taps,
                                       const float* bias, const float* coords0, float* flow_nchw, float* flow_nhwc,
                                       void* flow_f16_a, void* flow_f16_b, int f16_stride, int f16_offset, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(taps && bias && coords0, "woft_corr_lookup_gather: null pointer");
  WOFT_REQUIRE(!(flags & 64), "woft_corr_lookup_gather: not with the TMA variant (flag 64)");
  WOFT_REQUIRE((!flow_f16_a && !flow_f16_b) || (f16_stride % 2 == 0 && f16_offset % 2 == 0 &&
                                                  ((reinterpret_cast<uintptr_t>(flow_f16_a) | reinterpret_cast<uintptr_t>(flow_f16_b)) & 3) == 0),
               "woft_corr_lookup_gather: the fp16 flow copies need even strides / offsets");
  const LookupGather g{taps, bias, coords0, flow_nchw, flow_nhwc, flow_f16_a, flow_f16_b, f16_stride, f16_offset};
  return corr_lookup_impl(levels, num_levels, coords, batch, h1, w1, h2, w2, radius, out, flags, &g, stream);
}

static int woft::corr_lookup_impl(const float* const* levels, int num_levels, const float* coords, int batch, int h1, int w1,
                                  int h2, int w2, int radius, float* out, int flags, const LookupGather* g, void* stream) {
  WOFT_REQUIRE(levels && coords && out, "woft_corr_lookup: null pointer");
  WOFT_REQUIRE(num_levels >= 1 && num_levels <= kMaxLevels, "woft_corr_lookup: num_levels %d not in [1,%d]",
               num_levels, kMaxLevels);
  WOFT_REQUIRE(batch > 0 && h1 > 0 && w1 > 0 && h2 > 0 && w2 > 0, "woft_corr_lookup: non-positive extent");
  WOFT_REQUIRE(radius >= 1 && radius <= 4, "woft_corr_lookup: radius %d not in [1,4]", radius);
  WOFT_REQUIRE((long long)batch * h1 * w1 < (1ll << 31), "woft_corr_lookup: too many query pixels");
  LookupParams p{};
  int h = h2, w = w2;
  bool vec = (flags & 1) == 0;  // flags bit 0 forces the scalar load path (testing)
  const bool nhwc = (flags & 8) != 0;  // flags bit 3: channels-last output (B, H1, W1, L*P*P)
  const bool f16_levels = (flags & 32) != 0;   // the levels hold fp16 (WOFT_CORR_STORE_F16)
  WOFT_REQUIRE(!(f16_levels && (flags & 64)), "woft_corr_lookup: the TMA variant (flag 64) reads fp32 levels");
  for (int l = 0; l < num_levels; ++l) {
    // the reference divides by (W-1), (H-1) and returns NaN for 1-pixel maps (utils.py:63-64)
    WOFT_REQUIRE(h >= 2 && w >= 2, "woft_corr_lookup: level %d is %dx%d; every level must be at least 2x2", l, h, w);
    WOFT_REQUIRE(levels[l], "woft_corr_lookup: level %d pointer is null", l);
    p.level[l] = levels[l];
    p.lh[l] = h;
    p.lw[l] = w;
    vec = vec && (w % 4 == 0) && ((reinterpret_cast<uintptr_t>(levels[l]) & (f16_levels ? 7 : 15)) == 0);
    p.wide[l] = (f16_levels && w % 8 == 0 && (reinterpret_cast<uintptr_t>(levels[l]) & 15) == 0 && !(flags & 2)) ? 1 : 0;
    h /= 2;
    w /= 2;
  }
  p.num_levels = num_levels;
  p.coords = coords;
  p.out = out;
  {
    const int P = 2 * radius + 1, channels = num_levels * P * P;
    p.out_f16 = (flags & 16) != 0;
    WOFT_REQUIRE(!p.out_f16 || nhwc, "woft_corr_lookup: the fp16 output (flag 16) exists for the channels-last layout (flag 8)");
    p.out_stride = p.out_f16 ? (channels + 7) / 8 * 8 : channels;
  }
  p.nq = batch * h1 * w1;
  p.hw1 = h1 * w1;
  if (g) {
    p.g_taps = g->taps;
    p.g_bias = g->bias;
    p.g_coords = const_cast<float*>(coords);   // woft_corr_lookup_gather takes it mutable
    p.g_coords0 = g->coords0;
    p.g_flow_nchw = g->flow_nchw;
    p.g_flow_nhwc = g->flow_nhwc;
    p.g_h1 = static_cast<__half*>(g->f16_a);
    p.g_h2 = static_cast<__half*>(g->f16_b);
    p.g_stride = g->f16_stride;
    p.g_off = g->f16_offset;
    p.g_H = h1;
    p.g_W = w1;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (nhwc && radius == 4 && vec && (flags & 64)) {
    // flag 64: channels-last output, radius 4, 16-byte aligned level rows -- the footprints come in by TMA
    // (lookup_tma.cu).  Measured slower than the load/store-unit kernel below (20.2 vs 18.4 us per launch at 1280x720),
    // so it is opt-in
    const int rc = corr_lookup_tma(levels, num_levels, coords, batch, h1, w1, h2, w2, out, p.out_f16, p.out_stride, s);
    if (rc != WOFT_ERR_UNSUPPORTED) return rc;
  }
  switch (radius) {
    case 1: return launch_lookup<1>(p, vec, nhwc, f16_levels, s);
    case 2: return launch_lookup<2>(p, vec, nhwc, f16_levels, s);
    case 3: return launch_lookup<3>(p, vec, nhwc, f16_levels, s);
    default: return launch_lookup<4>(p, vec, nhwc, f16_levels, s);
  }
}
