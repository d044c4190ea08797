// K3: radius-r bilinear lookup into the 4-level correlation pyramid.
//
// Replaces CorrBlock.__call__ + bilinear_sampler of the reference
// (pytracking/external/RAFT/raft_core/corr.py:29-59, raft_core/utils/utils.py:59-73), which runs
// linspace/meshgrid + one grid_sample per level + cat + permute, 13 times per flow pass.
//
// Memory-bound gather.  Per query pixel and level the (2r+1)^2 bilinear taps share one fractional offset,
// so the kernel loads the (2r+2)^2 integer footprint once and forms all taps from it:
//   - a CTA owns 32 consecutive query pixels; warp w serves level w (loops if there are more levels);
//   - load phase: the warp walks the 32 footprints row segment by row segment, 32/(2r+2) segments per
//     instruction, so one load instruction touches 3 rows of 40 B (r=4) instead of 32 scattered lines;
//     footprints land in shared memory at an odd per-query stride (bank-conflict free for lane = query);
//   - compute phase: lane = query; horizontal then vertical lerp in registers, row by row, and every tap
//     is stored with lane = query, i.e. one coalesced 128-byte line per store instruction in the
//     reference's channel-first output layout.
// Algorithmic bytes per query pixel (r=4, 4 levels): 4*10*10*4 B read + 8 B coords + 324*4 B written = 2904 B.
#include "common.cuh"

namespace woft {

struct LookupParams {
  const float* level[kMaxLevels];
  int lh[kMaxLevels];
  int lw[kMaxLevels];
  int num_levels;
  const float* coords;  // (B, 2, H1, W1)
  float* out;           // (B, L*P*P, H1, W1)
  int nq;               // B*H1*W1
  int hw1;              // H1*W1
};

constexpr int kLookupWarps = 4;
constexpr int kQueriesPerCta = 32;

template <int R>
struct LookupGeom {
  static constexpr int P = 2 * R + 1;
  static constexpr int WIN = P + 1;
  static constexpr int STRIDE = WIN * WIN + 1;  // odd => lanes hit distinct banks
  static constexpr int SEG_PER_INSTR = 32 / WIN;
  static constexpr int NSEG = kQueriesPerCta * WIN;
  static constexpr size_t SMEM = size_t(kLookupWarps) * kQueriesPerCta * STRIDE * sizeof(float);
};

__device__ __forceinline__ void cp_async_4(float* smem_dst, const float* gsrc, bool pred) {
  unsigned dst = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
  int src_bytes = pred ? 4 : 0;  // 0 => hardware zero-fill, the source is not read
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(dst), "l"(gsrc), "r"(src_bytes) : "memory");
}

template <int R, bool CP_ASYNC>
__global__ void __launch_bounds__(kLookupWarps * 32) corr_lookup_kernel(const __grid_constant__ LookupParams p) {
  using G = LookupGeom<R>;
  constexpr int P = G::P, WIN = G::WIN, STRIDE = G::STRIDE;
  extern __shared__ float smem[];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * kQueriesPerCta;
  const int q = q0 + lane;
  const bool valid = q < p.nq;
  const int b = valid ? q / p.hw1 : 0;
  const int pix = valid ? q - b * p.hw1 : 0;
  float cx = 0.f, cy = 0.f;
  if (valid) {
    cx = __ldg(p.coords + (size_t)(2 * b) * p.hw1 + pix);
    cy = __ldg(p.coords + (size_t)(2 * b + 1) * p.hw1 + pix);
  }
  float* win = smem + warp * kQueriesPerCta * STRIDE;
  const int sub = lane / WIN;        // which of the SEG_PER_INSTR row segments this lane helps load
  const int col = lane - sub * WIN;  // column within the segment
  const bool loader = sub < G::SEG_PER_INSTR;
  const int channels = p.num_levels * P * P;

  for (int lvl = warp; lvl < p.num_levels; lvl += kLookupWarps) {
    const int H = p.lh[lvl], W = p.lw[lvl];
    const size_t plane = (size_t)H * W;
    const float* __restrict__ base = p.level[lvl];
    const float inv = 1.0f / float(1 << lvl);
    const float x = cx * inv, y = cy * inv;  // centroid_lvl = coords / 2**i (corr.py:41)
    float fx = floorf(x), fy = floorf(y);
    const float ax = x - fx, ay = y - fy;
    fx = fminf(fmaxf(fx, -1048576.f), 1048576.f);
    fy = fminf(fmaxf(fy, -1048576.f), 1048576.f);
    const int x0 = int(fx) - R, y0 = int(fy) - R;

    // ---- load phase: 32 footprints of WIN x WIN, SEG_PER_INSTR row segments per instruction
    constexpr int UNROLL = 8;
    constexpr int STEP = G::SEG_PER_INSTR * UNROLL;
#pragma unroll 1
    for (int s0 = 0; s0 < G::NSEG; s0 += STEP) {
      float v[UNROLL];
      int dst[UNROLL];
#pragma unroll
      for (int j = 0; j < UNROLL; ++j) {
        const int s = s0 + j * G::SEG_PER_INSTR + sub;
        const bool ok = loader && s < G::NSEG;
        const int ql = ok ? s / WIN : 0;
        const int r = s - ql * WIN;
        const int xx = __shfl_sync(0xffffffffu, x0, ql) + col;
        const int yy = __shfl_sync(0xffffffffu, y0, ql) + r;
        const bool inb = ok && (q0 + ql < p.nq) && xx >= 0 && xx < W && yy >= 0 && yy < H;
        const float* src = base + (size_t)(q0 + ql) * plane + (size_t)(inb ? yy * W + xx : 0);
        dst[j] = ok ? ql * STRIDE + r * WIN + col : -1;
        if (CP_ASYNC) {
          if (ok) cp_async_4(win + dst[j], inb ? src : base, inb);
        } else {
          v[j] = inb ? __ldg(src) : 0.f;
        }
      }
      if (!CP_ASYNC) {
#pragma unroll
        for (int j = 0; j < UNROLL; ++j)
          if (dst[j] >= 0) win[dst[j]] = v[j];
      }
    }
    if (CP_ASYNC) asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
    __syncwarp();

    // ---- compute phase: lane = query; taps formed row by row, stored coalesced across queries
    const float* my = win + lane * STRIDE;
    float* o = p.out + ((size_t)b * channels + (size_t)lvl * P * P) * p.hw1 + pix;
    const float bx = 1.f - ax, by = 1.f - ay;
    float prev[P], cur[P];
#pragma unroll
    for (int r = 0; r < WIN; ++r) {
      float v[WIN];
#pragma unroll
      for (int c = 0; c < WIN; ++c) v[c] = my[r * WIN + c];
#pragma unroll
      for (int xo = 0; xo < P; ++xo) cur[xo] = fmaf(ax, v[xo + 1], bx * v[xo]);
      if (r > 0 && valid) {
#pragma unroll
        for (int xo = 0; xo < P; ++xo)  // channel = (xoff+r)*P + (yoff+r): x walks slowest (corr.py:37-43)
          o[(size_t)(xo * P + (r - 1)) * p.hw1] = fmaf(ay, cur[xo], by * prev[xo]);
      }
#pragma unroll
      for (int xo = 0; xo < P; ++xo) prev[xo] = cur[xo];
    }
    __syncwarp();
  }
}

template <int R>
static int launch_lookup(const LookupParams& p, int flags, cudaStream_t stream) {
  using G = LookupGeom<R>;
  const int grid = ceil_div(p.nq, kQueriesPerCta);
  if (flags & 1) {
    WOFT_CUDA_CHECK(cudaFuncSetAttribute(corr_lookup_kernel<R, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)G::SMEM));
    corr_lookup_kernel<R, true><<<grid, kLookupWarps * 32, G::SMEM, stream>>>(p);
  } else {
    WOFT_CUDA_CHECK(cudaFuncSetAttribute(corr_lookup_kernel<R, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)G::SMEM));
    corr_lookup_kernel<R, false><<<grid, kLookupWarps * 32, G::SMEM, stream>>>(p);
  }
  return launch_status("corr_lookup_kernel");
}

}  // namespace woft

extern "C" int woft_corr_lookup(const float* const* levels, int num_levels, const float* coords, int batch, int h1,
                                int w1, int h2, int w2, int radius, float* out, int flags, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(levels && coords && out, "woft_corr_lookup: null pointer");
  WOFT_REQUIRE(num_levels >= 1 && num_levels <= kMaxLevels, "woft_corr_lookup: num_levels %d not in [1,%d]",
               num_levels, kMaxLevels);
  WOFT_REQUIRE(batch > 0 && h1 > 0 && w1 > 0 && h2 > 0 && w2 > 0, "woft_corr_lookup: non-positive extent");
  WOFT_REQUIRE(radius >= 1 && radius <= 4, "woft_corr_lookup: radius %d not in [1,4]", radius);
  WOFT_REQUIRE((long long)batch * h1 * w1 < (1ll << 31), "woft_corr_lookup: too many query pixels");
  LookupParams p{};
  int h = h2, w = w2;
  for (int l = 0; l < num_levels; ++l) {
    // the reference divides by (W-1), (H-1) and returns NaN for 1-pixel maps (utils.py:63-64)
    WOFT_REQUIRE(h >= 2 && w >= 2, "woft_corr_lookup: level %d is %dx%d; every level must be at least 2x2", l, h, w);
    WOFT_REQUIRE(levels[l], "woft_corr_lookup: level %d pointer is null", l);
    p.level[l] = levels[l];
    p.lh[l] = h;
    p.lw[l] = w;
    h /= 2;
    w /= 2;
  }
  p.num_levels = num_levels;
  p.coords = coords;
  p.out = out;
  p.nq = batch * h1 * w1;
  p.hw1 = h1 * w1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (radius) {
    case 1: return launch_lookup<1>(p, flags, s);
    case 2: return launch_lookup<2>(p, flags, s);
    case 3: return launch_lookup<3>(p, flags, s);
    default: return launch_lookup<4>(p, flags, s);
  }
}
