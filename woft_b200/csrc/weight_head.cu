// First layer of the weight head (raft_core/weighted_raft.py:319-339, 347-384): for every query pixel a 9 x 9 "image"
// with 5 channels -- the 4 pyramid levels of its correlation window (the lookup's channels read as (hp, wp, level),
// :267-272) and the mean response of its correlation plane broadcast over the patch (:358-361) -- goes through
// conv3x3(5 -> 128, pad 1) + ReLU.  That is 14 400 tiny images at 1280x720; cuDNN runs it as an fp32 direct
// convolution (206 us) after three ATen kernels have assembled the 5-channel input.  Here the assembly is the staging
// step of an implicit GEMM on the tensor cores (mma.sync m16n8k8, TF32 operands rounded to nearest, fp32 accumulation):
//   M = 81 pixels per image, images back to back; N = 128; K = 9 taps x 5 channels = 45, padded to 48.
// A CTA takes 64 consecutive (image, pixel) rows per step, which touch at most two images; those are staged as
// zero-bordered 11 x 11 x 5 tiles and the A fragments gathered through a k -> offset table.  The 48 x 128 weight matrix
// never leaves the registers: a warp owns 32 rows x 32 output channels and keeps its 6 x 4 B fragments for the whole
// kernel.  Output (n, 9, 9, 128) channels-last, 597 MB at 1280x720 -- the kernel is bound by writing it.
#include <cuda_fp16.h>

#include "common.cuh"
#include "ptx.cuh"

namespace woft {

constexpr int kWhP = 9, kWhPix = kWhP * kWhP, kWhL = 4, kWhCin = kWhL + 1, kWhCout = 128;
constexpr int kWhK = 9 * kWhCin, kWhKP = 48, kWhRows = 64;
constexpr int kWhTile = 11 * 11 * kWhCin;   // 605 floats per zero-bordered image

__device__ __forceinline__ float wh_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

__device__ __forceinline__ void wh_mma(float (&d)[4], const float (&a)[4], float b0, float b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(__float_as_uint(a[0])), "r"(__float_as_uint(a[1])), "r"(__float_as_uint(a[2])), "r"(__float_as_uint(a[3])),
        "r"(__float_as_uint(b0)), "r"(__float_as_uint(b1)));
}

// STACKED: the output is fp16 and every image takes 10 rows of 9 pixels, the tenth all zero -- the layout in which
// woft_conv_tc runs the next two layers (the zero row is the vertical padding between stacked images).
template <bool STACKED>
__global__ void __launch_bounds__(256, 2) weight_head_conv1_kernel(const float* __restrict__ window,   // (n, 81 * 4)
                                                                   const float* __restrict__ mean_resp, // (n)
                                                                   const float* __restrict__ weight,    // (128, 5, 3, 3)
                                                                   const float* __restrict__ bias,      // (128)
                                                                   void* __restrict__ out_any,          // (n * 81, 128)
                                                                   int n_img, int total_tiles) {
  __shared__ float tiles[2][2 * kWhTile];   // double buffered: the next step's images load while this one multiplies
  __shared__ int koff[kWhKP];
  // STACKED: a warp's 32 rows x 32 fp16 channels pass through 2 KB of staging so that a lane stores 16 contiguous bytes
  // (the accumulator fragment gives a lane 4 bytes in each of four 16-byte groups: 16-byte stores of which a quarter is
  // written made this kernel 193 us for 331 MB)
  __shared__ __align__(16) uint32_t stg[STACKED ? 8 : 1][STACKED ? 32 * 16 : 1];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int rg = warp >> 2, cq = warp & 3;   // 2 row groups of 32 rows x 4 channel quarters of 32
  for (int i = tid; i < 4 * kWhTile; i += 256) (&tiles[0][0])[i] = 0.f;   // borders stay zero for the whole kernel
  if (tid < kWhKP) {
    const int tap = tid / kWhCin, ci = tid - tap * kWhCin, ky = tap / 3, kx = tap - ky * 3;
    koff[tid] = tid < kWhK ? (ky * 11 + kx) * kWhCin + ci : 0;
  }
  // B fragments of this warp: b0 = W[k = ks*8 + t][n], b1 = W[k + 4][n], n = cq*32 + nt*8 + g; k = tap*5 + ci
  float bf[6][4][2];
#pragma unroll
  for (int ks = 0; ks < 6; ++ks)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int k = ks * 8 + t + 4 * h;
      const int tap = k / kWhCin, ci = k - tap * kWhCin;
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        const int n = cq * 32 + nt * 8 + g;
        bf[ks][nt][h] = k < kWhK ? wh_tf32(__ldg(weight + ((size_t)n * kWhCin + ci) * 9 + tap)) : 0.f;
      }
    }
  float bv[4][2];
#pragma unroll
  for (int nt = 0; nt < 4; ++nt) {
    bv[nt][0] = __ldg(bias + cq * 32 + nt * 8 + 2 * t);
    bv[nt][1] = __ldg(bias + cq * 32 + nt * 8 + 2 * t + 1);
  }
  const int total_rows = n_img * kWhPix;   // < 2^31, checked by the entry point

  // a step's input: 2 images x 81 pixels x 5 channels = 810 values, at most 4 per thread, held in registers between
  // the load (issued one step ahead) and the store into the step's buffer.  Which value a thread handles does not
  // depend on the step: source and destination offsets are worked out once.
  constexpr int kPer = (2 * kWhPix * kWhCin + 255) / 256;
  int src_off[kPer], dst_off[kPer], slot[kPer];   // src_off < 0: the mean response of image `slot`
#pragma unroll
  for (int u = 0; u < kPer; ++u) {
    const int i = tid + u * 256;
    const int sl = i / (kWhPix * kWhCin), r = i - sl * kWhPix * kWhCin;
    const int p = r / kWhCin, ci = r - p * kWhCin;
    const int y = p / kWhP, x = p - y * kWhP;
    slot[u] = i < 2 * kWhPix * kWhCin ? sl : -1;
    src_off[u] = ci < kWhL ? sl * (kWhPix * kWhL) + p * kWhL + ci : -1;
    dst_off[u] = sl * kWhTile + ((y + 1) * 11 + x + 1) * kWhCin + ci;
  }
  float pv[kPer];
  auto prefetch = [&](int tile) {
    const int img0 = (tile * kWhRows) / kWhPix;
#pragma unroll
    for (int u = 0; u < kPer; ++u) {
      pv[u] = 0.f;
      if (slot[u] >= 0 && img0 + slot[u] < n_img)
        pv[u] = src_off[u] >= 0 ? __ldg(window + (size_t)img0 * (kWhPix * kWhL) + src_off[u])
                                : __ldg(mean_resp + img0 + slot[u]);
    }
  };
  int buf = 0;
  if ((int)blockIdx.x < total_tiles) prefetch(blockIdx.x);
  __syncthreads();   // zeroed borders and the offset table are in place before any step writes its images
  for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, buf ^= 1) {
    const int row0 = tile * kWhRows;
    const int img0 = row0 / kWhPix;
    float* tl = tiles[buf];
#pragma unroll
    for (int u = 0; u < kPer; ++u)
      if (slot[u] >= 0) tl[dst_off[u]] = wh_tf32(pv[u]);
    __syncthreads();   // this buffer was last read two steps ago, before the previous step's barrier
    if (tile + (int)gridDim.x < total_tiles) prefetch(tile + gridDim.x);

    int base[2][2], grow[2][2];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        int r = row0 + rg * 32 + mt * 16 + g + 8 * h;
        grow[mt][h] = r;
        if (r >= total_rows) r = total_rows - 1;   // clamp for the gather, never stored
        const int img = r / kWhPix, p = r - img * kWhPix;
        const int y = p / kWhP, x = p - y * kWhP;
        base[mt][h] = (img - img0) * kWhTile + (y * 11 + x) * kWhCin;
      }
    float acc[2][4][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        acc[mt][nt][0] = bv[nt][0]; acc[mt][nt][1] = bv[nt][1];
        acc[mt][nt][2] = bv[nt][0]; acc[mt][nt][3] = bv[nt][1];
      }
#pragma unroll
    for (int ks = 0; ks < 6; ++ks) {
      const int o0 = koff[ks * 8 + t], o1 = koff[ks * 8 + t + 4];
      float a[2][4];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        a[mt][0] = tl[base[mt][0] + o0];
        a[mt][1] = tl[base[mt][1] + o0];
        a[mt][2] = tl[base[mt][0] + o1];
        a[mt][3] = tl[base[mt][1] + o1];
      }
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        wh_mma(acc[0][nt], a[0], bf[ks][nt][0], bf[ks][nt][1]);
        wh_mma(acc[1][nt], a[1], bf[ks][nt][0], bf[ks][nt][1]);
      }
    }
    if (STACKED) {
      uint32_t* sw = stg[warp];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int r = mt * 16 + h * 8 + g;   // row of the warp's 32; 16 words per row, 4-word pieces XOR-ed by the row
#pragma unroll
          for (int nt = 0; nt < 4; ++nt) {
            const __half2 v = __floats2half2_rn(fmaxf(acc[mt][nt][2 * h], 0.f), fmaxf(acc[mt][nt][2 * h + 1], 0.f));
            sw[r * 16 + (((nt ^ ((r >> 1) & 3)) << 2) | t)] = *reinterpret_cast<const uint32_t*>(&v);
          }
        }
      __syncwarp();
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int r = (lane >> 2) + 8 * i, q = lane & 3;
        const int gr = row0 + rg * 32 + r;
        if (gr < total_rows) {
          const uint4 v = *reinterpret_cast<const uint4*>(sw + r * 16 + ((q ^ ((r >> 1) & 3)) << 2));
          const int img = gr / kWhPix, pp = gr - img * kWhPix;
          *reinterpret_cast<uint4*>(reinterpret_cast<__half*>(out_any) + ((size_t)img * (kWhPix + kWhP) + pp) * kWhCout +
                                    cq * 32 + q * 8) = v;
        }
      }
      __syncwarp();
    }
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        if (grow[mt][h] >= total_rows) continue;
        if (STACKED) {
        } else {
          float* o = reinterpret_cast<float*>(out_any) + (size_t)grow[mt][h] * kWhCout + cq * 32 + 2 * t;
#pragma unroll
          for (int nt = 0; nt < 4; ++nt)
            *reinterpret_cast<float2*>(o + nt * 8) =
                make_float2(fmaxf(acc[mt][nt][2 * h], 0.f), fmaxf(acc[mt][nt][2 * h + 1], 0.f));
        }
      }
    if (STACKED) {
      // the image whose first pixel lies in this tile (at most one: 81 rows per image, 64 per tile) gets its zero row
      const int img_s = (row0 + kWhPix - 1) / kWhPix;
      if (img_s < n_img && img_s * kWhPix < row0 + kWhRows) {
        uint32_t* z = reinterpret_cast<uint32_t*>(reinterpret_cast<__half*>(out_any) +
                                                  ((size_t)img_s * (kWhPix + kWhP) + kWhPix) * kWhCout);
        for (int i = tid; i < kWhP * kWhCout / 2; i += 256) z[i] = 0u;
      }
    }
  }
}

// ---- the same layer on tcgen05 (stacked fp16 output) ------------------------------------------------------------------
// A tile is 128 consecutive output pixels (TMEM lanes) of the (image, y, x) sequence; K = 45 padded to 64 fp16 = ONE
// 128-byte K row, so the whole tile is four MMAs (128 x 128 x 16).  The A operand has no tensor to be loaded from -- it is
// the im2col of 9x9x5 images, the lookup channels re-read as (level, y, x) -- so 256 threads build it in shared memory
// (two per pixel, each gathering half of the 45 values from the zero-bordered images and writing four swizzled 16-byte
// pieces); the filters sit in shared memory for the whole kernel.  Against the mma.sync kernel above this reads every
// input value once per pixel instead of once per (pixel, channel quarter) and leaves the legacy tensor pipe.  Three CTAs
// per SM (128 TMEM columns and ~57 KB of shared memory each) overlap their gather, MMA and store phases.
constexpr int kTcRows = 128, kTcImgs = 3, kTcThreads = 256;
constexpr int kTcSmem = 1024 + 16384 + 16384 + 16384 + kTcImgs * kWhTile * 4 + 4 * kWhCout + 64;

__device__ int g_wh_abort = 0;

__device__ __forceinline__ uint64_t wh_desc(uint32_t saddr) {  // K-major, 128-byte swizzle, 8-row groups 1024 B apart
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

// offset of im2col column k in a zero-bordered 11 x 11 x 5 image, relative to the pixel's top-left neighbour
__host__ __device__ constexpr int wh_koff(int k) { return ((k / kWhCin) / 3 * 11 + (k / kWhCin) % 3) * kWhCin + k % kWhCin; }

template <int K0>
__device__ __forceinline__ void wh_build_half(const float* px, uint8_t* row_base, int r) {
  // 32 columns K0 .. K0 + 31 of one pixel's im2col row -> four swizzled 16-byte pieces
#pragma unroll
  for (int jj = 0; jj < 4; ++jj) {
    uint32_t w[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int k = K0 + jj * 8 + 2 * e;
      const float a = (k < kWhK && px) ? px[wh_koff(k < kWhK ? k : 0)] : 0.f;
      const float b = (k + 1 < kWhK && px) ? px[wh_koff(k + 1 < kWhK ? k + 1 : 0)] : 0.f;
      const __half2 h = __floats2half2_rn(a, b);
      w[e] = *reinterpret_cast<const uint32_t*>(&h);
    }
    const int j = K0 / 8 + jj;
    *reinterpret_cast<uint4*>(row_base + ((j ^ (r & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

__global__ void __launch_bounds__(kTcThreads, 3) weight_head_conv1_tc_kernel(const float* __restrict__ window,
                                                                             const float* __restrict__ mean_resp,
                                                                             const float* __restrict__ weight,
                                                                             const float* __restrict__ bias,
                                                                             __half* __restrict__ out, int n_img,
                                                                             int total_tiles) {
  extern __shared__ uint8_t wh_smem_raw[];
  uint8_t* smem = wh_smem_raw + ((1024u - (ptx::smem_u32(wh_smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;                       // 128 pixels x 128 bytes (K = 64 fp16), 128-byte swizzle
  uint8_t* sB = sA + 16384;                 // 128 output channels x 128 bytes
  uint8_t* s_stage = sB + 16384;            // 8 warps x 2 KB
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_stage + 16384);   // 8-byte aligned: everything before is a multiple of 1 KB
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bar + 1);
  float* s_bias = reinterpret_cast<float*>(bar + 2);
  float* tiles = s_bias + kWhCout;                                // kTcImgs zero-bordered images
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    ptx::mbar_init(bar, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 0) ptx::tmem_alloc(tmem_ptr, 128);
  for (int i = tid; i < kTcImgs * kWhTile; i += kTcThreads) tiles[i] = 0.f;   // borders stay zero for the whole kernel
  if (tid < kWhCout) s_bias[tid] = __ldg(bias + tid);
  // filters: B[n][k] = W[n][ci][tap], k = tap * 5 + ci, K-major rows of 64 fp16 (45 real), swizzled like A
  for (int i = tid; i < kWhCout * 8; i += kTcThreads) {
    const int n = i >> 3, j = i & 7;
    uint32_t w[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float v2[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int k = j * 8 + 2 * e + u;
        v2[u] = k < kWhK ? __ldg(weight + ((size_t)n * kWhCin + k % kWhCin) * 9 + k / kWhCin) : 0.f;
      }
      const __half2 h = __floats2half2_rn(v2[0], v2[1]);
      w[e] = *reinterpret_cast<const uint32_t*>(&h);
    }
    *reinterpret_cast<uint4*>(sB + n * 128 + ((j ^ (n & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr;
  const uint32_t idesc = (1u << 4) | (uint32_t(kWhCout >> 3) << 17) | (uint32_t(128 >> 4) << 24);   // f16 x f16 -> f32
  const uint64_t da = wh_desc(ptx::smem_u32(sA)), db = wh_desc(ptx::smem_u32(sB));
  const long long total_rows = (long long)n_img * kWhPix;
  uint32_t phase = 0;

  for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
    const long long row0 = (long long)tile * kTcRows;
    const int img0 = int(row0 / kWhPix);
    // ---- the images under this tile: 3 x 81 x 5 values, loads first, then the stores into the bordered images
    {
      constexpr int kPer = (kTcImgs * kWhPix * kWhCin + kTcThreads - 1) / kTcThreads;
      float pv[kPer];
#pragma unroll
      for (int u = 0; u < kPer; ++u) {
        const int i = tid + u * kTcThreads;
        const int sl = i / (kWhPix * kWhCin), r = i - sl * kWhPix * kWhCin;
        const int pp = r / kWhCin, ci = r - pp * kWhCin;
        pv[u] = 0.f;
        if (i < kTcImgs * kWhPix * kWhCin && img0 + sl < n_img)
          pv[u] = ci < kWhL ? __ldg(window + ((size_t)(img0 + sl) * kWhPix + pp) * kWhL + ci) : __ldg(mean_resp + img0 + sl);
      }
#pragma unroll
      for (int u = 0; u < kPer; ++u) {
        const int i = tid + u * kTcThreads;
        if (i >= kTcImgs * kWhPix * kWhCin) continue;
        const int sl = i / (kWhPix * kWhCin), r = i - sl * kWhPix * kWhCin;
        const int pp = r / kWhCin, ci = r - pp * kWhCin;
        const int y = pp / kWhP, x = pp - y * kWhP;
        tiles[sl * kWhTile + ((y + 1) * 11 + x + 1) * kWhCin + ci] = pv[u];
      }
    }
    __syncthreads();
    // ---- im2col row of pixel r, half (tid & 1) of its 64 columns
    {
      const int r = tid >> 1;
      const long long p = row0 + r;
      const float* px = nullptr;
      if (p < total_rows) {
        const int img = int(p / kWhPix), pp = int(p - (long long)img * kWhPix);
        const int y = pp / kWhP, x = pp - y * kWhP;
        px = tiles + (img - img0) * kWhTile + (y * 11 + x) * kWhCin;
      }
      if (tid & 1)
        wh_build_half<32>(px, sA + r * 128, r);
      else
        wh_build_half<0>(px, sA + r * 128, r);
    }
    ptx::fence_proxy_async_smem();
    __syncthreads();
    if (warp == 0 && ptx::elect_one()) {
      ptx::tc_fence_after_sync();
      ptx::mma_row4<true>(tmem_base, uint32_t(da), uint32_t(db), uint32_t(da >> 32), idesc, 0u);
      ptx::mma_commit(bar);
    }
    // ---- epilogue: two warps per TMEM lane quadrant, two 32-channel chunks each
    {
      uint32_t spins = 0;
      const uint64_t t0 = ptx::globaltimer_ns();
      while (!ptx::mbar_try_wait(bar, phase)) {
        if ((++spins & 255u) == 0 && (g_wh_abort != 0 || ptx::globaltimer_ns() - t0 > 2000000000ull)) {
          g_wh_abort = 1;
          break;
        }
      }
      phase ^= 1u;
      ptx::tc_fence_after_sync();
      const int q = warp & 3, chalf = warp >> 2;
      uint32_t* sw = reinterpret_cast<uint32_t*>(s_stage + warp * 2048);
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const int c = chalf * 2 + cc;
        float v[32];
        ptx::tmem_ld_32x32(tmem_base + (uint32_t(q * 32) << 16) + uint32_t(c * 32), v);
        // lane = pixel row of the quadrant: 32 channels -> 16 words, 4-word pieces XOR-ed by the row
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t w[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int ch = j * 8 + 2 * e;
            const __half2 h = __floats2half2_rn(fmaxf(v[ch] + s_bias[c * 32 + ch], 0.f), fmaxf(v[ch + 1] + s_bias[c * 32 + ch + 1], 0.f));
            w[e] = *reinterpret_cast<const uint32_t*>(&h);
          }
          *reinterpret_cast<uint4*>(sw + lane * 16 + ((j ^ ((lane >> 1) & 3)) << 2)) = make_uint4(w[0], w[1], w[2], w[3]);
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int r = (lane >> 2) + 8 * i, piece = lane & 3;
          const long long p = row0 + q * 32 + r;
          if (p < total_rows) {
            const uint4 t4 = *reinterpret_cast<const uint4*>(sw + r * 16 + ((piece ^ ((r >> 1) & 3)) << 2));
            const int img = int(p / kWhPix), pp = int(p - (long long)img * kWhPix);
            *reinterpret_cast<uint4*>(out + ((size_t)img * (kWhPix + kWhP) + pp) * kWhCout + c * 32 + piece * 8) = t4;
          }
        }
        __syncwarp();
      }
    }
    // the zero row behind every image that starts in this tile
    for (int img_s = int((row0 + kWhPix - 1) / kWhPix); img_s < n_img && (long long)img_s * kWhPix < row0 + kTcRows; ++img_s) {
      uint32_t* z = reinterpret_cast<uint32_t*>(out + ((size_t)img_s * (kWhPix + kWhP) + kWhPix) * kWhCout);
      for (int i = tid; i < kWhP * kWhCout / 2; i += kTcThreads) z[i] = 0u;
    }
    ptx::tc_fence_before_sync();
    __syncthreads();   // accumulator, A tile and image buffers are free for the next tile
    ptx::tc_fence_after_sync();
  }
  if (warp == 0) ptx::tmem_dealloc(tmem_base, 128);
}

}  // namespace woft

extern "C" int woft_weight_head_conv1(const float* window_nhwc, const float* mean_response, const float* weight,
                                      const float* bias, long long n_images, int patch, int levels, int channels_out,
                                      float* out_nhwc, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(window_nhwc && mean_response && weight && bias && out_nhwc, "woft_weight_head_conv1: null pointer");
  WOFT_REQUIRE(patch == kWhP && levels == kWhL && channels_out == kWhCout,
               "woft_weight_head_conv1: built for 9x9 patches, 4 levels and 128 output channels");
  WOFT_REQUIRE(n_images > 0 && n_images * kWhPix + kWhRows < (1LL << 31), "woft_weight_head_conv1: bad image count");
  const long long tiles = (n_images * kWhPix + kWhRows - 1) / kWhRows;
  WOFT_REQUIRE(tiles < (1LL << 31) && n_images < (1LL << 31), "woft_weight_head_conv1: too many images");
  const int grid = (int)(tiles < 148 * 2 ? tiles : 148 * 2);
  weight_head_conv1_kernel<false><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      window_nhwc, mean_response, weight, bias, out_nhwc, (int)n_images, (int)tiles);
  return launch_status("weight_head_conv1_kernel");
}

extern "C" int woft_weight_head_conv1_stacked_f16(const float* window_nhwc, const float* mean_response, const float* weight,
                                                  const float* bias, long long n_images, int patch, int levels,
                                                  int channels_out, void* out_stacked_f16, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(window_nhwc && mean_response && weight && bias && out_stacked_f16,
               "woft_weight_head_conv1_stacked_f16: null pointer");
  WOFT_REQUIRE(patch == kWhP && levels == kWhL && channels_out == kWhCout,
               "woft_weight_head_conv1_stacked_f16: built for 9x9 patches, 4 levels and 128 output channels");
  WOFT_REQUIRE(n_images > 0 && n_images * (kWhPix + kWhP) + kWhRows < (1LL << 31),
               "woft_weight_head_conv1_stacked_f16: bad image count");
  static const bool use_tc = [] {   // WOFT_HEAD_CONV1_TC=0: the mma.sync kernel (TF32 operands) instead of the tcgen05 one
    const char* e = getenv("WOFT_HEAD_CONV1_TC");
    return !(e && e[0] == '0');
  }();
  if (use_tc && (reinterpret_cast<uintptr_t>(out_stacked_f16) & 15) == 0) {
    const long long tiles = (n_images * kWhPix + kTcRows - 1) / kTcRows;
    const int grid = (int)(tiles < 148 * 3 ? tiles : 148 * 3);
    WOFT_CUDA_CHECK(cudaFuncSetAttribute(weight_head_conv1_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmem));
    weight_head_conv1_tc_kernel<<<grid, kTcThreads, kTcSmem, static_cast<cudaStream_t>(stream)>>>(
        window_nhwc, mean_response, weight, bias, static_cast<__half*>(out_stacked_f16), (int)n_images, (int)tiles);
    return launch_status("weight_head_conv1_tc_kernel");
  }
  const long long tiles = (n_images * kWhPix + kWhRows - 1) / kWhRows;
  const int grid = (int)(tiles < 148 * 2 ? tiles : 148 * 2);
  weight_head_conv1_kernel<true><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      window_nhwc, mean_response, weight, bias, out_stacked_f16, (int)n_images, (int)tiles);
  return launch_status("weight_head_conv1_kernel");
}
