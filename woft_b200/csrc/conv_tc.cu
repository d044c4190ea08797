// Convolutions of the flow network on the Blackwell tensor cores: an implicit GEMM on tcgen05/TMEM fed by TMA, with the
// element-wise stage that follows each convolution in the reference fused into the epilogue.
//
// Replaces, in the fused channels-last schedule (woft_b200/fast_net.py), the cuDNN convolutions of
//   BasicMotionEncoder  convc1, convc2, convf2, conv        (raft_core/update.py:79-97)
//   SepConvGRU          convz|convr stacked, convq, both halves, with the gate arithmetic (update.py:33-60)
//   FlowHead.conv1, the mask head                            (update.py:6-14, 122-125)
//   WeightHead          layers 2 and 3 with the 1x1 + patch mean (raft_core/weighted_raft.py:318-339, 347-384)
//
// Formulation.  Activations are channels-last (N, H, W, C) -- possibly a channel slice of a wider buffer: the tensor map
// carries the pixel stride.  For a tile of BW x BH = 128 output pixels and one filter tap (dy, dx), the A operand of the
// GEMM is the BW x BH box of input pixels shifted by (dy, dx), 32 fp32 (or 64 fp16) channels deep: ONE 4-D TMA box
// over (C, W, H, N) whose 128 rows land in shared memory K-major with the 128-byte swizzle tcgen05 reads.  Rows that
// fall outside the image are zero-filled by TMA -- that is the convolution's zero padding and the ragged edge of the
// tile grid at once, so there is no im2col buffer and no border code.  The B operand is the tap's (Cout, Cin) weight
// matrix, repacked once on the host to [tap][Cout][Cin] (rounded to tf32 / fp16).  The accumulator (128 pixels x up to
// 256 output channels, fp32) lives in TMEM across all taps and channel blocks: K = taps * Cin.
//
// Warp roles (384 threads, persistent CTAs over tiles), as in corr_tc.cu: warp 0 TMA producer, warp 1 MMA issuer (two
// accumulators in TMEM, so tile i+1 multiplies while tile i drains), warps 2-9 epilogue (one output pixel per thread,
// 32 channels per tcgen05.ld, the two warps of a TMEM lane quadrant taking alternate chunks), warps 10-11 two more TMA
// producers for launches whose CTAs own several tiles (the k-steps are dealt round-robin: one issuing thread's dependent
// instruction latency, not memory, bounds the k-step rate of a narrow tile).  Such launches with a filter tile of at
// most 128 channels take two pixel tiles per k-step (M = 256 per CTA, one filter tile for eight MMAs).  Epilogues:
//   BIAS_ACT  y = scale * act(acc + bias), fp32 or fp16, to one or two (strided) destinations -- writing straight into
//             the concatenation buffers of the next layer (torch.cat at update.py:96-97, 129 disappears)
//   GRU_ZR    z = sigmoid(acc[:, :C] + bz) -> z;  r*h = sigmoid(acc[:, C:] + br) * h -> the q convolution's input
//   GRU_Q     h' = (1 - z) * h + z * tanh(acc + bq), written over h
//   HEAD_MEAN relu(acc + bias) . w1x1, averaged over the 81 pixels of each 9x9 patch image (+ b1x1): the weight logit
//   FLOW_TAPS relu(acc + bias) times the second flow-head layer's 18 (tap, output) filters: per-pixel tap products
// Every mbarrier wait is bounded (2 s) and raises a status word instead of hanging (woft_conv_tc_check).
//
// Work per launch at 1280x720 (14 400 pixels): 2 * 14400 * taps * Cin * Cout flop; the 1x5 GRU gate pair
// (K = 5 * 384, N = 256) is 14.2 GFLOP.  Bound: tensor pipe (tf32: 4096 flop/clk/SM), on 113..115 of 148 SMs because
// 14 400 pixels are 112.5 tiles.
#include <cuda.h>
#include <cuda_fp16.h>

#include "common.cuh"
#include "ptx.cuh"

namespace woft {

constexpr int CT_THREADS = 384, CT_EPI_THREADS = 256;   // TMA warp, MMA warp, 8 epilogue warps, two more TMA warps
constexpr int CT_MAX_PRODUCERS = 3;
constexpr int CT_A_STAGE = 128 * 128;       // 128 pixel rows x one 128-byte K row (a_stage; more rows in halo mode)
constexpr int CT_MAX_STAGES = 8;
constexpr int CT_MAX_TAPS = 25;
constexpr int CT_SMEM_LIMIT = 232448;
constexpr int CT_STAT_C = 128;     // widest output for which the epilogue gathers instance-norm sums
__host__ __device__ constexpr int ct_stage_bytes(int) { return 2048; }   // per epilogue warp
constexpr int CT_MAX_BIAS = 640;   // padded Cout of one launch (the mask head's 576 is the largest)

enum { CT_EPI_BIAS_ACT = 0, CT_EPI_GRU_ZR = 1, CT_EPI_GRU_Q = 2, CT_EPI_HEAD_MEAN = 3, CT_EPI_FLOW_TAPS = 4 };
constexpr int CT_TAPS_W = 2 * 12;   // FLOW_TAPS: floats of second-layer filter per channel: [first 5 taps x 2 | last 4 taps x 2], padded to 12 each

struct ConvTcParams {
  CUtensorMap tmA;  // (Cin, W, H, N) box (k_elems, bw, bh, 1)
  CUtensorMap tmB;  // (Cin_pad, Cout_pad, taps) box (k_elems, bn, 1)
  int W, H, NB;
  int bw, bh;
  int x_tiles, y_tiles, n_tiles, total_tiles;
  int taps, k_chunks, k_elems, f16;
  int bn, stages, a_bytes, b_bytes;
  // halo mode: ONE activation box of pw x hh pixels per channel block serves every filter tap -- the tap's operand
  // starts tap_row[tap] rows into the box -- while the filters stream through their own ring; output pixel (i, j) of the
  // tile is GEMM row j * pw + i, rows with i >= bw are junk.  pw == bw, halo == 0 and one ring for both otherwise.
  int halo, pw, hh, a_stage, a_stages;
  // halo == 2 ("row halo"): one box per channel block AND filter column kx, holding the tile plus kh - 1 extra rows; the
  // taps of that column read it ky * bw rows further in.  With bw a multiple of 8 every operand starts on an 8-row
  // boundary of the swizzled tile (full MMA speed), and the activations are fetched kw instead of kh * kw times.
  int kw;
  // pair mode (cta_group::2): a cluster of two CTAs takes two pixel tiles that share the filters; each CTA fetches its own
  // activation box and HALF of the filter rows, the leader's MMA warp multiplies both tiles against both halves
  int pair, pix_tiles, pair_tiles;
  float* stats;      // BIAS_ACT with fp16 output, or null: per-CTA partial sums (sms, NB <= 2, 2, cout) of the stored
                     // values, for an instance norm that follows (extractor.py:30-32): removes its statistics pass
  int stride;        // 1 or 2: output pixel (x, y) reads input (stride * x + dx, stride * y + dy); W, H are OUTPUT extents
  const void* residual;      // BIAS_ACT: out = relu(residual + act(acc + bias)), fp16, laid out like `out` (the tail of a
  long long residual_stride; // residual block, raft_core/extractor.py:45-56), or null
  short tap_row[CT_MAX_TAPS];
  signed char dx[CT_MAX_TAPS], dy[CT_MAX_TAPS];
  int cout, relu;
  float scale;
  const float* bias;
  void* out;
  long long out_stride;
  int out_f16;
  void* out2;
  long long out2_stride;
  int out2_f16;
  const float* h;
  long long h_stride;
  float* z;
  long long z_stride;
  const float* w1x1;   // HEAD_MEAN: 1x1 filter (cout); FLOW_TAPS: second-layer filter (cout, CT_TAPS_W)
  float* head_out;     // HEAD_MEAN: logits; FLOW_TAPS: (pixels, 18) tap products
  float head_scale, head_bias;
  int row_period, row_valid;  // rows y with (y % row_period) >= row_valid are separators: stored as zeros, not averaged
  unsigned long long* trace;  // debug: 8 globaltimer stamps per CTA, or null
  int coalesce;               // every epilogue operand is 16-byte aligned with 16-byte multiple strides: staged accesses
  int dbg;                    // experiments (woft_conv_tc_args.debug_flags)
  int mt2, super_tiles;       // two pixel tiles per k-step share one filter tile (M = 256 per CTA): persistent launches, bn <= 128
  int producers;              // TMA-issuing warps of the plain mode (warp 0 and warps 10, 11): each takes every producers-th k-step
  int pdl;                    // host only: launch with programmatic stream serialization
};

__device__ __forceinline__ void ct_stamp(const ConvTcParams& p, int slot) {
  if (p.trace && !(p.dbg & 16)) p.trace[blockIdx.x * 8 + slot] = ptx::globaltimer_ns();
}

// debug_flags bit 16: SM-clock stamps of CTA 0's roles for its first 18 tiles (8 slots per tile) instead of the phase stamps
__device__ __forceinline__ void ct_tile_stamp(const ConvTcParams& p, int it, int slot) {
  if ((p.dbg & 16) && p.trace && blockIdx.x == 0 && it < 18 && (threadIdx.x & 31) == 0) p.trace[it * 8 + slot] = clock64();
}

// status word of the conv kernels (a pipeline wait timed out); read back by woft_conv_tc_check
__device__ int g_conv_abort = 0;

__device__ __forceinline__ bool ct_wait(uint64_t* bar, uint32_t parity) {
  if (ptx::mbar_try_wait(bar, parity)) return true;
  const uint64_t t0 = ptx::globaltimer_ns();
  uint32_t spins = 0;
  while (!ptx::mbar_try_wait(bar, parity)) {
    if ((++spins & 63u) == 0) {
      if (*reinterpret_cast<volatile int*>(&g_conv_abort) != 0) return false;
      if (ptx::globaltimer_ns() - t0 > 2000000000ull) {
        atomicExch(&g_conv_abort, 1);
        return false;
      }
    }
  }
  return true;
}

// the slow path alone, on a barrier address (the caller has tried once)
__device__ __noinline__ bool ct_wait_a(uint32_t bar, uint32_t parity) {
  const uint64_t t0 = ptx::globaltimer_ns();
  uint32_t spins = 0;
  while (!ptx::mbar_try_wait_a(bar, parity)) {
    if ((++spins & 63u) == 0) {
      if (*reinterpret_cast<volatile int*>(&g_conv_abort) != 0) return false;
      if (ptx::globaltimer_ns() - t0 > 2000000000ull) {
        atomicExch(&g_conv_abort, 1);
        return false;
      }
    }
  }
  return true;
}

__device__ __forceinline__ uint64_t ct_desc(uint32_t saddr) {  // K-major, 128-byte swizzle, 8-row groups 1024 B apart
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

struct CtTile {
  int n, x0, y0, c0;
};
__device__ __forceinline__ CtTile ct_decode(const ConvTcParams& p, int t) {
  CtTile c;
  const int nt = t % p.n_tiles;
  t /= p.n_tiles;
  const int xt = t % p.x_tiles;
  t /= p.x_tiles;
  const int yt = t % p.y_tiles;
  c.n = t / p.y_tiles;
  c.x0 = xt * p.bw;
  c.y0 = yt * p.bh;
  c.c0 = nt * p.bn;
  return c;
}

// the it-th tile of this CTA (false when there is none).  Pair mode: the cluster walks pair-tiles; rank r of the pair takes
// pixel tile 2 * pp + r, a tile past the end is a dummy outside the batch (its loads are zero-filled, nothing is stored)
__device__ __forceinline__ bool ct_tile_at(const ConvTcParams& p, int it, uint32_t rank, CtTile& c) {
  if (p.mt2) {   // the first pixel tile of the it-th super tile (n_tiles == 1); the second is ct_decode(p, 2 * st + 1),
    const int st = blockIdx.x + it * gridDim.x;   // which lies outside the batch when the tile count is odd
    if (st >= p.super_tiles) return false;
    c = ct_decode(p, 2 * st);
    return true;
  }
  if (!p.pair) {
    const int t = blockIdx.x + it * gridDim.x;
    if (t >= p.total_tiles) return false;
    c = ct_decode(p, t);
    return true;
  }
  const int pt = int(blockIdx.x >> 1) + it * int(gridDim.x >> 1);
  if (pt >= p.pair_tiles) return false;
  const int nt = pt % p.n_tiles, pix = 2 * (pt / p.n_tiles) + int(rank);
  c.c0 = nt * p.bn;
  if (pix >= p.pix_tiles) {
    c.n = p.NB;
    c.x0 = 0;
    c.y0 = 0;
  } else {
    const int xt = pix % p.x_tiles, r = pix / p.x_tiles;
    c.x0 = xt * p.bw;
    c.y0 = (r % p.y_tiles) * p.bh;
    c.n = r / p.y_tiles;
  }
  return true;
}

// exp through ex2.approx and the division through rcp.approx (about 2 ulp each): a handful of instructions where expf,
// tanhf and the IEEE division are dozens with a slow-path call -- the epilogue is issue-bound, not accuracy-bound
__device__ __forceinline__ float ct_sigmoid(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }
__device__ __forceinline__ float ct_tanh(float x) { return 1.0f - __fdividef(2.0f, 1.0f + __expf(2.0f * x)); }
// one MUFU instruction (relative error 2^-11, the rounding of the fp16 operands these values become); sigmoid through
// tanh: sigma(x) = 0.5 tanh(x / 2) + 0.5.  The GRU epilogues use them (p.dbg bit 64) unless WOFT_CONV_FAST_GATES=0:
// same end-point error against the CPU oracle (0.0046 px at 1280x720), 1 % more frames per second
__device__ __forceinline__ float ct_tanh_mufu(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float ct_sigmoid_mufu(float x) { return fmaf(0.5f, ct_tanh_mufu(0.5f * x), 0.5f); }

// 32 consecutive channels of one pixel, fp32 or fp16 destination; `nv` < 32 only in the last channel block
__device__ __forceinline__ void ct_store_row(void* base, long long stride, long long pix, int col, const float* v, int nv,
                                             bool f16) {
  if (f16) {
    __half* o = reinterpret_cast<__half*>(base) + pix * stride + col;
    if (nv == 32 && (reinterpret_cast<uintptr_t>(o) & 15) == 0) {
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        __half2 a = __floats2half2_rn(v[j], v[j + 1]), b = __floats2half2_rn(v[j + 2], v[j + 3]);
        __half2 c = __floats2half2_rn(v[j + 4], v[j + 5]), d = __floats2half2_rn(v[j + 6], v[j + 7]);
        uint4 u;
        u.x = *reinterpret_cast<uint32_t*>(&a);
        u.y = *reinterpret_cast<uint32_t*>(&b);
        u.z = *reinterpret_cast<uint32_t*>(&c);
        u.w = *reinterpret_cast<uint32_t*>(&d);
        *reinterpret_cast<uint4*>(o + j) = u;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (j < nv) o[j] = __float2half_rn(v[j]);
    }
  } else {
    float* o = reinterpret_cast<float*>(base) + pix * stride + col;
    if (nv == 32 && (reinterpret_cast<uintptr_t>(o) & 15) == 0) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (j < nv) o[j] = v[j];
    }
  }
}

__device__ __forceinline__ void ct_load_row(const float* base, long long stride, long long pix, int col, float* v) {
  const float4* s = reinterpret_cast<const float4*>(base + pix * stride + col);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float4 t = s[j];
    v[4 * j] = t.x;
    v[4 * j + 1] = t.y;
    v[4 * j + 2] = t.z;
    v[4 * j + 3] = t.w;
  }
}

// ---- coalescing through shared memory -------------------------------------------------------------------------------
// In the epilogue a thread owns one pixel (a TMEM lane), so a plain vector store touches 32 different 128-byte lines per
// warp instruction and the L1 spends 32 tag cycles on it: at 1280x720 that alone was 5-8 us per convolution.  Instead a
// warp passes blocks of 32 rows x 64 bytes through a private 2 KB staging buffer (16-byte pieces XOR-swizzled so that
// both the row-per-thread side and the transposed side are bank-conflict free) and touches global memory with 4 lanes
// per row: 8 lines per instruction, every byte of a 64-byte segment used.  rp[i] is the pixel index of row
// (lane >> 2) + 8 i of the warp's 32 rows, or -1 outside the image.
__device__ __forceinline__ uint8_t* ct_stage_own(uint8_t* stg, int lane, int j) {
  return stg + lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4);
}
__device__ __forceinline__ uint8_t* ct_stage_row(uint8_t* stg, int r, int q) {
  return stg + r * 64 + ((q ^ ((r >> 1) & 3)) << 4);
}

// `stat` (fp16 outputs only, or null): this WARP's shared-memory accumulators [sum (C) | sum of squares (C)] of the tile's
// image -- no atomics, so the sums are reproducible bit for bit; the values added are the rounded fp16 numbers that are
// stored, so a later normalisation sees exactly their statistics
__device__ __forceinline__ void ct_store_tile(void* base, long long stride, int col, const float* v, int nv, bool f16,
                                              uint8_t* stg, int lane, const int* rp, float* stat = nullptr,
                                              int stat_c = 0) {
  const int q = lane & 3;
  if (f16) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      __half2 a = __floats2half2_rn(v[8 * j], v[8 * j + 1]), b = __floats2half2_rn(v[8 * j + 2], v[8 * j + 3]);
      __half2 c = __floats2half2_rn(v[8 * j + 4], v[8 * j + 5]), d = __floats2half2_rn(v[8 * j + 6], v[8 * j + 7]);
      uint4 u;
      u.x = *reinterpret_cast<uint32_t*>(&a);
      u.y = *reinterpret_cast<uint32_t*>(&b);
      u.z = *reinterpret_cast<uint32_t*>(&c);
      u.w = *reinterpret_cast<uint32_t*>(&d);
      *reinterpret_cast<uint4*>(ct_stage_own(stg, lane, j)) = u;
    }
    __syncwarp();
    const int left = nv - q * 8;
    float ps[8], pq[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) ps[e] = pq[e] = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = (lane >> 2) + 8 * i;
      if (rp[i] >= 0 && left > 0) {
        const uint4 t = *reinterpret_cast<const uint4*>(ct_stage_row(stg, r, q));
        if (stat) {
          const __half2* hp = reinterpret_cast<const __half2*>(&t);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 f = __half22float2(hp[e]);
            ps[2 * e] += f.x;
            ps[2 * e + 1] += f.y;
            pq[2 * e] = fmaf(f.x, f.x, pq[2 * e]);
            pq[2 * e + 1] = fmaf(f.y, f.y, pq[2 * e + 1]);
          }
        }
        __half* o = reinterpret_cast<__half*>(base) + (long long)rp[i] * stride + col + q * 8;
        if (left >= 8) {
          *reinterpret_cast<uint4*>(o) = t;
        } else {
          const __half* hsrc = reinterpret_cast<const __half*>(&t);
#pragma unroll
          for (int e = 0; e < 8; ++e)
            if (e < left) o[e] = hsrc[e];
        }
      }
    }
    if (stat) {   // warp-uniform: the eight lanes that share a 16-byte piece add up, one of them updates shared memory
#pragma unroll
      for (int e = 0; e < 8; ++e) {
#pragma unroll
        for (int o = 4; o < 32; o <<= 1) {
          ps[e] += __shfl_xor_sync(0xffffffffu, ps[e], o);
          pq[e] += __shfl_xor_sync(0xffffffffu, pq[e], o);
        }
      }
      if (lane < 4 && left > 0) {   // `stat` is private to this warp: plain read-modify-write, a fixed order of additions
#pragma unroll
        for (int e = 0; e < 8; ++e)
          if (e < left) {
            stat[col + q * 8 + e] += ps[e];
            stat[stat_c + col + q * 8 + e] += pq[e];
          }
      }
    }
    __syncwarp();
  } else {
#pragma unroll
    for (int hb = 0; hb < 2; ++hb) {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        *reinterpret_cast<float4*>(ct_stage_own(stg, lane, j)) =
            make_float4(v[16 * hb + 4 * j], v[16 * hb + 4 * j + 1], v[16 * hb + 4 * j + 2], v[16 * hb + 4 * j + 3]);
      __syncwarp();
      const int left = nv - 16 * hb - q * 4;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int r = (lane >> 2) + 8 * i;
        if (rp[i] >= 0 && left > 0) {
          const float4 t = *reinterpret_cast<const float4*>(ct_stage_row(stg, r, q));
          float* o = reinterpret_cast<float*>(base) + (long long)rp[i] * stride + col + 16 * hb + q * 4;
          if (left >= 4) {
            *reinterpret_cast<float4*>(o) = t;
          } else {
            o[0] = t.x;
            if (left > 1) o[1] = t.y;
            if (left > 2) o[2] = t.z;
          }
        }
      }
      __syncwarp();
    }
  }
}

// ---- the GRU epilogues work in the coalesced layout itself ---------------------------------------------------------
// A 32-pixel x 32-channel fp32 block is 32 rows of 128 bytes.  Lane l owns the 16-byte piece l & 7 of rows
// (l >> 3) + 4 i, i = 0..7: a warp access covers four whole 128-byte lines.  Only the accumulator has to be brought into
// this layout (two 16-row passes through the warp's 2 KB buffer, 16-byte pieces XOR-swizzled by the row); h, z and the results
// are read and written by plain vector accesses whose latencies overlap, where the row-per-thread form needed a
// shared-memory round trip (and two global round trips) per operand.  rp8[i]: pixel index of row (l >> 3) + 4 i or -1.
__device__ __forceinline__ void ct_to_coalesced(float* v, uint8_t* stg, int lane) {
  float o[32];
#pragma unroll
  for (int hb = 0; hb < 2; ++hb) {   // rows 16 hb .. 16 hb + 15 through the 2 KB buffer
    if ((lane >> 4) == hb) {
      const int r = lane & 15;
#pragma unroll
      for (int j = 0; j < 8; ++j)
        *reinterpret_cast<float4*>(stg + r * 128 + ((j ^ (r & 7)) << 4)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = (lane >> 3) + 4 * i;
      const float4 t = *reinterpret_cast<const float4*>(stg + r * 128 + (((lane & 7) ^ (r & 7)) << 4));
      o[16 * hb + 4 * i] = t.x; o[16 * hb + 4 * i + 1] = t.y; o[16 * hb + 4 * i + 2] = t.z; o[16 * hb + 4 * i + 3] = t.w;
    }
    __syncwarp();
  }
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = o[j];
}
__device__ __forceinline__ void ct_ld_coalesced(const float* base, long long stride, int col, const int* rp8, int lane, float* o) {
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
    if (rp8[i] >= 0) t = *reinterpret_cast<const float4*>(base + (long long)rp8[i] * stride + col + (lane & 7) * 4);
    o[4 * i] = t.x; o[4 * i + 1] = t.y; o[4 * i + 2] = t.z; o[4 * i + 3] = t.w;
  }
}
__device__ __forceinline__ void ct_st_coalesced(void* base, long long stride, int col, const int* rp8, int lane, const float* v,
                                                bool f16) {
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    if (rp8[i] < 0) continue;
    if (f16) {
      const __half2 a = __floats2half2_rn(v[4 * i], v[4 * i + 1]), b = __floats2half2_rn(v[4 * i + 2], v[4 * i + 3]);
      uint2 u;
      u.x = *reinterpret_cast<const uint32_t*>(&a);
      u.y = *reinterpret_cast<const uint32_t*>(&b);
      *reinterpret_cast<uint2*>(reinterpret_cast<__half*>(base) + (long long)rp8[i] * stride + col + (lane & 7) * 4) = u;
    } else {
      *reinterpret_cast<float4*>(reinterpret_cast<float*>(base) + (long long)rp8[i] * stride + col + (lane & 7) * 4) =
          make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    }
  }
}

// 32 fp32 channels [col, col + 32) of the warp's 32 pixels, each thread ending up with its own pixel's values
__device__ __forceinline__ void ct_load_tile(const float* base, long long stride, int col, float* v, uint8_t* stg, int lane,
                                             const int* rp) {
  const int q = lane & 3;
#pragma unroll
  for (int hb = 0; hb < 2; ++hb) {
    float4 t[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
      t[i] = rp[i] >= 0 ? __ldg(reinterpret_cast<const float4*>(base + (long long)rp[i] * stride + col + 16 * hb + q * 4))
                        : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 4; ++i) *reinterpret_cast<float4*>(ct_stage_row(stg, (lane >> 2) + 8 * i, q)) = t[i];
    __syncwarp();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float4 u = *reinterpret_cast<const float4*>(ct_stage_own(stg, lane, j));
      v[16 * hb + 4 * j] = u.x;
      v[16 * hb + 4 * j + 1] = u.y;
      v[16 * hb + 4 * j + 2] = u.z;
      v[16 * hb + 4 * j + 3] = u.w;
    }
    __syncwarp();
  }
}

// PAIR instantiations contain cluster / cta_group::2 instructions and can only be launched as clusters of two
template <int EPI, bool PAIR, bool F16>
__global__ void __launch_bounds__(CT_THREADS, 1) conv_tc_kernel(const __grid_constant__ ConvTcParams p) {
  extern __shared__ uint8_t ct_smem_raw[];
  // 1 KiB alignment by an offset into the array (a cast through an integer would turn every later access generic)
  uint8_t* smem = ct_smem_raw + ((1024u - (ptx::smem_u32(ct_smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = sA + p.a_stages * p.a_stage;   // a_stages == stages unless the activations have their own ring (halo)
  float* s_bias = reinterpret_cast<float*>(sB + p.stages * p.b_bytes);   // CT_MAX_BIAS floats
  float* s_red = s_bias + CT_MAX_BIAS;                                   // 2 x 128 floats (HEAD_MEAN)
  uint8_t* s_stage = reinterpret_cast<uint8_t*>(s_red + 256);            // 8 warps x 2 KB
  float* s_stat = reinterpret_cast<float*>(s_stage + 8 * ct_stage_bytes(EPI));         // [8 warps][2 images][sum | squares][CT_STAT_C],
  // only when gathered; FLOW_TAPS: the second-layer filter lies over the (unused) staging buffer and 8 KB behind it
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_stat + (EPI == CT_EPI_FLOW_TAPS ? 2048 : p.stats ? 8 * 4 * CT_STAT_C : 0));
  float* s_w2 = reinterpret_cast<float*>(s_stage);
  uint64_t* full = bars;
  uint64_t* empty = bars + CT_MAX_STAGES;
  uint64_t* tfull = bars + 2 * CT_MAX_STAGES;
  uint64_t* tempty = tfull + 2;
  uint64_t* a_full = tempty + 2;   // [4] activation ring of the halo mode
  uint64_t* a_empty = a_full + 4;  // [4]
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(a_empty + 4);
  volatile int* epi_go = reinterpret_cast<volatile int*>(tmem_ptr + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    ct_stamp(p, 0);
    for (int s = 0; s < p.stages; ++s) {
      ptx::mbar_init(full + s, 1);   // pair: only the leader arrives (announcing both CTAs' bytes); the peer's loads just complete
      ptx::mbar_init(empty + s, 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(tfull + i, 1);
      ptx::mbar_init(tempty + i, PAIR ? 2 * CT_EPI_THREADS : CT_EPI_THREADS);   // pair: both CTAs' epilogues
    }
    for (int i = 0; i < 4; ++i) {
      ptx::mbar_init(a_full + i, 1);
      ptx::mbar_init(a_empty + i, 1);
    }
    ptx::fence_mbar_init();
    ptx::prefetch_tmap(&p.tmA);
    ptx::prefetch_tmap(&p.tmB);
  }
  const uint32_t rank = PAIR ? ptx::cluster_ctarank() : 0u;
  if (warp == 1) {
    if (PAIR)
      ptx::tmem_alloc_2sm(tmem_ptr, 512);
    else
      ptx::tmem_alloc(tmem_ptr, 512);
  }
  {
    const int nb = p.n_tiles * p.bn;
    for (int i = threadIdx.x; i < nb && i < CT_MAX_BIAS; i += CT_THREADS) {
      float b = (p.bias && i < p.cout) ? __ldg(p.bias + i) : 0.f;
      s_bias[i] = b;
    }
    if (p.stats)
      for (int i = threadIdx.x; i < 8 * 4 * CT_STAT_C; i += CT_THREADS) s_stat[i] = 0.f;
    if (EPI == CT_EPI_FLOW_TAPS)
      for (int i = threadIdx.x; i < p.cout * CT_TAPS_W; i += CT_THREADS) s_w2[i] = __ldg(p.w1x1 + i);
  }
  ptx::tc_fence_before_sync();
  if (PAIR)
    ptx::cluster_sync();   // the peer's barriers must be initialised before anything is signalled on them
  else
    __syncthreads();
  ptx::tc_fence_after_sync();
  // Everything above (barriers, TMEM, tensor-map prefetch, bias) is independent of the previous kernel in the stream; from
  // here on its output is read and buffers it may still be reading are written.  Launched with programmatic
  // serialization this kernel's set-up overlaps the predecessor's tail, and the successor's overlaps this kernel's.
  ptx::griddep_wait();
  ptx::griddep_launch();
  const uint32_t tmem_base = *tmem_ptr;
  const int k_steps = p.taps * p.k_chunks;
  if (threadIdx.x == 0) ct_stamp(p, 1);

  const int prod = warp == 0 ? 0 : warp - 9;   // warps 10 and 11: the second and third TMA warp of the plain mode
  if (warp == 0 || (warp >= 10 && prod < p.producers)) {
    // ================================ TMA producers ================================
    // One elected lane per warp.  A k-step costs the issuing thread a few hundred cycles of dependent latency (barrier
    // try-wait ~90, arrive, two TMA issues, coordinate arithmetic) -- more than the MMAs of a narrow tile take -- so in the
    // plain mode the k-steps are dealt round-robin to `producers` warps: the stage sequence is fixed, each warp walks
    // every producers-th stage of the ring.
    if (ptx::elect_one()) {
      bool alive = true;
      if (p.halo) {
        // two rings, filled in the order the MMA warp drains them: the activation box of a channel block, then the
        // filters of every tap for that block
        int stage = 0;
        uint32_t phase = 0;
        int a_st = 0;
        uint32_t a_ph = 0;
        CtTile tc;
        for (int it = 0; alive && ct_tile_at(p, it, rank, tc); ++it) {
          const int groups = p.halo == 2 ? p.kw : 1, tpg = p.taps / groups;
          for (int kg = 0; kg < p.k_chunks * groups && alive; ++kg) {
            const int kc = kg / groups, g = kg - kc * groups;   // g: filter column of a row-halo box (tap g is its first)
            if (!ct_wait(a_empty + a_st, a_ph ^ 1u)) {
              alive = false;
              break;
            }
            ptx::mbar_arrive_expect_tx(a_full + a_st, uint32_t(p.a_bytes));
            ptx::tma_load_4d(sA + a_st * p.a_stage, &p.tmA, a_full + a_st, kc * p.k_elems, tc.x0 + p.dx[g], tc.y0 + p.dy[g], tc.n);
            if (++a_st == p.a_stages) {
              a_st = 0;
              a_ph ^= 1u;
            }
            for (int i = 0; i < tpg; ++i) {
              const int tap = p.halo == 2 ? i * p.kw + g : i;
              if (!ct_wait(empty + stage, phase ^ 1u)) {
                alive = false;
                break;
              }
              ptx::mbar_arrive_expect_tx(full + stage, uint32_t(p.b_bytes));
              ptx::tma_load_3d(sB + stage * p.b_bytes, &p.tmB, full + stage, kc * p.k_elems, tc.c0, tap);
              if (++stage == p.stages) {
                stage = 0;
                phase ^= 1u;
              }
            }
          }
        }
      } else {
        const int np = p.producers, stages = p.stages, k_chunks = p.k_chunks, k_elems = p.k_elems, stride = p.stride;
        const int a_stage = p.a_stage, b_bytes = p.b_bytes;
        const bool load_a = !(p.dbg & 1), load_b = !(p.dbg & 2);   // dbg bits 1, 2: timing experiments without the loads
        const bool mt2 = p.mt2 != 0;
        const uint32_t tx = uint32_t((load_a ? (mt2 ? 2 : 1) * p.a_bytes : 0) + (load_b ? b_bytes : 0));
        int stage = prod;   // host: producers <= stages
        uint32_t phase = 0;
        int ks = prod;      // this warp's next k-step of the current tile
        CtTile tc;
        for (int it = 0; alive && ct_tile_at(p, it, rank, tc); ++it) {
          if (prod == 0) ct_tile_stamp(p, it, 0);
          int tap = ks / k_chunks, kc = ks - tap * k_chunks;
          const int xb = stride * tc.x0, yb = stride * tc.y0;
          CtTile tc2 = tc;
          if (mt2) tc2 = ct_decode(p, 2 * (int(blockIdx.x) + it * int(gridDim.x)) + 1);
          const int xb2 = stride * tc2.x0, yb2 = stride * tc2.y0;
          while (ks < k_steps) {
            const int xs = xb + p.dx[tap], ys = yb + p.dy[tap];
            if (!ct_wait(empty + stage, phase ^ 1u)) {
              alive = false;
              break;
            }
            if (PAIR) {
              // both CTAs' bytes are counted on the LEADER's barrier, which the leader arms with the sum.  The peer does
              // not arrive (a remote release-arrive per stage paced the whole pipeline at 0.74 us per step): its bytes may
              // land before the leader has armed the phase, which only drives the pending count negative for a while --
              // the phase cannot complete before the leader's arrival, and the peer cannot reach the next use of this
              // stage before the leader's MMAs have drained it
              const uint32_t lead_full = ptx::mapa(full + stage, 0);
              if (rank == 0) ptx::mbar_arrive_expect_tx(full + stage, 2u * uint32_t(p.a_bytes + b_bytes));
              ptx::tma_load_4d_2sm(sA + stage * a_stage, &p.tmA, lead_full, kc * k_elems, xs, ys, tc.n);
              ptx::tma_load_3d_2sm(sB + stage * b_bytes, &p.tmB, lead_full, kc * k_elems, tc.c0 + int(rank) * (p.bn >> 1), tap);
            } else {
              ptx::mbar_arrive_expect_tx(full + stage, tx);
              if (load_a) ptx::tma_load_4d(sA + stage * a_stage, &p.tmA, full + stage, kc * k_elems, xs, ys, tc.n);
              if (load_a && mt2)
                ptx::tma_load_4d(sA + stage * a_stage + CT_A_STAGE, &p.tmA, full + stage, kc * k_elems, xb2 + p.dx[tap],
                                 yb2 + p.dy[tap], tc2.n);
              if (load_b) ptx::tma_load_3d(sB + stage * b_bytes, &p.tmB, full + stage, kc * k_elems, tc.c0, tap);
            }
            stage += np;
            if (stage >= stages) {
              stage -= stages;
              phase ^= 1u;
            }
            ks += np;
            kc += np;
            while (kc >= k_chunks) {
              kc -= k_chunks;
              ++tap;
            }
          }
          ks -= k_steps;
          if (prod == 0) ct_tile_stamp(p, it, 1);
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    // pair mode: the leader CTA issues for both SMs
    if (rank == 0 && ptx::elect_one()) {   // one elected lane runs the whole loop
      const uint32_t fmt = F16 ? 0u : 2u;  // F16 = 0, TF32 = 2
      const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | (uint32_t(p.bn >> 3) << 17) |
                             (uint32_t((PAIR ? 256 : 128) >> 4) << 24);
      // this CTA's (pair's) tile count: the issuing warp has no use for tile coordinates
      const int my_tiles = p.pair ? (p.pair_tiles - int(blockIdx.x >> 1) + int(gridDim.x >> 1) - 1) / int(gridDim.x >> 1)
                                  : ((p.mt2 ? p.super_tiles : p.total_tiles) - int(blockIdx.x) + int(gridDim.x) - 1) / int(gridDim.x);
      int stage = 0;
      uint32_t phase = 0;
      bool alive = true;
      if (p.halo) {
        int a_st = 0;
        uint32_t a_ph = 0;
        for (int it = 0; alive && it < my_tiles; ++it) {
          const int buf = it & 1;
          const uint32_t use = uint32_t(it >> 1);
          if (!ct_wait(tempty + buf, (use & 1u) ^ 1u)) break;
          ptx::tc_fence_after_sync();
          const uint32_t d_tmem = tmem_base + uint32_t(buf * 256);
          const int groups = p.halo == 2 ? p.kw : 1, tpg = p.taps / groups;
          for (int kg = 0; kg < p.k_chunks * groups && alive; ++kg) {
            const int g = kg % groups;
            if (!ct_wait(a_full + a_st, a_ph)) {
              alive = false;
              break;
            }
            const uint32_t a_base = ptx::smem_u32(sA + a_st * p.a_stage);
            for (int i = 0; i < tpg; ++i) {
              const int tap = p.halo == 2 ? i * p.kw + g : i;
              if (!ct_wait(full + stage, phase)) {
                alive = false;
                break;
              }
              ptx::tc_fence_after_sync();
              if (kg == 0 && i == 0 && it == 0) ct_stamp(p, 2);
              // the tap's operand is the same box read from tap_row[tap] rows further in: the 128-byte swizzle is a
              // function of the shared-memory address, so a descriptor may start on any row of the tile
              const uint64_t da = ct_desc(a_base + uint32_t(p.tap_row[tap]) * 128u);
              const uint64_t db = ct_desc(ptx::smem_u32(sB + stage * p.b_bytes));
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const uint32_t acc = (kg > 0 || i > 0 || k > 0) ? 1u : 0u;
                if (F16)
                  ptx::mma_f16(d_tmem, da + uint64_t(2 * k), db + uint64_t(2 * k), idesc, acc);
                else
                  ptx::mma_tf32(d_tmem, da + uint64_t(2 * k), db + uint64_t(2 * k), idesc, acc);
              }
              ptx::mma_commit(empty + stage);
              if (++stage == p.stages) {
                stage = 0;
                phase ^= 1u;
              }
            }
            if (alive) ptx::mma_commit(a_empty + a_st);   // every MMA reading this activation box has been issued
            if (++a_st == p.a_stages) {
              a_st = 0;
              a_ph ^= 1u;
            }
          }
          if (alive) ptx::mma_commit(tfull + buf);
          if (it == 0) ct_stamp(p, 3);
        }
      } else {
        // Plain mode.  The issuing thread is latency bound (about 5 cycles per instruction, every one dependent on the
        // last through the loop state), so the k-step is kept to the barrier wait, four MMAs and the commit: descriptors
        // and barrier addresses are running 32-bit values (the descriptors' high words never change), and the ring wrap
        // is a loop boundary, not a select chain.
        const int stages = p.stages;
        const uint64_t da0 = ct_desc(ptx::smem_u32(sA)), db0 = ct_desc(ptx::smem_u32(sB));
        const uint32_t desc_hi = uint32_t(da0 >> 32);
        const uint32_t a_lo0 = uint32_t(da0), b_lo0 = uint32_t(db0);
        const uint32_t a_step = uint32_t(p.a_stage >> 4), b_step = uint32_t(p.b_bytes >> 4);
        const uint32_t full0 = ptx::smem_u32(full), empty0 = ptx::smem_u32(empty);
        const bool mt2 = p.mt2 != 0;   // second accumulator of the super tile: 128 columns further, second activation box
        uint32_t a_lo = a_lo0, b_lo = b_lo0, full_a = full0, empty_a = empty0;
        for (int it = 0; alive && it < my_tiles; ++it) {
          const int buf = it & 1;
          const uint32_t use = uint32_t(it >> 1);
          if (!ct_wait(tempty + buf, (use & 1u) ^ 1u)) break;
          ptx::tc_fence_after_sync();
          ct_tile_stamp(p, it, 2);
          const uint32_t d_tmem = tmem_base + uint32_t(buf * 256);
          uint32_t started = 0;   // 0: the tile's first MMA overwrites the accumulator
          int left = k_steps;
          while (left > 0 && alive) {
            int n = stages - stage;   // k-steps up to the end of the ring or of the tile
            if (n > left) n = left;
            left -= n;
            stage += n;
            for (; n > 0; --n) {
              if (!ptx::mbar_try_wait_a(full_a, phase) && !ct_wait_a(full_a, phase)) {
                alive = false;
                break;
              }
              ptx::tc_fence_after_sync();
              if (started == 0 && it == 0) ct_stamp(p, 2);
              if (PAIR) {
                const uint64_t da = (uint64_t(desc_hi) << 32) | a_lo, db = (uint64_t(desc_hi) << 32) | b_lo;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  if (F16)
                    ptx::mma_f16_2sm(d_tmem, da + uint64_t(2 * k), db + uint64_t(2 * k), idesc, k ? 1u : started);
                  else
                    ptx::mma_tf32_2sm(d_tmem, da + uint64_t(2 * k), db + uint64_t(2 * k), idesc, k ? 1u : started);
                }
                ptx::mma_commit_2sm(empty + (stage - n), 3);   // frees the stage in both CTAs
              } else {
                ptx::mma_row4<F16>(d_tmem, a_lo, b_lo, desc_hi, idesc, started);
                if (mt2) ptx::mma_row4<F16>(d_tmem + 128u, a_lo + uint32_t(CT_A_STAGE >> 4), b_lo, desc_hi, idesc, started);
                ptx::mma_commit_a(empty_a);
              }
              started = 1;
              a_lo += a_step;
              b_lo += b_step;
              full_a += 8;
              empty_a += 8;
            }
            if (stage == stages) {
              stage = 0;
              phase ^= 1u;
              a_lo = a_lo0;
              b_lo = b_lo0;
              full_a = full0;
              empty_a = empty0;
            }
          }
          if (alive) {
            if (PAIR)
              ptx::mma_commit_2sm(tfull + buf, 3);
            else
              ptx::mma_commit(tfull + buf);
          }
          if (it == 0) ct_stamp(p, 3);
          ct_tile_stamp(p, it, 3);
        }
      }
    }
  } else if (warp < 10) {   // warps 10, 11 that are not producers have nothing to do
    // ================================ epilogue ================================
    // Eight warps: two per TMEM lane quadrant (a warp reaches lanes 32 * (warp % 4) ...), the pair splitting the tile's
    // 32-column chunks by parity.  One warp per scheduler cannot hide the latency of its own dependent instructions;
    // two can, and each has half the work.
    const bool leader = (threadIdx.x == 64);
    const int half = (warp - 2) >> 2;                 // which chunks of the tile this warp takes
    const int m = (warp & 3) * 32 + lane;             // TMEM lane == pixel row of the tile
    const uint32_t lane_base = uint32_t((warp & 3) * 32) << 16;
    const int xi = m % p.pw, yj = m / p.pw;   // halo mode: GEMM rows run over the box pitch, columns >= bw are junk
    const int chunks = p.bn >> 5;
    int it = 0;
    CtTile tc;
    const uint32_t lead_tempty[2] = {PAIR ? ptx::mapa(tempty, 0) : 0u, PAIR ? ptx::mapa(tempty + 1, 0) : 0u};
    for (; ct_tile_at(p, it, rank, tc); ++it) {
      const int buf = it & 1;
      const uint32_t use = uint32_t(it >> 1);
      int x = tc.x0 + xi, y = tc.y0 + yj;
      bool valid = xi < p.bw && yj < p.bh && x < p.W && y < p.H && tc.n < p.NB;
      bool separator = p.row_period > 0 && (y % p.row_period) >= p.row_valid;
      long long pix = ((long long)tc.n * p.H + y) * p.W + x;
      uint8_t* stg = s_stage + (warp - 2) * ct_stage_bytes(EPI);
      const bool co = p.coalesce != 0;
      constexpr bool GRU = EPI == CT_EPI_GRU_ZR || EPI == CT_EPI_GRU_Q;
      // GRU epilogues, coalesced layout: the first chunk's h (and z) are requested before the accumulator is waited for
      int rp8[8];
      float hc[32], zc[32];
      const int hid = p.bn >> 1;
      // gates: this warp's chunks in the order r, r, z, z (the r chunks need h); blend: ascending
      const int first_c = EPI == CT_EPI_GRU_ZR ? half + 4 : half;
      if (GRU && co) {
#pragma unroll
        for (int i = 0; i < 8; ++i) rp8[i] = __shfl_sync(0xffffffffu, valid ? int(pix) : -1, (lane >> 3) + 4 * i);
        if (first_c < chunks && !(p.dbg & 32)) {   // dbg 32: experiment without the early request
          ct_ld_coalesced(p.h, p.h_stride, tc.c0 + first_c * 32 - (EPI == CT_EPI_GRU_ZR ? hid : 0), rp8, lane, hc);
          if (EPI == CT_EPI_GRU_Q) ct_ld_coalesced(p.z, p.z_stride, tc.c0 + first_c * 32, rp8, lane, zc);
        }
      }
      if (leader) *epi_go = ct_wait(tfull + buf, use & 1u) ? 1 : 0;
      if (leader) ct_tile_stamp(p, it, 4);
      ptx::named_bar_sync(1, CT_EPI_THREADS);
      if (*epi_go == 0) break;
      ptx::tc_fence_after_sync();
      if (leader) ct_tile_stamp(p, it, 5);
      if (leader && it == 0) ct_stamp(p, 4);
      if (GRU && co) {
        const int q4 = (lane & 7) * 4;
        const bool fast_gates = (p.dbg & 64) != 0;
        const int mine = (chunks - half + 1) >> 1;   // chunks of this warp
#pragma unroll 1
        for (int ci = 0; ci < mine; ++ci) {
          // gates with 8 chunks: order half+4, half+6, half, half+2; otherwise ascending
          const int c = (EPI == CT_EPI_GRU_ZR && chunks == 8) ? half + 2 * ((ci + 2) & 3) : half + 2 * ci;
          const int col = tc.c0 + c * 32;
          float v[32];
#define CT_ESTAMP(j) if ((p.dbg & 16) && p.trace && blockIdx.x == 0 && leader && it == 0) p.trace[64 + 8 * ci + (j)] = clock64();
          CT_ESTAMP(0)
          ptx::tmem_ld_32x32(tmem_base + lane_base + uint32_t(buf * 256 + c * 32), v);
          CT_ESTAMP(1)
          if (ci == mine - 1) {   // this warp's share of the accumulator is read, hand it back
            ptx::tc_fence_before_sync();
            if (PAIR)
              ptx::mbar_arrive_cluster(lead_tempty[buf]);
            else
              ptx::mbar_arrive(tempty + buf);
          }
          ct_to_coalesced(v, stg, lane);
          CT_ESTAMP(2)
          const float4 b4 = *reinterpret_cast<const float4*>(s_bias + col + q4);
          const float bb[4] = {b4.x, b4.y, b4.z, b4.w};
          if (EPI == CT_EPI_GRU_ZR) {
            if (col < hid) {
              if (fast_gates) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = ct_sigmoid_mufu(v[j] + bb[j & 3]);
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = ct_sigmoid(v[j] + bb[j & 3]);
              }
              ct_st_coalesced(p.z, p.z_stride, col, rp8, lane, v, false);
            } else {
              if (c != first_c || (p.dbg & 32)) ct_ld_coalesced(p.h, p.h_stride, col - hid, rp8, lane, hc);
              if (fast_gates) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = ct_sigmoid_mufu(v[j] + bb[j & 3]) * hc[j];
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = ct_sigmoid(v[j] + bb[j & 3]) * hc[j];
              }
              ct_st_coalesced(p.out, p.out_stride, col - hid, rp8, lane, v, p.out_f16 != 0);
            }
          } else {
            if (c != first_c || (p.dbg & 32)) {
              ct_ld_coalesced(p.h, p.h_stride, col, rp8, lane, hc);
              ct_ld_coalesced(p.z, p.z_stride, col, rp8, lane, zc);
            }
            if (__float_as_int(hc[0]) == 0x7fc12345 && __float_as_int(zc[31]) == 0x7fc12345) CT_ESTAMP(7)   // forces the loads to have landed
            CT_ESTAMP(3)
            if (fast_gates) {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = fmaf(zc[j], ct_tanh_mufu(v[j] + bb[j & 3]) - hc[j], hc[j]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = (1.f - zc[j]) * hc[j] + zc[j] * ct_tanh(v[j] + bb[j & 3]);
            }
            CT_ESTAMP(4)
            ct_st_coalesced(p.out, p.out_stride, col, rp8, lane, v, false);
            if (p.out2) ct_st_coalesced(p.out2, p.out2_stride, col, rp8, lane, v, p.out2_f16 != 0);
            CT_ESTAMP(5)
          }
        }
        if (mine == 0) {
          ptx::tc_fence_before_sync();
          if (PAIR)
            ptx::mbar_arrive_cluster(lead_tempty[buf]);
          else
            ptx::mbar_arrive(tempty + buf);
        }
        if (leader && it == 0) ct_stamp(p, 5);
        if (leader) ct_tile_stamp(p, it, 6);
        continue;
      }
      // mt2: the super tile's two pixel tiles one after the other, their accumulators 128 TMEM columns apart
      const int nsub = p.mt2 ? 2 : 1;
      for (int sub = 0; sub < nsub; ++sub) {
      if (sub == 1) {
        tc = ct_decode(p, 2 * (int(blockIdx.x) + it * int(gridDim.x)) + 1);
        x = tc.x0 + xi;
        y = tc.y0 + yj;
        valid = xi < p.bw && yj < p.bh && x < p.W && y < p.H && tc.n < p.NB;
        separator = p.row_period > 0 && (y % p.row_period) >= p.row_valid;
        pix = ((long long)tc.n * p.H + y) * p.W + x;
      }
      const uint32_t acc_col = uint32_t(buf * 256 + sub * 128);
      const bool last_sub = sub == nsub - 1;
      int rp[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) rp[i] = __shfl_sync(0xffffffffu, valid ? int(pix) : -1, (lane >> 2) + 8 * i);
      if (EPI == CT_EPI_FLOW_TAPS) {
        // FlowHead (update.py:6-14): relu(conv1) never leaves the SM.  The second, 3x3 -> 2-channel convolution is
        // linear, so every pixel's 18 (tap, output) products with its own 256 activations are formed here and a small
        // gather kernel adds the nine shifted products per output pixel.  Both warps of a lane quadrant read all
        // chunks; one forms taps 0-4, the other taps 5-8 (no exchange between them).
        float t[10];
#pragma unroll
        for (int k = 0; k < 10; ++k) t[k] = 0.f;
#pragma unroll 1
        for (int c = 0; c < chunks; ++c) {
          const int col = tc.c0 + c * 32;
          float v[32], bv[32];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 b4 = reinterpret_cast<const float4*>(s_bias + col)[j];
            bv[4 * j] = b4.x; bv[4 * j + 1] = b4.y; bv[4 * j + 2] = b4.z; bv[4 * j + 3] = b4.w;
          }
          ptx::tmem_ld_32x32(tmem_base + lane_base + uint32_t(buf * 256 + c * 32), v);
          if (c == chunks - 1) {
            ptx::tc_fence_before_sync();
            if (PAIR)
              ptx::mbar_arrive_cluster(lead_tempty[buf]);
            else
              ptx::mbar_arrive(tempty + buf);
          }
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float a = fmaxf(v[j] + bv[j], 0.f);
            const float4* w = reinterpret_cast<const float4*>(s_w2 + (col + j) * CT_TAPS_W + half * 12);
            const float4 w0 = w[0], w1 = w[1], w2 = w[2];
            t[0] = fmaf(a, w0.x, t[0]); t[1] = fmaf(a, w0.y, t[1]); t[2] = fmaf(a, w0.z, t[2]); t[3] = fmaf(a, w0.w, t[3]);
            t[4] = fmaf(a, w1.x, t[4]); t[5] = fmaf(a, w1.y, t[5]); t[6] = fmaf(a, w1.z, t[6]); t[7] = fmaf(a, w1.w, t[7]);
            t[8] = fmaf(a, w2.x, t[8]); t[9] = fmaf(a, w2.y, t[9]);
          }
        }
        if (valid) {
          float2* o = reinterpret_cast<float2*>(p.head_out + pix * 18 + half * 10);
          o[0] = make_float2(t[0], t[1]);
          o[1] = make_float2(t[2], t[3]);
          o[2] = make_float2(t[4], t[5]);
          o[3] = make_float2(t[6], t[7]);
          if (half == 0) o[4] = make_float2(t[8], t[9]);
        }
        if (leader && it == 0) ct_stamp(p, 5);
        continue;
      }
      float head_sum = 0.f;
#pragma unroll 1
      for (int c = half; c < chunks; c += 2) {
        const int col = tc.c0 + c * 32;  // output channel of v[0]
        float v[32], bv[32];
        // operands of this chunk that do not depend on the accumulator are requested before the TMEM read is waited for
        float hv[32], zv[32];
        if (EPI == CT_EPI_GRU_ZR) {
          if (col >= (p.bn >> 1)) {
            if (co)
              ct_load_tile(p.h, p.h_stride, col - (p.bn >> 1), hv, stg, lane, rp);
            else if (valid)
              ct_load_row(p.h, p.h_stride, pix, col - (p.bn >> 1), hv);
          }
        } else if (EPI == CT_EPI_GRU_Q) {
          if (co) {
            ct_load_tile(p.h, p.h_stride, col, hv, stg, lane, rp);
            ct_load_tile(p.z, p.z_stride, col, zv, stg, lane, rp);
          } else if (valid) {
            ct_load_row(p.h, p.h_stride, pix, col, hv);
            ct_load_row(p.z, p.z_stride, pix, col, zv);
          }
        } else if (EPI == CT_EPI_HEAD_MEAN) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 w4 = __ldg(reinterpret_cast<const float4*>(p.w1x1 + col) + j);
            hv[4 * j] = w4.x; hv[4 * j + 1] = w4.y; hv[4 * j + 2] = w4.z; hv[4 * j + 3] = w4.w;
          }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 b4 = reinterpret_cast<const float4*>(s_bias + col)[j];
          bv[4 * j] = b4.x; bv[4 * j + 1] = b4.y; bv[4 * j + 2] = b4.z; bv[4 * j + 3] = b4.w;
        }
        ptx::tmem_ld_32x32(tmem_base + lane_base + acc_col + uint32_t(c * 32), v);
        if (c + 2 >= chunks && last_sub) {  // this warp's last chunk: its share of the accumulator is read, hand it back
          ptx::tc_fence_before_sync();
          if (PAIR)
            ptx::mbar_arrive_cluster(lead_tempty[buf]);   // the leader's MMA warp waits for both CTAs' epilogues
          else
            ptx::mbar_arrive(tempty + buf);
        }
        if (EPI == CT_EPI_BIAS_ACT) {
          const int nv = min(32, p.cout - col);
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            float a = v[j] + bv[j];
            if (p.relu) a = fmaxf(a, 0.f);
            v[j] = separator ? 0.f : a * p.scale;
          }
          if (p.residual && valid && nv == 32) {   // fp16 residual, 64 contiguous bytes of this thread's pixel
            const uint4* rs = reinterpret_cast<const uint4*>(reinterpret_cast<const __half*>(p.residual) + pix * p.residual_stride + col);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint4 u = __ldg(rs + j);
              const __half2* hp = reinterpret_cast<const __half2*>(&u);
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float2 f = __half22float2(hp[e]);
                v[8 * j + 2 * e] = fmaxf(v[8 * j + 2 * e] + f.x, 0.f);
                v[8 * j + 2 * e + 1] = fmaxf(v[8 * j + 2 * e + 1] + f.y, 0.f);
              }
            }
          }
          if (p.dbg & 8) {   // timing experiment: nothing is stored
          } else if (co && nv > 0) {   // warp-uniform
            ct_store_tile(p.out, p.out_stride, col, v, nv, p.out_f16 != 0, stg, lane, rp,
                          (p.stats && tc.n < 2) ? s_stat + ((warp - 2) * 2 + tc.n) * 2 * CT_STAT_C : nullptr, CT_STAT_C);
            if (p.out2) ct_store_tile(p.out2, p.out2_stride, col, v, nv, p.out2_f16 != 0, stg, lane, rp);
          } else if (valid && nv > 0) {
            ct_store_row(p.out, p.out_stride, pix, col, v, nv, p.out_f16 != 0);
            if (p.out2) ct_store_row(p.out2, p.out2_stride, pix, col, v, nv, p.out2_f16 != 0);
          }
        } else if (EPI == CT_EPI_GRU_ZR) {
          const int hid = p.bn >> 1;
          if (col < hid) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = ct_sigmoid(v[j] + bv[j]);
            if (co)
              ct_store_tile(p.z, p.z_stride, col, v, 32, false, stg, lane, rp);
            else if (valid)
              ct_store_row(p.z, p.z_stride, pix, col, v, 32, false);
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = ct_sigmoid(v[j] + bv[j]) * hv[j];
            if (co)
              ct_store_tile(p.out, p.out_stride, col - hid, v, 32, p.out_f16 != 0, stg, lane, rp);
            else if (valid)
              ct_store_row(p.out, p.out_stride, pix, col - hid, v, 32, p.out_f16 != 0);
          }
        } else if (EPI == CT_EPI_GRU_Q) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = (1.f - zv[j]) * hv[j] + zv[j] * ct_tanh(v[j] + bv[j]);
          if (co) {
            ct_store_tile(p.out, p.out_stride, col, v, 32, false, stg, lane, rp);
            if (p.out2) ct_store_tile(p.out2, p.out2_stride, col, v, 32, p.out2_f16 != 0, stg, lane, rp);
          } else if (valid) {
            ct_store_row(p.out, p.out_stride, pix, col, v, 32, false);
            if (p.out2) ct_store_row(p.out2, p.out2_stride, pix, col, v, 32, p.out2_f16 != 0);
          }
        } else {  // CT_EPI_HEAD_MEAN
#pragma unroll
          for (int j = 0; j < 32; ++j) head_sum = fmaf(fmaxf(v[j] + bv[j], 0.f), hv[j], head_sum);
        }
      }
      if (chunks <= half && last_sub) {  // a one-chunk tile leaves the second warp of each pair without work
        ptx::tc_fence_before_sync();
        if (PAIR)
          ptx::mbar_arrive_cluster(lead_tempty[buf]);
        else
          ptx::mbar_arrive(tempty + buf);
      }
      if (EPI == CT_EPI_HEAD_MEAN) {
        // the two warps of a lane quadrant hold the even / odd chunks of the same pixels; the pixels of each image are
        // then summed in a fixed order
        s_red[half * 128 + m] = (valid && !separator) ? head_sum : 0.f;
        ptx::named_bar_sync(1, CT_EPI_THREADS);
        const int et = threadIdx.x - 64;
        const int img_first = tc.y0 / p.row_period;
        const int img = img_first + et;
        const int y_lo = max(img * p.row_period, tc.y0), y_hi = min(min(img * p.row_period + p.row_valid, tc.y0 + p.bh), p.H);
        if (et < 4 && y_lo < y_hi && tc.n < p.NB) {
          float s = 0.f;
          for (int r = (y_lo - tc.y0) * p.pw; r < (y_hi - tc.y0) * p.pw; ++r) s += s_red[r] + s_red[128 + r];
          s *= p.head_scale;
          if (y_lo == img * p.row_period && tc.c0 == 0) s += p.head_bias;  // one contributor brings the 1x1 bias
          // at most two tiles hold pixels of one image and the destination starts at zero: 0 + a + b == 0 + b + a
          atomicAdd(p.head_out + img, s);
        }
        ptx::named_bar_sync(1, CT_EPI_THREADS);
      }
      }   // sub
      if (leader && it == 0) ct_stamp(p, 5);
      if (leader) ct_tile_stamp(p, it, 6);
    }
  }

  ptx::tc_fence_before_sync();
  if (PAIR)
    ptx::cluster_sync();   // neither CTA may leave while the other can still read its shared memory or signal its barriers
  else
    __syncthreads();
  if (p.stats) {   // this CTA's partial sums, image by image: [n][sum (cout) | squares (cout)]
    const int nimg = p.NB < 2 ? p.NB : 2;
    for (int i = threadIdx.x; i < nimg * 2 * p.cout; i += CT_THREADS) {
      const int n = i / (2 * p.cout), r = i - n * 2 * p.cout, part = r / p.cout, c = r - part * p.cout;
      float acc = 0.f;
      for (int w = 0; w < 8; ++w) acc += s_stat[((w * 2 + n) * 2 + part) * CT_STAT_C + c];   // fixed order over the warps
      p.stats[(size_t)blockIdx.x * nimg * 2 * p.cout + i] = acc;
    }
  }
  if (warp == 1) {
    ptx::tc_fence_after_sync();
    if (PAIR)
      ptx::tmem_dealloc_2sm(tmem_base, 512);
    else
      ptx::tmem_dealloc(tmem_base, 512);
  }
  if (threadIdx.x == 0) ct_stamp(p, 6);
}

__global__ void fill_kernel(float* dst, float value, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = value;
}

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------
typedef CUresult (*CtEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static CtEncodeFn ct_encode_fn() {
  static CtEncodeFn cached = [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      fn = nullptr;
    return reinterpret_cast<CtEncodeFn>(fn);
  }();
  return cached;
}

template <int EPI>
static int ct_launch(const ConvTcParams& p, int grid, int smem, cudaStream_t s) {
  void (*kern)(const ConvTcParams) =
      p.pair ? (p.f16 ? conv_tc_kernel<EPI, true, true> : conv_tc_kernel<EPI, true, false>)
             : (p.f16 ? conv_tc_kernel<EPI, false, true> : conv_tc_kernel<EPI, false, false>);
  WOFT_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, CT_SMEM_LIMIT));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(CT_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (p.pair) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = 2;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  if (p.pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  WOFT_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, p));
  return launch_status("conv_tc_kernel");
}

}  // namespace woft

extern "C" int woft_conv_tc(const woft_conv_tc_args* a, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(a, "woft_conv_tc: null argument block");
  WOFT_REQUIRE(a->in && a->weight_packed, "woft_conv_tc: null input or weight pointer");
  WOFT_REQUIRE(a->batch > 0 && a->height > 0 && a->width > 0 && a->cin > 0 && a->cout > 0, "woft_conv_tc: bad extent");
  WOFT_REQUIRE(a->kh >= 1 && a->kw >= 1 && a->kh * a->kw <= CT_MAX_TAPS, "woft_conv_tc: at most %d filter taps", CT_MAX_TAPS);
  WOFT_REQUIRE(a->epilogue >= CT_EPI_BIAS_ACT && a->epilogue <= CT_EPI_FLOW_TAPS, "woft_conv_tc: unknown epilogue %d",
               a->epilogue);
  const int esz = a->in_f16 ? 2 : 4;
  const int k_elems = 128 / esz;
  const int cin_pad = (a->cin + k_elems - 1) / k_elems * k_elems;
  WOFT_REQUIRE(a->cin_pad == cin_pad, "woft_conv_tc: weights must be packed with Cin padded to %d (got %d)", cin_pad,
               a->cin_pad);
  WOFT_REQUIRE(a->cout_pad % 16 == 0 && a->cout_pad >= a->cout && a->cout_pad <= CT_MAX_BIAS,
               "woft_conv_tc: Cout must be padded to a multiple of 16 (at most %d)", CT_MAX_BIAS);
  WOFT_REQUIRE((reinterpret_cast<uintptr_t>(a->in) & 15) == 0 && ((long long)a->in_stride * esz) % 16 == 0,
               "woft_conv_tc: input must be 16-byte aligned with a pixel stride that is a multiple of 16 bytes");
  WOFT_REQUIRE(a->in_stride >= a->cin, "woft_conv_tc: input pixel stride smaller than Cin");
  int dev = 0, major = 0;
  WOFT_CUDA_CHECK(cudaGetDevice(&dev));
  WOFT_CUDA_CHECK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  if (major != 10) return fail(WOFT_ERR_UNSUPPORTED, "tcgen05 path needs sm_100; device is sm_%d*", major);
  CtEncodeFn enc = ct_encode_fn();
  if (!enc) return fail(WOFT_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");

  ConvTcParams p{};
  // N tile: the whole of Cout when it fits one accumulator, else the largest of 256/192/128/... dividing it
  int bn = a->cout_pad;
  if (bn > 256) {
    bn = 0;
    for (int c = 256; c >= 16; c -= 16)
      if (a->cout_pad % c == 0) {
        bn = c;
        break;
      }
  }
  const int taps = a->kh * a->kw;
  const int fixed = 1024 + (CT_MAX_BIAS + 256) * 4 + 8 * ct_stage_bytes(a->epilogue) +
                    (a->epilogue == CT_EPI_FLOW_TAPS ? 8192 : a->stats ? 8 * 4 * CT_STAT_C * 4 : 0) + 256;
  int a_stage = CT_A_STAGE;
  if (a->halo) {
    // one activation box per channel block serves all taps: it must hold the 128 GEMM rows behind the last tap's offset
    WOFT_REQUIRE(a->box_w > 0 && a->box_h > 0, "woft_conv_tc: halo mode takes the output tile in box_w, box_h");
    WOFT_REQUIRE(a->halo == 1 || (a->halo == 2 && a->box_w % 8 == 0), "woft_conv_tc: the row-halo mode needs box_w % 8 == 0");
    const int pw = a->halo == 2 ? a->box_w : a->box_w + a->kw - 1, hh = a->box_h + a->kh - 1;
    WOFT_REQUIRE(pw * a->box_h <= 128 && pw <= 256 && hh <= 256, "woft_conv_tc: halo tile %d x %d does not fit 128 GEMM rows",
                 a->box_w, a->box_h);
    const int last = (a->kh - 1) * pw + (a->halo == 2 ? 0 : a->kw - 1);   // row offset of the last tap's operand
    const int rows = 128 + last > pw * hh ? 128 + last : pw * hh;
    a_stage = (rows * 128 + 1023) / 1024 * 1024;
  }
  WOFT_REQUIRE(bn >= 16 && bn % 32 == 0, "woft_conv_tc: Cout tile %d (need a multiple of 32 dividing the padded Cout)", bn);
  p.bn = bn;
  p.n_tiles = a->cout_pad / bn;
  p.halo = a->halo;
  p.kw = a->kw;
  p.a_stage = a_stage;
  // pixel box: bw x bh = 128 (or the caller's own, e.g. 9 x 14 rows of stacked 9-wide images)
  int bw = a->box_w, bh = a->box_h;
  if (bw <= 0 || bh <= 0) {
    long long best = -1;
    for (int cand = 128; cand >= 8; cand >>= 1) {
      const int cw = cand, ch = 128 / cand;
      const int s_ = a->stride > 1 ? a->stride : 1;
      const long long tiles = (long long)ceil_div((a->width + 2 * a->pad_w - a->kw) / s_ + 1, cw) *
                              ceil_div((a->height + 2 * a->pad_h - a->kh) / s_ + 1, ch);
      if (best < 0 || tiles < best) {
        best = tiles;
        bw = cw;
        bh = ch;
      }
    }
  }
  WOFT_REQUIRE(bw >= 1 && bh >= 1 && bw <= 256 && bh <= 256 && bw * bh <= 128 && (a->halo || bw * bh >= 96),
               "woft_conv_tc: pixel box %d x %d must cover 96..128 pixels", bw, bh);
  p.bw = bw;
  p.bh = bh;
  p.pw = a->halo == 1 ? bw + a->kw - 1 : bw;
  p.hh = a->halo ? bh + a->kh - 1 : bh;
  const int stride = a->stride > 1 ? a->stride : 1;
  WOFT_REQUIRE(stride <= 2 && !(a->halo && stride != 1), "woft_conv_tc: stride 1 or 2 (1 in halo mode)");
  const int wo = (a->width + 2 * a->pad_w - a->kw) / stride + 1, ho = (a->height + 2 * a->pad_h - a->kh) / stride + 1;
  WOFT_REQUIRE(wo >= 1 && ho >= 1, "woft_conv_tc: empty output");
  p.stride = stride;
  p.W = wo;   // the tile grid, the validity tests and the pixel indices of the epilogue run over OUTPUT pixels
  p.H = ho;
  p.NB = a->batch;
  p.x_tiles = ceil_div(wo, bw);
  p.y_tiles = ceil_div(ho, bh);
  const long long total = (long long)p.x_tiles * p.y_tiles * a->batch * p.n_tiles;
  WOFT_REQUIRE(total < (1LL << 31), "woft_conv_tc: too many tiles");
  p.total_tiles = (int)total;
  // pair mode: two CTAs of a cluster share the filter tile, each loading half of its rows (cta_group::2 MMA, M = 256)
  p.pair = (a->pair && !p.halo && bn % 32 == 0 && !a->stats) ? 1 : 0;   // the statistics are gathered by single CTAs
  p.pix_tiles = p.x_tiles * p.y_tiles * a->batch;
  p.pair_tiles = (p.pix_tiles + 1) / 2 * p.n_tiles;
  p.taps = taps;
  for (int ky = 0; ky < a->kh; ++ky)
    for (int kx = 0; kx < a->kw; ++kx) {
      p.dx[ky * a->kw + kx] = (signed char)(kx - a->pad_w);
      p.dy[ky * a->kw + kx] = (signed char)(ky - a->pad_h);
      p.tap_row[ky * a->kw + kx] = (short)(a->halo == 2 ? ky * p.pw : ky * p.pw + kx);
    }
  p.k_elems = k_elems;
  p.k_chunks = cin_pad / k_elems;
  p.f16 = a->in_f16 ? 1 : 0;
  p.a_bytes = p.pw * p.hh * 128;
  p.b_bytes = (p.pair ? bn / 2 : bn) * 128;
  p.cout = a->cout;
  p.relu = a->relu;
  p.scale = a->scale;
  p.bias = a->bias;
  p.out = a->out;
  p.out_stride = a->out_stride;
  p.out_f16 = a->out_f16;
  p.out2 = a->out2;
  p.out2_stride = a->out2_stride;
  p.out2_f16 = a->out2_f16;
  p.h = a->h;
  p.h_stride = a->h_stride;
  p.z = a->z;
  p.z_stride = a->z_stride;
  p.w1x1 = a->w1x1;
  p.head_out = a->head_out;
  p.row_period = a->row_period;
  p.row_valid = a->row_valid;
  p.trace = static_cast<unsigned long long*>(a->debug_trace);
  p.stats = a->stats;
  p.residual = a->residual;
  p.residual_stride = a->residual_stride;
  p.dbg = a->debug_flags;
  static const bool fast_gates_env = [] {   // WOFT_CONV_FAST_GATES=0: ex2/rcp-based GRU gates instead of tanh.approx
    const char* e = getenv("WOFT_CONV_FAST_GATES");
    return !(e && e[0] == '0');
  }();
  if (fast_gates_env) p.dbg |= 64;
  p.pdl = (a->pdl && pdl_enabled()) ? 1 : 0;

  const int oesz = a->out_f16 ? 2 : 4;
  if (a->epilogue == CT_EPI_BIAS_ACT) {
    WOFT_REQUIRE(!a->out2 || a->out2_f16 == a->out_f16, "woft_conv_tc: both outputs of this epilogue share one type");
    WOFT_REQUIRE(!a->stats || (a->out_f16 && !a->out2 && a->batch <= 2 && a->stats_ctas >= 1 && a->cout_pad <= CT_STAT_C),
                 "woft_conv_tc: statistics are gathered for a single fp16 output of at most two images and %d channels", CT_STAT_C);
    WOFT_REQUIRE(!a->residual || ((reinterpret_cast<uintptr_t>(a->residual) & 15) == 0 && a->residual_stride % 8 == 0 &&
                                  a->cout % 32 == 0),
                 "woft_conv_tc: the residual is fp16, 16-byte aligned, with Cout a multiple of 32");
    // rows that are not 16-byte aligned (odd strides / offsets) are stored element by element
    WOFT_REQUIRE(a->out && (reinterpret_cast<uintptr_t>(a->out) % oesz) == 0 && a->out_stride >= a->cout,
                 "woft_conv_tc: output missing, misaligned or narrower than Cout");
    WOFT_REQUIRE(!a->out2 || ((reinterpret_cast<uintptr_t>(a->out2) % oesz) == 0 && a->out2_stride >= a->cout),
                 "woft_conv_tc: second output misaligned or narrower than Cout");
  } else if (a->epilogue == CT_EPI_GRU_ZR) {
    WOFT_REQUIRE(p.n_tiles == 1 && bn % 64 == 0 && a->cout == bn, "woft_conv_tc: GRU gate epilogue takes the stacked z|r "
                 "filters as one tile (Cout = 2 * hidden, a multiple of 64, at most 256)");
    WOFT_REQUIRE(a->h && a->z && a->out, "woft_conv_tc: GRU gate epilogue needs h, z and the r*h output");
    WOFT_REQUIRE(((reinterpret_cast<uintptr_t>(a->h) | reinterpret_cast<uintptr_t>(a->z)) & 15) == 0 &&
                 a->h_stride % 4 == 0 && a->z_stride % 4 == 0, "woft_conv_tc: GRU operands must be 16-byte aligned");
  } else if (a->epilogue == CT_EPI_GRU_Q) {
    WOFT_REQUIRE(p.n_tiles == 1 && a->cout == bn, "woft_conv_tc: GRU update epilogue takes Cout = hidden as one tile");
    WOFT_REQUIRE(a->h && a->z && a->out && !a->out_f16, "woft_conv_tc: GRU update epilogue needs h, z and the output (fp32)");
    WOFT_REQUIRE(((reinterpret_cast<uintptr_t>(a->h) | reinterpret_cast<uintptr_t>(a->z) | reinterpret_cast<uintptr_t>(a->out)) & 15) == 0 &&
                 a->h_stride % 4 == 0 && a->z_stride % 4 == 0 && a->out_stride % 4 == 0,
                 "woft_conv_tc: GRU operands must be 16-byte aligned");
  } else if (a->epilogue == CT_EPI_FLOW_TAPS) {
    WOFT_REQUIRE(p.n_tiles == 1 && a->cout == bn && a->cout * CT_TAPS_W * 4 <= 8 * 2048 + 8192 && !a->halo,
                 "woft_conv_tc: the flow-head epilogue takes Cout <= 256 as one tile");
    WOFT_REQUIRE(a->w1x1 && a->head_out && (reinterpret_cast<uintptr_t>(a->head_out) & 7) == 0 &&
                 (reinterpret_cast<uintptr_t>(a->w1x1) & 15) == 0,
                 "woft_conv_tc: the flow-head epilogue needs the second-layer filter (Cout x 24 floats) and the tap buffer");
  } else {
    WOFT_REQUIRE(a->w1x1 && a->head_out && a->row_period > 0 && a->row_valid > 0 && a->row_valid <= a->row_period,
                 "woft_conv_tc: patch-mean epilogue needs w1x1, head_out and the row period");
    // every image may receive at most two partial sums (their order then does not matter): either one Cout tile and
    // tiles that straddle images, or image-aligned tiles and two Cout tiles
    WOFT_REQUIRE(p.n_tiles == 1 || (p.n_tiles == 2 && bh == a->row_period),
                 "woft_conv_tc: patch-mean epilogue with two filter tiles needs one image per tile (box_h == row_period)");
    WOFT_REQUIRE(a->batch == 1 && a->height % a->row_period == 0, "woft_conv_tc: patch-mean epilogue takes the images "
                 "stacked along H (batch 1, H = images * row_period)");
    WOFT_REQUIRE(ceil_div(bh, a->row_period) + 1 <= 4, "woft_conv_tc: pixel box spans too many images");
    p.head_scale = 1.0f / float(a->row_valid * a->width);
    p.head_bias = a->head_bias;
  }
  if (a->bias == nullptr) p.bias = nullptr;
  {
    auto ok = [](const void* ptr, long long stride, int esz) {
      return ptr == nullptr || ((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (stride * esz) % 16 == 0);
    };
    p.coalesce = ok(a->out, a->out_stride, oesz) && ok(a->out2, a->out2_stride, a->out2_f16 ? 2 : 4) &&
                 ok(a->h, a->h_stride, 4) && ok(a->z, a->z_stride, 4);
    WOFT_REQUIRE((long long)a->batch * a->height * a->width < (1LL << 31), "woft_conv_tc: too many pixels");
  }

  const CUtensorMapDataType dt = a->in_f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  {
    cuuint64_t dims[4] = {(cuuint64_t)a->cin, (cuuint64_t)a->width, (cuuint64_t)a->height, (cuuint64_t)a->batch};
    cuuint64_t strides[3] = {(cuuint64_t)a->in_stride * esz, (cuuint64_t)a->in_stride * esz * a->width,
                             (cuuint64_t)a->in_stride * esz * a->width * a->height};
    // with an element stride the box extent is given in input pixels: every stride-th one is fetched
    cuuint32_t box[4] = {(cuuint32_t)k_elems, (cuuint32_t)(p.pw * stride), (cuuint32_t)(p.hh * stride), 1};
    cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
    if (enc(&p.tmA, dt, 4, const_cast<void*>(a->in), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return fail(WOFT_ERR_CUDA, "cuTensorMapEncodeTiled failed for the activation operand");
  }
  {
    cuuint64_t dims[3] = {(cuuint64_t)cin_pad, (cuuint64_t)a->cout_pad, (cuuint64_t)p.taps};
    cuuint64_t strides[2] = {(cuuint64_t)cin_pad * esz, (cuuint64_t)cin_pad * esz * a->cout_pad};
    cuuint32_t box[3] = {(cuuint32_t)k_elems, (cuuint32_t)(p.pair ? bn / 2 : bn), 1};
    cuuint32_t estr[3] = {1, 1, 1};
    if (enc(&p.tmB, dt, 3, const_cast<void*>(a->weight_packed), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return fail(WOFT_ERR_CUDA, "cuTensorMapEncodeTiled failed for the weight operand");
  }

  int sms = 0;
  WOFT_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  // Two pixel tiles per k-step (M = 256 per CTA) when the launch is persistent and the filter tile is narrow: one filter
  // tile, one barrier round trip and one commit serve eight MMAs instead of four.  The many-tile launches are bound by
  // L2 -> SM operand traffic (weight head: 32 KB per 384-cycle k-step on 148 SMs = 16.8 TB/s) and by the issue thread;
  // this takes a quarter of the bytes and half of the waits away.  debug_flags 1024 / 2048 force it off / on (tests).
  static const bool mt2_env = [] {
    const char* e = getenv("WOFT_CONV_MT2");
    return !(e && e[0] == '0');
  }();
  const bool mt2_fit = !p.halo && !p.pair && p.n_tiles == 1 && bn <= 128 &&
                       (a->epilogue == CT_EPI_BIAS_ACT || a->epilogue == CT_EPI_HEAD_MEAN);
  p.super_tiles = (p.total_tiles + 1) / 2;
  // ... unless the pairing costs a round: a CTA's share is ceil(tiles / SMs) single tiles or ceil(pairs / SMs) pairs
  // (encoder layer 1: 3 600 tiles = 25 rounds of one or 13 rounds of two -- measured 54.2 vs 53.8 us; the weight head's
  // 10 286 tiles: 70 rounds either way, 361 -> 329 us)
  const bool mt2_rounds = 2 * ceil_div(p.super_tiles, sms) <= ceil_div(p.total_tiles, sms);
  p.mt2 = (mt2_fit && !(a->debug_flags & 1024) &&
           ((mt2_env && p.total_tiles >= 2 * sms && mt2_rounds) || (a->debug_flags & 2048))) ? 1 : 0;
  if (p.mt2) {
    a_stage = 2 * CT_A_STAGE;
    p.a_stage = a_stage;
  }
  // halo mode: 3 activation stages (one box feeds `taps` MMA groups) and as many filter stages as fit beside them
  const int a_stages_halo = 3;
  int stages = p.halo ? (CT_SMEM_LIMIT - fixed - a_stages_halo * a_stage) / p.b_bytes : (CT_SMEM_LIMIT - fixed) / (a_stage + p.b_bytes);
  if (stages > CT_MAX_STAGES) stages = CT_MAX_STAGES;
  WOFT_REQUIRE(stages >= 2, "woft_conv_tc: tile does not fit shared memory");
  p.stages = stages;
  p.a_stages = p.halo ? a_stages_halo : stages;
  static const int producers_env = [] {   // WOFT_CONV_PRODUCERS: TMA-issuing warps of the plain mode (1..3)
    const char* e = getenv("WOFT_CONV_PRODUCERS");
    const int v = e ? atoi(e) : CT_MAX_PRODUCERS;
    return v < 1 ? 1 : v > CT_MAX_PRODUCERS ? CT_MAX_PRODUCERS : v;
  }();
  // several TMA warps pay off when a CTA walks many tiles (the issue latency of one thread bounds the k-step rate); with
  // one tile per CTA they only delay the first stage's arrival (three boxes share the start-up bandwidth): 0.3 us later
  p.producers = 1;   // decided below, once the grid is known
  // always more than half of the SM's shared memory: one CTA per SM, which owns all 512 TMEM columns
  int smem = p.a_stages * a_stage + stages * p.b_bytes + fixed;
  if (smem < 120 * 1024) smem = 120 * 1024;

  int grid = p.total_tiles < sms ? p.total_tiles : sms;
  if (p.mt2) grid = p.super_tiles < sms ? p.super_tiles : sms;
  if (p.pair) {
    const int pairs = p.pair_tiles < sms / 2 ? p.pair_tiles : sms / 2;
    grid = 2 * pairs;
  }
  WOFT_REQUIRE(grid >= 1, "woft_conv_tc: empty grid");
  if (!p.halo && (p.mt2 ? p.super_tiles : p.total_tiles) > grid) p.producers = producers_env < stages ? producers_env : stages;
  if (a->stats) {
    // every CTA writes its partial sums; the rows of CTAs that do not exist in this launch are cleared
    WOFT_REQUIRE(p.coalesce && !p.pair, "woft_conv_tc: statistics need 16-byte aligned outputs and single-CTA mode");
    WOFT_REQUIRE(grid <= a->stats_ctas, "woft_conv_tc: the statistics buffer holds %d CTA rows, the launch has %d", a->stats_ctas, grid);
    const size_t row = (size_t)a->batch * 2 * a->cout * sizeof(float);
    if (grid < a->stats_ctas)
      WOFT_CUDA_CHECK(cudaMemsetAsync(a->stats + (size_t)grid * a->batch * 2 * a->cout, 0, (a->stats_ctas - grid) * row,
                                      static_cast<cudaStream_t>(stream)));
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (a->epilogue) {
    case CT_EPI_BIAS_ACT:
      return ct_launch<CT_EPI_BIAS_ACT>(p, grid, smem, s);
    case CT_EPI_GRU_ZR:
      return ct_launch<CT_EPI_GRU_ZR>(p, grid, smem, s);
    case CT_EPI_GRU_Q:
      return ct_launch<CT_EPI_GRU_Q>(p, grid, smem, s);
    case CT_EPI_FLOW_TAPS:
      return ct_launch<CT_EPI_FLOW_TAPS>(p, grid, smem, s);
    default: {
      // the logits start at zero; the kernel adds each image's patch mean in (at most) two parts
      const long long n_img = a->height / a->row_period;
      fill_kernel<<<(unsigned)((n_img + 255) / 256), 256, 0, s>>>(a->head_out, 0.f, n_img);
      int rc = launch_status("fill_kernel");
      if (rc != WOFT_OK) return rc;
      return ct_launch<CT_EPI_HEAD_MEAN>(p, grid, smem, s);
    }
  }
}

extern "C" int woft_conv_tc_check(void* stream) {
  using namespace woft;
  int flag = 0;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  WOFT_CUDA_CHECK(cudaMemcpyFromSymbolAsync(&flag, g_conv_abort, sizeof(int), 0, cudaMemcpyDeviceToHost, s));
  WOFT_CUDA_CHECK(cudaStreamSynchronize(s));
  if (flag != 0) return fail(WOFT_ERR_CUDA, "convolution kernel aborted: a pipeline barrier timed out (flag=%d)", flag);
  return WOFT_OK;
}
