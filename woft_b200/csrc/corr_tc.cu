// K1 + K2 on the Blackwell tensor cores: all-pairs correlation as a tcgen05/TMEM GEMM fed by TMA, with the
// 4-level average-pooled pyramid formed in the epilogue.
//
// Replaces CorrBlock.corr (torch.matmul + a separate divide pass) and the three avg_pool2d passes of
// CorrBlock.__init__ (pytracking/external/RAFT/raft_core/corr.py:13-27, 61-69): the reference writes the
// 829 MB level-0 volume, re-reads and re-writes it for the 1/sqrt(C) scaling, and re-reads each level to pool
// the next.  Here every level is written exactly once and nothing is re-read.
//
// Data layout
//   fmaps arrive NCHW, i.e. both GEMM operands are (C, HW) with HW contiguous ("MN-major").  pack_kernel
//   transposes them once into K-major rows (HW, Kp) in the caller's workspace -- rounded to tf32 (one plane),
//   split into tf32 hi + lo planes (3xTF32), or converted to bf16 -- so TMA can fetch 128-byte swizzled K rows.
//   Output levels keep the reference layout (B*H*W, 1, H_l, W_l) fp32.
//
// Tiling
//   One output tile = 128 consecutive query pixels (UMMA M) x an 8 x 32 patch of the (h2, w2) plane
//   (UMMA N = 256).  The patch is fetched by ONE 4-D TMA box over (K, W2, H2, B) so its 256 rows sit in shared
//   memory in (h, w) order and whole 2x2 / 4x4 / 8x8 pooling cells live inside one accumulator row.
//   floor() semantics of avg_pool2d: a level-l cell exists iff its index is < H_l, W_l at that level; TMA
//   zero-fills reads past the map, and stores past it are clipped (columns, queries) or skipped (rows).
//
// Warp roles (192 threads, persistent CTAs, static round-robin over tiles)
//   warp 0      TMA producer: A box (Kblk x 128 queries) + B box (Kblk x 32 x 8) per stage, 3 stages of 48 KB
//   warp 1      allocates all 512 TMEM columns (2 accumulators of 128 lanes x 256 fp32 columns), issues
//               tcgen05.mma (kind::tf32 or kind::f16, M=128 N=256), commits stages back to the producer
//   warps 2-5   epilogue: tcgen05.ld one (h) row of 32 columns at a time, scale by 1/sqrt(C), stage the row in
//               a 3-slot shared-memory ring (128-byte swizzled) and TMA-store it (box 32 x 1 x 128 queries:
//               one full 128-byte line per query); 2x2 sums are carried in registers and the three pooled
//               levels leave through the same ring at the end of the tile.  Accumulator i+1 is being
//               computed while accumulator i drains.
// Every mbarrier wait has a 2 s timeout that raises an abort flag in the workspace instead of hanging the GPU.
//
// Algorithmic bytes per launch: 2 * B*C*HW*4 (fmaps in) + sum_l B*HW*H_l*W_l*4 (levels out); 2*B*HW^2*C flop
// (x3 for WOFT_CORR_TF32X3).  At 1280x720 (HW = 14 400, C = 256): 1.10 GB, 106 GFLOP -> HBM-write bound.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstdlib>

#include "common.cuh"
#include "ptx.cuh"

namespace woft {

constexpr int BM = 128;
constexpr int TH = 8, TW = 32;
constexpr int BN = TH * TW;
constexpr int A_STAGE = BM * 128;  // bytes: 128 rows of one 128-byte K row
constexpr int B_STAGE = BN * 128;
constexpr int MAX_STAGES = 4;
constexpr int SLOT = 16384;
constexpr int TC_THREADS = 192;
constexpr int EPI_THREADS = 128;
constexpr int TC_SMEM_LIMIT = 232448;  // 227 KB opt-in maximum per CTA
constexpr int kMaxTcLevels = 4;

struct CorrTcParams {
  CUtensorMap tmA;                   // (Kp, HW, B*planes), box (Kblk, 128, 1)
  CUtensorMap tmB;                   // (Kp, W, H, B*planes), box (Kblk, 32, 8, 1)
  CUtensorMap tmOut[kMaxTcLevels];   // (W_l, H_l, HW, B), box (32 >> l, 1, 128, 1)
  float* level[kMaxTcLevels];   // __half* when out_f16
  int lh[kMaxTcLevels];
  int lw[kMaxTcLevels];
  int tma_ok[kMaxTcLevels];
  int num_levels;
  int out_f16;   // levels stored as fp16 (WOFT_CORR_STORE_F16)
  int B, HW, H, W;
  int m_tiles, h_tiles, w_tiles, total_tiles;
  int k_blocks;   // 128-byte K rows per plane
  int k_elems;    // elements per 128-byte K row (32 tf32 / 64 bf16)
  int n_pairs;    // 1, or 3 for the hi/lo split
  int bf16;       // 16-bit operands (kind::f16): 1 = bf16, 2 = fp16; 0 = tf32 (kind::tf32)
  int a_resident; // 1: the A tile (all k blocks) stays in shared memory while a CTA walks `unit` consecutive N tiles
  int unit;       // tiles per work unit (consecutive tile ids share b and the M tile)
  int stages;     // operand stages (A+B, or B only when A is resident)
  int nslot;      // epilogue staging ring slots
  float scale;
  int* abort_flag;
  unsigned long long* trace;  // debug (WOFT_CORR_TRACE): [cta < 4][tile < 64][6] globaltimer stamps, else null
};

template <bool TRACE>
__device__ __forceinline__ void trace_stamp(const CorrTcParams& p, int it, int slot) {
  if (TRACE && p.trace && blockIdx.x < 4 && it < 64) p.trace[(blockIdx.x * 64 + it) * 6 + slot] = ptx::globaltimer_ns();
}

// ---------------------------------------------------------------------------------------------------
// pack: (B, C, HW) fp32 -> K-major (planes, B, HW, Kp)
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ float round_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

template <int MODE>  // 1 tf32, 2 tf32 hi/lo, 3 bf16, 4 fp16, 5 fp16 hi/lo
__global__ void __launch_bounds__(256) pack_kernel(const float* __restrict__ src1, const float* __restrict__ src2,
                                                   void* __restrict__ dst1, void* __restrict__ dst2, int B, int C,
                                                   int HW, int Kp, size_t plane_elems, int* flag) {
  __shared__ float tile[32][33];
  if (flag && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x == 0) *flag = 0;
  const bool second = (int)blockIdx.z >= B;  // grid.z = 2B: both feature maps in one launch
  const float* __restrict__ src = second ? src2 : src1;
  void* __restrict__ dst = second ? dst2 : dst1;
  const int b = second ? blockIdx.z - B : blockIdx.z;
  const int hw0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  const float* s = src + (size_t)b * C * HW;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = c0 + ty + i * 8, hw = hw0 + tx;
    tile[ty + i * 8][tx] = (c < C && hw < HW) ? __ldg(s + (size_t)c * HW + hw) : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int hw = hw0 + ty + i * 8, c = c0 + tx;
    if (hw >= HW || c >= Kp) continue;
    const float v = tile[tx][ty + i * 8];
    const size_t o = ((size_t)b * HW + hw) * Kp + c;
    if (MODE == 3) {
      reinterpret_cast<__nv_bfloat16*>(dst)[o] = __float2bfloat16_rn(v);
    } else if (MODE == 4 || MODE == 5) {
      const __half hi = __float2half_rn(fminf(fmaxf(v, -65504.f), 65504.f));  // saturate, no inf
      reinterpret_cast<__half*>(dst)[o] = hi;
      if (MODE == 5) reinterpret_cast<__half*>(dst)[plane_elems + o] = __float2half_rn(v - __half2float(hi));
    } else {
      const float hi = round_tf32(v);
      reinterpret_cast<float*>(dst)[o] = hi;
      if (MODE == 2) reinterpret_cast<float*>(dst)[plane_elems + o] = round_tf32(v - hi);
    }
  }
}

// The same packing for channels-last feature maps (B, HW, C): the K-major rows are already there, only the rounding /
// splitting and the padding of C to Kp remain.  One thread per packed element, consecutive threads along K.
template <int MODE>
__global__ void __launch_bounds__(256) pack_nhwc_kernel(const float* __restrict__ src1, const float* __restrict__ src2,
                                                        void* __restrict__ dst1, void* __restrict__ dst2, int C, int Kp,
                                                        size_t rows, size_t plane_elems, int* flag) {
  if (flag && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) *flag = 0;
  const float* __restrict__ src = blockIdx.y ? src2 : src1;
  void* __restrict__ dst = blockIdx.y ? dst2 : dst1;
  const size_t total = rows * Kp;
  for (size_t o = (size_t)blockIdx.x * 256 + threadIdx.x; o < total; o += (size_t)gridDim.x * 256) {
    const size_t row = o / Kp;
    const int c = int(o - row * Kp);
    const float v = c < C ? __ldg(src + row * C + c) : 0.f;
    if (MODE == 3) {
      reinterpret_cast<__nv_bfloat16*>(dst)[o] = __float2bfloat16_rn(v);
    } else if (MODE == 4 || MODE == 5) {
      const __half hi = __float2half_rn(fminf(fmaxf(v, -65504.f), 65504.f));
      reinterpret_cast<__half*>(dst)[o] = hi;
      if (MODE == 5) reinterpret_cast<__half*>(dst)[plane_elems + o] = __float2half_rn(v - __half2float(hi));
    } else {
      const float hi = round_tf32(v);
      reinterpret_cast<float*>(dst)[o] = hi;
      if (MODE == 2) reinterpret_cast<float*>(dst)[plane_elems + o] = round_tf32(v - hi);
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// GEMM + pooled epilogue
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool wait_or_abort(uint64_t* bar, uint32_t parity, int* abort_flag) {
  if (ptx::mbar_try_wait(bar, parity)) return true;
  const uint64_t t0 = ptx::globaltimer_ns();
  uint32_t spins = 0;
  while (!ptx::mbar_try_wait(bar, parity)) {
    if ((++spins & 63u) == 0) {
      if (*reinterpret_cast<volatile int*>(abort_flag) != 0) return false;
      if (ptx::globaltimer_ns() - t0 > 2000000000ull) {
        atomicExch(abort_flag, 1);
        return false;
      }
    }
  }
  return true;
}

__device__ __forceinline__ uint64_t make_kmajor_sw128_desc(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);  // start address, 16-byte units
  d |= (uint64_t)1 << 16;                    // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;          // stride byte offset: 8 rows x 128 B between row groups
  d |= (uint64_t)1 << 46;                    // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                    // SWIZZLE_128B
  return d;
}

// after committing a store group: with 3 slots one older group may still be reading its slot, with 2 none may
__device__ __forceinline__ void store_ring_wait(int nslot) {
  if (nslot >= 3)
    ptx::tma_store_wait_read<1>();
  else
    ptx::tma_store_wait_read<0>();
}

struct TcKnobs {
  int tma_levels;
  bool a_resident;
  int unit;
};
const TcKnobs& tc_knobs();
// Per-role timeline buffer (woft_corr_set_trace_buffer): caller-allocated device memory, null = tracing off.
static unsigned long long* g_trace_buffer = nullptr;
void corr_tc_set_trace_buffer(unsigned long long* buf) { __atomic_store_n(&g_trace_buffer, buf, __ATOMIC_RELEASE); }

struct TileCoord {
  int b, m0, h0, w0;
};
__device__ __forceinline__ TileCoord decode_tile(const CorrTcParams& p, int t) {
  TileCoord c;
  const int wt = t % p.w_tiles;
  t /= p.w_tiles;
  const int ht = t % p.h_tiles;
  t /= p.h_tiles;
  const int mt = t % p.m_tiles;
  c.b = t / p.m_tiles;
  c.m0 = mt * BM;
  c.h0 = ht * TH;
  c.w0 = wt * TW;
  return c;
}

// swizzled staging of one pooled/unpooled row of NV fp32 values (NV*4 bytes per query) for a TMA store whose
// box is (NV, 1, 128): 16-byte chunk index XOR-ed the way CU_TENSOR_MAP_SWIZZLE_{128,64,32}B expects.
// the same for fp16 storage: NV halves = NV*2 bytes per query, 16-byte chunks of 8 values
template <int NV>
__device__ __forceinline__ void stage_row_h(uint8_t* sub_slot, int m, const float* v) {
  constexpr int RB = NV * 2, CH = RB / 16;
  uint8_t* row = sub_slot + m * RB;
  const int x = (RB == 64) ? ((m >> 1) & 3) : (RB == 32) ? ((m >> 2) & 1) : 0;
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    uint4 u;
    __half2 h0 = __floats2half2_rn(v[8 * c], v[8 * c + 1]), h1 = __floats2half2_rn(v[8 * c + 2], v[8 * c + 3]);
    __half2 h2 = __floats2half2_rn(v[8 * c + 4], v[8 * c + 5]), h3 = __floats2half2_rn(v[8 * c + 6], v[8 * c + 7]);
    u.x = *reinterpret_cast<uint32_t*>(&h0);
    u.y = *reinterpret_cast<uint32_t*>(&h1);
    u.z = *reinterpret_cast<uint32_t*>(&h2);
    u.w = *reinterpret_cast<uint32_t*>(&h3);
    *reinterpret_cast<uint4*>(row + ((c ^ x) << 4)) = u;
  }
}
template <int NV>
__device__ __forceinline__ void direct_row_h(float* level_, int lh, int lw, int HW, int b, int m, int h, int w0,
                                             const float* v) {
  if (m >= HW || h >= lh) return;
  __half* dst = reinterpret_cast<__half*>(level_) + (((size_t)b * HW + m) * lh + h) * lw;
  if ((lw & 3) == 0 && (reinterpret_cast<uintptr_t>(level_) & 7) == 0) {
    // rows are 8-byte aligned and a 4-value chunk is wholly inside or outside the map
#pragma unroll
    for (int j = 0; j < NV; j += 4)
      if (w0 + j < lw) {
        __half2 a = __floats2half2_rn(v[j], v[j + 1]), c = __floats2half2_rn(v[j + 2], v[j + 3]);
        uint2 u;
        u.x = *reinterpret_cast<uint32_t*>(&a);
        u.y = *reinterpret_cast<uint32_t*>(&c);
        *reinterpret_cast<uint2*>(dst + w0 + j) = u;
      }
    return;
  }
#pragma unroll
  for (int j = 0; j < NV; ++j)
    if (w0 + j < lw) dst[w0 + j] = __float2half_rn(v[j]);
}

template <int NV>
__device__ __forceinline__ void stage_row(uint8_t* sub_slot, int m, const float* v) {
  constexpr int CH = NV / 4;  // 16-byte chunks per query row
  uint8_t* row = sub_slot + m * (NV * 4);
  const int x = (NV == 32) ? (m & 7) : (NV == 16) ? ((m >> 1) & 3) : (NV == 8) ? ((m >> 2) & 1) : 0;
#pragma unroll
  for (int c = 0; c < CH; ++c)
    *reinterpret_cast<float4*>(row + ((c ^ x) << 4)) = make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
}

template <int NV>
__device__ __forceinline__ void direct_row(float* level, int lh, int lw, int HW, int b, int m, int h, int w0,
                                           const float* v) {
  if (m >= HW || h >= lh) return;
  float* dst = level + (((size_t)b * HW + m) * lh + h) * lw;
  if ((lw & 3) == 0 && (reinterpret_cast<uintptr_t>(level) & 15) == 0) {
    // rows are 16-byte aligned and a 4-float chunk is wholly inside or outside the map
#pragma unroll
    for (int j = 0; j < NV; j += 4)
      if (w0 + j < lw) *reinterpret_cast<float4*>(dst + w0 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
    return;
  }
#pragma unroll
  for (int j = 0; j < NV; ++j)
    if (w0 + j < lw) dst[w0 + j] = v[j];
}

// DEBUG = false is the production instantiation: streamed A tile, 3 operand stages, 3 ring slots, no tracing, all of
// them compile-time so the role loops stay tight.  DEBUG = true takes stages / slots / the resident-A schedule /
// the trace buffer from the parameters (experiments described in DESIGN.md 4).
// threads: TMA warp, MMA warp, four epilogue warps per group.  TWO: two epilogue groups -- the instantiation for fp16
// storage, whose tile is bound by the chain of store events (255 -> 214 us for the call); with the fp32 volume the kernel
// sits at the HBM write rate and two groups on two-slot rings measured slower (266 -> 317 us), so it keeps one group.
__host__ __device__ constexpr int tc_threads(bool two) { return two ? TC_THREADS + EPI_THREADS : TC_THREADS; }
// ring bytes: one group, NSLOT slots; two groups, two slots each + the 8 KB level-2 exchange buffer
__host__ __device__ constexpr int tc_ring_bytes(bool two, int nslot) { return two ? 4 * SLOT + 8192 : nslot * SLOT; }

template <bool DEBUG, bool TWO = false>
__global__ void __launch_bounds__(tc_threads(TWO), 1) corr_tc_kernel(const __grid_constant__ CorrTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr int EPI_GROUPS = TWO ? 2 : 1;
  const int STAGES = DEBUG ? p.stages : 3, NSLOT = TWO ? 2 : DEBUG ? p.nslot : 3;
  const bool a_resident = DEBUG && p.a_resident;
  const int unit = DEBUG ? p.unit : 1;
  uint8_t* sA = smem;  // streamed: STAGES x 16 KB; resident: k_blocks x 16 KB
  uint8_t* sB = sA + (a_resident ? p.k_blocks : STAGES) * A_STAGE;
  uint8_t* ring = sB + STAGES * B_STAGE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(ring + tc_ring_bytes(TWO, NSLOT));
  uint64_t* full = bars;                     // [MAX_STAGES] TMA -> MMA
  uint64_t* empty = bars + MAX_STAGES;       // [MAX_STAGES] MMA -> TMA
  uint64_t* tfull = bars + 2 * MAX_STAGES;   // [2] MMA -> epilogue
  uint64_t* tempty = tfull + 2;              // [2] epilogue -> MMA
  uint64_t* a_full = tempty + 2;             // resident A tile landed
  uint64_t* a_empty = a_full + 1;            // every MMA reading the resident A tile has completed
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(a_empty + 1);
  volatile int* epi_go = reinterpret_cast<volatile int*>(tmem_ptr + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(full + s, 1);
      ptx::mbar_init(empty + s, 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(tfull + i, 1);
      ptx::mbar_init(tempty + i, EPI_GROUPS * EPI_THREADS);
    }
    ptx::mbar_init(a_full, 1);
    ptx::mbar_init(a_empty, 1);
    ptx::fence_mbar_init();
    ptx::prefetch_tmap(&p.tmA);
    ptx::prefetch_tmap(&p.tmB);
    for (int l = 0; l < p.num_levels; ++l)
      if (p.tma_ok[l]) ptx::prefetch_tmap(&p.tmOut[l]);
  }
  if (warp == 1) ptx::tmem_alloc(tmem_ptr, 512);
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr;

  const int k_steps = p.k_blocks * p.n_pairs;

  if (warp == 0) {
    // ================================ TMA producer ================================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      bool alive = true;
      int cur_a = -1;          // (b, M tile) whose A tile is resident
      uint32_t a_loads = 0;
      int pit = 0;
      for (int chunk = blockIdx.x; chunk * unit < p.total_tiles && alive; chunk += gridDim.x) {
        const int t_end = min((chunk + 1) * unit, p.total_tiles);
        for (int t = chunk * unit; t < t_end && alive; ++t, ++pit) {
          const TileCoord tc = decode_tile(p, t);
          trace_stamp<DEBUG>(p, pit, 0);
          const int a_id = tc.b * p.m_tiles + tc.m0 / BM;
          const bool reload = a_resident && a_id != cur_a;
          cur_a = a_id;
          // with a resident A tile, first queue as many B stages of this tile as the ring takes, then swap A in
          const int pre = reload ? min(STAGES, k_steps) : 0;
          for (int ks = 0; ks < k_steps; ++ks) {
            if (reload && ks == pre) {
              if (!wait_or_abort(a_empty, (a_loads & 1u) ^ 1u, p.abort_flag)) {
                alive = false;
                break;
              }
              ptx::mbar_arrive_expect_tx(a_full, uint32_t(p.k_blocks) * A_STAGE);
              for (int kb = 0; kb < p.k_blocks; ++kb)
                ptx::tma_load_3d(sA + kb * A_STAGE, &p.tmA, a_full, kb * p.k_elems, tc.m0, tc.b);
              ++a_loads;
            }
            const int kb = ks / p.n_pairs, pr = ks - kb * p.n_pairs;
            const int plane_a = (pr == 2) ? 1 : 0, plane_b = (pr == 1) ? 1 : 0;  // hi*hi, hi*lo, lo*hi
            if (!wait_or_abort(empty + stage, phase ^ 1u, p.abort_flag)) {
              alive = false;
              break;
            }
            if (a_resident) {
              ptx::mbar_arrive_expect_tx(full + stage, B_STAGE);
            } else {
              ptx::mbar_arrive_expect_tx(full + stage, A_STAGE + B_STAGE);
              ptx::tma_load_3d(sA + stage * A_STAGE, &p.tmA, full + stage, kb * p.k_elems, tc.m0, plane_a * p.B + tc.b);
            }
            ptx::tma_load_4d(sB + stage * B_STAGE, &p.tmB, full + stage, kb * p.k_elems, tc.w0, tc.h0,
                             plane_b * p.B + tc.b);
            if (++stage == STAGES) {
              stage = 0;
              phase ^= 1u;
            }
          }
          trace_stamp<DEBUG>(p, pit, 1);
          if (reload && alive && pre == k_steps) {  // k_steps <= STAGES: the A swap comes after all B loads
            if (!wait_or_abort(a_empty, (a_loads & 1u) ^ 1u, p.abort_flag)) {
              alive = false;
              break;
            }
            ptx::mbar_arrive_expect_tx(a_full, uint32_t(p.k_blocks) * A_STAGE);
            for (int kb = 0; kb < p.k_blocks; ++kb)
              ptx::tma_load_3d(sA + kb * A_STAGE, &p.tmA, a_full, kb * p.k_elems, tc.m0, tc.b);
            ++a_loads;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    if (lane == 0) {
      const uint32_t fmt = p.bf16 == 1 ? 1u : p.bf16 == 2 ? 0u : 2u;  // F16F32Format: F16 = 0, BF16 = 1, TF32 = 2
      const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | (uint32_t(BN >> 3) << 17) | (uint32_t(BM >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0;
      bool alive = true;
      int it = 0;
      int cur_a = -1;
      uint32_t a_loads = 0;
      for (int chunk = blockIdx.x; chunk * unit < p.total_tiles && alive; chunk += gridDim.x) {
        const int t_end = min((chunk + 1) * unit, p.total_tiles);
        for (int t = chunk * unit; t < t_end && alive; ++t, ++it) {
          if (a_resident) {
            const TileCoord tc = decode_tile(p, t);
            const int a_id = tc.b * p.m_tiles + tc.m0 / BM;
            if (a_id != cur_a) {
              if (cur_a >= 0) ptx::mma_commit(a_empty);  // everything issued so far has read the old A tile
              cur_a = a_id;
              if (!wait_or_abort(a_full, a_loads & 1u, p.abort_flag)) break;
              ++a_loads;
              ptx::tc_fence_after_sync();
            }
          }
          const int buf = it & 1;
          const uint32_t use = uint32_t(it >> 1);
          if (!wait_or_abort(tempty + buf, (use & 1u) ^ 1u, p.abort_flag)) {
            alive = false;
            break;
          }
          ptx::tc_fence_after_sync();
          trace_stamp<DEBUG>(p, it, 2);
          const uint32_t d_tmem = tmem_base + uint32_t(buf * BN);
          for (int ks = 0; ks < k_steps; ++ks) {
            if (!wait_or_abort(full + stage, phase, p.abort_flag)) {
              alive = false;
              break;
            }
            ptx::tc_fence_after_sync();
            const int a_slot = a_resident ? ks : stage;  // resident: k block ks of the A tile (n_pairs == 1)
            const uint64_t da = make_kmajor_sw128_desc(ptx::smem_u32(sA + a_slot * A_STAGE));
            const uint64_t db = make_kmajor_sw128_desc(ptx::smem_u32(sB + stage * B_STAGE));
#pragma unroll
            for (int k = 0; k < 4; ++k) {  // 4 x 32 bytes of K per 128-byte row: UMMA_K = 8 (tf32) / 16 (bf16)
              const uint32_t acc = (ks > 0 || k > 0) ? 1u : 0u;
              if (p.bf16)
                ptx::mma_f16(d_tmem, da + uint64_t(2 * k), db + uint64_t(2 * k), idesc, acc);
              else
                ptx::mma_tf32(d_tmem, da + uint64_t(2 * k), db + uint64_t(2 * k), idesc, acc);
            }
            ptx::mma_commit(empty + stage);  // frees the stage once these MMAs have read it
            if (++stage == STAGES) {
              stage = 0;
              phase ^= 1u;
            }
          }
          if (alive) ptx::mma_commit(tfull + buf);
          trace_stamp<DEBUG>(p, it, 3);
        }
      }
    }
  } else if (warp < 2 + 4 * EPI_GROUPS) {
    // ================================ epilogue ================================
    // EPI_GROUPS groups of four warps (one warp per TMEM lane quadrant).  A group owns RPG = TH / EPI_GROUPS consecutive
    // patch rows of every tile -- whole 2x2 and 4x4 pooling cells, so levels 1 and 2 need nothing from the other group --
    // with its own store ring, named barrier and issuing thread.  The tile costs a chain of store events (TMEM read,
    // staging, proxy fence, a 128-thread barrier, one TMA store), 11 of them with one group: with half the bytes (fp16
    // storage) the kernel took the same time, i.e. the chain, not HBM, bounds it; two groups walk 6 events each side by
    // side.  Only level 3 (one 8x8 cell row per tile) needs both groups' level-2 rows: group 1 hands its row over
    // through shared memory and a 256-thread barrier, once per tile.
    constexpr int G = EPI_GROUPS, RPG = TH / G;
    const int g = (warp - 2) >> 2;
    const int et = threadIdx.x - 64 - g * EPI_THREADS;   // 0..127 within the group
    const bool issuer = (et == 0);
    const int m = (warp & 3) * 32 + lane;         // TMEM lane == query row of the tile
    const uint32_t lane_base = uint32_t((warp & 3) * 32) << 16;
    const float scale = p.scale;
    const bool f16o = p.out_f16 != 0;
    const int NS = G == 1 ? NSLOT : 2;            // ring slots of this group
    uint8_t* gring = ring + g * NS * SLOT;
    const uint32_t bar_id = 1u + uint32_t(g);
    float* xchg = reinterpret_cast<float*>(ring + G * NS * SLOT);   // [2][128][8] floats, G == 2 only
    uint32_t ev = 0;  // ring event counter (identical in all threads of the group)
    int it = 0;
    bool go = true;
    for (int chunk = blockIdx.x; chunk * unit < p.total_tiles && go; chunk += gridDim.x) {
    const int t_end = min((chunk + 1) * unit, p.total_tiles);
    for (int t = chunk * unit; t < t_end; ++t, ++it) {
      const TileCoord tc = decode_tile(p, t);
      const int buf = it & 1;
      const uint32_t use = uint32_t(it >> 1);
      if (issuer) epi_go[g] = wait_or_abort(tfull + buf, use & 1u, p.abort_flag) ? 1 : 0;
      ptx::named_bar_sync(bar_id, EPI_THREADS);
      if (epi_go[g] == 0) {
        go = false;
        break;
      }
      if (issuer && g == 0) trace_stamp<DEBUG>(p, it, 4);
      ptx::tc_fence_after_sync();
      const int mq = tc.m0 + m;
      const int r0 = g * RPG;   // first patch row of this group

      float hs[16];             // horizontal pair sums of the even row
      float l1[RPG / 2][16];    // level-1 rows of this group
      // the TMEM read of the next patch row is in flight while this one is staged, fenced, synchronised and stored
      uint32_t ra[32], rb[32];
      ptx::tmem_ld_32x32_issue(tmem_base + lane_base + uint32_t(buf * BN + r0 * TW), ra);
#pragma unroll
      for (int rr = 0; rr < RPG; ++rr) {
        const int r = r0 + rr;
        float v[32];
        if ((rr & 1) == 0) {
          ptx::tmem_ld_32x32_wait(ra);
          if (rr + 1 < RPG) ptx::tmem_ld_32x32_issue(tmem_base + lane_base + uint32_t(buf * BN + (r + 1) * TW), rb);
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(ra[j]);
        } else {
          ptx::tmem_ld_32x32_wait(rb);
          if (rr + 1 < RPG) ptx::tmem_ld_32x32_issue(tmem_base + lane_base + uint32_t(buf * BN + (r + 1) * TW), ra);
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(rb[j]);
        }
        if (rr == RPG - 1) {  // this group's rows of the accumulator are read: hand them back to the MMA warp
          ptx::tc_fence_before_sync();
          ptx::mbar_arrive(tempty + buf);
        }
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] *= scale;
        uint8_t* slot = gring + (ev % NS) * SLOT;
        if (p.tma_ok[0]) {
          if (f16o)
            stage_row_h<32>(slot, m, v);
          else
            stage_row<32>(slot, m, v);
          ptx::fence_proxy_async_smem();
        } else if (f16o) {
          direct_row_h<32>(p.level[0], p.lh[0], p.lw[0], p.HW, tc.b, mq, tc.h0 + r, tc.w0, v);
        } else {
          direct_row<32>(p.level[0], p.lh[0], p.lw[0], p.HW, tc.b, mq, tc.h0 + r, tc.w0, v);
        }
        ptx::named_bar_sync(bar_id, EPI_THREADS);
        if (issuer) {
          if (p.tma_ok[0] && tc.h0 + r < p.lh[0]) ptx::tma_store_4d(&p.tmOut[0], slot, tc.w0, tc.h0 + r, tc.m0, tc.b);
          ptx::tma_store_commit();
          store_ring_wait(NS);
        }
        ++ev;
        // 2x2 mean in avg_pool2d's summation order: ((a00 + a01) + a10) + a11, then * 0.25
        if ((rr & 1) == 0) {
#pragma unroll
          for (int j = 0; j < 16; ++j) hs[j] = v[2 * j] + v[2 * j + 1];
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) l1[rr >> 1][j] = ((hs[j] + v[2 * j]) + v[2 * j + 1]) * 0.25f;
        }
      }

      if (p.num_levels > 1) {
#pragma unroll
        for (int half = 0; half < RPG / 4; ++half) {  // level-1 rows in pairs
          uint8_t* slot = gring + (ev % NS) * SLOT;
#pragma unroll
          for (int rr = 0; rr < 2; ++rr) {
            const int i1 = half * 2 + rr;                     // row within the group
            const int h1 = tc.h0 / 2 + g * (RPG / 2) + i1;    // row of the level-1 map
            if (p.tma_ok[1]) {
              if (f16o)
                stage_row_h<16>(slot + rr * 8192, m, l1[i1]);
              else
                stage_row<16>(slot + rr * 8192, m, l1[i1]);
            } else if (f16o) {
              direct_row_h<16>(p.level[1], p.lh[1], p.lw[1], p.HW, tc.b, mq, h1, tc.w0 / 2, l1[i1]);
            } else {
              direct_row<16>(p.level[1], p.lh[1], p.lw[1], p.HW, tc.b, mq, h1, tc.w0 / 2, l1[i1]);
            }
          }
          if (p.tma_ok[1]) {  // (kernel parameter: uniform) no ring event when the level goes out by direct stores
            ptx::fence_proxy_async_smem();
            ptx::named_bar_sync(bar_id, EPI_THREADS);
            if (issuer) {
#pragma unroll
              for (int rr = 0; rr < 2; ++rr) {
                const int h1 = tc.h0 / 2 + g * (RPG / 2) + half * 2 + rr;
                if (h1 < p.lh[1]) ptx::tma_store_4d(&p.tmOut[1], slot + rr * 8192, tc.w0 / 2, h1, tc.m0, tc.b);
              }
              ptx::tma_store_commit();
              store_ring_wait(NS);
            }
            ++ev;
          }
        }
      }
      if (p.num_levels > 2) {
        float l2[RPG / 4][8], l3[4];
#pragma unroll
        for (int i = 0; i < RPG / 4; ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j)
            l2[i][j] = ((l1[2 * i][2 * j] + l1[2 * i][2 * j + 1]) + l1[2 * i + 1][2 * j] + l1[2 * i + 1][2 * j + 1]) * 0.25f;
        bool have_l3 = false;
        if (p.num_levels > 3) {
          if (G == 1) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              l3[j] = ((l2[0][2 * j] + l2[0][2 * j + 1]) + l2[RPG / 4 - 1][2 * j] + l2[RPG / 4 - 1][2 * j + 1]) * 0.25f;
            have_l3 = true;
          } else {
            // group 1's level-2 row goes to group 0, which forms the level-3 row (the 8x8 cell spans both groups); the
            // exchange buffer alternates with the tile parity: a group cannot be more than one such barrier ahead
            float* x = xchg + ((it & 1) * BM + m) * 8;
            if (g == 1) {
#pragma unroll
              for (int j = 0; j < 8; j += 4) *reinterpret_cast<float4*>(x + j) = make_float4(l2[0][j], l2[0][j + 1], l2[0][j + 2], l2[0][j + 3]);
            }
            ptx::named_bar_sync(3, G * EPI_THREADS);
            if (g == 0) {
              float o[8];
#pragma unroll
              for (int j = 0; j < 8; j += 4) {
                const float4 t4 = *reinterpret_cast<const float4*>(x + j);
                o[j] = t4.x; o[j + 1] = t4.y; o[j + 2] = t4.z; o[j + 3] = t4.w;
              }
#pragma unroll
              for (int j = 0; j < 4; ++j) l3[j] = ((l2[0][2 * j] + l2[0][2 * j + 1]) + o[2 * j] + o[2 * j + 1]) * 0.25f;
              have_l3 = true;
            }
          }
        }
        uint8_t* slot = gring + (ev % NS) * SLOT;
#pragma unroll
        for (int i = 0; i < RPG / 4; ++i) {
          const int h2 = tc.h0 / 4 + g * (RPG / 4) + i;
          if (p.tma_ok[2]) {
            if (f16o)
              stage_row_h<8>(slot + i * 4096, m, l2[i]);
            else
              stage_row<8>(slot + i * 4096, m, l2[i]);
          } else if (f16o) {
            direct_row_h<8>(p.level[2], p.lh[2], p.lw[2], p.HW, tc.b, mq, h2, tc.w0 / 4, l2[i]);
          } else {
            direct_row<8>(p.level[2], p.lh[2], p.lw[2], p.HW, tc.b, mq, h2, tc.w0 / 4, l2[i]);
          }
        }
        if (have_l3) {   // uniform within a group
          if (p.tma_ok[3])   // never with fp16 storage: a 4-value row is 8 bytes, below TMA's 16-byte granule
            stage_row<4>(slot + 8192, m, l3);
          else if (f16o)
            direct_row_h<4>(p.level[3], p.lh[3], p.lw[3], p.HW, tc.b, mq, tc.h0 / 8, tc.w0 / 8, l3);
          else
            direct_row<4>(p.level[3], p.lh[3], p.lw[3], p.HW, tc.b, mq, tc.h0 / 8, tc.w0 / 8, l3);
        }
        if (p.tma_ok[2] || (have_l3 && p.tma_ok[3])) {
          ptx::fence_proxy_async_smem();
          ptx::named_bar_sync(bar_id, EPI_THREADS);
          if (issuer) {
            if (p.tma_ok[2]) {
#pragma unroll
              for (int i = 0; i < RPG / 4; ++i) {
                const int h2 = tc.h0 / 4 + g * (RPG / 4) + i;
                if (h2 < p.lh[2]) ptx::tma_store_4d(&p.tmOut[2], slot + i * 4096, tc.w0 / 4, h2, tc.m0, tc.b);
              }
            }
            if (have_l3 && p.tma_ok[3] && tc.h0 / 8 < p.lh[3])
              ptx::tma_store_4d(&p.tmOut[3], slot + 8192, tc.w0 / 8, tc.h0 / 8, tc.m0, tc.b);
            ptx::tma_store_commit();
            store_ring_wait(NS);
          }
          ++ev;
        }
      }
      if (issuer && g == 0) trace_stamp<DEBUG>(p, it, 5);
    }
    }
    if (issuer) ptx::tma_store_wait_all();
  }

  ptx::tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after_sync();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess ||
      qres != cudaDriverEntryPointSuccess)
    return nullptr;
  return reinterpret_cast<EncodeTiledFn>(fn);
}

static bool encode(EncodeTiledFn fn, CUtensorMap* map, CUtensorMapDataType dt, int elem_bytes, int rank, void* base,
                   const cuuint64_t* dims, const cuuint32_t* box, CUtensorMapSwizzle swz) {
  cuuint64_t strides[4];
  cuuint64_t s = elem_bytes;
  for (int i = 0; i < rank - 1; ++i) {
    s *= dims[i];
    strides[i] = s;
    if (s % 16 != 0) return false;
  }
  cuuint32_t estr[4] = {1, 1, 1, 1};
  return fn(map, dt, rank, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct TcLayout {
  int Kp, k_elems, planes, elem_bytes;
  size_t plane_bytes, off_mean, off_a, off_b, total;
};

TcLayout tc_layout(int B, int C, int HW, int mode) {
  TcLayout L{};
  L.elem_bytes = (mode == WOFT_CORR_BF16 || mode == WOFT_CORR_FP16 || mode == WOFT_CORR_FP16X3) ? 2 : 4;
  L.k_elems = 128 / L.elem_bytes;
  L.Kp = (C + L.k_elems - 1) / L.k_elems * L.k_elems;
  L.planes = (mode == WOFT_CORR_TF32X3 || mode == WOFT_CORR_FP16X3) ? 2 : 1;
  L.plane_bytes = (size_t)B * HW * L.Kp * L.elem_bytes;  // Kp*elem_bytes is a multiple of 128
  L.off_mean = 256;
  L.off_a = align_up(L.off_mean + (size_t)B * C * 4, 1024);
  L.off_b = L.off_a + L.planes * L.plane_bytes;
  L.total = align_up(L.off_b + L.planes * L.plane_bytes, 256);
  return L;
}

int corr_pyramid_tc(const float* fmap1, const float* fmap2, int B, int C, int H, int W, int num_levels,
                    float* const* levels, void* workspace, int mode, cudaStream_t stream) {
  const bool fmaps_nhwc = (mode & WOFT_CORR_FMAPS_NHWC) != 0;
  const bool store_f16 = (mode & WOFT_CORR_STORE_F16) != 0;
  const bool fmaps_f16 = (mode & WOFT_CORR_FMAPS_F16) != 0;   // the maps ARE the packed operands (checked by the caller)
  mode &= ~(WOFT_CORR_FMAPS_NHWC | WOFT_CORR_STORE_F16 | WOFT_CORR_FMAPS_F16);
  if (num_levels > kMaxTcLevels)
    return fail(WOFT_ERR_UNSUPPORTED, "tensor-core correlation supports at most %d levels (8x32 pooling tile)",
                kMaxTcLevels);
  int dev = 0, major = 0;
  WOFT_CUDA_CHECK(cudaGetDevice(&dev));
  WOFT_CUDA_CHECK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  if (major != 10) return fail(WOFT_ERR_UNSUPPORTED, "tcgen05 path needs sm_100; device is sm_%d*", major);
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return fail(WOFT_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");

  const int HW = H * W;
  const TcLayout L = tc_layout(B, C, HW, mode);
  uint8_t* ws = static_cast<uint8_t*>(workspace);
  int* flag = reinterpret_cast<int*>(ws);
  void* packA = fmaps_f16 ? const_cast<float*>(fmap1) : static_cast<void*>(ws + L.off_a);
  void* packB = fmaps_f16 ? const_cast<float*>(fmap2) : static_cast<void*>(ws + L.off_b);

  // ---- pack both fmaps into K-major planes (also clears the abort flag; the entry point has cleared it already)
  if (fmaps_f16) {
    // channels-last fp16 maps with C % 64 == 0: nothing to do
  } else if (fmaps_nhwc) {
    const size_t plane_elems = L.plane_bytes / L.elem_bytes, rows = (size_t)B * HW;
    dim3 grid(148 * 8, 2);
    if (mode == WOFT_CORR_BF16)
      pack_nhwc_kernel<3><<<grid, 256, 0, stream>>>(fmap1, fmap2, packA, packB, C, L.Kp, rows, plane_elems, flag);
    else if (mode == WOFT_CORR_FP16)
      pack_nhwc_kernel<4><<<grid, 256, 0, stream>>>(fmap1, fmap2, packA, packB, C, L.Kp, rows, plane_elems, flag);
    else if (mode == WOFT_CORR_FP16X3)
      pack_nhwc_kernel<5><<<grid, 256, 0, stream>>>(fmap1, fmap2, packA, packB, C, L.Kp, rows, plane_elems, flag);
    else if (mode == WOFT_CORR_TF32X3)
      pack_nhwc_kernel<2><<<grid, 256, 0, stream>>>(fmap1, fmap2, packA, packB, C, L.Kp, rows, plane_elems, flag);
    else
      pack_nhwc_kernel<1><<<grid, 256, 0, stream>>>(fmap1, fmap2, packA, packB, C, L.Kp, rows, plane_elems, flag);
    int rc = launch_status("pack_nhwc_kernel");
    if (rc != WOFT_OK) return rc;
  } else {
    dim3 grid(ceil_div(HW, 32), L.Kp / 32, 2 * B);
    const size_t plane_elems = L.plane_bytes / L.elem_bytes;
    if (mode == WOFT_CORR_BF16)
      pack_kernel<3><<<grid, 256, 0, stream>>>(fmap1, fmap2, packA, packB, B, C, HW, L.Kp, plane_elems, flag);
    else if (mode == WOFT_CORR_FP16)
      pack_kernel<4><<<grid, 256, 0, stream>>>(fmap1, fmap2, packA, packB, B, C, HW, L.Kp, plane_elems, flag);
    else if (mode == WOFT_CORR_FP16X3)
      pack_kernel<5><<<grid, 256, 0, stream>>>(fmap1, fmap2, packA, packB, B, C, HW, L.Kp, plane_elems, flag);
    else if (mode == WOFT_CORR_TF32X3)
      pack_kernel<2><<<grid, 256, 0, stream>>>(fmap1, fmap2, packA, packB, B, C, HW, L.Kp, plane_elems, flag);
    else
      pack_kernel<1><<<grid, 256, 0, stream>>>(fmap1, fmap2, packA, packB, B, C, HW, L.Kp, plane_elems, flag);
    int rc = launch_status("pack_kernel");
    if (rc != WOFT_OK) return rc;
  }

  CorrTcParams p{};
  const CUtensorMapDataType dt = (mode == WOFT_CORR_BF16)   ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                 : (mode == WOFT_CORR_FP16 || mode == WOFT_CORR_FP16X3) ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                                                            : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  // the hi and lo planes lie back to back, so they fold into the batch dimension of the operand maps
  {
    cuuint64_t dims[3] = {(cuuint64_t)L.Kp, (cuuint64_t)HW, (cuuint64_t)(B * L.planes)};
    cuuint32_t box[3] = {(cuuint32_t)L.k_elems, BM, 1};
    if (!encode(enc, &p.tmA, dt, L.elem_bytes, 3, packA, dims, box, CU_TENSOR_MAP_SWIZZLE_128B))
      return fail(WOFT_ERR_CUDA, "cuTensorMapEncodeTiled failed for the fmap1 operand");
  }
  {
    cuuint64_t dims[4] = {(cuuint64_t)L.Kp, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)(B * L.planes)};
    cuuint32_t box[4] = {(cuuint32_t)L.k_elems, TW, TH, 1};
    if (!encode(enc, &p.tmB, dt, L.elem_bytes, 4, packB, dims, box, CU_TENSOR_MAP_SWIZZLE_128B))
      return fail(WOFT_ERR_CUDA, "cuTensorMapEncodeTiled failed for the fmap2 operand");
  }
  int h = H, w = W;
  for (int l = 0; l < num_levels; ++l) {
    p.level[l] = levels[l];
    p.lh[l] = h;
    p.lw[l] = w;
    const int row_bytes = (TW >> l) * (store_f16 ? 2 : 4);   // one query's row of the tile at this level
    const CUtensorMapSwizzle swz = row_bytes == 128  ? CU_TENSOR_MAP_SWIZZLE_128B
                                   : row_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                   : row_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                                     : CU_TENSOR_MAP_SWIZZLE_NONE;
    cuuint64_t dims[4] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)HW, (cuuint64_t)B};
    cuuint32_t box[4] = {(cuuint32_t)(TW >> l), 1, BM, 1};
    const int tma_levels = tc_knobs().tma_levels;  // bit l set => level l may use TMA stores
    p.tma_ok[l] = (((tma_levels >> l) & 1) && h > 0 && w > 0 && row_bytes >= 16 &&
                   (reinterpret_cast<uintptr_t>(levels[l]) & 15) == 0 &&
                   encode(enc, &p.tmOut[l], store_f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
                          store_f16 ? 2 : 4, 4, levels[l], dims, box, swz))
                      ? 1
                      : 0;
    h /= 2;
    w /= 2;
  }
  p.num_levels = num_levels;
  p.out_f16 = store_f16 ? 1 : 0;
  p.B = B;
  p.HW = HW;
  p.H = H;
  p.W = W;
  p.m_tiles = ceil_div(HW, BM);
  p.h_tiles = ceil_div(H, TH);
  p.w_tiles = ceil_div(W, TW);
  p.total_tiles = B * p.m_tiles * p.h_tiles * p.w_tiles;
  p.k_blocks = L.Kp / L.k_elems;
  p.k_elems = L.k_elems;
  p.n_pairs = (mode == WOFT_CORR_TF32X3 || mode == WOFT_CORR_FP16X3) ? 3 : 1;
  p.bf16 = (mode == WOFT_CORR_BF16) ? 1 : (mode == WOFT_CORR_FP16 || mode == WOFT_CORR_FP16X3) ? 2 : 0;
  p.scale = 1.0f / sqrtf(float(C));
  p.abort_flag = flag;

  // Shared-memory plan.  Keeping the A tile resident removes its re-fetch from L2 for every N tile (the kernel is
  // L2-bandwidth bound on operand traffic); it needs k_blocks x 16 KB, which leaves room for 3 B
  // stages + 3 ring slots (bf16, K <= 256) or 2 + 2 (tf32, K <= 256).  The 3xTF32 split streams both operands.
  const int fixed = 256 + 1024;
  p.a_resident = 0;
  p.unit = 1;
  p.stages = 3;
  p.nslot = 3;
  const bool want_res = p.n_pairs == 1 && tc_knobs().a_resident;  // off by default: measured slower
  if (want_res) {
    const int a_bytes = p.k_blocks * A_STAGE;
    if (a_bytes + 3 * B_STAGE + 3 * SLOT + fixed <= TC_SMEM_LIMIT) {
      p.a_resident = 1;
    } else if (a_bytes + 2 * B_STAGE + 2 * SLOT + fixed <= TC_SMEM_LIMIT) {
      p.a_resident = 1;
      p.stages = 2;
      p.nslot = 2;
    }
    if (p.a_resident) {
      const int n_tiles = p.h_tiles * p.w_tiles;
      int unit = tc_knobs().unit;
      if (unit < 1) unit = 1;
      // units must not straddle M tiles more than necessary: prefer a divisor of the N-tile count near `unit`
      for (int u = unit; u >= 1; --u)
        if (n_tiles % u == 0) {
          unit = u;
          break;
        }
      p.unit = unit;
    }
  }
  const bool debug = p.a_resident || __atomic_load_n(&g_trace_buffer, __ATOMIC_ACQUIRE) != nullptr || p.stages != 3 ||
                     p.nslot != 3 || p.unit != 1;
  const bool two = store_f16 && !debug;   // two epilogue groups (see corr_tc_kernel)
  const int smem_bytes = (p.a_resident ? p.k_blocks : p.stages) * A_STAGE + p.stages * B_STAGE +
                         tc_ring_bytes(two, p.nslot) + fixed;

  int sms = 0;
  WOFT_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  WOFT_CUDA_CHECK(cudaFuncSetAttribute(corr_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  const int chunks = ceil_div(p.total_tiles, p.unit);
  const int grid = chunks < sms ? chunks : sms;
  // debug aid: per-role timestamps of the first 4 CTAs into a buffer the caller handed over beforehand
  // (4 * 64 * 6 unsigned long long, zeroed by the caller); the library itself never allocates
  p.trace = __atomic_load_n(&g_trace_buffer, __ATOMIC_ACQUIRE);
  if (debug) {
    WOFT_CUDA_CHECK(cudaFuncSetAttribute(corr_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    corr_tc_kernel<true><<<grid, tc_threads(false), smem_bytes, stream>>>(p);
  } else if (two) {
    WOFT_CUDA_CHECK(cudaFuncSetAttribute(corr_tc_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    corr_tc_kernel<false, true><<<grid, tc_threads(true), smem_bytes, stream>>>(p);
  } else {
    corr_tc_kernel<false><<<grid, tc_threads(false), smem_bytes, stream>>>(p);
  }
  int rc = launch_status("corr_tc_kernel");
  return rc;
}

// Debug knobs of the experiments in DESIGN.md 4, read from the environment ONCE per process (never on the launch path):
//   WOFT_CORR_POOL_TMA (bit l: level l may use TMA stores), WOFT_CORR_A_RESIDENT=1, WOFT_CORR_UNIT=<tiles per unit>
const TcKnobs& tc_knobs() {
  static const TcKnobs k = [] {
    TcKnobs v;
    const char* e = getenv("WOFT_CORR_POOL_TMA");
    v.tma_levels = e ? atoi(e) : 0xF;
    e = getenv("WOFT_CORR_A_RESIDENT");
    v.a_resident = e && e[0] == '1';
    e = getenv("WOFT_CORR_UNIT");
    v.unit = e ? atoi(e) : 6;
    return v;
  }();
  return k;
}

size_t corr_tc_workspace_bytes(int B, int C, int HW, int mode) { return tc_layout(B, C, HW, mode).total; }

}  // namespace woft
