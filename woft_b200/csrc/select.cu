// Correspondence selection on the device, without a host round trip.
//
// Replaces, for the tracker's global branch:
//   RAFTWrapper.compute_flow mode 'TC'      pixel grid + flow, flat weights   (pytracking/optical_flow/raft.py:185-218)
//   YAOFTrackerSingleControl._mask_coords   template mask & in bounds & pre-warp mask, boolean indexing
//                                           (pytracking/tracker/YAOF_tracker_single_control.py:287-312)
//   subsampler                              500 unscrambled 1-D Sobol positions -> a boolean mask built on the host,
//                                           copied back to the device (configs/..._wLSq.py:31-53)
// The reference does this with ~20 small kernels, a device->host->device trip of the mask and three
// synchronisations (nonzero).  Here:
//   select_flag_kernel    one thread per pixel: keep flag + per-1024-pixel block count
//   select_plan_kernel    one CTA: exclusive scan of the block counts -> N kept; the selected ranks
//                         r_k = rint(float(N) * u_k) in fp32, ties to even, exactly numpy's arithmetic in the reference;
//                         u sorted ascending so the ranks come out in raster order; duplicates dropped (the
//                         reference's mask[indices] = True); everything kept when N <= to_draw
//   select_gather_kernel  one CTA per selected rank: finds its 1024-pixel block, ranks the flags inside it with
//                         ballots, writes template point, current point (= pixel + flow) and weight
// The number of selected points stays in device memory; woft_homography_fit_n / woft_inlier_fraction read it there.
#include "common.cuh"

namespace woft {

constexpr int kSelBlock = 1024;   // pixels per counting block
constexpr int kSelMaxDraw = 1024; // upper bound on to_draw

__global__ void __launch_bounds__(kSelBlock) select_flag_kernel(const float* __restrict__ flow,       // (2, H, W)
                                                               const uint8_t* __restrict__ tmask,     // (H, W)
                                                               const uint8_t* __restrict__ pwmask,    // (H, W) or null
                                                               uint8_t* __restrict__ keep, int* __restrict__ counts,
                                                               int H, int W) {
  __shared__ int warp_cnt[kSelBlock / 32];
  const int i = blockIdx.x * kSelBlock + threadIdx.x;
  const int hw = H * W;
  bool k = false;
  if (i < hw) {
    const int y = i / W, x = i - y * W;
    const float cx = float(x) + __ldg(flow + i), cy = float(y) + __ldg(flow + hw + i);  // dst = src + flow (raft.py:195)
    // cur_coords.round().long(): round half to even; out of bounds if cur < 0 (as floats) or the rounded index >= W/H
    const float rx = rintf(cx), ry = rintf(cy);
    const bool oob = (cx < 0.f) || (cy < 0.f) || !(rx < float(W)) || !(ry < float(H));
    k = __ldg(tmask + i) != 0 && !oob;
    if (k && pwmask) k = __ldg(pwmask + (size_t)int(ry) * W + int(rx)) != 0;
    keep[i] = k ? 1 : 0;
  }
  const unsigned b = __ballot_sync(0xffffffffu, k);
  if ((threadIdx.x & 31) == 0) warp_cnt[threadIdx.x >> 5] = __popc(b);
  __syncthreads();
  if (threadIdx.x < 32) {
    int v = warp_cnt[threadIdx.x];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x == 0) counts[blockIdx.x] = v;
  }
}

// Lost branch (YAOF_tracker_single_control.py:314-327, _mask_coords_flow): the flow runs from the previous frame, and
// a pixel of it is kept iff the template mask, seen from the previous frame, covers it:
//   prev_mask = cv2.warpPerspective(template_mask, inv(prev_H2init), (W, H), flags=INTER_NEAREST) > 0
// i.e. prev_mask(x, y) = template_mask(round(prev_H2init (x, y, 1))), zero outside.  No bounds test on the targets.
struct LocalMap {
  double m[9];   // prev_H2init: previous-frame pixel -> template pixel
};

__global__ void __launch_bounds__(kSelBlock) select_flag_local_kernel(const uint8_t* __restrict__ tmask, const LocalMap map,
                                                                     uint8_t* __restrict__ keep,
                                                                     int* __restrict__ counts, int H, int W) {
  __shared__ int warp_cnt[kSelBlock / 32];
  const int i = blockIdx.x * kSelBlock + threadIdx.x;
  bool k = false;
  if (i < H * W) {
    const int y = i / W, x = i - y * W;
    // OpenCV's warpPerspective, nearest: fp64 homogeneous coordinates, division by W (0 if W == 0), cvRound
    const double X0 = map.m[0] * x + map.m[1] * y + map.m[2];
    const double Y0 = map.m[3] * x + map.m[4] * y + map.m[5];
    double Wd = map.m[6] * x + map.m[7] * y + map.m[8];
    Wd = Wd != 0.0 ? 1.0 / Wd : 0.0;
    const int sx = __double2int_rn(fmax(-2147483648.0, fmin(2147483647.0, X0 * Wd)));
    const int sy = __double2int_rn(fmax(-2147483648.0, fmin(2147483647.0, Y0 * Wd)));
    k = sx >= 0 && sx < W && sy >= 0 && sy < H && __ldg(tmask + (size_t)sy * W + sx) != 0;
    keep[i] = k ? 1 : 0;
  }
  const unsigned b = __ballot_sync(0xffffffffu, k);
  if ((threadIdx.x & 31) == 0) warp_cnt[threadIdx.x >> 5] = __popc(b);
  __syncthreads();
  if (threadIdx.x < 32) {
    int v = warp_cnt[threadIdx.x];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x == 0) counts[blockIdx.x] = v;
  }
}

// counts -> exclusive offsets (in place), n_kept, and the list of selected ranks (sorted, unique).
__global__ void __launch_bounds__(1024) select_plan_kernel(int* __restrict__ counts, int nblocks,
                                                           const float* __restrict__ u_sorted, int to_draw,
                                                           int* __restrict__ ranks, int* __restrict__ n_sel,
                                                           int* __restrict__ n_kept) {
  __shared__ int s_part[1024];
  __shared__ int s_total;
  // exclusive scan of up to nblocks counts, 1024 threads, several per thread
  const int per = (nblocks + 1023) / 1024;
  int local = 0;
  for (int j = 0; j < per; ++j) {
    const int idx = threadIdx.x * per + j;
    if (idx < nblocks) local += counts[idx];
  }
  s_part[threadIdx.x] = local;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {  // Hillis-Steele inclusive scan
    const int v = threadIdx.x >= o ? s_part[threadIdx.x - o] : 0;
    __syncthreads();
    s_part[threadIdx.x] += v;
    __syncthreads();
  }
  int run = s_part[threadIdx.x] - local;
  for (int j = 0; j < per; ++j) {
    const int idx = threadIdx.x * per + j;
    if (idx < nblocks) {
      const int c = counts[idx];
      counts[idx] = run;
      run += c;
    }
  }
  if (threadIdx.x == 1023) s_total = s_part[1023];
  __syncthreads();
  const int N = s_total;
  // selected ranks
  __shared__ int s_r[kSelMaxDraw];
  __shared__ int s_keep[kSelMaxDraw];
  const int k = threadIdx.x;
  const bool all = to_draw >= N;  // `if to_draw >= N_pts: return` everything (config :39-40)
  const int m_cand = all ? N : to_draw;
  if (k < kSelMaxDraw) {
    int r = -1;
    if (k < m_cand) r = all ? k : __float2int_rn(float(N) * __ldg(u_sorted + k));  // np.round(N * u) in float32
    s_r[k] = r;
  }
  __syncthreads();
  int flag = 0;
  if (k < m_cand) flag = (k == 0 || s_r[k] != s_r[k - 1]) && s_r[k] >= 0 && s_r[k] < N;
  if (k < kSelMaxDraw) s_keep[k] = flag;
  __syncthreads();
  // compact the unique ranks (<= 1024 entries: a serial prefix by one warp's worth of work is plenty)
  if (threadIdx.x == 0) {
    int m = 0;
    for (int j = 0; j < m_cand && j < kSelMaxDraw; ++j)
      if (s_keep[j]) ranks[m++] = s_r[j];
    *n_sel = m;
    *n_kept = N;
  }
}

__global__ void __launch_bounds__(kSelBlock) select_gather_kernel(const float* __restrict__ flow,
                                                                 const float* __restrict__ weights,  // (H*W) or null
                                                                 const uint8_t* __restrict__ keep,
                                                                 const int* __restrict__ offsets, int nblocks,
                                                                 const int* __restrict__ ranks,
                                                                 const int* __restrict__ n_sel, int H, int W,
                                                                 int apply_sigmoid, float* __restrict__ pts_template,
                                                                 float* __restrict__ pts_cur, float* __restrict__ w_out) {
  __shared__ int warp_cnt[kSelBlock / 32];
  __shared__ int s_blk;
  const int k = blockIdx.x;
  if (k >= *n_sel) return;
  const int r = ranks[k];
  if (threadIdx.x == 0) {  // last block whose offset <= r
    int lo = 0, hi = nblocks - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (offsets[mid] <= r) lo = mid; else hi = mid - 1;
    }
    s_blk = lo;
  }
  __syncthreads();
  const int blk = s_blk;
  const int want = r - offsets[blk];  // index among the kept pixels of this block
  const int i = blk * kSelBlock + threadIdx.x;
  const int hw = H * W;
  const bool kflag = i < hw && keep[i] != 0;
  const unsigned b = __ballot_sync(0xffffffffu, kflag);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) warp_cnt[warp] = __popc(b);
  __syncthreads();
  int before = 0;
  for (int wv = 0; wv < warp; ++wv) before += warp_cnt[wv];
  const int my_rank = before + __popc(b & ((1u << lane) - 1u));
  if (kflag && my_rank == want) {
    const int y = i / W, x = i - y * W;
    pts_template[2 * k] = float(x);
    pts_template[2 * k + 1] = float(y);
    pts_cur[2 * k] = float(x) + __ldg(flow + i);
    pts_cur[2 * k + 1] = float(y) + __ldg(flow + hw + i);
    float wv = weights ? __ldg(weights + i) : 1.0f;
    if (weights && apply_sigmoid) wv = 1.0f / (1.0f + expf(-wv));
    w_out[k] = wv;
  }
}

// fraction of correspondences with ||H a - b|| <= thr (configs/..._wLSq.py:14-21, least_squares_H.py:474-489)
__global__ void __launch_bounds__(256) inlier_fraction_kernel(const float* __restrict__ Hm, const float* __restrict__ a,
                                                              const float* __restrict__ b, const int* __restrict__ n_dev,
                                                              float thr, float min_frac, float* __restrict__ frac_out,
                                                              uint8_t* __restrict__ ok_out) {
  __shared__ int s_cnt[8];
  const int n = *n_dev;
  int cnt = 0;
  for (int i = threadIdx.x; i < n; i += 256) {
    const float x = a[2 * i], y = a[2 * i + 1];
    const float u = Hm[0] * x + Hm[1] * y + Hm[2], v = Hm[3] * x + Hm[4] * y + Hm[5], z = Hm[6] * x + Hm[7] * y + Hm[8];
    const float sc = fabsf(z) > 1e-8f ? 1.0f / (z + 1e-8f) : 1.0f;
    const float dx = u * sc - b[2 * i], dy = v * sc - b[2 * i + 1];
    cnt += sqrtf(dx * dx + dy * dy) <= thr ? 1 : 0;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  if ((threadIdx.x & 31) == 0) s_cnt[threadIdx.x >> 5] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < 8; ++w) t += s_cnt[w];
    const float f = n > 0 ? float(t) / float(n) : 0.f;
    frac_out[0] = f;
    frac_out[1] = float(n);
    *ok_out = (f > min_frac) ? 1 : 0;  // NaN homography -> comparisons false -> not ok
  }
}

}  // namespace woft

extern "C" size_t woft_select_scratch_bytes(int height, int width) {
  if (height <= 0 || width <= 0) return 0;
  const size_t hw = (size_t)height * width;
  const size_t nblocks = (hw + woft::kSelBlock - 1) / woft::kSelBlock;
  // keep flags | block counts/offsets | selected ranks | n_kept
  return ((hw + 255) / 256 * 256) + ((nblocks * 4 + 255) / 256 * 256) + woft::kSelMaxDraw * 4 + 256;
}

static int select_impl(const float* flow_up, const float* weights_up, const uint8_t* template_mask, const uint8_t* pw_mask,
                       const double* prev_to_template_host, int height, int width, const float* sobol_sorted, int to_draw,
                       int apply_sigmoid, float* pts_template, float* pts_cur, float* weights_out, int* n_selected,
                       void* scratch, size_t scratch_bytes, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(flow_up && template_mask && sobol_sorted && pts_template && pts_cur && weights_out && n_selected && scratch,
               "woft_select_correspondences: null pointer");
  WOFT_REQUIRE(height > 0 && width > 0 && (long long)height * width < (1ll << 30), "woft_select_correspondences: bad extent");
  WOFT_REQUIRE(to_draw >= 1 && to_draw <= kSelMaxDraw, "woft_select_correspondences: to_draw %d not in [1,%d]", to_draw,
               kSelMaxDraw);
  if (scratch_bytes < woft_select_scratch_bytes(height, width))
    return fail(WOFT_ERR_WORKSPACE, "woft_select_correspondences: scratch too small");
  const size_t hw = (size_t)height * width;
  const int nblocks = int((hw + kSelBlock - 1) / kSelBlock);
  WOFT_REQUIRE(nblocks <= 1024 * 64, "woft_select_correspondences: image too large");
  uint8_t* base = static_cast<uint8_t*>(scratch);
  uint8_t* keep = base;
  int* counts = reinterpret_cast<int*>(base + (hw + 255) / 256 * 256);
  int* ranks = reinterpret_cast<int*>(reinterpret_cast<uint8_t*>(counts) + ((size_t)nblocks * 4 + 255) / 256 * 256);
  int* n_kept = ranks + kSelMaxDraw;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (prev_to_template_host) {
    LocalMap map;
    for (int k = 0; k < 9; ++k) map.m[k] = prev_to_template_host[k];
    select_flag_local_kernel<<<nblocks, kSelBlock, 0, s>>>(template_mask, map, keep, counts, height, width);
  } else {
    select_flag_kernel<<<nblocks, kSelBlock, 0, s>>>(flow_up, template_mask, pw_mask, keep, counts, height, width);
  }
  select_plan_kernel<<<1, 1024, 0, s>>>(counts, nblocks, sobol_sorted, to_draw, ranks, n_selected, n_kept);
  select_gather_kernel<<<to_draw, kSelBlock, 0, s>>>(flow_up, weights_up, keep, counts, nblocks, ranks, n_selected, height,
                                                     width, apply_sigmoid, pts_template, pts_cur, weights_out);
  return launch_status("select kernels");
}

extern "C" int woft_select_correspondences(const float* flow_up, const float* weights_up, const uint8_t* template_mask,
                                           const uint8_t* pw_mask, int height, int width, const float* sobol_sorted,
                                           int to_draw, int apply_sigmoid, float* pts_template, float* pts_cur,
                                           float* weights_out, int* n_selected, void* scratch, size_t scratch_bytes,
                                           void* stream) {
  return select_impl(flow_up, weights_up, template_mask, pw_mask, nullptr, height, width, sobol_sorted, to_draw,
                     apply_sigmoid, pts_template, pts_cur, weights_out, n_selected, scratch, scratch_bytes, stream);
}

extern "C" int woft_select_correspondences_local(const float* flow_up, const float* weights_up,
                                                 const uint8_t* template_mask, const double* prev_to_template_host,
                                                 int height, int width, const float* sobol_sorted, int to_draw,
                                                 int apply_sigmoid, float* pts_prev, float* pts_cur, float* weights_out,
                                                 int* n_selected, void* scratch, size_t scratch_bytes, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(prev_to_template_host, "woft_select_correspondences_local: null matrix");
  return select_impl(flow_up, weights_up, template_mask, nullptr, prev_to_template_host, height, width, sobol_sorted,
                     to_draw, apply_sigmoid, pts_prev, pts_cur, weights_out, n_selected, scratch, scratch_bytes, stream);
}

extern "C" int woft_inlier_fraction(const float* H, const float* pts_a, const float* pts_b, const int* n_points_dev,
                                    float threshold, float min_fraction, float* fraction_out, uint8_t* ok_out,
                                    void* stream) {
  using namespace woft;
  WOFT_REQUIRE(H && pts_a && pts_b && n_points_dev && fraction_out && ok_out, "woft_inlier_fraction: null pointer");
  inlier_fraction_kernel<<<1, 256, 0, static_cast<cudaStream_t>(stream)>>>(H, pts_a, pts_b, n_points_dev, threshold,
                                                                          min_fraction, fraction_out, ok_out);
  return launch_status("inlier_fraction_kernel");
}
