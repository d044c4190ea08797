// Convex upsampling of the flow and of the weight logits, 1/8 resolution -> full resolution, in one kernel
// (SURVEY.md 8a rows a10 and a12, 8f row 4).
//
// Replaces WeightedRAFT.upsample_flow applied twice (pytracking/external/RAFT/raft_core/weighted_raft.py:92-103,
// 249-255, 285-288): softmax over the 9 taps of the (N, 9, 8, 8, H, W) mask, unfold of 8*field with a zero-padded
// 3x3 window, weighted sum, pixel shuffle -- run once for the 2-channel flow and once more (softmax included) for
// the 1-channel weights, whose result is then divided by 8 -- and, when asked, the sigmoid of
// optical_flow/raft.py:158-159, 274-280.
// One CTA of 64 threads per coarse pixel; thread (i, j) owns fine pixel (8y+i, 8x+j): 9 logits -> softmax once ->
// three convex combinations.  mask: (B, 576, H, W) with channel k*64 + i*8 + j, NCHW or channels-last.
#include "common.cuh"

namespace woft {

__global__ void __launch_bounds__(64) convex_upsample_kernel(const float* __restrict__ flow,     // (B,2,H,W)
                                                             const float* __restrict__ wlow,     // (B,1,H,W) or null
                                                             const float* __restrict__ mask, int mask_nhwc,
                                                             float* __restrict__ flow_up,        // (B,2,8H,8W)
                                                             float* __restrict__ w_up,           // (B,1,8H,8W) or null
                                                             int H, int W, int apply_sigmoid) {
  const int x = blockIdx.x, y = blockIdx.y, b = blockIdx.z;
  const int ij = threadIdx.x, i = ij >> 3, j = ij & 7;
  const size_t HW = (size_t)H * W;
  float m[9];
  float mx = -3.4e38f;
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    m[k] = mask_nhwc ? __ldg(mask + (((size_t)b * H + y) * W + x) * 576 + k * 64 + ij)
                     : __ldg(mask + ((size_t)b * 576 + k * 64 + ij) * HW + (size_t)y * W + x);
    mx = fmaxf(mx, m[k]);
  }
  float den = 0.f;
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    m[k] = expf(m[k] - mx);
    den += m[k];
  }
  const float inv = 1.0f / den;
  float fx = 0.f, fy = 0.f, fw = 0.f;
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    const int yy = y + k / 3 - 1, xx = x + k % 3 - 1;  // unfold(kernel 3, padding 1): zero outside the map
    if (yy < 0 || yy >= H || xx < 0 || xx >= W) continue;
    const size_t o = (size_t)yy * W + xx;
    const float wk = m[k] * inv;
    fx = fmaf(wk, 8.0f * __ldg(flow + (size_t)(b * 2) * HW + o), fx);
    fy = fmaf(wk, 8.0f * __ldg(flow + (size_t)(b * 2 + 1) * HW + o), fy);
    if (wlow) fw = fmaf(wk, 8.0f * __ldg(wlow + (size_t)b * HW + o), fw);
  }
  const size_t W8 = (size_t)W * 8, HW8 = HW * 64;
  const size_t fo = (size_t)(y * 8 + i) * W8 + x * 8 + j;
  flow_up[(size_t)(b * 2) * HW8 + fo] = fx;
  flow_up[(size_t)(b * 2 + 1) * HW8 + fo] = fy;
  if (w_up) {
    float v = fw / 8.0f;  // "flow upsampling multiplies the values by 8, we have to revert that" (:284-288)
    if (apply_sigmoid) v = 1.0f / (1.0f + expf(-v));
    w_up[(size_t)b * HW8 + fo] = v;
  }
}

}  // namespace woft

extern "C" int woft_convex_upsample(const float* flow_low, const float* weights_low, const float* mask, int mask_nhwc,
                                    int batch, int height, int width, int apply_sigmoid, float* flow_up,
                                    float* weights_up, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(flow_low && mask && flow_up, "woft_convex_upsample: null pointer");
  WOFT_REQUIRE((weights_low == nullptr) == (weights_up == nullptr),
               "woft_convex_upsample: weights_low and weights_up must both be given or both be NULL");
  WOFT_REQUIRE(batch > 0 && height > 0 && width > 0 && height < 65536 && batch < 65536,
               "woft_convex_upsample: bad extent");
  dim3 grid(width, height, batch);
  convex_upsample_kernel<<<grid, 64, 0, static_cast<cudaStream_t>(stream)>>>(flow_low, weights_low, mask, mask_nhwc,
                                                                            flow_up, weights_up, height, width,
                                                                            apply_sigmoid);
  return launch_status("convex_upsample_kernel");
}
