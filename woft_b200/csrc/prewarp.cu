// Device-side pre-warp of the incoming frame.
//
// Replaces the two cv2.warpPerspective calls at the top of YAOFTrackerSingleControl.track
// (pytracking/tracker/YAOF_tracker_single_control.py:89-95): the frame warped by the last good homography with
// bilinear interpolation and a zero border, and a float64 image of ones warped the same way whose `> 0` is the
// validity mask of the pre-warp.  On the host those cost 10-35 ms per 1280x720 frame; here it is one pass over
// the frame on the device, fed by the raw uint8 upload.
//
// The arithmetic is OpenCV's, integer for integer, so the result is bit-identical to cv2.warpPerspective(...,
// INTER_LINEAR) on uint8 (imgwarp.cpp WarpPerspectiveInvoker + remapBilinear<FixPtCast<int, uchar, 15>>):
//   - destination pixels are walked in blocks of 64 columns (min(1024 / min(16, h), w) in general); inside a block the fp64 numerators are formed as
//     (M0*x_block + M1*y + M2) + M0*x1 -- the same operations in the same order, no contraction into FMAs;
//   - source coordinates are scaled by INTER_TAB_SIZE = 32, rounded to nearest even (cvRound), split into the
//     integer pixel (>> 5) and two 5-bit fractions;
//   - the four bilinear weights are OpenCV's 15-bit table: 32*(32-fy)*(32-fx), ... (exact in its float arithmetic),
//     except the entry for fx = fy = 0, which its short saturation and sum fix-up turn into {32767, 0, 0, 1};
//   - taps outside the image contribute the constant border value 0; the pixel is (sum + 2^14) >> 15.
// The mask is the `> 0` of a float64 image of ones warped alike: 1 where a tap with non-zero (float-table) weight lies
// inside the image.
#include "common.cuh"

namespace woft {

struct WarpParams {
  double m[9];  // inverse map: destination (x, y, 1) -> source
  const uint8_t* src;
  uint8_t* dst;
  uint8_t* mask;  // may be null
  int h, w, channels;
  int block_w;  // OpenCV's tile width: min(1024 / min(16, h), w), 64 for any frame at least 16 x 64
};

__global__ void __launch_bounds__(256) warp_perspective_u8_kernel(const WarpParams p) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y;
  if (x >= p.w) return;
  const int xblock = x / p.block_w * p.block_w;
  const double xb = double(xblock), x1 = double(x - xblock), yd = double(y);
  // __dmul_rn / __dadd_rn: never contracted, so every intermediate rounds as in OpenCV's scalar and SSE code
  const double X0 = __dadd_rn(__dadd_rn(__dmul_rn(p.m[0], xb), __dmul_rn(p.m[1], yd)), p.m[2]);
  const double Y0 = __dadd_rn(__dadd_rn(__dmul_rn(p.m[3], xb), __dmul_rn(p.m[4], yd)), p.m[5]);
  const double W0 = __dadd_rn(__dadd_rn(__dmul_rn(p.m[6], xb), __dmul_rn(p.m[7], yd)), p.m[8]);
  double Wd = __dadd_rn(W0, __dmul_rn(p.m[6], x1));
  Wd = Wd != 0.0 ? __ddiv_rn(32.0, Wd) : 0.0;
  const double fx = fmax(-2147483648.0, fmin(2147483647.0, __dmul_rn(__dadd_rn(X0, __dmul_rn(p.m[0], x1)), Wd)));
  const double fy = fmax(-2147483648.0, fmin(2147483647.0, __dmul_rn(__dadd_rn(Y0, __dmul_rn(p.m[3], x1)), Wd)));
  const int X = __double2int_rn(fx), Y = __double2int_rn(fy);  // cvRound: nearest, ties to even
  const int sx = max(-32768, min(32767, X >> 5)), sy = max(-32768, min(32767, Y >> 5));  // saturate_cast<short>
  const int ax = X & 31, ay = Y & 31;
  int w00 = 32 * (32 - ay) * (32 - ax), w01 = 32 * (32 - ay) * ax, w10 = 32 * ay * (32 - ax), w11 = 32 * ay * ax;
  if ((ax | ay) == 0) {
    w00 = 32767;
    w11 = 1;
  }
  const bool in_x0 = sx >= 0 && sx < p.w, in_x1 = sx + 1 >= 0 && sx + 1 < p.w;
  const bool in_y0 = sy >= 0 && sy < p.h, in_y1 = sy + 1 >= 0 && sy + 1 < p.h;
  const int c = p.channels;
  const uint8_t* r0 = p.src + ((size_t)(in_y0 ? sy : 0) * p.w) * c;
  const uint8_t* r1 = p.src + ((size_t)(in_y1 ? sy + 1 : 0) * p.w) * c;
  uint8_t* o = p.dst + ((size_t)y * p.w + x) * c;
  for (int k = 0; k < c; ++k) {
    int v = 1 << 14;
    if (in_y0 && in_x0) v += w00 * int(r0[(size_t)sx * c + k]);
    if (in_y0 && in_x1) v += w01 * int(r0[(size_t)(sx + 1) * c + k]);
    if (in_y1 && in_x0) v += w10 * int(r1[(size_t)sx * c + k]);
    if (in_y1 && in_x1) v += w11 * int(r1[(size_t)(sx + 1) * c + k]);
    o[k] = (uint8_t)(v >> 15);  // weights sum to 2^15, pixels <= 255: the result cannot leave [0, 255]
  }
  if (p.mask) {
    // float table of the 64-bit remap: (1-fy)(1-fx), (1-fy)fx, fy(1-fx), fy*fx, no saturation quirk
    const bool any = (in_y0 && in_x0) || (in_y0 && in_x1 && ax > 0) || (in_y1 && in_x0 && ay > 0) ||
                     (in_y1 && in_x1 && ax > 0 && ay > 0);
    p.mask[(size_t)y * p.w + x] = any ? 1 : 0;
  }
}

}  // namespace woft

extern "C" int woft_warp_perspective_u8(const uint8_t* src, int height, int width, int channels,
                                        const double* inv_matrix_host, uint8_t* dst, uint8_t* valid_mask, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(src && dst && inv_matrix_host, "woft_warp_perspective_u8: null pointer");
  WOFT_REQUIRE(height > 0 && width > 0 && channels >= 1 && channels <= 4, "woft_warp_perspective_u8: bad extent");
  WOFT_REQUIRE(height < 65536, "woft_warp_perspective_u8: image too tall");
  WarpParams p;
  for (int i = 0; i < 9; ++i) p.m[i] = inv_matrix_host[i];
  p.src = src;
  p.dst = dst;
  p.mask = valid_mask;
  p.h = height;
  p.w = width;
  p.channels = channels;
  p.block_w = min(1024 / min(16, height), width);
  dim3 grid(ceil_div(width, 256), height);
  warp_perspective_u8_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  return launch_status("warp_perspective_u8_kernel");
}
