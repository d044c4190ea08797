// Device-side pre-warp of the incoming frame.
//
// Replaces the two cv2.warpPerspective calls at the top of YAOFTrackerSingleControl.track
// (pytracking/tracker/YAOF_tracker_single_control.py:89-95): the frame warped by the last good homography with
// bilinear interpolation and a zero border, and a float64 image of ones warped the same way whose `> 0` is the
// validity mask of the pre-warp.  On the host those cost 10-35 ms per 1280x720 frame; here it is one pass over
// the frame on the device, fed by the raw uint8 upload.
//
// The arithmetic follows OpenCV's remap: source coordinates in fp64 from the inverse matrix, quantised to 1/32
// pixel (INTER_TAB_SIZE), bilinear weights from the quantised fractions, taps outside the image contribute zero.
// OpenCV blends 8-bit pixels with 15-bit fixed-point weights; this kernel blends in fp32 and rounds, which differs
// from it by at most one grey level.  The mask is 1 where any tap with a non-zero weight lies inside the image.
#include "common.cuh"

namespace woft {

struct WarpParams {
  double m[9];  // inverse map: destination (x, y, 1) -> source
  const uint8_t* src;
  uint8_t* dst;
  uint8_t* mask;  // may be null
  int h, w, channels;
};

__global__ void __launch_bounds__(256) warp_perspective_u8_kernel(const WarpParams p) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y;
  if (x >= p.w) return;
  const double X0 = p.m[0] * x + p.m[1] * y + p.m[2];
  const double Y0 = p.m[3] * x + p.m[4] * y + p.m[5];
  double Wd = p.m[6] * x + p.m[7] * y + p.m[8];
  Wd = Wd != 0.0 ? 32.0 / Wd : 0.0;
  const double fx = fmax(-2147483648.0, fmin(2147483647.0, X0 * Wd));
  const double fy = fmax(-2147483648.0, fmin(2147483647.0, Y0 * Wd));
  const int X = __double2int_rn(fx), Y = __double2int_rn(fy);  // cvRound: nearest, ties to even
  const int sx = X >> 5, sy = Y >> 5;
  const float a = float(X & 31) * (1.0f / 32.0f), b = float(Y & 31) * (1.0f / 32.0f);
  const float w00 = (1.f - a) * (1.f - b), w01 = a * (1.f - b), w10 = (1.f - a) * b, w11 = a * b;
  const bool in_x0 = sx >= 0 && sx < p.w, in_x1 = sx + 1 >= 0 && sx + 1 < p.w;
  const bool in_y0 = sy >= 0 && sy < p.h, in_y1 = sy + 1 >= 0 && sy + 1 < p.h;
  const int c = p.channels;
  const uint8_t* r0 = p.src + ((size_t)(in_y0 ? sy : 0) * p.w) * c;
  const uint8_t* r1 = p.src + ((size_t)(in_y1 ? sy + 1 : 0) * p.w) * c;
  uint8_t* o = p.dst + ((size_t)y * p.w + x) * c;
  for (int k = 0; k < c; ++k) {
    float v = 0.f;
    if (in_y0 && in_x0) v += w00 * float(r0[(size_t)sx * c + k]);
    if (in_y0 && in_x1) v += w01 * float(r0[(size_t)(sx + 1) * c + k]);
    if (in_y1 && in_x0) v += w10 * float(r1[(size_t)sx * c + k]);
    if (in_y1 && in_x1) v += w11 * float(r1[(size_t)(sx + 1) * c + k]);
    o[k] = (uint8_t)fminf(fmaxf(rintf(v), 0.f), 255.f);
  }
  if (p.mask) {
    const bool any = (in_y0 && in_x0 && w00 > 0.f) || (in_y0 && in_x1 && w01 > 0.f) || (in_y1 && in_x0 && w10 > 0.f) ||
                     (in_y1 && in_x1 && w11 > 0.f);
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
  dim3 grid(ceil_div(width, 256), height);
  warp_perspective_u8_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  return launch_status("warp_perspective_u8_kernel");
}
