/* woft_b200.h -- C ABI of libwoft_b200.so: the per-frame hot path of the WOFT planar tracker on B200 (sm_100a).
 *
 * The reference (serycjon/WOFT) is Python/PyTorch; its only FFI is the pybind11 module `alt_cuda_corr`
 * (pytracking/external/RAFT/alt_cuda_corr/correlation.cpp:23-54).  Every entry point below therefore cites
 * the reference *operator* it replaces; INTEGRATION.md shows the ctypes binding a maintainer adds on the
 * reference side.
 *
 * Conventions
 *   - all pointers are DEVICE pointers unless the parameter comment says "host"; tensors are dense, row-major
 *     (C-contiguous) fp32 in exactly the layout the reference uses at that seam;
 *   - the caller (PyTorch in this repo) owns every buffer, including workspaces; the library never allocates,
 *     frees or keeps a pointer after the call returns (two stated exceptions: woft_corr_pyramid_static uses a
 *     fixed scratch array inside the library, woft_corr_set_trace_buffer keeps the debug pointer it is given);
 *   - `stream` is a cudaStream_t passed as void*; work is enqueued on it and the call returns without
 *     synchronising;
 *   - return value: WOFT_OK or a negative-free error code below; `woft_last_error()` gives a thread-local,
 *     human-readable reason.  No C++ exception crosses the boundary.
 */
#ifndef WOFT_B200_H
#define WOFT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WOFT_ABI_VERSION 1

enum woft_status {
  WOFT_OK = 0,
  WOFT_ERR_INVALID_ARGUMENT = 1, /* bad shape, null pointer, unsupported radius / level count */
  WOFT_ERR_CUDA = 2,             /* a CUDA runtime / driver call failed; see woft_last_error() */
  WOFT_ERR_UNSUPPORTED = 3,      /* valid request this build cannot serve (e.g. tensor-core path off sm_100) */
  WOFT_ERR_WORKSPACE = 4         /* workspace pointer null or smaller than woft_corr_workspace_bytes() */
};

/* Arithmetic of the all-pairs correlation (K1).  The reference multiplies in fp32 (corr.py:67). */
enum woft_corr_mode {
  WOFT_CORR_FP32_SIMT = 0, /* fp32 FMA on CUDA cores: same precision class as the reference, slowest */
  WOFT_CORR_TF32 = 1,      /* tcgen05 kind::tf32, operands rounded to tf32 once, fp32 accumulate in TMEM */
  WOFT_CORR_TF32X3 = 2,    /* tcgen05, error-compensated 3xTF32 split (hi*hi + hi*lo + lo*hi): fp32-class result */
  WOFT_CORR_BF16 = 3,      /* tcgen05 kind::f16 with bf16 operands, fp32 accumulate */
  WOFT_CORR_FP16 = 4,      /* tcgen05 kind::f16 with fp16 operands (tf32's 10-bit mantissa at half the bytes; inputs
                              are saturated to +-65504), fp32 accumulate */
  WOFT_CORR_FP16X3 = 5     /* tcgen05 kind::f16, error-compensated split into fp16 hi + lo (hi*hi + hi*lo + lo*hi): the
                              fp32-class result of WOFT_CORR_TF32X3 at half its operand traffic */
};
/* OR-ed into `mode` of woft_corr_pyramid[_static] (tensor-core modes only): fmap1 / fmap2 are channels-last
 * (B, H, W, C) instead of (B, C, H, W), as a channels-last encoder leaves them; the result is the same volume. */
#define WOFT_CORR_FMAPS_NHWC 0x100
/* OR-ed into `mode` (tensor-core modes only): the pyramid levels are written as fp16 (IEEE half), each value rounded
 * once from its fp32 accumulator / fp32 pooled mean -- half the bytes K1 writes and K3 gathers.  The `levels` pointers
 * then address __half arrays of the same shapes; read them with woft_corr_lookup flag 32.  A stated-tolerance storage
 * format (relative rounding 2^-11 per stored value), not the reference's fp32 volume. */
#define WOFT_CORR_STORE_F16 0x200
/* OR-ed into `mode` (WOFT_CORR_FP16 only, channels a multiple of 64): fmap1 / fmap2 are channels-last fp16 tensors
 * (B, H, W, C) -- already the K-major fp16 operands of the GEMM, as an fp16 encoder epilogue leaves them -- so the
 * packing pass disappears.  Values beyond +-65504 must not occur (the packing pass of fp32 maps saturates them). */
#define WOFT_CORR_FMAPS_F16 0x400

/* Estimators of pytracking/utils/least_squares_H.py served by woft_homography_fit (K4). */
enum woft_h_method {
  WOFT_H_NONHOMOGENEOUS = 0, /* find_homography_nonhomogeneous_QR (:142-210) and ..._nonhomogeneous (:56-139):
                                8-parameter system, rows scaled by w, least squares */
  WOFT_H_DLT = 1,            /* find_homography_dlt (:213-265): 9x9 Gram A^T (w A), smallest eigenvector */
  WOFT_H_IRLS_L1 = 2,        /* find_homography_IRLSq_QR (:280-346) with IRLSq_L1 (:268-269) */
  WOFT_H_IRLS_HUBER = 3      /* find_homography_IRLSq_QR with IRLSq_Huber(k) (:272-277) */
};

int woft_abi_version(void);
const char* woft_last_error(void);

/* Bytes of scratch woft_corr_pyramid needs for this shape and mode (0 if an extent is not positive). */
size_t woft_corr_workspace_bytes(int batch, int channels, int height, int width, int mode);

/* K1 + K2.  Replaces CorrBlock.__init__ + CorrBlock.corr
 * (pytracking/external/RAFT/raft_core/corr.py:13-27, 61-69): all-pairs correlation
 *   level0[b*H*W + i][j] = sum_c fmap1[b][c][i] * fmap2[b][c][j] / sqrt(C)
 * and `num_levels-1` successive 2x2 average poolings over the (H, W) plane of every query (floor on odd sizes).
 *   fmap1, fmap2 : (B, C, H, W)
 *   levels       : HOST array of `num_levels` device pointers; level l is (B*H*W, 1, H_l, W_l), H_l = H_{l-1}/2
 *   row_mean     : (B*H*W) mean of level 0 over its (H, W) plane, what WeightHead.forward
 *                  (weighted_raft.py:358-361) reduces from the volume; may be NULL
 *   workspace    : >= woft_corr_workspace_bytes(...) bytes, 256-byte aligned
 */
int woft_corr_pyramid(const float* fmap1, const float* fmap2, int batch, int channels, int height, int width,
                      int num_levels, float* const* levels, float* row_mean, void* workspace,
                      size_t workspace_bytes, int mode, void* stream);

/* woft_corr_pyramid for callers that cannot allocate.  No counterpart in the reference (there PyTorch allocates
 * whatever an op needs).  It is meant for code that only holds the input and output buffers it was handed and
 * has no device allocator to turn to -- a plug-in inside a host application that owns all GPU memory, a thin
 * binding from another language -- or that may not pay for cudaMalloc/cudaFree inside its frame loop.
 * Arguments are those of woft_corr_pyramid without the workspace: the scratch lives in a fixed 32 MiB device
 * array inside the library (held by every GPU that loads the library, until the process exits).
 *   - Shapes/modes whose woft_corr_workspace_bytes() exceeds 32 MiB are refused with WOFT_ERR_WORKSPACE; one
 *     1280x720 frame pair (90x160x256 feature maps) fits in the fp32, tf32, bf16, fp16 and fp16x3 modes (14.1 to
 *     28.2 MiB), 3xTF32 (56.3 MiB) does not.
 *   - The first call on each device looks up the array's device address (cudaGetSymbolAddress); make that first
 *     call with the device current and outside any stream capture.
 *   - All calls on a device share the one array.  Calls whose GPU work can overlap (different streams, or
 *     different host threads) race on it: serialise them.  Nothing is carried over between calls: the status
 *     word is cleared and every scratch byte that is read has been rewritten by the same call.
 *   - No cudaMalloc, cudaFree or other allocating call is made. */
int woft_corr_pyramid_static(const float* fmap1, const float* fmap2, int batch, int channels, int height, int width,
                             int num_levels, float* const* levels, float* row_mean, int mode, void* stream);

/* Synchronises `stream` and reports whether the last woft_corr_pyramid on `workspace` ran to completion.  The
 * tensor-core kernel bounds every pipeline wait (2 s) and raises a flag in the workspace instead of hanging;
 * this reads the flag back (one 4-byte device-to-host copy).  No counterpart in the reference.
 * woft_corr_check_static is the same check after woft_corr_pyramid_static (whose scratch the caller never sees),
 * on the current device. */
int woft_corr_check(const void* workspace, void* stream);
int woft_corr_check_static(void* stream);

/* Debug aid, no counterpart in the reference: per-role timestamps (TMA producer / MMA issuer / epilogue) of the first
 * four CTAs of the tensor-core correlation kernel.  `device_buffer` is caller-allocated, zeroed device memory of
 * 4 * 64 * 6 unsigned long long (tile-major, nanoseconds of %globaltimer), or NULL to switch tracing off; the library
 * keeps the pointer until it is replaced.  Every later woft_corr_pyramid[_static] on any stream writes into it. */
void woft_corr_set_trace_buffer(unsigned long long* device_buffer);

/* K3.  Replaces CorrBlock.__call__ + bilinear_sampler (corr.py:29-59, raft_core/utils/utils.py:59-73):
 * for every query pixel and level l a (2r+1)^2 window of level l, centred at coords / 2^l, bilinear,
 * zero outside the map.
 *   levels : HOST array of `num_levels` device pointers as written by woft_corr_pyramid (h2 x w2 at level 0)
 *   coords : (B, 2, H1, W1), channel 0 = x, channel 1 = y
 *   out    : (B, num_levels*(2r+1)^2, H1, W1); channel = l*(2r+1)^2 + (xoff+r)*(2r+1) + (yoff+r)
 *   Flag 64 (with the channels-last output, radius 4, at most 4 levels, level rows that are multiples of 16 bytes): the
 *   footprints are fetched by TMA, one 16 x 10 box per query and level (lookup_tma.cu) -- same results bit for bit,
 *   measured slower than the default kernel; woft_corr_lookup_check synchronises the stream and reports whether such a
 *   launch gave up on a load (2 s bound).
 *   flags  : bit 0 = scalar loads (testing); bit 3 (8) = write `out` channels-last, (B, H1, W1, channels), for an
 *            NHWC consumer (same values); bit 4 (16, with bit 3) = `out` holds fp16 with the channels of a query padded
 *            to a multiple of 8 (328 for 4 levels of radius 4; the padding is not written): the operand of an fp16
 *            tensor-core convolution, whose TMA loads need 16-byte multiples between pixels; bit 5 (32) = the levels hold
 *            fp16 (written with WOFT_CORR_STORE_F16): interpolation in fp32 as before; levels whose width is a multiple
 *            of 8 are then read in 16-byte pieces (2-3 loads per 20-byte footprint row instead of 3-4) -- bit 1 (2) keeps
 *            the 8-byte pieces (testing)
 * Level maps narrower than 2 pixels make the reference return NaN (division by W-1 = 0, utils.py:63-64);
 * this call rejects them with WOFT_ERR_INVALID_ARGUMENT instead.
 */
int woft_corr_lookup(const float* const* levels, int num_levels, const float* coords, int batch, int h1, int w1,
                     int h2, int w2, int radius, float* out, int flags, void* stream);
int woft_corr_lookup_check(void* stream);
/* woft_corr_lookup preceded, in the same launch, by woft_flow_head_gather for the launch's own queries (the step between
 * two refinement iterations, raft_core/weighted_raft.py:244-246 then :232-233): `coords` (B,2,H1,W1) is advanced IN PLACE
 * by the flow head's delta (bias + the nine shifted tap products of `taps`, (B*H1*W1, 9, 2)), the new flow
 * coords - coords0 goes to flow_nchw / flow_nhwc / the two fp16 operand buffers (each optional), and the lookup is taken at
 * the new coordinates.  Bit for bit what the two separate calls give; one launch and one pass over the coordinates less
 * per iteration.  Not with flag 64. */
int woft_corr_lookup_gather(const float* const* levels, int num_levels, float* coords, int batch, int h1, int w1, int h2,
                            int w2, int radius, float* out, int flags, const float* taps, const float* bias,
                            const float* coords0, float* flow_nchw, float* flow_nhwc, void* flow_f16_a, void* flow_f16_b,
                            int f16_stride, int f16_offset, void* stream);

/* A''.  Replaces alt_cuda_corr.forward (alt_cuda_corr/correlation.cpp:23-33, correlation_kernel.cu:18-119,
 * 260-286), the reference's one native op: on-the-fly local correlation without a materialised volume.
 *   fmap1 : (B, H1, W1, C) channels-last      fmap2 : (B, H2, W2, C) channels-last
 *   coords: (B, N, H1, W1, 2) as (x, y)        out   : (B, N, (2r+1)^2, H1, W1), UNSCALED dot products,
 *   channel = (yoff+r) + (2r+1)*(xoff+r).  Unlike the reference op the output is caller-allocated.
 */
int woft_alt_corr_forward(const float* fmap1, const float* fmap2, const float* coords, int batch, int h1, int w1,
                          int h2, int w2, int channels, int n_coords, int radius, float* out, void* stream);

/* K4.  Replaces find_homography_nonhomogeneous_QR / _nonhomogeneous / _dlt / _IRLSq_QR
 * (pytracking/utils/least_squares_H.py:56-346) including the Hartley normalisation they call
 * (kornia.geometry.epipolar.normalize_points) and the final de-normalisation and division by H33 + 1e-8.
 *   points1, points2 : (B, N, 2)      weights : (B, N) or NULL      H_out : (B, 3, 3), maps points1 -> points2
 *   n_iter, huber_k  : IRLS only (reference defaults 5 and 1); n_iter+1 solves are made
 * One launch per call: normalisation, Gram accumulation in fp64, in-kernel 8x8 Cholesky or 9x9 Jacobi.
 */
int woft_homography_fit(const float* points1, const float* points2, const float* weights, int batch, int n_points,
                        int method, int n_iter, float huber_k, float* H_out, void* stream);

/* Replaces torch_proj_errors (least_squares_H.py:474-489): L2 distance between H*pts_a and pts_b.
 *   H : (B, 3, 3)    pts_a, pts_b : (B, 2, N)    err : (B, N)
 */
int woft_proj_errors(const float* H, const float* pts_a, const float* pts_b, int batch, int n_points, float* err,
                     void* stream);

/* Fused element-wise stages between the update block's cuDNN convolutions.  Replace the gate arithmetic of
 * SepConvGRU.forward (pytracking/external/RAFT/raft_core/update.py:45-60), per half step:
 *   woft_gru_gates : zr = bias-free conv of [h, x] with convz and convr weights stacked, (B, 2C, HW);
 *                    z = sigmoid(zr[:, :C] + bz), rh = sigmoid(zr[:, C:] + br) * h      -> (B, C, HW) each
 *   woft_gru_update: h_new = (1 - z) * h + z * tanh(q + bq), q = bias-free convq([rh, x])
 * NCHW fp32 contiguous, H*W a multiple of 4; bias_zr is [bz; br] (2C). */
int woft_gru_gates(const float* zr, const float* bias_zr, const float* h, int batch, int channels, int hw, float* z,
                   float* rh, void* stream);
int woft_gru_update(const float* q, const float* bias_q, const float* z, const float* h, int batch, int channels,
                    int hw, float* h_new, void* stream);

/* Channels-last (NHWC) forms of the two GRU kernels, and the concatenation that feeds the update block's
 * convolutions (torch.cat at update.py:47,51,54,58,96-97,129).  Tensors are (pixels, channels) with the channels of a
 * pixel contiguous; `*_stride` is the distance in floats between consecutive pixels, so an operand can be a channel
 * slice of a wider buffer -- the GRU keeps [h | x] and [r*h | x] in two persistent buffers and never concatenates.
 *   woft_gru_gates_nhwc : zr (pixels, 2C) dense; h, rh strided; z (pixels, C) dense
 *   woft_gru_gates_nhwc_split: the same with the z and r pre-activations in two (pixels, C) operands of stride
 *                         pre_stride and their biases apart -- for a schedule that runs convz and convr as two
 *                         convolutions (cuDNN has no sm_100 kernel for the stacked 1x5 384->256 shape)
 *   woft_gru_update_nhwc: h_out may alias h; h_dense (pixels, C) is an optional second, dense copy of the result
 *   woft_concat_nhwc    : up to 4 dense sources (HOST arrays of pointers / channel counts) written side by side at
 *                         channel offset dst_offset of dst (and of dst2 when not NULL) */
int woft_gru_gates_nhwc(const float* zr, const float* bias_zr, const float* h, int h_stride, long long pixels,
                        int channels, float* z, float* rh, int rh_stride, void* stream);
int woft_gru_gates_nhwc_split(const float* z_pre, const float* r_pre, int pre_stride, const float* bias_z,
                              const float* bias_r, const float* h, int h_stride, long long pixels, int channels,
                              float* z, float* rh, int rh_stride, void* stream);
int woft_gru_update_nhwc(const float* q, const float* bias_q, const float* z, const float* h, int h_stride,
                         long long pixels, int channels, float* h_out, int h_out_stride, float* h_dense, void* stream);
int woft_concat_nhwc(const float* const* srcs, const int* channels, int n_src, long long pixels, float* dst,
                     int dst_stride, int dst_offset, float* dst2, int dst2_stride, int dst2_offset, void* stream);
/* The same with a pixel stride per source (HOST array, floats, >= channels[i]): a source may be the leading channels
 * of a wider tensor, e.g. the 126 motion features in front of the two padding channels of the encoder's last conv. */
int woft_concat_nhwc_strided(const float* const* srcs, const int* channels, const int* src_strides, int n_src,
                             long long pixels, float* dst, int dst_stride, int dst_offset, float* dst2, int dst2_stride,
                             int dst2_offset, void* stream);

/* Two thin convolutions of the update block, 12 calls each per flow pass, fp32 on CUDA cores:
 *   woft_flow_conv7x7_relu : BasicMotionEncoder.convf1 + ReLU (update.py:85, 92): 7x7, pad 3, 2 -> 128 channels.
 *                            flow (B,2,H,W) NCHW, weight (128,2,7,7), bias (128) -> out (B,H,W,128) channels-last
 *   woft_flow_out3x3       : FlowHead.conv2 (update.py:10, 14): 3x3, pad 1, C -> 2 channels + bias.
 *                            in (B,H,W,C) channels-last, weight (2,C,3,3), bias (2) -> out (B,2,H,W) NCHW */
int woft_flow_conv7x7_relu(const float* flow, const float* weight, const float* bias, int batch, int height, int width,
                           float* out_nhwc, void* stream);
int woft_flow_out3x3(const float* in_nhwc, const float* weight, const float* bias, int batch, int height, int width,
                     int channels, float* out_nchw, void* stream);
/* woft_flow_conv7x7_relu with an fp16 output tensor (B,H,W,128): the operand of an fp16 tensor-core convolution. */
int woft_flow_conv7x7_relu_f16(const float* flow, const float* weight, const float* bias, int batch, int height,
                               int width, void* out_nhwc_f16, void* stream);
/* fp32 channels-last values (pixels, src_stride), the first `channels` of every pixel, rounded to fp16 into channel
 * slice [dst_offset, dst_offset + channels) of one or two fp16 tensors (pixels, dst_stride): fills the fp16 operand
 * buffers of woft_conv_tc from fp32 producers (context features, hidden state, the flow behind the motion features,
 * update.py:97).  Even channel counts, strides and offsets. */
int woft_cast_nhwc_f16(const float* src, int src_stride, int channels, long long pixels, void* dst, int dst_stride,
                       int dst_offset, void* dst2, int dst2_stride, int dst2_offset, void* stream);

/* FlowHead.conv2 and the step of the refinement loop that follows it (raft_core/weighted_raft.py:244-246:
 * coords1 = coords1 + delta_flow; next iteration's flow = coords1 - coords0) in two launches that read every input
 * value once: per input pixel the 18 (tap, output) products, then per output pixel the sum of 9 shifted products.
 *   in (B,H,W,C) channels-last, 16-byte aligned, C a multiple of 32 and <= 320; weight (2,C,3,3); bias (2)
 *   scratch: woft_flow_head_scratch_bytes(B,H,W) bytes
 *   delta_nchw (B,2,H,W) or NULL: the convolution's result, as woft_flow_out3x3 gives it
 *   coords (B,2,H,W) or NULL: updated IN PLACE by the delta; with it coords0 (B,2,H,W) is required and
 *   flow_nchw (B,2,H,W) / flow_nhwc (B,H,W,2), each optional, receive coords - coords0 */
size_t woft_flow_head_scratch_bytes(int batch, int height, int width);
/* The second half of woft_flow_head_step alone, for tap products formed elsewhere (woft_conv_tc, WOFT_CONV_FLOW_TAPS):
 * taps (B*H*W, 9, 2).  flow_f16_a / flow_f16_b (optional, with coords): fp16 channels-last tensors (pixels, f16_stride)
 * that also receive the new flow, at channel offset f16_offset (the GRU operand buffers, update.py:97). */
int woft_flow_head_gather(const float* taps, const float* bias, int batch, int height, int width, float* delta_nchw,
                          float* coords, const float* coords0, float* flow_nchw, float* flow_nhwc, void* flow_f16_a,
                          void* flow_f16_b, int f16_stride, int f16_offset, void* stream);
int woft_flow_head_step(const float* in_nhwc, const float* weight, const float* bias, int batch, int height, int width,
                        int channels, void* scratch, size_t scratch_bytes, float* delta_nchw, float* coords,
                        const float* coords0, float* flow_nchw, float* flow_nhwc, void* stream);

/* Replaces nn.InstanceNorm2d + ReLU after each convolution of the feature encoder
 * (raft_core/extractor.py:30-32, 49-50, 127, 180-181): per (n, c) plane y = (x - mean) / sqrt(var + eps), biased
 * variance, no affine parameters, followed by ReLU when `relu` is non-zero (the 1x1 projection branch has none).
 * x, y: (planes, HW) fp32, may alias. */
int woft_instance_norm_relu(const float* x, int planes, int hw, float eps, int relu, float* y, void* stream);

/* Replaces the two cv2.warpPerspective calls of YAOFTrackerSingleControl.track
 * (pytracking/tracker/YAOF_tracker_single_control.py:89-95): bilinear warp of an 8-bit frame with a zero border,
 * plus the validity mask the reference gets from warping an image of ones (`> 0`).
 *   src, dst   : (H, W, C) uint8, C <= 4           valid_mask : (H, W) uint8 0/1, may be NULL
 *   inv_matrix_host : HOST pointer to 9 doubles, row-major, mapping destination (x, y, 1) to source -- the inverse of
 *                     the matrix cv2.warpPerspective is given (it inverts it unless WARP_INVERSE_MAP is set)
 *                     (formed by cv::invert's cofactor formula, woft_b200.tracker.invert3x3_opencv, for identical bits)
 * Integer arithmetic of OpenCV's remap throughout (1/32-px coordinates, 15-bit weight table, (v + 2^14) >> 15):
 * frame and mask are bit-identical to cv2's. */
int woft_warp_perspective_u8(const uint8_t* src, int height, int width, int channels, const double* inv_matrix_host,
                             uint8_t* dst, uint8_t* valid_mask, void* stream);

/* Correspondence selection of the tracker's global branch without a host round trip.  Replaces the 'TC' assembly of
 * RAFTWrapper.compute_flow (pytracking/optical_flow/raft.py:185-218), YAOFTrackerSingleControl._mask_coords
 * (pytracking/tracker/YAOF_tracker_single_control.py:287-312) and the Sobol subsampler of the default config
 * (pytracking/configs/YAOFT_single_control_repRAFT_sub500_noreliableinl_wLSq.py:31-53):
 *   keep pixel i  iff  template_mask[i] and cur = pixel + flow is not < 0 and round(cur) is inside the frame and
 *                      pw_mask[round(cur)] (pw_mask may be NULL);
 *   of the N kept pixels, in raster order, those of rank rint(float(N) * u_k) for the first `to_draw` points u_k of
 *   the unscrambled 1-D Sobol sequence (duplicates once; all of them when N <= to_draw).
 *   flow_up (2,H,W); weights_up (H*W) logits or probabilities, or NULL (weights 1); masks (H,W) uint8;
 *   sobol_sorted: DEVICE array of the `to_draw` (<= 1024) Sobol points sorted ascending;
 *   outputs pts_template / pts_cur (to_draw, 2) as (x, y), weights_out (to_draw), n_selected (1 int), all DEVICE.
 * scratch: woft_select_scratch_bytes(H, W) bytes.  woft_homography_fit_n is woft_homography_fit for one problem
 * whose point count is read from device memory (fewer than 4 points give a NaN matrix instead of the reference's
 * AssertionError); woft_inlier_fraction is the default config's redet_success_fn (:14-21): the fraction of points
 * with |H a - b| <= threshold (fraction_out[0]; fraction_out[1] receives the point count) and whether it exceeds
 * min_fraction (ok_out). */
size_t woft_select_scratch_bytes(int height, int width);
int woft_select_correspondences(const float* flow_up, const float* weights_up, const uint8_t* template_mask,
                                const uint8_t* pw_mask, int height, int width, const float* sobol_sorted, int to_draw,
                                int apply_sigmoid, float* pts_template, float* pts_cur, float* weights_out,
                                int* n_selected, void* scratch, size_t scratch_bytes, void* stream);
/* The same for the tracker's lost branch (YAOF_tracker_single_control.py:178-204, 314-327): flow_up runs from the
 * previous frame to the current one and a previous-frame pixel is kept iff the template mask covers it when seen
 * through prev_H2init (HOST, 9 doubles, row-major): template_mask[round(prev_H2init (x, y, 1))], zero outside -- what
 * cv2.warpPerspective(template_mask, inv(prev_H2init), flags=INTER_NEAREST) > 0 gives.  No bounds test on the targets.
 * pts_prev receives the previous-frame pixels, pts_cur = pixel + flow. */
int woft_select_correspondences_local(const float* flow_up, const float* weights_up, const uint8_t* template_mask,
                                      const double* prev_to_template_host, int height, int width,
                                      const float* sobol_sorted, int to_draw, int apply_sigmoid, float* pts_prev,
                                      float* pts_cur, float* weights_out, int* n_selected, void* scratch,
                                      size_t scratch_bytes, void* stream);
int woft_homography_fit_n(const float* points1, const float* points2, const float* weights, const int* n_points_dev,
                          int max_points, int method, int n_iter, float huber_k, float* H_out, void* stream);
int woft_inlier_fraction(const float* H, const float* pts_a, const float* pts_b, const int* n_points_dev, float threshold,
                         float min_fraction, float* fraction_out, uint8_t* ok_out, void* stream);

/* Replaces WeightedRAFT.upsample_flow for the flow and for the weight logits in one pass
 * (raft_core/weighted_raft.py:92-103, 249-255, 285-288) and, optionally, the sigmoid of
 * pytracking/optical_flow/raft.py:158-159: convex combination of the 3x3 coarse neighbourhood of 8*field with softmax
 * weights from `mask`, pixel-shuffled to full resolution; the weights are divided by 8 again.
 *   flow_low (B,2,H,W), weights_low (B,1,H,W) or NULL, mask (B,576,H,W) NCHW or channels-last (mask_nhwc != 0)
 *   -> flow_up (B,2,8H,8W), weights_up (B,1,8H,8W) or NULL */
int woft_convex_upsample(const float* flow_low, const float* weights_low, const float* mask, int mask_nhwc, int batch,
                         int height, int width, int apply_sigmoid, float* flow_up, float* weights_up, void* stream);

/* The same instance norm (+ ReLU) on channels-last tensors x, y: (batch, HW, channels).  `scratch` holds the per-strip
 * partial sums and the per-channel mean / rstd: woft_instance_norm_nhwc_scratch_bytes(batch, channels) bytes. */
size_t woft_instance_norm_nhwc_scratch_bytes(int batch, int channels);
int woft_instance_norm_relu_nhwc(const float* x, int batch, int hw, int channels, float eps, int relu, float* y,
                                 void* scratch, size_t scratch_bytes, void* stream);

/* The tail of a residual block in one pass (raft_core/extractor.py:45-56: y = relu(norm2(conv2(y))); return relu(x+y)):
 *   woft_instance_norm_relu_add_nhwc : y = relu(residual + relu(instance_norm(x))), all (batch, HW, channels)
 *   woft_add_relu                    : y = relu(a + b) over n floats (the batch-norm encoder, whose norm is folded
 *                                      into the convolution); y may alias a or b */
int woft_instance_norm_relu_add_nhwc(const float* x, int batch, int hw, int channels, float eps, const float* residual,
                                     float* y, void* scratch, size_t scratch_bytes, void* stream);
int woft_add_relu(const float* a, const float* b, long long n, float* y, void* stream);
/* The channels-last instance norm on fp16 tensors (statistics and arithmetic in fp32): the form the encoders take when
 * their convolutions run on woft_conv_tc with fp16 operands.  y = act(norm(x)), act = ReLU when `relu`; with `residual`
 * (fp16, same shape; relu must be set) y = relu(residual + relu(norm(x))).  x, y: (batch, HW, channels) fp16. */
int woft_instance_norm_relu_nhwc_f16(const void* x, int batch, int hw, int channels, float eps, int relu,
                                     const void* residual, void* y, void* scratch, size_t scratch_bytes, void* stream);
/* The same norm with the statistics taken from the partial sums woft_conv_tc gathered while it wrote x (`stats` of its
 * argument block: (rows, batch, 2, channels) fp32): two launches (combine, apply) instead of three, and x is read once.
 * scratch: batch * 2 * channels floats. */
int woft_instance_norm_from_sums_f16(const void* x, const float* partial, int rows, int batch, int hw, int channels,
                                     float eps, int relu, const void* residual, void* y, void* scratch,
                                     size_t scratch_bytes, void* stream);
/* x = scale * (x + bias[c]) in place on a channels-last (pixels, channels) tensor, optional ReLU: the bias of the
 * convolutions that have no activation behind them (encoder outputs, extractor.py:150; downsampling 1x1s, :36-37; the
 * upsampling-mask head with its 0.25, update.py:122-125, 135). */
/* Tail of the context encoder (raft_core/weighted_raft.py:215-219): y (pixels, 2C) is cnet's output without its bias;
 * h = tanh(y[:, :C] + b[:C]) goes to h_dense (pixels, C) and to channels [0, C) of buf1, the context
 * relu(y[:, C:] + b[C:]) to channels [C, 2C) of buf1 and buf2 (both (pixels, buf_stride)). */
int woft_context_split_nhwc(const float* y, const float* bias, long long pixels, int channels, float* h_dense, float* buf1,
                            float* buf2, int buf_stride, void* stream);
int woft_bias_act_nhwc(float* x, const float* bias, long long pixels, int channels, float scale, int relu, void* stream);

/* The flow network's input from the two uint8 frames in one launch: BGR -> RGB, 2 * (v / 255) - 1
 * (pytracking/optical_flow/raft.py:113-120, raft_core/weighted_raft.py:197-198), both frames stacked as the batch the
 * feature encoder takes (extractor.py:171-174), channels-last.
 *   img1_bgr, img2_bgr (H,W,3) uint8 -> out (2,H,W,3) fp32, i.e. a (2,3,H,W) tensor in channels-last memory format */
int woft_images_to_input(const uint8_t* img1_bgr, const uint8_t* img2_bgr, int height, int width, float* out_nhwc,
                         void* stream);

/* First layer of the two encoders (raft_core/extractor.py:131, 171-178): 7x7, stride 2, pad 3, 3 -> 64 channels, as an
 * implicit GEMM on the tensor cores (mma.sync, TF32 operands rounded to nearest, fp32 accumulation).
 *   in (B,H,W,3) fp32 channels-last; weight (64,3,7,7); bias (64) or NULL; relu != 0 applies ReLU
 *   -> out (B,(H-1)/2+1,(W-1)/2+1,64) fp32 channels-last */
int woft_stem_conv7x7(const float* in_nhwc, const float* weight, const float* bias, int batch, int height, int width,
                      int relu, float* out_nhwc, void* stream);
/* The same with an fp16 output tensor, for encoders whose residual stages run on woft_conv_tc with fp16 operands. */
int woft_stem_conv7x7_f16(const float* in_nhwc, const float* weight, const float* bias, int batch, int height, int width,
                          int relu, void* out_nhwc_f16, void* stream);

/* First layer of the weight head (raft_core/weighted_raft.py:319-339, 347-384) with its input assembly: per query pixel
 * the 9x9x4 correlation window (the channels-last lookup output read as (hp, wp, level), :267-272) and the mean
 * response broadcast as a fifth channel (:358-361) -> conv3x3(5 -> 128, pad 1) + bias + ReLU, on the tensor cores
 * (mma.sync, TF32 operands, fp32 accumulation).
 *   window_nhwc (n, 9*9*4); mean_response (n); weight (128,5,3,3); bias (128)
 *   -> out (n, 9, 9, 128) channels-last.  patch, levels, channels_out must be 9, 4, 128 (the shipped head). */
int woft_weight_head_conv1(const float* window_nhwc, const float* mean_response, const float* weight, const float* bias,
                           long long n_images, int patch, int levels, int channels_out, float* out_nhwc, void* stream);
/* The same layer writing fp16 in the layout woft_conv_tc takes for layers 2 and 3: (n, 10, 9, 128), every image's 9 rows
 * of 9 pixels followed by one all-zero row (the vertical zero padding between the images stacked along H). */
int woft_weight_head_conv1_stacked_f16(const float* window_nhwc, const float* mean_response, const float* weight,
                                       const float* bias, long long n_images, int patch, int levels, int channels_out,
                                       void* out_stacked_f16, void* stream);

/* The wide convolutions of the update block and of the weight head as ONE implicit-GEMM kernel on the tcgen05 tensor
 * cores, with the element-wise stage that follows each of them in the reference fused into the epilogue.  Replaces the
 * nn.Conv2d calls (stride 1 or 2, zero padding) of the two encoders' residual stages (raft_core/extractor.py:6-57,
 * 118-192), of BasicMotionEncoder, SepConvGRU, FlowHead.conv1 and the mask head
 * (pytracking/external/RAFT/raft_core/update.py:6-14, 33-60, 79-97, 122-125) and layers 2-3 + the 1x1 and patch mean of
 * WeightHead (raft_core/weighted_raft.py:318-339, 347-384).
 *   in            (N, H, W, in_stride) channels-last, fp32 or fp16 (in_f16); the first `cin` channels of every pixel
 *                 are convolved, so `in` may be a channel slice of a wider buffer
 *   weight_packed [kh*kw][cout_pad][cin_pad] in the input's type, tap = ky*kw + kx, zero padded; cin_pad = cin rounded
 *                 up to 32 (fp32) / 64 (fp16), cout_pad a multiple of 32.  fp32 operands are used as TF32 (the tensor
 *                 core reads the top 19 bits): round the weights to TF32 beforehand for round-to-nearest
 *   box_w, box_h  tile of output pixels, box_w * box_h in [96, 128]; 0 = chosen by the library
 *   epilogue      WOFT_CONV_BIAS_ACT : out[pix][c] = scale * act(acc + bias[c]), act = ReLU when `relu`; fp32 or fp16
 *                                      (out_f16); `out2` (optional) receives the same values; strides in elements
 *                 WOFT_CONV_GRU_ZR   : cout = 2C stacked convz|convr filters; z[pix][c] = sigmoid(acc[c] + bias[c]);
 *                                      out[pix][c] = sigmoid(acc[C+c] + bias[C+c]) * h[pix][c]      (update.py:47-52)
 *                 WOFT_CONV_GRU_Q    : cout = C; out[pix][c] = (1 - z) * h + z * tanh(acc + bias), fp32, may alias h;
 *                                      out2 optional second copy, fp32 or fp16 (out2_f16)            (update.py:53-54)
 *                 (h and z are fp32; in GRU_ZR the r*h output may be fp16 (out_f16): it is a convolution operand)
 *                 WOFT_CONV_HEAD_MEAN: images of row_valid x W pixels stacked along H every row_period rows (batch 1);
 *                                      head_out[image] = head_bias + mean over the image of w1x1 . relu(acc + bias)
 *                 WOFT_CONV_FLOW_TAPS: FlowHead.conv1 + ReLU with the products of FlowHead.conv2 (update.py:6-14):
 *                                      head_out[pix][tap][o] = sum_c relu(acc[c] + bias[c]) * w2[o][c][tap] for the 9 taps and
 *                                      2 outputs; `w1x1` holds w2 as (cout, 24) floats, [taps 0-4 x 2 | pad | taps 5-8 x 2 |
 *                                      pad]; woft_flow_head_gather adds the nine shifted products per output pixel.  The
 *                                      256-channel activation is never written
 *   row_period, row_valid: when row_period > 0, rows y with y % row_period >= row_valid separate stacked images: the
 *                 BIAS_ACT epilogue stores zeros there (they are the next layer's zero padding between images)
 * One launch (two for HEAD_MEAN: the logits are cleared first).  woft_conv_tc_check synchronises the stream and reports
 * whether any convolution kernel gave up on a pipeline wait (2 s bound) since the library was loaded. */
enum woft_conv_epilogue { WOFT_CONV_BIAS_ACT = 0, WOFT_CONV_GRU_ZR = 1, WOFT_CONV_GRU_Q = 2, WOFT_CONV_HEAD_MEAN = 3,
                          WOFT_CONV_FLOW_TAPS = 4 };
typedef struct woft_conv_tc_args {
  const void* in;
  long long in_stride;
  int in_f16;
  int batch, height, width, cin, cin_pad;
  const void* weight_packed;
  int cout, cout_pad, kh, kw, pad_h, pad_w;
  int box_w, box_h;
  int epilogue, relu;
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
  const float* w1x1;
  float* head_out;
  float head_bias;
  int row_period, row_valid;
  void* debug_trace; /* NULL, or device memory of 8 * 148 unsigned long long: per-CTA %globaltimer stamps (kernel entry,
                        set-up done, first operands landed, last MMA issued, accumulator ready, epilogue done, exit) */
  int debug_flags;   /* 0; experiments only */
  int halo;          /* 1: fetch ONE activation box per tile and channel block -- the output tile plus its filter halo --
                        which every filter tap reads at its own row offset, instead of one shifted box per tap (kh * kw
                        times less activation traffic; the filters stream as before).  box_w x box_h is then the tile of
                        OUTPUT pixels, with (box_w + kw - 1) * box_h <= 128.
                        2 ("row halo", box_w % 8 == 0, box_w * box_h <= 128): one box per channel block and filter COLUMN,
                        with kh - 1 extra rows; the taps of that column read it ky * box_w rows further in, so every MMA
                        operand starts on an 8-row boundary: kw instead of kh * kw activation fetches at full MMA speed */
  int pair;          /* 1: clusters of two CTAs share each filter tile (tcgen05 cta_group::2, M = 256): every CTA fetches its
                        own activation box and half of the filter rows; ignored in halo mode */
  int pdl;           /* 1: programmatic dependent launch -- the kernel may be scheduled while its predecessor in the stream
                        still runs; it sets up (barriers, TMEM, filters' tensor map) and then waits for the predecessor
                        to complete before touching any tensor */
  int stride;        /* 0 / 1, or 2: output pixel (x, y) reads input pixels (2x + kx - pad_w, 2y + ky - pad_h); height and
                        width are the INPUT extents, the outputs have (H + 2 pad - k) / stride + 1 rows / columns */
  const void* residual; /* BIAS_ACT only, or NULL: fp16 tensor laid out like `out`; out = relu(residual + act(acc + bias)),
                           the tail of a residual block (raft_core/extractor.py:45-56) */
  long long residual_stride;
  float* stats;      /* BIAS_ACT with one fp16 output and at most two images, or NULL: (stats_ctas, batch, 2, cout) fp32; CTA k
                        writes the sums and sums of squares of the values it stored, per image and channel, into row k
                        (rows of CTAs the launch does not have are cleared): the statistics of an instance norm that
                        follows (woft_instance_norm_from_sums_f16), without a pass over the tensor */
  int stats_ctas;    /* rows of `stats`; at least the SM count (148) */
} woft_conv_tc_args;
int woft_conv_tc(const woft_conv_tc_args* args, void* stream);
int woft_conv_tc_check(void* stream);

#ifdef __cplusplus
}
#endif
#endif /* WOFT_B200_H */
