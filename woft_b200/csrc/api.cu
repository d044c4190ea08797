// C-ABI glue: error reporting, woft_corr_pyramid dispatch over the arithmetic modes, workspace sizing.
#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace woft {

char* last_error_buffer() {
  static thread_local char buf[512] = "";
  return buf;
}

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(last_error_buffer(), 512, fmt, ap);
  va_end(ap);
  return code;
}

int corr_pyramid_simt(const float*, const float*, int, int, int, int, int, float* const*, cudaStream_t);
int corr_pyramid_tc(const float*, const float*, int, int, int, int, int, float* const*, void*, int, cudaStream_t);
size_t corr_tc_workspace_bytes(int, int, int, int);
void corr_tc_set_trace_buffer(unsigned long long*);
int launch_row_mean(const float*, const float*, int, int, int, float*, float*, float*, bool, cudaStream_t);

bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("WOFT_PDL");
    return !(e && e[0] == '0');
  }();
  return on;
}

static size_t simt_workspace_bytes(int B, int C) { return 256 + ((size_t)B * C * 4 + 255) / 256 * 256; }

}  // namespace woft

extern "C" int woft_abi_version(void) { return WOFT_ABI_VERSION; }

extern "C" const char* woft_last_error(void) { return woft::last_error_buffer(); }

extern "C" size_t woft_corr_workspace_bytes(int batch, int channels, int height, int width, int mode) {
  if (batch <= 0 || channels <= 0 || height <= 0 || width <= 0) return 0;
  mode &= ~(WOFT_CORR_FMAPS_NHWC | WOFT_CORR_STORE_F16 | WOFT_CORR_FMAPS_F16);
  if (mode == WOFT_CORR_FP32_SIMT) return woft::simt_workspace_bytes(batch, channels);
  return woft::corr_tc_workspace_bytes(batch, channels, height * width, mode);
}

extern "C" int woft_corr_pyramid(const float* fmap1, const float* fmap2, int batch, int channels, int height,
                                 int width, int num_levels, float* const* levels, float* row_mean, void* workspace,
                                 size_t workspace_bytes, int mode, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(fmap1 && fmap2 && levels, "woft_corr_pyramid: null pointer");
  WOFT_REQUIRE(batch > 0 && channels > 0 && height > 0 && width > 0, "woft_corr_pyramid: non-positive extent");
  WOFT_REQUIRE(num_levels >= 1 && num_levels <= kMaxLevels, "woft_corr_pyramid: num_levels %d not in [1,%d]",
               num_levels, kMaxLevels);
  const bool fmaps_f16 = (mode & WOFT_CORR_FMAPS_F16) != 0;
  const bool fmaps_nhwc = (mode & WOFT_CORR_FMAPS_NHWC) != 0 || fmaps_f16;
  const bool store_f16 = (mode & WOFT_CORR_STORE_F16) != 0;
  const int mode_flags = mode;
  mode &= ~(WOFT_CORR_FMAPS_NHWC | WOFT_CORR_STORE_F16 | WOFT_CORR_FMAPS_F16);
  WOFT_REQUIRE(!fmaps_f16 || (mode == WOFT_CORR_FP16 && channels % 64 == 0 &&
                              ((reinterpret_cast<uintptr_t>(fmap1) | reinterpret_cast<uintptr_t>(fmap2)) & 15) == 0),
               "woft_corr_pyramid: fp16 feature maps need mode WOFT_CORR_FP16, channels %% 64 == 0 and 16-byte alignment");
  WOFT_REQUIRE(mode >= WOFT_CORR_FP32_SIMT && mode <= WOFT_CORR_FP16X3, "woft_corr_pyramid: unknown mode %d", mode);
  WOFT_REQUIRE(!((fmaps_nhwc || store_f16) && mode == WOFT_CORR_FP32_SIMT),
               "woft_corr_pyramid: channels-last feature maps and fp16 storage need one of the tensor-core modes");
  WOFT_REQUIRE((long long)batch * height * width < (1ll << 31), "woft_corr_pyramid: too many query pixels");
  for (int l = 0, h = height, w = width; l < num_levels; ++l, h /= 2, w /= 2) {
    WOFT_REQUIRE(h >= 1 && w >= 1, "woft_corr_pyramid: level %d is empty", l);
    WOFT_REQUIRE(levels[l], "woft_corr_pyramid: level %d pointer is null", l);
  }
  const size_t need = woft_corr_workspace_bytes(batch, channels, height, width, mode);
  if (!workspace || workspace_bytes < need)
    return fail(WOFT_ERR_WORKSPACE, "woft_corr_pyramid: workspace of %zu bytes given, %zu needed", workspace_bytes,
                need);
  WOFT_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "woft_corr_pyramid: workspace not 256-byte aligned");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  WOFT_CUDA_CHECK(cudaMemsetAsync(workspace, 0, sizeof(int), s));  // the abort flag woft_corr_check reads
  int rc;
  if (mode == WOFT_CORR_FP32_SIMT) {
    rc = corr_pyramid_simt(fmap1, fmap2, batch, channels, height, width, num_levels, levels, s);
  } else {
    rc = corr_pyramid_tc(fmap1, fmap2, batch, channels, height, width, num_levels, levels, workspace, mode_flags, s);
  }
  if (rc != WOFT_OK) return rc;
  if (row_mean) {
    float* mean2 = reinterpret_cast<float*>(static_cast<uint8_t*>(workspace) + 256);
    // channels-last maps: the per-strip partial sums go where the packed operands were (the GEMM is done with them);
    // that region starts at the first 1 KiB boundary behind mean2 and is far larger than strips * B * C floats
    float* scratch = nullptr;
    if (fmaps_nhwc)
      scratch = reinterpret_cast<float*>(static_cast<uint8_t*>(workspace) +
                                         (256 + (size_t)batch * channels * 4 + 1023) / 1024 * 1024);
    rc = launch_row_mean(fmap1, fmap2, batch, channels, height * width, row_mean, mean2, scratch, fmaps_f16, s);
  }
  return rc;
}

extern "C" void woft_corr_set_trace_buffer(unsigned long long* device_buffer) {
  woft::corr_tc_set_trace_buffer(device_buffer);
}

extern "C" int woft_corr_check(const void* workspace, void* stream) {
  using namespace woft;
  WOFT_REQUIRE(workspace, "woft_corr_check: null workspace");
  int flag = 0;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  WOFT_CUDA_CHECK(cudaMemcpyAsync(&flag, workspace, sizeof(int), cudaMemcpyDeviceToHost, s));
  WOFT_CUDA_CHECK(cudaStreamSynchronize(s));
  if (flag != 0) return fail(WOFT_ERR_CUDA, "correlation kernel aborted: a pipeline barrier timed out (flag=%d)", flag);
  return WOFT_OK;
}
