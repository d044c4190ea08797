// woft_corr_pyramid_static: the correlation pyramid for callers that cannot hand the library a workspace.
//
// woft_corr_pyramid needs scratch memory (the K-major repack of both feature maps, the per-channel means and
// a status word) and asks the caller to allocate it.  Some callers have no way to do that: a plug-in that is
// handed its input and output buffers by a host application which owns all device memory, a binding from a
// language runtime that has no device allocator of its own, or a frame loop with a latency budget in which a
// cudaMalloc / cudaFree pair (both of which may synchronise the whole device) is not acceptable.  For them this
// entry point keeps the scratch inside the library, in a file-scope device array of fixed size.  The price is
// stated once here: every GPU on which the library gets loaded carries kStaticScratchBytes of device memory
// until the process exits, whether or not this function is ever called, and larger problems are refused.
#include "common.cuh"

namespace woft {

// 32 MiB: enough for one 1280x720 frame pair (90x160 feature maps, 256 channels) in the single-pass tensor-core
// modes (tf32: 2 x 14.06 MiB of repacked operands) and for the fp32 CUDA-core mode at any size.
constexpr size_t kStaticScratchBytes = size_t(32) << 20;
constexpr int kMaxDevices = 64;

static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

// In host code `g_static_scratch` is only a handle: the device address has to be asked for, per device.
static void* g_scratch_addr[kMaxDevices];  // written once per device by the first call on that device

static int static_scratch_pointer(void** out) {
  int dev = -1;
  WOFT_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail(WOFT_ERR_UNSUPPORTED, "device ordinal %d out of range", dev);
  void* p = __atomic_load_n(&g_scratch_addr[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    // set-up, once per device, with that device current
    WOFT_CUDA_CHECK(cudaGetSymbolAddress(&p, g_static_scratch));
    if (p == nullptr) return fail(WOFT_ERR_CUDA, "cudaGetSymbolAddress returned a null scratch address");
    __atomic_store_n(&g_scratch_addr[dev], p, __ATOMIC_RELEASE);  // racing first calls store the same value
  }
  *out = p;
  return WOFT_OK;
}

}  // namespace woft

extern "C" int woft_corr_pyramid_static(const float* fmap1, const float* fmap2, int batch, int channels, int height,
                                        int width, int num_levels, float* const* levels, float* row_mean, int mode,
                                        void* stream) {
  using namespace woft;
  WOFT_REQUIRE(batch > 0 && channels > 0 && height > 0 && width > 0, "woft_corr_pyramid_static: non-positive extent");
  const size_t need = woft_corr_workspace_bytes(batch, channels, height, width, mode);
  if (need == 0 || need > kStaticScratchBytes)
    return fail(WOFT_ERR_WORKSPACE,
                "woft_corr_pyramid_static: this shape and mode need %zu bytes of scratch but the built-in scratch "
                "holds %zu; use woft_corr_pyramid with a caller-allocated workspace",
                need, kStaticScratchBytes);
  void* scratch = nullptr;
  const int rc = static_scratch_pointer(&scratch);
  if (rc != WOFT_OK) return rc;
  // Same code path as the caller-allocated variant from here on.  It clears the status word at the head of the
  // scratch and overwrites every other byte it later reads, so nothing left behind by the previous call is seen.
  return woft_corr_pyramid(fmap1, fmap2, batch, channels, height, width, num_levels, levels, row_mean, scratch,
                           kStaticScratchBytes, mode, stream);
}

// The tensor-core kernel bounds its pipeline waits and raises a status word at the head of its scratch instead of
// hanging; woft_corr_check reads that word back but needs the scratch pointer, which a caller of the _static variant
// never sees.  This is the same check on the built-in scratch of the current device.
extern "C" int woft_corr_check_static(void* stream) {
  using namespace woft;
  void* scratch = nullptr;
  const int rc = static_scratch_pointer(&scratch);
  if (rc != WOFT_OK) return rc;
  return woft_corr_check(scratch, stream);
}
