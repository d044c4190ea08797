// Latency of the per-k-step synchronisation instructions on one elected thread (scratch probe).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ bool try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ bool test_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void arrive(uint64_t* bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory"); }
__device__ __forceinline__ void arrive_tx(uint64_t* bar, uint32_t tx) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(tx) : "memory"); }
__device__ __forceinline__ void commit(uint64_t* bar) { asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__global__ void probe(long long* out) {
  __shared__ uint64_t bars[16];
  if (threadIdx.x == 0) {
    for (int i = 0; i < 16; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bars + i)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int N = 64;
  if (threadIdx.x < 32 && elect_one()) {
    long long t0, t1;
    // (0) try_wait on a phase that completed long ago (parity 1 of a fresh barrier)
    t0 = clock64();
    for (int i = 0; i < N; ++i) try_wait(bars + (i & 7), 1);
    t1 = clock64();
    out[0] = (t1 - t0) / N;
    // (1) test_wait likewise
    t0 = clock64();
    for (int i = 0; i < N; ++i) test_wait(bars + (i & 7), 1);
    t1 = clock64();
    out[1] = (t1 - t0) / N;
    // (2) arrive then try_wait on the same barrier (phase flips each time)
    uint32_t ph = 0;
    t0 = clock64();
    for (int i = 0; i < N; ++i) { arrive(bars + 8); while (!try_wait(bars + 8, ph)) {} ph ^= 1; }
    t1 = clock64();
    out[2] = (t1 - t0) / N;
    // (3) arrive only (8 barriers round robin)
    t0 = clock64();
    for (int i = 0; i < N; ++i) arrive(bars + (i & 7));
    t1 = clock64();
    out[3] = (t1 - t0) / N;
    // (4) arrive.expect_tx(0)
    t0 = clock64();
    for (int i = 0; i < N; ++i) arrive_tx(bars + (i & 7), 0);
    t1 = clock64();
    out[4] = (t1 - t0) / N;
    // (5) tcgen05.commit then try_wait on it
    ph = 0;
    t0 = clock64();
    for (int i = 0; i < N; ++i) { commit(bars + 9); while (!try_wait(bars + 9, ph)) {} ph ^= 1; }
    t1 = clock64();
    out[5] = (t1 - t0) / N;
    // (6) tcgen05.commit only, round robin
    t0 = clock64();
    for (int i = 0; i < N; ++i) commit(bars + (i & 7));
    t1 = clock64();
    out[6] = (t1 - t0) / N;
    // (7) fence::after_thread_sync
    t0 = clock64();
    for (int i = 0; i < N; ++i) fence_after();
    t1 = clock64();
    out[7] = (t1 - t0) / N;
    // (8) try_wait (completed) + fence + commit: the skeleton k-step of the MMA thread
    for (int i = 0; i < 8; ++i) while (!try_wait(bars + i, (N / 8 * 3) & 1 ? 0 : 0)) { break; }
    t0 = clock64();
    for (int i = 0; i < N; ++i) { try_wait(bars + 10, 1); fence_after(); commit(bars + 11 + (i & 3)); }
    t1 = clock64();
    out[8] = (t1 - t0) / N;
  }
}
int main() {
  long long* d;
  cudaMalloc(&d, 16 * 8);
  cudaMemset(d, 0, 16 * 8);
  for (int r = 0; r < 2; ++r) probe<<<1, 128>>>(d);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[16];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  const char* names[] = {"try_wait, phase complete", "test_wait, phase complete", "arrive + try_wait (same thread)", "arrive", "arrive.expect_tx(0)",
                         "tcgen05.commit + try_wait", "tcgen05.commit", "tcgen05.fence::after_thread_sync", "try_wait + fence + commit"};
  printf("status %s\n", cudaGetErrorString(e));
  for (int i = 0; i < 9; ++i) printf("%-36s %lld cycles\n", names[i], h[i]);
  return 0;
}
