// Stand-alone microbenchmark (not part of libacx.so): cycles per kind::tf32 128x128x8 tcgen05.mma for the operand
// sources the kernels use.  One CTA; thread 0 issues `iters` groups of 6 MMAs (one 16-row chunk, 3xTF32) and waits.
//   mode 0: A in tensor memory, B un-swizzled K-major core matrices (forward / chain kernels)
//   mode 1: A and B in shared memory, 128-byte swizzle (weight-gradient kernel)
//   mode 2: A in tensor memory, B 128-byte swizzled
#include "acx_umma.cuh"
using namespace acxumma;

__global__ void __launch_bounds__(128) mma_rate_test(int mode, int iters, int n_out, long long* out) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* base = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  float* sA = reinterpret_cast<float*>(base);            // 128 x 32 (16 KB)
  float* sB = sA + 128 * 32;                             // 128 x 32
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 2 * 128 * 32; i += 128) sA[i] = 1.0f;
  if (warp == 0) tmem_alloc(&tmem_base, 512);
  if (tid == 0) { acx_mbar_init(&bar, 1); acx_mbar_fence_init(); }
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t t = tmem_base;
  if (tid == 0) {
    const uint32_t idesc = idesc_tf32(128, n_out, false, false);
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int ks = 0; ks < 2; ++ks) {
        if (mode == 0) {
          const uint64_t bh = smem_desc(sB + ks * 64, 128, 512), bl = smem_desc(sB + 2048 + ks * 64, 128, 512);
          mma_tf32_ts(t + 256, t + 128 + ks * 8, bh, idesc, true);
          mma_tf32_ts(t + 256, t + ks * 8, bl, idesc, true);
          mma_tf32_ts(t + 256, t + ks * 8, bh, idesc, true);
        } else if (mode == 1) {
          const uint64_t ah = smem_desc_sw128(sA + ks * 8), bh = smem_desc_sw128(sB + ks * 8);
          mma_tf32(t + 256, ah, bh, idesc, true);
          mma_tf32(t + 256, ah, bh, idesc, true);
          mma_tf32(t + 256, ah, bh, idesc, true);
        } else {
          const uint64_t bh = smem_desc_sw128(sB + ks * 8);
          mma_tf32_ts(t + 256, t + 128 + ks * 8, bh, idesc, true);
          mma_tf32_ts(t + 256, t + ks * 8, bh, idesc, true);
          mma_tf32_ts(t + 256, t + ks * 8, bh, idesc, true);
        }
      }
    }
    const long long t1 = clock64();
    commit(&bar);
    acx_mbar_wait(&bar, 0);
    const long long t2 = clock64();
    out[0] = t1 - t0;
    out[1] = t2 - t0;
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(t, 512);
}

extern "C" int acx_mma_rate_test(int mode, int iters, int n_out, long long* out_dev, int ctas) {
  cudaFuncSetAttribute(mma_rate_test, cudaFuncAttributeMaxDynamicSharedMemorySize, 40 * 1024);
  mma_rate_test<<<ctas, 128, 34 * 1024>>>(mode, iters, n_out, out_dev);
  return (int)cudaDeviceSynchronize();
}
void acx_set_error(const char*, ...) {}
