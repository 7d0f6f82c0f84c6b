// Stand-alone check of the TMA-tensor / 128-byte-swizzle plumbing used by the weight-gradient kernel
// (not part of libacx.so):  D[128][N] = A[128][K] * B[N][K]^T, A and B row-major fp32 in global memory
// (K contiguous = K-major), loaded as [rows x 32] boxes with CU_TENSOR_MAP_SWIZZLE_128B and consumed through
// SWIZZLE_128B shared-memory descriptors.
//   mode 0: one kind::tf32 pass on the raw fp32 words (exact for small integers)
//   mode 1: 3xTF32 with hi = the raw word (the MMA ignores the low 13 mantissa bits), lo = rna(x - trunc(x))
//   mode 2: 3xTF32 with hi = rna(x) written back, lo = rna(x - hi)
#include <cuda.h>
#include "acx_umma.cuh"
#include "acx_tmap.cuh"
using namespace acxumma;

__global__ void __launch_bounds__(128) tma_gemm_test(const __grid_constant__ CUtensorMap mapA,
                                                     const __grid_constant__ CUtensorMap mapB, float* __restrict__ D,
                                                     int K, int N, int mode) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* base = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int nk = K / 32;                            // 32-column (128-byte) sub-tiles
  float* sA = reinterpret_cast<float*>(base);       // nk x [128 rows][32]
  float* sB = sA + (size_t)nk * 128 * 32;           // nk x [N rows][32]   (N = 128 keeps every sub-tile 1024-aligned)
  float* sAlo = sB + (size_t)nk * 128 * 32;
  float* sBlo = sAlo + (size_t)nk * 128 * 32;
  __shared__ __align__(8) uint64_t bar_full, bar_done;
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (warp == 0) tmem_alloc(&tmem_base, 128);
  if (tid == 0) {
    acx_mbar_init(&bar_full, 1);
    acx_mbar_init(&bar_done, 1);
    acx_mbar_fence_init();
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t taddr = tmem_base;
  if (tid == 0) {
    acx_mbar_expect_tx(&bar_full, (uint32_t)(nk * (128 + N) * 32 * 4));
    for (int kk = 0; kk < nk; ++kk) {
      tma_load_2d(sA + (size_t)kk * 128 * 32, &mapA, kk * 32, 0, &bar_full);
      tma_load_2d(sB + (size_t)kk * 128 * 32, &mapB, kk * 32, 0, &bar_full);
    }
  }
  acx_mbar_wait(&bar_full, 0);
  if (mode >= 1) {
    for (int i = tid; i < nk * 128 * 32; i += 128) {
      float x = sA[i], y = sB[i];
      if (mode == 1) {
        sAlo[i] = round_tf32(x - trunc_tf32(x));
        sBlo[i] = round_tf32(y - trunc_tf32(y));
      } else {
        float h, l;
        split_tf32(x, h, l); sA[i] = h; sAlo[i] = l;
        split_tf32(y, h, l); sB[i] = h; sBlo[i] = l;
      }
    }
    fence_proxy_async();
  }
  __syncthreads();
  if (tid == 0) {
    fence_after_sync();
    const uint32_t idesc = idesc_tf32(128, N, false, false);
    bool acc = false;
    for (int kk = 0; kk < nk; ++kk)
      for (int ks = 0; ks < 4; ++ks) {
        const size_t off = (size_t)kk * 128 * 32 + ks * 8;
        const uint64_t ad = smem_desc_sw128(sA + off), bd = smem_desc_sw128(sB + off);
        mma_tf32(taddr, ad, bd, idesc, acc);
        acc = true;
        if (mode >= 1) {
          mma_tf32(taddr, smem_desc_sw128(sAlo + off), bd, idesc, true);
          mma_tf32(taddr, ad, smem_desc_sw128(sBlo + off), idesc, true);
        }
      }
    commit(&bar_done);
  }
  acx_mbar_wait(&bar_done, 0);
  fence_after_sync();
  const int row = warp * 32 + (tid & 31);
  for (int c0 = 0; c0 < N; c0 += 32) {
    float v[32];
    tmem_ld32(taddr + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
#pragma unroll
    for (int i = 0; i < 32; ++i) D[(size_t)row * N + c0 + i] = v[i];
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(taddr, 128);
}

extern "C" int acx_tma_gemm_test(const float* A, const float* B, float* D, int K, int N, int mode, void* stream) {
  CUtensorMap mA, mB;
  if (acx_tmap_2d(&mA, A, (uint64_t)K, 128, (uint64_t)K * 4, 32, 128)) return -2;
  if (acx_tmap_2d(&mB, B, (uint64_t)K, (uint64_t)N, (uint64_t)K * 4, 32, (uint32_t)N)) return -3;
  const size_t bytes = (size_t)(K / 32) * 128 * 32 * 4 * 4 + 2048;
  cudaFuncSetAttribute(tma_gemm_test, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  tma_gemm_test<<<1, 128, bytes, (cudaStream_t)stream>>>(mA, mB, D, K, N, mode);
  return (int)cudaGetLastError();
}
void acx_set_error(const char*, ...) {}
