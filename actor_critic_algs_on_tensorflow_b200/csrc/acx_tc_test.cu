// Stand-alone check of the tcgen05 plumbing in acx_umma.cuh (not part of libacx.so):
// D[128][N] = A[128][K] * B[N][K]^T with operands packed on the host in the canonical layouts.
//   mode 0: A MN-major, B MN-major      mode 1: A K-major, B K-major
//   mode 2: like 0, but fp32 operands split on the device into TF32 hi/lo, 3 MMAs per K block (3xTF32)
//   mode 4: A operand in tensor memory (row-major A[128][K] from global, written with tcgen05.st), B K-major
//   mode 5: like 4 with the 3xTF32 split
#include "acx_umma.cuh"
using namespace acxumma;

__global__ void __launch_bounds__(128) tc_gemm_test(const float* __restrict__ Ap, const float* __restrict__ Bp,
                                                    float* __restrict__ D, int K, int N, int mode) {
  extern __shared__ __align__(1024) unsigned char smem[];
  float* sA = reinterpret_cast<float*>(smem);      // K x 128
  float* sB = sA + (size_t)K * 128;                // K x N
  float* sAlo = sB + (size_t)K * N;                // mode 2
  float* sBlo = sAlo + (size_t)K * 128;
  __shared__ __align__(8) uint64_t mbar;
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < K * 128; i += 128) {
    float x = Ap[i];
    if (mode == 2) { float h, l; split_tf32(x, h, l); sA[i] = h; sAlo[i] = l; } else sA[i] = x;
  }
  for (int i = tid; i < K * N; i += 128) {
    float x = Bp[i];
    if (mode == 2 || mode == 5) { float h, l; split_tf32(x, h, l); sB[i] = h; sBlo[i] = l; } else sB[i] = x;
  }
  if (warp == 0) tmem_alloc(&tmem_base, 512);
  if (tid == 0) { acx_mbar_init(&mbar, 1); acx_mbar_fence_init(); }
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t taddr = tmem_base;
  if (mode >= 4) {
    // A (row-major [128][K] in global for these modes) -> TMEM columns 128.. (hi) and 128+K.. (lo): thread = row
    const int arow = warp * 32 + (tid & 31);
    for (int c0 = 0; c0 < K; c0 += 32) {
      float hi[32], lo[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const float x = Ap[(size_t)arow * K + c0 + i];
        if (mode == 5) split_tf32(x, hi[i], lo[i]); else { hi[i] = x; lo[i] = 0.f; }
      }
      tmem_st32(taddr + ((uint32_t)(warp * 32) << 16) + 128u + (uint32_t)c0, hi);
      tmem_st32(taddr + ((uint32_t)(warp * 32) << 16) + 128u + (uint32_t)(K + c0), lo);
    }
    tmem_wait_st();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    if (tid == 0) {
      const uint32_t idesc = idesc_tf32(128, N, false, false);
      for (int kb = 0; kb < K / 8; ++kb) {
        const uint64_t bd = smem_desc(sB + (size_t)kb * 2 * 32, 128, (K / 4) * 128);
        const uint64_t bdl = smem_desc(sBlo + (size_t)kb * 2 * 32, 128, (K / 4) * 128);
        mma_tf32_ts(taddr, taddr + 128u + (uint32_t)(kb * 8), bd, idesc, kb > 0);
        if (mode == 5) {
          mma_tf32_ts(taddr, taddr + 128u + (uint32_t)(kb * 8), bdl, idesc, true);
          mma_tf32_ts(taddr, taddr + 128u + (uint32_t)(K + kb * 8), bd, idesc, true);
        }
      }
      commit(&mbar);
    }
  } else if (tid == 0) {
    const bool mn = (mode != 1);
    const bool swap = (mode == 3);
    const uint32_t idesc = idesc_tf32(128, N, mn, mn);
    // [K/8][MN/4][8][4] (MN-major) : SBO = 128 (next 4 of MN), LBO = (MN/4)*128 (next 8 of K)
    // [MN/8][K/4][8][4] (K-major)  : SBO = (K/4)*128 (next 8 of MN), LBO = 128 (next 4 of K); one MMA = 2 chunks
    for (int kb = 0; kb < K / 8; ++kb) {
      uint64_t ad, bd, adl = 0, bdl = 0;
      if (mn) {
        const uint32_t la = swap ? 128 : 32 * 128, sa = swap ? 32 * 128 : 128;
        const uint32_t lb = swap ? 128 : (N / 4) * 128, sb = swap ? (N / 4) * 128 : 128;
        ad = smem_desc(sA + (size_t)kb * 32 * 32, la, sa);
        bd = smem_desc(sB + (size_t)kb * (N / 4) * 32, lb, sb);
        if (mode == 2) {
          adl = smem_desc(sAlo + (size_t)kb * 32 * 32, la, sa);
          bdl = smem_desc(sBlo + (size_t)kb * (N / 4) * 32, lb, sb);
        }
      } else {
        ad = smem_desc(sA + (size_t)kb * 2 * 32, 128, (K / 4) * 128);
        bd = smem_desc(sB + (size_t)kb * 2 * 32, 128, (K / 4) * 128);
      }
      mma_tf32(taddr, ad, bd, idesc, kb > 0);
      if (mode == 2) {
        mma_tf32(taddr, ad, bdl, idesc, true);
        mma_tf32(taddr, adl, bd, idesc, true);
      }
    }
    commit(&mbar);
  }
  acx_mbar_wait(&mbar, 0);
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
  if (warp == 0) tmem_dealloc(taddr, 512);
}

extern "C" int acx_tc_gemm_test(const float* Ap, const float* Bp, float* D, int K, int N, int mode, void* stream) {
  const size_t bytes = (size_t)(K * 128 + K * N) * 4 * (mode == 2 || mode == 5 ? 2 : 1) + 1024;
  cudaFuncSetAttribute(tc_gemm_test, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  tc_gemm_test<<<1, 128, bytes, (cudaStream_t)stream>>>(Ap, Bp, D, K, N, mode);
  return (int)cudaGetLastError();
}
void acx_set_error(const char*, ...) {}
