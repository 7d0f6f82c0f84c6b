// Shared device/host helpers for libacx (sm_100a only).
//
// Nothing here allocates; every buffer is owned by the caller.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/acx.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libacx is written for sm_100a (B200) only"
#endif

// ------------------------------------------------------------------ host-side error plumbing
void acx_set_error(const char* fmt, ...);

#define ACX_REQUIRE(cond, ...)                 \
  do {                                         \
    if (!(cond)) {                             \
      acx_set_error(__VA_ARGS__);              \
      return ACX_ERR_INVALID;                  \
    }                                          \
  } while (0)

#define ACX_CUDA_OK(expr)                                                              \
  do {                                                                                 \
    cudaError_t e__ = (expr);                                                          \
    if (e__ != cudaSuccess) {                                                          \
      acx_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, \
                    __LINE__);                                                         \
      return ACX_ERR_CUDA;                                                             \
    }                                                                                  \
  } while (0)

#define ACX_LAUNCH_OK(name)                                                     \
  do {                                                                          \
    cudaError_t e__ = cudaGetLastError();                                       \
    if (e__ != cudaSuccess) {                                                   \
      acx_set_error("launch of %s failed: %s", name, cudaGetErrorString(e__)); \
      return ACX_ERR_CUDA;                                                      \
    }                                                                           \
  } while (0)

static inline bool acx_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
int acx_num_sms();   // SM count of the current device (cached per device)

// ------------------------------------------------------------------ constants
#define ACX_A_H1 128
#define ACX_A_H2 128
#define ACX_A_H3 64
#define ACX_C_H1 256
#define ACX_C_H2 128
#define ACX_C_H3 128
#define ACX_LOG_STD_MIN (-2.5f)
#define ACX_LOG_STD_MAX (2.5f)
#define ACX_HALF_LOG_2PI 0.9189385332046727f
#define ACX_STREAM_RESET 1u
#define ACX_STREAM_ACTION 2u

__host__ __device__ inline int acx_al4(int x) { return (x + 3) & ~3; }

// flat parameter layout (floats); identical to oracle/policies.py::{actor,critic}_layout
struct AcxNetLayout {
  int W1, b1, W2, b2, W3, b3, Wh, bh, log_std, size;
  int H1, H2, H3, ob, nout;
};
__host__ __device__ inline AcxNetLayout acx_actor_layout(int ob, int num_ac, int cont) {
  AcxNetLayout L;
  L.H1 = ACX_A_H1; L.H2 = ACX_A_H2; L.H3 = ACX_A_H3; L.ob = ob; L.nout = num_ac;
  int o = 0;
  L.W1 = o; o = acx_al4(o + ob * L.H1);
  L.b1 = o; o = acx_al4(o + L.H1);
  L.W2 = o; o = acx_al4(o + L.H1 * L.H2);
  L.b2 = o; o = acx_al4(o + L.H2);
  L.W3 = o; o = acx_al4(o + L.H2 * L.H3);
  L.b3 = o; o = acx_al4(o + L.H3);
  L.Wh = o; o = acx_al4(o + L.H3 * num_ac);
  L.bh = o; o = acx_al4(o + num_ac);
  L.log_std = -1;
  if (cont) { L.log_std = o; o = acx_al4(o + num_ac); }
  L.size = o;
  return L;
}
__host__ __device__ inline AcxNetLayout acx_critic_layout(int ob) {
  AcxNetLayout L;
  L.H1 = ACX_C_H1; L.H2 = ACX_C_H2; L.H3 = ACX_C_H3; L.ob = ob; L.nout = 1;
  int o = 0;
  L.W1 = o; o = acx_al4(o + ob * L.H1);
  L.b1 = o; o = acx_al4(o + L.H1);
  L.W2 = o; o = acx_al4(o + L.H1 * L.H2);
  L.b2 = o; o = acx_al4(o + L.H2);
  L.W3 = o; o = acx_al4(o + L.H2 * L.H3);
  L.b3 = o; o = acx_al4(o + L.H3);
  L.Wh = o; o = acx_al4(o + L.H3);
  L.bh = o; o = acx_al4(o + 1);
  L.log_std = -1;
  L.size = o;
  return L;
}

// ------------------------------------------------------------------ Philox4x32-10 (matches oracle/philox.py)
struct AcxRng4 { uint32_t x, y, z, w; };

__host__ __device__ inline AcxRng4 acx_philox(uint64_t index, uint32_t step, uint32_t stream, uint64_t seed) {
  uint32_t c0 = (uint32_t)index, c1 = (uint32_t)(index >> 32), c2 = step, c3 = stream;
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  AcxRng4 o; o.x = c0; o.y = c1; o.z = c2; o.w = c3;
  return o;
}
__host__ __device__ inline float acx_u01(uint32_t x) { return (float)(x >> 8) * 5.9604644775390625e-08f; }

#ifdef __CUDACC__
// standard normal from two words, Box-Muller cosine branch (oracle/philox.py::normals)
__device__ inline float acx_normal(uint32_t a, uint32_t b) {
  float u1 = __fmul_rn(__fadd_rn((float)(a >> 8), 1.0f), 5.9604644775390625e-08f);
  float u2 = acx_u01(b);
  float r = sqrtf(__fmul_rn(-2.0f, logf(u1)));
  return __fmul_rn(r, cosf(__fmul_rn(6.283185307179586f, u2)));
}

// ------------------------------------------------------------------ mbarrier + bulk-TMA (cp.async.bulk, SASS UBLKCP)
__device__ inline uint32_t acx_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ inline void acx_mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(acx_smem_u32(bar)), "r"(count) : "memory");
}
__device__ inline void acx_mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ inline void acx_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(acx_smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ inline bool acx_mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(acx_smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// bounded wait: a protocol bug traps (launch error) instead of hanging the GPU
__device__ inline void acx_mbar_wait(uint64_t* bar, uint32_t parity) {
  for (uint32_t spin = 0; spin < (1u << 24); ++spin)
    if (acx_mbar_try_wait(bar, parity)) return;
  __trap();
}
// 1-D bulk copy global -> shared, completion counted in bytes on an mbarrier (TMA engine)
__device__ inline void acx_bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   acx_smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(acx_smem_u32(bar))
               : "memory");
}

__device__ inline float acx_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ inline double acx_warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
#endif  // __CUDACC__
