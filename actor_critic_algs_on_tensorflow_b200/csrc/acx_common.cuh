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
// bounded wait: a protocol bug traps (launch error) instead of hanging the GPU.  The bound is wall time
// (%globaltimer, ns), not a poll count: under ncu replay, compute-sanitizer or a throttled clock a legitimate
// wait can outlast any fixed number of polls, and a trap is sticky for the whole CUDA context.
#ifndef ACX_MBAR_TIMEOUT_NS
#define ACX_MBAR_TIMEOUT_NS 20000000000ull      // 20 s
#endif
__device__ inline uint64_t acx_globaltimer() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ inline void acx_mbar_wait(uint64_t* bar, uint32_t parity) {
  for (uint32_t spin = 0; spin < (1u << 16); ++spin)
    if (acx_mbar_try_wait(bar, parity)) return;
  const uint64_t t0 = acx_globaltimer();          // slow path only: the common case never reads the timer
  for (;;) {
    for (uint32_t spin = 0; spin < (1u << 12); ++spin)
      if (acx_mbar_try_wait(bar, parity)) return;
    if (acx_globaltimer() - t0 > ACX_MBAR_TIMEOUT_NS) __trap();
  }
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

// ------------------------------------------------------------------ grid-wide sums without atomics on the data
// Every block leaves its K partial sums in ws->partial[K * blockIdx.x ..]; the block that takes the last ticket folds
// them (thread j takes blocks j, j + blockDim, ... in index order, then a fixed butterfly and a fixed walk over the
// warps): the result depends on the grid size only, never on block timing, and the fold is a handful of dependent
// L2 reads long.  Returns true in thread 0 of that last block, with the totals in tot[]; the ticket is back at zero
// when it does.  blockDim.x: a multiple of 32, at most 256.  ws: 16-byte aligned, ticket zero before the first use.
struct AcxFoldWs {
  unsigned int ticket;
  unsigned int pad[3];
  double partial[1];         // [K * gridDim.x]
};
template <int K>
__device__ inline bool acx_fold_blocks(const double (&m)[K], AcxFoldWs* ws, double (&tot)[K]) {
  __shared__ double sh[K][8];
  __shared__ bool last;
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (int)(blockDim.x >> 5);
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const double r = acx_warp_sum_d(m[k]);
    if (l == 0) sh[k][w] = r;
  }
  __syncthreads();
  if (threadIdx.x < K) {
    double a = 0;
    for (int i = 0; i < nw; ++i) a += sh[threadIdx.x][i];
    ws->partial[(size_t)K * blockIdx.x + threadIdx.x] = a;
    __threadfence();
  }
  __syncthreads();
  if (threadIdx.x == 0) last = (atomicAdd(&ws->ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!last) return false;
  __threadfence();
  double a[K];
#pragma unroll
  for (int k = 0; k < K; ++k) a[k] = 0;
  for (unsigned i = threadIdx.x; i < gridDim.x; i += blockDim.x) {
#pragma unroll
    for (int k = 0; k < K; ++k) a[k] += ((volatile double*)ws->partial)[(size_t)K * i + k];
  }
  __syncthreads();           // sh[] is reused
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const double r = acx_warp_sum_d(a[k]);
    if (l == 0) sh[k][w] = r;
  }
  __syncthreads();
  if (threadIdx.x != 0) return false;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    tot[k] = 0;
    for (int i = 0; i < nw; ++i) tot[k] += sh[k][i];
  }
  ws->ticket = 0;
  return true;
}

// np.mean(((x - mean x) / std x) * ((y - mean y) / std y)) from the five raw sums over c samples -- var_accounted_for,
// A3C/util.py:5-13 (population std; nan for a constant input, like numpy's 0/0)
__device__ inline double acx_corr_from_sums(double sx, double sxx, double sy, double syy, double sxy, double c) {
  const double mx = sx / c, my = sy / c;
  double vx = sxx / c - mx * mx, vy = syy / c - my * my;
  if (vx < 0) vx = 0;
  if (vy < 0) vy = 0;
  return (sxy / c - mx * my) / (sqrt(vx) * sqrt(vy));
}
#endif  // __CUDACC__
