// tcgen05 (5th-gen tensor core) helpers for sm_100a: TMEM allocation, shared-memory matrix
// descriptors for the un-swizzled canonical layouts, kind::tf32 MMA issue, commit, TMEM loads.
//
// Canonical un-swizzled operand layouts (bit layout per CUTLASS cute/arch/mma_sm100_desc.hpp and
// cute/atom/mma_traits_sm100.hpp::make_umma_desc; 4-byte elements, T = 4 elements per 16 bytes):
//   "core matrix" = 8 rows x 16 bytes, stored as 128 contiguous bytes.
//   MN-major (the MN index is contiguous):  elem(mn, k) at  (mn/4)*SBO + (k/8)*LBO + (k%8)*16 + (mn%4)*4
//   K-major  (the K  index is contiguous):  elem(mn, k) at  (mn/8)*SBO + (k/4)*LBO + (mn%8)*16 + (k%4)*4
// One kind::tf32 instruction consumes K = 8 (32 bytes).  The same buffer [a/8][b/4][8][4] is an MN-major
// operand over (mn = b, k = a) and a K-major operand over (mn = a, k = b) with SBO and LBO exchanged.
#pragma once
#include "acx_common.cuh"

namespace acxumma {

__device__ inline void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {   // one full warp; ncols power of 2 >= 32
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(acx_smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ inline void tmem_dealloc(uint32_t taddr, uint32_t ncols) {      // the same warp that allocated
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ inline void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ inline void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// generic-proxy shared-memory writes -> visible to the async proxy (tcgen05.mma operand reads)
__device__ inline void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// 64-bit shared-memory matrix descriptor, no swizzle (layout_type 0), Blackwell version field = 1
__device__ inline uint64_t smem_desc(const void* ptr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  const uint64_t addr = (uint64_t)((acx_smem_u32(ptr) & 0x3FFFFu) >> 4);
  return addr | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) | ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) |
         (1ull << 46);
}

// 128-byte-swizzled K-major operand (what a TMA tensor load with CU_TENSOR_MAP_SWIZZLE_128B produces from a box of
// 32 fp32 = 128 bytes per row): rows of 128 bytes, 8-row groups of 1024 bytes (tile base 1024-byte aligned), the
// 16-byte chunk c of row r stored at chunk c ^ (r % 8).  SBO = 1024 (next 8 rows), LBO unused, layout_type 2.
// A K step inside the row is taken by advancing the start address (+32 bytes per kind::tf32 instruction).
__device__ inline uint64_t smem_desc_sw128(const void* ptr) {
  const uint64_t addr = (uint64_t)((acx_smem_u32(ptr) & 0x3FFFFu) >> 4);
  return addr | ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
}

// 32-bit instruction descriptor: D = F32, A = B = TF32, dense, no negate
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N, bool a_mn_major, bool b_mn_major) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((a_mn_major ? 1u : 0u) << 15) | ((b_mn_major ? 1u : 0u) << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// D[tmem] (+)= A[smem] * B[smem]^T ; one thread issues
__device__ inline void mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, bool accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"((uint32_t)accumulate)
      : "memory");
}
// same, with the A operand read from tensor memory (lane = row of A, 8 consecutive 32-bit columns = one K step)
__device__ inline void mma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, bool accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"((uint32_t)accumulate)
      : "memory");
}
// 32 lanes x 32 consecutive 32-bit columns, registers -> tensor memory
__device__ inline void tmem_st32(uint32_t taddr, const float (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
        "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])),
        "r"(__float_as_uint(v[7])), "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])),
        "r"(__float_as_uint(v[11])), "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])),
        "r"(__float_as_uint(v[14])), "r"(__float_as_uint(v[15])), "r"(__float_as_uint(v[16])),
        "r"(__float_as_uint(v[17])), "r"(__float_as_uint(v[18])), "r"(__float_as_uint(v[19])),
        "r"(__float_as_uint(v[20])), "r"(__float_as_uint(v[21])), "r"(__float_as_uint(v[22])),
        "r"(__float_as_uint(v[23])), "r"(__float_as_uint(v[24])), "r"(__float_as_uint(v[25])),
        "r"(__float_as_uint(v[26])), "r"(__float_as_uint(v[27])), "r"(__float_as_uint(v[28])),
        "r"(__float_as_uint(v[29])), "r"(__float_as_uint(v[30])), "r"(__float_as_uint(v[31]))
      : "memory");
}
__device__ inline void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// arrive on an mbarrier once every previously issued tcgen05.mma of this thread has completed
__device__ inline void commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(acx_smem_u32(bar))
               : "memory");
}

// 32 lanes x 32 consecutive fp32 columns: thread `lane` of warp w reads row 32*(w%4)+lane
__device__ inline void tmem_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// fp32 -> (hi, lo) with hi, lo exactly representable in TF32 (10 explicit mantissa bits): x ~= hi + lo to ~2^-22 |x|
__device__ inline float round_tf32(float x) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
  return __uint_as_float(u);
}
// what kind::tf32 sees of a raw fp32 word: the low 13 mantissa bits are ignored
__device__ inline float trunc_tf32(float x) { return __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }
// lo part when the raw word itself serves as hi: x - trunc(x) is exact; adding half an ulp of the kept 10 bits to its
// bit pattern makes the MMA's own truncation a round-to-nearest (ties away), i.e. what cvt.rna.tf32 would keep
__device__ inline float lo_trunc_tf32(float x) {
  return __uint_as_float(__float_as_uint(x - trunc_tf32(x)) + 0x1000u);
}
__device__ inline void split_tf32(float x, float& hi, float& lo) {
  hi = round_tf32(x);
  lo = round_tf32(x - hi);      // x - hi is exact in fp32
}

}  // namespace acxumma
