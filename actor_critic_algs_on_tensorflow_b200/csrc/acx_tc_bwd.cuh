// Pieces shared by the tensor-core backward kernels (acx_mlp_tc_train.cu: one-kernel form, acx_mlp_tc_chain.cu:
// chain-only form): tensor-memory map, the issuer's weight-chunk ring, the hi/lo A-operand store.
//   TMEM columns: A operand hi 0..127, lo 128..255 (lane = sample, column = reduction index), accumulators D0 256..383,
//   D1 384..511.  Packed backward weights: acx_tc_pack_weights_bwd (16 KB chunks = 16 reduction rows x 128 outputs,
//   hi plane then lo plane, K-major core matrices).
#pragma once
#include "acx_mlp.cuh"
#include "acx_umma.cuh"

namespace acxtcbwd {
using namespace acxmlp;
using namespace acxumma;

constexpr int TT_WORK = 256;
constexpr int TT_NT = TT_WORK + 64;      // + issuer warp (8) + producer warp (9)
constexpr int TT_SLOT_FLOATS = 4096;
constexpr int TT_NSLOT = 4;
constexpr uint32_t TT_A_HI = 0, TT_A_LO = 128, TT_D0 = 256, TT_D1 = 384;

__device__ inline void work_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
__device__ inline void mbar_arrive1(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(acx_smem_u32(bar)) : "memory");
}

// Weight-chunk ring.  Warp 9 (one lane) produces: bulk-TMA of every chunk of the CTA, in order, as slots free up.
// Warp 8 (one lane) consumes: MMAs on the landed chunk, slot released by tcgen05.commit.  Two warps, because a thread
// parked in mbarrier.try_wait parks its whole warp: a producer waiting for a free slot in the issuer's own warp
// leaves the tensor pipe idle between chunks (measured: 8.1k instead of 3.1k cycles for 48 MMAs).
template <class SM> struct TtPipe {      // SM: slot[TT_NSLOT][TT_SLOT_FLOATS], full[], empty[]
  SM* sm;
  const float* packed;
  unsigned int g, total;
  int chunks_per_tile;
  __device__ inline void produce() {
    for (unsigned int c = 0; c < total; ++c) {
      const unsigned int q = c % TT_NSLOT;
      if (c >= TT_NSLOT) acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm->empty[q]), ((c / TT_NSLOT) - 1u) & 1u);
      const float* src = packed + (size_t)(c % (unsigned)chunks_per_tile) * TT_SLOT_FLOATS;
      acx_mbar_expect_tx(reinterpret_cast<uint64_t*>(&sm->full[q]), TT_SLOT_FLOATS * 4);
      acx_bulk_g2s(sm->slot[q], src, TT_SLOT_FLOATS * 4, reinterpret_cast<uint64_t*>(&sm->full[q]));
    }
  }
  // 16 reduction rows (A columns k0..k0+15) x 128 outputs, accumulate into tmem_d
  __device__ inline void mma_chunk(uint32_t tmem, int k0, uint32_t tmem_d, bool first) {
    const unsigned int q = g % TT_NSLOT;
    acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm->full[q]), (g / TT_NSLOT) & 1u);
    fence_after_sync();
    const float* b_hi = sm->slot[q];
    const float* b_lo = b_hi + 128 * 16;
    const uint32_t idesc = idesc_tf32(128, 128, false, false);
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
      const uint64_t bh = smem_desc(b_hi + ks * 64, 128, 512), bl = smem_desc(b_lo + ks * 64, 128, 512);
      const uint32_t ah = tmem + TT_A_HI + (uint32_t)(k0 + ks * 8), al = tmem + TT_A_LO + (uint32_t)(k0 + ks * 8);
      mma_tf32_ts(tmem_d, al, bh, idesc, !(first && ks == 0));
      mma_tf32_ts(tmem_d, ah, bl, idesc, true);
      mma_tf32_ts(tmem_d, ah, bh, idesc, true);
    }
    commit(reinterpret_cast<uint64_t*>(&sm->empty[q]));
    ++g;
  }
};

__device__ inline void a_store32_tt(uint32_t tmem_lane, int j0, const float (&v)[32]) {
  // hi = the raw word (kind::tf32 ignores its low 13 mantissa bits), lo = the rounded remainder -- the forward kernels'
  // split: an AND, a subtract and an integer add per value instead of two cvt.rna.tf32 (conversion-pipe instructions,
  // 4 % of the chain kernels' warp samples) and a subtract
  float lo[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) lo[i] = lo_trunc_tf32(v[i]);
  tmem_st32(tmem_lane + TT_A_HI + (uint32_t)j0, v);
  tmem_st32(tmem_lane + TT_A_LO + (uint32_t)j0, lo);
}

}  // namespace acxtcbwd
