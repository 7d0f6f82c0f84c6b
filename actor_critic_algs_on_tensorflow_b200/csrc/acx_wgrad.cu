// Weight-gradient products of K4 on the tensor cores (tcgen05, 3xTF32), one launch per network and PHASE (template
// parameter PH: 0 = dW3 / db3, 1 = dW2 / db2, each with its own compile-time ring geometry; -1 = both in one launch):
//     dW3[k][j] = sum_s X2[k][s] * dPre3[j][s]          (A3C/policies.py:82,155: compute_gradients of the
//     dW2[k][j] = sum_s X1[k][s] * dPre2[j][s]           third / second dense layer's kernel and bias)
//     db3[j] = sum_s dPre3[j][s],   db2[j] = sum_s dPre2[j][s]
// summed over the whole [rows x n] batch.  Operands are the tile-blocked [tile][feature][128] arrays the chain kernel
// wrote (dPre3, dPre2) and the activation cache (X2): the sample index is contiguous, i.e. both operands are K-major
// as they lie in HBM.  X1 = act(obs W1 + b1) is not read but recomputed (ob_dim <= 8 FMAs per element) straight into the swizzled
// operand layout.  Each persistent CTA owns the 128-sample tiles  blockIdx.x, +gridDim.x, ...  and keeps
// its partial dW3 / dW2 in tensor memory (fp32).  The tensor core's accumulate TRUNCATES, once per instruction: a bias of
// about half an ulp of the accumulator per accumulating MMA, towards zero.  Where the per-sample terms cancel it hides
// (7.7e-5 of the result over a CTA's whole share of ~3,300 MMAs, 8.4e-6 with a flush every 8 tiles); where they have one
// sign it adds up linearly in the number of MMAs per accumulation -- measured against an fp64 product of the same
// operands on one-signed data (tests/test_gpu_a_fullsize.py, tests/tc/run_flush_ab.py): 2.2e-5 of the result at 8 tiles
// per accumulation, 1.1e-5 at 4, 5.4e-6 at 2.  So the accumulators are double-buffered and flushed every WG_FLUSH = 2
// tiles (96 accumulating MMAs): while the MMAs fill one buffer, warps 2-5 add the other into the CTA's scratch row with
// ordinary fp32 (round-to-nearest) adds; acx_grad_reduce then sums the rows in fixed order (bit-reproducible run to run).
// Cost against 8 tiles: +5 % of the two kernels' time (critic pair 2.11 -> 2.21 ms); it buys the 1e-5 parity bar.
//   warp 0     producer: TMA tensor loads of [features x 32 samples] boxes, 128-byte swizzle, 2- or 3-deep ring (WgCfg)
//   warp 1     issuer:   per box 4 K-steps x 3 kind::tf32 MMAs (lo*hi, hi*lo, hi*hi) per 128-row block
//   warps 2-9  splitters: lo = rna(x - trunc(x)) of every landed word into the stage's second buffer (the MMA itself
//              ignores the low 13 mantissa bits of the raw word, which therefore serves as hi); they also generate the
//              X1 boxes, keep the bias-gradient sums of the dPre rows they pass over, and warps 2-5 run the epilogue
//              (tcgen05.ld -> scratch row).
// HBM-bound: 4 bytes per streamed operand element, every element read once.
#include "acx_mlp.cuh"
#include "acx_wgrad.cuh"
#include "acx_tmap.cuh"
#include "acx_umma.cuh"

using namespace acxumma;

#ifdef ACX_WG_PROF
// dev build (make prof): issuer of CTA 0 -- [0] phase-0 cycles, [1] phase-1 cycles, [2]/[3] of which waiting for split
// boxes, [4]/[5] boxes per phase; splitter warp 6 (no flush duty), thread 0, per phase p: [8+4p] wait for the free slot,
// [9+4p] X1 generation, [10+4p] wait for the TMA box, [11+4p] split + arrive
__device__ unsigned long long acx_wg_prof[24];     // [16+p]: of "wait for the free slot", the part before the wait itself
#define WG_SPLIT_MARK(i) do { if (sprof_) { const long long n_ = clock64(); sp_[i] += n_ - sl_; sl_ = n_; } } while (0)
#else
#define WG_SPLIT_MARK(i) do { } while (0)
#endif

namespace {

constexpr int WG_THREADS = 320;
constexpr int BOX = 32;                    // samples per box row (128 bytes)
constexpr int BLK = 128 * BOX;             // floats of a 128-row box
#ifndef ACX_WG_FLUSH
#define ACX_WG_FLUSH 2
#endif
constexpr int WG_FLUSH = ACX_WG_FLUSH;                // tiles (256 samples, 96 accumulating MMAs) per tensor-memory accumulation

struct WgBars {
  unsigned long long full[3], split[3], empty[3], dfull[2], dfree[2];
  unsigned int tmem_base;
  float W1[acxmlp::MAX_OB * 256];
  float b1[256];
};

// PH = -1: one kernel runs both phases on one ring geometry; PH = 0 / 1: a kernel of its own per phase, each with the
// geometry that suits it (the critic: phase 0 moves two operands from HBM per box and is bound by the latency of those
// loads -- it gets three 2-block stages instead of two 3-block ones; the geometry is a compile-time constant either way).
template <int H1, int PH> struct WgCfg {
  static constexpr int NB = H1 / 128;
  static constexpr int ABLK = PH == 0 ? 1 : NB;             // 128-row A blocks of a stage
  static constexpr int STAGE = (ABLK + 1) * BLK;            // floats: A blocks, then the B box
  static constexpr int NSTAGE = ABLK == 1 ? 3 : 2;
  static constexpr size_t SMEM = (size_t)NSTAGE * STAGE * 2 * sizeof(float) + 1024 + sizeof(WgBars);
};

__device__ inline void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(acx_smem_u32(bar)) : "memory");
}

// float4 slot of (row r, 16-byte chunk c) inside a 128-row, 128-byte-swizzled box
__device__ inline int sw128_f4(int r, int c) { return (r >> 3) * 64 + (r & 7) * 8 + (c ^ (r & 7)); }

__device__ inline void red_add_v4(float4* p, const float4 a) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w) : "memory");
}

__device__ inline float4 lo_of(const float4 x) {
  return make_float4(lo_trunc_tf32(x.x), lo_trunc_tf32(x.y), lo_trunc_tf32(x.z), lo_trunc_tf32(x.w));
}

template <int H1, int H3, bool LRELU, int OBM, int PH>
__global__ void __launch_bounds__(WG_THREADS, 1)
acx_wgrad_kernel(const __grid_constant__ CUtensorMap mapX2, const __grid_constant__ CUtensorMap mapD3,
                 const __grid_constant__ CUtensorMap mapD2, const float* __restrict__ obs,
                 const float* __restrict__ params, int ob, int offW1, int offb1, int rows, int n,
                 float* __restrict__ scratch, int scr_stride, int offW3, int offW2, int offb3, int offb2) {
  using Cfg = WgCfg<H1, PH>;
  constexpr int NB = Cfg::NB, NSTAGE = Cfg::NSTAGE, ABLK = Cfg::ABLK;
  constexpr int PH0 = PH < 0 ? 0 : PH, PH1 = PH < 0 ? 2 : PH + 1;          // the phases this kernel runs
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  // 1024-byte alignment for the swizzled boxes as an offset added to the __shared__ array: a round trip through
  // uintptr_t loses the address space and every ring access becomes a generic LD / ST.  (In the two-phase kernel, which
  // spilled, the generic path measured FASTER; in the per-phase kernels LDS / STS win: step 5.64 -> 5.54 ms.)
  unsigned char* base = smem_raw + ((1024u - (acx_smem_u32(smem_raw) & 1023u)) & 1023u);
  float* raw = reinterpret_cast<float*>(base);                         // [NSTAGE][STAGE]
  float* lo = raw + (size_t)NSTAGE * Cfg::STAGE;                       // [NSTAGE][STAGE]
  WgBars& B = *reinterpret_cast<WgBars*>(lo + (size_t)NSTAGE * Cfg::STAGE);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tpr = (n + 127) / 128;
  const int ntiles = rows * tpr;

  if (warp == 1) tmem_alloc(&B.tmem_base, 512);
  for (int i = tid; i < ob * H1; i += WG_THREADS) B.W1[i] = params[offW1 + i];
  for (int i = tid; i < H1; i += WG_THREADS) B.b1[i] = params[offb1 + i];
  if (tid == 0) {
    for (int i = 0; i < NSTAGE; ++i) {
      acx_mbar_init(reinterpret_cast<uint64_t*>(&B.full[i]), 1);
      acx_mbar_init(reinterpret_cast<uint64_t*>(&B.split[i]), 256);
      acx_mbar_init(reinterpret_cast<uint64_t*>(&B.empty[i]), 1);
    }
    for (int i = 0; i < 2; ++i) {
      acx_mbar_init(reinterpret_cast<uint64_t*>(&B.dfull[i]), 1);
      acx_mbar_init(reinterpret_cast<uint64_t*>(&B.dfree[i]), 128);
    }
    acx_mbar_fence_init();
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = B.tmem_base;

  // every role walks the same sequence of boxes: phase 0 = layer 3 (A = X2, B = dPre3), phase 1 = layer 2 (A = the NB
  // 128-row blocks of X1, B = dPre2); per tile one box per started 32 samples
  if (warp == 0) {
    if (lane == 0) {
      unsigned int it = 0;
      for (int phase = PH0; phase < PH1; ++phase) {
        const int brows = phase == 0 ? H3 : 128;
        const CUtensorMap* mB = phase == 0 ? &mapD3 : &mapD2;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
          const int row = tile / tpr, n0 = (tile % tpr) * 128;
          const int nsub = (min(128, n - n0) + BOX - 1) / BOX;
          for (int kk = 0; kk < nsub; ++kk, ++it) {
            const unsigned int slot = it % NSTAGE;
            if (it >= (unsigned)NSTAGE)
              acx_mbar_wait(reinterpret_cast<uint64_t*>(&B.empty[slot]), ((it / NSTAGE) - 1u) & 1u);
            uint64_t* full = reinterpret_cast<uint64_t*>(&B.full[slot]);
            acx_mbar_expect_tx(full, (uint32_t)(((phase == 0 ? 128 : 0) + brows) * BOX * 4));
            float* dst = raw + (size_t)slot * Cfg::STAGE;
            (void)row;
            if (phase == 0) tma_load_2d(dst, &mapX2, kk * BOX, tile * 128, full);          // phase 1: X1 is generated
            tma_load_2d(dst + ABLK * BLK, mB, kk * BOX, tile * brows, full);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      unsigned int it = 0, group = 0, used[2] = {0u, 0u};
#ifdef ACX_WG_PROF
      long long pt_ = clock64(), pw_ = 0;
      unsigned int pit_ = 0;
#endif
      for (int phase = PH0; phase < PH1; ++phase) {
        // phase 1 with two X1 blocks (critic): the roles are swapped -- A = the dPre2 box (128 rows), B = both X1 blocks
        // (256 rows, contiguous under the 128-byte swizzle), ONE N = 256 instruction per split product; the accumulator
        // holds dW2 transposed (lane = j, column = k).  Per K step the tensor core then reads 3 x (4 + 8) KB of shared
        // memory instead of 6 x (4 + 4) (ncu: tensor-core reads 39 % + load/store wavefronts 37 % of the shared-memory
        // data pipe's peak in this kernel).
        const bool wide = phase == 1 && NB == 2;
        const int nblk = (phase == 0 || wide) ? 1 : NB;
        const uint32_t idesc = idesc_tf32(128, phase == 0 ? H3 : (wide ? 256 : 128), false, false);
        if (phase == 1)                              // phase-1 buffers overlap phase-0's: every flush must be through
          for (int b = 0; b < 2; ++b)
            if (used[b]) acx_mbar_wait(reinterpret_cast<uint64_t*>(&B.dfree[b]), (used[b] - 1u) & 1u);
        int in_group = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
          const int n0 = (tile % tpr) * 128;
          const int nsub = (min(128, n - n0) + BOX - 1) / BOX;
          const unsigned int buf = group & 1u;
          const uint32_t dbase = tmem + (phase == 0 ? buf * 128u : buf * 256u);
          if (in_group == 0 && used[buf]) {          // the buffer's previous contents have been added to the scratch row
            acx_mbar_wait(reinterpret_cast<uint64_t*>(&B.dfree[buf]), (used[buf] - 1u) & 1u);
            fence_after_sync();
          }
          for (int kk = 0; kk < nsub; ++kk, ++it) {
            const unsigned int slot = it % NSTAGE;
#ifdef ACX_WG_PROF
            const long long w0_ = clock64();
#endif
            acx_mbar_wait(reinterpret_cast<uint64_t*>(&B.split[slot]), (it / NSTAGE) & 1u);
#ifdef ACX_WG_PROF
            pw_ += clock64() - w0_;
#endif
            fence_after_sync();
            const float* r = raw + (size_t)slot * Cfg::STAGE;
            const float* l = lo + (size_t)slot * Cfg::STAGE;
            const bool fresh = in_group == 0 && kk == 0;
#pragma unroll
            for (int ks = 0; ks < BOX / 8; ++ks) {
              const uint64_t bh = smem_desc_sw128(r + ABLK * BLK + ks * 8), bl = smem_desc_sw128(l + ABLK * BLK + ks * 8);
              if (wide) {
                const uint64_t xh = smem_desc_sw128(r + ks * 8), xl = smem_desc_sw128(l + ks * 8);
                mma_tf32(dbase, bl, xh, idesc, !(fresh && ks == 0));
                mma_tf32(dbase, bh, xl, idesc, true);
                mma_tf32(dbase, bh, xh, idesc, true);
                continue;
              }
              for (int b = 0; b < nblk; ++b) {
                const uint64_t ah = smem_desc_sw128(r + b * BLK + ks * 8), al = smem_desc_sw128(l + b * BLK + ks * 8);
                const uint32_t d = dbase + 128u * (uint32_t)b;
                mma_tf32(d, al, bh, idesc, !(fresh && ks == 0));
                mma_tf32(d, ah, bl, idesc, true);
                mma_tf32(d, ah, bh, idesc, true);
              }
            }
            commit(reinterpret_cast<uint64_t*>(&B.empty[slot]));
#ifdef ACX_WG_PROF_DRAIN      // measurement only: block until this box's MMAs are through (serialises the ring)
            {
              const long long d0_ = clock64();
              acx_mbar_wait(reinterpret_cast<uint64_t*>(&B.empty[slot]), (it / NSTAGE) & 1u);
              if (blockIdx.x == 0) acx_wg_prof[6 + phase] += clock64() - d0_;
            }
#endif
          }
          if (++in_group == WG_FLUSH || tile + (int)gridDim.x >= ntiles) {       // group complete: hand the buffer over
            commit(reinterpret_cast<uint64_t*>(&B.dfull[buf]));
            ++used[buf];
            ++group;
            in_group = 0;
          }
        }
#ifdef ACX_WG_PROF
        if (blockIdx.x == 0) {
          const long long now_ = clock64();
          acx_wg_prof[phase] = now_ - pt_; acx_wg_prof[2 + phase] = pw_; acx_wg_prof[4 + phase] = it - pit_;
          pt_ = now_; pw_ = 0; pit_ = it;
        }
#endif
      }
    }
  } else {
    const int t = tid - 64;                       // 0..255
    const int c4 = t & 7, r0 = t >> 3;            // this thread's 16-byte chunk and its rows r0 + 32 m of every box
    unsigned int it = 0;
    // warps 2-5 also flush completed accumulation groups: tensor memory -> (+)= this CTA's scratch row
    unsigned int group = 0, fl_used[2] = {0u, 0u};
    int pend_phase = -1, pend_first = 0;
    unsigned int pend_buf = 0;
    float* scr_row = scratch + (size_t)blockIdx.x * scr_stride;
    // One 32-column slice per call (so that a flush is spread over the following boxes instead of stalling the ring);
    // the partials are kept lane-major -- element (r, c) of the region at ((c / 4) * 128 + r) * 4 + c % 4 -- so that a
    // warp's float4 accesses are 512 contiguous bytes.  Returns true when the group is through.
    int pend_c0 = 0;
    auto flush_slice = [&]() {
      if (pend_c0 == 0) {
        acx_mbar_wait(reinterpret_cast<uint64_t*>(&B.dfull[pend_buf]), fl_used[pend_buf] & 1u);
        ++fl_used[pend_buf];
        fence_after_sync();
      }
      const int r = (warp & 3) * 32 + lane;                   // TMEM lane = row of the A block
      const uint32_t tl = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (pend_phase == 0 ? pend_buf * 128u : pend_buf * 256u);
      const int ncol = pend_phase == 0 ? H3 : NB * 128;
      float v[32];
      tmem_ld32(tl + (uint32_t)pend_c0, v);
      float4* p = reinterpret_cast<float4*>(scr_row + (pend_phase == 0 ? offW3 : offW2)) + (pend_c0 >> 2) * 128 + r;
      // the group's first flush stores, the later ones ADD with vector reductions (REDG.ADD.F32x4, round-to-nearest): the
      // row belongs to this CTA and every address is touched by this one thread, in program order, so the sum is the same
      // ((p + a1) + a2) ... a load-add-store gives, run to run -- without the L2 round trip of the load on the splitters'
      // path (with a flush every 2 tiles that round trip cost 5 % of the two kernels)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float4 a = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        if (pend_first) p[i * 128] = a;
        else red_add_v4(p + i * 128, a);
      }
      pend_c0 += 32;
      if (pend_c0 < ncol) return false;
      fence_before_sync();
      mbar_arrive(&B.dfree[pend_buf]);
      pend_phase = -1;
      pend_c0 = 0;
      return true;
    };
    auto flush = [&]() { while (!flush_slice()) {} };
    // observations of this thread's 4 samples of box (tile, kk); zeros beyond the batch
    // (row, n0) of the tile come from the caller: one division per tile, not one per box
    auto load_obs = [&](bool tile_ok, int row, int n0, int kk, float4 (&o)[OBM]) {
      const int col = n0 + kk * BOX + 4 * c4;
#pragma unroll
      for (int i = 0; i < OBM; ++i)
        o[i] = (i < ob && tile_ok && col < n)
                   ? __ldg(reinterpret_cast<const float4*>(obs + ((size_t)row * ob + i) * n + col))
                   : make_float4(0.f, 0.f, 0.f, 0.f);
    };
    for (int phase = PH0; phase < PH1; ++phase) {
      const int brows = phase == 0 ? H3 : 128;
      float bacc[4] = {0.f, 0.f, 0.f, 0.f};       // bias-gradient partials of rows r0 + 32 m (chunk c4 of every box)
      float4 o[OBM], o_next[OBM];
      // phase 1, at most 4 network inputs: this thread's rows of W1 and b1 (H1 / 32 rows, the same for every box) stay in
      // registers instead of being re-read from shared memory for every box (40 registers for the critic; the 8-input
      // build already sits at the register cap and keeps the loads)
      constexpr bool W1REG = OBM <= 4;
      float w1r[W1REG ? H1 / 32 : 1][W1REG ? OBM : 1], b1r[W1REG ? H1 / 32 : 1];
      if (W1REG && phase == 1) {
#pragma unroll
        for (int m = 0; m < H1 / 32; ++m) {
          b1r[W1REG ? m : 0] = B.b1[r0 + 32 * m];
#pragma unroll
          for (int i = 0; i < OBM; ++i) w1r[W1REG ? m : 0][W1REG ? i : 0] = i < ob ? B.W1[i * H1 + r0 + 32 * m] : 0.f;
        }
      }
#ifdef ACX_WG_PROF
      const bool sprof_ = blockIdx.x == 0 && tid == 64 + 128;
      long long sp_[5] = {0, 0, 0, 0, 0}, sl_ = clock64();
#endif
      if (phase == 1) load_obs((int)blockIdx.x < ntiles, (int)blockIdx.x / tpr, ((int)blockIdx.x % tpr) * 128, 0, o_next);
      int in_group = 0, groups_done = 0;          // within this phase
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int n0 = (tile % tpr) * 128;
        const int nsub = (min(128, n - n0) + BOX - 1) / BOX;
        const int trow = tile / tpr, nxt = tile + (int)gridDim.x, nrow = nxt / tpr, nn0 = (nxt % tpr) * 128;
        for (int kk = 0; kk < nsub; ++kk, ++it) {
          const unsigned int slot = it % NSTAGE;
          float4* r4 = reinterpret_cast<float4*>(raw + (size_t)slot * Cfg::STAGE);
          float4* l4 = reinterpret_cast<float4*>(lo + (size_t)slot * Cfg::STAGE);
          if (phase == 1) {
            // A = X1 blocks, generated: X1[k][s] = act(sum_i obs[i][s] W1[i][k] + b1[k])  (layer1_block's arithmetic).
            // Needs no loaded data, only the slot: start as soon as the MMAs that read it last are done.
#pragma unroll
            for (int i = 0; i < OBM; ++i) o[i] = o_next[i];
            if (kk + 1 < nsub) load_obs(true, trow, n0, kk + 1, o_next); else load_obs(nxt < ntiles, nrow, nn0, 0, o_next);
            WG_SPLIT_MARK(4);
            if (it >= (unsigned)NSTAGE)
              acx_mbar_wait(reinterpret_cast<uint64_t*>(&B.empty[slot]), ((it / NSTAGE) - 1u) & 1u);
            WG_SPLIT_MARK(0);
#pragma unroll
            for (int m = 0; m < H1 / 32; ++m) {
              const int k = r0 + 32 * m;
              float4 sacc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
              for (int i = 0; i < OBM; ++i)
                if (i < ob) {
                  const float w = W1REG ? w1r[W1REG ? m : 0][W1REG ? i : 0] : B.W1[i * H1 + k];
                  sacc.x = fmaf(o[i].x, w, sacc.x); sacc.y = fmaf(o[i].y, w, sacc.y);
                  sacc.z = fmaf(o[i].z, w, sacc.z); sacc.w = fmaf(o[i].w, w, sacc.w);
                }
              const float bk = W1REG ? b1r[W1REG ? m : 0] : B.b1[k];
              float4 x;
              x.x = acxmlp::act_fn<LRELU>(__fadd_rn(sacc.x, bk)); x.y = acxmlp::act_fn<LRELU>(__fadd_rn(sacc.y, bk));
              x.z = acxmlp::act_fn<LRELU>(__fadd_rn(sacc.z, bk)); x.w = acxmlp::act_fn<LRELU>(__fadd_rn(sacc.w, bk));
              const int idx = (k >> 7) * (BLK / 4) + sw128_f4(k & 127, c4);
              r4[idx] = x;
              l4[idx] = lo_of(x);
            }
          }
          WG_SPLIT_MARK(1);
          acx_mbar_wait(reinterpret_cast<uint64_t*>(&B.full[slot]), (it / NSTAGE) & 1u);
          WG_SPLIT_MARK(2);
          if (phase == 0) {
#pragma unroll
            for (int m = 0; m < 4; ++m) {          // A = X2 box: rows r0 + 32 m, stored chunk c4 (any bijection works)
              const int idx = (r0 + 32 * m) * 8 + c4;
              l4[idx] = lo_of(r4[idx]);
            }
          }
          // B = dPre box: split, and add this thread's 4 samples of rows r0 + 32 m to the bias-gradient partials
#pragma unroll
          for (int m = 0; m < 4; ++m)
            if (r0 + 32 * m < brows) {
              const int idx = ABLK * (BLK / 4) + (r0 + 32 * m) * 8 + c4;
              const float4 x = r4[idx];
              l4[idx] = lo_of(x);
              bacc[m] += (x.x + x.y) + (x.z + x.w);
            }
          fence_proxy_async();
          mbar_arrive(&B.split[slot]);
          WG_SPLIT_MARK(3);
          if (warp < 6 && pend_phase >= 0) flush_slice();     // a slice of the previous group per box
        }
        if (++in_group == WG_FLUSH || tile + (int)gridDim.x >= ntiles) {
          if (warp < 6) {
            if (pend_phase >= 0) flush();
            pend_phase = phase; pend_buf = group & 1u; pend_first = groups_done == 0 ? 1 : 0;
          }
          ++group;
          ++groups_done;
          in_group = 0;
        }
      }
      // the phase's last group now (the next phase's accumulators overlap this phase's: the issuer waits for it, and the
      // ring is too short to feed a sliced flush meanwhile)
      if (warp < 6 && pend_phase >= 0) flush();
#ifdef ACX_WG_PROF
      if (sprof_) {
        for (int i = 0; i < 4; ++i) acx_wg_prof[8 + 4 * phase + i] = sp_[i];
        acx_wg_prof[16 + phase] = sp_[4];
      }
#endif
      // the 8 threads of a row (consecutive lanes) fold their partials; fixed order -> reproducible
      float* scr = scratch + (size_t)blockIdx.x * scr_stride + (phase == 0 ? offb3 : offb2);
#pragma unroll
      for (int m = 0; m < 4; ++m) {
        float v = bacc[m];
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        if (c4 == 0 && r0 + 32 * m < brows) scr[r0 + 32 * m] = v;
      }
    }
    if (warp < 6 && pend_phase >= 0) flush();                 // the last group
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

template <int H1, int H3, bool LRELU, int OBM, int PH>
int launch_wgrad_ph(const AcxWgradArgs& a, const CUtensorMap& mX2, const CUtensorMap& mD3, const CUtensorMap& mD2,
                    cudaStream_t st) {
  auto kern = acx_wgrad_kernel<H1, H3, LRELU, OBM, PH>;
  ACX_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WgCfg<H1, PH>::SMEM));
  kern<<<a.grid, WG_THREADS, WgCfg<H1, PH>::SMEM, st>>>(mX2, mD3, mD2, a.obs, a.params, a.ob, a.offW1, a.offb1, a.rows,
                                                       a.n, a.scratch, a.scr_stride, a.offW3, a.offW2, a.offb3, a.offb2);
  ACX_LAUNCH_OK("acx_wgrad_kernel");
  return ACX_OK;
}

template <int H1, int H3, bool LRELU, int OBM>
int launch_wgrad(const AcxWgradArgs& a, cudaStream_t st) {
  CUtensorMap mX2, mD3, mD2;
  // tile-blocked arrays (activation cache, dPre streams): [tiles * F rows][128 samples]
  const uint64_t tiles = (uint64_t)a.rows * (uint64_t)((a.n + 127) / 128);
  int rc = acx_tmap_2d(&mX2, a.x2, 128, tiles * 128, 512, BOX, 128);
  if (!rc) rc = acx_tmap_2d(&mD3, a.d3, 128, tiles * H3, 512, BOX, H3);
  if (!rc) rc = acx_tmap_2d(&mD2, a.d2, 128, tiles * 128, 512, BOX, 128);
  ACX_REQUIRE(rc == 0, "acx_wgrad: cuTensorMapEncodeTiled failed (%d) for rows=%d n=%d", rc, a.rows, a.n);
  rc = launch_wgrad_ph<H1, H3, LRELU, OBM, 0>(a, mX2, mD3, mD2, st);                        // dW3, db3
  if (rc) return rc;
  return launch_wgrad_ph<H1, H3, LRELU, OBM, 1>(a, mX2, mD3, mD2, st);                      // dW2, db2
}

}  // namespace

int acx_wgrad_launch(const AcxWgradArgs& a, void* stream) {
  ACX_REQUIRE((a.n & 3) == 0, "acx_wgrad: n must be a multiple of 4 (16-byte row pitch for the TMA tensor maps), got %d", a.n);
  ACX_REQUIRE(acx_aligned16(a.x2) && acx_aligned16(a.d3) && acx_aligned16(a.obs) && acx_aligned16(a.d2) &&
                  acx_aligned16(a.scratch), "acx_wgrad: operands must be 16-byte aligned");
  ACX_REQUIRE(a.ob >= 1 && a.ob <= acxmlp::MAX_OB, "acx_wgrad: ob_dim %d outside [1,%d]", a.ob, acxmlp::MAX_OB);
  cudaStream_t st = (cudaStream_t)stream;
  if (a.H1 == 128 && a.H3 == 64 && a.lrelu)
    return a.ob <= 4 ? launch_wgrad<128, 64, true, 4>(a, st) : launch_wgrad<128, 64, true, 8>(a, st);
  if (a.H1 == 256 && a.H3 == 128 && !a.lrelu)
    return a.ob <= 4 ? launch_wgrad<256, 128, false, 4>(a, st) : launch_wgrad<256, 128, false, 8>(a, st);
  ACX_REQUIRE(false, "acx_wgrad: unsupported layer widths H1=%d H3=%d", a.H1, a.H3);
  return ACX_OK;
}

#ifdef ACX_WG_PROF
extern "C" int acx_wg_prof_read(unsigned long long* out16) {
  return (int)cudaMemcpyFromSymbol(out16, acx_wg_prof, sizeof(unsigned long long) * 24);
}
#endif
