// Tensor-core (tcgen05, 3xTF32) variant of the forward passes of K2: policy sample, policy log-prob / KL,
// value.  Same math as acx_mlp.cu (A3C/policies.py:42-72,110-115,146-150,176-177), same heads; the two
// 128-wide hidden GEMMs run on the 5th-generation tensor cores with fp32-class accuracy:
//     x ~= hi + lo  (both exactly representable in TF32),   A.B ~= Alo.Bhi + Ahi.Blo + Ahi.Bhi
// accumulated in fp32 in tensor memory (measured 2-7e-7 relative on random operands, tests/tc).
//
// One persistent CTA per SM walks 128-sample tiles; TMEM lane = sample.
//   * Activations never touch shared memory: they are the **A operand in tensor memory** (columns 0..127 hi,
//     128..255 lo), written by the epilogue threads with tcgen05.st; accumulators D2 / D3 sit in columns
//     256..383 / 384..511.  The epilogue thread of sample s reads its accumulator row (tcgen05.ld 32x32b),
//     applies bias + activation, optionally writes the fp32 activation cache for the training kernels, splits
//     into hi/lo and stores the next layer's A operand.
//   * Shared memory therefore belongs to the weights (B operand, un-swizzled K-major core-matrix layout
//     [n/8][k/4][8][4], hi and lo planes, packed once per parameter update by acx_tc_pack_weights): twelve
//     16 KB slots.  The actor's 12 chunks stay resident for the whole kernel; the critic's 24 chunks per tile
//     stream through the slots as a ring (bulk-TMA, full barriers; tcgen05.commit frees a slot).
//   * Warp 8 is the issuer: one elected thread runs TMA prefetch + tcgen05.mma + commits; warps 0-7 compute
//     layer 1 (ob_dim <= 8, FFMA), the epilogues and the heads, and hand A over through mbarriers.
// Only K-major / TMEM operands are used: un-swizzled MN-major TF32 operands return zeros on this part
// (tests/tc/run_tc_test.py), which is why the weight-gradient GEMMs stay on the FFMA engine.
#include "acx_mlp.cuh"
#include "acx_umma.cuh"
#include "acx_env_dev.cuh"
#include <stdlib.h>

using namespace acxmlp;
using namespace acxumma;

// acx_mlp_tc2.cu: the evaluation-mode passes with two tiles in flight per SM.  Parity-tested like this file's kernel,
// but measured SLOWER on B200 (value pass 1.00 vs 0.93 ms, log-prob pass 0.68 vs 0.55 ms at 65,536 x 20: both kernels
// are bound by the workers' instruction issue, not by tensor-pipe idle time -- profiles/r02_forward_experiments.md), so
// it runs only when ACX_TC_TWO_TILES=1 is set in the environment.
int acx_tc2_launch(int head, const Args& a, const float* packed, int grid, void* stream);
static bool tc_two_tiles() {
  static const bool on = [] { const char* e = getenv("ACX_TC_TWO_TILES"); return e && e[0] == '1'; }();
  return on;
}

#ifdef ACX_FWD_PROF
__device__ unsigned long long acx_fwd_prof[16];
#define FPROF(i) do { if (tid == 0) { const long long now_ = clock64(); prof_[i] += now_ - last_; last_ = now_; } } while (0)
#else
#define FPROF(i) do { } while (0)
#endif

namespace {

constexpr int TC_EPI = 256;                      // epilogue threads (warps 0-7)
constexpr int TC_NT = TC_EPI + 64;               // + issuer warp (8) + producer warp (9)
constexpr int TC_SLOT_FLOATS = 4096;             // 16 KB: hi + lo planes of one weight chunk
constexpr int TC_NSLOT = 12;
constexpr int TC_KC2 = 16;                       // layer-2 input rows per chunk (128 outputs)
constexpr uint32_t TC_A_HI = 0, TC_A_LO = 128, TC_D2 = 256, TC_D3 = 384;

struct TcSmem {
  float slot[TC_NSLOT][TC_SLOT_FLOATS];          // 196,608 B
  float obs[MAX_OB * MT];
  float W1[MAX_OB * 256];
  float b1[256];
  float b2[128];
  float b3[128];
  float Wh[128 * MAX_OUT];
  float bh[MAX_OUT];
  float log_std[MAX_OUT];
  float pre[MAX_OUT * MT];                       // partial head sums of the second warpgroup
  float4 W1p[256];                               // ob_dim <= 3: {W1[0][k], W1[1][k], W1[2][k], b1[k]}
  float obs_x[MAX_OB * MT];                      // rollout: the observations the env step just produced
  double red[8];
  unsigned long long full[TC_NSLOT], empty[TC_NSLOT], a_ready[2], stage_done;
  unsigned int tmem_base;
};

__device__ inline void epi_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }     // warps 0-7 only
__device__ inline void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(acx_smem_u32(bar)) : "memory");
}

// The issuer's view of the weight stream.  Chunk c of a tile: c < n2 -> layer 2 (TC_KC2 rows x 128), else layer 3
// (kc3 rows x H3).  All chunks are 16 KB.  resident: chunks_per_tile == TC_NSLOT, loaded once.
struct TcPipe {
  TcSmem* sm;
  const float* packed;
  unsigned int g, total;
  int chunks_per_tile;
  bool resident;

  // producer (its own warp, see acx_tc_bwd.cuh): streams the CTA's chunks in order as slots free up (resident: the
  // first TC_NSLOT chunks only -- they are the whole network and stay)
  __device__ inline void produce() {
    const unsigned int n = resident ? (total < (unsigned)TC_NSLOT ? total : (unsigned)TC_NSLOT) : total;
    for (unsigned int c = 0; c < n; ++c) {
      const unsigned int q = c % TC_NSLOT;
      if (c >= (unsigned)TC_NSLOT)
        acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm->empty[q]), ((c / TC_NSLOT) - 1u) & 1u);
      const float* src = packed + (size_t)(c % (unsigned)chunks_per_tile) * TC_SLOT_FLOATS;
      acx_mbar_expect_tx(reinterpret_cast<uint64_t*>(&sm->full[q]), TC_SLOT_FLOATS * 4);
      acx_bulk_g2s(sm->slot[q], src, TC_SLOT_FLOATS * 4, reinterpret_cast<uint64_t*>(&sm->full[q]));
    }
  }
  // one chunk: kc input rows starting at A column k0, n_out outputs, accumulate into tmem_d
  __device__ inline void mma_chunk(uint32_t tmem, int k0, int kc, int n_out, uint32_t tmem_d, bool first) {
    const unsigned int q = g % TC_NSLOT;
    acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm->full[q]), resident ? 0u : ((g / TC_NSLOT) & 1u));
    fence_after_sync();
    const float* b_hi = sm->slot[q];
    const float* b_lo = b_hi + n_out * kc;
    const uint32_t idesc = idesc_tf32(128, n_out, false, false);
    const uint32_t sbo = (uint32_t)(kc / 4) * 128u;           // next 8 outputs
    for (int ks = 0; ks < kc / 8; ++ks) {
      const uint64_t bh = smem_desc(b_hi + ks * 64, 128, sbo), bl = smem_desc(b_lo + ks * 64, 128, sbo);
      const uint32_t ah = tmem + TC_A_HI + (uint32_t)(k0 + ks * 8), al = tmem + TC_A_LO + (uint32_t)(k0 + ks * 8);
      mma_tf32_ts(tmem_d, al, bh, idesc, !(first && ks == 0));
      mma_tf32_ts(tmem_d, ah, bl, idesc, true);
      mma_tf32_ts(tmem_d, ah, bh, idesc, true);
    }
    if (!resident) commit(reinterpret_cast<uint64_t*>(&sm->empty[q]));
    ++g;
  }
};

// bit e = (v[e] > 0): the integer image i of a float is > 0 exactly when the float is (+0 -> 0, negatives and -0 -> < 0),
// and the sign bit of -i is set exactly when i > 0 (or i = INT_MIN = -0.0f, which neither activation produces: relu gives
// +0, lrelu's 0.8 max(x, 0) + 0.2 x is +0 only from x = +-0 and rounds to +0) -- one negate and one funnel shift per
// element instead of compare + select + or
__device__ inline unsigned int pos_mask32(const float (&v)[32]) {
  unsigned int bits = 0u;
#pragma unroll
  for (int e = 31; e >= 0; --e) bits = __funnelshift_l((unsigned int)(-__float_as_int(v[e])), bits, 1);
  return bits;
}

// 32 activations of sample `s` (features j0..j0+31) -> A operand columns, hi and lo
__device__ inline void a_store32(uint32_t tmem_lane, int j0, const float (&v)[32]) {
  // hi = the raw word (kind::tf32 ignores its low 13 mantissa bits), lo = the rounded remainder
  float lo[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) lo[i] = lo_trunc_tf32(v[i]);
  tmem_st32(tmem_lane + TC_A_HI + (uint32_t)j0, v);
  tmem_st32(tmem_lane + TC_A_LO + (uint32_t)j0, lo);
}

// ENVR >= 0 (MODE_SAMPLE only): persistent rollout of environment ENVR.  A.rows = T time steps; a CTA keeps its column
// tiles for all T steps, the sampling head steps the environment (acx_env_dev.cuh, the arithmetic of acx_env_step) with
// the state in registers, and the next step's observation reaches layer 1 through shared memory instead of HBM.
template <int H1, int H3, bool LRELU, int HEAD, int MODE, int ENVR = -1>
__global__ void __launch_bounds__(TC_NT, 1) acx_mlp_tc_kernel(const Args A, const float* __restrict__ packed) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  TcSmem& sm = *reinterpret_cast<TcSmem*>(smem_raw);
  constexpr int NB = H1 / 128;                   // 128-feature blocks of layer 1
  constexpr int KC3 = 2048 / H3;                 // layer-3 rows per 16 KB chunk: 16 (critic) / 32 (actor)
  constexpr int N3C = 128 / KC3;                 // layer-3 chunks
  const AcxNetLayout L = A.L;
  const int ob = L.ob, nout = L.nout, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const float* __restrict__ P = A.params;

  for (int i = tid; i < ob * H1; i += TC_NT) sm.W1[i] = P[L.W1 + i];
  for (int i = tid; i < H1; i += TC_NT) sm.b1[i] = P[L.b1 + i];
  if (ob <= 3)
    for (int k = tid; k < H1; k += TC_NT)
      sm.W1p[k] = make_float4(P[L.W1 + k], ob > 1 ? P[L.W1 + H1 + k] : 0.f, ob > 2 ? P[L.W1 + 2 * H1 + k] : 0.f, P[L.b1 + k]);
  for (int i = tid; i < H2; i += TC_NT) sm.b2[i] = P[L.b2 + i];
  for (int i = tid; i < H3; i += TC_NT) sm.b3[i] = P[L.b3 + i];
  for (int i = tid; i < H3 * nout; i += TC_NT) sm.Wh[i] = P[L.Wh + i];
  if (tid < nout) {
    sm.bh[tid] = P[L.bh + tid];
    if (HEAD == HEAD_GAUSS) sm.log_std[tid] = P[L.log_std + tid];
  }
  constexpr bool ROLL = ENVR >= 0;
  static_assert(!ROLL || (H3 == 64 && H1 == 128), "the rollout keeps two 64-column layer-3 accumulators (D3 buffer = unit & 1)");
  const int tpr = (A.n + MT - 1) / MT;
  const int ntiles = A.rows * tpr;
  const int nouter = ROLL ? tpr : ntiles;         // what a CTA iterates over: column tiles (x T steps each) or tiles
  const int T = ROLL ? A.rows : 1;
  const unsigned int mine =
      ((int)blockIdx.x < nouter) ? (unsigned)((nouter - blockIdx.x + gridDim.x - 1) / gridDim.x) * (unsigned)T : 0u;

  if (warp == 8) tmem_alloc(&sm.tmem_base, 512);
  if (tid == 0) {
    for (int i = 0; i < TC_NSLOT; ++i) {
      acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.full[i]), 1);
      acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.empty[i]), 1);
    }
    for (int i = 0; i < 2; ++i) acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.a_ready[i]), TC_EPI);
    acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.stage_done), 1);
    acx_mbar_fence_init();
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = sm.tmem_base;

  TcPipe pipe;
  pipe.sm = &sm; pipe.packed = packed; pipe.g = 0;
  pipe.chunks_per_tile = H1 / TC_KC2 + N3C;
  pipe.resident = (pipe.chunks_per_tile == TC_NSLOT);
  pipe.total = mine * pipe.chunks_per_tile;
  if (warp == 9) {
    // ================================================================ producer of the weight-chunk ring
    if (lane == 0) pipe.produce();
  } else if (warp == 8) {
    // ================================================================ issuer
    if (lane == 0) {
      unsigned int ready = 0;
      for (unsigned int t = 0; t < mine; ++t) {
        for (int blk = 0; blk < NB; ++blk) {                     // layer 2, one 128-feature block of X1 at a time
          // the A operand arrives in two halves of 64 features (a_ready[0], then [1]): the MMAs over the first half
          // run while the epilogue warps still produce the second
          for (int sub = 0; sub < 2; ++sub) {
            acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.a_ready[sub]), ready & 1u);
            fence_after_sync();
            for (int kc = sub * (64 / TC_KC2); kc < (sub + 1) * (64 / TC_KC2); ++kc)
              pipe.mma_chunk(tmem, kc * TC_KC2, TC_KC2, 128, tmem + TC_D2, blk == 0 && kc == 0);
          }
          ++ready;
          commit(reinterpret_cast<uint64_t*>(&sm.stage_done));
        }
        for (int sub = 0; sub < 2; ++sub) {                                                   // X2 likewise
          acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.a_ready[sub]), ready & 1u);
          fence_after_sync();
          for (int kc = sub * (N3C / 2); kc < (sub + 1) * (N3C / 2); ++kc)
            pipe.mma_chunk(tmem, kc * KC3, KC3, H3, tmem + TC_D3 + (ROLL ? 64u * (t & 1u) : 0u), kc == 0);
        }
        ++ready;
        commit(reinterpret_cast<uint64_t*>(&sm.stage_done));
      }
    }
  } else {
    // ================================================================ epilogue warps
    const int s = (warp & 3) * 32 + lane;          // sample = TMEM lane
    const int half = warp >> 2;                    // column half handled by this warpgroup
    const uint32_t tl = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    const unsigned int step0 = A.step0 + ((MODE == MODE_SAMPLE && A.step_base) ? *A.step_base : 0u);
    const bool basic = (HEAD == HEAD_VALUE) && A.basic_head;
    double kl_sum = 0.0;
    unsigned int done = 0;                         // completed waits on stage_done

    // this thread's observation of tile `tile` (coalesced; zeros beyond the batch), fetched one tile ahead
    auto load_obs = [&](int tile, float (&o)[MAX_OB]) {
      const int row = tile / tpr, c = (tile % tpr) * MT + s;
#pragma unroll
      for (int i = 0; i < MAX_OB; ++i)
        o[i] = (i < ob && tile < ntiles && c < A.n) ? __ldg(A.obs + ((size_t)row * ob + i) * A.n + c) : 0.f;
    };
    float o_next[MAX_OB];
    if (!ROLL) load_obs(blockIdx.x, o_next);
#ifdef ACX_FWD_PROF
    long long prof_[16] = {0}, last_ = clock64();
#endif
    // rollout: this sample's environment (first warpgroup), in registers for all T steps
    constexpr int ES = AcxEnv<ROLL ? ENVR : 0>::S, EOB = AcxEnv<ROLL ? ENVR : 0>::OB;   // (unused unless ROLL)
    float est[ES], est_b[ES];                      // (est_b / eel_b: the second tile of a pair, see the pair loop)
    int eel = 0, eel_b = 0;
    const int emax = ROLL ? (A.max_steps > 0 ? A.max_steps : AcxEnv<ROLL ? ENVR : 0>::MAX_STEPS) : 0;
    // one environment transition after the action of step `row` (the arithmetic and order of acx_env_step_kernel)
    auto env_step = [&](float act, int row, size_t col, float (&est)[ES], int& eel, float* __restrict__ obsx) {
      if (!ROLL) return;
      float rew;
      bool term;
      float ov[EOB];
      acx_env_transition<ROLL ? ENVR : 0>(est, act, rew, term, ov);
      int e = eel + 1;
      const bool dn = term || (e >= emax);
      if (dn) {
        float ru[ES];
        if (A.reset_u != nullptr) {
#pragma unroll
          for (int k = 0; k < ES; ++k) ru[k] = A.reset_u[((size_t)row * ES + k) * A.n + col];
        } else {
          const AcxRng4 g = acx_philox(A.index0 + col, step0 + (unsigned)row, ACX_STREAM_RESET, A.seed);
          const float u4[4] = {acx_u01(g.x), acx_u01(g.y), acx_u01(g.z), acx_u01(g.w)};
#pragma unroll
          for (int k = 0; k < ES; ++k) ru[k] = u4[k];
        }
        AcxEnv<ROLL ? ENVR : 0>::reset(ru, est);
        e = 0;
      }
      acx_env_observe_after<ROLL ? ENVR : 0>(est, dn, ov);
      eel = e;
      A.rew_out[(size_t)row * A.n + col] = rew;
      A.done_out[(size_t)row * A.n + col] = dn ? 1 : 0;
#pragma unroll
      for (int k = 0; k < EOB; ++k) {
        A.obs_rw[((size_t)(row + 1) * EOB + k) * A.n + col] = ov[k];
        obsx[k * MT + s] = ov[k];
      }
    };

    // wait for the layer-2 MMAs, then X2 = act(D2 + b2) -> A operand (+ cache, masks) in two hand-overs
    auto ph_x2 = [&](int tile, bool live, float (&pre)[MAX_OUT]) {
      // ---------------------------------------------- layer-2 epilogue: X2 = act(D2 + b2) -> A operand (+ cache)
      acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.stage_done), done & 1u); ++done;
      fence_after_sync();
      FPROF(4);
#pragma unroll 1
      for (int gq = 0; gq < 2; ++gq) {
        float v[32];
        const int j0 = gq * 64 + half * 32;
        tmem_ld32(tl + TC_D2 + (uint32_t)j0, v);
#pragma unroll
        for (int e = 0; e < 32; ++e) v[e] = act_fn<LRELU>(__fadd_rn(v[e], sm.b2[j0 + e]));
        a_store32(tl, j0, v);
        tmem_wait_st();
        fence_before_sync();
        mbar_arrive(&sm.a_ready[gq]);
        // everything else this thread does with the 32 values comes AFTER the hand-over: the cache / mask stores and
        // the Basic_AC head sums then run while the tensor cores form layer 3
        if (A.mask_out != nullptr) {               // act'(X2)
          unsigned int bits = 0u;
          bits = pos_mask32(v);
          A.mask_out[mask_tile_off(tile, H1) + (size_t)(H1 / 32 + gq * 2 + half) * MT + s] = bits;
        }
        if (A.x2_out != nullptr) {                 // tile-blocked cache; zeros beyond the batch
          float* g2 = A.x2_out + cache_tile_off(tile, H2) + (size_t)j0 * MT + s;
#pragma unroll
          for (int e = 0; e < 32; ++e) g2[e * MT] = live ? v[e] : 0.f;
        }
        if (basic) {
#pragma unroll
          for (int e = 0; e < 32; ++e) pre[0] = fmaf(v[e], sm.Wh[j0 + e], pre[0]);
        }
      }
      FPROF(5);
    };
    // X3 = act(D3 + b3) (+ cache) from the accumulator at column d3, head partial sums, exchange, head: outputs of the
    // sample; pact = the action the env step takes (rollout)
    auto ph_x3 = [&](int tile, int row, size_t col, bool live, uint32_t d3, float (&pre)[MAX_OUT], float& pact) {
      (void)tile;
#pragma unroll 1
      for (int gq = 0; gq < H3 / 64; ++gq) {       // this thread: H3/2 features
        float v[32];
        const int j0 = half * (H3 / 2) + gq * 32;
        tmem_ld32(tl + d3 + (uint32_t)j0, v);
#pragma unroll
        for (int e = 0; e < 32; ++e) v[e] = act_fn<LRELU>(__fadd_rn(v[e], sm.b3[j0 + e]));
        if (A.x3_out != nullptr) {
          float* g3 = A.x3_out + cache_tile_off(tile, H3) + (size_t)j0 * MT + s;
#pragma unroll
          for (int e = 0; e < 32; ++e) g3[e * MT] = live ? v[e] : 0.f;
        }
        if (!basic) {
          if (nout == 1) {
            // one output (value head, Pendulum's mean): four independent FMA chains over 16-byte weight loads instead of
            // one 32-deep chain with a predicated 4-byte load per term (9 % of the value pass's warp samples, ncu)
            float p0 = 0.f, p1 = 0.f, p2 = 0.f, p3 = 0.f;
#pragma unroll
            for (int e = 0; e < 32; e += 4) {
              const float4 w = *reinterpret_cast<const float4*>(&sm.Wh[j0 + e]);
              p0 = fmaf(v[e], w.x, p0); p1 = fmaf(v[e + 1], w.y, p1);
              p2 = fmaf(v[e + 2], w.z, p2); p3 = fmaf(v[e + 3], w.w, p3);
            }
            pre[0] = __fadd_rn(pre[0], __fadd_rn(__fadd_rn(p0, p1), __fadd_rn(p2, p3)));
          } else {
#pragma unroll
            for (int e = 0; e < 32; ++e) {
#pragma unroll
              for (int a = 0; a < MAX_OUT; ++a)
                if (a < nout) pre[a] = fmaf(v[e], sm.Wh[(j0 + e) * nout + a], pre[a]);
            }
          }
        }
      }
      if (half == 1) {
#pragma unroll
        for (int a = 0; a < MAX_OUT; ++a)
          if (a < nout) sm.pre[a * MT + s] = pre[a];
      }
      FPROF(7);
      epi_sync();
      FPROF(8);
      if (half == 0) {
#pragma unroll
        for (int a = 0; a < MAX_OUT; ++a)
          if (a < nout) pre[a] = __fadd_rn(__fadd_rn(pre[a], sm.pre[a * MT + s]), sm.bh[a]);
        const size_t s_idx = (size_t)row * A.n + col;
        if (live) {
          if (HEAD == HEAD_VALUE) {
            A.v_out[s_idx] = pre[0];
          } else if (HEAD == HEAD_GAUSS) {
            float lp = 0.f, ent = 0.f, ac_env = 0.f;
            AcxRng4 g = {0u, 0u, 0u, 0u};
            if (MODE == MODE_SAMPLE && A.noise == nullptr)
              g = acx_philox(A.index0 + col, step0 + (unsigned)row, ACX_STREAM_ACTION, A.seed);
#pragma unroll
            for (int d = 0; d < MAX_OUT; ++d)
              if (d < nout) {
                const float mu = __fmul_rn(tanhf(pre[d]), A.ac_scale);
                const float ls = fminf(fmaxf(sm.log_std[d], ACX_LOG_STD_MIN), ACX_LOG_STD_MAX);
                const float sd = expf(ls);
                const size_t a_idx = ((size_t)row * nout + d) * A.n + col;
                float ac;
                if (MODE == MODE_SAMPLE) {
                  float eps;
                  if (A.noise) eps = A.noise[a_idx];
                  else eps = (d == 0) ? acx_normal(g.x, g.y) : ((d == 1) ? acx_normal(g.z, g.w) : 0.f);
                  ac = __fadd_rn(mu, __fmul_rn(sd, eps));
                  A.ac_out_f[a_idx] = ac;
                  if (d == 0) ac_env = ac;
                } else {
                  ac = A.ac_in_f[a_idx];
                }
                if (A.head_out) A.head_out[a_idx] = mu;
                const float z = __fdiv_rn(__fsub_rn(ac, mu), sd);
                lp += __fsub_rn(__fsub_rn(__fmul_rn(__fmul_rn(-0.5f, z), z), ls), ACX_HALF_LOG_2PI);
                ent += __fadd_rn(__fadd_rn(0.5f, ACX_HALF_LOG_2PI), ls);
              }
            if (A.logp_out) A.logp_out[s_idx] = lp;
            if (A.ent_out) A.ent_out[s_idx] = ent;
            if (MODE == MODE_EVAL && A.logp_old) {
              const double df = (double)A.logp_old[s_idx] - (double)lp;
              kl_sum += df * df;
            }
            pact = ac_env;
          } else {   // HEAD_CAT
            float l1[MAX_OUT], logps[MAX_OUT], p[MAX_OUT];
            float mx1 = -INFINITY, mx2 = -INFINITY;
#pragma unroll
            for (int a = 0; a < MAX_OUT; ++a)
              if (a < nout) {
                l1[a] = __fadd_rn(pre[a], 1e-8f);
                mx1 = fmaxf(mx1, l1[a]); mx2 = fmaxf(mx2, pre[a]);
              }
            float se1 = 0.f, se2 = 0.f;
#pragma unroll
            for (int a = 0; a < MAX_OUT; ++a)
              if (a < nout) {
                se1 += expf(l1[a] - mx1);
                p[a] = expf(pre[a] - mx2);
                se2 += p[a];
              }
            const float lse = mx1 + logf(se1);
            float ent = 0.f;
#pragma unroll
            for (int a = 0; a < MAX_OUT; ++a)
              if (a < nout) {
                logps[a] = l1[a] - lse;
                p[a] = p[a] / se2;
                ent -= p[a] * logps[a];
              }
            int ac;
            if (MODE == MODE_SAMPLE) {
              float u;
              if (A.noise) u = A.noise[s_idx];
              else u = acx_u01(acx_philox(A.index0 + col, step0 + (unsigned)row, ACX_STREAM_ACTION, A.seed).x);
              double cdf[MAX_OUT], run = 0.0;
#pragma unroll
              for (int a = 0; a < MAX_OUT; ++a)
                if (a < nout) { run += exp((double)pre[a] - (double)mx2); cdf[a] = run; }
              const double thr = (double)u * run;
              ac = 0;
#pragma unroll
              for (int a = 0; a < MAX_OUT; ++a)
                if (a < nout && cdf[a] <= thr) ++ac;
              if (ac > nout - 1) ac = nout - 1;
              A.ac_out_i[s_idx] = ac;
            } else {
              ac = A.ac_in_i[s_idx];
            }
            float lp = 0.f;
#pragma unroll
            for (int a = 0; a < MAX_OUT; ++a)
              if (a < nout && a == ac) lp = logps[a];
            if (A.head_out) {
#pragma unroll
              for (int a = 0; a < MAX_OUT; ++a)
                if (a < nout) A.head_out[((size_t)row * nout + a) * A.n + col] = pre[a];
            }
            if (A.logp_out) A.logp_out[s_idx] = lp;
            if (A.ent_out) A.ent_out[s_idx] = ent;
            if (MODE == MODE_EVAL && A.logp_old) {
              const double df = (double)A.logp_old[s_idx] - (double)lp;
              kl_sum += df * df;
            }
            pact = (float)ac;
          }
        }
      }
    };
    // rollout: the env step of the sample (first warpgroup) -> reward, done, next observation (HBM and obsx)
    auto tail_env = [&](float pact, int row, size_t col, bool live, float (&est_x)[ES], int& eel_x, float* __restrict__ obsx) {
      if (half != 0) return;
      if (live) {
        env_step(pact, row, col, est_x, eel_x, obsx);
      } else {
#pragma unroll
        for (int k = 0; k < EOB; ++k) obsx[k * MT + s] = 0.f;
      }
    };
    unsigned int units = 0;                        // rollout: (tile, step) units done; unit u accumulates layer 3 in D3 buffer u & 1
    // Layer 1 of one 128-feature block of tile `tile_l` from the observation in o[] -> A operand (two hand-overs of 64
    // features).  In the evaluation modes block 0 of the NEXT tile is produced right after the wait for this tile's
    // layer-3 MMAs, before this tile's X3 epilogue and head: the tensor cores then form the next tile's layer 2 while
    // the workers finish this tile (the A operand is free once layer 3 has been read; D3 is not touched by layer 2).
    // The rollout cannot do that: its next observation comes out of this step's head.
    float o[MAX_OB];
    auto l1_block = [&](int blk, int tile_l) {
#pragma unroll 1
      for (int gq = 0; gq < 2; ++gq) {           // this thread: features gq*64 + half*32 .. +31 of the block
        float v[32];
        const int j0 = gq * 64 + half * 32;
        if (ob <= 3) {
#pragma unroll
          for (int e = 0; e < 32; ++e) {
            const float4 w = sm.W1p[blk * 128 + j0 + e];
            float acc = fmaf(o[0], w.x, 0.f);
            if (ob > 1) acc = fmaf(o[1], w.y, acc);
            if (ob > 2) acc = fmaf(o[2], w.z, acc);
            v[e] = act_fn<LRELU>(__fadd_rn(acc, w.w));
          }
        } else {
#pragma unroll
          for (int e = 0; e < 32; ++e) {
            const int k = blk * 128 + j0 + e;
            float acc = 0.f;
#pragma unroll
            for (int i = 0; i < MAX_OB; ++i)
              if (i < ob) acc = fmaf(o[i], sm.W1[i * H1 + k], acc);
            v[e] = act_fn<LRELU>(__fadd_rn(acc, sm.b1[k]));
          }
        }
        if (blk > 0 && gq == 0) {                // the MMAs reading the previous block must be done
          FPROF(1);
          acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.stage_done), done & 1u); ++done;
          fence_after_sync();
          FPROF(2);
        }
        a_store32(tl, j0, v);
        tmem_wait_st();
        fence_before_sync();
        mbar_arrive(&sm.a_ready[gq]);            // features gq*64 .. +63 of the A operand are in place
        if (A.mask_out != nullptr) {             // act'(X1) for the chain kernel (after the hand-over)
          unsigned int bits = 0u;
          bits = pos_mask32(v);
          A.mask_out[mask_tile_off(tile_l, H1) + (size_t)(blk * 4 + gq * 2 + half) * MT + s] = bits;
        }
      }
      FPROF(blk == 0 ? 0 : 3);
    };
    bool l1_ahead = false;                         // block 0 of the tile the loop is about to start is already in flight
    for (int outer = blockIdx.x; outer < nouter; outer += gridDim.x)
    for (int t = 0; t < T; ++t) {
      const int tile = ROLL ? t * tpr + outer : outer;
      const int row = tile / tpr, n0 = (tile % tpr) * MT;
      const int valid = min(MT, A.n - n0);
      const bool live = s < valid;
      const size_t col = (size_t)n0 + s;
      if (!ROLL) {
        if (!l1_ahead) {
#pragma unroll
          for (int i = 0; i < MAX_OB; ++i) o[i] = o_next[i];
          load_obs(tile + gridDim.x, o_next);
        }
      } else if (t == 0) {
        load_obs(tile, o);
        if (half == 0 && live) {
#pragma unroll
          for (int k = 0; k < ES; ++k) est[k] = A.env_state[(size_t)k * A.n + col];
          eel = A.env_elapsed[col];
        }
      } else {
#pragma unroll
        for (int i = 0; i < MAX_OB; ++i) o[i] = i < ob ? sm.obs_x[i * MT + s] : 0.f;
      }

      // ---------------------------------------------- layer 1 -> A operand, one 128-feature block per stage
      for (int blk = 0; blk < NB; ++blk) {
        if (blk == 0 && l1_ahead) continue;
        l1_block(blk, tile);
      }
      float pre[MAX_OUT] = {0.f, 0.f, 0.f, 0.f};
      ph_x2(tile, live, pre);
      // ---------------------------------------------- layer-3 epilogue + head
      acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.stage_done), done & 1u); ++done;
      fence_after_sync();
      FPROF(6);
      if (!ROLL) {                                 // the next tile's layer-1 block 0, ahead of this tile's epilogue
        const int nxt = tile + (int)gridDim.x;
        l1_ahead = nxt < ntiles;
        if (l1_ahead) {
#pragma unroll
          for (int i = 0; i < MAX_OB; ++i) o[i] = o_next[i];
          load_obs(nxt + (int)gridDim.x, o_next);
          l1_block(0, nxt);
        }
        FPROF(10);
      }
      float pact = 0.f;
      ph_x3(tile, row, col, live, ROLL ? TC_D3 + 64u * (units & 1u) : TC_D3, pre, pact);
      if (ROLL) {
        tail_env(pact, row, col, live, est, eel, sm.obs_x);
        ++units;
      }
      if (ROLL) {
        if (t == T - 1 && half == 0 && live) {
#pragma unroll
          for (int k = 0; k < ES; ++k) A.env_state[(size_t)k * A.n + col] = est[k];
          A.env_elapsed[col] = eel;
        }
        epi_sync();              // the next step's observations (obs_x) are visible to both warpgroups
      }
      // (no tcgen05 fence here: every a_ready arrive is preceded by its own fence::before_thread_sync, and this tile's
      // tcgen05.ld have completed -- tmem_ld32 waits for them)
      FPROF(9);
    }
#ifdef ACX_FWD_PROF
    if (tid == 0 && blockIdx.x == 0)
      for (int i = 0; i < 16; ++i) acx_fwd_prof[i] = (unsigned long long)prof_[i];
#endif
    if (MODE == MODE_EVAL && HEAD != HEAD_VALUE && A.kl_partial != nullptr) {
      const double v = acx_warp_sum_d(kl_sum);
      if (lane == 0) sm.red[warp] = v;
    }
  }

  // ---------------------------------------------------- teardown
  fence_before_sync();
  __syncthreads();
  if (warp == 8) tmem_dealloc(tmem, 512);
  if (MODE == MODE_EVAL && HEAD != HEAD_VALUE && A.kl_partial != nullptr && tid == 0)
    A.kl_partial[blockIdx.x] = ((sm.red[0] + sm.red[1]) + (sm.red[2] + sm.red[3])) +
                               ((sm.red[4] + sm.red[5]) + (sm.red[6] + sm.red[7]));
}

// packed: 16 KB chunks.  Layer 2 (H1 x 128): TC_KC2 = 16 input rows per chunk; layer 3 (128 x H3): 2048 / H3
// rows per chunk.  Inside a chunk: hi plane [n/8][kc/4][8][4] then the lo plane.
__global__ void __launch_bounds__(256) acx_tc_pack_kernel(const float* __restrict__ P, AcxNetLayout L,
                                                          float* __restrict__ packed) {
  const int n2 = L.H1 * 128, n3 = 128 * L.H3;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n2 + n3; i += gridDim.x * blockDim.x) {
    const bool l2 = i < n2;
    const int e = l2 ? i : i - n2;
    const int N = l2 ? 128 : L.H3;
    const int kc = l2 ? TC_KC2 : 2048 / L.H3;
    const int k = e / N, n = e % N;                       // W[k][n]
    const float w = P[(l2 ? L.W2 : L.W3) + e];
    float hi, lo;
    split_tf32(w, hi, lo);
    const int chunk = (l2 ? 0 : L.H1 / TC_KC2) + k / kc, kk = k % kc;
    const size_t base = (size_t)chunk * TC_SLOT_FLOATS;
    const int off = ((n >> 3) * (kc / 4) + (kk >> 2)) * 32 + (n & 7) * 4 + (kk & 3);
    packed[base + off] = hi;
    packed[base + (size_t)N * kc + off] = lo;
  }
}

template <int H1, int H3, bool LRELU, int HEAD, int MODE, int ENVR = -1>
int launch_tc(const Args& a, const float* packed, int grid, cudaStream_t st) {
  auto kern = acx_mlp_tc_kernel<H1, H3, LRELU, HEAD, MODE, ENVR>;
  ACX_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TcSmem) + 1024));
  kern<<<grid, TC_NT, sizeof(TcSmem) + 1024, st>>>(a, packed);
  ACX_LAUNCH_OK("acx_mlp_tc_kernel");
  return ACX_OK;
}

int tc_grid(int rows, int n) {
  const long long tiles = (long long)rows * ((n + MT - 1) / MT);
  const int sms = acx_num_sms();
  return (int)(tiles < sms ? tiles : sms);
}

int tc_check(const char* fn, const float* params, const float* packed, int ob_dim, int nout, int rows, int n,
             const float* obs) {
  ACX_REQUIRE(params && packed && obs, "%s: null params/packed/obs", fn);
  ACX_REQUIRE(acx_aligned16(packed), "%s: packed weights must be 16-byte aligned (bulk-TMA source)", fn);
  ACX_REQUIRE(ob_dim >= 1 && ob_dim <= MAX_OB, "%s: ob_dim %d outside [1,%d]", fn, ob_dim, MAX_OB);
  ACX_REQUIRE(nout >= 1 && nout <= MAX_OUT, "%s: output width %d outside [1,%d]", fn, nout, MAX_OUT);
  ACX_REQUIRE(rows >= 0 && n >= 0, "%s: negative batch (rows=%d, n=%d)", fn, rows, n);
  return ACX_OK;
}

}  // namespace

extern "C" size_t acx_tc_packed_floats(int net, int ob_dim, int num_ac, int cont) {
  (void)ob_dim; (void)num_ac; (void)cont;
  return net == 0 ? (size_t)2 * (ACX_A_H1 * 128 + 128 * ACX_A_H3) : (size_t)2 * (ACX_C_H1 * 128 + 128 * ACX_C_H3);
}

extern "C" int acx_tc_pack_weights(int net, const float* params, int ob_dim, int num_ac, int cont, float* packed,
                                   void* stream) {
  ACX_REQUIRE(net == 0 || net == 1, "acx_tc_pack_weights: net must be 0 (actor) or 1 (critic), got %d", net);
  ACX_REQUIRE(params && packed, "acx_tc_pack_weights: null buffer");
  const AcxNetLayout L = net == 0 ? acx_actor_layout(ob_dim, num_ac, cont) : acx_critic_layout(ob_dim);
  acx_tc_pack_kernel<<<96, 256, 0, (cudaStream_t)stream>>>(params, L, packed);
  ACX_LAUNCH_OK("acx_tc_pack_kernel");
  return ACX_OK;
}

extern "C" int acx_policy_forward_sample_gauss_tc(const float* actor, const float* packed, int ob_dim, int ac_dim,
                                                  float ac_scale, int rows, int n, const float* obs, const float* eps,
                                                  uint64_t seed, uint64_t index0, uint32_t step0,
                                                  const uint32_t* step_base, float* ac_out, float* logp_out,
                                                  float* ent_out, float* mu_out, float* x2_out, float* x3_out,
                                                  void* stream) {
  int rc = tc_check("acx_policy_forward_sample_gauss_tc", actor, packed, ob_dim, ac_dim, rows, n, obs);
  if (rc) return rc;
  if (rows == 0 || n == 0) return ACX_OK;
  ACX_REQUIRE(ac_out, "acx_policy_forward_sample_gauss_tc: ac_out is null");
  ACX_REQUIRE(eps || ac_dim <= 2, "acx_policy_forward_sample_gauss_tc: device RNG supports ac_dim <= 2 (got %d)", ac_dim);
  ACX_REQUIRE((x2_out == nullptr) == (x3_out == nullptr), "acx_policy_forward_sample_gauss_tc: give both caches or none");
  Args a; memset(&a, 0, sizeof(a));
  a.params = actor; a.L = acx_actor_layout(ob_dim, ac_dim, 1); a.rows = rows; a.n = n; a.obs = obs;
  a.ac_scale = ac_scale; a.noise = eps; a.seed = seed; a.index0 = index0; a.step0 = step0; a.step_base = step_base;
  a.ac_out_f = ac_out; a.logp_out = logp_out; a.ent_out = ent_out; a.head_out = mu_out;
  a.x2_out = x2_out; a.x3_out = x3_out;
  return launch_tc<128, 64, true, HEAD_GAUSS, MODE_SAMPLE>(a, packed, tc_grid(rows, n), (cudaStream_t)stream);
}

extern "C" int acx_policy_forward_sample_cat_tc(const float* actor, const float* packed, int ob_dim, int n_act,
                                                int rows, int n, const float* obs, const float* u, uint64_t seed,
                                                uint64_t index0, uint32_t step0, const uint32_t* step_base,
                                                int* ac_out, float* logp_out, float* ent_out, float* logits_out,
                                                float* x2_out, float* x3_out, void* stream) {
  int rc = tc_check("acx_policy_forward_sample_cat_tc", actor, packed, ob_dim, n_act, rows, n, obs);
  if (rc) return rc;
  if (rows == 0 || n == 0) return ACX_OK;
  ACX_REQUIRE(ac_out, "acx_policy_forward_sample_cat_tc: ac_out is null");
  ACX_REQUIRE((x2_out == nullptr) == (x3_out == nullptr), "acx_policy_forward_sample_cat_tc: give both caches or none");
  Args a; memset(&a, 0, sizeof(a));
  a.params = actor; a.L = acx_actor_layout(ob_dim, n_act, 0); a.rows = rows; a.n = n; a.obs = obs;
  a.noise = u; a.seed = seed; a.index0 = index0; a.step0 = step0; a.step_base = step_base;
  a.ac_out_i = ac_out; a.logp_out = logp_out; a.ent_out = ent_out; a.head_out = logits_out;
  a.x2_out = x2_out; a.x3_out = x3_out;
  return launch_tc<128, 64, true, HEAD_CAT, MODE_SAMPLE>(a, packed, tc_grid(rows, n), (cudaStream_t)stream);
}

extern "C" int acx_policy_logp_gauss_tc(const float* actor, const float* packed, int ob_dim, int ac_dim,
                                        float ac_scale, int rows, int n, const float* obs, const float* acs,
                                        float* logp_out, float* ent_out, float* mu_out, const float* logp_old,
                                        double* kl_partial, int* kl_blocks, void* stream) {
  int rc = tc_check("acx_policy_logp_gauss_tc", actor, packed, ob_dim, ac_dim, rows, n, obs);
  if (rc) return rc;
  const int grid = (rows == 0 || n == 0) ? 0 : tc_grid(rows, n);
  if (kl_blocks) *kl_blocks = grid;
  if (grid == 0) return ACX_OK;
  ACX_REQUIRE(acs, "acx_policy_logp_gauss_tc: acs is null");
  ACX_REQUIRE((logp_old == nullptr) == (kl_partial == nullptr),
              "acx_policy_logp_gauss_tc: logp_old and kl_partial must be given together");
  Args a; memset(&a, 0, sizeof(a));
  a.params = actor; a.L = acx_actor_layout(ob_dim, ac_dim, 1); a.rows = rows; a.n = n; a.obs = obs;
  a.ac_scale = ac_scale; a.ac_in_f = acs; a.logp_out = logp_out; a.ent_out = ent_out; a.head_out = mu_out;
  a.logp_old = logp_old; a.kl_partial = kl_partial;
  if (tc_two_tiles()) return acx_tc2_launch(HEAD_GAUSS, a, packed, grid, stream);
  return launch_tc<128, 64, true, HEAD_GAUSS, MODE_EVAL>(a, packed, grid, (cudaStream_t)stream);
}

extern "C" int acx_policy_logp_cat_tc(const float* actor, const float* packed, int ob_dim, int n_act, int rows, int n,
                                      const float* obs, const int* acs, float* logp_out, float* ent_out,
                                      float* logits_out, const float* logp_old, double* kl_partial, int* kl_blocks,
                                      void* stream) {
  int rc = tc_check("acx_policy_logp_cat_tc", actor, packed, ob_dim, n_act, rows, n, obs);
  if (rc) return rc;
  const int grid = (rows == 0 || n == 0) ? 0 : tc_grid(rows, n);
  if (kl_blocks) *kl_blocks = grid;
  if (grid == 0) return ACX_OK;
  ACX_REQUIRE(acs, "acx_policy_logp_cat_tc: acs is null");
  ACX_REQUIRE((logp_old == nullptr) == (kl_partial == nullptr),
              "acx_policy_logp_cat_tc: logp_old and kl_partial must be given together");
  Args a; memset(&a, 0, sizeof(a));
  a.params = actor; a.L = acx_actor_layout(ob_dim, n_act, 0); a.rows = rows; a.n = n; a.obs = obs;
  a.ac_in_i = acs; a.logp_out = logp_out; a.ent_out = ent_out; a.head_out = logits_out;
  a.logp_old = logp_old; a.kl_partial = kl_partial;
  if (tc_two_tiles()) return acx_tc2_launch(HEAD_CAT, a, packed, grid, stream);
  return launch_tc<128, 64, true, HEAD_CAT, MODE_EVAL>(a, packed, grid, (cudaStream_t)stream);
}

extern "C" int acx_value_forward_tc(const float* critic, const float* packed, int ob_dim, int rows, int n,
                                    const float* obs, float* v_out, int basic_head, float* x2_out, float* x3_out,
                                    unsigned int* mask_out,
                                    void* stream) {
  int rc = tc_check("acx_value_forward_tc", critic, packed, ob_dim, 1, rows, n, obs);
  if (rc) return rc;
  if (rows == 0 || n == 0) return ACX_OK;
  ACX_REQUIRE(v_out, "acx_value_forward_tc: v_out is null");
  ACX_REQUIRE((x2_out == nullptr) == (x3_out == nullptr), "acx_value_forward_tc: give both caches or none");
  Args a; memset(&a, 0, sizeof(a));
  a.params = critic; a.L = acx_critic_layout(ob_dim); a.rows = rows; a.n = n; a.obs = obs;
  a.v_out = v_out; a.basic_head = basic_head; a.x2_out = x2_out; a.x3_out = x3_out; a.mask_out = mask_out;
  if (tc_two_tiles()) return acx_tc2_launch(HEAD_VALUE, a, packed, tc_grid(rows, n), stream);
  return launch_tc<256, 128, false, HEAD_VALUE, MODE_EVAL>(a, packed, tc_grid(rows, n), (cudaStream_t)stream);
}

extern "C" int acx_rollout_tc(int env, const float* actor, const float* packed, float ac_scale, int T, int n, float* obs,
                              float* state, int* elapsed, const float* noise, const float* reset_u, uint64_t seed,
                              uint64_t index0, const uint32_t* step_base, int max_steps, void* ac_out, float* logp_out,
                              float* ent_out, float* rew_out, unsigned char* done_out, float* x2_out, float* x3_out,
                              unsigned int* mask_out,
                              void* stream) {
  ACX_REQUIRE(env == ACX_ENV_PENDULUM || env == ACX_ENV_CARTPOLE || env == ACX_ENV_ACROBOT,
              "acx_rollout_tc: unknown env id %d", env);
  const int ob_dim = env == ACX_ENV_PENDULUM ? 3 : (env == ACX_ENV_CARTPOLE ? 4 : 6);
  const int cont = env == ACX_ENV_PENDULUM, num_ac = cont ? 1 : (env == ACX_ENV_CARTPOLE ? 2 : 3);
  int rc = tc_check("acx_rollout_tc", actor, packed, ob_dim, num_ac, T, n, obs);
  if (rc) return rc;
  if (T == 0 || n == 0) return ACX_OK;
  ACX_REQUIRE(state && elapsed && ac_out && rew_out && done_out, "acx_rollout_tc: null state / elapsed / ac_out / rew_out / done_out");
  ACX_REQUIRE((x2_out == nullptr) == (x3_out == nullptr), "acx_rollout_tc: give both caches or none");
  Args a; memset(&a, 0, sizeof(a));
  a.params = actor; a.L = acx_actor_layout(ob_dim, num_ac, cont); a.rows = T; a.n = n; a.obs = obs; a.obs_rw = obs;
  a.ac_scale = ac_scale; a.noise = noise; a.seed = seed; a.index0 = index0; a.step0 = 0; a.step_base = step_base;
  a.ac_out_f = cont ? reinterpret_cast<float*>(ac_out) : nullptr;
  a.ac_out_i = cont ? nullptr : reinterpret_cast<int*>(ac_out);
  a.logp_out = logp_out; a.ent_out = ent_out; a.x2_out = x2_out; a.x3_out = x3_out; a.mask_out = mask_out;
  a.env_state = state; a.env_elapsed = elapsed; a.rew_out = rew_out; a.done_out = done_out; a.reset_u = reset_u;
  a.max_steps = max_steps;
  const int grid = tc_grid(1, n);                      // column tiles: a CTA keeps its tiles for all T steps
  cudaStream_t st = (cudaStream_t)stream;
  if (env == ACX_ENV_PENDULUM) return launch_tc<128, 64, true, HEAD_GAUSS, MODE_SAMPLE, ACX_ENV_PENDULUM>(a, packed, grid, st);
  if (env == ACX_ENV_CARTPOLE) return launch_tc<128, 64, true, HEAD_CAT, MODE_SAMPLE, ACX_ENV_CARTPOLE>(a, packed, grid, st);
  return launch_tc<128, 64, true, HEAD_CAT, MODE_SAMPLE, ACX_ENV_ACROBOT>(a, packed, grid, st);
}

#ifdef ACX_FWD_PROF
extern "C" int acx_fwd_prof_read(unsigned long long* out16) {
  return (int)cudaMemcpyFromSymbol(out16, acx_fwd_prof, sizeof(unsigned long long) * 16);
}
#endif
