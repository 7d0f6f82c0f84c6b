// Evaluation-mode forward passes of K2 on the tensor cores with TWO 128-sample tiles in flight per SM:
// acx_value_forward_tc (A3C/policies.py:146-150,176-177) and acx_policy_logp_{gauss,cat}_tc (:60,69,106-108).
// Same math, same operand split (3xTF32), same heads, same cache / mask outputs as acx_mlp_tc_kernel; what changes is
// the schedule.  There, tensor memory holds one tile -- a 128-feature A operand (hi + lo = 256 columns) and two
// accumulators -- so a tile's chain  layer 1 -> MMA -> epilogue -> MMA -> epilogue  is serial and the tensor pipe idles
// while the 256 worker threads run an epilogue, and the workers idle while it multiplies (~30 % of a tile's cycles).
// Here a tile takes HALF of tensor memory:
//     columns 256 t + [0, 64) A hi, [64, 128) A lo   -- the A operand of 64 reduction features at a time
//     columns 256 t + [128, 256) D                    -- ONE accumulator: layer 2's, then (same columns) layer 3's
// and two tiles t = 0, 1 alternate phase by phase on the same workers: while the tensor cores run the 24 MMAs of tile
// 0's round, the workers produce tile 1's operand, and vice versa.  A tile's phases:
//     Q_q  (q < H1 / 64)  layer-1 features 64 q .. 64 q + 63 -> A operand            -> round: D (+)= A . W2[64 q ..]
//     X2a  D -> registers: the thread's 64 layer-2 values (bias, activation; cache + masks), features 0..63 -> A
//                                                                                     -> round: D  = A . W3[0 .. 63]
//     X2b  features 64..127 (kept in 32 registers across the other tile's phase) -> A -> round: D += A . W3[64 ..]
//     X3   D -> layer-3 epilogue (+ cache), head
// Layer 3 may overwrite layer 2's accumulator because X2a reads ALL of it into registers before the first layer-3 MMA is
// issued (the hand-over barrier counts all 256 workers).  The critic's weight chunks stream through the 12-slot ring once
// per PAIR of tiles: a landed chunk serves tile 0's round, then tile 1's, and is released after the second.
#include "acx_mlp.cuh"
#include "acx_umma.cuh"

using namespace acxmlp;
using namespace acxumma;

namespace {

constexpr int T2_EPI = 256;
constexpr int T2_NT = T2_EPI + 64;               // + issuer warp (8) + producer warp (9)
constexpr int T2_SLOT_FLOATS = 4096;             // 16 KB: hi + lo planes of one weight chunk
constexpr int T2_NSLOT = 12;
constexpr int T2_KC2 = 16;                       // layer-2 input rows per chunk (128 outputs)
constexpr uint32_t T2_A_HI = 0, T2_A_LO = 64, T2_D = 128, T2_TILE_COLS = 256;

struct Tc2Smem {
  float slot[T2_NSLOT][T2_SLOT_FLOATS];          // 196,608 B
  float W1[MAX_OB * 256];
  float b1[256];
  float b2[128];
  float b3[128];
  float Wh[128 * MAX_OUT];
  float bh[MAX_OUT];
  float log_std[MAX_OUT];
  float pre[2][MAX_OUT * MT];                    // per tile slot: head partial sums of the second warpgroup
  float4 W1p[256];                               // ob_dim <= 3: {W1[0][k], W1[1][k], W1[2][k], b1[k]}
  double red[8];
  unsigned long long full[T2_NSLOT], empty[T2_NSLOT], a_ready[2], stage_done[2];
  unsigned int tmem_base;
};

__device__ inline void epi_sync2() { asm volatile("bar.sync 1, 256;" ::: "memory"); }     // warps 0-7 only
__device__ inline void mbar_arrive2(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(acx_smem_u32(bar)) : "memory");
}

// 32 values of sample `s` -> A operand columns c0 .. c0 + 31 of the tile's 64-feature buffer, hi and lo
__device__ inline void a2_store32(uint32_t tile_lane, int c0, const float (&v)[32]) {
  float lo[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) lo[i] = lo_trunc_tf32(v[i]);
  tmem_st32(tile_lane + T2_A_HI + (uint32_t)c0, v);
  tmem_st32(tile_lane + T2_A_LO + (uint32_t)c0, lo);
}

template <int H1, int H3, bool LRELU, int HEAD>
__global__ void __launch_bounds__(T2_NT, 1) acx_mlp_tc2_kernel(const Args A, const float* __restrict__ packed) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  Tc2Smem& sm = *reinterpret_cast<Tc2Smem*>(smem_raw);
  constexpr int NQ = H1 / 64;                    // layer-1 quarters (64 features each): 4 critic, 2 actor
  constexpr int KC3 = 2048 / H3;                 // layer-3 rows per 16 KB chunk: 16 (critic) / 32 (actor)
  constexpr int C2Q = 64 / T2_KC2;               // chunks of a layer-2 round (4)
  constexpr int C3H = 64 / KC3;                  // chunks of a layer-3 round: 4 (critic) / 2 (actor)
  constexpr int CPT = NQ * C2Q + 2 * C3H;        // chunks per tile (= per pair): 24 critic, 12 actor
  constexpr bool RESIDENT = CPT == T2_NSLOT;
  const AcxNetLayout L = A.L;
  const int ob = L.ob, nout = L.nout, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const float* __restrict__ P = A.params;

  for (int i = tid; i < ob * H1; i += T2_NT) sm.W1[i] = P[L.W1 + i];
  for (int i = tid; i < H1; i += T2_NT) sm.b1[i] = P[L.b1 + i];
  if (ob <= 3)
    for (int k = tid; k < H1; k += T2_NT)
      sm.W1p[k] = make_float4(P[L.W1 + k], ob > 1 ? P[L.W1 + H1 + k] : 0.f, ob > 2 ? P[L.W1 + 2 * H1 + k] : 0.f, P[L.b1 + k]);
  for (int i = tid; i < H2; i += T2_NT) sm.b2[i] = P[L.b2 + i];
  for (int i = tid; i < H3; i += T2_NT) sm.b3[i] = P[L.b3 + i];
  for (int i = tid; i < H3 * nout; i += T2_NT) sm.Wh[i] = P[L.Wh + i];
  if (tid < nout) {
    sm.bh[tid] = P[L.bh + tid];
    if (HEAD == HEAD_GAUSS) sm.log_std[tid] = P[L.log_std + tid];
  }
  const int tpr = (A.n + MT - 1) / MT;
  const int ntiles = A.rows * tpr;
  const int mine = ((int)blockIdx.x < ntiles) ? (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  const int npairs = (mine + 1) / 2;

  if (warp == 8) tmem_alloc(&sm.tmem_base, 512);
  if (tid == 0) {
    for (int i = 0; i < T2_NSLOT; ++i) {
      acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.full[i]), 1);
      acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.empty[i]), 1);
    }
    for (int i = 0; i < 2; ++i) {
      acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.a_ready[i]), T2_EPI);
      acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.stage_done[i]), 1);
    }
    acx_mbar_fence_init();
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = sm.tmem_base;

  if (warp == 9) {
    // ================================================================ producer of the weight-chunk ring
    if (lane == 0) {
      const unsigned int total = RESIDENT ? (npairs > 0 ? (unsigned)T2_NSLOT : 0u) : (unsigned)npairs * (unsigned)CPT;
      for (unsigned int c = 0; c < total; ++c) {
        const unsigned int q = c % T2_NSLOT;
        if (c >= (unsigned)T2_NSLOT)
          acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.empty[q]), ((c / T2_NSLOT) - 1u) & 1u);
        const float* src = packed + (size_t)(c % (unsigned)CPT) * T2_SLOT_FLOATS;
        acx_mbar_expect_tx(reinterpret_cast<uint64_t*>(&sm.full[q]), T2_SLOT_FLOATS * 4);
        acx_bulk_g2s(sm.slot[q], src, T2_SLOT_FLOATS * 4, reinterpret_cast<uint64_t*>(&sm.full[q]));
      }
    }
  } else if (warp == 8) {
    // ================================================================ issuer
    if (lane == 0) {
      unsigned int ready[2] = {0u, 0u};
      unsigned int g = 0;                         // chunks consumed so far (per pair, not per tile)
      for (int p = 0; p < npairs; ++p) {
        const int nact = (2 * p + 1 < mine) ? 2 : 1;
#pragma unroll 1
        for (int r = 0; r < NQ + 2; ++r) {        // rounds: NQ of layer 2, then the two halves of layer 3
          const bool l3 = r >= NQ;
          const int nc = l3 ? C3H : C2Q, kc = l3 ? KC3 : T2_KC2, n_out = l3 ? H3 : 128;
          const bool first = (r == 0) || (r == NQ);          // a fresh accumulator: layer 2's, then layer 3's
          const uint32_t idesc = idesc_tf32(128, n_out, false, false);
          const uint32_t sbo = (uint32_t)(kc / 4) * 128u;     // next 8 outputs
          for (int t = 0; t < nact; ++t) {
            acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.a_ready[t]), ready[t] & 1u);
            ++ready[t];
            fence_after_sync();
            const uint32_t tb = tmem + (uint32_t)t * T2_TILE_COLS;
            for (int j = 0; j < nc; ++j) {
              const unsigned int gi = g + (unsigned)j, q = gi % T2_NSLOT;
              if (t == 0) {
                acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.full[q]), RESIDENT ? 0u : ((gi / T2_NSLOT) & 1u));
                fence_after_sync();
              }
              const float* b_hi = sm.slot[q];
              const float* b_lo = b_hi + n_out * kc;
              for (int ks = 0; ks < kc / 8; ++ks) {
                const uint64_t bh = smem_desc(b_hi + ks * 64, 128, sbo), bl = smem_desc(b_lo + ks * 64, 128, sbo);
                const uint32_t kcol = (uint32_t)(j * kc + ks * 8);
                const uint32_t ah = tb + T2_A_HI + kcol, al = tb + T2_A_LO + kcol;
                mma_tf32_ts(tb + T2_D, al, bh, idesc, !(first && j == 0 && ks == 0));
                mma_tf32_ts(tb + T2_D, ah, bl, idesc, true);
                mma_tf32_ts(tb + T2_D, ah, bh, idesc, true);
              }
              if (!RESIDENT && t == nact - 1) commit(reinterpret_cast<uint64_t*>(&sm.empty[q]));
            }
            commit(reinterpret_cast<uint64_t*>(&sm.stage_done[t]));
          }
          if (!RESIDENT) g += (unsigned)nc;
          else g = (g + (unsigned)nc) % (unsigned)T2_NSLOT;
        }
      }
    }
  } else {
    // ================================================================ workers (layer 1, epilogues, heads)
    const int s = (warp & 3) * 32 + lane;          // sample = TMEM lane
    const int half = warp >> 2;                    // which 32 of a phase's 64 features this warpgroup handles
    const uint32_t tl = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    const bool basic = (HEAD == HEAD_VALUE) && A.basic_head;
    double kl_sum = 0.0;
    unsigned int done[2] = {0u, 0u};               // completed waits on stage_done, per tile slot
    float o[2][MAX_OB];                            // the sample's observation, per tile slot
    float v1[2][32];                               // layer-2 features 64 + 32 half .. of the sample, between X2a and X2b
    float pre[2][MAX_OUT];                         // head partial sums
    int tile[2];
    bool act[2], live[2];

    auto load_obs = [&](int tile_i, float (&ov)[MAX_OB]) {
      const int row = tile_i / tpr, c = (tile_i % tpr) * MT + s;
#pragma unroll
      for (int i = 0; i < MAX_OB; ++i)
        ov[i] = (i < ob && tile_i < ntiles && c < A.n) ? __ldg(A.obs + ((size_t)row * ob + i) * A.n + c) : 0.f;
    };
    auto wait_done = [&](int t) {
      acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.stage_done[t]), done[t] & 1u);
      ++done[t];
      fence_after_sync();
    };
    auto hand_over = [&](int t) {
      tmem_wait_st();
      fence_before_sync();
      mbar_arrive2(&sm.a_ready[t]);
    };
#pragma unroll
    for (int t = 0; t < 2; ++t) load_obs((int)blockIdx.x + t * (int)gridDim.x, o[t]);

    for (int p = 0; p < npairs; ++p) {
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        tile[t] = (int)blockIdx.x + (2 * p + t) * (int)gridDim.x;
        act[t] = 2 * p + t < mine;
        live[t] = act[t] && s < min(MT, A.n - (tile[t] % tpr) * MT);
#pragma unroll
        for (int a = 0; a < MAX_OUT; ++a) pre[t][a] = 0.f;
      }
      // ---------------------------------------------- layer 1, a quarter (64 features) per round
#pragma unroll 1
      for (int q = 0; q < NQ; ++q) {
#pragma unroll
        for (int t = 0; t < 2; ++t) {
          if (!act[t]) continue;
          const uint32_t tt = tl + (uint32_t)t * T2_TILE_COLS;
          float v[32];
          const int j0 = q * 64 + half * 32;
          if (ob <= 3) {
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              const float4 w = sm.W1p[j0 + e];
              float acc = fmaf(o[t][0], w.x, 0.f);
              if (ob > 1) acc = fmaf(o[t][1], w.y, acc);
              if (ob > 2) acc = fmaf(o[t][2], w.z, acc);
              v[e] = act_fn<LRELU>(__fadd_rn(acc, w.w));
            }
          } else {
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              const int k = j0 + e;
              float acc = 0.f;
#pragma unroll
              for (int i = 0; i < MAX_OB; ++i)
                if (i < ob) acc = fmaf(o[t][i], sm.W1[i * H1 + k], acc);
              v[e] = act_fn<LRELU>(__fadd_rn(acc, sm.b1[k]));
            }
          }
          if (q > 0) wait_done(t);                 // the MMAs reading the previous quarter of this tile are done
          a2_store32(tt, half * 32, v);
          hand_over(t);
          if (A.mask_out != nullptr) {             // act'(X1) for the chain kernel (after the hand-over)
            unsigned int bits = 0u;
#pragma unroll
            for (int e = 0; e < 32; ++e) bits |= (v[e] > 0.f ? 1u : 0u) << e;
            A.mask_out[mask_tile_off(tile[t], H1) + (size_t)(q * 2 + half) * MT + s] = bits;
          }
        }
      }
      // ---------------------------------------------- X2a: the whole layer-2 accumulator -> registers, first half -> A
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        if (!act[t]) continue;
        const uint32_t tt = tl + (uint32_t)t * T2_TILE_COLS;
        wait_done(t);
        float v0[32];
        tmem_ld32(tt + T2_D + (uint32_t)(half * 32), v0);
        tmem_ld32(tt + T2_D + (uint32_t)(64 + half * 32), v1[t]);
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          v0[e] = act_fn<LRELU>(__fadd_rn(v0[e], sm.b2[half * 32 + e]));
          v1[t][e] = act_fn<LRELU>(__fadd_rn(v1[t][e], sm.b2[64 + half * 32 + e]));
        }
        a2_store32(tt, half * 32, v0);
        hand_over(t);                              // all 256 workers have read D: layer 3 may overwrite it
        if (A.mask_out != nullptr) {               // act'(X2)
          unsigned int b0 = 0u, b1 = 0u;
#pragma unroll
          for (int e = 0; e < 32; ++e) {
            b0 |= (v0[e] > 0.f ? 1u : 0u) << e;
            b1 |= (v1[t][e] > 0.f ? 1u : 0u) << e;
          }
          unsigned int* mo = A.mask_out + mask_tile_off(tile[t], H1) + (size_t)(H1 / 32) * MT + s;
          mo[(size_t)half * MT] = b0;
          mo[(size_t)(2 + half) * MT] = b1;
        }
        if (A.x2_out != nullptr) {                 // tile-blocked cache; zeros beyond the batch
          float* g2 = A.x2_out + cache_tile_off(tile[t], H2) + (size_t)(half * 32) * MT + s;
#pragma unroll
          for (int e = 0; e < 32; ++e) g2[e * MT] = live[t] ? v0[e] : 0.f;
          g2 += (size_t)64 * MT;
#pragma unroll
          for (int e = 0; e < 32; ++e) g2[e * MT] = live[t] ? v1[t][e] : 0.f;
        }
        if (basic) {
#pragma unroll
          for (int e = 0; e < 32; ++e) pre[t][0] = fmaf(v0[e], sm.Wh[half * 32 + e], pre[t][0]);
#pragma unroll
          for (int e = 0; e < 32; ++e) pre[t][0] = fmaf(v1[t][e], sm.Wh[64 + half * 32 + e], pre[t][0]);
        }
      }
      // ---------------------------------------------- X2b: second half -> A
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        if (!act[t]) continue;
        wait_done(t);                              // the MMAs over the first half are done: the A buffer is free
        a2_store32(tl + (uint32_t)t * T2_TILE_COLS, half * 32, v1[t]);
        hand_over(t);
      }
      // ---------------------------------------------- X3: layer-3 epilogue + head
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        if (!act[t]) continue;
        const uint32_t tt = tl + (uint32_t)t * T2_TILE_COLS;
        const int row = tile[t] / tpr;
        const size_t col = (size_t)(tile[t] % tpr) * MT + s;
        load_obs(tile[t] + 2 * (int)gridDim.x, o[t]);          // the next pair's observation of this slot
        wait_done(t);
#pragma unroll 1
        for (int gq = 0; gq < H3 / 64; ++gq) {     // this thread: H3/2 features
          float v[32];
          const int j0 = half * (H3 / 2) + gq * 32;
          tmem_ld32(tt + T2_D + (uint32_t)j0, v);
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] = act_fn<LRELU>(__fadd_rn(v[e], sm.b3[j0 + e]));
          if (A.x3_out != nullptr) {
            float* g3 = A.x3_out + cache_tile_off(tile[t], H3) + (size_t)j0 * MT + s;
#pragma unroll
            for (int e = 0; e < 32; ++e) g3[e * MT] = live[t] ? v[e] : 0.f;
          }
          if (!basic) {
#pragma unroll
            for (int e = 0; e < 32; ++e) {
#pragma unroll
              for (int a = 0; a < MAX_OUT; ++a)
                if (a < nout) pre[t][a] = fmaf(v[e], sm.Wh[(j0 + e) * nout + a], pre[t][a]);
            }
          }
        }
        fence_before_sync();                       // this tile's tcgen05.ld precede its next hand-over
        if (half == 1) {
#pragma unroll
          for (int a = 0; a < MAX_OUT; ++a)
            if (a < nout) sm.pre[t][a * MT + s] = pre[t][a];
        }
        epi_sync2();
        if (half == 0 && live[t]) {
          float hp[MAX_OUT];
#pragma unroll
          for (int a = 0; a < MAX_OUT; ++a)
            hp[a] = a < nout ? __fadd_rn(__fadd_rn(pre[t][a], sm.pre[t][a * MT + s]), sm.bh[a]) : 0.f;
          const size_t s_idx = (size_t)row * A.n + col;
          if (HEAD == HEAD_VALUE) {
            A.v_out[s_idx] = hp[0];
          } else if (HEAD == HEAD_GAUSS) {
            float lp = 0.f, ent = 0.f;
#pragma unroll
            for (int d = 0; d < MAX_OUT; ++d)
              if (d < nout) {
                const float mu = __fmul_rn(tanhf(hp[d]), A.ac_scale);
                const float ls = fminf(fmaxf(sm.log_std[d], ACX_LOG_STD_MIN), ACX_LOG_STD_MAX);
                const float sd = expf(ls);
                const size_t a_idx = ((size_t)row * nout + d) * A.n + col;
                const float ac = A.ac_in_f[a_idx];
                if (A.head_out) A.head_out[a_idx] = mu;
                const float z = __fdiv_rn(__fsub_rn(ac, mu), sd);
                lp += __fsub_rn(__fsub_rn(__fmul_rn(__fmul_rn(-0.5f, z), z), ls), ACX_HALF_LOG_2PI);
                ent += __fadd_rn(__fadd_rn(0.5f, ACX_HALF_LOG_2PI), ls);
              }
            if (A.logp_out) A.logp_out[s_idx] = lp;
            if (A.ent_out) A.ent_out[s_idx] = ent;
            if (A.logp_old) {
              const double df = (double)A.logp_old[s_idx] - (double)lp;
              kl_sum += df * df;
            }
          } else {   // HEAD_CAT
            float l1[MAX_OUT], logps[MAX_OUT], pr[MAX_OUT];
            float mx1 = -INFINITY, mx2 = -INFINITY;
#pragma unroll
            for (int a = 0; a < MAX_OUT; ++a)
              if (a < nout) {
                l1[a] = __fadd_rn(hp[a], 1e-8f);
                mx1 = fmaxf(mx1, l1[a]); mx2 = fmaxf(mx2, hp[a]);
              }
            float se1 = 0.f, se2 = 0.f;
#pragma unroll
            for (int a = 0; a < MAX_OUT; ++a)
              if (a < nout) {
                se1 += expf(l1[a] - mx1);
                pr[a] = expf(hp[a] - mx2);
                se2 += pr[a];
              }
            const float lse = mx1 + logf(se1);
            float ent = 0.f;
#pragma unroll
            for (int a = 0; a < MAX_OUT; ++a)
              if (a < nout) {
                logps[a] = l1[a] - lse;
                pr[a] = pr[a] / se2;
                ent -= pr[a] * logps[a];
              }
            const int ac = A.ac_in_i[s_idx];
            float lp = 0.f;
#pragma unroll
            for (int a = 0; a < MAX_OUT; ++a)
              if (a < nout && a == ac) lp = logps[a];
            if (A.head_out) {
#pragma unroll
              for (int a = 0; a < MAX_OUT; ++a)
                if (a < nout) A.head_out[((size_t)row * nout + a) * A.n + col] = hp[a];
            }
            if (A.logp_out) A.logp_out[s_idx] = lp;
            if (A.ent_out) A.ent_out[s_idx] = ent;
            if (A.logp_old) {
              const double df = (double)A.logp_old[s_idx] - (double)lp;
              kl_sum += df * df;
            }
          }
        }
      }
    }
    if (HEAD != HEAD_VALUE && A.kl_partial != nullptr) {
      const double v = acx_warp_sum_d(kl_sum);
      if (lane == 0) sm.red[warp] = v;
    }
  }

  // ---------------------------------------------------- teardown
  fence_before_sync();
  __syncthreads();
  if (warp == 8) tmem_dealloc(tmem, 512);
  if (HEAD != HEAD_VALUE && A.kl_partial != nullptr && tid == 0)
    A.kl_partial[blockIdx.x] = ((sm.red[0] + sm.red[1]) + (sm.red[2] + sm.red[3])) +
                               ((sm.red[4] + sm.red[5]) + (sm.red[6] + sm.red[7]));
}

template <int H1, int H3, bool LRELU, int HEAD>
int launch_tc2_t(const Args& a, const float* packed, int grid, cudaStream_t st) {
  auto kern = acx_mlp_tc2_kernel<H1, H3, LRELU, HEAD>;
  ACX_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Tc2Smem) + 1024));
  kern<<<grid, T2_NT, sizeof(Tc2Smem) + 1024, st>>>(a, packed);
  ACX_LAUNCH_OK("acx_mlp_tc2_kernel");
  return ACX_OK;
}

}  // namespace

// head: HEAD_VALUE (critic 256-128-128), HEAD_GAUSS / HEAD_CAT (actor 128-128-64, log-prob of stored actions)
int acx_tc2_launch(int head, const Args& a, const float* packed, int grid, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (head == HEAD_VALUE) return launch_tc2_t<256, 128, false, HEAD_VALUE>(a, packed, grid, st);
  if (head == HEAD_GAUSS) return launch_tc2_t<128, 64, true, HEAD_GAUSS>(a, packed, grid, st);
  return launch_tc2_t<128, 64, true, HEAD_CAT>(a, packed, grid, st);
}
