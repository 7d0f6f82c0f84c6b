// K4 with the back-propagation chain on the tensor cores (tcgen05, 3xTF32) and the weight-gradient GEMMs on
// the CUDA cores, concurrently.  Same contract and reference lines as acx_loss_backward_{gauss,cat}_cached
// (A3C/policies.py:74-84,152-158); needs the activation cache (layers 2 and 3 are never recomputed).
//
// Per 128-sample tile, one persistent CTA per SM, warps 0-7 = workers, warp 8 = issuer:
//   workers  TMA-load the cached X2 / X3 tiles into shared memory ([feature][sample], fp32), run the head
//            (loss terms, seed gradients), form dPre3 in place AND as the hi/lo A operand in tensor memory
//            (thread = sample = TMEM lane), then -- while the issuer's MMAs run --
//              dW3 += X2^T dPre3            (G stage, 8 x 8 register tiles, FFMA)
//            read dX2 from TMEM, dPre2 = dX2 * act'(X2) in place + as the next A operand,
//              dW2 += X1^T dPre2            (X1 recomputed per 128-row block)
//            read dX1 from TMEM, dPre1 = dX1 * act'(X1), dW1 / db1.
//   issuer   dX2[s][k] = sum_j dPre3[s][j] W3[k][j]   and   dX1[s][k] = sum_j dPre2[s][j] W2[k][j]:
//            A from TMEM, B = the weight rows themselves (K-major in j), pre-split and pre-arranged by
//            acx_tc_pack_weights_bwd, streamed through a 4 x 16 KB ring (bulk-TMA), accumulators in TMEM.
// Gradients are accumulated in the per-CTA scratch rows and reduced in fixed order exactly as in acx_mlp.cu.
//
// This one-kernel form serves callers without room for the dPre streams (or with n % 4 != 0).  With the larger
// workspace the entry points below run the two-kernel form instead: acx_mlp_tc_chain_kernel (acx_mlp_tc_chain.cu: the
// chain only, dPre3 / dPre2 streamed to HBM) + acx_wgrad_kernel (acx_wgrad.cu: every weight-gradient product on the
// tensor cores).
#include "acx_mlp.cuh"
#include "acx_head.cuh"
#include "acx_umma.cuh"
#include "acx_wgrad.cuh"
#include "acx_tc_bwd.cuh"

using namespace acxmlp;
using namespace acxumma;
using namespace acxtcbwd;

namespace {

struct TtSmem {
  float X2[H2 * MS];
  float X3[128 * MS];          // X3 / dPre3, later the recomputed X1 block / dPre1 block
  float slot[TT_NSLOT][TT_SLOT_FLOATS];
  static constexpr int OBW = MAX_OB;
  float obs[MAX_OB * MT];
  float W1[MAX_OB * 256];
  float b1[256];
  float Wh[128 * MAX_OUT];
  float bh[MAX_OUT];
  float log_std[MAX_OUT];
  float dHead[MAX_OUT * MT];
  double red[8][4];
  unsigned long long full[TT_NSLOT], empty[TT_NSLOT], a_ready, stage_done[3], bar_act;
  unsigned int tmem_base;
};

template <int H1, int H3, bool LRELU, int HEAD>
__global__ void __launch_bounds__(TT_NT, 1) acx_mlp_tc_train_kernel(const Args A, const float* __restrict__ packed) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  TtSmem& sm = *reinterpret_cast<TtSmem*>(smem_raw);
  constexpr int TJ3 = H3 / 16, NB = H1 / 128, N3C = H3 / 16;
  const AcxNetLayout L = A.L;
  const int ob = L.ob, nout = L.nout, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const float* __restrict__ P = A.params;

  for (int i = tid; i < ob * H1; i += TT_NT) sm.W1[i] = P[L.W1 + i];
  for (int i = tid; i < H1; i += TT_NT) sm.b1[i] = P[L.b1 + i];
  for (int i = tid; i < H3 * nout; i += TT_NT) sm.Wh[i] = P[L.Wh + i];
  if (tid < nout) {
    sm.bh[tid] = P[L.bh + tid];
    if (HEAD == HEAD_GAUSS) sm.log_std[tid] = P[L.log_std + tid];
  }
  const int tpr = (A.n + MT - 1) / MT;
  const int ntiles = A.rows * tpr;
  const unsigned int mine = ((int)blockIdx.x < ntiles) ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

  if (warp == 8) tmem_alloc(&sm.tmem_base, 512);
  if (tid == 0) {
    for (int i = 0; i < TT_NSLOT; ++i) {
      acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.full[i]), 1);
      acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.empty[i]), 1);
    }
    acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.a_ready), TT_WORK);
    for (int i = 0; i < 3; ++i) acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.stage_done[i]), 1);
    acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.bar_act), 1);
    acx_mbar_fence_init();
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = sm.tmem_base;

  TtPipe<TtSmem> pipe;
  pipe.sm = &sm; pipe.packed = packed; pipe.g = 0;
  pipe.chunks_per_tile = N3C + 8 * NB;
  pipe.total = mine * pipe.chunks_per_tile;
  if (warp == 9) {
    // ================================================================ producer of the weight-chunk ring
    if (lane == 0) pipe.produce();
  } else if (warp == 8) {
    // ================================================================ issuer
    if (lane == 0) {
      unsigned int ready = 0;
      for (unsigned int t = 0; t < mine; ++t) {
        acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.a_ready), ready & 1u); ++ready;       // dPre3 is the A operand
        fence_after_sync();
        for (int cc = 0; cc < N3C; ++cc) pipe.mma_chunk(tmem, cc * 16, tmem + TT_D0, cc == 0);
        commit(reinterpret_cast<uint64_t*>(&sm.stage_done[0]));
        acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.a_ready), ready & 1u); ++ready;       // dPre2 is the A operand
        fence_after_sync();
        for (int h = 0; h < NB; ++h) {
          for (int cc = 0; cc < 8; ++cc) pipe.mma_chunk(tmem, cc * 16, tmem + (h == 0 ? TT_D0 : TT_D1), cc == 0);
          commit(reinterpret_cast<uint64_t*>(&sm.stage_done[1 + h]));
        }
      }
    }
  } else {
    // ================================================================ workers
    const Coord c;
    const int s = (warp & 3) * 32 + lane;          // sample = TMEM lane
    const int half = warp >> 2;
    const uint32_t tl = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    HeadScalars hs;
    hs.beta = hs.gam = 0.f; hs.adv_mean = 0.f; hs.adv_inv = 1.f; hs.step0 = 0u;
    hs.invB = 1.0 / ((double)A.rows * (double)A.n);
    if (HEAD != HEAD_VALUE) {
      hs.beta = A.opt[ACX_OPT_BETA_KL]; hs.gam = A.opt[ACX_OPT_GAMMA_ENT];
      if (A.adv_stats) { hs.adv_mean = A.adv_stats[0]; hs.adv_inv = A.adv_stats[1]; }
    }
    double lsum[4] = {0.0, 0.0, 0.0, 0.0};
    float dls[MAX_OUT] = {0.f, 0.f, 0.f, 0.f};
    float* __restrict__ scr = A.scratch + (size_t)blockIdx.x * L.size;
    bool first = true;
    unsigned int act_phase = 0, tiles_done = 0;
    if (tid < L.size - L.bh) scr[L.bh + tid] = 0.f;
    const bool basic = (HEAD == HEAD_VALUE) && A.basic_head;
    const bool vec_ok = (A.n & 3) == 0;

    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++tiles_done) {
      const int row = tile / tpr, n0 = (tile % tpr) * MT;
      const int valid = min(MT, A.n - n0);
      const unsigned int par = tiles_done & 1u;
      const bool bulk = valid == MT;                 // full tile: bulk-TMA row copies (the cache is tile-blocked)
      work_sync();                                   // previous tile is done with every shared buffer
      for (int idx = tid; idx < ob * MT; idx += TT_WORK) {
        const int i = idx / MT, m = idx % MT;
        sm.obs[idx] = (m < valid) ? A.obs[((size_t)row * ob + i) * A.n + n0 + m] : 0.f;
      }
      {
        const float* g2 = A.x2_in + cache_tile_off(tile, H2);
        const float* g3 = A.x3_in + cache_tile_off(tile, H3);
        if (bulk) {
          uint64_t* bar = reinterpret_cast<uint64_t*>(&sm.bar_act);
          if (tid == 0) acx_mbar_expect_tx(bar, (uint32_t)((H2 + H3) * MT * 4));
          work_sync();
          if (tid < H2) acx_bulk_g2s(sm.X2 + tid * MS, g2 + (size_t)tid * MT, MT * 4, bar);
          else if (tid - H2 < H3) acx_bulk_g2s(sm.X3 + (tid - H2) * MS, g3 + (size_t)(tid - H2) * MT, MT * 4, bar);
          acx_mbar_wait(bar, act_phase & 1u);
          ++act_phase;
        } else {
          load_tile(sm.X2, g2, (size_t)MT, H2, valid, true);
          load_tile(sm.X3, g3, (size_t)MT, H3, valid, true);
        }
      }
      work_sync();

      // ---------------------------------------------- head
      const float* Xh = basic ? sm.X2 : sm.X3;
      if (tid < MT) {
        float pre[MAX_OUT];
        head_pre<H3>(sm, Xh, tid, nout, pre);
        float dh[MAX_OUT] = {0.f, 0.f, 0.f, 0.f};
        head_apply<HEAD, MODE_TRAIN>(A, sm.log_std, nout, pre, tid < valid, row, (size_t)n0 + tid, hs, lsum, dls, dh);
#pragma unroll
        for (int a = 0; a < MAX_OUT; ++a)
          if (a < nout) sm.dHead[a * MT + tid] = dh[a];
      }
      work_sync();
      {   // head weights and bias
        if (tid < H3) {
          for (int a = 0; a < nout; ++a) {
            float acc = 0.f;
            for (int m4 = 0; m4 < MT / 4; ++m4) {
              const float4 x = *reinterpret_cast<const float4*>(&Xh[tid * MS + 4 * m4]);
              const float4 d = *reinterpret_cast<const float4*>(&sm.dHead[a * MT + 4 * m4]);
              acc = fmaf(x.x, d.x, acc); acc = fmaf(x.y, d.y, acc); acc = fmaf(x.z, d.z, acc); acc = fmaf(x.w, d.w, acc);
            }
            float* p = scr + L.Wh + tid * nout + a;
            *p = first ? acc : *p + acc;
          }
        } else if (tid >= 128 && tid < 128 + nout) {
          const int a = tid - 128;
          float acc = 0.f;
          for (int m = 0; m < MT; ++m) acc += sm.dHead[a * MT + m];
          float* p = scr + L.bh + a;
          *p = first ? acc : *p + acc;
        }
      }
      work_sync();
      // ---------------------------------------------- dPre3 (in place, and as the A operand), column-wise
#pragma unroll 1
      for (int gq = 0; gq < H3 / 64; ++gq) {
        float v[32];
        const int j0 = half * (H3 / 2) + gq * 32;
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const int j = j0 + e;
          float acc = 0.f;
          for (int a = 0; a < nout; ++a) acc = fmaf(sm.Wh[j * nout + a], sm.dHead[a * MT + s], acc);
          float* x = &sm.X3[j * MS + s];
          const float d = basic ? 0.f : acc * act_grad<LRELU>(*x);
          *x = d;
          v[e] = d;
        }
        a_store32_tt(tl, j0, v);
      }
      tmem_wait_st();
      fence_before_sync();
      mbar_arrive1(&sm.a_ready);
      work_sync();
      // ---------------------------------------------- dW3, db3 on the CUDA cores while dX2 runs on the tensor cores
      g_stage<8, TJ3>(sm.X2, sm.X3, scr + L.W3, H3, first, c);
      bias_grad(sm.X3, H3, scr + L.b3, first);
      work_sync();
      // ---------------------------------------------- dPre2 = dX2 * act'(X2): in place + next A operand
      acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.stage_done[0]), par);
      fence_after_sync();
#pragma unroll 1
      for (int gq = 0; gq < 2; ++gq) {
        float v[32];
        const int k0 = half * 64 + gq * 32;
        tmem_ld32(tl + TT_D0 + (uint32_t)k0, v);
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          float* x = &sm.X2[(k0 + e) * MS + s];
          float d = v[e];
          if (basic) d = fmaf(sm.Wh[k0 + e], sm.dHead[s], d);
          d *= act_grad<LRELU>(*x);
          *x = d;
          v[e] = d;
        }
        a_store32_tt(tl, k0, v);
      }
      tmem_wait_st();
      fence_before_sync();
      mbar_arrive1(&sm.a_ready);
      work_sync();
      bias_grad(sm.X2, H2, scr + L.b2, first);
      // ---------------------------------------------- layers 1/2: dW2 on the CUDA cores, dX1 on the tensor cores
      for (int h = 0; h < NB; ++h) {
        layer1_block<H1, LRELU>(sm, h, ob, sm.X3);
        work_sync();
        g_stage<8, 8>(sm.X3, sm.X2, scr + L.W2 + (size_t)h * 128 * H2, H2, first, c);
        work_sync();
        acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.stage_done[1 + h]), par);
        fence_after_sync();
#pragma unroll 1
        for (int gq = 0; gq < 2; ++gq) {
          float v[32];
          const int k0 = half * 64 + gq * 32;
          tmem_ld32(tl + (h == 0 ? TT_D0 : TT_D1) + (uint32_t)k0, v);
#pragma unroll
          for (int e = 0; e < 32; ++e) {
            float* x = &sm.X3[(k0 + e) * MS + s];
            *x = v[e] * act_grad<LRELU>(*x);
          }
        }
        fence_before_sync();
        work_sync();
        layer1_grad<H1>(sm, h, ob, sm.X3, scr + L.W1, scr + L.b1, first);
        work_sync();
      }
      first = false;
    }

    // ---------------------------------------------------- per-CTA scalar sums (head threads = warps 0..3)
    work_sync();
    double v4[4] = {lsum[0], lsum[1], lsum[2], 0.0};
#pragma unroll
    for (int q = 0; q < 4; ++q) v4[q] = acx_warp_sum_d(v4[q]);
    float dl[MAX_OUT];
#pragma unroll
    for (int d = 0; d < MAX_OUT; ++d) dl[d] = acx_warp_sum(dls[d]);
    if (lane == 0 && warp < 4) {
#pragma unroll
      for (int q = 0; q < 4; ++q) sm.red[warp][q] = v4[q];
#pragma unroll
      for (int d = 0; d < MAX_OUT; ++d) sm.dHead[warp * MAX_OUT + d] = dl[d];
    }
    work_sync();
    if (tid == 0) {
#pragma unroll
      for (int q = 0; q < 4; ++q)
        A.loss_partial[(size_t)blockIdx.x * 4 + q] = (sm.red[0][q] + sm.red[1][q]) + (sm.red[2][q] + sm.red[3][q]);
      if (HEAD == HEAD_GAUSS)
        for (int d = 0; d < nout; ++d)
          scr[L.log_std + d] = (sm.dHead[d] + sm.dHead[MAX_OUT + d]) + (sm.dHead[2 * MAX_OUT + d] + sm.dHead[3 * MAX_OUT + d]);
    }
  }

  fence_before_sync();
  __syncthreads();
  if (warp == 8) tmem_dealloc(tmem, 512);
}

// packed backward operands, 16 KB chunks of 16 reduction rows (j) x 128 outputs (k), hi plane then lo plane,
// each plane [k/8][4][8][4]:  chunks 0 .. H3/16-1: W3[k][j] ; then per 128-row block h of layer 1: W2[128h + k][j]
__global__ void __launch_bounds__(256) acx_tc_pack_bwd_kernel(const float* __restrict__ P, AcxNetLayout L,
                                                              float* __restrict__ packed) {
  const int n3 = 128 * L.H3, n2 = L.H1 * 128;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n3 + n2; i += gridDim.x * blockDim.x) {
    const bool l3 = i < n3;
    const int e = l3 ? i : i - n3;
    int k, j, chunk;
    float w;
    if (l3) { k = e / L.H3; j = e % L.H3; w = P[L.W3 + e]; chunk = j / 16; }
    else { const int kg = e / 128; j = e % 128; w = P[L.W2 + e]; k = kg & 127; chunk = L.H3 / 16 + (kg >> 7) * 8 + j / 16; }
    float hi, lo;
    split_tf32(w, hi, lo);
    const int jj = j & 15;
    const int off = ((k >> 3) * 4 + (jj >> 2)) * 32 + (k & 7) * 4 + (jj & 3);
    packed[(size_t)chunk * TT_SLOT_FLOATS + off] = hi;
    packed[(size_t)chunk * TT_SLOT_FLOATS + 2048 + off] = lo;
  }
}

__global__ void __launch_bounds__(256) acx_grad_reduce_tc_kernel(const float* __restrict__ scratch, int ncta, int size,
                                                                 float* __restrict__ grad,
                                                                 const double* __restrict__ loss_partial,
                                                                 float* __restrict__ stats, const float* __restrict__ opt,
                                                                 int is_actor, double count, int lane_major, int offW3,
                                                                 int H3, int offW2, int H1) {
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < size; p += gridDim.x * blockDim.x) {
    // lane_major: acx_wgrad_kernel keeps its dW3 / dW2 partials as it reads them from tensor memory (row r of a 128-row
    // block = lane, column c): element (r, c) at ((c / 4) * 128 + r) * 4 + c % 4 of the region, so that its flushes are
    // 512-byte coalesced
    int q = p;
    if (lane_major && p >= offW3 && p < offW3 + 128 * H3) {
      const int r = (p - offW3) / H3, c = (p - offW3) % H3;
      q = offW3 + ((c >> 2) * 128 + r) * 4 + (c & 3);
    } else if (lane_major && p >= offW2 && p < offW2 + H1 * 128) {
      // H1 = 128: lane = k, column = j; H1 = 256: the kernel accumulates dW2 transposed (lane = j, column = k)
      const int k1 = (p - offW2) / 128, j = (p - offW2) % 128;
      const int r = H1 == 256 ? j : k1, c = H1 == 256 ? k1 : j;
      q = offW2 + ((c >> 2) * 128 + r) * 4 + (c & 3);
    }
    // fixed order cta = 0, 1, ...; the loads of sixteen rows are issued ahead of the dependent adds
    float acc = 0.f;
    int cta = 0;
    for (; cta + 16 <= ncta; cta += 16) {
      float x[16];
#pragma unroll
      for (int u = 0; u < 16; ++u) x[u] = scratch[(size_t)(cta + u) * size + q];
#pragma unroll
      for (int u = 0; u < 16; ++u) acc += x[u];
    }
    for (; cta < ncta; ++cta) acc += scratch[(size_t)cta * size + q];
    grad[p] = acc;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && stats != nullptr) {
    double s0 = 0, s1 = 0, s2 = 0;
    for (int cta = 0; cta < ncta; ++cta) {
      s0 += loss_partial[4 * cta]; s1 += loss_partial[4 * cta + 1]; s2 += loss_partial[4 * cta + 2];
    }
    if (is_actor) {
      const double rew_loss = -s0 / count, kl = s1 / count, ent = s2 / count;
      stats[ACX_STAT_ACTOR_LOSS] = (float)(rew_loss + (double)opt[ACX_OPT_BETA_KL] * kl - (double)opt[ACX_OPT_GAMMA_ENT] * ent);
      stats[ACX_STAT_KL_PRE] = (float)kl;
      stats[ACX_STAT_ENT_MEAN] = (float)ent;
      stats[ACX_STAT_REW_LOSS] = (float)rew_loss;
    } else {
      stats[ACX_STAT_CRITIC_LOSS] = (float)(s0 / count);
    }
  }
}

// split == nullptr: everything in one kernel (weight gradients FFMA); else chain kernel + acx_wgrad_kernel
template <int H1, int H3, bool LRELU, int HEAD>
int launch_tt(const Args& a, const float* packed, float* split, int grid, cudaStream_t st) {
  if (split == nullptr) {
    auto kern = acx_mlp_tc_train_kernel<H1, H3, LRELU, HEAD>;
    ACX_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TtSmem) + 1024));
    kern<<<grid, TT_NT, sizeof(TtSmem) + 1024, st>>>(a, packed);
    ACX_LAUNCH_OK("acx_mlp_tc_train_kernel");
    return ACX_OK;
  }
  float* d3 = split;                                   // tile-blocked like the activation cache
  float* d2 = d3 + cache_floats(a.rows, H3, a.n);
  int rc = acx_tc_chain_launch(a, packed, d3, d2, HEAD, grid, st);
  if (rc) return rc;
  AcxWgradArgs w;
  w.x2 = a.x2_in; w.d3 = d3; w.d2 = d2; w.obs = a.obs; w.params = a.params; w.ob = a.L.ob; w.offW1 = a.L.W1;
  w.offb1 = a.L.b1; w.lrelu = LRELU ? 1 : 0; w.rows = a.rows; w.n = a.n; w.H1 = H1; w.H3 = H3;
  w.scratch = a.scratch; w.scr_stride = a.L.size; w.offW3 = a.L.W3; w.offW2 = a.L.W2; w.offb3 = a.L.b3;
  w.offb2 = a.L.b2; w.grid = grid;
  return acx_wgrad_launch(w, st);
}

int loss_backward_tc(int cont, const float* actor, const float* critic, const float* actor_packed,
                     const float* critic_packed, int ob_dim, int num_ac, float ac_scale, int rows, int n,
                     const float* obs, const void* acs, const float* adv, const float* adv_stats,
                     const float* logp_old, const float* targets, const float* opt, float* grad_actor,
                     float* grad_critic, float* stats, const float* ax2, const float* ax3, const float* cx2,
                     const float* cx3, const unsigned int* amask, const unsigned int* cmask, void* workspace,
                     size_t workspace_bytes, cudaStream_t st) {
  const char* fn = cont ? "acx_loss_backward_gauss_tc" : "acx_loss_backward_cat_tc";
  ACX_REQUIRE(actor || critic, "%s: neither actor nor critic given", fn);
  ACX_REQUIRE(rows >= 1 && n >= 1, "%s: empty batch (rows=%d, n=%d): the mean loss is undefined", fn, rows, n);
  ACX_REQUIRE(ob_dim >= 1 && ob_dim <= MAX_OB && num_ac >= 1 && num_ac <= MAX_OUT, "%s: bad ob_dim/num_ac", fn);
  ACX_REQUIRE(workspace && workspace_bytes >= acx_loss_backward_workspace_bytes(ob_dim, num_ac, cont),
              "%s: workspace too small (%zu < %zu)", fn, workspace_bytes,
              acx_loss_backward_workspace_bytes(ob_dim, num_ac, cont));
  ACX_REQUIRE(acx_aligned16(workspace) && obs, "%s: workspace must be 16-byte aligned, obs non-null", fn);
  const int sms = acx_num_sms();
  const long long tiles = (long long)rows * ((n + MT - 1) / MT);
  const int grid = (int)(tiles < sms ? tiles : sms);
  const AcxNetLayout La = acx_actor_layout(ob_dim, num_ac, cont), Lc = acx_critic_layout(ob_dim);
  unsigned char* wsb = reinterpret_cast<unsigned char*>(workspace);
  double* lossp_a = reinterpret_cast<double*>(wsb);
  double* lossp_c = lossp_a + (size_t)sms * 4;
  float* wT_a = reinterpret_cast<float*>(wsb + 2 * (size_t)sms * 4 * sizeof(double) + 128);     // same carve-up as acx_mlp.cu
  float* scr_a = wT_a + (size_t)(ACX_A_H3 + ACX_A_H1 + ACX_C_H3 + ACX_C_H1) * 128;
  float* scr_c = scr_a + (size_t)sms * La.size;
  const double count = (double)rows * (double)n;
  // room for dPre3 / dPre2 / X1 of one network at a time behind the base workspace -> two-kernel (split) path
  float* split = nullptr;
  {
    const size_t base_bytes = (acx_loss_backward_workspace_bytes(ob_dim, num_ac, cont) + 255) & ~(size_t)255;
    if ((n & 3) == 0 && workspace_bytes >= acx_loss_backward_tc_workspace_bytes(ob_dim, num_ac, cont, rows, n))
      split = reinterpret_cast<float*>(wsb + base_bytes);
  }
  if (actor) {
    ACX_REQUIRE(actor_packed && ax2 && ax3 && acs && adv && logp_old && opt && grad_actor && acx_aligned16(actor_packed),
                "%s: null / misaligned actor-side buffer (packed weights and the activation cache are required)", fn);
    Args a; memset(&a, 0, sizeof(a));
    a.params = actor; a.L = La; a.rows = rows; a.n = n; a.obs = obs; a.ac_scale = ac_scale;
    a.ac_in_f = cont ? reinterpret_cast<const float*>(acs) : nullptr;
    a.ac_in_i = cont ? nullptr : reinterpret_cast<const int*>(acs);
    a.adv = adv; a.adv_stats = adv_stats; a.logp_old = logp_old; a.opt = opt;
    a.scratch = scr_a; a.loss_partial = lossp_a; a.x2_in = ax2; a.x3_in = ax3; a.mask_in = amask;
    int rc = cont ? launch_tt<128, 64, true, HEAD_GAUSS>(a, actor_packed, split, grid, st)
                  : launch_tt<128, 64, true, HEAD_CAT>(a, actor_packed, split, grid, st);
    if (rc) return rc;
    acx_grad_reduce_tc_kernel<<<(La.size + 255) / 256, 256, 0, st>>>(scr_a, grid, La.size, grad_actor, lossp_a, stats,
                                                                    opt, 1, count, split != nullptr, La.W3, ACX_A_H3,
                                                                    La.W2, ACX_A_H1);
    ACX_LAUNCH_OK("acx_grad_reduce_tc_kernel(actor)");
  }
  if (critic) {
    ACX_REQUIRE(critic_packed && cx2 && cx3 && targets && grad_critic && acx_aligned16(critic_packed),
                "%s: null / misaligned critic-side buffer (packed weights and the activation cache are required)", fn);
    Args a; memset(&a, 0, sizeof(a));
    a.params = critic; a.L = Lc; a.rows = rows; a.n = n; a.obs = obs;
    a.targets = targets; a.scratch = scr_c; a.loss_partial = lossp_c; a.x2_in = cx2; a.x3_in = cx3; a.mask_in = cmask;
    int rc = launch_tt<256, 128, false, HEAD_VALUE>(a, critic_packed, split, grid, st);
    if (rc) return rc;
    acx_grad_reduce_tc_kernel<<<(Lc.size + 255) / 256, 256, 0, st>>>(scr_c, grid, Lc.size, grad_critic, lossp_c, stats,
                                                                    opt, 0, count, split != nullptr, Lc.W3, ACX_C_H3,
                                                                    Lc.W2, ACX_C_H1);
    ACX_LAUNCH_OK("acx_grad_reduce_tc_kernel(critic)");
  }
  return ACX_OK;
}

}  // namespace

extern "C" size_t acx_activation_cache_floats(int rows, int features, int n) {
  return cache_floats(rows, features > 0 ? features : 0, n > 0 ? n : 0);
}

extern "C" size_t acx_loss_backward_tc_workspace_bytes(int ob_dim, int num_ac, int cont, int rows, int n) {
  const size_t base_bytes = (acx_loss_backward_workspace_bytes(ob_dim, num_ac, cont) + 255) & ~(size_t)255;
  const size_t feat = (size_t)(ACX_C_H3 > ACX_A_H3 ? ACX_C_H3 : ACX_A_H3) + H2;      // dPre3 + dPre2 of one network
  return base_bytes + cache_floats(rows, (int)feat, n) * sizeof(float);
}

extern "C" size_t acx_act_mask_words(int net, int rows, int n) {
  const int h1 = net == 0 ? ACX_A_H1 : ACX_C_H1;
  return (size_t)(rows > 0 ? rows : 0) * (size_t)(((n > 0 ? n : 0) + MT - 1) / MT) * (size_t)(h1 / 32 + 4) * MT;
}

extern "C" size_t acx_tc_packed_bwd_floats(int net, int ob_dim, int num_ac, int cont) {
  (void)ob_dim; (void)num_ac; (void)cont;
  const int h1 = net == 0 ? ACX_A_H1 : ACX_C_H1, h3 = net == 0 ? ACX_A_H3 : ACX_C_H3;
  return (size_t)(h3 / 16 + 8 * (h1 / 128)) * TT_SLOT_FLOATS;
}

extern "C" int acx_tc_pack_weights_bwd(int net, const float* params, int ob_dim, int num_ac, int cont, float* packed,
                                       void* stream) {
  ACX_REQUIRE(net == 0 || net == 1, "acx_tc_pack_weights_bwd: net must be 0 (actor) or 1 (critic), got %d", net);
  ACX_REQUIRE(params && packed, "acx_tc_pack_weights_bwd: null buffer");
  const AcxNetLayout L = net == 0 ? acx_actor_layout(ob_dim, num_ac, cont) : acx_critic_layout(ob_dim);
  acx_tc_pack_bwd_kernel<<<96, 256, 0, (cudaStream_t)stream>>>(params, L, packed);
  ACX_LAUNCH_OK("acx_tc_pack_bwd_kernel");
  return ACX_OK;
}

extern "C" int acx_loss_backward_gauss_tc(const float* actor, const float* critic, const float* actor_packed_bwd,
                                          const float* critic_packed_bwd, int ob_dim, int ac_dim, float ac_scale,
                                          int rows, int n, const float* obs, const float* acs, const float* adv,
                                          const float* adv_stats, const float* logp_old, const float* targets,
                                          const float* opt, float* grad_actor, float* grad_critic, float* stats,
                                          const float* actor_x2, const float* actor_x3, const float* critic_x2,
                                          const float* critic_x3, const unsigned int* actor_masks,
                                          const unsigned int* critic_masks, void* workspace, size_t workspace_bytes,
                                          void* stream) {
  return loss_backward_tc(1, actor, critic, actor_packed_bwd, critic_packed_bwd, ob_dim, ac_dim, ac_scale, rows, n, obs,
                          acs, adv, adv_stats, logp_old, targets, opt, grad_actor, grad_critic, stats, actor_x2,
                          actor_x3, critic_x2, critic_x3, actor_masks, critic_masks, workspace, workspace_bytes,
                          (cudaStream_t)stream);
}

extern "C" int acx_loss_backward_cat_tc(const float* actor, const float* critic, const float* actor_packed_bwd,
                                        const float* critic_packed_bwd, int ob_dim, int n_act, int rows, int n,
                                        const float* obs, const int* acs, const float* adv, const float* adv_stats,
                                        const float* logp_old, const float* targets, const float* opt,
                                        float* grad_actor, float* grad_critic, float* stats, const float* actor_x2,
                                        const float* actor_x3, const float* critic_x2, const float* critic_x3,
                                        const unsigned int* actor_masks, const unsigned int* critic_masks,
                                        void* workspace, size_t workspace_bytes, void* stream) {
  return loss_backward_tc(0, actor, critic, actor_packed_bwd, critic_packed_bwd, ob_dim, n_act, 0.f, rows, n, obs, acs,
                          adv, adv_stats, logp_old, targets, opt, grad_actor, grad_critic, stats, actor_x2, actor_x3,
                          critic_x2, critic_x3, actor_masks, critic_masks, workspace, workspace_bytes,
                          (cudaStream_t)stream);
}
