// K2 (fused policy/value MLP forward + action sampling) and K4 (fused loss + backward).
//
// Replaces, for a whole [rows x n] batch per launch:
//   Actor.act            A3C/policies.py:42-72,110-115      -> acx_policy_forward_sample_{gauss,cat}
//   Actor.get_kl / logp  A3C/policies.py:60,69,75,106-108   -> acx_policy_logp_{gauss,cat}
//   Critic.value         A3C/policies.py:146-150,176-177    -> acx_value_forward
//   Actor/Critic loss + compute_gradients
//                        A3C/policies.py:74-85,152-158      -> acx_loss_backward_{gauss,cat}
// See acx_mlp.cuh for the tile engine.  Head math follows oracle/policies.py line by line.
#include "acx_mlp.cuh"
#include "acx_head.cuh"

using namespace acxmlp;

namespace {

template <int H1, int H3, bool LRELU, int HEAD, int MODE, bool REUSE = false>
__global__ void __launch_bounds__(NT, 1) acx_mlp_kernel(const Args A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
  constexpr bool TRAIN = MODE == MODE_TRAIN;
  constexpr int TJ3 = H3 / 16, NB = H1 / 128;
  const AcxNetLayout L = A.L;
  const int ob = L.ob, nout = L.nout, tid = threadIdx.x;
  const float* __restrict__ P = A.params;

  for (int i = tid; i < ob * H1; i += NT) sm.W1[i] = P[L.W1 + i];
  for (int i = tid; i < H1; i += NT) sm.b1[i] = P[L.b1 + i];
  for (int i = tid; i < H2; i += NT) sm.b2[i] = P[L.b2 + i];
  for (int i = tid; i < H3; i += NT) sm.b3[i] = P[L.b3 + i];
  for (int i = tid; i < H3 * nout; i += NT) sm.Wh[i] = P[L.Wh + i];
  if (tid < nout) {
    sm.bh[tid] = P[L.bh + tid];
    if (HEAD == HEAD_GAUSS) sm.log_std[tid] = P[L.log_std + tid];
  }

  const int tpr = (A.n + MT - 1) / MT;
  const int ntiles = A.rows * tpr;
  const unsigned int mine = ((int)blockIdx.x < ntiles) ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  Ring<H1, H3, TRAIN, REUSE> ring;
  ring.init(&sm, P + L.W2, P + L.W3, A.wT, mine);     // __syncthreads inside: small params visible after it
  const Coord c;

  const unsigned int step0 = A.step0 + ((MODE == MODE_SAMPLE && A.step_base) ? *A.step_base : 0u);
  const double invB = 1.0 / ((double)A.rows * (double)A.n);
  double lsum[4] = {0.0, 0.0, 0.0, 0.0};        // per-thread loss sums (head threads)
  float dls[MAX_OUT] = {0.f, 0.f, 0.f, 0.f};    // d loss / d log_std (gauss train)
  float beta = 0.f, gam = 0.f, adv_mean = 0.f, adv_inv = 1.f;
  if (TRAIN && HEAD != HEAD_VALUE) {
    beta = A.opt[ACX_OPT_BETA_KL]; gam = A.opt[ACX_OPT_GAMMA_ENT];
    if (A.adv_stats) { adv_mean = A.adv_stats[0]; adv_inv = A.adv_stats[1]; }
  }
  HeadScalars hs;
  hs.beta = beta; hs.gam = gam; hs.adv_mean = adv_mean; hs.adv_inv = adv_inv; hs.invB = invB; hs.step0 = step0;
  float* __restrict__ scr = TRAIN ? A.scratch + (size_t)blockIdx.x * L.size : nullptr;
  bool first = true;
  unsigned int act_phase = 0;
  if (TRAIN && tid < L.size - L.bh) scr[L.bh + tid] = 0.f;     // head bias / log_std slots incl. alignment padding

  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int row = tile / tpr, n0 = (tile % tpr) * MT;
    const int valid = min(MT, A.n - n0);
    for (int idx = tid; idx < ob * MT; idx += NT) {
      const int i = idx / MT, m = idx % MT;
      sm.obs[idx] = (m < valid) ? A.obs[((size_t)row * ob + i) * A.n + n0 + m] : 0.f;
    }
    __syncthreads();

    const bool vec_ok = (A.n & 3) == 0;          // tile rows start on 16-byte boundaries of the cache
    if (REUSE) {
      // ---------------------------------------------- layers 2 and 3 come from the activation cache
      const float* g2 = A.x2_in + cache_tile_off(tile, H2);
      const float* g3 = A.x3_in + cache_tile_off(tile, H3);
      if (valid == MT) {
        // full tile: every row is 512 contiguous, 16-byte aligned bytes -> one bulk-TMA copy per row, all in flight
        uint64_t* bar = reinterpret_cast<uint64_t*>(&sm.bar_act);
        if (tid == 0) acx_mbar_expect_tx(bar, (uint32_t)((H2 + H3) * MT * 4));
        __syncthreads();      // the expect_tx is registered before any copy can complete; X2/X3 are idle by now
        if (tid < H2) acx_bulk_g2s(sm.X2 + tid * MS, g2 + (size_t)tid * MT, MT * 4, bar);
        else if (tid - H2 < H3) acx_bulk_g2s(sm.X3 + (tid - H2) * MS, g3 + (size_t)(tid - H2) * MT, MT * 4, bar);
        acx_mbar_wait(bar, act_phase & 1u);
        ++act_phase;
      } else {
        load_tile(sm.X2, g2, (size_t)MT, H2, valid, true);
        load_tile(sm.X3, g3, (size_t)MT, H3, valid, true);
      }
      __syncthreads();
    } else {
    // ------------------------------------------------ forward: layers 1+2 (layer 1 recomputed per 128-row block)
    {
      float acc[8][8];
      f_zero<8>(acc);
      for (int h = 0; h < NB; ++h) {
        layer1_block<H1, LRELU>(sm, h, ob, sm.X3);
        __syncthreads();
        for (int cc = 0; cc < 4; ++cc) {
          const float* slot = ring.begin();
          f_stage<8>(slot, sm.X3 + cc * CH * MS, acc, c);
          __syncthreads();
        }
      }
      f_store<8, LRELU>(sm.X2, sm.b2, acc, c, A.x2_out ? A.x2_out + cache_tile_off(tile, H2) : nullptr, (size_t)MT, valid,
                        true);
    }
    __syncthreads();
    // ------------------------------------------------ forward: layer 3
    {
      float acc[TJ3][8];
      f_zero<TJ3>(acc);
      for (int c3 = 0; c3 < 4; ++c3) {
        const float* slot = ring.begin();
        f_stage<TJ3>(slot, sm.X2 + c3 * CH * MS, acc, c);
        __syncthreads();
      }
      f_store<TJ3, LRELU>(sm.X3, sm.b3, acc, c, A.x3_out ? A.x3_out + cache_tile_off(tile, H3) : nullptr, (size_t)MT,
                          valid, true);
    }
    __syncthreads();
    }

    // ------------------------------------------------ head (one thread per sample)
    const bool basic = (HEAD == HEAD_VALUE) && A.basic_head;
    const float* Xh = basic ? sm.X2 : sm.X3;
    if (tid < MT) {
      const int m = tid;
      const bool live = m < valid;
      const size_t col = (size_t)n0 + m;
      float pre[MAX_OUT];
      head_pre<H3>(sm, Xh, m, nout, pre);
      float dh[MAX_OUT] = {0.f, 0.f, 0.f, 0.f};
      head_apply<HEAD, MODE>(A, sm.log_std, nout, pre, live, row, col, hs, lsum, dls, dh);
      if (TRAIN) {
#pragma unroll
        for (int a = 0; a < MAX_OUT; ++a)
          if (a < nout) sm.dHead[a * MT + m] = dh[a];
      }
    }
    if (!TRAIN) {
      __syncthreads();
      continue;
    }

    // ------------------------------------------------ backward
    __syncthreads();
    {   // head weights: dWh[k][a] = sum_m Xh[k][m] dHead[a][m] ; dbh[a] = sum_m dHead[a][m]
      if (tid < H3) {
        for (int a = 0; a < nout; ++a) {
          float s = 0.f;
          for (int m4 = 0; m4 < MT / 4; ++m4) {
            const float4 x = *reinterpret_cast<const float4*>(&Xh[tid * MS + 4 * m4]);
            const float4 d = *reinterpret_cast<const float4*>(&sm.dHead[a * MT + 4 * m4]);
            s = fmaf(x.x, d.x, s); s = fmaf(x.y, d.y, s); s = fmaf(x.z, d.z, s); s = fmaf(x.w, d.w, s);
          }
          float* p = scr + L.Wh + tid * nout + a;
          *p = first ? s : *p + s;
        }
      } else if (tid >= 128 && tid < 128 + nout) {
        const int a = tid - 128;
        float s = 0.f;
        for (int m = 0; m < MT; ++m) s += sm.dHead[a * MT + m];
        float* p = scr + L.bh + a;
        *p = first ? s : *p + s;
      }
    }
    __syncthreads();
    // dPre3 = (Wh . dHead) * act'(X3), in place
    for (int idx = tid; idx < H3 * MT; idx += NT) {
      const int k = idx >> 7, m = idx & 127;
      float s = 0.f;
      for (int a = 0; a < nout; ++a) s = fmaf(sm.Wh[k * nout + a], sm.dHead[a * MT + m], s);
      float* x = &sm.X3[k * MS + m];
      // Basic_AC critic (policies.py:134-135): the value head reads layer 2, so nothing flows into layer 3
      *x = basic ? 0.f : s * act_grad<LRELU>(*x);
    }
    __syncthreads();
    g_stage<8, TJ3>(sm.X2, sm.X3, scr + L.W3, H3, first, c);
    bias_grad(sm.X3, H3, scr + L.b3, first);
    __syncthreads();
    {   // dPre2 = (W3 . dPre3) * act'(X2), in place: F-type over the transposed copy W3T[j][k]
      float acc[8][8];
      f_zero<8>(acc);
      for (int c3 = 0; c3 < H3 / CH; ++c3) {
        const float* slot = ring.begin();
        f_stage<8>(slot, sm.X3 + c3 * CH * MS, acc, c);
        __syncthreads();
      }
      if (basic) {   // ... and the head's gradient enters layer 2 directly: dX2[k][m] = Wv[k] * dv[m]
#pragma unroll
        for (int jj = 0; jj < 8; ++jj)
#pragma unroll
          for (int mm = 0; mm < 8; ++mm)
            acc[jj][mm] = fmaf(sm.Wh[f_col(jj, c)], sm.dHead[4 * c.b + (mm & 3) + 64 * (mm >> 2)], acc[jj][mm]);
      }
      fb_store<LRELU>(sm.X2, acc, c);
    }
    __syncthreads();
    bias_grad(sm.X2, H2, scr + L.b2, first);
    for (int h = 0; h < NB; ++h) {
      layer1_block<H1, LRELU>(sm, h, ob, sm.X3);        // X1 rows 128h.. (dPre3 is dead by now)
      __syncthreads();
      g_stage<8, 8>(sm.X3, sm.X2, scr + L.W2 + (size_t)h * 128 * H2, H2, first, c);
      float acc[8][8];
      f_zero<8>(acc);
      for (int cc = 0; cc < 4; ++cc) {                  // dX1 block = W2T_h . dPre2
        const float* slot = ring.begin();
        f_stage<8>(slot, sm.X2 + cc * CH * MS, acc, c);
        __syncthreads();                                // (also: every g_stage read of X3 is done before fb_store)
      }
      fb_store<LRELU>(sm.X3, acc, c);                   // dPre1 block, in place over the recomputed X1 block
      __syncthreads();
      layer1_grad<H1>(sm, h, ob, sm.X3, scr + L.W1, scr + L.b1, first);
      __syncthreads();
    }
    first = false;
  }

  // ---------------------------------------------------- per-CTA scalar sums
  if (MODE == MODE_SAMPLE) return;
  if (MODE == MODE_EVAL && (HEAD == HEAD_VALUE || A.kl_partial == nullptr)) return;
  __syncthreads();
  double v[4] = {lsum[0], lsum[1], lsum[2], 0.0};
  if (TRAIN && HEAD == HEAD_GAUSS) v[3] = 0.0;
#pragma unroll
  for (int q = 0; q < 4; ++q) v[q] = acx_warp_sum_d(v[q]);
  float dl[MAX_OUT];
#pragma unroll
  for (int d = 0; d < MAX_OUT; ++d) dl[d] = acx_warp_sum(dls[d]);
  const int w = tid >> 5;
  if ((tid & 31) == 0 && w < 4) {      // the head threads are warps 0..3
#pragma unroll
    for (int q = 0; q < 4; ++q) sm.red[w][q] = v[q];
#pragma unroll
    for (int d = 0; d < MAX_OUT; ++d) sm.dHead[w * MAX_OUT + d] = dl[d];
  }
  __syncthreads();
  if (tid == 0) {
    if (TRAIN) {
#pragma unroll
      for (int q = 0; q < 4; ++q) A.loss_partial[(size_t)blockIdx.x * 4 + q] = (sm.red[0][q] + sm.red[1][q]) + (sm.red[2][q] + sm.red[3][q]);
      if (HEAD == HEAD_GAUSS)
        for (int d = 0; d < nout; ++d) scr[L.log_std + d] = (sm.dHead[d] + sm.dHead[MAX_OUT + d]) + (sm.dHead[2 * MAX_OUT + d] + sm.dHead[3 * MAX_OUT + d]);
    } else {
      A.kl_partial[blockIdx.x] = (sm.red[0][1] + sm.red[1][1]) + (sm.red[2][1] + sm.red[3][1]);
    }
  }
}

// wT = [ W3T[j][k] = W3[k][j]  (H3 x 128) | W2T[h][j][kk] = W2[128h + kk][j]  (H1/128 x 128 x 128) ]
__global__ void __launch_bounds__(256) acx_transpose_weights_kernel(const float* __restrict__ P, AcxNetLayout L,
                                                                    float* __restrict__ wT) {
  const int n3 = L.H3 * 128, n2 = L.H1 * 128;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n3 + n2; i += gridDim.x * blockDim.x) {
    if (i < n3) {
      const int j = i >> 7, k = i & 127;
      wT[i] = P[L.W3 + k * L.H3 + j];
    } else {
      const int r = i - n3, h = r / (128 * 128), j = (r >> 7) & 127, kk = r & 127;
      wT[i] = P[L.W2 + (h * 128 + kk) * 128 + j];
    }
  }
}

// grad[p] = sum over CTAs (fixed order) of scratch[cta][p]; block 0 also folds the loss sums
__global__ void __launch_bounds__(256) acx_grad_reduce_kernel(const float* __restrict__ scratch, int ncta, int size,
                                                              float* __restrict__ grad,
                                                              const double* __restrict__ loss_partial,
                                                              float* __restrict__ stats, const float* __restrict__ opt,
                                                              int is_actor, double count) {
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < size; p += gridDim.x * blockDim.x) {
    // fixed order cta = 0, 1, ...; the loads of eight rows are issued ahead of the dependent adds
    float s = 0.f;
    int cta = 0;
    for (; cta + 8 <= ncta; cta += 8) {
      float x[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) x[u] = scratch[(size_t)(cta + u) * size + p];
#pragma unroll
      for (int u = 0; u < 8; ++u) s += x[u];
    }
    for (; cta < ncta; ++cta) s += scratch[(size_t)cta * size + p];
    grad[p] = s;
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

// ------------------------------------------------------------------ host side
template <int H1, int H3, bool LRELU, int HEAD, int MODE, bool REUSE = false>
int launch(const Args& a, int grid, cudaStream_t st) {
  auto kern = acx_mlp_kernel<H1, H3, LRELU, HEAD, MODE, REUSE>;
  static bool configured[64] = {false};
  int dev = 0;
  ACX_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 64 && !configured[dev]) {
    ACX_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem)));
    configured[dev] = true;
  } else if (dev >= 64) {
    ACX_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem)));
  }
  kern<<<grid, NT, sizeof(Smem), st>>>(a);
  ACX_LAUNCH_OK("acx_mlp_kernel");
  return ACX_OK;
}

int tile_grid(int rows, int n) {
  const long long tiles = (long long)rows * ((n + MT - 1) / MT);
  const int sms = acx_num_sms();
  return (int)(tiles < sms ? tiles : sms);
}

int check_common(const char* fn, const float* params, int ob_dim, int nout, int rows, int n, const float* obs) {
  ACX_REQUIRE(params && obs, "%s: null params/obs", fn);
  ACX_REQUIRE(acx_aligned16(params), "%s: params must be 16-byte aligned (bulk-TMA source)", fn);
  ACX_REQUIRE(ob_dim >= 1 && ob_dim <= MAX_OB_WIDE, "%s: ob_dim %d outside [1,%d]", fn, ob_dim, MAX_OB_WIDE);
  ACX_REQUIRE(nout >= 1 && nout <= MAX_OUT, "%s: output width %d outside [1,%d]", fn, nout, MAX_OUT);
  ACX_REQUIRE(rows >= 0 && n >= 0, "%s: negative batch (rows=%d, n=%d)", fn, rows, n);
  return ACX_OK;
}

}  // namespace

static int sample_gauss(const float* actor, int ob_dim, int ac_dim, float ac_scale, int rows, int n, const float* obs,
                        const float* eps, uint64_t seed, uint64_t index0, uint32_t step0, const uint32_t* step_base,
                        float* ac_out, float* logp_out, float* ent_out, float* mu_out, float* x2_out, float* x3_out,
                        void* stream) {
  int rc = check_common("acx_policy_forward_sample_gauss", actor, ob_dim, ac_dim, rows, n, obs);
  if (rc) return rc;
  if (rows == 0 || n == 0) return ACX_OK;
  ACX_REQUIRE(ac_out, "acx_policy_forward_sample_gauss: ac_out is null");
  ACX_REQUIRE(eps || ac_dim <= 2, "acx_policy_forward_sample_gauss: device RNG supports ac_dim <= 2 (got %d)", ac_dim);
  Args a; memset(&a, 0, sizeof(a));
  a.params = actor; a.L = acx_actor_layout(ob_dim, ac_dim, 1); a.rows = rows; a.n = n; a.obs = obs;
  a.ac_scale = ac_scale; a.noise = eps; a.seed = seed; a.index0 = index0; a.step0 = step0; a.step_base = step_base;
  a.ac_out_f = ac_out; a.logp_out = logp_out; a.ent_out = ent_out; a.head_out = mu_out;
  a.x2_out = x2_out; a.x3_out = x3_out;
  return launch<128, 64, true, HEAD_GAUSS, MODE_SAMPLE>(a, tile_grid(rows, n), (cudaStream_t)stream);
}

extern "C" int acx_policy_forward_sample_gauss(const float* actor, int ob_dim, int ac_dim, float ac_scale, int rows,
                                               int n, const float* obs, const float* eps, uint64_t seed,
                                               uint64_t index0, uint32_t step0, const uint32_t* step_base, float* ac_out,
                                               float* logp_out, float* ent_out, float* mu_out, void* stream) {
  return sample_gauss(actor, ob_dim, ac_dim, ac_scale, rows, n, obs, eps, seed, index0, step0, step_base, ac_out,
                      logp_out, ent_out, mu_out, nullptr, nullptr, stream);
}

extern "C" int acx_policy_forward_sample_gauss_cached(const float* actor, int ob_dim, int ac_dim, float ac_scale,
                                                      int rows, int n, const float* obs, const float* eps,
                                                      uint64_t seed, uint64_t index0, uint32_t step0,
                                                      const uint32_t* step_base, float* ac_out, float* logp_out,
                                                      float* ent_out, float* mu_out, float* x2_out, float* x3_out,
                                                      void* stream) {
  ACX_REQUIRE(x2_out && x3_out, "acx_policy_forward_sample_gauss_cached: null activation cache");
  return sample_gauss(actor, ob_dim, ac_dim, ac_scale, rows, n, obs, eps, seed, index0, step0, step_base, ac_out,
                      logp_out, ent_out, mu_out, x2_out, x3_out, stream);
}

static int sample_cat(const float* actor, int ob_dim, int n_act, int rows, int n, const float* obs, const float* u,
                      uint64_t seed, uint64_t index0, uint32_t step0, const uint32_t* step_base, int* ac_out,
                      float* logp_out, float* ent_out, float* logits_out, float* x2_out, float* x3_out, void* stream) {
  int rc = check_common("acx_policy_forward_sample_cat", actor, ob_dim, n_act, rows, n, obs);
  if (rc) return rc;
  if (rows == 0 || n == 0) return ACX_OK;
  ACX_REQUIRE(ac_out, "acx_policy_forward_sample_cat: ac_out is null");
  Args a; memset(&a, 0, sizeof(a));
  a.params = actor; a.L = acx_actor_layout(ob_dim, n_act, 0); a.rows = rows; a.n = n; a.obs = obs;
  a.noise = u; a.seed = seed; a.index0 = index0; a.step0 = step0; a.step_base = step_base;
  a.ac_out_i = ac_out; a.logp_out = logp_out; a.ent_out = ent_out; a.head_out = logits_out;
  a.x2_out = x2_out; a.x3_out = x3_out;
  return launch<128, 64, true, HEAD_CAT, MODE_SAMPLE>(a, tile_grid(rows, n), (cudaStream_t)stream);
}

extern "C" int acx_policy_forward_sample_cat(const float* actor, int ob_dim, int n_act, int rows, int n,
                                             const float* obs, const float* u, uint64_t seed, uint64_t index0,
                                             uint32_t step0, const uint32_t* step_base, int* ac_out, float* logp_out,
                                             float* ent_out, float* logits_out, void* stream) {
  return sample_cat(actor, ob_dim, n_act, rows, n, obs, u, seed, index0, step0, step_base, ac_out, logp_out, ent_out,
                    logits_out, nullptr, nullptr, stream);
}

extern "C" int acx_policy_forward_sample_cat_cached(const float* actor, int ob_dim, int n_act, int rows, int n,
                                                    const float* obs, const float* u, uint64_t seed, uint64_t index0,
                                                    uint32_t step0, const uint32_t* step_base, int* ac_out,
                                                    float* logp_out, float* ent_out, float* logits_out, float* x2_out,
                                                    float* x3_out, void* stream) {
  ACX_REQUIRE(x2_out && x3_out, "acx_policy_forward_sample_cat_cached: null activation cache");
  return sample_cat(actor, ob_dim, n_act, rows, n, obs, u, seed, index0, step0, step_base, ac_out, logp_out, ent_out,
                    logits_out, x2_out, x3_out, stream);
}

extern "C" size_t acx_kl_workspace_bytes(void) { return sizeof(double) * 1024; }

extern "C" int acx_policy_logp_gauss(const float* actor, int ob_dim, int ac_dim, float ac_scale, int rows, int n,
                                     const float* obs, const float* acs, float* logp_out, float* ent_out,
                                     float* mu_out, const float* logp_old, double* kl_partial, int* kl_blocks,
                                     void* stream) {
  int rc = check_common("acx_policy_logp_gauss", actor, ob_dim, ac_dim, rows, n, obs);
  if (rc) return rc;
  const int grid = (rows == 0 || n == 0) ? 0 : tile_grid(rows, n);
  if (kl_blocks) *kl_blocks = grid;
  if (grid == 0) return ACX_OK;
  ACX_REQUIRE(acs, "acx_policy_logp_gauss: acs is null");
  ACX_REQUIRE((logp_old == nullptr) == (kl_partial == nullptr),
              "acx_policy_logp_gauss: logp_old and kl_partial must be given together");
  Args a; memset(&a, 0, sizeof(a));
  a.params = actor; a.L = acx_actor_layout(ob_dim, ac_dim, 1); a.rows = rows; a.n = n; a.obs = obs;
  a.ac_scale = ac_scale; a.ac_in_f = acs; a.logp_out = logp_out; a.ent_out = ent_out; a.head_out = mu_out;
  a.logp_old = logp_old; a.kl_partial = kl_partial;
  return launch<128, 64, true, HEAD_GAUSS, MODE_EVAL>(a, grid, (cudaStream_t)stream);
}

extern "C" int acx_policy_logp_cat(const float* actor, int ob_dim, int n_act, int rows, int n, const float* obs,
                                   const int* acs, float* logp_out, float* ent_out, float* logits_out,
                                   const float* logp_old, double* kl_partial, int* kl_blocks, void* stream) {
  int rc = check_common("acx_policy_logp_cat", actor, ob_dim, n_act, rows, n, obs);
  if (rc) return rc;
  const int grid = (rows == 0 || n == 0) ? 0 : tile_grid(rows, n);
  if (kl_blocks) *kl_blocks = grid;
  if (grid == 0) return ACX_OK;
  ACX_REQUIRE(acs, "acx_policy_logp_cat: acs is null");
  ACX_REQUIRE((logp_old == nullptr) == (kl_partial == nullptr),
              "acx_policy_logp_cat: logp_old and kl_partial must be given together");
  Args a; memset(&a, 0, sizeof(a));
  a.params = actor; a.L = acx_actor_layout(ob_dim, n_act, 0); a.rows = rows; a.n = n; a.obs = obs;
  a.ac_in_i = acs; a.logp_out = logp_out; a.ent_out = ent_out; a.head_out = logits_out;
  a.logp_old = logp_old; a.kl_partial = kl_partial;
  return launch<128, 64, true, HEAD_CAT, MODE_EVAL>(a, grid, (cudaStream_t)stream);
}

static int value_forward(const float* critic, int ob_dim, int rows, int n, const float* obs, float* v_out,
                         int basic_head, float* x2_out, float* x3_out, void* stream) {
  int rc = check_common("acx_value_forward", critic, ob_dim, 1, rows, n, obs);
  if (rc) return rc;
  if (rows == 0 || n == 0) return ACX_OK;
  ACX_REQUIRE(v_out, "acx_value_forward: v_out is null");
  Args a; memset(&a, 0, sizeof(a));
  a.params = critic; a.L = acx_critic_layout(ob_dim); a.rows = rows; a.n = n; a.obs = obs;
  a.v_out = v_out; a.basic_head = basic_head; a.x2_out = x2_out; a.x3_out = x3_out;
  return launch<256, 128, false, HEAD_VALUE, MODE_EVAL>(a, tile_grid(rows, n), (cudaStream_t)stream);
}

extern "C" int acx_value_forward(const float* critic, int ob_dim, int rows, int n, const float* obs, float* v_out,
                                 int basic_head, void* stream) {
  return value_forward(critic, ob_dim, rows, n, obs, v_out, basic_head, nullptr, nullptr, stream);
}

extern "C" int acx_value_forward_cached(const float* critic, int ob_dim, int rows, int n, const float* obs,
                                        float* v_out, int basic_head, float* x2_out, float* x3_out, void* stream) {
  ACX_REQUIRE(x2_out && x3_out, "acx_value_forward_cached: null activation cache");
  return value_forward(critic, ob_dim, rows, n, obs, v_out, basic_head, x2_out, x3_out, stream);
}

extern "C" size_t acx_loss_backward_workspace_bytes(int ob_dim, int num_ac, int cont) {
  const size_t sms = (size_t)acx_num_sms();
  const size_t pa = (size_t)acx_actor_layout(ob_dim, num_ac, cont).size, pc = (size_t)acx_critic_layout(ob_dim).size;
  const size_t wt = (size_t)(ACX_A_H3 + ACX_A_H1 + ACX_C_H3 + ACX_C_H1) * 128;      // transposed weight copies
  return sms * (pa + pc) * sizeof(float) + 2 * sms * 4 * sizeof(double) + wt * sizeof(float) + 512;
}

// set per call by acx_loss_backward_basic_critic (thread-local: the C ABI is one host thread per device)
static thread_local int g_basic_critic_head = 0;
// activation caches of the current call (acx_loss_backward_*_cached); null = recompute layers 2 and 3
static thread_local const float* g_cache[4] = {nullptr, nullptr, nullptr, nullptr};   // actor x2, x3, critic x2, x3

static int loss_backward(int cont, const float* actor, const float* critic, int ob_dim, int num_ac, float ac_scale,
                         int rows, int n, const float* obs, const void* acs, const float* adv, const float* adv_stats,
                         const float* logp_old, const float* targets, const float* opt, float* grad_actor,
                         float* grad_critic, float* stats, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  const char* fn = cont ? "acx_loss_backward_gauss" : "acx_loss_backward_cat";
  ACX_REQUIRE(actor || critic, "%s: neither actor nor critic given", fn);
  ACX_REQUIRE(rows >= 1 && n >= 1, "%s: empty batch (rows=%d, n=%d): the mean loss is undefined", fn, rows, n);
  ACX_REQUIRE(workspace && workspace_bytes >= acx_loss_backward_workspace_bytes(ob_dim, num_ac, cont),
              "%s: workspace too small (%zu < %zu)", fn, workspace_bytes,
              acx_loss_backward_workspace_bytes(ob_dim, num_ac, cont));
  ACX_REQUIRE(acx_aligned16(workspace), "%s: workspace must be 16-byte aligned", fn);
  const int sms = acx_num_sms();
  const int grid = tile_grid(rows, n);
  const AcxNetLayout La = acx_actor_layout(ob_dim, num_ac, cont), Lc = acx_critic_layout(ob_dim);
  unsigned char* wsb = reinterpret_cast<unsigned char*>(workspace);
  double* lossp_a = reinterpret_cast<double*>(wsb);
  double* lossp_c = lossp_a + (size_t)sms * 4;
  float* wT_a = reinterpret_cast<float*>(wsb + 2 * (size_t)sms * 4 * sizeof(double) + 128);
  float* wT_c = wT_a + (size_t)(ACX_A_H3 + ACX_A_H1) * 128;
  float* scr_a = wT_c + (size_t)(ACX_C_H3 + ACX_C_H1) * 128;
  float* scr_c = scr_a + (size_t)sms * La.size;
  const double count = (double)rows * (double)n;
  if (actor) {
    int rc = check_common(fn, actor, ob_dim, num_ac, rows, n, obs);
    if (rc) return rc;
    ACX_REQUIRE(acs && adv && logp_old && opt && grad_actor, "%s: null actor-side buffer", fn);
    Args a; memset(&a, 0, sizeof(a));
    a.params = actor; a.L = La; a.rows = rows; a.n = n; a.obs = obs; a.ac_scale = ac_scale;
    a.ac_in_f = cont ? reinterpret_cast<const float*>(acs) : nullptr;
    a.ac_in_i = cont ? nullptr : reinterpret_cast<const int*>(acs);
    a.adv = adv; a.adv_stats = adv_stats; a.logp_old = logp_old; a.opt = opt;
    a.scratch = scr_a; a.loss_partial = lossp_a; a.wT = wT_a;
    acx_transpose_weights_kernel<<<48, 256, 0, st>>>(actor, La, wT_a);
    ACX_LAUNCH_OK("acx_transpose_weights_kernel(actor)");
    a.x2_in = g_cache[0]; a.x3_in = g_cache[1];
    if (a.x2_in && a.x3_in)
      rc = cont ? launch<128, 64, true, HEAD_GAUSS, MODE_TRAIN, true>(a, grid, st)
                : launch<128, 64, true, HEAD_CAT, MODE_TRAIN, true>(a, grid, st);
    else
      rc = cont ? launch<128, 64, true, HEAD_GAUSS, MODE_TRAIN>(a, grid, st)
                : launch<128, 64, true, HEAD_CAT, MODE_TRAIN>(a, grid, st);
    if (rc) return rc;
    acx_grad_reduce_kernel<<<(La.size + 255) / 256, 256, 0, st>>>(scr_a, grid, La.size, grad_actor, lossp_a, stats, opt,
                                                                 1, count);
    ACX_LAUNCH_OK("acx_grad_reduce_kernel(actor)");
  }
  if (critic) {
    int rc = check_common(fn, critic, ob_dim, 1, rows, n, obs);
    if (rc) return rc;
    ACX_REQUIRE(targets && grad_critic, "%s: null critic-side buffer", fn);
    Args a; memset(&a, 0, sizeof(a));
    a.params = critic; a.L = Lc; a.rows = rows; a.n = n; a.obs = obs;
    a.targets = targets; a.scratch = scr_c; a.loss_partial = lossp_c; a.wT = wT_c;
    a.basic_head = g_basic_critic_head;
    acx_transpose_weights_kernel<<<96, 256, 0, st>>>(critic, Lc, wT_c);
    ACX_LAUNCH_OK("acx_transpose_weights_kernel(critic)");
    a.x2_in = g_cache[2]; a.x3_in = g_cache[3];
    rc = (a.x2_in && a.x3_in) ? launch<256, 128, false, HEAD_VALUE, MODE_TRAIN, true>(a, grid, st)
                              : launch<256, 128, false, HEAD_VALUE, MODE_TRAIN>(a, grid, st);
    if (rc) return rc;
    acx_grad_reduce_kernel<<<(Lc.size + 255) / 256, 256, 0, st>>>(scr_c, grid, Lc.size, grad_critic, lossp_c, stats,
                                                                 opt, 0, count);
    ACX_LAUNCH_OK("acx_grad_reduce_kernel(critic)");
  }
  return ACX_OK;
}

extern "C" int acx_loss_backward_gauss(const float* actor, const float* critic, int ob_dim, int ac_dim, float ac_scale,
                                       int rows, int n, const float* obs, const float* acs, const float* adv,
                                       const float* adv_stats, const float* logp_old, const float* targets,
                                       const float* opt, float* grad_actor, float* grad_critic, float* stats,
                                       void* workspace, size_t workspace_bytes, void* stream) {
  return loss_backward(1, actor, critic, ob_dim, ac_dim, ac_scale, rows, n, obs, acs, adv, adv_stats, logp_old,
                       targets, opt, grad_actor, grad_critic, stats, workspace, workspace_bytes, (cudaStream_t)stream);
}

extern "C" int acx_loss_backward_cat(const float* actor, const float* critic, int ob_dim, int n_act, int rows, int n,
                                     const float* obs, const int* acs, const float* adv, const float* adv_stats,
                                     const float* logp_old, const float* targets, const float* opt, float* grad_actor,
                                     float* grad_critic, float* stats, void* workspace, size_t workspace_bytes,
                                     void* stream) {
  return loss_backward(0, actor, critic, ob_dim, n_act, 0.f, rows, n, obs, acs, adv, adv_stats, logp_old, targets,
                       opt, grad_actor, grad_critic, stats, workspace, workspace_bytes, (cudaStream_t)stream);
}

extern "C" int acx_loss_backward_basic_critic(const float* critic, int ob_dim, int rows, int n, const float* obs,
                                              const float* targets, float* grad_critic, float* stats, void* workspace,
                                              size_t workspace_bytes, void* stream) {
  g_basic_critic_head = 1;
  const int rc = loss_backward(1, nullptr, critic, ob_dim, 1, 1.f, rows, n, obs, nullptr, nullptr, nullptr, nullptr,
                               targets, nullptr, nullptr, grad_critic, stats, workspace, workspace_bytes,
                               (cudaStream_t)stream);
  g_basic_critic_head = 0;
  return rc;
}

// Same as acx_loss_backward_{gauss,cat}, but layers 2 and 3 of each network are read from the activation cache the
// *_cached forward entry points filled (x2[rows][128][n], x3[rows][H3][n]) instead of being recomputed.  Only valid
// while the parameters are the ones those forwards ran with -- which is the reference's schedule: one update per
// batch, computed on the weights that produced it (A3C/process.py:221-258).
extern "C" int acx_loss_backward_gauss_cached(const float* actor, const float* critic, int ob_dim, int ac_dim,
                                              float ac_scale, int rows, int n, const float* obs, const float* acs,
                                              const float* adv, const float* adv_stats, const float* logp_old,
                                              const float* targets, const float* opt, float* grad_actor,
                                              float* grad_critic, float* stats, const float* actor_x2,
                                              const float* actor_x3, const float* critic_x2, const float* critic_x3,
                                              void* workspace, size_t workspace_bytes, void* stream) {
  g_cache[0] = actor_x2; g_cache[1] = actor_x3; g_cache[2] = critic_x2; g_cache[3] = critic_x3;
  const int rc = loss_backward(1, actor, critic, ob_dim, ac_dim, ac_scale, rows, n, obs, acs, adv, adv_stats, logp_old,
                               targets, opt, grad_actor, grad_critic, stats, workspace, workspace_bytes,
                               (cudaStream_t)stream);
  g_cache[0] = g_cache[1] = g_cache[2] = g_cache[3] = nullptr;
  return rc;
}

extern "C" int acx_loss_backward_cat_cached(const float* actor, const float* critic, int ob_dim, int n_act, int rows,
                                            int n, const float* obs, const int* acs, const float* adv,
                                            const float* adv_stats, const float* logp_old, const float* targets,
                                            const float* opt, float* grad_actor, float* grad_critic, float* stats,
                                            const float* actor_x2, const float* actor_x3, const float* critic_x2,
                                            const float* critic_x3, void* workspace, size_t workspace_bytes,
                                            void* stream) {
  g_cache[0] = actor_x2; g_cache[1] = actor_x3; g_cache[2] = critic_x2; g_cache[3] = critic_x3;
  const int rc = loss_backward(0, actor, critic, ob_dim, n_act, 0.f, rows, n, obs, acs, adv, adv_stats, logp_old,
                               targets, opt, grad_actor, grad_critic, stats, workspace, workspace_bytes,
                               (cudaStream_t)stream);
  g_cache[0] = g_cache[1] = g_cache[2] = g_cache[3] = nullptr;
  return rc;
}
