// Tile engine for the policy / value MLPs (K2 forward+sample, K4 loss+backward).
//
// One persistent CTA (256 threads) per SM walks 128-sample tiles.  Activations live in shared
// memory as [feature][sample] (row stride 132 floats, == 4 mod 32, so that "consecutive rows
// across lanes" 16-byte accesses are bank-conflict free).  Weight matrices are streamed from L2
// in 32-row chunks (16 KB, ONE bulk copy each) through a 3-slot shared-memory ring by the TMA
// engine (cp.async.bulk + mbarrier complete_tx; SASS UBLKCP), prefetched two stages ahead, so
// neither the 99 KB actor nor the 197.5 KB critic has to fit next to the activations
// (SURVEY.md 0.4 / 7 hard part 2).  All math is fp32 FFMA; every output neuron is a sequential
// k-ascending FMA chain, bias added last (tf.layers.dense = matmul + bias_add).
//
// Shared-memory bandwidth is the binding resource of an FFMA GEMM on this machine (every
// LDS.128 costs four wavefronts, broadcast or not; ncu, profiles/r01_*): with T_a x T_b outputs
// per thread a k-step costs (T_a + T_b) wavefronts per T_a*T_b FFMA, and the FMA pipes issue four
// warp-FFMA per cycle against one wavefront per cycle -> 8 x 8 tiles are the break-even, and both
// block GEMMs below use them:
//   F  OUT[j][m] += sum_k W[k][j] * X[k][m]     W chunk = 32 rows of k.  Forward layers, and
//      back-propagation dX[k][m] = sum_j WT[j][k] * D[j][m] with a transposed weight copy WT
//      (made once per call by acx_transpose_weights_kernel), so that every ring chunk is read
//      row-wise by all lanes (no padding, no bank games).
//   G  dW[k][j] += sum_m X[k][m] * D[j][m]      accumulated into a per-CTA global scratch row
//      (L2 resident), reduced across CTAs in fixed order afterwards.
// Layer 1 (K = ob_dim <= 8) is recomputed in 128-row blocks into the (then idle) X3 buffer.
#pragma once
#include "acx_common.cuh"

namespace acxmlp {

constexpr int MT = 128;             // samples per tile
constexpr int MS = 132;             // activation row stride (floats)
constexpr int NT = 256;             // threads per CTA
constexpr int CH = 32;              // weight rows per ring chunk
constexpr int SLOT_FLOATS = CH * 128;
constexpr int NSLOT = 3;
constexpr int H2 = 128;
constexpr int MAX_OB = 8;           // network inputs of the tensor-core kernels (shared memory is full of weights there)
constexpr int MAX_OB_WIDE = 24;     // of the FFMA engine: 3 stacked frames of Acrobot's 6 observations (--frames 3, the
                                    // reference's process_fn default, A3C/process.py:156-157)
constexpr int MAX_OUT = 4;

enum { HEAD_GAUSS = 0, HEAD_CAT = 1, HEAD_VALUE = 2 };
enum { MODE_SAMPLE = 0, MODE_EVAL = 1, MODE_TRAIN = 2 };

struct Args {
  const float* params;       // flat net parameters (16-byte aligned)
  AcxNetLayout L;
  int rows, n;               // batch = rows x n ; obs is [rows][ob][n]
  const float* obs;
  float ac_scale;
  // sampling noise: eps[rows][nout][n] (gauss) / u[rows][n] (cat); NULL -> Philox(seed, index0+col, step0+row)
  const float* noise;
  unsigned long long seed, index0;
  unsigned int step0;
  const unsigned int* step_base;   // nullable device counter added to step0
  float* ac_out_f;           // gauss sample: [rows][nout][n]
  int* ac_out_i;             // cat sample:   [rows][n]
  const float* ac_in_f;      // eval/train
  const int* ac_in_i;
  float* logp_out;           // [rows][n]
  float* ent_out;            // [rows][n] (nullable)
  float* head_out;           // mu [rows][nout][n] / logits [rows][nout][n] (nullable)
  const float* logp_old;     // eval (nullable) / train
  double* kl_partial;        // eval: per-CTA sum (lp_old - lp)^2 (nullable)
  float* v_out;              // value forward [rows][n]
  int basic_head;            // Basic_AC critic: value head reads the second layer
  // train
  const float* adv;          // raw advantages [rows][n]
  const float* adv_stats;    // device {mean, 1/(1e-8+std)}; NULL -> advantages used as given
  const float* targets;      // critic targets [rows][n]
  const float* opt;          // device opt block (beta_kl, gamma_ent)
  float* scratch;            // per-CTA gradient partials [grid][L.size]
  double* loss_partial;      // per-CTA loss sums [grid][4]
  const float* wT;           // transposed weights: W3T [H3][128] then W2T [H1/128][128][128]
  // activation cache, tile-blocked: x2[rows][tiles][128][128], x3[rows][tiles][H3][128] (tiles = ceil(n / 128) blocks of
  // 128 samples, a block is one contiguous F x 512-byte region; samples beyond n hold zeros).  Forward kernels may
  // write it, the training kernels may read it instead of recomputing layers 2 and 3 (valid while the weights are
  // unchanged).  A tile's rows lie 512 bytes apart instead of 4 n bytes: one DRAM page / TLB entry per tile.
  float* x2_out;
  float* x3_out;
  const float* x2_in;
  const float* x3_in;
  // activation sign masks (tensor-core path, optional): per tile NW = H1/32 + 4 words x 128 samples, u32; word w < H1/32:
  // bit e = X1[32 w + e] > 0, word H1/32 + w: bit e = X2[32 w + e] > 0.  Written by the forward kernels, read by the
  // chain kernel instead of re-deriving act'(X2) from the cache and act'(X1) from the observations.
  unsigned int* mask_out;
  const unsigned int* mask_in;
  // persistent rollout (acx_rollout_tc): the env step runs in the sampling head, rows = T time steps of the same envs
  float* env_state;          // [S][n], read before the first step, written after the last
  int* env_elapsed;          // [n]
  float* obs_rw;             // = obs, [rows + 1][ob][n]: row 0 is read, rows 1..rows are written
  float* rew_out;            // [rows][n]
  unsigned char* done_out;   // [rows][n]
  const float* reset_u;      // [rows][S][n] uniforms for resets, or NULL (Philox)
  int max_steps;
};

struct Smem {
  static constexpr int OBW = MAX_OB_WIDE;
  float slot[NSLOT][SLOT_FLOATS];
  float X2[H2 * MS];
  float X3[128 * MS];        // layer-3 output / dPre3; also the staging area of a 128-row block of layer 1
  float obs[MAX_OB_WIDE * MT];
  float W1[MAX_OB_WIDE * 256];
  float b1[256];
  float b2[128];
  float b3[128];
  float Wh[128 * MAX_OUT];
  float bh[MAX_OUT];
  float log_std[MAX_OUT];
  float dHead[MAX_OUT * MT];
  double red[8][4];
  unsigned long long bar[NSLOT];
  unsigned long long bar_act;      // arrival of a cached activation tile (bulk-TMA row copies)
};

// float offset of tile `tile` (= row * tiles_per_row + tile_in_row) of a tile-blocked [..][F][128] array
__host__ __device__ inline size_t cache_tile_off(int tile, int F) { return (size_t)tile * F * MT; }
__host__ __device__ inline size_t mask_tile_off(int tile, int H1) { return (size_t)tile * (H1 / 32 + 4) * MT; }
__host__ __device__ inline size_t cache_floats(int rows, int F, int n) {
  return (size_t)(rows > 0 ? rows : 0) * (size_t)((n + MT - 1) / MT) * F * MT;
}

template <bool LRELU> __device__ inline float act_fn(float x) {
  // lrelu exactly as A3C/policies.py:9-10: 0.8*relu(x) + 0.2*x
  return LRELU ? __fadd_rn(__fmul_rn(0.8f, fmaxf(x, 0.0f)), __fmul_rn(0.2f, x)) : fmaxf(x, 0.0f);
}
template <bool LRELU> __device__ inline float act_grad(float y) {   // from the activation's output
  return y > 0.0f ? 1.0f : (LRELU ? 0.2f : 0.0f);
}

// ---------------------------------------------------------------- weight ring
// Stage order per tile: W2 (4 chunks per 128-row block of layer 1), W3 (4 chunks); training adds
// W3T (H3/32 chunks) and W2T (4 chunks per block).
template <int H1, int H3, bool TRAIN, bool REUSE = false> struct Ring {
  static constexpr int NB = H1 / 128;
  static constexpr int NF = REUSE ? 0 : 4 * NB + 4;       // forward stages (none when activations come from the cache)
  static constexpr int P = TRAIN ? NF + H3 / CH + 4 * NB : NF;
  const float* W2;
  const float* W3;
  const float* W3T;
  const float* W2T;
  Smem* sm;
  unsigned int g;        // running stage counter (same in every thread)
  unsigned int total;    // stages this CTA will run

  __device__ inline void desc(unsigned int s, const float*& src, unsigned int& bytes) const {
    s %= P;
    if (!REUSE) {
      if (s < 4 * NB) { src = W2 + (size_t)s * CH * H2; bytes = CH * H2 * 4; return; }
      s -= 4 * NB;
      if (s < 4) { src = W3 + (size_t)s * CH * H3; bytes = CH * H3 * 4; return; }
      s -= 4;
    }
    if (s < H3 / CH) { src = W3T + (size_t)s * CH * 128; bytes = CH * 128 * 4; return; }
    s -= H3 / CH;
    src = W2T + (size_t)s * CH * 128; bytes = CH * 128 * 4;
  }
  // thread 0 only: one bulk copy per chunk
  __device__ inline void issue(unsigned int s) const {
    if (s >= total) return;
    const float* src; unsigned int bytes;
    desc(s, src, bytes);
    uint64_t* bar = reinterpret_cast<uint64_t*>(&sm->bar[s % NSLOT]);
    acx_mbar_expect_tx(bar, bytes);
    acx_bulk_g2s(sm->slot[s % NSLOT], src, bytes, bar);
  }
  __device__ inline void init(Smem* s, const float* w2, const float* w3, const float* wT, unsigned int ntiles) {
    sm = s; W2 = w2; W3 = w3; W3T = wT; W2T = wT ? wT + (size_t)H3 * 128 : nullptr; g = 0; total = ntiles * P;
    if (threadIdx.x == 0) {
      for (int i = 0; i < NSLOT; ++i) acx_mbar_init(reinterpret_cast<uint64_t*>(&sm->bar[i]), 1);
      acx_mbar_init(reinterpret_cast<uint64_t*>(&sm->bar_act), 1);
      acx_mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) { issue(0); issue(1); }
  }
  // every thread; returns the slot holding this stage's chunk.  The slot refilled here was last
  // read in stage g-1, which ended with a __syncthreads.
  __device__ inline const float* begin() {
    if (threadIdx.x == 0) issue(g + 2);
    acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm->bar[g % NSLOT]), (g / NSLOT) & 1u);
    const float* s = sm->slot[g % NSLOT];
    ++g;
    return s;
  }
};

// thread coordinates shared by the block GEMMs
struct Coord {
  int a;    // 0..15 : (warp&3)*4 + (lane&3)   -> j-quad (F), k row (G)
  int b;    // 0..15 : (warp>>2)*8 + (lane>>2) -> m-quad (F), j row (G)
  __device__ inline Coord() {
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    a = (w & 3) * 4 + (l & 3);
    b = (w >> 2) * 8 + (l >> 2);
  }
};

// output feature owned by accumulator row jj: 4a + (jj&3) (+64 for jj >= 4)
__device__ inline int f_col(int jj, const Coord& c) { return 4 * c.a + (jj & 3) + 64 * (jj >> 2); }

// F: acc[jj][mm] += sum_{k<32} W[k][f_col(jj)] * X[k][m(mm)],  m(mm) = 4b + (mm&3) + 64*(mm>>2)
template <int TJ>
__device__ inline void f_stage(const float* __restrict__ slot, const float* __restrict__ X, float (&acc)[TJ][8],
                               const Coord& c) {
  constexpr int J = 16 * TJ;
  const float* wp = slot + 4 * c.a;
  const float* xp = X + 4 * c.b;
#pragma unroll 4
  for (int k = 0; k < CH; ++k) {
    const float4 x0 = *reinterpret_cast<const float4*>(xp + k * MS);
    const float4 x1 = *reinterpret_cast<const float4*>(xp + k * MS + 64);
    const float x[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
    float w[TJ];
    {
      const float4 t = *reinterpret_cast<const float4*>(wp + k * J);
      w[0] = t.x; w[1] = t.y; w[2] = t.z; w[3] = t.w;
    }
    if (TJ == 8) {
      const float4 t = *reinterpret_cast<const float4*>(wp + k * J + 64);
      w[TJ - 4] = t.x; w[TJ - 3] = t.y; w[TJ - 2] = t.z; w[TJ - 1] = t.w;
    }
#pragma unroll
    for (int j = 0; j < TJ; ++j)
#pragma unroll
      for (int m = 0; m < 8; ++m) acc[j][m] = fmaf(w[j], x[m], acc[j][m]);
  }
}

template <int TJ> __device__ inline void f_zero(float (&acc)[TJ][8]) {
#pragma unroll
  for (int j = 0; j < TJ; ++j)
#pragma unroll
    for (int m = 0; m < 8; ++m) acc[j][m] = 0.f;
}

// forward epilogue: OUT[j][m] = act(acc + bias[j])
// gout (nullable): global copy of the tile, row j at gout + j * gld (column 0 = first sample of the tile)
template <int TJ, bool LRELU>
__device__ inline void f_store(float* __restrict__ OUT, const float* __restrict__ bias, float (&acc)[TJ][8],
                               const Coord& c, float* __restrict__ gout = nullptr, size_t gld = 0, int valid = MT,
                               bool vec_ok = false) {
#pragma unroll
  for (int jj = 0; jj < TJ; ++jj) {
    const int j = f_col(jj, c);
    const float bj = bias[j];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float4 o;
      o.x = act_fn<LRELU>(__fadd_rn(acc[jj][4 * h + 0], bj));
      o.y = act_fn<LRELU>(__fadd_rn(acc[jj][4 * h + 1], bj));
      o.z = act_fn<LRELU>(__fadd_rn(acc[jj][4 * h + 2], bj));
      o.w = act_fn<LRELU>(__fadd_rn(acc[jj][4 * h + 3], bj));
      const int m = 4 * c.b + 64 * h;
      *reinterpret_cast<float4*>(OUT + j * MS + m) = o;
      if (gout != nullptr) {          // cache tile (always whole and 16-byte aligned): zeros beyond the batch
        (void)vec_ok;
        float4 z = o;
        if (m + 0 >= valid) z.x = 0.f;
        if (m + 1 >= valid) z.y = 0.f;
        if (m + 2 >= valid) z.z = 0.f;
        if (m + 3 >= valid) z.w = 0.f;
        *reinterpret_cast<float4*>(gout + (size_t)j * gld + m) = z;
      }
    }
  }
}

// cached activations -> shared tile: DST[r][m] = src[r * gld + m] for r < rows_, m < valid (zero beyond)
__device__ inline void load_tile(float* __restrict__ DST, const float* __restrict__ src, size_t gld, int rows_,
                                 int valid, bool vec_ok) {
  for (int idx = threadIdx.x; idx < rows_ * (MT / 4); idx += NT) {
    const int r = idx >> 5, m = (idx & 31) * 4;
    const float* g = src + (size_t)r * gld + m;
    float4 v;
    if (vec_ok && m + 3 < valid) {
      v = __ldg(reinterpret_cast<const float4*>(g));
    } else {
      v.x = (m + 0 < valid) ? __ldg(g + 0) : 0.f;
      v.y = (m + 1 < valid) ? __ldg(g + 1) : 0.f;
      v.z = (m + 2 < valid) ? __ldg(g + 2) : 0.f;
      v.w = (m + 3 < valid) ? __ldg(g + 3) : 0.f;
    }
    *reinterpret_cast<float4*>(DST + r * MS + m) = v;
  }
}

// back-prop epilogue, in place: Y[k][m] = acc * act'(Y[k][m]) over the thread's 8 x 8 outputs
template <bool LRELU>
__device__ inline void fb_store(float* __restrict__ Y, float (&acc)[8][8], const Coord& c) {
#pragma unroll
  for (int jj = 0; jj < 8; ++jj) {
    const int j = f_col(jj, c);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float4* p = reinterpret_cast<float4*>(Y + j * MS + 4 * c.b + 64 * h);
      const float4 y = *p;
      float4 o;
      o.x = acc[jj][4 * h + 0] * act_grad<LRELU>(y.x);
      o.y = acc[jj][4 * h + 1] * act_grad<LRELU>(y.y);
      o.z = acc[jj][4 * h + 2] * act_grad<LRELU>(y.z);
      o.w = acc[jj][4 * h + 3] * act_grad<LRELU>(y.w);
      *p = o;
    }
  }
}

// G: dW[a+16i][b+16i'] (+)= sum_{m<128} X[a+16i][m] * D[b+16i'][m]; accumulated in the CTA's global scratch
template <int TK, int TJ>
__device__ inline void g_stage(const float* __restrict__ X, const float* __restrict__ D, float* __restrict__ gscr,
                               int ldg, bool first, const Coord& c) {
  float acc[TK][TJ];
#pragma unroll
  for (int i = 0; i < TK; ++i)
#pragma unroll
    for (int j = 0; j < TJ; ++j) acc[i][j] = 0.0f;
  const float* xp = X + c.a * MS;
  const float* dp = D + c.b * MS;
#pragma unroll 2
  for (int m4 = 0; m4 < MT / 4; ++m4) {
    float4 x[TK], d[TJ];
#pragma unroll
    for (int i = 0; i < TK; ++i) x[i] = *reinterpret_cast<const float4*>(xp + 16 * i * MS + 4 * m4);
#pragma unroll
    for (int j = 0; j < TJ; ++j) d[j] = *reinterpret_cast<const float4*>(dp + 16 * j * MS + 4 * m4);
#pragma unroll
    for (int i = 0; i < TK; ++i)
#pragma unroll
      for (int j = 0; j < TJ; ++j) {
        float s = acc[i][j];
        s = fmaf(x[i].x, d[j].x, s);
        s = fmaf(x[i].y, d[j].y, s);
        s = fmaf(x[i].z, d[j].z, s);
        s = fmaf(x[i].w, d[j].w, s);
        acc[i][j] = s;
      }
  }
#pragma unroll
  for (int i = 0; i < TK; ++i)
#pragma unroll
    for (int j = 0; j < TJ; ++j) {
      float* p = gscr + (size_t)(c.a + 16 * i) * ldg + (c.b + 16 * j);
      *p = first ? acc[i][j] : (*p + acc[i][j]);
    }
}

// db[j] (+)= sum_m D[j][m] for j < rows   (threads j < rows)
__device__ inline void bias_grad(const float* __restrict__ D, int rows, float* __restrict__ gscr, bool first) {
  const int j = threadIdx.x;
  if (j < rows) {
    float s = 0.0f;
#pragma unroll 8
    for (int m4 = 0; m4 < MT / 4; ++m4) {
      const float4 d = *reinterpret_cast<const float4*>(D + j * MS + 4 * m4);
      s += d.x; s += d.y; s += d.z; s += d.w;
    }
    gscr[j] = first ? s : gscr[j] + s;
  }
}

// ---------------------------------------------------------------- small pieces shared by the MLP kernels
// (SM: the kernel's shared-memory struct; needs obs, W1, b1, Wh, bh)
template <int H1, bool LRELU, class SM>
__device__ inline void layer1_block(SM& sm, int h, int ob, float* __restrict__ dst) {
  // dst[r][m] = X1[128h + r][m] = act(sum_i obs[i][m] * W1[i][k] + b1[k]), r < 128, m < 128
  const int r = threadIdx.x >> 1, q = (threadIdx.x & 1) * 64, k = h * 128 + r;
  const float bk = sm.b1[k];
#pragma unroll 4
  for (int m4 = 0; m4 < 16; ++m4) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i = 0; i < ob; ++i) {
      const float w = sm.W1[i * H1 + k];
      const float4 o = *reinterpret_cast<const float4*>(&sm.obs[i * MT + q + 4 * m4]);
      s.x = fmaf(o.x, w, s.x); s.y = fmaf(o.y, w, s.y); s.z = fmaf(o.z, w, s.z); s.w = fmaf(o.w, w, s.w);
    }
    float4 o;
    o.x = act_fn<LRELU>(__fadd_rn(s.x, bk)); o.y = act_fn<LRELU>(__fadd_rn(s.y, bk));
    o.z = act_fn<LRELU>(__fadd_rn(s.z, bk)); o.w = act_fn<LRELU>(__fadd_rn(s.w, bk));
    *reinterpret_cast<float4*>(dst + r * MS + q + 4 * m4) = o;
  }
}

// dW1[i][128h + j] (+)= sum_m obs[i][m] * D[j][m], db1 likewise, from a 128-row block D of dPre1
template <int H1, class SM>
__device__ inline void layer1_grad(SM& sm, int h, int ob, const float* __restrict__ D, float* __restrict__ scrW1,
                                   float* __restrict__ scrb1, bool first) {
  const int j = threadIdx.x;
  if (j < 128) {
    constexpr int OBW = SM::OBW;
    float sb = 0.f, sw[OBW];
#pragma unroll
    for (int i = 0; i < OBW; ++i) sw[i] = 0.f;
#pragma unroll 4
    for (int m4 = 0; m4 < MT / 4; ++m4) {
      const float4 d = *reinterpret_cast<const float4*>(D + j * MS + 4 * m4);
      sb += d.x; sb += d.y; sb += d.z; sb += d.w;
#pragma unroll
      for (int i = 0; i < OBW; ++i)
        if (i < ob) {
          const float4 o = *reinterpret_cast<const float4*>(&sm.obs[i * MT + 4 * m4]);
          sw[i] = fmaf(o.x, d.x, sw[i]); sw[i] = fmaf(o.y, d.y, sw[i]);
          sw[i] = fmaf(o.z, d.z, sw[i]); sw[i] = fmaf(o.w, d.w, sw[i]);
        }
    }
    const int k = h * 128 + j;
    scrb1[k] = first ? sb : scrb1[k] + sb;
#pragma unroll
    for (int i = 0; i < OBW; ++i)
      if (i < ob) {
        float* p = scrW1 + i * H1 + k;
        *p = first ? sw[i] : *p + sw[i];
      }
  }
}

// pre-activation of the head for sample m: pre[a] = sum_k Xh[k][m] * Wh[k][a] + bh[a]
template <int HH, class SM>
__device__ inline void head_pre(const SM& sm, const float* __restrict__ Xh, int m, int nout, float (&pre)[MAX_OUT]) {
#pragma unroll
  for (int a = 0; a < MAX_OUT; ++a) pre[a] = 0.f;
  for (int k0 = 0; k0 < HH; k0 += 8) {        // loads batched ahead of the k-ascending FMA chain
    float x[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) x[u] = Xh[(k0 + u) * MS + m];
#pragma unroll
    for (int u = 0; u < 8; ++u)
#pragma unroll
      for (int a = 0; a < MAX_OUT; ++a)
        if (a < nout) pre[a] = fmaf(x[u], sm.Wh[(k0 + u) * nout + a], pre[a]);
  }
#pragma unroll
  for (int a = 0; a < MAX_OUT; ++a)
    if (a < nout) pre[a] = __fadd_rn(pre[a], sm.bh[a]);
}

}  // namespace acxmlp
