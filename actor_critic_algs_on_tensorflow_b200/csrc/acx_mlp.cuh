// Tile engine for the policy / value MLPs (K2 forward+sample, K4 loss+backward).
//
// One persistent CTA (256 threads) per SM walks 64-sample tiles.  Activations live in shared
// memory as [feature][sample] (row stride 68 floats, == 4 mod 32, so that both "rows across
// lanes" and "4 consecutive samples per lane" 16-byte accesses are bank-conflict free).
// Weight matrices are streamed from L2 in 64-row chunks through a 3-slot shared-memory ring
// by the TMA engine (cp.async.bulk + mbarrier complete_tx; SASS UBLKCP; eight 4 KB copies per
// chunk -- 8-row groups, each followed by 16 bytes of padding), prefetched two stages ahead, so neither the 99 KB actor nor the 197.5 KB critic has to fit next to the
// activations (SURVEY.md 0.4 / 7 hard part 2).  All math is fp32 FFMA; every output neuron is
// a sequential k-ascending FMA chain, bias added last (tf.layers.dense = matmul + bias_add).
//
// Three register-tiled block GEMMs, all on [feature][sample] operands:
//   F  (forward)   OUT[j][m] += sum_k W[k][j] * X[k][m]        W chunk = 64 rows of k
//   B  (backprop)  OUT[k][m]  = sum_j W[k][j] * D[j][m]        W chunk = 64 rows of k
//   G  (weight grad) dW[k][j] += sum_m X[k][m] * D[j][m]       accumulated into a per-CTA
//                                                              global scratch (L2 resident)
#pragma once
#include "acx_common.cuh"

namespace acxmlp {

constexpr int MT = 64;              // samples per tile
constexpr int MS = 68;              // activation row stride (floats)
constexpr int NT = 256;             // threads per CTA
constexpr int CH = 64;              // weight rows per ring chunk
constexpr int SLOT_FLOATS = CH * 128 + 32;   // 8 groups of (8 rows x J floats + 4 floats of padding), J <= 128
constexpr int NSLOT = 3;
constexpr int H2 = 128;
constexpr int MAX_OB = 8;
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
};

struct Smem {
  float slot[NSLOT][SLOT_FLOATS];
  float X2[H2 * MS];
  float X3[128 * MS];
  float bufA[CH * MS];
  float obs[MAX_OB * MT];
  float W1[MAX_OB * 256];
  float b1[256];
  float b2[128];
  float b3[128];
  float Wh[128 * MAX_OUT];
  float bh[MAX_OUT];
  float log_std[MAX_OUT];
  float dHead[MAX_OUT * MT];
  double red[8][4];
  unsigned long long bar[NSLOT];
};

template <bool LRELU> __device__ inline float act_fn(float x) {
  // lrelu exactly as A3C/policies.py:9-10: 0.8*relu(x) + 0.2*x
  return LRELU ? __fadd_rn(__fmul_rn(0.8f, fmaxf(x, 0.0f)), __fmul_rn(0.2f, x)) : fmaxf(x, 0.0f);
}
template <bool LRELU> __device__ inline float act_grad(float y) {   // from the activation's output
  return y > 0.0f ? 1.0f : (LRELU ? 0.2f : 0.0f);
}

// ---------------------------------------------------------------- weight ring
template <int H1, int H3, bool TRAIN> struct Ring {
  static constexpr int NC1 = H1 / CH;                       // chunks of W2 (rows of H1)
  static constexpr int P = TRAIN ? 2 * (NC1 + 2) : (NC1 + 2);   // ring stages per tile
  const float* W2;
  const float* W3;
  Smem* sm;
  unsigned int g;        // running stage counter (same in every thread)
  unsigned int total;    // stages this CTA will run

  __device__ inline void desc(unsigned int s, const float*& src, int& J) const {
    s %= P;
    if (s < NC1) { src = W2 + (size_t)s * CH * H2; J = H2; return; }
    s -= NC1;
    if (s < 2) { src = W3 + (size_t)s * CH * H3; J = H3; return; }
    s -= 2;
    if (s < 2) { src = W3 + (size_t)s * CH * H3; J = H3; return; }
    s -= 2;
    src = W2 + (size_t)s * CH * H2; J = H2;
  }
  // warp 0 only
  __device__ inline void issue(unsigned int s) const {
    if (s >= total) return;
    const float* src; int J;
    desc(s, src, J);
    const int q = s % NSLOT, lane = threadIdx.x & 31;
    unsigned long long* bar = &sm->bar[q];
    if (lane == 0) acx_mbar_expect_tx(reinterpret_cast<uint64_t*>(bar), (uint32_t)(CH * J * 4));
    __syncwarp();
    // eight 8-row groups, one bulk copy each (4 KB for J = 128); a group is followed by 4 floats of padding
    float* dst = sm->slot[q];
    if (lane < CH / 8)
      acx_bulk_g2s(dst + lane * (8 * J + 4), src + (size_t)lane * 8 * J, (uint32_t)(8 * J * 4),
                   reinterpret_cast<uint64_t*>(bar));
  }
  __device__ inline void init(Smem* s, const float* w2, const float* w3, unsigned int ntiles) {
    sm = s; W2 = w2; W3 = w3; g = 0; total = ntiles * P;
    if (threadIdx.x == 0) {
      for (int i = 0; i < NSLOT; ++i) acx_mbar_init(reinterpret_cast<uint64_t*>(&sm->bar[i]), 1);
      acx_mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x < 32) { issue(0); issue(1); }
  }
  // every thread; returns the slot holding this stage's chunk.  The slot refilled here was last
  // read in stage g-1, which ended with a __syncthreads.
  __device__ inline const float* begin() {
    if (threadIdx.x < 32) issue(g + 2);
    acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm->bar[g % NSLOT]), (g / NSLOT) & 1u);
    const float* s = sm->slot[g % NSLOT];
    ++g;
    return s;
  }
};

// thread coordinates shared by the three GEMMs
struct Coord {
  int a;    // 0..15 : (warp&3)*4 + (lane&3)   -> j-quad (F), k row (B, G)
  int b;    // 0..15 : (warp>>2)*8 + (lane>>2) -> m-quad (F, B), j row (G)
  int lj, wj;   // (lane&3), (warp&3)
  __device__ inline Coord() {
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    lj = l & 3; wj = w & 3;
    a = wj * 4 + lj;
    b = (w >> 2) * 8 + (l >> 2);
  }
  // B stage: row of the 64-row weight chunk owned by accumulator i (0..3).  The four lanes that differ in
  // lj read rows of four different 8-row groups, i.e. 16-byte chunks 4 banks apart: conflict free.
  __device__ inline int brow(int i) const { return 8 * (lj + 4 * (i & 1)) + 2 * wj + (i >> 1); }
};

// offset (floats) of row r of a 64-row chunk with J columns inside a ring slot
template <int J> __device__ inline int slot_row(int r) { return (r >> 3) * (8 * J + 4) + (r & 7) * J; }

// F: acc[jj][mm] += sum_{k<64} W[k][j(jj)] * X[k][4b+mm];  j(jj) = 4a + (jj&3) + 64*(jj>>2)
template <int TJ>
__device__ inline void f_stage(const float* __restrict__ slot, const float* __restrict__ X, float (&acc)[TJ][4],
                               const Coord& c) {
  constexpr int J = 16 * TJ;
  const float* wp = slot + 4 * c.a;
  const float* xp = X + 4 * c.b;
#pragma unroll 8
  for (int k = 0; k < CH; ++k) {
    const float4 x = *reinterpret_cast<const float4*>(xp + k * MS);
    const float* wk = wp + slot_row<J>(k);
    float w[TJ];
    {
      const float4 t = *reinterpret_cast<const float4*>(wk);
      w[0] = t.x; w[1] = t.y; w[2] = t.z; w[3] = t.w;
    }
    if (TJ == 8) {
      const float4 t = *reinterpret_cast<const float4*>(wk + 64);
      w[TJ - 4] = t.x; w[TJ - 3] = t.y; w[TJ - 2] = t.z; w[TJ - 1] = t.w;
    }
#pragma unroll
    for (int j = 0; j < TJ; ++j) {
      acc[j][0] = fmaf(w[j], x.x, acc[j][0]);
      acc[j][1] = fmaf(w[j], x.y, acc[j][1]);
      acc[j][2] = fmaf(w[j], x.z, acc[j][2]);
      acc[j][3] = fmaf(w[j], x.w, acc[j][3]);
    }
  }
}
template <int TJ> __device__ inline int f_col(int jj, const Coord& c) { return 4 * c.a + (jj & 3) + 64 * (jj >> 2); }

// F epilogue: OUT[j][m] = act(acc + bias[j])
template <int TJ, bool LRELU>
__device__ inline void f_store(float* __restrict__ OUT, const float* __restrict__ bias, float (&acc)[TJ][4],
                               const Coord& c) {
#pragma unroll
  for (int jj = 0; jj < TJ; ++jj) {
    const int j = f_col<TJ>(jj, c);
    const float bj = bias[j];
    float4 o;
    o.x = act_fn<LRELU>(__fadd_rn(acc[jj][0], bj));
    o.y = act_fn<LRELU>(__fadd_rn(acc[jj][1], bj));
    o.z = act_fn<LRELU>(__fadd_rn(acc[jj][2], bj));
    o.w = act_fn<LRELU>(__fadd_rn(acc[jj][3], bj));
    *reinterpret_cast<float4*>(OUT + j * MS + 4 * c.b) = o;
  }
}

// B: out[i][mm] = sum_{j<J} W[brow(i)][j] * D[j][4b+mm]   (rows brow(i) of the 64-row chunk)
template <int J>
__device__ inline void b_stage(const float* __restrict__ slot, const float* __restrict__ D, float (&acc)[4][4],
                               const Coord& c) {
  const float* wp[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) wp[i] = slot + slot_row<J>(c.brow(i));
  const float* dp = D + 4 * c.b;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int m = 0; m < 4; ++m) acc[i][m] = 0.0f;
#pragma unroll 2
  for (int j4 = 0; j4 < J / 4; ++j4) {
    float4 d[4], w[4];
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) d[jj] = *reinterpret_cast<const float4*>(dp + (4 * j4 + jj) * MS);
#pragma unroll
    for (int i = 0; i < 4; ++i) w[i] = *reinterpret_cast<const float4*>(wp[i] + 4 * j4);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float wv[4] = {w[i].x, w[i].y, w[i].z, w[i].w};
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        acc[i][0] = fmaf(wv[jj], d[jj].x, acc[i][0]);
        acc[i][1] = fmaf(wv[jj], d[jj].y, acc[i][1]);
        acc[i][2] = fmaf(wv[jj], d[jj].z, acc[i][2]);
        acc[i][3] = fmaf(wv[jj], d[jj].w, acc[i][3]);
      }
    }
  }
}
// B epilogue, in place: Y[row0 + brow(i)][m] = acc * act'(Y[..][m])
template <bool LRELU>
__device__ inline void b_store(float* __restrict__ Y, float (&acc)[4][4], const Coord& c) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float4* p = reinterpret_cast<float4*>(Y + c.brow(i) * MS + 4 * c.b);
    const float4 y = *p;
    float4 o;
    o.x = acc[i][0] * act_grad<LRELU>(y.x);
    o.y = acc[i][1] * act_grad<LRELU>(y.y);
    o.z = acc[i][2] * act_grad<LRELU>(y.z);
    o.w = acc[i][3] * act_grad<LRELU>(y.w);
    *p = o;
  }
}

// G: dW[a+16i][b+16i'] (+)= sum_m X[a+16i][m] * D[b+16i'][m]; accumulated in the CTA's global scratch
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
#pragma unroll
    for (int m4 = 0; m4 < MT / 4; ++m4) {
      const float4 d = *reinterpret_cast<const float4*>(D + j * MS + 4 * m4);
      s += d.x; s += d.y; s += d.z; s += d.w;
    }
    gscr[j] = first ? s : gscr[j] + s;
  }
}

}  // namespace acxmlp
