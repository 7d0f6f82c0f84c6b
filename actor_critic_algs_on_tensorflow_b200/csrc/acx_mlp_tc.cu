// Tensor-core (tcgen05, 3xTF32) variant of the forward passes of K2: policy sample, policy log-prob / KL,
// value.  Same math as acx_mlp.cu (A3C/policies.py:42-72,110-115,146-150,176-177), same heads; the two
// 128-wide hidden GEMMs run on the 5th-generation tensor cores with fp32-class accuracy:
//     x ~= hi + lo  (both exactly representable in TF32),   A.B ~= Ahi.Bhi + Ahi.Blo + Alo.Bhi
// accumulated in fp32 in tensor memory (error ~2^-21 per product; the dropped lo.lo term is ~2^-22).
//
// One persistent 256-thread CTA per SM walks 128-sample tiles; TMEM lane = sample.
//   * A operand (activations): two shared-memory sub-buffers [128 samples][64 features], hi and lo planes,
//     un-swizzled K-major core-matrix layout [s/8][k/4][8][4] (csrc/acx_umma.cuh).  The epilogue thread of
//     sample s reads its accumulator row from TMEM (tcgen05.ld 32x32b), applies bias + activation, splits,
//     and writes 16-byte chunks straight into that layout (conflict free), in place over the buffer the
//     finished MMAs have consumed.
//   * B operand (weights): packed once per parameter update by acx_tc_pack_weights into K-major hi/lo
//     chunks of 16 input rows (16 KB for 128 outputs), streamed through a 3-slot ring by bulk-TMA
//     (full barriers) and released by tcgen05.commit (empty barriers).
//   * One elected thread issues TMA + MMA; everybody else computes layer 1 (ob_dim <= 8, plain FFMA),
//     the epilogues and the heads.  Layer 1 and the heads are bit-identical to the FFMA engine.
// Only K-major operands are used: un-swizzled MN-major TF32 operands return zeros on this part
// (tests/tc/run_tc_test.py), which is also why the backward pass (whose weight-gradient GEMMs need the
// transposed operands) stays on the FFMA engine for now.
#include "acx_mlp.cuh"
#include "acx_umma.cuh"

using namespace acxmlp;
using namespace acxumma;

namespace {

constexpr int TC_NT = 256;
constexpr int TC_KC = 16;                        // input rows per ring chunk
constexpr int TC_SLOT_FLOATS = 2 * 128 * TC_KC;  // hi + lo planes for up to 128 outputs
constexpr int TC_NSLOT = 3;
constexpr int TC_SUB_FLOATS = 128 * 64;          // one plane of one A sub-buffer
constexpr int TC_TMEM_COLS = 256;

struct TcSmem {
  float A[2][2][TC_SUB_FLOATS];       // [sub-buffer][hi|lo][128 x 64]         131,072 B
  float slot[TC_NSLOT][TC_SLOT_FLOATS];  //                                     49,152 B
  float obs[MAX_OB * MT];
  float W1[MAX_OB * 256];
  float b1[256];
  float b2[128];
  float b3[128];
  float Wh[128 * MAX_OUT];
  float bh[MAX_OUT];
  float log_std[MAX_OUT];
  double red[8];
  unsigned long long full[TC_NSLOT], empty[TC_NSLOT], layer_done, sub_free[2];
  unsigned int tmem_base;
};

// byte offset of element (s, k) inside one plane of a sub-buffer [s/8][k/4][8][4]
__device__ inline int a_off(int s, int k) { return ((s >> 3) * 16 + (k >> 2)) * 32 + (s & 7) * 4 + (k & 3); }

struct TcPipe {
  TcSmem* sm;
  const float* packed;        // chunk stream of this network (global)
  unsigned int g;             // chunks consumed so far (running)
  unsigned int issued;        // chunks whose TMA has been issued
  unsigned int total;
  int chunks_per_tile;
  int n2_chunks;              // chunks of layer 2 (N = 128 each, 4096 floats per chunk)
  int n3_floats;              // floats per layer-3 chunk (2 * H3 * 16)

  __device__ inline void chunk_src(unsigned int i, const float*& src, unsigned int& bytes) const {
    const int c = (int)(i % (unsigned)chunks_per_tile);
    if (c < n2_chunks) { src = packed + (size_t)c * TC_SLOT_FLOATS; bytes = TC_SLOT_FLOATS * 4; }
    else { src = packed + (size_t)n2_chunks * TC_SLOT_FLOATS + (size_t)(c - n2_chunks) * n3_floats; bytes = n3_floats * 4; }
  }
  // thread 0: keep the TMA two chunks ahead of consumption
  __device__ inline void prefetch() {
    while (issued < total && issued < g + TC_NSLOT) {
      const unsigned int q = issued % TC_NSLOT;
      if (issued >= TC_NSLOT) acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm->empty[q]), ((issued / TC_NSLOT) - 1u) & 1u);
      const float* src; unsigned int bytes;
      chunk_src(issued, src, bytes);
      acx_mbar_expect_tx(reinterpret_cast<uint64_t*>(&sm->full[q]), bytes);
      acx_bulk_g2s(sm->slot[q], src, bytes, reinterpret_cast<uint64_t*>(&sm->full[q]));
      ++issued;
    }
  }
  // thread 0: consume one chunk = TC_KC input rows: 2 K-steps x 3 MMAs.
  //   a_hi/a_lo: start of these 16 k inside the A planes; n_out: 128 or 64; accumulate into tmem_d
  __device__ inline void mma_chunk(const float* a_hi, const float* a_lo, int n_out, uint32_t tmem_d, bool first) {
    prefetch();
    const unsigned int q = g % TC_NSLOT;
    acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm->full[q]), (g / TC_NSLOT) & 1u);
    fence_after_sync();
    const float* b_hi = sm->slot[q];
    const float* b_lo = b_hi + n_out * TC_KC;
    const uint32_t idesc = idesc_tf32(128, n_out, false, false);
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
      // A: [s/8][16 chunks of 4 k][8][4]: LBO = 128 B (next 4 k), SBO = 2048 B (next 8 samples); a K-step = 8 k = 256 B
      // B: [n/8][ 4 chunks of 4 k][8][4]: LBO = 128 B,            SBO =  512 B (next 8 outputs)
      const uint64_t ah = smem_desc(a_hi + ks * 64, 128, 2048), al = smem_desc(a_lo + ks * 64, 128, 2048);
      const uint64_t bh = smem_desc(b_hi + ks * 64, 128, 512), bl = smem_desc(b_lo + ks * 64, 128, 512);
      mma_tf32(tmem_d, al, bh, idesc, !(first && ks == 0));
      mma_tf32(tmem_d, ah, bl, idesc, true);
      mma_tf32(tmem_d, ah, bh, idesc, true);
    }
    commit(reinterpret_cast<uint64_t*>(&sm->empty[q]));
    ++g;
  }
};

// write 4 consecutive features (k0..k0+3, k0 % 4 == 0) of sample s into the hi/lo planes
__device__ inline void a_store4(float* hi_plane, float* lo_plane, int s, int k0, const float (&v)[4]) {
  float4 h, l;
  split_tf32(v[0], h.x, l.x); split_tf32(v[1], h.y, l.y); split_tf32(v[2], h.z, l.z); split_tf32(v[3], h.w, l.w);
  const int off = a_off(s, k0);
  *reinterpret_cast<float4*>(hi_plane + off) = h;
  *reinterpret_cast<float4*>(lo_plane + off) = l;
}

template <int H1, int H3, bool LRELU, int HEAD, int MODE>
__global__ void __launch_bounds__(TC_NT, 1) acx_mlp_tc_kernel(const Args A, const float* __restrict__ packed) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  TcSmem& sm = *reinterpret_cast<TcSmem*>(smem_raw);
  const AcxNetLayout L = A.L;
  const int ob = L.ob, nout = L.nout, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int s = (warp & 3) * 32 + lane;          // sample (TMEM lane) of this thread
  const int half = warp >> 2;                    // which half of the columns this thread's warpgroup handles
  const float* __restrict__ P = A.params;

  for (int i = tid; i < ob * H1; i += TC_NT) sm.W1[i] = P[L.W1 + i];
  for (int i = tid; i < H1; i += TC_NT) sm.b1[i] = P[L.b1 + i];
  for (int i = tid; i < H2; i += TC_NT) sm.b2[i] = P[L.b2 + i];
  for (int i = tid; i < H3; i += TC_NT) sm.b3[i] = P[L.b3 + i];
  for (int i = tid; i < H3 * nout; i += TC_NT) sm.Wh[i] = P[L.Wh + i];
  if (tid < nout) {
    sm.bh[tid] = P[L.bh + tid];
    if (HEAD == HEAD_GAUSS) sm.log_std[tid] = P[L.log_std + tid];
  }
  const int tpr = (A.n + MT - 1) / MT;
  const int ntiles = A.rows * tpr;
  const unsigned int mine = ((int)blockIdx.x < ntiles) ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

  TcPipe pipe;
  pipe.sm = &sm; pipe.packed = packed; pipe.g = 0; pipe.issued = 0;
  pipe.n2_chunks = H1 / TC_KC; pipe.n3_floats = 2 * H3 * TC_KC;
  pipe.chunks_per_tile = H1 / TC_KC + H2 / TC_KC;
  pipe.total = mine * pipe.chunks_per_tile;
  if (warp == 0) tmem_alloc(&sm.tmem_base, TC_TMEM_COLS);
  if (tid == 0) {
    for (int i = 0; i < TC_NSLOT; ++i) {
      acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.full[i]), 1);
      acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.empty[i]), 1);
    }
    acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.layer_done), 1);
    acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.sub_free[0]), 1);
    acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.sub_free[1]), 1);
    acx_mbar_fence_init();
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_d2 = sm.tmem_base, tmem_d3 = sm.tmem_base + 128;
  const uint32_t lane_sel = (uint32_t)((warp & 3) * 32) << 16;
  if (tid == 0) pipe.prefetch();

  const unsigned int step0 = A.step0 + ((MODE == MODE_SAMPLE && A.step_base) ? *A.step_base : 0u);
  double kl_sum = 0.0;
  unsigned int layer_uses = 0;            // completed waits on layer_done (parity)
  unsigned int sub_uses[2] = {0u, 0u};    // completed waits on sub_free[i]

  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int row = tile / tpr, n0 = (tile % tpr) * MT;
    const int valid = min(MT, A.n - n0);
    for (int idx = tid; idx < ob * MT; idx += TC_NT) {
      const int i = idx / MT, m = idx % MT;
      sm.obs[idx] = (m < valid) ? A.obs[((size_t)row * ob + i) * A.n + n0 + m] : 0.f;
    }
    __syncthreads();
    float o[MAX_OB];
#pragma unroll
    for (int i = 0; i < MAX_OB; ++i) o[i] = (i < ob) ? sm.obs[i * MT + s] : 0.f;

    // ---------------------------------------------- layers 1+2: X1 in 64-feature chunks -> sub-buffer (chunk & 1)
    constexpr int NX = H1 / 64;
    for (int xc = 0; xc < NX; ++xc) {
      const int sb = xc & 1;
      if (xc >= 2) {     // the MMAs that read this sub-buffer two chunks ago must have completed
        acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.sub_free[sb]), sub_uses[sb] & 1u);
        ++sub_uses[sb];
      }
      // this thread: 32 of the 64 features of the chunk, for its own sample
#pragma unroll 2
      for (int f4 = 0; f4 < 8; ++f4) {
        float v[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int kl = half * 32 + f4 * 4 + e, k = xc * 64 + kl;
          float acc = 0.f;
#pragma unroll
          for (int i = 0; i < MAX_OB; ++i)
            if (i < ob) acc = fmaf(o[i], sm.W1[i * H1 + k], acc);
          v[e] = act_fn<LRELU>(__fadd_rn(acc, sm.b1[k]));
        }
        a_store4(sm.A[sb][0], sm.A[sb][1], s, half * 32 + f4 * 4, v);
      }
      fence_proxy_async();
      __syncthreads();
      if (tid == 0) {
        fence_after_sync();
        for (int kc = 0; kc < 4; ++kc)
          pipe.mma_chunk(sm.A[sb][0] + kc * 128, sm.A[sb][1] + kc * 128, 128, tmem_d2, xc == 0 && kc == 0);
        if (xc + 2 < NX) commit(reinterpret_cast<uint64_t*>(&sm.sub_free[sb]));
        if (xc == NX - 1) commit(reinterpret_cast<uint64_t*>(&sm.layer_done));
      }
    }
    // ---------------------------------------------- layer-2 epilogue: X2 = act(D2 + b2) -> both sub-buffers
    acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.layer_done), layer_uses & 1u);
    ++layer_uses;
    fence_after_sync();
#pragma unroll 1
    for (int cb = 0; cb < 2; ++cb) {          // this thread's 64 columns: features half*64 + cb*32 ...
      float v[32];
      const int j0 = half * 64 + cb * 32;
      tmem_ld32(tmem_d2 + lane_sel + (uint32_t)j0, v);
#pragma unroll
      for (int f4 = 0; f4 < 8; ++f4) {
        float w[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) w[e] = act_fn<LRELU>(__fadd_rn(v[f4 * 4 + e], sm.b2[j0 + f4 * 4 + e]));
        a_store4(sm.A[half][0], sm.A[half][1], s, cb * 32 + f4 * 4, w);     // feature j -> sub-buffer j / 64
      }
    }
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    // ---------------------------------------------- layer 3
    if (tid == 0) {
      fence_after_sync();
      for (int kc = 0; kc < 8; ++kc)
        pipe.mma_chunk(sm.A[kc >> 2][0] + (kc & 3) * 128, sm.A[kc >> 2][1] + (kc & 3) * 128, H3, tmem_d3, kc == 0);
      commit(reinterpret_cast<uint64_t*>(&sm.layer_done));
    }
    acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.layer_done), layer_uses & 1u);
    ++layer_uses;
    fence_after_sync();

    // ---------------------------------------------- head (warpgroup 0, one thread per sample)
    if (half == 0) {
      float pre[MAX_OUT] = {0.f, 0.f, 0.f, 0.f};
      const bool basic = (HEAD == HEAD_VALUE) && A.basic_head;
#pragma unroll 1
      for (int cb = 0; cb < (basic ? H2 : H3) / 32; ++cb) {
        float v[32];
        tmem_ld32((basic ? tmem_d2 : tmem_d3) + lane_sel + (uint32_t)(cb * 32), v);
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const int k = cb * 32 + e;
          const float x = basic ? fmaxf(__fadd_rn(v[e], sm.b2[k]), 0.f) : act_fn<LRELU>(__fadd_rn(v[e], sm.b3[k]));
#pragma unroll
          for (int a = 0; a < MAX_OUT; ++a)
            if (a < nout) pre[a] = fmaf(x, sm.Wh[k * nout + a], pre[a]);
        }
      }
#pragma unroll
      for (int a = 0; a < MAX_OUT; ++a)
        if (a < nout) pre[a] = __fadd_rn(pre[a], sm.bh[a]);
      const int m = s;
      const bool live = m < valid;
      const size_t col = (size_t)n0 + m;
      const size_t s_idx = (size_t)row * A.n + col;
      if (live) {
        if (HEAD == HEAD_VALUE) {
          A.v_out[s_idx] = pre[0];
        } else if (HEAD == HEAD_GAUSS) {
          float lp = 0.f, ent = 0.f;
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
        }
      }
    }
    fence_before_sync();
    __syncthreads();      // D2/D3 fully read before the next tile's MMAs overwrite them; obs reusable
    fence_after_sync();
  }

  // ---------------------------------------------------- teardown + KL partial
  if (MODE == MODE_EVAL && HEAD != HEAD_VALUE && A.kl_partial != nullptr) {
    const double v = acx_warp_sum_d(kl_sum);
    if (lane == 0) sm.red[warp] = v;
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(sm.tmem_base, TC_TMEM_COLS);
  if (MODE == MODE_EVAL && HEAD != HEAD_VALUE && A.kl_partial != nullptr && tid == 0)
    A.kl_partial[blockIdx.x] = (sm.red[0] + sm.red[1]) + (sm.red[2] + sm.red[3]);
}

// packed[chunk][hi|lo][n/8][4][8][4]: chunk = 16 input rows; layer 2 (H1 x 128) first, then layer 3 (128 x H3)
__global__ void __launch_bounds__(256) acx_tc_pack_kernel(const float* __restrict__ P, AcxNetLayout L,
                                                          float* __restrict__ packed) {
  const int n2 = L.H1 * 128, n3 = 128 * L.H3;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n2 + n3; i += gridDim.x * blockDim.x) {
    const bool l2 = i < n2;
    const int e = l2 ? i : i - n2;
    const int N = l2 ? 128 : L.H3;
    const int k = e / N, n = e % N;                       // W[k][n]
    const float w = P[(l2 ? L.W2 : L.W3) + e];
    float hi, lo;
    split_tf32(w, hi, lo);
    const int chunk = k / TC_KC, kk = k % TC_KC;
    const size_t base = (l2 ? 0 : (size_t)2 * n2) + (size_t)chunk * 2 * N * TC_KC;
    const int off = ((n >> 3) * 4 + (kk >> 2)) * 32 + (n & 7) * 4 + (kk & 3);
    packed[base + off] = hi;
    packed[base + (size_t)N * TC_KC + off] = lo;
  }
}

template <int H1, int H3, bool LRELU, int HEAD, int MODE>
int launch_tc(const Args& a, const float* packed, int grid, cudaStream_t st) {
  auto kern = acx_mlp_tc_kernel<H1, H3, LRELU, HEAD, MODE>;
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
                                                  float* ent_out, float* mu_out, void* stream) {
  int rc = tc_check("acx_policy_forward_sample_gauss_tc", actor, packed, ob_dim, ac_dim, rows, n, obs);
  if (rc) return rc;
  if (rows == 0 || n == 0) return ACX_OK;
  ACX_REQUIRE(ac_out, "acx_policy_forward_sample_gauss_tc: ac_out is null");
  ACX_REQUIRE(eps || ac_dim <= 2, "acx_policy_forward_sample_gauss_tc: device RNG supports ac_dim <= 2 (got %d)", ac_dim);
  Args a; memset(&a, 0, sizeof(a));
  a.params = actor; a.L = acx_actor_layout(ob_dim, ac_dim, 1); a.rows = rows; a.n = n; a.obs = obs;
  a.ac_scale = ac_scale; a.noise = eps; a.seed = seed; a.index0 = index0; a.step0 = step0; a.step_base = step_base;
  a.ac_out_f = ac_out; a.logp_out = logp_out; a.ent_out = ent_out; a.head_out = mu_out;
  return launch_tc<128, 64, true, HEAD_GAUSS, MODE_SAMPLE>(a, packed, tc_grid(rows, n), (cudaStream_t)stream);
}

extern "C" int acx_policy_forward_sample_cat_tc(const float* actor, const float* packed, int ob_dim, int n_act,
                                                int rows, int n, const float* obs, const float* u, uint64_t seed,
                                                uint64_t index0, uint32_t step0, const uint32_t* step_base,
                                                int* ac_out, float* logp_out, float* ent_out, float* logits_out,
                                                void* stream) {
  int rc = tc_check("acx_policy_forward_sample_cat_tc", actor, packed, ob_dim, n_act, rows, n, obs);
  if (rc) return rc;
  if (rows == 0 || n == 0) return ACX_OK;
  ACX_REQUIRE(ac_out, "acx_policy_forward_sample_cat_tc: ac_out is null");
  Args a; memset(&a, 0, sizeof(a));
  a.params = actor; a.L = acx_actor_layout(ob_dim, n_act, 0); a.rows = rows; a.n = n; a.obs = obs;
  a.noise = u; a.seed = seed; a.index0 = index0; a.step0 = step0; a.step_base = step_base;
  a.ac_out_i = ac_out; a.logp_out = logp_out; a.ent_out = ent_out; a.head_out = logits_out;
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
  return launch_tc<128, 64, true, HEAD_CAT, MODE_EVAL>(a, packed, grid, (cudaStream_t)stream);
}

extern "C" int acx_value_forward_tc(const float* critic, const float* packed, int ob_dim, int rows, int n,
                                    const float* obs, float* v_out, int basic_head, void* stream) {
  int rc = tc_check("acx_value_forward_tc", critic, packed, ob_dim, 1, rows, n, obs);
  if (rc) return rc;
  if (rows == 0 || n == 0) return ACX_OK;
  ACX_REQUIRE(v_out, "acx_value_forward_tc: v_out is null");
  Args a; memset(&a, 0, sizeof(a));
  a.params = critic; a.L = acx_critic_layout(ob_dim); a.rows = rows; a.n = n; a.obs = obs;
  a.v_out = v_out; a.basic_head = basic_head;
  return launch_tc<256, 128, false, HEAD_VALUE, MODE_EVAL>(a, packed, tc_grid(rows, n), (cudaStream_t)stream);
}
