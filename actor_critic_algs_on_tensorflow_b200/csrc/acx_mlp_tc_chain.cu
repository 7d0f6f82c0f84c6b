// K4, chain part of the two-kernel tensor-core form (A3C/policies.py:74-85,152-158 on the cached activations):
// per 128-sample tile the head (loss terms, seed gradients, head-layer gradients), then back-propagation
//     dPre3 = (Wh dHead) * act'(X3)          dPre2 = (dPre3 W3^T) * act'(X2)          dPre1 = (dPre2 W2^T) * act'(X1)
// with both products on tcgen05 (3xTF32, A operand = dPre in tensor memory, B = packed weight chunks through a bulk-TMA
// ring, accumulators in tensor memory) and the first-layer gradients dW1 / db1.  dPre3 and dPre2 are streamed to HBM
// (tile-blocked like the activation cache) for acx_wgrad_kernel, which forms dW3 / dW2 / db3 / db2 from them.
//
// One persistent CTA per SM; warps 0-7 workers (thread = sample x half of the feature range = TMEM lane), warp 8 issues
// the MMAs, warp 9 feeds the weight-chunk ring.
//  * The activation cache and the dPre streams are tile-blocked ([tile][feature][128]): a tile is one contiguous block,
//    one DRAM page and TLB entry instead of `feature` rows 4 n bytes apart (worth 17 % of this kernel).  X3 tiles
//    arrive one tile ahead as ONE bulk-TMA copy each (two dense staging buffers); dPre3 and dPre2 are staged dense in
//    the tile's buffer and leave as one bulk-TMA store each, off the load/store pipe; after dPre2 the buffer is the
//    transposition staging of the dPre1 blocks (thread = sample writes columns, thread = neuron reads rows; padded
//    [feature][132] layout).  The A operand reaches the issuer in two halves of 64 reduction features.
//  * X2 is needed only as act'(X2): each thread reads its 64 values straight from the cache (coalesced) into two
//    32-bit masks, the loads issued a phase ahead of their use.
//  * act'(X1) is recomputed per element from the thread's own observation (ob_dim FMAs, layer1_block's arithmetic).
#include "acx_head.cuh"
#include "acx_tc_bwd.cuh"
#include "acx_wgrad.cuh"

using namespace acxmlp;
using namespace acxumma;
using namespace acxtcbwd;

#ifdef ACX_CHAIN_PROF
__device__ unsigned long long acx_chain_prof[16];
#define PROF_MARK(i) do { if (tid == 0) { const long long now_ = clock64(); prof_[i] += now_ - last_; last_ = now_; } } while (0)
#else
#define PROF_MARK(i) do { } while (0)
#endif

namespace {

struct TcSmem {
  float S[2][128 * MS];
  float slot[TT_NSLOT][TT_SLOT_FLOATS];
  float obs[MAX_OB * MT];
  float W1[MAX_OB * 256];
  float b1[256];
  float Wh[128 * MAX_OUT];
  float bh[MAX_OUT];
  float log_std[MAX_OUT];
  float dHead[MAX_OUT * MT];
  float4 W1p[256];             // ob_dim <= 3: {W1[0][k], W1[1][k], W1[2][k], b1[k]} -- one 16-byte load per neuron
  float l1p[128 * (MAX_OB + 1)];   // layer1 gradient partials of the second half of the samples
  double red[8][4];
  unsigned long long full[TT_NSLOT], empty[TT_NSLOT], a_ready[2], stage_done[3], bar_act[2];
  unsigned int tmem_base;
};

// MASKS: act'(X2) / act'(X1) come as the sign-mask words the forward kernels stored (A.mask_in) instead of being
// re-derived from the X2 cache (64 loads per thread) and the observations (ob_dim FMAs per element).
template <int H1, int H3, bool LRELU, int HEAD, bool MASKS>
__global__ void __launch_bounds__(TT_NT, 1)
acx_mlp_tc_chain_kernel(const Args A, const float* __restrict__ packed, float* __restrict__ d3_out,
                        float* __restrict__ d2_out) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  TcSmem& sm = *reinterpret_cast<TcSmem*>(smem_raw);
  constexpr int NB = H1 / 128, N3C = H3 / 16;
  const AcxNetLayout L = A.L;
  const int ob = L.ob, nout = L.nout, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const float* __restrict__ P = A.params;

  for (int i = tid; i < ob * H1; i += TT_NT) sm.W1[i] = P[L.W1 + i];
  for (int i = tid; i < H1; i += TT_NT) sm.b1[i] = P[L.b1 + i];
  for (int i = tid; i < H3 * nout; i += TT_NT) sm.Wh[i] = P[L.Wh + i];
  if (ob <= 3)
    for (int k = tid; k < H1; k += TT_NT)
      sm.W1p[k] = make_float4(P[L.W1 + k], ob > 1 ? P[L.W1 + H1 + k] : 0.f, ob > 2 ? P[L.W1 + 2 * H1 + k] : 0.f, P[L.b1 + k]);
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
    for (int i = 0; i < 2; ++i) acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.a_ready[i]), TT_WORK);
    for (int i = 0; i < 3; ++i) acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.stage_done[i]), 1);
    for (int i = 0; i < 2; ++i) acx_mbar_init(reinterpret_cast<uint64_t*>(&sm.bar_act[i]), 1);
    acx_mbar_fence_init();
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = sm.tmem_base;

  TtPipe<TcSmem> pipe;
  pipe.sm = &sm; pipe.packed = packed; pipe.g = 0;
  pipe.chunks_per_tile = N3C + 8 * NB;
  pipe.total = mine * pipe.chunks_per_tile;
  if (warp == 9) {
    // ================================================================ producer of the weight-chunk ring
    if (lane == 0) pipe.produce();
  } else if (warp == 8) {
    // ================================================================ issuer
    if (lane == 0) {
      // The A operand (dPre3, then dPre2) is handed over in halves of 64 reduction features, a_ready[0] then [1]: the
      // MMAs over the first half run while the workers still produce the second (H3 = 64: a single hand-over)
      unsigned int cnt[2] = {0u, 0u};
      constexpr int SUB3 = H3 / 64;
      for (unsigned int t = 0; t < mine; ++t) {
        for (int sub = 0; sub < SUB3; ++sub) {
          acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.a_ready[sub]), cnt[sub] & 1u); ++cnt[sub];
          fence_after_sync();
          for (int cc = sub * 4; cc < sub * 4 + 4; ++cc) pipe.mma_chunk(tmem, cc * 16, tmem + TT_D0, cc == 0);
        }
        commit(reinterpret_cast<uint64_t*>(&sm.stage_done[0]));
        // dX1 block 0 goes to D1: D0 still holds the half of dX2 the workers are turning into dPre2
        for (int sub = 0; sub < 2; ++sub) {
          acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.a_ready[sub]), cnt[sub] & 1u); ++cnt[sub];
          fence_after_sync();
          for (int cc = sub * 4; cc < sub * 4 + 4; ++cc) pipe.mma_chunk(tmem, cc * 16, tmem + TT_D1, cc == 0);
        }
        commit(reinterpret_cast<uint64_t*>(&sm.stage_done[1]));
        for (int h = 1; h < NB; ++h) {
          for (int cc = 0; cc < 8; ++cc) pipe.mma_chunk(tmem, cc * 16, tmem + TT_D0, cc == 0);
          commit(reinterpret_cast<uint64_t*>(&sm.stage_done[1 + h]));
        }
      }
    }
  } else {
    // ================================================================ workers
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
    // gradients this thread owns, summed over all of the CTA's tiles in registers and stored once at the end:
    // head weights (thread = feature k < H3; thread 128 + a: head bias a), layer 1 (thread = (neuron j, sample half))
    float gWh[MAX_OUT] = {0.f, 0.f, 0.f, 0.f}, gbh = 0.f;
    float gW1[NB][MAX_OB], gb1[NB];
#pragma unroll
    for (int h = 0; h < NB; ++h) {
      gb1[h] = 0.f;
#pragma unroll
      for (int i = 0; i < MAX_OB; ++i) gW1[h][i] = 0.f;
    }
    unsigned int act_phase[2] = {0u, 0u}, tiles_done = 0;
    if (tid < L.size - L.bh) scr[L.bh + tid] = 0.f;
    const size_t ld = (size_t)A.n;
    constexpr float SLOPE = LRELU ? 0.2f : 0.0f;

    // X3 tile -> staging buffer `buf` by bulk-TMA (full, 16-byte aligned tiles only)
    // the activation cache and the dPre streams are tile-blocked ([tile][feature][128], zeros beyond the batch): every
    // tile is whole, its rows 512 bytes apart
    // X3, dPre3 and dPre2 are staged DENSE ([feature][128], exactly the tile's block in HBM), so each moves with ONE
    // bulk copy (128 per-row copies cost 2-3 k cycles of TMA issue per tile and direction); only the dPre1 blocks,
    // which are read back by rows, use the padded [feature][132] layout in the same buffer
    auto issue_x3 = [&](int tile, int buf) {
      if (tid != 0) return;
      uint64_t* bar = reinterpret_cast<uint64_t*>(&sm.bar_act[buf]);
      acx_mbar_expect_tx(bar, (uint32_t)(H3 * MT * 4));
      acx_bulk_g2s(sm.S[buf], A.x3_in + cache_tile_off(tile, H3), H3 * MT * 4, bar);
    };
    if ((int)blockIdx.x < ntiles) issue_x3(blockIdx.x, 0);
    // act'(X2) of this thread's 64 features comes straight from the cache (coalesced loads) as two 32-bit masks.  The
    // loads are software-pipelined: the upper 32 of a tile are issued during the previous tile's layer-1 phase, the
    // lower 32 once dPre3 is handed over (they have until dX2 is done).
    float xm[32];
    auto load_x2 = [&](int tile, int q) {
      const float* g2 = A.x2_in + cache_tile_off(tile, H2) + (size_t)(q * 64 + half * 32) * MT + s;
      const bool ok = tile < ntiles;                 // (the cache holds zeros beyond the batch)
      // volatile: the loads must be ISSUED here (the compiler would otherwise sink them to the pack, next to their use)
#pragma unroll
      for (int e = 0; e < 32; ++e) {
        xm[e] = 0.f;
        if (ok) asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(xm[e]) : "l"(g2 + e * MT));
      }
    };
    auto pack_x2 = [&]() {
      unsigned int m = 0u;
#pragma unroll
      for (int e = 0; e < 32; ++e) m |= (xm[e] > 0.f ? 1u : 0u) << e;
      return m;
    };
    if (!MASKS) load_x2(blockIdx.x, 1);

    // staged dense tile (nrows x 128 floats) -> its block in HBM: one bulk-TMA store, off the load/store pipe (as plain
    // stores the 64 + 64 KB per tile queued in front of every later global load of the SM)
    auto bulk_rows_out = [&](const float* src, float* dst, int nrows) {
      if (tid == 0) {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(acx_smem_u32(src)),
                     "r"(nrows * MT * 4)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    };
    auto bulk_rows_read_done = [&]() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); };
#ifdef ACX_CHAIN_PROF
    long long prof_[16] = {0}, last_ = clock64();
#endif
    // Head of one tile: loss terms, seed gradients dHead (shared memory) and the head-layer gradients (registers) from the
    // tile's X3 block in staging buffer Sx.  It needs nothing from the tensor cores, so it runs one tile AHEAD: the
    // head of tile t+1 fills the wait for tile t's dX2 product (its X3 block was prefetched into the other buffer at the
    // start of tile t, and dHead of tile t has been consumed by then), its head-layer gradients the wait for dX1; only
    // the CTA's first tile pays for its head up front.
    auto do_head = [&](const float* __restrict__ Sx, int row_h, size_t col_h, bool live_h) {
      // ---------------------------------------------- head
      if (tid < MT) {
        float pre[MAX_OUT];
#pragma unroll
        for (int a = 0; a < MAX_OUT; ++a) pre[a] = 0.f;
        for (int k0 = 0; k0 < H3; k0 += 8) {        // head_pre's k-ascending FMA chain over the dense tile
          float x[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) x[u] = Sx[(k0 + u) * MT + tid];
          if (nout == 1) {                           // same chain, weights by 16-byte loads, no per-output predication
            const float4 w0 = *reinterpret_cast<const float4*>(&sm.Wh[k0]), w1 = *reinterpret_cast<const float4*>(&sm.Wh[k0 + 4]);
            float p = pre[0];
            p = fmaf(x[0], w0.x, p); p = fmaf(x[1], w0.y, p); p = fmaf(x[2], w0.z, p); p = fmaf(x[3], w0.w, p);
            p = fmaf(x[4], w1.x, p); p = fmaf(x[5], w1.y, p); p = fmaf(x[6], w1.z, p); p = fmaf(x[7], w1.w, p);
            pre[0] = p;
            continue;
          }
#pragma unroll
          for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int a = 0; a < MAX_OUT; ++a)
              if (a < nout) pre[a] = fmaf(x[u], sm.Wh[(k0 + u) * nout + a], pre[a]);
        }
#pragma unroll
        for (int a = 0; a < MAX_OUT; ++a)
          if (a < nout) pre[a] = __fadd_rn(pre[a], sm.bh[a]);
        float dh[MAX_OUT] = {0.f, 0.f, 0.f, 0.f};
        head_apply<HEAD, MODE_TRAIN>(A, sm.log_std, nout, pre, live_h, row_h, col_h, hs, lsum, dls, dh);
#pragma unroll
        for (int a = 0; a < MAX_OUT; ++a)
          if (a < nout) sm.dHead[a * MT + tid] = dh[a];
      }
      work_sync();
      PROF_MARK(1);
    };
    // ... and its head-layer gradients, which fill the wait for dX1 (they read the same X3 block and dHead)
    auto do_head_grads = [&](const float* __restrict__ Sx) {
      {   // head weights and bias.  Thread k < H3 owns row k of the dense tile; rows are 512 bytes apart, so every thread
          // starts at its own 16-byte column (k mod 32) and wraps: a warp's accesses are then 512 distinct bytes
        if (tid < H3) {
          for (int a = 0; a < nout; ++a) {
            float acc = 0.f;
#pragma unroll 8
            for (int m4 = 0; m4 < MT / 4; ++m4) {
              const int mm = (m4 + tid) & 31;
              const float4 x = *reinterpret_cast<const float4*>(&Sx[tid * MT + 4 * mm]);
              const float4 d = *reinterpret_cast<const float4*>(&sm.dHead[a * MT + 4 * mm]);
              acc = fmaf(x.x, d.x, acc); acc = fmaf(x.y, d.y, acc); acc = fmaf(x.z, d.z, acc); acc = fmaf(x.w, d.w, acc);
            }
            gWh[a] += acc;
          }
        } else if (tid >= 128 && tid < 128 + nout) {
          const int a = tid - 128;
          float acc = 0.f;
          for (int m = 0; m < MT; ++m) acc += sm.dHead[a * MT + m];
          gbh += acc;
        }
      }
      PROF_MARK(2);
      work_sync();                                  // head-layer gradients have read their X3 rows
    };
    if ((int)blockIdx.x < ntiles) {                 // the first tile's head
      const int t0 = blockIdx.x, n00 = (t0 % tpr) * MT;
      acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.bar_act[0]), act_phase[0] & 1u);
      ++act_phase[0];
      work_sync();
      do_head(sm.S[0], t0 / tpr, (size_t)n00 + s, s < min(MT, A.n - n00));
      do_head_grads(sm.S[0]);
    }
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++tiles_done) {
      const int row = tile / tpr, n0 = (tile % tpr) * MT;
      const int valid = min(MT, A.n - n0);
      const unsigned int par = tiles_done & 1u;
      const int buf = tiles_done & 1;
      float* __restrict__ S = sm.S[buf];
      const bool live = s < valid;
      const size_t col = (size_t)n0 + s;
      // the other buffer is idle since the end of the previous tile: fetch the next tile's X3 into it
      {
        const int next = tile + gridDim.x;
        if (next < ntiles) issue_x3(next, buf ^ 1);
        if (next < ntiles && tid < MT && (next % tpr) * MT + tid < A.n)
          head_train_prefetch<HEAD>(A, nout, next / tpr, (size_t)(next % tpr) * MT + tid);
      }
      unsigned int m2[2], m1[NB][2];
      if (MASKS) {      // this thread's mask words of the tile: X2 features gq*64 + half*32.., X1 features half*64 + gq*32..
        const unsigned int* mk = A.mask_in + mask_tile_off(tile, H1) + s;
#pragma unroll
        for (int gq = 0; gq < 2; ++gq) {
          m2[gq] = __ldg(mk + (H1 / 32 + gq * 2 + half) * MT);
#pragma unroll
          for (int h = 0; h < NB; ++h) m1[h][gq] = __ldg(mk + (h * 4 + half * 2 + gq) * MT);
        }
      } else {
        m2[1] = pack_x2();
      }
      PROF_MARK(12);

      // ---------------------------------------------- dPre3: A operand + HBM stream (staged in place over X3)
      {
        float dh[MAX_OUT];
#pragma unroll
        for (int a = 0; a < MAX_OUT; ++a) dh[a] = a < nout ? sm.dHead[a * MT + s] : 0.f;
#pragma unroll 1
        for (int gq = 0; gq < H3 / 64; ++gq) {
          const int j0 = gq * 64 + half * 32;
          float v[32];
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] = S[(j0 + e) * MT + s];
          if (nout == 1) {                           // one output: 16-byte weight loads, one product per feature
#pragma unroll
            for (int e = 0; e < 32; e += 4) {
              const float4 w = *reinterpret_cast<const float4*>(&sm.Wh[j0 + e]);
              v[e] = __fmul_rn(w.x, dh[0]) * act_grad<LRELU>(v[e]);
              v[e + 1] = __fmul_rn(w.y, dh[0]) * act_grad<LRELU>(v[e + 1]);
              v[e + 2] = __fmul_rn(w.z, dh[0]) * act_grad<LRELU>(v[e + 2]);
              v[e + 3] = __fmul_rn(w.w, dh[0]) * act_grad<LRELU>(v[e + 3]);
            }
          } else {
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              float acc = 0.f;
#pragma unroll
              for (int a = 0; a < MAX_OUT; ++a)
                if (a < nout) acc = fmaf(sm.Wh[(j0 + e) * nout + a], dh[a], acc);
              v[e] = acc * act_grad<LRELU>(v[e]);
            }
          }
          a_store32_tt(tl, j0, v);
#pragma unroll
          for (int e = 0; e < 32; ++e) S[(j0 + e) * MT + s] = v[e];
          tmem_wait_st();
          fence_before_sync();
          mbar_arrive1(&sm.a_ready[gq]);           // reduction features gq*64 .. +63 of the A operand are in place
        }
      }
      if (!MASKS) load_x2(tile, 0);                 // lower 32 act'(X2) values: in flight while dX2 is formed (not earlier:
                                                    // 32 more live registers across the head spill in the Gaussian kernel)
      fence_proxy_async();
      work_sync();
      bulk_rows_out(S, d3_out + cache_tile_off(tile, H3), H3);
      PROF_MARK(3);
      {   // the next tile's head, while the tensor cores form dX2 (dHead of this tile was consumed above)
        const int next = tile + gridDim.x;
        if (next < ntiles) {
          const int n0n = (next % tpr) * MT;
          acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.bar_act[buf ^ 1]), act_phase[buf ^ 1] & 1u);
          ++act_phase[buf ^ 1];
          PROF_MARK(13);
          do_head(sm.S[buf ^ 1], next / tpr, (size_t)n0n + s, s < min(MT, A.n - n0n));
        }
      }
      // ---------------------------------------------- dPre2 = dX2 * act'(X2): next A operand + HBM stream
      acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.stage_done[0]), par);
      fence_after_sync();
      PROF_MARK(4);
      bulk_rows_read_done();                        // dPre3 has left S: it takes dPre2 next
      work_sync();
      if (!MASKS) m2[0] = pack_x2();
#pragma unroll 1
      for (int gq = 0; gq < 2; ++gq) {
        const unsigned int bits = m2[gq];
        float v[32];
        const int k0 = gq * 64 + half * 32;
        tmem_ld32(tl + TT_D0 + (uint32_t)k0, v);
#pragma unroll
        for (int e = 0; e < 32; ++e) v[e] *= ((bits >> e) & 1u) ? 1.0f : SLOPE;
        a_store32_tt(tl, k0, v);
#pragma unroll
        for (int e = 0; e < 32; ++e) S[(k0 + e) * MT + s] = v[e];
        tmem_wait_st();
        fence_before_sync();
        mbar_arrive1(&sm.a_ready[gq]);
      }
      PROF_MARK(14);
      fence_proxy_async();
      work_sync();
      bulk_rows_out(S, d2_out + cache_tile_off(tile, H2), H2);
      PROF_MARK(15);
      if (tile + (int)gridDim.x < ntiles) do_head_grads(sm.S[buf ^ 1]);      // while the tensor cores form dX1
      if (!MASKS) load_x2(tile + gridDim.x, 1);     // next tile's upper 32 (zeros past the last tile)
      // this thread's observation (registers; also the shared copy layer1_grad reads), while dX1 is formed
      float ob_r[MAX_OB];
#pragma unroll
      for (int i = 0; i < MAX_OB; ++i) {
        ob_r[i] = (i < ob && live) ? __ldg(A.obs + ((size_t)row * ob + i) * ld + col) : 0.f;
        if (i < ob && half == 0) sm.obs[i * MT + s] = ob_r[i];
      }
      PROF_MARK(5);
      // ---------------------------------------------- layer 1: dPre1 = dX1 * act'(X1) per 128-row block, dW1, db1
#pragma unroll
      for (int h = 0; h < NB; ++h) {
        acx_mbar_wait(reinterpret_cast<uint64_t*>(&sm.stage_done[1 + h]), par);
        fence_after_sync();
        if (h == 0) {                               // dPre2 has left S: it takes the dPre1 blocks next
          bulk_rows_read_done();
          work_sync();
        }
        PROF_MARK(6 + 3 * h);
#pragma unroll 1
        for (int gq = 0; gq < 2; ++gq) {
          float v[32];
          const int k0 = half * 64 + gq * 32;
          tmem_ld32(tl + (h == 0 ? TT_D1 : TT_D0) + (uint32_t)k0, v);
          if (MASKS) {
            const unsigned int bits = m1[h][gq];
#pragma unroll
            for (int e = 0; e < 32; ++e) v[e] *= ((bits >> e) & 1u) ? 1.0f : SLOPE;
          } else if (ob <= 3) {
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              const float4 w = sm.W1p[h * 128 + k0 + e];
              float pre = fmaf(ob_r[0], w.x, 0.f);
              if (ob > 1) pre = fmaf(ob_r[1], w.y, pre);
              if (ob > 2) pre = fmaf(ob_r[2], w.z, pre);
              const float x1 = act_fn<LRELU>(__fadd_rn(pre, w.w));
              v[e] *= act_grad<LRELU>(x1);
            }
          } else {
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              const int k = h * 128 + k0 + e;
              float pre = 0.f;
#pragma unroll
              for (int i = 0; i < MAX_OB; ++i)
                if (i < ob) pre = fmaf(ob_r[i], sm.W1[i * H1 + k], pre);
              const float x1 = act_fn<LRELU>(__fadd_rn(pre, sm.b1[k]));
              v[e] *= act_grad<LRELU>(x1);
            }
          }
#pragma unroll
          for (int e = 0; e < 32; ++e) S[(k0 + e) * MS + s] = v[e];
        }
        fence_before_sync();
        work_sync();
        PROF_MARK(7 + 3 * h);
        {
          // dW1[i][128h + j] += sum_m obs[i][m] dPre1[j][m], db1 likewise: thread (j, sample half), in registers
          const int j = tid & 127, hs = tid >> 7;
          // two independent chains per sum (even / odd sample pairs): with one chain per input the 64 dependent FMAs of a
          // block were latency-bound (15 % of the critic kernel's warp samples, ncu)
          float sb[2] = {0.f, 0.f}, sw[MAX_OB][2];
#pragma unroll
          for (int i = 0; i < MAX_OB; ++i) sw[i][0] = sw[i][1] = 0.f;
#pragma unroll 4
          for (int m4 = 16 * hs; m4 < 16 * hs + 16; ++m4) {
            const float4 d = *reinterpret_cast<const float4*>(S + j * MS + 4 * m4);
            sb[0] += d.x; sb[1] += d.y; sb[0] += d.z; sb[1] += d.w;
#pragma unroll
            for (int i = 0; i < MAX_OB; ++i)
              if (i < ob) {
                const float4 o = *reinterpret_cast<const float4*>(&sm.obs[i * MT + 4 * m4]);
                sw[i][0] = fmaf(o.x, d.x, sw[i][0]); sw[i][1] = fmaf(o.y, d.y, sw[i][1]);
                sw[i][0] = fmaf(o.z, d.z, sw[i][0]); sw[i][1] = fmaf(o.w, d.w, sw[i][1]);
              }
          }
          gb1[h] += sb[0] + sb[1];
#pragma unroll
          for (int i = 0; i < MAX_OB; ++i)
            if (i < ob) gW1[h][i] += sw[i][0] + sw[i][1];
          fence_proxy_async();                      // S is next written by the async proxy (X3 prefetch)
          work_sync();
        }
        PROF_MARK(8 + 3 * h);
      }
    }

#ifdef ACX_CHAIN_PROF
    if (tid == 0 && blockIdx.x == 0)
      for (int i = 0; i < 16; ++i) acx_chain_prof[i] = (unsigned long long)prof_[i];
#endif
    // ---------------------------------------------------- the register-held gradients -> this CTA's scratch row
    if (tid < H3)
      for (int a = 0; a < nout; ++a) scr[L.Wh + tid * nout + a] = gWh[a];
    else if (tid >= 128 && tid < 128 + nout)
      scr[L.bh + tid - 128] = gbh;
    {
      const int j = tid & 127, hs = tid >> 7;
#pragma unroll
      for (int h = 0; h < NB; ++h) {
        work_sync();
        if (hs == 1) {                              // second sample half -> shared memory, first half adds and stores
          sm.l1p[j] = gb1[h];
#pragma unroll
          for (int i = 0; i < MAX_OB; ++i)
            if (i < ob) sm.l1p[(i + 1) * 128 + j] = gW1[h][i];
        }
        work_sync();
        if (hs == 0) {
          const int k = h * 128 + j;
          scr[L.b1 + k] = gb1[h] + sm.l1p[j];
#pragma unroll
          for (int i = 0; i < MAX_OB; ++i)
            if (i < ob) scr[L.W1 + i * H1 + k] = gW1[h][i] + sm.l1p[(i + 1) * 128 + j];
        }
      }
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

template <int H1, int H3, bool LRELU, int HEAD, bool MASKS>
int launch_chain1(const Args& a, const float* packed, float* d3, float* d2, int grid, cudaStream_t st) {
  auto kern = acx_mlp_tc_chain_kernel<H1, H3, LRELU, HEAD, MASKS>;
  ACX_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TcSmem) + 1024));
  kern<<<grid, TT_NT, sizeof(TcSmem) + 1024, st>>>(a, packed, d3, d2);
  ACX_LAUNCH_OK("acx_mlp_tc_chain_kernel");
  return ACX_OK;
}

template <int H1, int H3, bool LRELU, int HEAD>
int launch_chain(const Args& a, const float* packed, float* d3, float* d2, int grid, cudaStream_t st) {
  return a.mask_in != nullptr ? launch_chain1<H1, H3, LRELU, HEAD, true>(a, packed, d3, d2, grid, st)
                              : launch_chain1<H1, H3, LRELU, HEAD, false>(a, packed, d3, d2, grid, st);
}

}  // namespace

int acx_tc_chain_launch(const Args& a, const float* packed, float* d3, float* d2, int head, int grid, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  ACX_REQUIRE(a.x2_in && a.x3_in && packed && d3 && d2, "acx_tc_chain: null activation cache / packed weights / dPre buffer");
  ACX_REQUIRE(acx_aligned16(a.x2_in) && acx_aligned16(a.x3_in) && acx_aligned16(d3) && acx_aligned16(d2),
              "acx_tc_chain: the activation cache and the dPre arrays must be 16-byte aligned (bulk-TMA rows)");
  if (head == HEAD_GAUSS) return launch_chain<128, 64, true, HEAD_GAUSS>(a, packed, d3, d2, grid, st);
  if (head == HEAD_CAT) return launch_chain<128, 64, true, HEAD_CAT>(a, packed, d3, d2, grid, st);
  return launch_chain<256, 128, false, HEAD_VALUE>(a, packed, d3, d2, grid, st);
}

#ifdef ACX_CHAIN_PROF
extern "C" int acx_chain_prof_read(unsigned long long* out16) {
  return (int)cudaMemcpyFromSymbol(out16, acx_chain_prof, sizeof(unsigned long long) * 16);
}
#endif
