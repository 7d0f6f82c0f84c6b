// Head math shared by every MLP kernel: turns the head pre-activations of ONE sample into actions /
// log-probs / entropies / values (A3C/policies.py:50-69,150) and, in MODE_TRAIN, into the per-sample loss
// terms and the seed gradients d loss / d pre (A3C/policies.py:74-80,152).  Follows oracle/policies.py.
#pragma once
#include "acx_mlp.cuh"

namespace acxmlp {

struct HeadScalars {          // per-kernel constants of the training head
  float beta, gam, adv_mean, adv_inv;
  double invB;
  unsigned int step0;
};

// MODE_TRAIN reads per-sample inputs (targets | adv, logp_old, actions) that nothing else touches: a kernel that knows
// its next tile can pull them towards the SM ahead of time
template <int HEAD>
__device__ inline void head_train_prefetch(const Args& A, int nout, int row, size_t col) {
  const size_t s_idx = (size_t)row * A.n + col;
  if (HEAD == HEAD_VALUE) {
    asm volatile("prefetch.global.L1 [%0];" ::"l"(A.targets + s_idx));
  } else {
    asm volatile("prefetch.global.L1 [%0];" ::"l"(A.adv + s_idx));
    asm volatile("prefetch.global.L1 [%0];" ::"l"(A.logp_old + s_idx));
    if (HEAD == HEAD_GAUSS) {
      for (int d = 0; d < nout; ++d)
        asm volatile("prefetch.global.L1 [%0];" ::"l"(A.ac_in_f + ((size_t)row * nout + d) * A.n + col));
    } else {
      asm volatile("prefetch.global.L1 [%0];" ::"l"(A.ac_in_i + s_idx));
    }
  }
}

// pre[a]: head pre-activation (bias added).  Outputs are written to the Args buffers at (row, col).
// lsum: {sum adv*lp | sum err^2, sum (lp_old-lp)^2, sum entropy, -}; dls: d loss / d log_std; dh: d loss / d pre.
template <int HEAD, int MODE>
__device__ inline void head_apply(const Args& A, const float* __restrict__ log_std, int nout, float (&pre)[MAX_OUT],
                                  bool live, int row, size_t col, const HeadScalars& hs, double (&lsum)[4],
                                  float (&dls)[MAX_OUT], float (&dh)[MAX_OUT]) {
  const size_t s_idx = (size_t)row * A.n + col;
  const unsigned int step0 = hs.step0;
  const float beta = hs.beta, gam = hs.gam, adv_mean = hs.adv_mean, adv_inv = hs.adv_inv;
  const double invB = hs.invB;
  if (HEAD == HEAD_VALUE) {
    if (live) {
      if (MODE != MODE_TRAIN) {
        A.v_out[s_idx] = pre[0];
      } else {
        const float err = pre[0] - A.targets[s_idx];
        lsum[0] += (double)err * (double)err;
        dh[0] = (float)(2.0 * (double)err * invB);
      }
    }
  } else if (HEAD == HEAD_GAUSS) {
    if (live) {
      float lp = 0.f, ent = 0.f, dlp = 0.f, dent = 0.f;
      float z[MAX_OUT], sd[MAX_OUT], th[MAX_OUT], ls[MAX_OUT];
      AcxRng4 g = {0u, 0u, 0u, 0u};
      if (MODE == MODE_SAMPLE && A.noise == nullptr)
        g = acx_philox(A.index0 + col, step0 + (unsigned)row, ACX_STREAM_ACTION, A.seed);
#pragma unroll
      for (int d = 0; d < MAX_OUT; ++d)
        if (d < nout) {
          th[d] = tanhf(pre[d]);
          const float mu = __fmul_rn(th[d], A.ac_scale);
          ls[d] = fminf(fmaxf(log_std[d], ACX_LOG_STD_MIN), ACX_LOG_STD_MAX);
          sd[d] = expf(ls[d]);
          const size_t a_idx = ((size_t)row * nout + d) * A.n + col;
          float ac;
          if (MODE == MODE_SAMPLE) {
            float eps;
            if (A.noise) eps = A.noise[a_idx];
            else eps = (d == 0) ? acx_normal(g.x, g.y) : ((d == 1) ? acx_normal(g.z, g.w) : 0.f);
            ac = __fadd_rn(mu, __fmul_rn(sd[d], eps));
            A.ac_out_f[a_idx] = ac;
          } else {
            ac = A.ac_in_f[a_idx];
          }
          if (A.head_out) A.head_out[a_idx] = mu;
          z[d] = __fdiv_rn(__fsub_rn(ac, mu), sd[d]);
          lp += __fsub_rn(__fsub_rn(__fmul_rn(__fmul_rn(-0.5f, z[d]), z[d]), ls[d]), ACX_HALF_LOG_2PI);
          ent += __fadd_rn(__fadd_rn(0.5f, ACX_HALF_LOG_2PI), ls[d]);
        }
      if (MODE != MODE_TRAIN) {
        if (A.logp_out) A.logp_out[s_idx] = lp;
        if (A.ent_out) A.ent_out[s_idx] = ent;
        if (MODE == MODE_EVAL && A.logp_old) {
          const double df = (double)A.logp_old[s_idx] - (double)lp;
          lsum[1] += df * df;
        }
      } else {
        const float advn = (A.adv[s_idx] - adv_mean) * adv_inv;
        const float diff = A.logp_old[s_idx] - lp;
        lsum[0] += (double)advn * (double)lp;
        lsum[1] += (double)diff * (double)diff;
        lsum[2] += (double)ent;
        dlp = (float)((-(double)advn - 2.0 * (double)beta * (double)diff) * invB);
        dent = (float)(-(double)gam * invB);
#pragma unroll
        for (int d = 0; d < MAX_OUT; ++d)
          if (d < nout) {
            const float dmu = dlp * z[d] / sd[d];
            dh[d] = dmu * A.ac_scale * (1.0f - th[d] * th[d]);
            const bool inside = log_std[d] >= ACX_LOG_STD_MIN && log_std[d] <= ACX_LOG_STD_MAX;
            if (inside) dls[d] += dlp * (z[d] * z[d] - 1.0f) + dent;
          }
      }
    }
  } else {   // HEAD_CAT
    if (live) {
      float l1[MAX_OUT], logps[MAX_OUT], p[MAX_OUT];
      float mx1 = -INFINITY, mx2 = -INFINITY;
#pragma unroll
      for (int a = 0; a < MAX_OUT; ++a)
        if (a < nout) {
          l1[a] = __fadd_rn(pre[a], 1e-8f);                 // policies.py:65
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
        // tf.multinomial: inverse CDF in double over exp(logit - max)
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
      if (MODE != MODE_TRAIN) {
        if (A.logp_out) A.logp_out[s_idx] = lp;
        if (A.ent_out) A.ent_out[s_idx] = ent;
        if (MODE == MODE_EVAL && A.logp_old) {
          const double df = (double)A.logp_old[s_idx] - (double)lp;
          lsum[1] += df * df;
        }
      } else {
        const float advn = (A.adv[s_idx] - adv_mean) * adv_inv;
        const float diff = A.logp_old[s_idx] - lp;
        lsum[0] += (double)advn * (double)lp;
        lsum[1] += (double)diff * (double)diff;
        lsum[2] += (double)ent;
        const float dlp = (float)((-(double)advn - 2.0 * (double)beta * (double)diff) * invB);
        const float dent = (float)(-(double)gam * invB);
        float dlogps[MAX_OUT], dpp[MAX_OUT], sdl = 0.f, sdp = 0.f;
#pragma unroll
        for (int a = 0; a < MAX_OUT; ++a)
          if (a < nout) {
            dlogps[a] = (a == ac ? dlp : 0.f) - dent * p[a];
            dpp[a] = -dent * logps[a];
            sdl += dlogps[a];
            sdp += dpp[a] * p[a];
          }
#pragma unroll
        for (int a = 0; a < MAX_OUT; ++a)
          if (a < nout) dh[a] = dlogps[a] - expf(logps[a]) * sdl + p[a] * (dpp[a] - sdp);
      }
    }
  }
}

}  // namespace acxmlp
