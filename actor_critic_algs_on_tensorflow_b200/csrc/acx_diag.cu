// Training diagnostics the reference logs every iteration (A3C/process.py:243,261,280-283), computed on the device
// from buffers the hot path already holds, so that the host never has to pull a rollout to print a log row:
//   * acx_episode_returns     avg_rew = mean return of the episodes that ended in the batch (process.py:234,243)
//   * acx_explained_variance  ev_after = var_accounted_for(target, pred) (util.py:5-13, process.py:261)
// (ev_before falls out of the scan kernel, acx_scan.cu.)  Also the phase stagger of the batched rollout.
#include "acx_common.cuh"

#define ACX_STREAM_PHASE 3u
constexpr int DIAG_MAX_BLOCKS = 1024;

static int diag_grid(long long n, int per_block) {
  long long g = (n + per_block - 1) / per_block;
  int cap = acx_num_sms() * 8;
  if (cap > DIAG_MAX_BLOCKS) cap = DIAG_MAX_BLOCKS;
  if (g > cap) g = cap;
  return g < 1 ? 1 : (int)g;
}

extern "C" size_t acx_diag_workspace_bytes(void) { return sizeof(AcxFoldWs) + sizeof(double) * 5 * DIAG_MAX_BLOCKS; }

// ------------------------------------------------------------------ episode returns
// One thread per env column walks the T rows forward: ret += rew; where done, the finished episode's return goes into
// the batch sums and ret restarts at 0.  ep_ret[n] carries the return of the episode still running across calls.  A NaN
// in ep_ret marks an episode whose beginning was not seen (a column that starts mid-episode): it is dropped, not
// counted, when it ends.
__global__ void __launch_bounds__(256) acx_episode_returns_kernel(int T, int N, const float* __restrict__ rew,
                                                                  const unsigned char* __restrict__ done,
                                                                  float* __restrict__ ep_ret, AcxFoldWs* ws,
                                                                  float* __restrict__ stats) {
  double m[2] = {0, 0};
  for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < N; n += gridDim.x * blockDim.x) {
    float ret = ep_ret[n];
    for (int t0 = 0; t0 < T; t0 += 8) {
      float r[8];
      unsigned char d[8];
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (t0 + i < T) {
          r[i] = __ldg(rew + (size_t)(t0 + i) * N + n);
          d[i] = __ldg(done + (size_t)(t0 + i) * N + n);
        }
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (t0 + i < T) {
          ret += r[i];
          if (d[i]) {
            if (ret == ret) { m[0] += (double)ret; m[1] += 1.0; }
            ret = 0.f;
          }
        }
    }
    ep_ret[n] = ret;
  }
  double t[2];
  if (!acx_fold_blocks<2>(m, ws, t)) return;
  stats[ACX_STAT_EP_COUNT] = (float)t[1];
  if (t[1] > 0) stats[ACX_STAT_EP_REW_MEAN] = (float)(t[0] / t[1]);      // no episode ended: keep the last mean
}

extern "C" int acx_episode_returns(int T, int n, const float* rew, const unsigned char* done, float* ep_ret,
                                   float* stats, void* workspace, size_t workspace_bytes, void* stream) {
  ACX_REQUIRE(T >= 0 && n >= 0, "acx_episode_returns: negative size (T=%d, n=%d)", T, n);
  if (T == 0 || n == 0) return ACX_OK;
  ACX_REQUIRE(rew && done && ep_ret && stats && workspace, "acx_episode_returns: null buffer");
  ACX_REQUIRE(workspace_bytes >= acx_diag_workspace_bytes() && acx_aligned16(workspace),
              "acx_episode_returns: workspace too small or unaligned (%zu < %zu)", workspace_bytes,
              acx_diag_workspace_bytes());
  acx_episode_returns_kernel<<<diag_grid(n, 256), 256, 0, (cudaStream_t)stream>>>(
      T, n, rew, done, ep_ret, reinterpret_cast<AcxFoldWs*>(workspace), stats);
  ACX_LAUNCH_OK("acx_episode_returns_kernel");
  return ACX_OK;
}

// ------------------------------------------------------------------ explained variance
__global__ void __launch_bounds__(256) acx_explained_variance_kernel(long long count, const float* __restrict__ target,
                                                                     const float* __restrict__ pred, AcxFoldWs* ws,
                                                                     float* __restrict__ out) {
  double m[5] = {0, 0, 0, 0, 0};
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
    const double x = (double)__ldg(target + i), y = (double)__ldg(pred + i);
    m[0] += x; m[1] += x * x; m[2] += y; m[3] += y * y; m[4] += x * y;
  }
  double t[5];
  if (!acx_fold_blocks<5>(m, ws, t)) return;
  *out = (float)acx_corr_from_sums(t[0], t[1], t[2], t[3], t[4], (double)count);
}

extern "C" int acx_explained_variance(long long count, const float* target, const float* pred, float* out,
                                      void* workspace, size_t workspace_bytes, void* stream) {
  ACX_REQUIRE(count >= 0, "acx_explained_variance: negative count");
  if (count == 0) return ACX_OK;
  ACX_REQUIRE(target && pred && out && workspace, "acx_explained_variance: null buffer");
  ACX_REQUIRE(workspace_bytes >= acx_diag_workspace_bytes() && acx_aligned16(workspace),
              "acx_explained_variance: workspace too small or unaligned (%zu < %zu)", workspace_bytes,
              acx_diag_workspace_bytes());
  acx_explained_variance_kernel<<<diag_grid(count, 1024), 256, 0, (cudaStream_t)stream>>>(
      count, target, pred, reinterpret_cast<AcxFoldWs*>(workspace), out);
  ACX_LAUNCH_OK("acx_explained_variance_kernel");
  return ACX_OK;
}

// ------------------------------------------------------------------ phase stagger
// N environments reset together reach their TimeLimit together: a 20-step batch then holds one phase of the episode,
// and (Pendulum never ends early) every done flag of the run falls into the same iteration.  The reference batches
// whole episodes, so each of its updates sees every phase.  This spreads the time-limit counters over [0, max_steps):
// elapsed[i] = Philox(index0 + i, step 0, stream 3) mod max_steps -- the state stays the fresh reset state, only the
// moment the limit fires moves.
__global__ void __launch_bounds__(256) acx_env_stagger_kernel(int n, int* __restrict__ elapsed, uint64_t seed,
                                                              uint64_t index0, int max_steps) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    elapsed[i] = (int)(acx_philox(index0 + (uint64_t)i, 0u, ACX_STREAM_PHASE, seed).x % (uint32_t)max_steps);
}

extern "C" int acx_env_stagger(int env, int n, int* elapsed, uint64_t seed, uint64_t index0, int max_steps,
                               void* stream) {
  ACX_REQUIRE(n >= 0, "acx_env_stagger: n must be >= 0 (got %d)", n);
  if (n == 0) return ACX_OK;
  ACX_REQUIRE(elapsed, "acx_env_stagger: null buffer");
  if (max_steps <= 0) {
    int rc = acx_env_dims(env, nullptr, nullptr, nullptr, &max_steps);
    if (rc != ACX_OK) return rc;
  }
  acx_env_stagger_kernel<<<diag_grid(n, 256), 256, 0, (cudaStream_t)stream>>>(n, elapsed, seed, index0, max_steps);
  ACX_LAUNCH_OK("acx_env_stagger_kernel");
  return ACX_OK;
}

// ------------------------------------------------------------------ per-iteration log ring
// The reference logs one row per iteration (A3C/process.py:280-281).  Reading the stats block back every iteration
// would put a host synchronisation into every step; instead each iteration leaves its stats and opt blocks in row
// (opt[ITER] - 1) mod rows of a device ring (opt[ITER] has just been advanced by acx_kl_lr_update), and the host reads
// the ring once every `rows` iterations or less.  Row layout: float[32] = stats[16] then opt[16].
__global__ void acx_log_push_kernel(float* ring, int rows, const float* __restrict__ stats,
                                    const float* __restrict__ opt) {
  const int it = (int)opt[ACX_OPT_ITER];
  const int row = ((it - 1) % rows + rows) % rows;
  if (threadIdx.x < ACX_STATS_SIZE) ring[row * 32 + threadIdx.x] = stats[threadIdx.x];
  else if (threadIdx.x < ACX_STATS_SIZE + ACX_OPT_SIZE) ring[row * 32 + threadIdx.x] = opt[threadIdx.x - ACX_STATS_SIZE];
}

extern "C" int acx_log_push(float* ring, int rows, const float* stats, const float* opt, void* stream) {
  ACX_REQUIRE(ring && stats && opt && rows >= 1, "acx_log_push: null buffer or rows < 1");
  acx_log_push_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(ring, rows, stats, opt);
  ACX_LAUNCH_OK("acx_log_push_kernel");
  return ACX_OK;
}
