// K3 -- windowed discounted-return / advantage scan over the rollout horizon, with the
// batch moments needed for advantage normalisation.
//
// Replaces PathAdv.__call__ (A3C/process.py:54-66 == Basic_AC/run_AC.py:63-75) and the
// moments of A3C/process.py:247.  Buffers are time-major [T][N]; done[t][n] = 1 closes an
// episode segment (terminal, no bootstrap); the tail segment of a column is cut by the
// horizon and bootstraps from V[T] (the reference's max_path_length cut, process.py:63).
//
//   S_t   = r_t + gamma * (done_t ? 0 : S_{t+1})                    (segment suffix sum)
//   e_t   = exclusive end of the segment that contains t
//   t+L <  e_t :  G_t = S_t - gamma^L S_{t+L} + gamma^L V[t+L]      (window cut inside the segment)
//   t+L >= e_t :  G_t = S_t + (terminal ? 0 : gamma^(e_t - t) V[e_t])
//   adv_t = G_t - V[t]
//
// Two kernels compute the same thing:
//   * acx_scan_cols_kernel   one thread per env column, serial in time, loads coalesced over
//                            the env axis -- the HBM-bound path for wide batches (17 B/sample);
//   * acx_scan_warp_kernel   one warp per column, a warp-level segmented (affine) scan over
//                            time with shuffles -- for few, long columns (the reference's own
//                            shape: 1 env x ~1200 steps).
// Recurrences run in fp64 like the reference's numpy code.
#include <mutex>

#include "acx_common.cuh"

// per-block partial sums kept for the batch statistics: adv, adv^2 (advantage normalisation, A3C/process.py:247) and
// G, G^2, V, V^2, G*V over the T x N samples (ev_before = var_accounted_for(target = G, pred = V), A3C/util.py:5-13,
// A3C/process.py:111-116)
constexpr int SCAN_NMOM = 7;
constexpr int SCAN_MAX_BLOCKS = 2048;      // grid cap == partial[] capacity of acx_returns_adv_workspace_bytes()
typedef AcxFoldWs ScanWs;    // caller-provided workspace, see acx_returns_adv_workspace_bytes

__device__ inline void scan_finish_moments(const double (&m)[SCAN_NMOM], ScanWs* ws, float* stats, long long count) {
  double t[SCAN_NMOM];
  if (!acx_fold_blocks<SCAN_NMOM>(m, ws, t)) return;
  const double c = (double)count;
  const double mean = t[0] / c;
  double var = t[1] / c - mean * mean;
  if (var < 0) var = 0;
  stats[ACX_STAT_ADV_MEAN] = (float)mean;
  stats[ACX_STAT_ADV_INV_STD] = (float)(1.0 / (1e-8 + sqrt(var)));      // A3C/process.py:247
  stats[ACX_STAT_ADV_STD] = (float)sqrt(var);
  stats[ACX_STAT_EV_BEFORE] = (float)acx_corr_from_sums(t[2], t[3], t[4], t[5], t[6], c);
}

// the seven running sums of one sample
#define SCAN_ACC(m, a, g, v)                                                         \
  do {                                                                               \
    const double a__ = (double)(a), g__ = (double)(g), v__ = (double)(v);            \
    m[0] += a__; m[1] += a__ * a__; m[2] += g__; m[3] += g__ * g__; m[4] += v__;     \
    m[5] += v__ * v__; m[6] += g__ * v__;                                            \
  } while (0)

constexpr int SCAN_CHUNK = 8;

__global__ void __launch_bounds__(256) acx_scan_cols_kernel(int T, int N, const float* __restrict__ rew,
                                                            const unsigned char* __restrict__ done,
                                                            const float* __restrict__ V, double gamma, int L,
                                                            float* __restrict__ G, float* __restrict__ adv,
                                                            ScanWs* ws, float* __restrict__ stats) {
  double m[SCAN_NMOM] = {0, 0, 0, 0, 0, 0, 0};
  const double gL = pow(gamma, (double)L);
  for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < N; n += gridDim.x * blockDim.x) {
    const bool windowed = T > L;
    // ---- reverse pass: S_t (and, when the window can cut a segment, e_t)
    double carry = 0.0;           // S_{t+1} within the segment, 0 past its end
    double boot = (double)V[(size_t)T * N + n];   // gamma-discounted bootstrap carried along (plain mode)
    int e = T;
    for (int t1 = T; t1 > 0; t1 -= SCAN_CHUNK) {
      const int cnt = t1 < SCAN_CHUNK ? t1 : SCAN_CHUNK;
      float r[SCAN_CHUNK], v[SCAN_CHUNK];
      unsigned char d[SCAN_CHUNK];
#pragma unroll
      for (int i = 0; i < SCAN_CHUNK; ++i) {
        if (i < cnt) {
          const size_t off = (size_t)(t1 - 1 - i) * N + n;
          r[i] = __ldg(rew + off); d[i] = __ldg(done + off); v[i] = __ldg(V + off);
        }
      }
#pragma unroll
      for (int i = 0; i < SCAN_CHUNK; ++i) {
        if (i < cnt) {
          const int t = t1 - 1 - i;
          const size_t off = (size_t)t * N + n;
          if (d[i]) { carry = 0.0; boot = 0.0; e = t + 1; }
          if (!windowed) {
            // G_t = r_t + gamma * G_{t+1}, with G_{e} = V[e] for a horizon-cut tail, 0 after a terminal
            carry = (double)r[i] + gamma * (carry + boot);
            boot = 0.0;
            const float g = (float)carry, a = (float)(carry - (double)v[i]);
            G[off] = g; adv[off] = a;
            SCAN_ACC(m, a, g, v[i]);
          } else {
            carry = (double)r[i] + gamma * carry;
            G[off] = (float)carry;
            reinterpret_cast<int*>(adv)[off] = e;
          }
        }
      }
    }
    if (windowed) {
      // ---- forward fix-up, in place: S_{t+L} is read before its own slot is rewritten
      for (int t = 0; t < T; ++t) {
        const size_t off = (size_t)t * N + n;
        const int et = reinterpret_cast<const int*>(adv)[off];
        double g = (double)G[off];
        if (t + L < et) {
          g += gL * ((double)V[(size_t)(t + L) * N + n] - (double)G[(size_t)(t + L) * N + n]);
        } else {
          const bool terminal = done[(size_t)(et - 1) * N + n] != 0;
          if (!terminal) g += pow(gamma, (double)(et - t)) * (double)V[(size_t)et * N + n];
        }
        const float vt = V[off];
        const float a = (float)(g - (double)vt);
        G[off] = (float)g; adv[off] = a;
        SCAN_ACC(m, a, (float)g, vt);
      }
    }
  }
  scan_finish_moments(m, ws, stats, (long long)T * N);
}

// Plain (T <= L) column scan, VEC adjacent env columns per thread (VEC = 4: every row access of a warp is one 512-byte
// burst -- float4 rew / V / G / adv, uchar4 done; VEC = 1: 128-byte bursts, four times the threads).  Same arithmetic as
// acx_scan_cols_kernel.  CHUNK rows are loaded before the first is used.
//   <4, 4, 256>   wide batches (>= 2 CTAs per SM at 4 columns per thread): HBM-bound, 66 % (1M columns) to 78 % (8M) of the
//                 measured copy bandwidth (ncu dram bytes / duration, profiles/r02_ncu_hbm.md);
//   <1, 10, 128>  narrow ones.  At 65,536 columns 4 columns per thread are 16,384 threads = 3.5 warps per SM, and the
//                 kernel is bound by the instruction latency of those few warps (20 us for 22 MB); one column per thread
//                 gives 14 warps per SM and half a 20-row column's loads in flight per thread.
template <int VEC, int CHUNK, int BLOCK>
__global__ void __launch_bounds__(BLOCK) acx_scan_plain_kernel(int T, int N, const float* __restrict__ rew,
                                                               const unsigned char* __restrict__ done,
                                                               const float* __restrict__ V, double gamma,
                                                               float* __restrict__ G, float* __restrict__ adv,
                                                               ScanWs* ws, float* __restrict__ stats) {
  double m[SCAN_NMOM] = {0, 0, 0, 0, 0, 0, 0};
  const int NV = N / VEC;
  for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < NV; q += gridDim.x * blockDim.x) {
    const int n = q * VEC;
    // fp32 recurrence: T <= look_ahead steps of r + gamma * x carry ~T * 2^-24 of relative error, far inside the
    // 1e-5 budget, and keep the kernel on the HBM roofline (the fp64 pipe and its converts are ~10x slower)
    const float gf = (float)gamma, glo = (float)(gamma - (double)(float)gamma);   // gamma = gf + glo to ~2^-48
    float carry[VEC], boot[VEC];
    if (VEC == 4) {
      const float4 vT = __ldg(reinterpret_cast<const float4*>(V + (size_t)T * N + n));
      boot[0] = vT.x; boot[VEC > 1 ? 1 : 0] = vT.y; boot[VEC > 2 ? 2 : 0] = vT.z; boot[VEC > 3 ? 3 : 0] = vT.w;
    } else {
      boot[0] = __ldg(V + (size_t)T * N + n);
    }
#pragma unroll
    for (int c = 0; c < VEC; ++c) carry[c] = 0.f;
    for (int t1 = T; t1 > 0; t1 -= CHUNK) {
      const int cnt = t1 < CHUNK ? t1 : CHUNK;
      float r[CHUNK][VEC], v[CHUNK][VEC];
      unsigned char d[CHUNK][VEC];
#pragma unroll
      for (int i = 0; i < CHUNK; ++i)
        if (i < cnt) {
          const size_t off = (size_t)(t1 - 1 - i) * N + n;
          if (VEC == 4) {
            const float4 r4 = __ldg(reinterpret_cast<const float4*>(rew + off));
            const float4 v4 = __ldg(reinterpret_cast<const float4*>(V + off));
            const uchar4 d4 = __ldg(reinterpret_cast<const uchar4*>(done + off));
            r[i][0] = r4.x; r[i][VEC > 1 ? 1 : 0] = r4.y; r[i][VEC > 2 ? 2 : 0] = r4.z; r[i][VEC > 3 ? 3 : 0] = r4.w;
            v[i][0] = v4.x; v[i][VEC > 1 ? 1 : 0] = v4.y; v[i][VEC > 2 ? 2 : 0] = v4.z; v[i][VEC > 3 ? 3 : 0] = v4.w;
            d[i][0] = d4.x; d[i][VEC > 1 ? 1 : 0] = d4.y; d[i][VEC > 2 ? 2 : 0] = d4.z; d[i][VEC > 3 ? 3 : 0] = d4.w;
          } else {
            r[i][0] = __ldg(rew + off); v[i][0] = __ldg(V + off); d[i][0] = __ldg(done + off);
          }
        }
      float p[SCAN_NMOM] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};      // this chunk's CHUNK x VEC samples, fp32
#pragma unroll
      for (int i = 0; i < CHUNK; ++i)
        if (i < cnt) {
          const size_t off = (size_t)(t1 - 1 - i) * N + n;
          float g[VEC], a[VEC];
#pragma unroll
          for (int c = 0; c < VEC; ++c) {
            if (d[i][c]) { carry[c] = 0.f; boot[c] = 0.f; }
            const float x = carry[c] + boot[c];
            carry[c] = fmaf(gf, x, fmaf(glo, x, r[i][c]));
            boot[c] = 0.f;
            g[c] = carry[c];
            a[c] = carry[c] - v[i][c];
            p[0] += a[c]; p[1] = fmaf(a[c], a[c], p[1]);
            p[2] += g[c]; p[3] = fmaf(g[c], g[c], p[3]);
            p[4] += v[i][c]; p[5] = fmaf(v[i][c], v[i][c], p[5]);
            p[6] = fmaf(g[c], v[i][c], p[6]);
          }
          if (VEC == 4) {
            *reinterpret_cast<float4*>(G + off) = make_float4(g[0], g[VEC > 1 ? 1 : 0], g[VEC > 2 ? 2 : 0], g[VEC > 3 ? 3 : 0]);
            *reinterpret_cast<float4*>(adv + off) = make_float4(a[0], a[VEC > 1 ? 1 : 0], a[VEC > 2 ? 2 : 0], a[VEC > 3 ? 3 : 0]);
          } else {
            G[off] = g[0]; adv[off] = a[0];
          }
        }
#pragma unroll
      for (int k = 0; k < SCAN_NMOM; ++k) m[k] += (double)p[k];
    }
  }
  scan_finish_moments(m, ws, stats, (long long)T * N);
}

// affine element of the reverse recurrence x_t = b + a * x_{t+1}; `e` rides along as a min-scan
struct Aff { double a, b; int e; };
__device__ inline Aff aff_combine(const Aff& lo, const Aff& hi) {   // lo is earlier in time, applied last
  Aff o; o.a = lo.a * hi.a; o.b = lo.b + lo.a * hi.b; o.e = lo.e < hi.e ? lo.e : hi.e;
  return o;
}

__global__ void __launch_bounds__(256) acx_scan_warp_kernel(int T, int N, const float* __restrict__ rew,
                                                            const unsigned char* __restrict__ done,
                                                            const float* __restrict__ V, double gamma, int L,
                                                            float* __restrict__ G, float* __restrict__ adv,
                                                            ScanWs* ws, float* __restrict__ stats) {
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  double m[SCAN_NMOM] = {0, 0, 0, 0, 0, 0, 0};
  const double gL = pow(gamma, (double)L);
  const bool windowed = T > L;
  for (int n = blockIdx.x * warps_per_block + (threadIdx.x >> 5); n < N; n += gridDim.x * warps_per_block) {
    // carry from later time: value of S (or G) just after the current 32-step block, and its segment end
    double carry = windowed ? 0.0 : (double)V[(size_t)T * N + n];
    int carry_e = T;
    for (int hi = T; hi > 0; hi -= 32) {
      const int t = hi - 1 - lane;            // lane 0 holds the latest step of the block
      Aff x; x.a = 1.0; x.b = 0.0; x.e = 0x7fffffff;
      float v = 0.f;
      if (t >= 0) {
        const size_t off = (size_t)t * N + n;
        const bool d = done[off] != 0;
        x.a = d ? 0.0 : gamma; x.b = (double)rew[off]; x.e = d ? t + 1 : 0x7fffffff;
        v = V[off];
      }
      // inclusive scan toward later time: after it, lane k holds steps t_k .. hi-1 composed
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        Aff y;
        y.a = __shfl_up_sync(0xffffffffu, x.a, o);
        y.b = __shfl_up_sync(0xffffffffu, x.b, o);
        y.e = __shfl_up_sync(0xffffffffu, x.e, o);
        if (lane >= o) x = aff_combine(x, y);
      }
      const double val = x.b + x.a * carry;
      const int e = x.e < carry_e ? x.e : carry_e;
      if (t >= 0) {
        const size_t off = (size_t)t * N + n;
        if (!windowed) {
          const float a = (float)(val - (double)v);
          G[off] = (float)val; adv[off] = a;
          SCAN_ACC(m, a, (float)val, v);
        } else {
          G[off] = (float)val;
          reinterpret_cast<int*>(adv)[off] = e;
        }
      }
      // new carry = value at the earliest step of this block (highest active lane)
      const int src = (hi >= 32) ? 31 : hi - 1;
      carry = __shfl_sync(0xffffffffu, val, src);
      carry_e = __shfl_sync(0xffffffffu, e, src);
    }
    if (windowed) {
      __syncwarp();
      __threadfence_block();
      // S_{t+L} must be read for the whole 32-step block before any slot of it is rewritten
      for (int lo = 0; lo < T; lo += 32) {
        const int t = lo + lane;
        double g = 0; float a = 0, vt = 0; bool live = t < T;
        if (live) {
          const size_t off = (size_t)t * N + n;
          const int et = reinterpret_cast<const int*>(adv)[off];
          g = (double)G[off];
          if (t + L < et) {
            g += gL * ((double)V[(size_t)(t + L) * N + n] - (double)G[(size_t)(t + L) * N + n]);
          } else {
            const bool terminal = done[(size_t)(et - 1) * N + n] != 0;
            if (!terminal) g += pow(gamma, (double)(et - t)) * (double)V[(size_t)et * N + n];
          }
          vt = V[off];
          a = (float)(g - (double)vt);
        }
        __syncwarp();
        if (live) {
          const size_t off = (size_t)t * N + n;
          G[off] = (float)g; adv[off] = a;
          SCAN_ACC(m, a, (float)g, vt);
        }
        __syncwarp();
        __threadfence_block();
      }
    }
  }
  scan_finish_moments(m, ws, stats, (long long)T * N);
}

static int scan_grid(int N, int per_block) {
  int g = (N + per_block - 1) / per_block;
  int cap = acx_num_sms() * 8;
  if (cap > SCAN_MAX_BLOCKS) cap = SCAN_MAX_BLOCKS;      // what the workspace has room for, on any part
  if (g > cap) g = cap;
  return g < 1 ? 1 : g;
}

extern "C" size_t acx_returns_adv_workspace_bytes(int T, int n) {
  (void)T; (void)n;
  return sizeof(ScanWs) + sizeof(double) * SCAN_NMOM * (size_t)SCAN_MAX_BLOCKS;
}

static int scan_launch(int T, int n, const float* rew, const unsigned char* done, const float* V, double gamma,
                       int look_ahead, float* G_out, float* adv_out, float* stats_out, ScanWs* ws, int mode,
                       int max_blocks, cudaStream_t st) {
  auto grid = [&](int cols, int per_block) {
    const int g = scan_grid(cols, per_block);
    return g > max_blocks ? max_blocks : g;
  };
  if (mode == ACX_SCAN_WARP)
    acx_scan_warp_kernel<<<grid(n, 8), 256, 0, st>>>(T, n, rew, done, V, gamma, look_ahead, G_out, adv_out, ws,
                                                     stats_out);
  else if (T <= look_ahead && (n & 3) == 0 && acx_aligned16(rew) && acx_aligned16(V) && acx_aligned16(G_out) &&
           acx_aligned16(adv_out) && (reinterpret_cast<uintptr_t>(done) & 3u) == 0 && n / 4 >= 2 * 256 * acx_num_sms()) {
    acx_scan_plain_kernel<4, 4, 256><<<grid(n / 4, 256), 256, 0, st>>>(T, n, rew, done, V, gamma, G_out, adv_out, ws,
                                                                       stats_out);
  } else if (T <= look_ahead) {
    acx_scan_plain_kernel<1, 10, 128><<<grid(n, 128), 128, 0, st>>>(T, n, rew, done, V, gamma, G_out, adv_out, ws,
                                                                    stats_out);
  } else
    acx_scan_cols_kernel<<<grid(n, 256), 256, 0, st>>>(T, n, rew, done, V, gamma, look_ahead, G_out, adv_out, ws,
                                                       stats_out);
  ACX_LAUNCH_OK("acx_scan kernel");
  return ACX_OK;
}

extern "C" int acx_returns_adv_scan(int T, int n, const float* rew, const unsigned char* done, const float* V,
                                    double gamma, int look_ahead, float* G_out, float* adv_out, float* stats_out,
                                    void* workspace, size_t workspace_bytes, int mode, void* stream) {
  ACX_REQUIRE(T >= 0 && n >= 0, "acx_returns_adv_scan: negative size (T=%d, n=%d)", T, n);
  ACX_REQUIRE(look_ahead >= 1, "acx_returns_adv_scan: look_ahead must be >= 1 (got %d)", look_ahead);
  if (T == 0 || n == 0) return ACX_OK;
  ACX_REQUIRE(rew && done && V && G_out && adv_out && stats_out && workspace, "acx_returns_adv_scan: null buffer");
  ACX_REQUIRE(workspace_bytes >= acx_returns_adv_workspace_bytes(T, n),
              "acx_returns_adv_scan: workspace too small (%zu < %zu)", workspace_bytes,
              acx_returns_adv_workspace_bytes(T, n));
  ACX_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 15u) == 0, "acx_returns_adv_scan: workspace must be 16-byte aligned");
  if (mode == ACX_SCAN_AUTO) mode = (n < 2048 && T >= 64) ? ACX_SCAN_WARP : ACX_SCAN_COLS;
  ACX_REQUIRE(mode == ACX_SCAN_COLS || mode == ACX_SCAN_WARP, "acx_returns_adv_scan: bad mode %d", mode);
  return scan_launch(T, n, rew, done, V, gamma, look_ahead, G_out, adv_out, stats_out,
                     reinterpret_cast<ScanWs*>(workspace), mode, SCAN_MAX_BLOCKS, (cudaStream_t)stream);
}

// ------------------------------------------------------------------ allocation-free variant
// Same kernels, but the partial-moment scratch is a fixed array that lives in the library's
// device image instead of a caller workspace.  Every GPU that loads libacx.so carries these
// ACX_SCAN_STATIC_SCRATCH_BYTES until the process exits.
#define ACX_SCAN_STATIC_MAX_BLOCKS 512
#define ACX_SCAN_STATIC_SCRATCH_BYTES (16 + 8 * 7 * ACX_SCAN_STATIC_MAX_BLOCKS)
static __device__ __align__(256) unsigned char g_scan_static_scratch[ACX_SCAN_STATIC_SCRATCH_BYTES];

// In host code `g_scan_static_scratch` is only a handle for the runtime, not an address a kernel can
// use: the device address has to be asked for, once per device, while that device is current.  The
// look-up is serialised (two host threads may make their first call at the same time); the cached
// address belongs to the device's primary context and is stale after cudaDeviceReset().
static int scan_static_scratch(ScanWs** out) {
  static std::mutex mu;
  static void* addr[64] = {nullptr};
  int dev = 0;
  ACX_CUDA_OK(cudaGetDevice(&dev));
  ACX_REQUIRE(dev >= 0 && dev < 64, "acx_returns_adv_scan_static: device ordinal %d not supported (max 63)", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (addr[dev] == nullptr) {
    void* p = nullptr;
    ACX_CUDA_OK(cudaGetSymbolAddress(&p, g_scan_static_scratch));
    ACX_REQUIRE(p != nullptr, "acx_returns_adv_scan_static: no device address for the static scratch");
    addr[dev] = p;
  }
  *out = reinterpret_cast<ScanWs*>(addr[dev]);
  return ACX_OK;
}

extern "C" int acx_returns_adv_scan_static(int T, int n, const float* rew, const unsigned char* done, const float* V,
                                           double gamma, int look_ahead, float* G_out, float* adv_out,
                                           float* stats_out, int mode, void* stream) {
  ACX_REQUIRE(T >= 0 && n >= 0, "acx_returns_adv_scan_static: negative size (T=%d, n=%d)", T, n);
  ACX_REQUIRE(look_ahead >= 1, "acx_returns_adv_scan_static: look_ahead must be >= 1 (got %d)", look_ahead);
  ACX_REQUIRE(n <= ACX_SCAN_STATIC_MAX_COLUMNS,
              "acx_returns_adv_scan_static: n = %d env columns exceeds the built-in scratch (maximum %d); use "
              "acx_returns_adv_scan with a caller workspace for wider batches", n, ACX_SCAN_STATIC_MAX_COLUMNS);
  if (T == 0 || n == 0) return ACX_OK;
  ACX_REQUIRE(rew && done && V && G_out && adv_out && stats_out, "acx_returns_adv_scan_static: null buffer");
  if (mode == ACX_SCAN_AUTO) mode = (n < 2048 && T >= 64) ? ACX_SCAN_WARP : ACX_SCAN_COLS;
  ACX_REQUIRE(mode == ACX_SCAN_COLS || mode == ACX_SCAN_WARP, "acx_returns_adv_scan_static: bad mode %d", mode);
  ScanWs* ws = nullptr;
  int rc = scan_static_scratch(&ws);
  if (rc != ACX_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // the array keeps whatever the previous call left: the "last block" ticket has to start at zero
  ACX_CUDA_OK(cudaMemsetAsync(ws, 0, 16, st));
  int grid = mode == ACX_SCAN_WARP ? (n + 7) / 8 : (n + 255) / 256;
  if (grid > ACX_SCAN_STATIC_MAX_BLOCKS) grid = ACX_SCAN_STATIC_MAX_BLOCKS;
  if (mode == ACX_SCAN_WARP)
    acx_scan_warp_kernel<<<grid, 256, 0, st>>>(T, n, rew, done, V, gamma, look_ahead, G_out, adv_out, ws, stats_out);
  else
    acx_scan_cols_kernel<<<grid, 256, 0, st>>>(T, n, rew, done, V, gamma, look_ahead, G_out, adv_out, ws, stats_out);
  ACX_LAUNCH_OK("acx_scan kernel (static scratch)");
  return ACX_OK;
}
