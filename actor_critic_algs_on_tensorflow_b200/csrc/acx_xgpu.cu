// The exchange step of a synchronous multi-GPU update fused with the optimiser step, over NVLink peer memory:
//
//     ONE kernel =  sum over ranks of [flat gradient ‖ previous update's KL]   (replaces the workers' pushes onto the
//                   ps variables and the pull, A3C/policies.py:88-98, A3C/process.py:257-260)
//                +  the KL-driven learning-rate rule on the rank-averaged KL   (A3C/process.py:262-270)
//                +  clip, Adam on the rank-averaged gradient                   (A3C/policies.py:85,88-89,160-161)
//
// Every rank keeps its [gradient ‖ tail ‖ flags] block in a buffer that all ranks of the node have mapped (the caller
// maps it: torch symmetric memory, CUDA IPC, cuMemMap -- the library allocates nothing).  Block b of the grid owns
// elements [1024 b, 1024 b + 1024) on EVERY rank, so all synchronisation is between same-numbered blocks:
//
//   entry   block b of rank r stores its launch generation g into slot in[b][r] of every rank (st.release.sys): "my
//           gradient is complete" -- it is, the kernels that wrote it precede this one on the stream; then it waits
//           until its own in[b][*] all hold g (ld.acquire.sys);
//   reduce  thread i adds element i of ranks 0, 1, ..., W-1 in that fixed order (plain 16-byte loads over NVLink,
//           L1 bypassed): every rank forms bit-identical sums, so replicated parameters stay bit-identical;
//   exit    the block tells every rank "I have read your block b" (out[b][r] = g), runs the Adam step on its 1024
//           elements from registers, and only then waits for its own out[b][*]: nobody overwrites a gradient a peer
//           is still reading, and the wait hides behind the optimiser arithmetic.
//
// No NCCL call, no host round trip, graph-capturable (the generation lives in device memory).  Loads from peer memory
// are plain ld.global here: TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load
// from another GPU's memory hung the GPU.
//
// A rank that never arrives would leave the others spinning: every wait is bounded by wall time (ACX_XGPU_TIMEOUT_NS,
// 10 s), then the kernel reports which rank it waited for and traps (a launch error on the host, not a hung GPU).
#include "acx_optim.cuh"

#ifndef ACX_XGPU_TIMEOUT_NS
#define ACX_XGPU_TIMEOUT_NS 10000000000ull
#endif
#define XG_THREADS 256

namespace {

struct XgArgs {
  const float* peer[ACX_XGPU_MAX_RANKS];   // rank r's exchange buffer as mapped on THIS device (peer[rank] is local)
  int world, rank;
  int n4a, n4c, n4t;                       // float4 counts: actor gradient, critic gradient, tail
  size_t flag_off;                         // uint32 offset of the flags within a buffer
  float *actor, *ma, *va, *critic, *mc, *vc;
  const float* opt;
  float* grad_sum;                         // local, [4 (n4a + n4c + n4t)]: the summed block (nullable)
  uint32_t* gen;                           // local, [ACX_XGPU_MAX_BLOCKS]
  float scale;
  int apply_lr_rule;
};

__device__ inline void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ inline uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ inline float4 ld_peer4(const float* p) {      // system-scope, never served from this SM's L1
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ inline float ld_peer1(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ inline uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// polls with relaxed loads (an acquire load invalidates L1 on every poll), then ONE acquire load orders what follows
__device__ inline void wait_flag(const uint32_t* p, uint32_t g, int my_rank, int other, int what) {
  if ((int32_t)(ld_relaxed_sys(p) - g) < 0) {
    const uint64_t t0 = acx_globaltimer();
    for (bool seen = false; !seen;) {
      for (int spin = 0; spin < 64 && !seen; ++spin) seen = (int32_t)(ld_relaxed_sys(p) - g) >= 0;
      if (!seen && acx_globaltimer() - t0 > ACX_XGPU_TIMEOUT_NS) {
        printf("acx_allreduce_adam: rank %d block %d waited %llu s for rank %d (%s, generation %u, flag %u)\n", my_rank,
               (int)blockIdx.x, (unsigned long long)(ACX_XGPU_TIMEOUT_NS / 1000000000ull), other,
               what ? "exit: peer has not read this rank's gradient" : "entry: peer's gradient not ready", g,
               ld_relaxed_sys(p));
        __trap();
      }
    }
  }
  (void)ld_acquire_sys(p);
}

__global__ void __launch_bounds__(XG_THREADS) acx_allreduce_adam_kernel(const XgArgs a) {
  __shared__ uint32_t g_sh;
  __shared__ float kl_sh[ACX_XGPU_MAX_RANKS];
  const int tid = threadIdx.x, b = blockIdx.x, W = a.world;
  if (tid == 0) g_sh = ++a.gen[b];
  __syncthreads();
  const uint32_t g = g_sh;
  const size_t in_slot = a.flag_off + (size_t)b * ACX_XGPU_MAX_RANKS;
  const size_t out_slot = a.flag_off + (size_t)(ACX_XGPU_MAX_BLOCKS + b) * ACX_XGPU_MAX_RANKS;
  uint32_t* mine = reinterpret_cast<uint32_t*>(const_cast<float*>(a.peer[a.rank]));

  // ---- entry: every rank's gradient is complete
  if (tid < W) {
    st_release_sys(reinterpret_cast<uint32_t*>(const_cast<float*>(a.peer[tid])) + in_slot + a.rank, g);
    wait_flag(mine + in_slot + tid, g, a.rank, tid, 0);
  }
  __syncthreads();

  // ---- reduce: all W loads of a thread are issued before the first add (one NVLink round trip, not W), the adds run
  // in the fixed rank order
  const int n4p = a.n4a + a.n4c, n4 = n4p + a.n4t;
  const int i = b * XG_THREADS + tid;
  const bool rule = a.apply_lr_rule && a.n4t > 0;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  if (i < n4) {
    float4 v[ACX_XGPU_MAX_RANKS];
#pragma unroll
    for (int r = 0; r < ACX_XGPU_MAX_RANKS; ++r)
      if (r < W) v[r] = ld_peer4(a.peer[r] + 4 * (size_t)i);
    if (rule && tid < W) kl_sh[tid] = ld_peer1(a.peer[tid] + 4 * (size_t)n4p);      // tail word 0 of rank tid
#pragma unroll
    for (int r = 0; r < ACX_XGPU_MAX_RANKS; ++r)
      if (r < W) {
        s.x = __fadd_rn(s.x, v[r].x); s.y = __fadd_rn(s.y, v[r].y);
        s.z = __fadd_rn(s.z, v[r].z); s.w = __fadd_rn(s.w, v[r].w);
      }
    if (a.grad_sum != nullptr) reinterpret_cast<float4*>(a.grad_sum)[i] = s;
  } else if (rule && tid < W) {
    kl_sh[tid] = ld_peer1(a.peer[tid] + 4 * (size_t)n4p);
  }
  __syncthreads();      // all loads of this block have returned (they fed the adds / the shared words above)

  // ---- exit, first half: peers may overwrite their block b
  if (tid < W)
    st_release_sys(reinterpret_cast<uint32_t*>(const_cast<float*>(a.peer[tid])) + out_slot + a.rank, g);

  // ---- Adam on this block's elements (the arithmetic of acx_adam_kernel)
  if (i < n4p) {
    // the actor's learning rate of THIS step: the rule on the rank-averaged KL of the previous update (tail word 0),
    // formed by every thread from the same words in the same order.  The opt block is not written here (a block
    // reading it after another had updated it would apply the rule twice): acx_kl_lr_update_split(flags = 1) on
    // grad_sum's tail, launched right after this kernel, records the same value.
    float lr_a = a.opt[ACX_OPT_ACTOR_LR];
    if (rule) {
      float kl = 0.f;
      for (int r = 0; r < W; ++r) kl = __fadd_rn(kl, kl_sh[r]);
      lr_a = acx_lr_rule(lr_a, __fmul_rn(kl, a.scale), a.opt);
    }
    const float corr = acx_adam_corr(a.opt);
    const bool is_a = i < a.n4a;
    const int j = is_a ? i : i - a.n4a;
    const float alpha = __fmul_rn(is_a ? lr_a : a.opt[ACX_OPT_CRITIC_LR], corr);
    const float clip = a.opt[ACX_OPT_CLIP];
    float4* th = reinterpret_cast<float4*>(is_a ? a.actor : a.critic) + j;
    float4* m = reinterpret_cast<float4*>(is_a ? a.ma : a.mc) + j;
    float4* v = reinterpret_cast<float4*>(is_a ? a.va : a.vc) + j;
    float4 gr = s;
    gr.x *= a.scale; gr.y *= a.scale; gr.z *= a.scale; gr.w *= a.scale;
    if (is_a && clip > 0.f) gr = acx_clip4(gr, clip);
    float4 t = *th, mm = *m, vv = *v;
    acx_adam_one(t.x, gr.x, mm.x, vv.x, alpha); acx_adam_one(t.y, gr.y, mm.y, vv.y, alpha);
    acx_adam_one(t.z, gr.z, mm.z, vv.z, alpha); acx_adam_one(t.w, gr.w, mm.w, vv.w, alpha);
    *th = t; *m = mm; *v = vv;
  }

  // ---- exit, second half: every peer has read this rank's block b
  if (tid < W) wait_flag(mine + out_slot + tid, g, a.rank, tid, 1);
}

}  // namespace

extern "C" size_t acx_xgpu_flag_offset_bytes(int n_floats) {
  if (n_floats < 0) return 0;
  return ((size_t)n_floats * sizeof(float) + 255) / 256 * 256;
}

extern "C" size_t acx_xgpu_buffer_bytes(int n_floats) {
  if (n_floats < 0) return 0;
  return acx_xgpu_flag_offset_bytes(n_floats) + 2ull * ACX_XGPU_MAX_BLOCKS * ACX_XGPU_MAX_RANKS * sizeof(uint32_t);
}

extern "C" int acx_allreduce_adam(const void* const* peer_bufs, int world, int rank, int n_actor, int n_critic,
                                  int n_tail, float* actor, float* m_actor, float* v_actor, float* critic,
                                  float* m_critic, float* v_critic, const float* opt, float* grad_sum,
                                  unsigned int* generation, float scale, int apply_lr_rule, void* stream) {
  ACX_REQUIRE(peer_bufs, "acx_allreduce_adam: peer_bufs is null");
  ACX_REQUIRE(world >= 1 && world <= ACX_XGPU_MAX_RANKS, "acx_allreduce_adam: world %d outside [1,%d]", world,
              ACX_XGPU_MAX_RANKS);
  ACX_REQUIRE(rank >= 0 && rank < world, "acx_allreduce_adam: rank %d outside [0,%d)", rank, world);
  ACX_REQUIRE(n_actor >= 0 && n_critic >= 0 && n_tail >= 0 && (n_actor % 4) == 0 && (n_critic % 4) == 0 &&
                  (n_tail % 4) == 0,
              "acx_allreduce_adam: counts must be non-negative multiples of 4 (got %d, %d, %d)", n_actor, n_critic,
              n_tail);
  ACX_REQUIRE(opt && generation, "acx_allreduce_adam: opt / generation is null");
  for (int r = 0; r < world; ++r)
    ACX_REQUIRE(peer_bufs[r] && acx_aligned16(peer_bufs[r]), "acx_allreduce_adam: buffer of rank %d is null or not "
                "16-byte aligned", r);
  if (n_actor > 0)
    ACX_REQUIRE(actor && m_actor && v_actor && acx_aligned16(actor) && acx_aligned16(m_actor) && acx_aligned16(v_actor),
                "acx_allreduce_adam: actor buffers must be non-null and 16-byte aligned");
  if (n_critic > 0)
    ACX_REQUIRE(critic && m_critic && v_critic && acx_aligned16(critic) && acx_aligned16(m_critic) &&
                    acx_aligned16(v_critic),
                "acx_allreduce_adam: critic buffers must be non-null and 16-byte aligned");
  ACX_REQUIRE(grad_sum == nullptr || acx_aligned16(grad_sum), "acx_allreduce_adam: grad_sum must be 16-byte aligned");
  const int n4 = (n_actor + n_critic + n_tail) / 4;
  if (n4 == 0) return ACX_OK;
  const int grid = (n4 + XG_THREADS - 1) / XG_THREADS;
  // one element per thread and all blocks resident at once: a block waits for the same-numbered blocks of its peers
  ACX_REQUIRE(grid <= ACX_XGPU_MAX_BLOCKS && grid <= acx_num_sms(),
              "acx_allreduce_adam: %d floats need %d blocks, more than %d (or than this GPU's SMs)",
              n_actor + n_critic + n_tail, grid, ACX_XGPU_MAX_BLOCKS);
  XgArgs a;
  memset(&a, 0, sizeof(a));
  for (int r = 0; r < world; ++r) a.peer[r] = static_cast<const float*>(peer_bufs[r]);
  a.world = world; a.rank = rank;
  a.n4a = n_actor / 4; a.n4c = n_critic / 4; a.n4t = n_tail / 4;
  a.flag_off = acx_xgpu_flag_offset_bytes(n_actor + n_critic + n_tail) / sizeof(uint32_t);
  a.actor = actor; a.ma = m_actor; a.va = v_actor; a.critic = critic; a.mc = m_critic; a.vc = v_critic;
  a.opt = opt; a.grad_sum = grad_sum; a.gen = generation; a.scale = scale; a.apply_lr_rule = apply_lr_rule;
  acx_allreduce_adam_kernel<<<grid, XG_THREADS, 0, (cudaStream_t)stream>>>(a);
  ACX_LAUNCH_OK("acx_allreduce_adam_kernel");
  return ACX_OK;
}
