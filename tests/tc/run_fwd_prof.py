"""Dev: per-phase cycle counts of the forward tensor-core kernel (CTA 0 / thread 0) from `make prof` (libacx_prof.so)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402
from actor_critic_algs_on_tensorflow_b200 import _lib as L  # noqa: E402

L.LIB_PATH = os.environ.get("ACX_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libacx_prof.so")
from actor_critic_algs_on_tensorflow_b200.trainer import A2CTrainer  # noqa: E402

tr = A2CTrainer("Pendulum-v0", n_envs=65536, horizon=20, seed=12321, use_graph=False)
for _ in range(2):
    tr.iteration()
torch.cuda.synchronize()
lib = L.load()
lib.acx_fwd_prof_read.argtypes = [C.c_void_p]
names = ["layer 1 block 0 -> A (first tile only: later tiles at [10])", "layer 1 block 1 compute", "wait MMA block 0",
         "layer 1 block 1 -> A", "wait MMA layer 2", "X2 epilogue (+cache) -> A", "wait MMA layer 3",
         "X3 epilogue (+cache), head partial", "sync", "head", "layer 1 block 0 of the NEXT tile -> A (ahead of the epilogue)"]
buf = (C.c_ulonglong * 16)()
s = L.cur_stream()
T, N = tr.T, tr.N
for which in ("value", "logp", "rollout"):
    if which == "value":
        L.call("acx_value_forward_tc", L.ptr(tr.critic), L.ptr(tr.packed_c), tr.ob, T + 1, N, L.ptr(tr.obs_buf), L.ptr(tr.V),
               0, L.ptr(tr.cx2), L.ptr(tr.cx3), L.ptr(tr.cmask), s)
        tiles = (T + 1) * 512 / 148
    elif which == "rollout":
        st3, el3 = tr.state.clone(), tr.elapsed.clone()
        L.call("acx_rollout_tc", tr.env, L.ptr(tr.actor), L.ptr(tr.packed_a), tr.ac_scale, T, N, L.ptr(tr.obs_buf), L.ptr(st3),
               L.ptr(el3), None, None, tr.seed, 0, None, tr.max_steps, L.ptr(tr.acs), L.ptr(tr.logps), L.ptr(tr.ent),
               L.ptr(tr.rew), L.ptr(tr.done), L.ptr(tr.ax2), L.ptr(tr.ax3), L.ptr(tr.amask), s)
        tiles = T * 4          # CTA 0: four column tiles x T steps
    else:
        L.call("acx_policy_logp_gauss_tc", L.ptr(tr.actor), L.ptr(tr.packed_a), tr.ob, tr.num_ac, tr.ac_scale, T, N,
               L.ptr(tr.obs_buf), L.ptr(tr.acs), None, None, None, L.ptr(tr.logps), L.ptr(tr.kl_part),
               C.byref(tr.kl_blocks), s)
        tiles = T * 512 / 148
    torch.cuda.synchronize()
    lib.acx_fwd_prof_read(buf)
    tot = sum(buf[i] for i in range(11))
    print(which, "total cycles", tot, "tiles ~ %.1f" % tiles)
    for i, nm in enumerate(names):
        print("  %-40s %9d  %5.1f%%  per tile %7.0f" % (nm, buf[i], 100.0 * buf[i] / max(tot, 1), buf[i] / tiles))
