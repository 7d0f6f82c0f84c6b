"""Dev: per-phase cycle counts of the chain kernel (CTA 0, thread 0) from a -DACX_CHAIN_PROF build (libacx_prof.so)."""
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
names = ["(unused)", "head of the NEXT tile (inside the dX2 wait)", "head-layer gradients of the NEXT tile (inside the dX1 wait)",
         "dPre3 + sync + bulk store", "wait MMA3 (what the head did not cover)", "obs loads",
         "wait MMA2 b0 (what the head gradients did not cover)", "dPre1 b0", "l1grad b0", "wait MMA2 b1", "dPre1 b1", "l1grad b1",
         "tile start: X3 prefetch issue, mask loads", "wait for the next tile's X3 block", "dPre2 loop",
         "dPre2: st-wait, arrive, bulk store"]
buf = (C.c_ulonglong * 16)()
s = L.cur_stream()
for which in ("actor", "critic"):
    ga, gc = tr.grads[:tr.Pa], tr.grads[tr.Pa:]
    common = (L.ptr(tr.obs_buf), L.ptr(tr.acs), L.ptr(tr.adv), L.ptr(tr.stats), L.ptr(tr.logps), L.ptr(tr.G), L.ptr(tr.opt),
              L.ptr(ga), L.ptr(gc), L.ptr(tr.stats), L.ptr(tr.ax2), L.ptr(tr.ax3), L.ptr(tr.cx2), L.ptr(tr.cx3),
              L.ptr(tr.amask), L.ptr(tr.cmask), L.ptr(tr.ws_bwd), tr.ws_bwd_bytes, s)
    L.call("acx_loss_backward_gauss_tc", L.ptr(tr.actor) if which == "actor" else None,
           L.ptr(tr.critic) if which == "critic" else None, L.ptr(tr.packed_bwd_a), L.ptr(tr.packed_bwd_c), tr.ob,
           tr.num_ac, tr.ac_scale, tr.T, tr.N, *common)
    torch.cuda.synchronize()
    lib.acx_chain_prof_read.argtypes = [C.c_void_p]
    lib.acx_chain_prof_read(buf)
    if hasattr(lib, "acx_wg_prof_read"):
        wg = (C.c_ulonglong * 24)()
        lib.acx_wg_prof_read.argtypes = [C.c_void_p]
        lib.acx_wg_prof_read(wg)
        for ph in (0, 1):
            print("  wgrad %s phase %d: %d boxes, %.0f cycles per box, of which %.0f waiting for the split box"
                  % (which, ph, wg[4 + ph], wg[ph] / max(wg[4 + ph], 1), wg[2 + ph] / max(wg[4 + ph], 1)))
            nb = max(wg[4 + ph], 1)
            if wg[6 + ph]:
                print("        issuer: commit -> MMAs complete %.0f cycles per box (ACX_WG_PROF_DRAIN build)" % (wg[6 + ph] / nb))
            print("        splitter warp 6 per box: loop + obs prefetch %.0f, wait slot %.0f, X1 generation %.0f, wait TMA box %.0f, "
                  "split + arrive %.0f" % ((wg[16 + ph] / nb,) + tuple(wg[8 + 4 * ph + i] / nb for i in range(4))))
    tot = sum(buf[i] for i in range(16))
    print(which, "total cycles", tot, "tiles ~", 20 * 512 / 148)
    for i, nm in enumerate(names):
        print("  %-44s %9d  %5.1f%%  per tile %7.0f" % (nm, buf[i], 100.0 * buf[i] / tot, buf[i] / (20 * 512 / 148)))
