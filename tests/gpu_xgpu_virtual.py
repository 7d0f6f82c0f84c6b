"""acx_allreduce_adam on ONE GPU with W "virtual ranks": W launches on W streams, each with its own exchange buffer,
parameters, moments and opt block, all pointing at the same W buffers as their peers.  Run as a script by
tests/test_gpu_xgpu.py (a subprocess: a rank that never arrives ends in a device trap, which is sticky for the process).

Expected values come from the library's own single-GPU entry points on the fixed-order sum:
    gsum = ((b_0 + b_1) + ... + b_{W-1})  in fp32;   acx_kl_lr_update_split(flags=1, gsum tail);   acx_adam_apply(gsum)
and every virtual rank must match them bit for bit, over several generations (flag reuse)."""
import ctypes
import os
import sys

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")      # one hardware queue per stream: the launches must overlap
# a kernel's first launch loads its module lazily, which can wait for running kernels: rank 0's launch would then spin
# for a rank 1 whose launch the host cannot issue (separate processes in real use -- no such coupling there)
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from actor_critic_algs_on_tensorflow_b200 import _lib as L  # noqa: E402


def run(world, ob, num_ac, cont, generations=4, seed=0):
    dev = torch.device("cuda:0")
    lib = L.load()
    _, Pa = L.params_layout(0, ob, num_ac, cont)
    _, Pc = L.params_layout(1, ob)
    P, TAIL = Pa + Pc, 4
    nflt = lib.acx_xgpu_buffer_bytes(P + TAIL) // 4
    gen = torch.Generator(device="cpu").manual_seed(seed)
    bufs = [torch.zeros(nflt, device=dev) for _ in range(world)]
    peers = (ctypes.c_void_p * world)(*[b.data_ptr() for b in bufs])
    theta0 = torch.randn(P, generator=gen).to(dev)

    def fresh():
        opt = torch.zeros(L.OPT_SIZE, device=dev)
        L.call("acx_opt_init", L.ptr(opt), 0.005, 0.001, 1.0, 0.01, 0.1, 0.002, 1e-6, 0.1, 1, L.cur_stream())
        return dict(params=theta0.clone(), m=torch.zeros(P, device=dev), v=torch.zeros(P, device=dev), opt=opt,
                    stats=torch.zeros(L.STATS_SIZE, device=dev), gsum=torch.zeros(P + TAIL, device=dev),
                    gen=torch.zeros(128, device=dev, dtype=torch.int32))

    ranks = [fresh() for _ in range(world)]
    ref = fresh()
    streams = [torch.cuda.Stream(device=dev) for _ in range(world)]
    scale = 1.0 / world
    for g in range(generations):
        rule = int(g > 0)                       # the first update has no previous KL
        # gradients with entries on both sides of the clip, KL tails that move the lr both ways over the generations
        kl = [0.0, 1e-5, 0.05, 0.002][g % 4]
        for r in range(world):
            bufs[r][:P] = (torch.randn(P, generator=gen) * 0.2).to(dev)
            bufs[r][P:P + TAIL] = torch.tensor([kl * (1 + 0.1 * r), 0, 0, 0], device=dev)
        torch.cuda.synchronize()
        for r in range(world):
            R = ranks[r]
            with torch.cuda.stream(streams[r]):
                s = L.cur_stream()
                L.call("acx_allreduce_adam", peers, world, r, Pa, Pc, TAIL, L.ptr(R["params"][:Pa]), L.ptr(R["m"][:Pa]),
                       L.ptr(R["v"][:Pa]), L.ptr(R["params"][Pa:]), L.ptr(R["m"][Pa:]), L.ptr(R["v"][Pa:]), L.ptr(R["opt"]),
                       L.ptr(R["gsum"]), L.ptr(R["gen"]), scale, rule, s)
                if rule:
                    L.call("acx_kl_lr_update_split", L.ptr(R["opt"]), L.ptr(R["stats"]), R["gsum"].data_ptr() + 4 * P,
                           scale, 1, s)
                L.call("acx_adam_tick", L.ptr(R["opt"]), s)
        torch.cuda.synchronize()
        # expected: fixed-order fp32 sum, then the single-GPU entry points
        want = bufs[0][:P + TAIL].clone()
        for r in range(1, world):
            want = want + bufs[r][:P + TAIL]
        s = L.cur_stream()
        if rule:
            L.call("acx_kl_lr_update_split", L.ptr(ref["opt"]), L.ptr(ref["stats"]), want.data_ptr() + 4 * P, scale, 1, s)
        L.call("acx_adam_apply", L.ptr(ref["params"][:Pa]), L.ptr(want[:Pa]), L.ptr(ref["m"][:Pa]), L.ptr(ref["v"][:Pa]), Pa,
               L.ptr(ref["params"][Pa:]), want.data_ptr() + 4 * Pa, L.ptr(ref["m"][Pa:]), L.ptr(ref["v"][Pa:]), Pc,
               L.ptr(ref["opt"]), scale, s)
        L.call("acx_adam_tick", L.ptr(ref["opt"]), s)
        torch.cuda.synchronize()
        for r in range(world):
            R = ranks[r]
            assert torch.equal(R["gsum"], want), "generation %d rank %d: summed block differs" % (g, r)
            for k in ("params", "m", "v", "opt"):
                assert torch.equal(R[k], ref[k]), "generation %d rank %d: %s differs" % (g, r, k)
            assert int(R["gen"][0]) == g + 1
        # nothing but the flags changed in the exchange buffers' tails, and the gradients are untouched
    lrs = float(ref["opt"][L.OPT_ACTOR_LR])
    print("ok world=%d P=%d generations=%d final lr %.6g" % (world, P, generations, lrs), flush=True)


if __name__ == "__main__":
    world = int(sys.argv[1])
    run(world, 3, 1, True)           # Pendulum
    run(world, 4, 2, False, seed=1)  # CartPole
    run(world, 18, 3, False, generations=2, seed=2)   # Acrobot x 3 frames: the largest layouts
