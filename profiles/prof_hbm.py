"""The HBM-bound kernels (K1 env step x 3 envs, K3 scan, K5 Adam, the diagnostics) at 65,536 / 1M / 8M envs, one launch
each after a warm-up launch, for `ncu --set full` captures (dram__bytes_read/write.sum, dram__throughput):

    ncu --set full --clock-control none -k regex:'acx_env_step|acx_scan|acx_adam_kernel|acx_episode' \
        -o gpurun_out/r02_hbm python profiles/prof_hbm.py

Every launch works on buffers of its own (written once by torch before), so nothing is L2-resident by accident at the
8M size; at 65,536 envs the working set (3-22 MB) fits the 126 MB L2 whatever one does -- what ncu reports there is the
replayed, cache-controlled launch."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from actor_critic_algs_on_tensorflow_b200 import _lib as L  # noqa: E402
from actor_critic_algs_on_tensorflow_b200.envs import VecEnv  # noqa: E402

dev = torch.device("cuda")
s = L.cur_stream()
T = 20
for n in (65536, 1 << 20, 8 << 20):
    for env in ("Pendulum-v0", "CartPole-v0", "Acrobot-v1"):
        ve = VecEnv(env, n, seed=1)
        ve.reset()
        act = torch.zeros(n, device=dev) if ve.n_actions == 0 else torch.ones(n, device=dev, dtype=torch.int32)
        for _ in range(2):
            ve.step(act)
        del ve, act
    rew, done = torch.randn((T, n), device=dev), (torch.rand((T, n), device=dev) < 0.005).to(torch.uint8)
    V, G, A = torch.randn((T + 1, n), device=dev), torch.empty((T, n), device=dev), torch.empty((T, n), device=dev)
    stats = torch.zeros(16, device=dev)
    ws = torch.zeros(L.load().acx_returns_adv_workspace_bytes(T, n), device=dev, dtype=torch.uint8)
    for _ in range(2):
        L.call("acx_returns_adv_scan", T, n, L.ptr(rew), L.ptr(done), L.ptr(V), 0.98, 40, L.ptr(G), L.ptr(A), L.ptr(stats),
               L.ptr(ws), ws.numel(), 0, s)
    ep = torch.zeros(n, device=dev)
    wd = torch.zeros(L.load().acx_diag_workspace_bytes(), device=dev, dtype=torch.uint8)
    for _ in range(2):
        L.call("acx_episode_returns", T, n, L.ptr(rew), L.ptr(done), L.ptr(ep), L.ptr(stats), L.ptr(wd), wd.numel(), s)
    del rew, done, V, G, A
    torch.cuda.synchronize()
_, Pa = L.params_layout(0, 3, 1, True)
_, Pc = L.params_layout(1, 3)
P = Pa + Pc
p, g, m, v = (torch.randn(P, device=dev) for _ in range(4))
v.abs_()
opt = torch.zeros(16, device=dev)
L.call("acx_opt_init", L.ptr(opt), 0.005, 0.001, 1.0, 0.01, 0.1, 0.002, 1e-6, 0.1, 1, s)
for _ in range(2):
    L.call("acx_adam_apply", L.ptr(p[:Pa]), L.ptr(g[:Pa]), L.ptr(m[:Pa]), L.ptr(v[:Pa]), Pa, L.ptr(p[Pa:]), L.ptr(g[Pa:]),
           L.ptr(m[Pa:]), L.ptr(v[Pa:]), Pc, L.ptr(opt), 1.0, s)
torch.cuda.synchronize()
print("ok")
