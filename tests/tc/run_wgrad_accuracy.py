"""Dev: accuracy of acx_wgrad_kernel at the bench size against an fp64 product of the very arrays it reads."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from actor_critic_algs_on_tensorflow_b200 import _lib as lib  # noqa: E402
from oracle import policies as P  # noqa: E402
import test_gpu_backward_tc as T  # noqa: E402

dev = torch.device("cuda")
rows, n = 20, 65536
k = T.Case(lib, dev, True, 3, 1, rows, n, 77)
ga, gc, st = k.run("tc", actor=False)
fa, fc, fs = k.run("ffma", actor=False)
base = (lib.load().acx_loss_backward_workspace_bytes(3, 1, 1) + 255) & ~255
split = k.ws[base:].view(torch.float32)
bn = rows * n                                              # (n is a multiple of 128: no padding)
d3 = split[:bn * 128].view(rows, n // 128, 128, 128).double()           # tile-blocked [rows][tiles][feature][128]
x2 = k.cx2.double()
ref_w3 = torch.einsum("rtks,rtjs->kj", x2, d3)             # dW3[k][j] = sum X2[k][s] dPre3[j][s]  (dPre already carries 1/B)
lay = k.clay
o, s = lay["W3"]
for name, g in (("tc", gc), ("ffma", fc)):
    w3 = torch.tensor(g[o:o + 128 * 128].reshape(128, 128), device=dev).double()
    print(name, "dW3 max rel err vs fp64 of the same operands: %.3g" % ((w3 - ref_w3).abs().max() / ref_w3.abs().max()).item())
o, s = lay["b3"]
ref_b3 = d3.sum(dim=(0, 1, 3))
for name, g in (("tc", gc), ("ffma", fc)):
    b3 = torch.tensor(g[o:o + 128], device=dev).double()
    print(name, "db3 max rel err: %.3g" % ((b3 - ref_b3).abs().max() / ref_b3.abs().max()).item())
