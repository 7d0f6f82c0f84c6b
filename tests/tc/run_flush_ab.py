"""Dev: accuracy (against fp64 of the same operands, one-signed data) and time of the tensor-core backward for the library
named by ACX_LIB_PATH -- used to choose ACX_WG_FLUSH."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from actor_critic_algs_on_tensorflow_b200 import _lib as lib  # noqa: E402
from gpu_util import dev_t  # noqa: E402
import test_gpu_backward_tc as T  # noqa: E402

dev = torch.device("cuda")
rows, n = 20, 65536
k = T.Case(lib, dev, True, 3, 1, rows, n, 77)
k.targets = k.targets + np.float32(12.0)
k.tg_d = dev_t(k.targets, dev)
_, gc, _ = k.run("tc", actor=False)
base = (lib.load().acx_loss_backward_workspace_bytes(3, 1, 1) + 255) & ~255
split = k.ws[base:].view(torch.float32)
bn = rows * n
d3 = split[:bn * 128].view(rows, n // 128, 128, 128).double()
ref_w3 = torch.einsum("rtks,rtjs->kj", k.cx2.double(), d3)
o = k.clay["W3"][0]
w3 = torch.tensor(gc[o:o + 128 * 128].reshape(128, 128), device=dev).double()
err = ((w3 - ref_w3).abs().max() / ref_w3.abs().max()).item()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ts = []
for net in ("critic", "actor"):
    tot = 0.0
    for _ in range(6):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        k.run("tc", actor=(net == "actor"), critic=(net == "critic"))
        b.record()
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    ts.append(tot / 6 / 2)          # run() calls the entry point twice
print("%s: one-signed dW3 err %.2e of max vs fp64 of the same operands; backward pair critic %.3f ms, actor %.3f ms" % (
    os.environ.get("ACX_LIB_PATH", "libacx.so"), err, ts[0], ts[1]), flush=True)
