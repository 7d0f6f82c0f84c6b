"""Dev: cycles per kind::tf32 128xNx8 MMA by operand source (csrc/acx_mma_rate_test.cu)."""
import ctypes as C
import os

import torch

lib = C.CDLL(os.path.join(os.path.dirname(os.path.abspath(__file__)), "libacx_mma_rate.so"))
lib.acx_mma_rate_test.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
out = torch.zeros(2, dtype=torch.int64, device="cuda")
for ctas in (1, 148):
    for n_out in (128, 64, 256):
        for mode, name in ((0, "A tmem, B no-swizzle"), (1, "A,B smem sw128"), (2, "A tmem, B sw128")):
            for iters in (8, 64, 512):
                rc = lib.acx_mma_rate_test(mode, iters, n_out, out.data_ptr(), ctas)
                o = out.cpu().tolist()
                n = iters * 6
                print("ctas=%3d N=%3d %-22s iters=%4d rc=%d issue %.1f cyc/mma, done %.1f cyc/mma (%.0f MAC/clk)" % (
                    ctas, n_out, name, iters, rc, o[0] / n, o[1] / n, 128 * n_out * 8 * n / o[1]), flush=True)
