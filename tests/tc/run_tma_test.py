"""Dev check of the TMA-tensor + 128-byte-swizzle operand path (csrc/acx_tma_test.cu) on a B200."""
import ctypes as C
import os

import numpy as np
import torch

here = os.path.dirname(os.path.abspath(__file__))
lib = C.CDLL(os.path.join(here, "libacx_tma_test.so"))
lib.acx_tma_gemm_test.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]


def run(K, N, mode, exact):
    rng = np.random.default_rng(K + N + mode)
    if exact:
        A = rng.integers(-4, 5, (128, K)).astype(np.float32)
        B = rng.integers(-4, 5, (N, K)).astype(np.float32)
    else:
        A = rng.normal(size=(128, K)).astype(np.float32)
        B = rng.normal(size=(N, K)).astype(np.float32)
    a, b = torch.tensor(A, device="cuda"), torch.tensor(B, device="cuda")
    d = torch.full((128, N), -777.0, device="cuda")
    rc = lib.acx_tma_gemm_test(a.data_ptr(), b.data_ptr(), d.data_ptr(), K, N, mode, None)
    torch.cuda.synchronize()
    got = d.cpu().numpy()
    want = A.astype(np.float64) @ B.astype(np.float64).T
    err = np.abs(got - want).max() / np.abs(want).max()
    print("K=%d N=%d mode=%d exact=%s rc=%d rel_err=%.3g" % (K, N, mode, exact, rc, err), flush=True)
    if err > 1e-2:
        bad = np.argwhere(np.abs(got - want) > 1e-2 * np.abs(want).max())
        print("  bad entries: %d of %d; first:" % (len(bad), got.size), bad[:6].tolist())
        print("  got[0,:8]", got[0, :8], "\n  want[0,:8]", want[0, :8])
    return err


if __name__ == "__main__":
    for mode in (0, 1, 2):
        for K, N in ((32, 128), (64, 128), (96, 128), (64, 64)):
            run(K, N, mode, True)
            run(K, N, mode, False)
