"""Dev check of the tcgen05 plumbing (csrc/acx_umma.cuh) on a B200: python tests/tc/run_tc_test.py [modes]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

here = os.path.dirname(os.path.abspath(__file__))
lib = C.CDLL(os.path.join(here, "libacx_tc_test.so"))
lib.acx_tc_gemm_test.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]


def pack_mn(X):      # X[mn][k] -> [k/8][mn/4][8][4]
    MN, K = X.shape
    return np.ascontiguousarray(X.reshape(MN // 4, 4, K // 8, 8).transpose(2, 0, 3, 1)).reshape(-1)


def pack_k(X):       # X[mn][k] -> [mn/8][k/4][8][4]
    MN, K = X.shape
    return np.ascontiguousarray(X.reshape(MN // 8, 8, K // 4, 4).transpose(0, 2, 1, 3)).reshape(-1)


def run(K, N, mode, exact):
    rng = np.random.default_rng(K + N + mode)
    if exact:      # small integers: exactly representable in TF32, products/sums exact in fp32
        A = rng.integers(-4, 5, (128, K)).astype(np.float32)
        B = rng.integers(-4, 5, (N, K)).astype(np.float32)
    else:
        A = rng.normal(size=(128, K)).astype(np.float32)
        B = rng.normal(size=(N, K)).astype(np.float32)
    pk = pack_k if mode in (1, 4, 5) else pack_mn
    a = torch.tensor(A.reshape(-1) if mode >= 4 else pk(A), device="cuda")
    b = torch.tensor(pk(B), device="cuda")
    d = torch.full((128, N), -777.0, device="cuda")
    rc = lib.acx_tc_gemm_test(a.data_ptr(), b.data_ptr(), d.data_ptr(), K, N, mode, None)
    torch.cuda.synchronize()
    got = d.cpu().numpy()
    want = A.astype(np.float64) @ B.astype(np.float64).T
    err = np.abs(got - want).max() / np.abs(want).max()
    print("K=%d N=%d mode=%d exact=%s rc=%d rel_err=%.3g" % (K, N, mode, exact, rc, err), flush=True)
    if mode in (2, 5) and not exact:
        assert err < 2e-6, err
    if err > 1e-2:
        bad = np.argwhere(np.abs(got - want) > 1e-2 * np.abs(want).max())
        print("  bad entries: %d of %d; first:" % (len(bad), got.size), bad[:6].tolist())
        print("  got[0,:8]", got[0, :8], "\n  want[0,:8]", want[0, :8])
    return err


if __name__ == "__main__":
    modes = [int(x) for x in sys.argv[1:]] or [0, 1, 2]
    for mode in modes:
        for K, N in ((8, 128), (64, 128), (64, 64), (128, 128)):
            run(K, N, mode, True)
            run(K, N, mode, False)
