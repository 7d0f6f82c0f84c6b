"""The workspace-free scan entry point, acx_returns_adv_scan_static, against the regular one.

Some hosts of this library cannot give it a workspace: a worker embedded in another framework's
process whose memory pool already owns the GPU, a loop that cannot tolerate the synchronisation an
allocation brings, a plugin thread that only ever sees its input and output buffers.  For them the
scan can run on a small scratch array built into the library.  This file checks, through the C ABI,
that it returns what acx_returns_adv_scan returns (same kernels, so the returns and advantages must
agree to the tolerance the scan tests use, 1e-5 relative) on two calls in a row with different
inputs -- the second call sees the first call's leftovers in the shared scratch -- and that a batch
wider than the built-in scratch is refused with a message.  It is named to be collected last, so a
`pytest -x` run has exercised everything else before it; it passes on its own in a fresh process too.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(T, N, pdone, seed, dev):
    rng = np.random.default_rng(seed)
    rew = torch.tensor((rng.normal(size=(T, N)) * 2).astype(np.float32), device=dev)
    V = torch.tensor((rng.normal(size=(T + 1, N)) * 5).astype(np.float32), device=dev)
    done = torch.tensor((rng.random((T, N)) < pdone).astype(np.uint8), device=dev)
    return rew, done, V


def _regular(lib, dev, T, N, rew, done, V, L, mode):
    G, A, stats = torch.empty((T, N), device=dev), torch.empty((T, N), device=dev), torch.zeros(16, device=dev)
    nbytes = lib.load().acx_returns_adv_workspace_bytes(T, N)
    ws = torch.zeros(nbytes, device=dev, dtype=torch.uint8)
    lib.call("acx_returns_adv_scan", T, N, lib.ptr(rew), lib.ptr(done), lib.ptr(V), 0.98, L, lib.ptr(G), lib.ptr(A),
             lib.ptr(stats), lib.ptr(ws), nbytes, mode, lib.cur_stream())
    torch.cuda.synchronize()
    return G.cpu().numpy(), A.cpu().numpy(), stats.cpu().numpy()


def _static(lib, dev, T, N, rew, done, V, L, mode):
    G, A, stats = torch.empty((T, N), device=dev), torch.empty((T, N), device=dev), torch.zeros(16, device=dev)
    lib.call("acx_returns_adv_scan_static", T, N, lib.ptr(rew), lib.ptr(done), lib.ptr(V), 0.98, L, lib.ptr(G),
             lib.ptr(A), lib.ptr(stats), mode, lib.cur_stream())
    torch.cuda.synchronize()
    return G.cpu().numpy(), A.cpu().numpy(), stats.cpu().numpy()


def test_static_scratch_scan_matches_workspace_scan_twice_in_a_row(lib, dev):
    cases = [(20, 65536, 0.02, 40, 0),      # config-3 shape, column kernel, plain scan
             (200, 33, 0.01, 40, 0)]        # few long columns: warp kernel, windowed
    for i, (T, N, pdone, L, mode) in enumerate(cases):
        rew, done, V = _inputs(T, N, pdone, 100 + i, dev)
        want = _regular(lib, dev, T, N, rew, done, V, L, mode)
        got = _static(lib, dev, T, N, rew, done, V, L, mode)
        scale = np.abs(want[0]).max()
        np.testing.assert_allclose(got[0], want[0], rtol=1e-5, atol=2e-6 * scale, err_msg="returns, call %d" % i)
        np.testing.assert_allclose(got[1], want[1], rtol=1e-5, atol=4e-6 * scale, err_msg="advantages, call %d" % i)
        np.testing.assert_allclose(got[2][:3], want[2][:3], rtol=1e-5, atol=1e-6, err_msg="moments, call %d" % i)


def test_static_scratch_scan_refuses_oversized_batches(lib, dev):
    t = torch.zeros(64, device=dev)
    f = lib.load().acx_returns_adv_scan_static
    too_wide = lib.SCAN_STATIC_MAX_COLUMNS + 1
    rc = f(2, too_wide, t.data_ptr(), t.data_ptr(), t.data_ptr(), 0.98, 40, t.data_ptr(), t.data_ptr(), t.data_ptr(),
           0, None)
    assert rc == -1
    msg = lib.load().acx_last_error().decode()
    assert "maximum" in msg and str(lib.SCAN_STATIC_MAX_COLUMNS) in msg
    assert f(0, 4, None, None, None, 0.98, 40, None, None, None, 0, None) == 0      # empty batch: nothing to do
