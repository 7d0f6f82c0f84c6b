"""Philox4x32-10 counter-based RNG in numpy (oracle side; test infrastructure).

The reference seeds only numpy/gym (``A3C/process.py:179-180``) and never seeds the
TF graph, so its trajectories are irreproducible; parity is "same uniforms /
normals in -> same actions out".  The CUDA library can draw its own uniforms and
normals from Philox4x32-10 keyed on (seed, stream, env index, step); this module is
the bit-identical host restatement (Salmon et al., SC'11, the published algorithm)
so that even the device-RNG mode is checkable.

Counter layout (must match ``csrc/acx_common.cuh::acx_philox``):
    ctr = (index_lo, index_hi, step, stream)      key = (seed_lo, seed_hi)
Uniform:  u = (x >> 8) * 2^-24           in [0, 1)      (exact in fp32)
Normal:   Box-Muller on (x0, x1):  u1 = ((x0 >> 8) + 1) * 2^-24 in (0, 1],
          z = sqrt(-2 ln u1) * cos(2 pi u2)   (fp32 arithmetic)
"""
import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = 0x9E3779B9
W1 = 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)

# stream ids shared with the device code
STREAM_RESET = 1     # env reset uniforms
STREAM_ACTION = 2    # action noise (normal eps / categorical uniform)
STREAM_PHASE = 3     # time-limit phase stagger of the batched trainer (word 0 mod max_episode_steps)


def philox4x32(index, step, stream, seed):
    """index: uint64 array (any shape). Returns uint32 array [..., 4]."""
    index = np.asarray(index, dtype=np.uint64)
    c0 = index & MASK
    c1 = (index >> np.uint64(32)) & MASK
    c2 = np.full_like(c0, np.uint64(int(step) & 0xFFFFFFFF))
    c3 = np.full_like(c0, np.uint64(int(stream) & 0xFFFFFFFF))
    k0 = int(seed) & 0xFFFFFFFF
    k1 = (int(seed) >> 32) & 0xFFFFFFFF
    for _ in range(10):
        p0 = M0 * c0
        p1 = M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK
        n0 = hi1 ^ c1 ^ np.uint64(k0)
        n1 = lo1
        n2 = hi0 ^ c3 ^ np.uint64(k1)
        n3 = lo0
        c0, c1, c2, c3 = n0, n1, n2, n3
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return np.stack([c0, c1, c2, c3], axis=-1).astype(np.uint32)


def u01(x):
    """uint32 -> float32 uniform in [0,1)."""
    return ((x >> np.uint32(8)).astype(np.float32)) * np.float32(2.0 ** -24)


def uniforms(index, step, stream, seed):
    """float32 [..., 4] uniforms in [0,1)."""
    return u01(philox4x32(index, step, stream, seed))


def normals(index, step, stream, seed):
    """float32 [..., 2] standard normals (Box-Muller, cosine branch on pairs (0,1), (2,3))."""
    x = philox4x32(index, step, stream, seed)
    f32 = np.float32
    u1 = ((x[..., 0::2] >> np.uint32(8)).astype(np.float32) + f32(1.0)) * f32(2.0 ** -24)
    u2 = u01(x[..., 1::2])
    r = np.sqrt(f32(-2.0) * np.log(u1)).astype(np.float32)
    th = (f32(6.283185307179586) * u2).astype(np.float32)
    return (r * np.cos(th)).astype(np.float32)
