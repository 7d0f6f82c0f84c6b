"""Restatement of ``PathAdv`` (``A3C/process.py:46-71`` == ``Basic_AC/run_AC.py:55-80``),
``Framer`` (``A3C/process.py:14-43``), ``LinearSchedule`` and ``var_accounted_for``
(``A3C/util.py:5-42``).  TEST INFRASTRUCTURE (see ``oracle/__init__``).

PINNED: ``tests/golden/pathadv_*.npz`` etc. were produced by the reference's own
classes (imported from /root/reference under tensorflow/gym stubs by
``tests/golden/make_golden.py``); ``tests/test_oracle.py`` checks this file against them.
"""
import numpy as np


def path_adv_episode(rews, vals, terminal, gamma=0.98, look_ahead=40):
    """One episode, exactly the reference's semantics (A3C/process.py:54-66), float64:
    G_i = sum_{k<min(L, n-i)} gamma^k r_{i+k};  h = min(i+L, n);
    if not terminal or h != n:  G_i += gamma^(h-i) * vals[h];   adv_i = G_i - vals[i]."""
    rews = np.asarray(rews, dtype=np.float64)
    vals = np.asarray(vals, dtype=np.float64)
    n = len(rews)
    assert len(vals) == n + 1
    G = np.zeros(n)
    for i in range(n):
        kmax = min(look_ahead, n - i)
        G[i] = np.sum(gamma ** np.arange(kmax) * rews[i:i + kmax])
        h = min(i + look_ahead, n)
        if (not terminal) or h != n:
            G[i] += gamma ** (h - i) * vals[h]
    return G, G - vals[:n]


def path_adv_columns(rew, done, V, gamma=0.98, look_ahead=40, dtype=np.float64):
    """Batched layout used by the CUDA path: rew[T,N], done[T,N] (1 = the transition at t
    ended an episode, terminal), V[T+1,N] (value of the observation *entering* step t;
    V[T] = value of the last observation).  Each column holds several episode segments
    (auto-reset); the tail segment is cut by the horizon and is non-terminal -> bootstraps
    from V[T], exactly the reference's max_path_length cut (A3C/process.py:63).
    Returns (G[T,N], adv[T,N])."""
    f = dtype
    rew = np.asarray(rew).astype(f)
    V = np.asarray(V).astype(f)
    done = np.asarray(done).astype(bool)
    T, N = rew.shape
    # e[t,n] = exclusive end index of the segment containing t; term[t,n] = that segment is terminal
    e = np.zeros((T, N), dtype=np.int64)
    term = np.zeros((T, N), dtype=bool)
    cur_e = np.full(N, T, dtype=np.int64)
    cur_t = np.zeros(N, dtype=bool)
    for t in range(T - 1, -1, -1):
        cur_e = np.where(done[t], t + 1, cur_e)
        cur_t = np.where(done[t], True, cur_t)
        e[t], term[t] = cur_e, cur_t
    G = np.zeros((T, N), dtype=f)
    cols = np.arange(N)
    for t in range(T):
        acc = np.zeros(N, dtype=f)
        for k in range(min(look_ahead, T - t)):
            live = (t + k) < e[t]
            acc = acc + np.where(live, f(gamma) ** f(k) * rew[t + k], f(0))
        h = np.minimum(t + look_ahead, e[t])
        boot = (~term[t]) | (h != e[t])
        acc = acc + np.where(boot, f(gamma) ** (h - t).astype(f) * V[h, cols], f(0))
        G[t] = acc
    return G, (G - V[:T]).astype(f)


def framer_full(obs, frame_num):
    """Framer.full (A3C/process.py:36-43): stack the last frame_num observations, first one
    repeated as left padding."""
    obs = list(obs)
    ext = [obs[0]] * (frame_num - 1) + obs
    return [np.concatenate([ext[i + j] for j in range(frame_num)]) for i in range(len(ext) - frame_num + 1)]


def framer_last(obs, frame_num):
    """Framer.last (A3C/process.py:31-34)."""
    obs = list(obs)
    ext = [obs[0]] * (frame_num - 1) + obs
    return np.concatenate([ext[i] for i in range(-frame_num, 0)])


def linear_schedule_val(t, init_t, end_t, init_val, end_val):
    """LinearSchedule.val (A3C/util.py:34-39)."""
    if t < init_t:
        return init_val
    if t > end_t:
        return end_val
    return ((t - init_t) * end_val + (end_t - t) * init_val) / float(end_t - init_t)


def var_accounted_for(target, pred):
    """A3C/util.py:5-13."""
    pred, target = np.asarray(pred, dtype=np.float64).reshape(-1), np.asarray(target, dtype=np.float64).reshape(-1)
    pred = (pred - np.mean(pred)) / np.std(pred)
    target = (target - np.mean(target)) / np.std(target)
    return np.mean(target * pred)
