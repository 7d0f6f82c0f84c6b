"""K1 parity: env step / reset through the C ABI vs the oracle (oracle/envs.py), same inputs.

Tolerance: 1e-5 relative (+1e-6 absolute for values near zero) in fp32, per north_star; done
flags and elapsed counters bit-exact.  Single transitions are teacher-forced from identical
states, so only sin/cos rounding can differ for Pendulum/CartPole.
"""
import numpy as np
import pytest
import torch

from oracle import envs as E
from oracle import philox as R

pytestmark = pytest.mark.gpu
RTOL, ATOL = 1e-5, 2e-6


def _rand_state(env, n, rng):
    if env == E.PENDULUM:
        return np.stack([rng.uniform(-10, 10, n), rng.uniform(-8, 8, n)]).astype(np.float32)
    if env == E.CARTPOLE:
        return np.stack([rng.uniform(-2.5, 2.5, n), rng.uniform(-2, 2, n), rng.uniform(-0.22, 0.22, n),
                         rng.uniform(-2, 2, n)]).astype(np.float32)
    # velocities well inside the MAX_VEL clips: at |v| ~ 9 pi the RK4 step over dt = 0.2 is ill-conditioned and
    # amplifies fp32 rounding (FMA contraction vs numpy op-by-op) to ~1e-3; gym trajectories stay below ~6 rad/s
    return np.stack([rng.uniform(-np.pi, np.pi, n), rng.uniform(-np.pi, np.pi, n), rng.uniform(-3, 3, n),
                     rng.uniform(-6, 6, n)]).astype(np.float32)


def _rand_action(env, n, rng):
    if env == E.PENDULUM:
        return rng.uniform(-3, 3, n).astype(np.float32)
    return rng.integers(0, E.N_ACTIONS[env], n).astype(np.int32)


def _step_gpu(lib, dev, env, state, elapsed, action, reset_u, seed=0, index0=0, step=0, max_steps=0):
    n = state.shape[1]
    st = torch.tensor(state, device=dev).contiguous()
    el = torch.tensor(elapsed, device=dev)
    ac = torch.tensor(action, device=dev)
    ru = None if reset_u is None else torch.tensor(reset_u, device=dev).contiguous()
    obs = torch.empty((E.OBS_DIM[env], n), device=dev)
    rew = torch.empty(n, device=dev)
    done = torch.empty(n, device=dev, dtype=torch.uint8)
    lib.call("acx_env_step", env, n, lib.ptr(st), lib.ptr(el), lib.ptr(ac), lib.ptr(obs), lib.ptr(rew), lib.ptr(done),
             lib.ptr(ru), seed, index0, step, None, max_steps, lib.cur_stream())
    torch.cuda.synchronize()
    return st.cpu().numpy(), el.cpu().numpy(), obs.cpu().numpy(), rew.cpu().numpy(), done.cpu().numpy()


@pytest.mark.parametrize("env", [E.PENDULUM, E.CARTPOLE, E.ACROBOT])
@pytest.mark.parametrize("n", [1, 7, 64, 4096, 100003])
def test_env_step_matches_oracle(lib, dev, env, n):
    rng = np.random.default_rng(100 + env * 10 + n % 7)
    state = _rand_state(env, n, rng)
    elapsed = rng.integers(0, E.MAX_EPISODE_STEPS[env], n).astype(np.int32)
    elapsed[:: 5] = E.MAX_EPISODE_STEPS[env] - 1          # force TimeLimit hits
    action = _rand_action(env, n, rng)
    reset_u = rng.random((E.STATE_DIM[env], n)).astype(np.float32)
    got = _step_gpu(lib, dev, env, state, elapsed, action, reset_u)
    a = action[None, :] if env == E.PENDULUM else action
    want = E.step(env, state, elapsed, a, reset_u, dtype=np.float32)
    want64 = E.step(env, state.astype(np.float64), elapsed, a, reset_u.astype(np.float64), dtype=np.float64)
    np.testing.assert_array_equal(got[4], want[4])                      # done flags bit-exact
    np.testing.assert_array_equal(got[1], want[1])                      # elapsed bit-exact
    assert got[4].sum() > 0 or n < 5
    for g, w, w64 in zip((got[0], got[2], got[3]), (want[0], want[2], want[3]), (want64[0], want64[2], want64[3])):
        np.testing.assert_allclose(g, w, rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(g, w64, rtol=1e-4, atol=1e-4)        # vs gym's own float64 arithmetic


@pytest.mark.parametrize("env", [E.PENDULUM, E.CARTPOLE, E.ACROBOT])
def test_env_reset_and_philox_reset(lib, dev, env):
    n, seed, index0, step = 1000, 12321, 5000, 77
    S = E.STATE_DIM[env]
    st = torch.empty((S, n), device=dev)
    el = torch.full((n,), 5, device=dev, dtype=torch.int32)
    obs = torch.empty((E.OBS_DIM[env], n), device=dev)
    lib.call("acx_env_reset", env, n, lib.ptr(st), lib.ptr(el), lib.ptr(obs), None, seed, index0, step, lib.cur_stream())
    torch.cuda.synchronize()
    u = R.uniforms(np.arange(n, dtype=np.uint64) + np.uint64(index0), step, R.STREAM_RESET, seed).T[:S]
    want = E.reset_from_uniforms(env, u, np.float32)
    np.testing.assert_array_equal(st.cpu().numpy(), want)               # uniforms exact, affine map non-contracted
    np.testing.assert_allclose(obs.cpu().numpy(), E.observe(env, want, np.float32), rtol=RTOL, atol=ATOL)
    assert int(el.sum()) == 0
    # a step that terminates every env (TimeLimit) must reset from the same Philox stream
    el.fill_(E.MAX_EPISODE_STEPS[env] - 1)
    state0 = st.cpu().numpy().copy()
    action = _rand_action(env, n, np.random.default_rng(3))
    got = _step_gpu(lib, dev, env, state0, el.cpu().numpy(), action, None, seed=seed, index0=index0, step=step + 1)
    u2 = R.uniforms(np.arange(n, dtype=np.uint64) + np.uint64(index0), step + 1, R.STREAM_RESET, seed).T[:S]
    np.testing.assert_array_equal(got[0], E.reset_from_uniforms(env, u2, np.float32))
    assert got[4].all() and not got[1].any()


def test_pendulum_trajectory_drift_is_small(lib, dev):
    """200 free-running steps (no teacher forcing): fp32 kernel vs float64 gym arithmetic."""
    n = 512
    rng = np.random.default_rng(5)
    u = rng.random((2, n)).astype(np.float32)
    s32 = E.reset_from_uniforms(E.PENDULUM, u, np.float32)
    s64 = s32.astype(np.float64)
    el = np.zeros(n, np.int32)
    el64 = el.copy()
    ret32 = np.zeros(n)
    ret64 = np.zeros(n)
    for t in range(199):
        a = rng.uniform(-2, 2, n).astype(np.float32)
        ru = rng.random((2, n)).astype(np.float32)
        s32, el, _, r, d = _step_gpu(lib, dev, E.PENDULUM, s32, el, a, ru)
        s64, el64, _, r64, d64 = E.step(E.PENDULUM, s64, el64, a[None], ru.astype(np.float64), dtype=np.float64)
        ret32 += r
        ret64 += r64
        assert not d.any() and not d64.any()
    # random torques make the pendulum chaotic near the top: a few envs diverge visibly over 199 steps
    rel = np.abs(ret32 - ret64) / np.abs(ret64)
    assert np.quantile(rel, 0.9) < 1e-4 and rel.max() < 0.1, (np.quantile(rel, 0.9), rel.max())


def test_env_edge_cases(lib, dev):
    # n == 0 is a no-op; a null buffer and an unknown env id are rejected with a message
    assert lib.load().acx_env_step(0, 0, None, None, None, None, None, None, None, 0, 0, 0, None, 0, None) == 0
    assert lib.load().acx_env_step(0, 4, None, None, None, None, None, None, None, 0, 0, 0, None, 0, None) == -1
    assert b"null" in lib.load().acx_last_error()
    t = torch.zeros(16, device=dev)
    rc = lib.load().acx_env_step(9, 4, t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(),
                                 t.data_ptr(), None, 0, 0, 0, None, 0, None)
    assert rc == -1 and b"unknown env" in lib.load().acx_last_error()


@pytest.mark.parametrize("ob,frames,n", [(3, 2, 1000), (4, 2, 77), (2, 4, 513), (6, 1, 64)])
def test_frame_push_is_the_reference_framer(lib, dev, ob, frames, n):
    """acx_frame_push over a sequence with random episode ends == Framer.full (A3C/process.py:36-43) per episode."""
    from oracle.pathadv import framer_full
    from gpu_util import dev_t
    rng = np.random.default_rng(ob * 100 + frames)
    T = 9
    obs = rng.normal(size=(T + 1, ob, n)).astype(np.float32)
    done = (rng.uniform(size=(T, n)) < 0.3).astype(np.uint8)
    o_d, d_d = dev_t(obs, dev), dev_t(done, dev)
    out = torch.full((T + 1, ob * frames, n), -7.0, device=dev)
    s = lib.cur_stream()
    lib.call("acx_frame_push", n, ob, frames, None, lib.ptr(o_d[0]), None, lib.ptr(out[0]), s)
    for t in range(T):
        lib.call("acx_frame_push", n, ob, frames, lib.ptr(out[t]), lib.ptr(o_d[t + 1]), lib.ptr(d_d[t]), lib.ptr(out[t + 1]), s)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    for e in range(0, n, max(1, n // 25)):                  # per env: split into episodes, stack with the reference rule
        want, start = [], 0
        for t in range(T + 1):
            if t > 0 and done[t - 1, e]:
                start = t
            want.append(framer_full([obs[u, :, e] for u in range(start, t + 1)], frames)[-1])
        np.testing.assert_array_equal(got[:, :, e], np.stack(want))
    f = lib.load().acx_frame_push
    assert f(8, ob, frames, out[0].data_ptr(), o_d[0].data_ptr(), None, out[0].data_ptr(), None) == -1      # aliasing
    assert f(0, ob, frames, None, None, None, None, None) == 0


# ------------------------------------------------------------------ the kernels against the physics (tests/test_env_physics.py)
def _gpu_dynamics(lib, dev, env):
    """(state, action) -> (next state, reward, terminal) through acx_env_step with the time limit out of reach."""
    def dyn(s, a):
        s32 = np.ascontiguousarray(np.asarray(s, dtype=np.float32))
        n = s32.shape[1]
        a = np.asarray(a)
        a = a.astype(np.float32) if env == E.PENDULUM else a.astype(np.int32)
        ru = np.full((E.STATE_DIM[env], n), 0.5, np.float32)
        st, _, _, rew, done = _step_gpu(lib, dev, env, s32, np.zeros(n, np.int32), a, ru, max_steps=1 << 30)
        return st.astype(np.float64), rew.astype(np.float64), done.astype(bool)
    return dyn


def test_env_kernels_obey_the_equations_of_motion(lib, dev):
    """The CUDA transitions satisfy the Lagrangian equations of the three systems with the documented constants,
    integrator and time step -- the same property checks as the oracle, at fp32 resolution (an acceleration recovered
    from a velocity difference over dt carries eps * |v| / dt ~ 1e-4)."""
    import test_env_physics as PH
    PH.check_pendulum_equation_of_motion(_gpu_dynamics(lib, dev, E.PENDULUM), atol=3e-4)
    PH.check_pendulum_energy_and_period(_gpu_dynamics(lib, dev, E.PENDULUM), tol_scale=1.02)
    PH.check_cartpole_equations_of_motion(_gpu_dynamics(lib, dev, E.CARTPOLE), atol=5e-4, theta_range=0.15)
    dsdt = lambda s, a: E._acrobot_dsdt(s, a, np.float64)
    PH.check_acrobot_step(_gpu_dynamics(lib, dev, E.ACROBOT), dsdt, tol=2e-3, terminal_states=False)


def test_env_kernel_time_limits_fire_on_the_documented_step(lib, dev):
    import test_env_physics as PH

    def step_fn(env, state, elapsed, ac, ru):
        a = np.asarray(ac).reshape(-1)
        a = a.astype(np.float32) if env == E.PENDULUM else a.astype(np.int32)
        st, el, obs, rew, done = _step_gpu(lib, dev, env, np.ascontiguousarray(state, dtype=np.float32),
                                           np.asarray(elapsed, np.int32), a, np.ascontiguousarray(ru, dtype=np.float32))
        return st.astype(np.float64), el, obs, rew, done
    PH.check_time_limits(step_fn)


def test_categorical_kernel_draws_from_softmax(lib, dev):
    """acx_policy_forward_sample_cat (the tf.multinomial restatement, A3C/policies.py:64): 10^7 draws per logit vector
    follow softmax(logits) (chi^2 at p ~ 1e-7), with caller uniforms and with the device's Philox stream; u = 0 and the
    largest u < 1 give the first and the last category; the action index is monotone in u."""
    import test_env_physics as PH
    from oracle import policies as P

    def sampler(seed_box=[1]):
        def sample(lg, u):
            n, k = lg.shape
            lay = P.actor_layout(4, k, False)
            flat = np.zeros(lay["_size"], np.float32)
            o, _ = lay["bh"]
            flat[o:o + k] = lg[0]                       # zero weights: logits = head bias for every observation
            a = torch.tensor(flat, device=dev)
            obs = torch.zeros((1, 4, n), device=dev)
            ac = torch.empty((1, n), device=dev, dtype=torch.int32)
            ud = None if u is None else torch.tensor(np.ascontiguousarray(u, dtype=np.float32).reshape(1, n), device=dev)
            seed_box[0] += 1
            lib.call("acx_policy_forward_sample_cat", lib.ptr(a), 4, k, 1, n, lib.ptr(obs), lib.ptr(ud), 1000 + seed_box[0], 0, 0,
                     None, lib.ptr(ac), None, None, None, lib.cur_stream())
            torch.cuda.synchronize()
            return ac.cpu().numpy().reshape(-1)
        return sample
    PH.check_sampler_distribution(sampler(), n=10_000_000, given_u=True)
    PH.check_sampler_distribution(sampler(), n=10_000_000, given_u=False)
