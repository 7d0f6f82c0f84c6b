"""Independent pins for the parts of the oracle that restate third-party code from its published description: the gym
classic-control dynamics behind ``env.step`` (A3C/process.py:79,91,104) and ``tf.multinomial`` (A3C/policies.py:64).
gym and TensorFlow are not installable here, so neither restatement can be compared with its original.  What CAN be
checked without them is everything a transcription error breaks: the restated step has to satisfy the equations of
motion of the mechanical system it claims to integrate -- derived here from the Lagrangian with the documented masses,
lengths and gravity, not from gym's formulas -- with the documented integrator and time step; the time limit has to
fire on exactly the documented step; the sampler has to draw from softmax(logits).

  Pendulum-v0   uniform rod (m = l = 1, g = 10) about a pivot, theta from upright:  (m l^2 / 3) th'' = (m g l / 2) sin th + u
                semi-implicit Euler, dt = 0.05: energy error bounded, no drift; small-oscillation period 2 pi / sqrt(15)
  CartPole-v0   cart (M = 1) + uniform rod pole (m = 0.1, half length l = 0.5), theta from upright, force F = +-10:
                (M + m) x'' + m l (th'' cos th - th'^2 sin th) = F ;   (4/3) l th'' + x'' cos th = g sin th,   g = 9.8
  Acrobot-v1    two links (m = l = 1, lc = 0.5, I = 1, g = 9.8), torque a on the second joint:
                d/dt [ 1/2 q'^T M(q) q' + V(q) ] = a * th2'   along the restated dsdt;  RK4, dt = 0.2
These run against the numpy oracle here (CPU) and against the CUDA kernels in tests/test_gpu_env.py.

What stays unpinnable without gym: the order of CartPole's explicit Euler updates, the termination thresholds
(2.4 m, 12 degrees, Acrobot's height 1), the reward conventions, the reset ranges and the velocity clips -- they are
conventions of gym's code, not consequences of the physics."""
import numpy as np
import pytest

from oracle import envs as E
from oracle import policies as P

G_PEND, G_OTHER = 10.0, 9.8


# ------------------------------------------------------------------ the physics, written from the Lagrangians
def pendulum_accel(th, u):
    m = l = 1.0
    return ((m * G_PEND * l / 2.0) * np.sin(th) + u) / (m * l * l / 3.0)


def pendulum_energy(th, thdot):
    m = l = 1.0
    return 0.5 * (m * l * l / 3.0) * thdot ** 2 + (m * G_PEND * l / 2.0) * np.cos(th)


def cartpole_residuals(s, force, xacc, thacc):
    M, m, l, g = 1.0, 0.1, 0.5, G_OTHER
    _, _, th, thd = s
    r1 = (M + m) * xacc + m * l * (thacc * np.cos(th) - thd ** 2 * np.sin(th)) - force
    r2 = (4.0 / 3.0) * l * thacc + xacc * np.cos(th) - g * np.sin(th)
    return r1, r2


def acrobot_energy(s):
    m1 = m2 = l1 = 1.0
    lc1 = lc2 = 0.5
    I1 = I2 = 1.0
    g = G_OTHER
    t1, t2, w1, w2 = s
    # kinetic energy of two rigid links: centre-of-mass velocities + rotation
    v1sq = (lc1 * w1) ** 2
    v2sq = (l1 * w1) ** 2 + (lc2 * (w1 + w2)) ** 2 + 2.0 * l1 * lc2 * w1 * (w1 + w2) * np.cos(t2)
    T = 0.5 * m1 * v1sq + 0.5 * I1 * w1 ** 2 + 0.5 * m2 * v2sq + 0.5 * I2 * (w1 + w2) ** 2
    # heights below the pivot, theta1 from hanging down
    V = -m1 * g * lc1 * np.cos(t1) - m2 * g * (l1 * np.cos(t1) + lc2 * np.cos(t1 + t2))
    return T + V


# ------------------------------------------------------------------ stepping helpers (oracle or any other stepper)
def oracle_dynamics(env):
    return lambda s, a: E.dynamics(env, s, a, dtype=np.float64)


def check_pendulum_equation_of_motion(dyn, atol):
    rng = np.random.default_rng(0)
    th = rng.uniform(-3 * np.pi, 3 * np.pi, 4096)
    thd = rng.uniform(-6.0, 6.0, 4096)                   # |th' + th'' dt| stays below the max_speed clip of 8
    u = rng.uniform(-2.0, 2.0, 4096)
    ns, rew, term = dyn(np.stack([th, thd]), u)
    dt = 0.05
    acc = (ns[1] - thd) / dt
    np.testing.assert_allclose(acc, pendulum_accel(th, u), atol=atol, rtol=0)
    np.testing.assert_allclose(ns[0] - th, ns[1] * dt, atol=atol * dt, rtol=0)      # theta advances with the NEW speed
    an = np.mod(th + np.pi, 2 * np.pi) - np.pi
    np.testing.assert_allclose(rew, -(an ** 2 + 0.1 * thd ** 2 + 0.001 * u ** 2), atol=max(atol, 1e-9) * 40, rtol=0)
    assert not np.asarray(term).any()
    # torque and speed limits
    ns2, _, _ = dyn(np.stack([th, thd]), 10.0 * np.sign(u))
    np.testing.assert_allclose((ns2[1] - thd) / dt, pendulum_accel(th, 2.0 * np.sign(u)), atol=atol, rtol=0)
    ns3, _, _ = dyn(np.stack([np.full(8, np.pi / 2), np.full(8, 7.9)]), np.full(8, 2.0))
    np.testing.assert_allclose(ns3[1], 8.0, atol=1e-6)


def check_pendulum_energy_and_period(dyn, tol_scale=1.0):
    # zero torque, 1000 steps: the semi-implicit update keeps the energy error bounded (an explicit Euler step would
    # pump energy in: at dt = 0.05 it grows by a few per cent per swing)
    s = np.stack([np.array([np.pi - 1.0, np.pi - 0.5, np.pi - 2.0]), np.zeros(3)])
    e0 = pendulum_energy(s[0], s[1])
    dev = []
    for _ in range(1000):
        s, _, _ = dyn(s, np.zeros(3))
        dev.append(np.abs(pendulum_energy(s[0], s[1]) - e0))
    dev = np.array(dev)
    scale = G_PEND / 2.0                                  # m g l / 2
    assert dev.max() < 0.15 * scale * tol_scale, dev.max()             # measured 0.6 = 0.12 (explicit Euler: 92)
    assert (dev[800:].max(0) < 1.05 * dev[:200].max(0) + 1e-4 * tol_scale).all()      # no secular drift
    # small oscillation about the hanging position: period 2 pi / sqrt(3 g / (2 l)) = 1.6223 s = 32.446 steps
    s = np.stack([np.array([np.pi - 0.01]), np.zeros(1)])
    x, prev, cross = [], None, []
    for k in range(4000):
        s, _, _ = dyn(s, np.zeros(1))
        d = float(s[0][0] - np.pi)
        if prev is not None and prev < 0 <= d:
            cross.append(k - 1 + (-prev) / (d - prev))
        prev = d
    period = (cross[-1] - cross[0]) / (len(cross) - 1) * 0.05
    want = 2 * np.pi / np.sqrt(1.5 * G_PEND)
    assert abs(period / want - 1.0) < 4e-3, (period, want)        # g = 9.8 would be 1.0 % off


def check_cartpole_equations_of_motion(dyn, atol, theta_range=0.25):
    """theta_range 0.25: some transitions leave the 12-degree band (the done rule is checked too); 0.15: none does
    (for a stepper that replaces a finished episode's state at once)."""
    rng = np.random.default_rng(1)
    n = 4096
    s = np.stack([rng.uniform(-2, 2, n), rng.uniform(-2, 2, n), rng.uniform(-theta_range, theta_range, n),
                  rng.uniform(-2, 2, n)])
    a = rng.integers(0, 2, n)
    ns, rew, done = dyn(s, a)
    tau = 0.02
    xacc, thacc = (ns[1] - s[1]) / tau, (ns[3] - s[3]) / tau
    force = np.where(a == 1, 10.0, -10.0)
    r1, r2 = cartpole_residuals(s, force, xacc, thacc)
    np.testing.assert_allclose(r1, 0, atol=atol)
    np.testing.assert_allclose(r2, 0, atol=atol)
    np.testing.assert_allclose(ns[0] - s[0], tau * s[1], atol=atol * tau)          # explicit Euler on the positions
    np.testing.assert_allclose(ns[2] - s[2], tau * s[3], atol=atol * tau)
    np.testing.assert_array_equal(np.asarray(rew), 1.0)
    want_done = (np.abs(ns[0]) > 2.4) | (np.abs(ns[2]) > 12 * 2 * np.pi / 360)
    np.testing.assert_array_equal(np.asarray(done).astype(bool), want_done)
    assert (want_done.any() and not want_done.all()) or theta_range < 0.2


def check_acrobot_power_balance(dsdt, rtol):
    """dE/dt along the restated derivative = torque * theta2' (central difference of the Lagrangian energy)."""
    rng = np.random.default_rng(2)
    n = 4096
    s = np.stack([rng.uniform(-np.pi, np.pi, n), rng.uniform(-np.pi, np.pi, n), rng.uniform(-4, 4, n), rng.uniform(-6, 6, n)])
    for torque in (-1.0, 0.0, 1.0):
        ds = dsdt(s, np.full(n, torque))
        np.testing.assert_allclose(ds[:2], s[2:], rtol=0, atol=1e-12)
        eps = 1e-6
        dE = (acrobot_energy(s + eps * ds) - acrobot_energy(s - eps * ds)) / (2 * eps)
        np.testing.assert_allclose(dE, torque * s[3], rtol=rtol, atol=rtol * 50)


def rk4_reference(dsdt, s, torque, dt, sub):
    h = dt / sub
    for _ in range(sub):
        k1 = dsdt(s, torque); k2 = dsdt(s + h / 2 * k1, torque); k3 = dsdt(s + h / 2 * k2, torque); k4 = dsdt(s + h * k3, torque)
        s = s + h / 6 * (k1 + 2 * k2 + 2 * k3 + k4)
    return s


def check_acrobot_step(dyn, dsdt64, tol, terminal_states=True):
    """One step = one classical RK4 step of 0.2 s: against a 400-substep integration of the same derivative the local
    error is O(dt^5) (a forward-Euler step of 0.2 s would be off by ~0.2, a dt of 0.1 by half the motion)."""
    rng = np.random.default_rng(3)
    n = 2048
    s = np.stack([rng.uniform(-1, 1, n), rng.uniform(-1, 1, n), rng.uniform(-1.5, 1.5, n), rng.uniform(-2, 2, n)])
    a = rng.integers(0, 3, n)
    ns, rew, term = dyn(s, a)
    fine = rk4_reference(dsdt64, s, a - 1.0, 0.2, 400)
    wrap = lambda x: np.mod(x + np.pi, 2 * np.pi) - np.pi
    np.testing.assert_allclose(wrap(ns[0] - fine[0]), 0, atol=tol)
    np.testing.assert_allclose(wrap(ns[1] - fine[1]), 0, atol=tol)
    np.testing.assert_allclose(ns[2:], fine[2:], atol=4 * tol)
    euler = s + 0.2 * dsdt64(s, a - 1.0)
    assert np.abs(euler[2:] - fine[2:]).max() > 50 * tol                      # the check can tell them apart
    # termination: the tip is more than one link length above the pivot; reward -1 until then, 0 on that step
    height = -np.cos(ns[0]) - np.cos(ns[0] + ns[1])
    np.testing.assert_array_equal(np.asarray(term).astype(bool), height > 1.0)
    np.testing.assert_array_equal(np.asarray(rew), np.where(height > 1.0, 0.0, -1.0))
    if not terminal_states:
        return
    # states on both sides of the line, one step of rest at (almost) zero velocity
    t1 = np.array([np.pi * 0.75, np.pi * 0.9, np.pi, 2.0, 2.2])
    s2 = np.stack([t1, np.zeros(5), np.zeros(5), np.zeros(5)])
    ns2, _, term2 = dyn(s2, np.ones(5, dtype=np.int64))
    h2 = -np.cos(ns2[0]) - np.cos(ns2[0] + ns2[1])
    np.testing.assert_array_equal(np.asarray(term2).astype(bool), h2 > 1.0)
    assert np.asarray(term2).any() and not np.asarray(term2).all()


def check_time_limits(step_fn):
    """done fires on exactly step 200 / 200 / 500 of an episode that does not end earlier, and elapsed restarts."""
    for env, limit in ((E.PENDULUM, 200), (E.CARTPOLE, 200), (E.ACROBOT, 500)):
        n = 4
        u0 = np.full((4, n), 0.5)
        state = E.reset_from_uniforms(env, u0[:E.STATE_DIM[env]], np.float64)          # the centre of the reset box
        if env == E.CARTPOLE:
            state = state + np.array([[0.0], [0.0], [0.001], [0.0]]) * np.arange(n)[None, :]
        elapsed = np.zeros(n, np.int32)
        fired = []
        for k in range(1, limit + 2):
            if env == E.PENDULUM:
                ac = np.zeros((1, n))
            elif env == E.CARTPOLE:                       # bang-bang balance: push towards the side the pole leans to
                ac = ((state[2] + 0.5 * state[3] + 0.05 * state[0] + 0.1 * state[1]) > 0).astype(np.int32)
            else:
                ac = np.ones(n, np.int32)                  # zero torque: the resting acrobot stays down
            state, elapsed, _, _, done = step_fn(env, state, elapsed, ac, u0[:E.STATE_DIM[env]])
            if np.asarray(done).any():
                fired.append(k)
                assert np.asarray(done).all() and (np.asarray(elapsed) == 0).all()
        assert fired == [limit], (env, fired)


# ------------------------------------------------------------------ the oracle against the physics (CPU)
def test_pendulum_oracle_obeys_the_rod_pendulum_equation():
    check_pendulum_equation_of_motion(oracle_dynamics(E.PENDULUM), atol=1e-9)
    check_pendulum_energy_and_period(oracle_dynamics(E.PENDULUM))


def test_cartpole_oracle_obeys_the_cart_and_rod_equations():
    check_cartpole_equations_of_motion(oracle_dynamics(E.CARTPOLE), atol=1e-9)


def test_acrobot_oracle_conserves_energy_and_steps_rk4():
    dsdt = lambda s, a: E._acrobot_dsdt(s, a, np.float64)
    check_acrobot_power_balance(dsdt, rtol=1e-6)
    check_acrobot_step(oracle_dynamics(E.ACROBOT), dsdt, tol=2e-3)
    # convergence order of the integrator on the energy: zero torque, total time 4 s, dt -> dt / 2 cuts the error ~32x/2
    s0 = np.array([[1.0], [0.5], [0.0], [0.0]])
    errs = []
    for sub in (1, 2, 4):
        s = s0.copy()
        for _ in range(20 * sub):
            s = rk4_reference(dsdt, s, 0.0, 0.2 / sub, 1)
        errs.append(abs(float(acrobot_energy(s)[0] - acrobot_energy(s0)[0])))
    assert errs[0] / errs[1] > 12 and errs[1] / errs[2] > 12, errs       # O(dt^4) global: ratio ~16 (Euler: 2)


def test_time_limits_fire_on_the_documented_step():
    check_time_limits(lambda env, s, el, ac, ru: E.step(env, s, el, ac, ru, None, np.float64))


# ------------------------------------------------------------------ tf.multinomial restatement
def chi2_threshold(dof):
    """p ~ 1e-7 upper quantile of chi^2(dof), Wilson-Hilferty."""
    z = 5.2
    return dof * (1 - 2.0 / (9 * dof) + z * np.sqrt(2.0 / (9 * dof))) ** 3


def check_sampler_distribution(sample, n=1_000_000, given_u=True):
    """sample(logits[B, k], u[B]) -> actions: counts follow softmax(logits) (chi^2), u = 0 gives the first category,
    u -> 1 the last one."""
    rng = np.random.default_rng(4)
    for logits in ([0.3, -1.2], [2.0, 0.0, -1.0], [-0.5, -0.5, 3.0, 1.0]):
        lg = np.broadcast_to(np.asarray(logits, np.float32), (n, len(logits)))
        u = rng.random(n).astype(np.float32) if given_u else None        # None: the sampler draws its own uniforms
        a = np.asarray(sample(lg, u))
        p = np.exp(np.asarray(logits, np.float64))
        p /= p.sum()
        cnt = np.bincount(a, minlength=len(logits)).astype(np.float64)
        chi2 = float(np.sum((cnt - n * p) ** 2 / (n * p)))
        assert chi2 < chi2_threshold(len(logits) - 1), (logits, chi2, cnt / n, p)
        if not given_u:
            continue
        edge = np.asarray(sample(lg[:2], np.array([0.0, 1.0 - 2.0 ** -24], np.float32)))
        assert edge[0] == 0 and edge[1] == len(logits) - 1
    if not given_u:
        return
    # monotone in u: the action index never decreases as u grows (inverse CDF)
    lg = np.tile(np.asarray([0.1, 0.7, -0.3, 0.2], np.float32), (1001, 1))
    a = np.asarray(sample(lg, np.linspace(0, 1, 1001, endpoint=False).astype(np.float32)))
    assert (np.diff(a) >= 0).all() and a[0] == 0 and a[-1] == 3


def test_categorical_sampler_draws_from_softmax():
    check_sampler_distribution(P.cat_sample)
