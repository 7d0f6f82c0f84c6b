"""Batched numpy restatement of the gym classic-control dynamics the reference steps
through ``env.reset()`` / ``env.step(ac)`` (``A3C/process.py:79,91``; env built by
``gym.make`` at ``A3C/process.py:104``).  TEST INFRASTRUCTURE (see ``oracle/__init__``).

gym is a third-party, un-vendored, un-pinned dependency of the reference
(``README.md:6`` names "OpenAI gym" with no version; the 4-tuple ``step`` API and
``env.seed`` put it at gym ~0.9.x, Sept 2017) -> **parity unpinned**: the formulas
below restate gym's published ``classic_control`` sources.

State is structure-of-arrays ``state[k, n]`` (k = state component, n = env), the
layout the CUDA kernel uses.  ``dtype=np.float64`` is gym's arithmetic;
``dtype=np.float32`` mirrors the kernel op-by-op (no FMA contraction) for Pendulum and
CartPole so that only sin/cos rounding can differ.

Auto-reset contract shared with the kernel (the reference restarts an episode by
calling ``rollout`` again, ``A3C/process.py:221-222``): ``step`` returns the reward and
``done`` of the transition; where ``done`` the state is replaced by a fresh reset
state drawn from the supplied uniforms, ``elapsed`` returns to 0 and the returned
observation is the first observation of the new episode.  ``done`` includes gym's
``TimeLimit`` (``elapsed >= max_episode_steps``), which the reference treats as a true
terminal (``A3C/process.py:63``).
"""
import numpy as np

PENDULUM, CARTPOLE, ACROBOT = 0, 1, 2
ENV_IDS = {"Pendulum-v0": PENDULUM, "CartPole-v0": CARTPOLE, "Acrobot-v1": ACROBOT}
STATE_DIM = {PENDULUM: 2, CARTPOLE: 4, ACROBOT: 4}
OBS_DIM = {PENDULUM: 3, CARTPOLE: 4, ACROBOT: 6}
MAX_EPISODE_STEPS = {PENDULUM: 200, CARTPOLE: 200, ACROBOT: 500}
N_ACTIONS = {PENDULUM: 0, CARTPOLE: 2, ACROBOT: 3}


# ------------------------------------------------------------------ reset
def reset_from_uniforms(env, u, dtype=np.float64):
    """u[k, n] uniforms in [0,1) -> initial state[k, n] (gym ``reset``)."""
    f = dtype
    u = np.asarray(u).astype(f)
    if env == PENDULUM:   # th ~ U(-pi, pi), thdot ~ U(-1, 1)
        pi = f(np.pi)
        return np.stack([f(-1.0) * pi + (f(2.0) * pi) * u[0], f(-1.0) + f(2.0) * u[1]]).astype(f)
    if env == CARTPOLE:   # U(-0.05, 0.05)^4
        return (f(-0.05) + f(0.1) * u[:4]).astype(f)
    if env == ACROBOT:    # U(-0.1, 0.1)^4
        return (f(-0.1) + f(0.2) * u[:4]).astype(f)
    raise ValueError(env)


def observe(env, s, dtype=np.float64):
    f = dtype
    s = s.astype(f)
    if env == PENDULUM:
        return np.stack([np.cos(s[0]), np.sin(s[0]), s[1]]).astype(f)
    if env == CARTPOLE:
        return s.copy()
    if env == ACROBOT:
        return np.stack([np.cos(s[0]), np.sin(s[0]), np.cos(s[1]), np.sin(s[1]), s[2], s[3]]).astype(f)
    raise ValueError(env)


# ------------------------------------------------------------------ dynamics
def _pendulum(s, a, f):
    th, thdot = s[0], s[1]
    max_speed, max_torque, dt, g, m, l = f(8.0), f(2.0), f(0.05), f(10.0), f(1.0), f(1.0)
    pi = f(np.pi)
    u = np.clip(a.astype(f), -max_torque, max_torque)
    two_pi = f(2.0) * pi
    an = np.remainder(th + pi, two_pi) - pi          # angle_normalize, floored mod
    cost = an * an + f(0.1) * (thdot * thdot) + f(0.001) * (u * u)
    c1 = f(-3.0) * g / (f(2.0) * l)                  # -15
    c2 = f(3.0) / (m * l * l)                        # 3
    newthdot = thdot + (c1 * np.sin(th + pi) + c2 * u) * dt
    newth = th + newthdot * dt
    newthdot = np.clip(newthdot, -max_speed, max_speed)
    ns = np.stack([newth, newthdot]).astype(f)
    return ns, (-cost).astype(f), np.zeros(th.shape, dtype=bool)


def _cartpole(s, a, f):
    x, x_dot, theta, theta_dot = s
    gravity, masscart, masspole = f(9.8), f(1.0), f(0.1)
    total_mass = f(masspole + masscart) if f is np.float64 else f(np.float32(0.1) + np.float32(1.0))
    length = f(0.5)
    polemass_length = f(masspole * length)
    force_mag, tau = f(10.0), f(0.02)
    theta_thr = f(12 * 2 * np.pi / 360)
    x_thr = f(2.4)
    force = np.where(np.asarray(a) == 1, force_mag, -force_mag).astype(f)
    costheta, sintheta = np.cos(theta), np.sin(theta)
    temp = (force + polemass_length * theta_dot * theta_dot * sintheta) / total_mass
    thetaacc = (gravity * sintheta - costheta * temp) / (
        length * (f(4.0) / f(3.0) - masspole * costheta * costheta / total_mass))
    xacc = temp - polemass_length * thetaacc * costheta / total_mass
    x = x + tau * x_dot
    x_dot = x_dot + tau * xacc
    theta = theta + tau * theta_dot
    theta_dot = theta_dot + tau * thetaacc
    done = (x < -x_thr) | (x > x_thr) | (theta < -theta_thr) | (theta > theta_thr)
    ns = np.stack([x, x_dot, theta, theta_dot]).astype(f)
    return ns, np.ones(x.shape, dtype=f), done


def _acrobot_dsdt(s, a, f):
    m1 = m2 = l1 = f(1.0)
    lc1 = lc2 = f(0.5)
    I1 = I2 = f(1.0)
    g = f(9.8)
    pi = f(np.pi)
    theta1, theta2, dtheta1, dtheta2 = s
    d1 = m1 * lc1 ** 2 + m2 * (l1 ** 2 + lc2 ** 2 + f(2.0) * l1 * lc2 * np.cos(theta2)) + I1 + I2
    d2 = m2 * (lc2 ** 2 + l1 * lc2 * np.cos(theta2)) + I2
    phi2 = m2 * lc2 * g * np.cos(theta1 + theta2 - pi / f(2.0))
    phi1 = (-m2 * l1 * lc2 * dtheta2 ** 2 * np.sin(theta2)
            - f(2.0) * m2 * l1 * lc2 * dtheta2 * dtheta1 * np.sin(theta2)
            + (m1 * lc1 + m2 * l1) * g * np.cos(theta1 - pi / f(2.0)) + phi2)
    ddtheta2 = (a + d2 / d1 * phi1 - m2 * l1 * lc2 * dtheta1 ** 2 * np.sin(theta2) - phi2) / (
        m2 * lc2 ** 2 + I2 - d2 ** 2 / d1)
    ddtheta1 = -(d2 * ddtheta2 + phi1) / d1
    return np.stack([dtheta1, dtheta2, ddtheta1, ddtheta2]).astype(f)


def _wrap(x, m, M):
    diff = M - m
    x = x.copy()
    for _ in range(64):      # gym: while x > M: x -= diff ; while x < m: x += diff
        hi = x > M
        if not hi.any():
            break
        x[hi] -= diff
    for _ in range(64):
        lo = x < m
        if not lo.any():
            break
        x[lo] += diff
    return x


def _acrobot(s, a, f):
    dt = f(0.2)
    pi = f(np.pi)
    torque = (np.asarray(a).astype(f) - f(1.0))          # AVAIL_TORQUE = [-1, 0, +1]
    s = s.astype(f)
    dt2 = dt / f(2.0)
    k1 = _acrobot_dsdt(s, torque, f)
    k2 = _acrobot_dsdt(s + dt2 * k1, torque, f)
    k3 = _acrobot_dsdt(s + dt2 * k2, torque, f)
    k4 = _acrobot_dsdt(s + dt * k3, torque, f)
    ns = s + dt / f(6.0) * (k1 + f(2.0) * k2 + f(2.0) * k3 + k4)
    ns = ns.astype(f)
    ns[0] = _wrap(ns[0], -pi, pi)
    ns[1] = _wrap(ns[1], -pi, pi)
    ns[2] = np.clip(ns[2], f(-4.0) * pi, f(4.0) * pi)
    ns[3] = np.clip(ns[3], f(-9.0) * pi, f(9.0) * pi)
    terminal = (-np.cos(ns[0]) - np.cos(ns[1] + ns[0])) > f(1.0)
    rew = np.where(terminal, f(0.0), f(-1.0)).astype(f)
    return ns, rew, terminal


_DYN = {PENDULUM: _pendulum, CARTPOLE: _cartpole, ACROBOT: _acrobot}


def dynamics(env, state, action, dtype=np.float64):
    """One raw transition, no TimeLimit / auto-reset: (next_state, reward, terminal)."""
    a = np.asarray(action)
    if env == PENDULUM and a.ndim == 2:
        a = a[0]
    return _DYN[env](np.asarray(state).astype(dtype), a, dtype)


def step(env, state, elapsed, action, reset_u, max_steps=None, dtype=np.float64):
    """Transition + TimeLimit + auto-reset (contract in the module docstring).

    Returns (state', elapsed', obs', reward, done) with done as uint8.
    """
    if max_steps is None:
        max_steps = MAX_EPISODE_STEPS[env]
    ns, rew, term = dynamics(env, state, action, dtype)
    el = np.asarray(elapsed).astype(np.int32) + 1
    done = term | (el >= max_steps)
    rs = reset_from_uniforms(env, reset_u, dtype)
    ns = np.where(done[None, :], rs, ns).astype(dtype)
    el = np.where(done, 0, el).astype(np.int32)
    return ns, el, observe(env, ns, dtype), rew.astype(dtype), done.astype(np.uint8)
