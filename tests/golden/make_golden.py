"""Generate the golden vectors under tests/golden/ from the REFERENCE'S OWN CODE.

Run in the build container only (``/root/reference`` does not exist on the GPU box):

    python tests/golden/make_golden.py

It imports ``/root/reference/A3C/{util,process}.py`` with ``tensorflow``/``gym``/``scipy``
replaced by MagicMock stubs (their module-level uses are absorbed by the stubs) and runs the
reference's ``PathAdv``, ``Framer``, ``LinearSchedule``, ``var_accounted_for``, ``make_np``
and ``Logger`` on seeded inputs.  Outputs:

  pathadv.npz        per-episode PathAdv cases (rews, vals, terminal, gamma, L) -> (G, adv)
  pathadv_cols.npz   [T,N] column batches with done flags, assembled from per-episode
                     reference calls exactly as A3C/process.py:221-235 does
  framer.npz, schedule.npz, vaf.npz, advnorm.npz
  logger_golden.log  text file written by the reference's Logger
  model-Pendulum_a3c.{index,data-00000-of-00001}   copy of the reference's shipped checkpoint
                     fixture (demos/), the only numeric fixture the reference owns

Nothing here is imported at test time; tests read only the files it wrote.
"""
import os
import shutil
import sys
from unittest.mock import MagicMock

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def load_reference():
    for m in ("tensorflow", "gym", "scipy", "scipy.signal"):
        sys.modules[m] = MagicMock()
    sys.path.insert(0, os.path.join(REF, "A3C"))
    import process as P   # noqa
    import util as U      # noqa
    return P, U


def main():
    P, U = load_reference()
    rng = np.random.default_rng(12321)

    # ---- PathAdv, per episode
    cases = [(20, 40, False), (20, 40, True), (200, 40, True), (200, 40, False), (57, 5, False), (5, 5, True),
             (6, 5, True), (1, 40, True), (1, 40, False), (41, 40, True), (40, 40, False), (3, 1, True)]
    out = {"n_cases": len(cases)}
    for c, (n, L, term) in enumerate(cases):
        gamma = [0.98, 0.9, 0.99][c % 3]
        rews = rng.normal(size=n) * 3
        vals = rng.normal(size=n + 1) * 10
        G, adv = P.PathAdv(gamma=gamma, look_ahead=L)(rews=list(rews), vals=vals.astype(np.float32), terminal=term)
        out.update({"rews_%d" % c: rews, "vals_%d" % c: vals.astype(np.float32), "term_%d" % c: term,
                    "gamma_%d" % c: gamma, "L_%d" % c: L, "G_%d" % c: np.array(G), "adv_%d" % c: np.array(adv)})
    np.savez(os.path.join(HERE, "pathadv.npz"), **out)

    # ---- PathAdv over [T,N] columns with done flags (auto-reset batches)
    colcases = [(20, 64, 40, 0.98, 0.0), (20, 64, 40, 0.98, 0.08), (64, 32, 40, 0.98, 0.03), (200, 8, 40, 0.98, 0.01),
                (33, 40, 5, 0.9, 0.2), (7, 16, 1, 0.99, 0.3), (1, 16, 40, 0.98, 0.5)]
    out = {"n_cases": len(colcases)}
    for c, (T, N, L, gamma, pdone) in enumerate(colcases):
        rew = rng.normal(size=(T, N)) * 2
        V = (rng.normal(size=(T + 1, N)) * 5).astype(np.float32)
        done = (rng.random(size=(T, N)) < pdone)
        G = np.zeros((T, N))
        adv = np.zeros((T, N))
        pa = P.PathAdv(gamma=gamma, look_ahead=L)
        for n in range(N):
            s = 0
            while s < T:
                e = s
                while e < T and not done[e, n]:
                    e += 1
                term = e < T            # a done flag closed this segment
                e = min(e + 1, T) if term else T
                g, a = pa(rews=list(rew[s:e, n]), vals=V[s:e + 1, n], terminal=term)
                G[s:e, n], adv[s:e, n] = g, a
                s = e
        out.update({"rew_%d" % c: rew, "V_%d" % c: V, "done_%d" % c: done.astype(np.uint8), "L_%d" % c: L,
                    "gamma_%d" % c: gamma, "G_%d" % c: G, "adv_%d" % c: adv})
    np.savez(os.path.join(HERE, "pathadv_cols.npz"), **out)

    # ---- Framer
    obs = [rng.normal(size=3) for _ in range(7)]
    out = {"obs": np.array(obs)}
    for k in (1, 2, 3):
        fr = P.Framer(frame_num=k)
        out["full_%d" % k] = np.array(fr.full(obs))
        out["last_%d" % k] = np.array(fr.last(obs))
        out["last3_%d" % k] = np.array(fr.last(obs[:3]))
    np.savez(os.path.join(HERE, "framer.npz"), **out)

    # ---- LinearSchedule (the two schedules of A3C/process.py:173-174)
    ts = np.arange(0, 3500, 7)
    g = U.LinearSchedule(init_t=100, end_t=3000, init_val=-2, end_val=-8, update_every_t=100)
    b = U.LinearSchedule(init_t=100, end_t=3000, init_val=0, end_val=-4, update_every_t=100)
    np.savez(os.path.join(HERE, "schedule.npz"), ts=ts, gamma=np.array([g.val(int(t)) for t in ts], dtype=np.float64),
             beta=np.array([b.val(int(t)) for t in ts], dtype=np.float64),
             upd=np.array([g.update_time(int(t)) for t in ts]))

    # ---- var_accounted_for + advantage normalisation (A3C/process.py:247, evaluated with the reference's make_np)
    tgt = rng.normal(size=500) * 4 + 1
    pred = tgt * 0.7 + rng.normal(size=500)
    np.savez(os.path.join(HERE, "vaf.npz"), target=tgt, pred=pred, vaf=U.var_accounted_for(target=tgt, pred=pred))
    advs = list(rng.normal(size=1200) * 3 + 0.5)
    (ep_advs,) = U.make_np(advs)
    np.savez(os.path.join(HERE, "advnorm.npz"), advs=np.array(advs),
             norm=(ep_advs - np.mean(ep_advs)) / (1e-8 + np.std(ep_advs)))

    # ---- Logger text format
    path = os.path.join(HERE, "logger_golden.log")
    lg = U.Logger(logfile=path)
    import io
    import contextlib
    with contextlib.redirect_stdout(io.StringIO()) as so:
        for t in range(5):
            lg(t, act_loss=0.1 * t - 0.3, circ_loss=1.5 + t, kl_dist=0.002 * t, avg_rew=-1000.0 + 13.37 * t,
               print_tog=(t % 2 == 0), act_lr=0.005, avg_ent=1.4189 - 0.01 * t, worker_id=1, ev_before=0.1 * t,
               ev_after=0.2 * t)
        lg.flush()
        for t in range(5, 8):
            lg(t, act_loss=0.5, circ_loss=2.5, kl_dist=0.01, avg_rew=-900.0, print_tog=False, act_lr=0.0075,
               avg_ent=1.3, worker_id=1)
        lg.flush()
    lg.f.close()
    with open(os.path.join(HERE, "logger_golden.stdout"), "w") as fh:
        fh.write(so.getvalue())

    # ---- the reference's shipped checkpoint fixture
    for ext in (".index", ".data-00000-of-00001"):
        shutil.copyfile(os.path.join(REF, "demos", "model-Pendulum_a3c" + ext),
                        os.path.join(HERE, "model-Pendulum_a3c" + ext))
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
