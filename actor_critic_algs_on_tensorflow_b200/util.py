"""Host-side utilities with the reference's names and behaviour (``A3C/util.py``):
``var_accounted_for`` (:5-13), ``LinearSchedule`` (:16-42), ``make_np`` (:45-49) and ``Logger``
(:55-111, same stdout block and the same 8-column text file, including the quirk that the step
column restarts after every flush because ``last_write`` is assigned, not accumulated, :111).
Pure host logic: nothing here touches the GPU.
"""
import os

import numpy as np


def var_accounted_for(target, pred):
    """Correlation of the standardised predictions and targets, in [-1, 1] (logging only)."""
    p = np.asarray(pred, dtype=np.float64).reshape(-1)
    t = np.asarray(target, dtype=np.float64).reshape(-1)
    p = (p - p.mean()) / p.std()
    t = (t - t.mean()) / t.std()
    return np.mean(t * p)


class LinearSchedule(object):
    """Piecewise-linear annealing: ``init_val`` up to ``init_t``, ``end_val`` after ``end_t``, linear in
    between; ``update_time(t)`` says whether step ``t`` is one of the every-``update_every_t`` refreshes."""

    def __init__(self, init_t, end_t, init_val, end_val, update_every_t):
        self.init_t, self.end_t = init_t, end_t
        self.init_val, self.end_val = init_val, end_val
        self.update_every_t = update_every_t

    def val(self, t):
        if t < self.init_t:
            return self.init_val
        if t > self.end_t:
            return self.end_val
        span = float(self.end_t - self.init_t)
        return ((t - self.init_t) * self.end_val + (self.end_t - t) * self.init_val) / span

    def update_time(self, t):
        return t % self.update_every_t == 0


def make_np(*lists):
    """Generator of ``np.array(x)`` for every argument (the reference returns a generator too)."""
    return (np.array(x) for x in lists)


_COLUMNS = ("rews", "ev_before", "ev_after", "act_loss", "circ_loss", "kl_dist", "ents")


class Logger(object):
    """Collects per-iteration training statistics, echoes a block to stdout when asked, and appends
    rows to ``logfile`` on ``flush()``."""
    HEADER = "step avg_rew ev_before ev_after act_loss crit_loss kl_dist avg_ent\n"
    ROW = "%d %.4f %.4f  %.4f %.4f  %.4f %.4f %.4f\n"

    def __init__(self, logfile):
        self.logfile = logfile
        folder = os.path.dirname(logfile)
        if folder != "" and not os.path.exists(folder):
            os.makedirs(folder, exist_ok=True)
        self.f = open(logfile, "w")
        self.last_write = 0
        self.f.write(self.HEADER)
        self._reset()

    def _reset(self):
        for c in _COLUMNS:
            setattr(self, c, [])

    def __call__(self, t, act_loss, circ_loss, kl_dist, avg_rew, print_tog, act_lr, avg_ent, worker_id=0,
                 ev_before=-1, ev_after=-1):
        if print_tog:
            print("\nIteration %d" % t)
            print("EpRewMean %.4f " % avg_rew)
            print("EV Before %f" % ev_before)
            print("EV After %f" % ev_after)
            print("Act losses %.4f  " % act_loss)
            print("Critic loss  %.4f" % circ_loss)
            print("Actor lr %f" % act_lr)
            print("KL dist %.4f" % kl_dist)
            print("Avg Ent %.4f" % avg_ent)
            print("Performed by worker %d" % worker_id)
        for c, v in zip(_COLUMNS, (avg_rew, ev_before, ev_after, act_loss, circ_loss, kl_dist, avg_ent)):
            getattr(self, c).append(v)

    def flush(self):
        n = len(self.rews)
        for i in range(n):
            self.f.write(self.ROW % (i + self.last_write, self.rews[i], self.ev_before[i], self.ev_after[i],
                                     self.act_loss[i], self.circ_loss[i], self.kl_dist[i], self.ents[i]))
        self.f.flush()
        self._reset()
        self.last_write = +n          # reference behaviour (A3C/util.py:111): assigned, not accumulated

    def __del__(self):
        try:
            self.f.close()
        except Exception:
            pass
