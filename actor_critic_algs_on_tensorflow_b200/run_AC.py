"""``python -m actor_critic_algs_on_tensorflow_b200.run_AC [flags]`` -- the single-process actor-critic of
``Basic_AC/run_AC.py`` with its flag set (``Basic_AC/run_AC.py:11-22``) on the fused GPU path.

The reference collects at most 7 whole episodes / EP_LENGTH_STOP steps from ONE environment per update
(``Basic_AC/run_AC.py:172,212``).  Here one update consumes ``--n_envs`` environments x ``--horizon`` steps, rolled out,
scanned and differentiated by ``trainer.A2CTrainer(basic=True)`` inside CUDA graphs; what makes the algorithm Basic_AC
rather than A3C is kept:

* the critic's value head reads the SECOND hidden layer, the third never gets a gradient (``Basic_AC/policies.py:134-135``);
* actor gradients clipped to +-1 (``Basic_AC/policies.py:81``), actor lr within [1e-6, 1.0] (``Basic_AC/run_AC.py:178``);
* ``mu_layer`` starts at 0.1 x xavier (``Basic_AC/policies.py:5-6,48``);
* the log row of every update (``Basic_AC/run_AC.py:277-278``; stdout block every 20), flushed at i % 100 == 50;
* a checkpoint ``<checkpoint_dir>-<Env>-<episodes so far>`` with the ``Actor/...`` / ``Critic/...`` variables every
  ``--save_every`` updates (``Basic_AC/run_AC.py:171,282-284``).

With ``--n_envs 8 --horizon 100`` an update holds the 800 steps of the reference's own batch (BASELINE.json configs[0]).
"""
import argparse
import os

import numpy as np

from . import util as U
from .process import _emit_log_rows, trainer_checkpoint_tensors
from .tf_bundle import write_bundle

MAX_LR, MIN_LR, GRAD_CLIP = 1.0, 1e-6, 1.0


def build_parser():
    parser = argparse.ArgumentParser(formatter_class=argparse.ArgumentDefaultsHelpFormatter)
    parser.add_argument("--animate", default=False, action='store_true')      # accepted; no graphics stack here
    parser.add_argument("--env", default='Pendulum-v0')
    parser.add_argument("--seed", default=12321, type=int)
    parser.add_argument("--tboard", default=False, action='store_true')
    parser.add_argument("--save_every", default=600, type=int)
    parser.add_argument("--outdir", default='log.txt')
    parser.add_argument("--checkpoint_dir", default=os.path.join('tmp', 'checkpoints'))
    parser.add_argument("--frames", default=1, type=int)
    parser.add_argument("--mode", choices=["train", "debug"], default="train")
    parser.add_argument("--desired_kl", default=0.002, type=float)
    parser.add_argument("--iters", default=5000000, type=int, help="(not in the reference) stop after this many updates")
    parser.add_argument("--n_envs", default=1024, type=int, help="(not in the reference) environments per update")
    parser.add_argument("--horizon", default=20, type=int, help="(not in the reference) steps per environment and update")
    return parser


def basic_initial_weights(env_ob, num_ac, cont, seed):
    """glorot-uniform kernels, zero biases; the Gaussian head's kernel scaled by 0.1 (Basic_AC/policies.py:5-6,48)."""
    from . import _lib as L
    from .trainer import glorot_flat
    gen = np.random.default_rng(seed)
    actor = glorot_flat(0, env_ob, num_ac, cont, gen)
    critic = glorot_flat(1, env_ob, 1, False, gen)
    if cont:
        off, size = L.params_layout(0, env_ob, num_ac, cont)[0]["Wh"]
        actor[off:off + size] *= np.float32(0.1)
    return actor, critic


def main(argv=None):
    from .trainer import A2CTrainer, env_dims, ENV_IDS
    args = build_parser().parse_args(argv)
    prefix = args.checkpoint_dir + '-' + args.env.split('-')[0]
    _, env_ob, n_act, _ = env_dims(ENV_IDS[args.env])
    cont = n_act == 0
    print('Continuous Action Space. Action Scale is {}.'.format(2.0) if cont else
          'Discrete Action Space. Numer of actions is {}.'.format(n_act))
    actor0, critic0 = basic_initial_weights(env_ob * args.frames, 1 if cont else n_act, cont, args.seed)
    tr = A2CTrainer(args.env, n_envs=args.n_envs, horizon=args.horizon, seed=args.seed, desired_kl=args.desired_kl,
                    min_lr=MIN_LR, max_lr=MAX_LR, clip=GRAD_CLIP, frames=args.frames, basic=True, diagnostics=True,
                    actor_params=actor0, critic_params=critic0)
    logger = U.Logger(logfile=args.outdir)
    window = min(20, tr.log_rows)
    logged, episodes = 0, 0
    for i in range(args.iters):
        tr.iteration()
        if i + 1 - logged >= window or i + 1 == args.iters:
            rows = tr.read_log(logged, i + 1 - logged)
            _emit_log_rows(tr, logger, None, logged, i + 1 - logged, 20, rows=rows)
            if logged <= (i // args.save_every) * args.save_every <= i:        # a save_every multiple fell into the window
                at, ct = trainer_checkpoint_tensors(tr, "Actor", "Critic")
                at.update(ct)
                write_bundle("%s-%d" % (prefix, episodes), at)
            episodes += sum(r['episodes'] for r in rows)
            logged = i + 1
    logger.flush()
    return tr


if __name__ == '__main__':
    main()
