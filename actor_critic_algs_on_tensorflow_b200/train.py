"""``python -m actor_critic_algs_on_tensorflow_b200.train {ps,worker} TASK [flags]`` -- the A3C launcher
with the flag set of ``A3C/train.py:8-25``.  On one NVSwitch box there is no parameter-server process:
``ps`` jobs exit at once, each ``worker`` is one rank on one GPU (NCCL when ``--worker_num`` > 1; start
one process per worker as ``runner.sh`` does, or use torchrun -- RANK/WORLD_SIZE are honoured).
Extra flags (not in the reference): ``--n_envs``, ``--horizon``, ``--max_iters``, ``--async_mode``.
"""
import argparse
import os

from . import util as U
from .process import process_fn


def build_parser():
    parser = argparse.ArgumentParser(formatter_class=argparse.ArgumentDefaultsHelpFormatter)
    parser.add_argument("job", choices=["ps", "worker"])
    parser.add_argument("task", type=int)
    parser.add_argument("--animate", default=False, action='store_true')      # accepted; no graphics stack here
    parser.add_argument("--env", default='Pendulum-v0')
    parser.add_argument("--seed", default=12321, type=int)
    parser.add_argument("--tboard", default=False)                            # parsed and dropped, as in the reference
    parser.add_argument("--worker_num", default=4, type=int)
    parser.add_argument("--ps_num", default=2, type=int)
    parser.add_argument("--initport", default=2849, type=int)
    parser.add_argument("--stdout_freq", default=20, type=int)
    parser.add_argument("--save_every", default=600, type=int)
    parser.add_argument("--outdir", default=os.path.join('tmp', 'logs'))
    parser.add_argument("--checkpoint_dir", default=os.path.join('tmp', 'checkpoints'))
    parser.add_argument("--frames", default=1, type=int)
    parser.add_argument("--mode", choices=["train", "debug-light", "debug-full"], default="train")
    parser.add_argument("--desired_kl", default=0.002, type=float)
    parser.add_argument("--n_envs", default=4096, type=int, help="environments per worker (GPU batch)")
    parser.add_argument("--horizon", default=20, type=int, help="steps per env per update")
    parser.add_argument("--max_iters", default=None, type=int, help="stop after this many global steps")
    parser.add_argument("--async_mode", default=False, action='store_true',
                        help="stale-gradient parameter-server semantics instead of synchronous A2C")
    return parser


def main(argv=None):
    args = build_parser().parse_args(argv)
    animate = args.animate and args.task == 0 and args.job == 'worker'
    cluster = {'ps': ['localhost:{}'.format(args.initport + i) for i in range(args.ps_num)],
               'worker': ['localhost:{}'.format(i + args.ps_num + args.initport) for i in range(args.worker_num)]}
    log_file = os.path.join(args.outdir, 'worker_{}.log'.format(args.task)) if args.job == 'worker' else 'N/A'
    random_seed = args.seed + args.task
    checkpoint_basename = 'model' + '-' + args.env.split('-')[0]
    logger = U.Logger(logfile=log_file) if args.job == 'worker' else None
    print("Starting {} {} with log at {}".format(args.job, args.task, log_file))
    return process_fn(cluster=cluster, task_id=args.task, job=args.job, logger=logger, env_id=args.env, animate=animate,
                      random_seed=random_seed, save_path=args.checkpoint_dir, stack_frames=args.frames,
                      save_every=args.save_every, run_mode=args.mode, desired_kl=args.desired_kl,
                      checkpoint_basename=checkpoint_basename, stdout_freq=args.stdout_freq, n_envs=args.n_envs,
                      horizon=args.horizon, max_iters=args.max_iters, mode='async' if args.async_mode else 'sync')


if __name__ == '__main__':
    main()
