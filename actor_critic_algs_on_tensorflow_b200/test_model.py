"""``python -m actor_critic_algs_on_tensorflow_b200.test_model ENV MODEL_PATH [flags]`` -- restore a TF V2
checkpoint and roll out episodes, printing what ``A3C/test_model.py`` prints (sampled actions, as the
reference does).  ``--batched N`` evaluates N episodes in lock-step on the GPU instead (``--greedy``: action
= mu), which is BASELINE.json configs[4]."""
import argparse

import numpy as np

from .process import evaluate_batched, test_process


def main(argv=None):
    parser = argparse.ArgumentParser(formatter_class=argparse.ArgumentDefaultsHelpFormatter)
    parser.add_argument("env")
    parser.add_argument("model_path")
    parser.add_argument("--animate_not", default=False, action='store_true')
    parser.add_argument("--seed", default=12321, type=int)
    parser.add_argument("--frames", default=1, type=int)
    parser.add_argument("--num_episodes", default=3, help="For Test mode. How many episodes to render.", type=int)
    parser.add_argument("--batched", default=0, type=int, help="evaluate this many episodes in one GPU batch")
    parser.add_argument("--greedy", default=False, action='store_true', help="with --batched: action = mu")
    args = parser.parse_args(argv)
    print("\n************Test Mode**********\nUsing model path {}\n\n".format(args.model_path))
    if args.batched:
        ret = evaluate_batched(args.env, args.model_path, args.batched, greedy=args.greedy, seed=args.seed)
        print('Average reward over {} was {} (std {})'.format(args.batched, float(np.mean(ret)), float(np.std(ret))))
        return ret
    return test_process(random_seed=args.seed, animate=not args.animate_not, env_id=args.env, model_path=args.model_path,
                        stack_frames=args.frames, num_episodes=args.num_episodes)


if __name__ == '__main__':
    main()
