"""``python -m actor_critic_algs_on_tensorflow_b200.run_AC [flags]`` -- the single-process actor-critic of
``Basic_AC/run_AC.py`` (same flags, :11-22) in its reference shape: whole episodes from one environment,
batch-1 ``act`` per step, at most 7 rollouts / EP_LENGTH_STOP steps per update (:172,212), critic update then
actor update (:250-251), KL-driven lr with max_lr = 1.0 (:178,261-266), beta/gamma schedules (:268-275),
checkpoint ``<checkpoint_dir>-<Env>-<tot_rolls>`` every ``save_every`` iterations (:171,282-283).
Every arithmetic step runs in the CUDA library through the ``Actor`` / ``Critic`` / ``PathAdv`` / env objects.
The Basic_AC quirks are kept: gradient clip +-1 (policies.py:81), 0.1 x xavier ``mu_layer`` (:5-6,48), and the
critic whose value head reads the second layer (:134-135; its third layer never receives a gradient).
"""
import argparse
import os

import numpy as np

from . import policies as pol
from . import util as U
from .process import Framer, PathAdv, _space_params, get_roll_params, rollout, train_actor
from .tf_bundle import write_bundle


def build_parser():
    parser = argparse.ArgumentParser(formatter_class=argparse.ArgumentDefaultsHelpFormatter)
    parser.add_argument("--animate", default=False, action='store_true')
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
    return parser


def train_ciritic(critic, sess, obs, targets):
    """Basic_AC variant: also returns ev_after from a third forward (run_AC.py:109-116)."""
    assert len(obs) == len(targets)
    ev_before = U.var_accounted_for(targets, critic.value(obs, sess=sess))
    loss, _ = critic.optimize(obs=obs, targets=targets, sess=sess)
    ev_after = U.var_accounted_for(targets, critic.value(obs, sess=sess))
    return loss, ev_before, ev_after


def main(argv=None):
    args = build_parser().parse_args(argv)
    checkpoint_path = args.checkpoint_dir + '-' + args.env.split('-')[0]
    MAX_ROLLS = 7
    env, MAX_PATH_LENGTH, EP_LENGTH_STOP = get_roll_params(args.env, basic=True)
    print('\nMAX PATH LENGTH, EP LENGTH STEP: {}, {}\n'.format(MAX_PATH_LENGTH, EP_LENGTH_STOP))
    desired_kl = args.desired_kl
    max_lr, min_lr = 1., 1e-6
    framer = Framer(frame_num=args.frames)
    log_gamma_schedule = U.LinearSchedule(init_t=100, end_t=3000, init_val=-2, end_val=-8, update_every_t=100)
    log_beta_schedule = U.LinearSchedule(init_t=100, end_t=3000, init_val=0, end_val=-4, update_every_t=100)
    rew_to_advs = PathAdv(gamma=0.98, look_ahead=40)
    logger = U.Logger(logfile=args.outdir)
    np.random.seed(args.seed)
    env.seed(args.seed)
    act_type, ac_dim, ac_scale = _space_params(env)
    if act_type == 'disc':
        print('Discrete Action Space. Numer of actions is {}.'.format(ac_dim))
    else:
        print('Continuous Action Space. Action Scale is {}.'.format(ac_scale))
    ob_dim = env.observation_space.shape[0] * args.frames
    critic = pol.Critic(num_ob_feat=ob_dim, seed=args.seed, basic_head=True)
    actor = pol.Actor(num_ob_feat=ob_dim, ac_dim=ac_dim, act_type=act_type, ac_scale=ac_scale, seed=args.seed + 1,
                      grad_clip=1.0, head_init="basic")
    tot_rolls = 0
    for i in range(args.iters):
        ep_obs, ep_advs, ep_logps, ep_target_vals, ep_acs, ep_rews = [], [], [], [], [], []
        tot_rews, tot_ent, rolls = 0, 0, 0
        while len(ep_rews) < EP_LENGTH_STOP and rolls < MAX_ROLLS:
            path = rollout(env=env, sess=None, policy=actor.act, max_path_length=MAX_PATH_LENGTH, framer=framer,
                           render=False)
            obs_aug = framer.full(path['obs'])
            ep_obs += obs_aug[:-1]
            ep_logps += path['logps']
            ep_acs += path['acs']
            tot_ent += path['entropy']
            obs_vals = critic.value(obs=obs_aug, sess=None).reshape(-1)
            target_val, advs = rew_to_advs(rews=path['rews'], terminal=path['terminated'], vals=obs_vals)
            ep_target_vals += list(target_val)
            ep_advs += list(advs)
            ep_rews += path['rews']
            tot_rews += sum(path['rews'])
            rolls += 1
        avg_rew = float(tot_rews) / rolls
        avg_ent = tot_ent / float(len(ep_logps))
        ep_obs, ep_advs, ep_logps, ep_target_vals, ep_acs, ep_rews = U.make_np(ep_obs, ep_advs, ep_logps, ep_target_vals,
                                                                               ep_acs, ep_rews)
        ep_advs = (ep_advs - np.mean(ep_advs)) / (1e-8 + np.std(ep_advs))
        cir_loss, ev_before, ev_after = train_ciritic(critic=critic, sess=None, obs=ep_obs, targets=ep_target_vals)
        act_loss = train_actor(actor=actor, sess=None, obs=ep_obs, advs=ep_advs, acs=ep_acs, logps=ep_logps)
        act_lr, cur_beta, cur_gamma = actor.get_opt_param(None)
        kl_dist = actor.get_kl(sess=None, obs=ep_obs, logp_feeds=ep_logps, acs=ep_acs)
        if kl_dist < desired_kl / 4:
            actor.set_opt_param(sess=None, new_lr=min(max_lr, act_lr * 1.5))
        elif kl_dist > desired_kl * 4:
            actor.set_opt_param(sess=None, new_lr=max(min_lr, act_lr / 1.5))
        if log_gamma_schedule.update_time(i):
            new_gamma = np.power(10., log_gamma_schedule.val(i))
            actor.set_opt_param(sess=None, new_gamma=new_gamma)
            print('\nUpdated gamma from %.4f to %.4f.' % (cur_gamma, new_gamma))
        if log_beta_schedule.update_time(i):
            new_beta = np.power(10., log_beta_schedule.val(i))
            actor.set_opt_param(sess=None, new_beta=new_beta)
            print('Updated beta from %.4f to %.4f.' % (cur_beta, new_beta))
        logger(i, act_loss=act_loss, circ_loss=np.sqrt(cir_loss), avg_rew=avg_rew, ev_before=ev_before, ev_after=ev_after,
               act_lr=act_lr, print_tog=(i % 20) == 0, kl_dist=kl_dist, avg_ent=avg_ent)
        if i % 100 == 50:
            logger.flush()
        if i % args.save_every == 0:
            t = actor.checkpoint_tensors("Actor")
            t.update(critic.checkpoint_tensors("Critic"))
            write_bundle("%s-%d" % (checkpoint_path, tot_rolls), t)
        tot_rolls += rolls
    logger.flush()
    return actor, critic


if __name__ == '__main__':
    main()
