#!/usr/bin/env python
"""Train / evaluate a SoftTreeMax agent with the B200-native hot path — the flow of the reference's
main.py (flags of utils.py:60-95 that matter here), without wandb.

    python examples/train.py --env_name BreakoutNoFrameskip-v4 --tree_depth 2 --total_timesteps 2048
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np
import torch

from softtreemax_b200.envs import CuleEnv
from softtreemax_b200.policy import ActorCriticCnnTSPolicy
from softtreemax_b200.ppo import PPO, get_linear_fn


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--total_timesteps", type=int, default=2048)
    ap.add_argument("--learning_rate", type=float, default=2.5e-4)
    ap.add_argument("--seed", type=int, default=4)
    ap.add_argument("--env_name", default="BreakoutNoFrameskip-v4")
    ap.add_argument("--gamma", type=float, default=0.99)
    ap.add_argument("--tree_depth", type=int, default=2)
    ap.add_argument("--clip_reward", type=int, default=1)
    ap.add_argument("--learn_alpha", type=int, default=0)
    ap.add_argument("--learn_beta", type=int, default=1)
    ap.add_argument("--use_leaves_v", type=int, default=0)
    ap.add_argument("--is_cumulative_mode", type=int, default=0)
    ap.add_argument("--regularization", type=float, default=0.001)
    ap.add_argument("--hash_buffer_size", type=int, default=-1)
    ap.add_argument("--model_filename", default=None)
    args = ap.parse_args()
    torch.manual_seed(args.seed)
    np.random.seed(args.seed)
    env_kwargs = dict(env_name=args.env_name, color_mode="gray", repeat_prob=0.0, rescale=True, episodic_life=True,
                      frameskip=4)
    env = CuleEnv(device="cuda:0", env_kwargs=env_kwargs, clip_reward=bool(args.clip_reward), fire_reset=True)
    print("Environment:", args.env_name, "Num actions:", env.action_space.n, "Tree depth:", args.tree_depth)
    policy_kwargs = dict(step_env=env, gamma=args.gamma, tree_depth=args.tree_depth,
                         buffer_size=max(args.hash_buffer_size, 128), learn_alpha=bool(args.learn_alpha),
                         learn_beta=bool(args.learn_beta), max_width=-1, use_leaves_v=bool(args.use_leaves_v),
                         is_cumulative_mode=bool(args.is_cumulative_mode), regularization=args.regularization)
    model = PPO(ActorCriticCnnTSPolicy, env, learning_rate=get_linear_fn(args.learning_rate, 0, 1), n_epochs=3, gamma=0.99,
                n_steps=128, batch_size=32, ent_coef=0.01, vf_coef=1.0, gae_lambda=0.95,
                clip_range=get_linear_fn(0.1, 0, 1), policy_kwargs=policy_kwargs, verbose=1, seed=args.seed)
    model.learn(args.total_timesteps)
    if args.model_filename:
        model.policy.save(args.model_filename)
        print("Saved", args.model_filename)


if __name__ == "__main__":
    main()
