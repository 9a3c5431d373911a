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

from softtreemax_b200.callbacks import TrainingCallback
from softtreemax_b200.envs import CuleEnv, CuleEnvMultiple
from softtreemax_b200.policy import ActorCriticCnnPolicyDepth0, ActorCriticCnnTSPolicy
from softtreemax_b200.ppo import PPO, evaluate_policy, get_linear_fn


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
    ap.add_argument("--max_width", type=int, default=-1)
    ap.add_argument("--n_envs", type=int, default=256, help="depth-0 arm only (CuleEnvMultiple)")
    ap.add_argument("--run_type", default="train", choices=["train", "evaluate"])
    ap.add_argument("--n_eval_episodes", type=int, default=10)
    ap.add_argument("--log_jsonl", default=None, help="episode metrics (what the reference sends to wandb), one JSON per line")
    args = ap.parse_args()
    torch.manual_seed(args.seed)
    np.random.seed(args.seed)
    env_kwargs = dict(env_name=args.env_name, color_mode="gray", repeat_prob=0.0, rescale=True, episodic_life=True,
                      frameskip=4)
    fire_reset = args.env_name not in ["AsterixNoFrameskip-v4", "CrazyClimberNoFrameskip-v4", "FreewayNoFrameskip-v4",
                                       "MsPacmanNoFrameskip-v4", "SkiingNoFrameskip-v4", "TutankhamNoFrameskip-v4"]   # main.py:39-41
    depth0_train = args.tree_depth == 0 and args.run_type == "train"
    if depth0_train:                                                                              # main.py:42-45
        env = CuleEnvMultiple(device="cuda:0", env_kwargs=env_kwargs, clip_reward=bool(args.clip_reward),
                              fire_reset=fire_reset, n_envs=args.n_envs)
    else:
        env = CuleEnv(device="cuda:0", env_kwargs=env_kwargs, clip_reward=bool(args.clip_reward), fire_reset=fire_reset)
    print("Environment:", args.env_name, "Num actions:", env.action_space.n, "Tree depth:", args.tree_depth)
    ppo_params = dict(learning_rate=get_linear_fn(args.learning_rate, 0, 1), n_epochs=3, gamma=0.99, n_steps=128,
                      batch_size=32, ent_coef=0.01, vf_coef=1.0, gae_lambda=0.95, clip_range=get_linear_fn(0.1, 0, 1))
    if depth0_train:
        model = PPO(ActorCriticCnnPolicyDepth0, env, verbose=1, seed=args.seed, **ppo_params)
    else:
        max_width = int(args.max_width / env.action_space.n) if args.max_width != -1 else -1     # main.py:63
        policy_kwargs = dict(step_env=env, gamma=args.gamma, tree_depth=args.tree_depth,
                             buffer_size=max(args.hash_buffer_size, 128), learn_alpha=bool(args.learn_alpha),
                             learn_beta=bool(args.learn_beta), max_width=max_width, use_leaves_v=bool(args.use_leaves_v),
                             is_cumulative_mode=bool(args.is_cumulative_mode), regularization=args.regularization)
        model = PPO(ActorCriticCnnTSPolicy, env, policy_kwargs=policy_kwargs, verbose=1, seed=args.seed, **ppo_params)
    if args.run_type == "train":
        import time
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        model.learn(args.total_timesteps, callback=TrainingCallback(path=args.log_jsonl))      # callbacks.py:11-57
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print(f"learn: {model.num_timesteps} env steps (rollouts + PPO updates) in {dt:.2f} s = {model.num_timesteps / dt:.0f} steps/s")
        if args.model_filename:
            model.policy.save(args.model_filename)
            print("Saved", args.model_filename)
    else:                                                                                         # main.py:87-99
        if args.model_filename is None:
            raise ValueError("Model filename missing. Please specify using model_filename argument.")
        if args.tree_depth == 0:
            model.policy = ActorCriticCnnPolicyDepth0.load(args.model_filename)
        else:
            model.policy = ActorCriticCnnTSPolicy.load(args.model_filename, lr_schedule=ppo_params["learning_rate"], env=env)
        scores, lengths = evaluate_policy(model, env, n_eval_episodes=args.n_eval_episodes, return_episode_rewards=True,
                                          deterministic=False)
        print("Scores of all episodes: ", scores)
        print("Lengths of all episodes: ", lengths)
        print("Average episode score:", np.mean(scores), "Average episode length: ", np.mean(lengths))


if __name__ == "__main__":
    main()
