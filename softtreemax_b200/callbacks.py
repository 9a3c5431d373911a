"""Training callback — the metrics of the reference's callbacks.py:11-57 (WandbTrainingCallback) without
wandb: episodic reward / length, alpha, beta, the variance of the policy-head gradients over the last
updates (the paper's gradient-variance figure), trimmed gradient history.  Records go to `self.records`
and, if a path is given, to a JSON-lines file; pass them to wandb.log yourself when a server is reachable.
"""
import json

import numpy as np
import torch


class TrainingCallback:
    def __init__(self, path=None, verbose=0):
        self.total_rewards, self.episode_length, self.prev_life = 0.0, 0, 0
        self.warm_start = 0
        self.records, self.path, self.verbose = [], path, verbose

    def on_step(self, model, infos, dones):
        """Once per environment step (callbacks.py:19-57 `_on_step`)."""
        if isinstance(infos, dict):
            infos = [infos]
        self.total_rewards += float(np.mean(infos[0]["orig_reward"])) if "orig_reward" in infos[0] else 0.0
        self.episode_length += 1
        info = infos[0]
        if "ale.lives" in info:                                   # an episode ends when the life counter jumps back up (:28-33)
            lives = info["ale.lives"]
            lives = int(lives.reshape(-1)[0]) if hasattr(lives, "reshape") else int(lives[0] if isinstance(lives, list) else lives)
            done = lives > self.prev_life
            self.prev_life = lives
        else:
            done = bool(np.asarray(dones).reshape(-1)[0])
        if not done:
            return True
        pol = model.policy
        rec = {"step": model.num_timesteps + self.warm_start, "num_steps": model.num_timesteps,
               "train/episodic_reward": self.total_rewards, "train/episodic_length": self.episode_length}
        for name in ("alpha", "beta"):
            if hasattr(pol, name):
                v = getattr(pol, name)
                rec["train/" + name] = float(v.detach()) if hasattr(v, "detach") else float(v)
        grad_var = 0.0
        for name, hist in pol.gradients_history.items():          # :46-50
            if len(hist) > 1:
                grad_var += torch.mean(torch.var(torch.stack(hist), dim=0)).item()
            pol.gradients_history[name] = hist[-100:]
        rec["train/policy_weights_grad_var"] = grad_var
        if model.log:
            rec.update({"train/" + k: v for k, v in model.log[-1].items() if k in ("loss", "pg_loss", "value_loss", "entropy")})
        self.records.append(rec)
        if self.path:
            with open(self.path, "a") as f:
                f.write(json.dumps(rec) + "\n")
        if self.verbose:
            print(rec, flush=True)
        self.total_rewards, self.episode_length = 0.0, 0
        return True

    def __call__(self, model, rec):                                # per-iteration hook of ppo.PPO.learn: nothing to add
        return True
