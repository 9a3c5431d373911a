"""SoftTreeMaxPolicy — host-side mirror of policies/actor_critic_ts.py:15-273
(ActorCriticCnnTSPolicy) and of its parent policies/actor_critic_depth0.py:9-35.

Same constructor keywords (main.py:64-67), same methods and return tuples: forward, evaluate_actions,
hash_obs, compute_value, compute_value_with_root, predict, save, load, add_gradients_history.
Stable-Baselines3 is not installable here, so the SB3 base class is restated as a plain nn.Module
with SB3's parameter names (features_extractor.cnn.{0,2,4}, features_extractor.linear.0,
action_net, value_net) — a checkpoint's state_dict loads either way.

Where the work happens:
  forward (rollout, no grad): CuleBFS.bfs -> stm_bfs (K1-K3); leaves + root through stm_cnn_forward
      (K4, tcgen05); SoftTreeMax logits through stm_logits_fwd (K5); Categorical in torch.
  evaluate_actions (PPO update, needs grad): cached u8 leaves through the SAME network in torch
      autograd (K4's backward is SURVEY §8f row 2, not built yet); logits through K5's autograd
      function (stm_logits_fwd/bwd), batched over roots — no Python loop over the minibatch
      (the reference loops, actor_critic_ts.py:155-178).
"""
import math

import torch as th
import torch.nn as nn
import torch.nn.functional as F

from . import ops
from .cule_bfs import CuleBFS


class NatureCNN(nn.Module):
    """SB3 1.4.0 NatureCNN (restated; SB3 is a third-party dependency of the reference)."""

    def __init__(self, n_input_channels=4, features_dim=512):
        super().__init__()
        self.cnn = nn.Sequential(nn.Conv2d(n_input_channels, 32, 8, 4), nn.ReLU(), nn.Conv2d(32, 64, 4, 2), nn.ReLU(),
                                 nn.Conv2d(64, 64, 3, 1), nn.ReLU(), nn.Flatten())
        self.linear = nn.Sequential(nn.Linear(3136, features_dim), nn.ReLU())

    def forward(self, obs):
        return self.linear(self.cnn(obs))


class ActorCriticCnnPolicyDepth0(nn.Module):
    """policies/actor_critic_depth0.py — the depth-0 CNN policy plus the gradient history hook."""

    def __init__(self, observation_space, action_space, lr_schedule, optimizer_class=th.optim.Adam,
                 optimizer_kwargs=None, device="cuda:0", **kwargs):
        super().__init__()
        self.observation_space, self.action_space = observation_space, action_space
        self.features_extractor = NatureCNN(4, 512)
        self.action_net = nn.Linear(512, action_space.n)
        self.value_net = nn.Linear(512, 1)
        for mod, gain in [(self.features_extractor, math.sqrt(2)), (self.action_net, 0.01), (self.value_net, 1.0)]:
            for m in mod.modules():
                if isinstance(m, (nn.Linear, nn.Conv2d)):
                    nn.init.orthogonal_(m.weight, gain=gain)
                    nn.init.zeros_(m.bias)
        self.optimizer_class = optimizer_class
        self.optimizer_kwargs = {"eps": 1e-5} if optimizer_kwargs is None else optimizer_kwargs
        self.lr_schedule = lr_schedule
        self.device = th.device(device)
        self.to(self.device)
        self.optimizer = self.optimizer_class(self.parameters(), lr=lr_schedule(1), **self.optimizer_kwargs)
        self.gradients_history = {}

    def extract_features(self, obs):
        return self.features_extractor(obs.float() / 255.0)              # SB3 preprocess_obs for image spaces

    def add_gradients_history(self):                                      # actor_critic_depth0.py:27-35
        for name, param in self.named_parameters():
            if not name.startswith("action_net"):
                continue
            if param.grad is None:
                break
            self.gradients_history.setdefault(name, []).append(param.grad.data.detach().clone())

    # what SB3's ActorCriticCnnPolicy gives the depth-0 policy (forward / evaluate_actions / predict), batched
    def _logits_values(self, obs):
        feats = self.extract_features(th.as_tensor(obs).to(self.device))
        return self.action_net(feats), self.value_net(feats)

    def forward(self, obs, deterministic=False):
        logits, values = self._logits_values(obs)
        dist = th.distributions.Categorical(logits=logits)
        actions = logits.argmax(dim=1) if deterministic else dist.sample()
        return actions, values, dist.log_prob(actions)

    def evaluate_actions(self, obs, actions):
        self.add_gradients_history()
        logits, values = self._logits_values(obs)
        dist = th.distributions.Categorical(logits=logits)
        return values, dist.log_prob(th.as_tensor(actions).to(self.device).reshape(-1)), dist.entropy()

    def predict(self, obs, state=None, mask=None, deterministic=False):
        with th.no_grad():
            return self.forward(th.as_tensor(obs), deterministic)[0].cpu().numpy(), None

    def save(self, path):                                                 # SB3 BasePolicy.save: constructor data + weights
        th.save({"state_dict": self.state_dict(),
                 "data": dict(observation_space=self.observation_space, action_space=self.action_space)}, path)

    @classmethod
    def load(cls, path, device="cuda:0", **_):                            # main.py:91
        saved = th.load(path, map_location=device, weights_only=False)
        model = cls(**saved["data"], lr_schedule=lambda _: 0.0, device=device)
        model.load_state_dict(saved["state_dict"])
        return model

    def kernel_param_list(self):
        """The twelve parameters themselves (not detached), in Engine.PARAM_KEYS order: what ops.cnn_heads differentiates."""
        fe = self.features_extractor
        mods = (fe.cnn[0], fe.cnn[2], fe.cnn[4], fe.linear[0], self.action_net, self.value_net)
        return [t for m in mods for t in (m.weight, m.bias)]

    def kernel_params(self):
        sd = self.state_dict()
        m = {"cnn.0": "features_extractor.cnn.0", "cnn.2": "features_extractor.cnn.2", "cnn.4": "features_extractor.cnn.4",
             "linear.0": "features_extractor.linear.0", "action_net": "action_net", "value_net": "value_net"}
        return {f"{k}.{p}": sd[f"{v}.{p}"] for k, v in m.items() for p in ("weight", "bias")}


class SoftTreeMaxPolicy(ActorCriticCnnPolicyDepth0):
    def __init__(self, observation_space, action_space, lr_schedule, tree_depth, gamma, step_env, buffer_size,
                 learn_alpha, learn_beta, max_width, use_leaves_v, is_cumulative_mode, regularization, **kwargs):
        super().__init__(observation_space, action_space, lr_schedule, **kwargs)
        self.cule_bfs = CuleBFS(step_env, tree_depth, gamma, self.compute_value_kernel, max_width)  # :19
        self.eng = self.cule_bfs.eng
        self.time_step = 0
        self.obs2leaves_dict, self.timestep2obs_dict, self.obs2timestep_dict = {}, {}, {}            # :21-23
        self.buffer_size = buffer_size
        self.learn_alpha, self.learn_beta = learn_alpha, learn_beta
        self.tree_depth, self.max_width = tree_depth, max_width
        self.is_cumulative_mode, self.regularization = is_cumulative_mode, regularization
        self.alpha = th.tensor(0.5 if learn_alpha else 1.0, device=self.device)                      # :31
        self.beta = th.tensor(1.0, device=self.device)                                               # :32
        if learn_alpha:
            self.alpha = nn.Parameter(self.alpha)
        if learn_beta:
            self.beta = nn.Parameter(self.beta)
        if learn_alpha or learn_beta:                                                                # :34-39
            self.optimizer = self.optimizer_class(self.parameters(), lr=lr_schedule(1), **self.optimizer_kwargs)
        self.use_leaves_v = use_leaves_v
        self.update_backend = kwargs.get("update_backend", "kernel")          # "kernel" (K4 forward_train + backward) | "torch"

    # ------------------------------------------------------------------------------------------
    def _sync_kernel_weights(self):
        """K4 holds a bf16 re-tiled copy of the parameters: refresh it when they changed."""
        # The tiled copy lives in the Engine (one per cartridge and GPU, shared by every policy on it): the
        # Engine remembers whose parameters, and which version of them, it holds.
        ver = (id(self), tuple(p._version for p in self.parameters()))
        if getattr(self.eng, "_cnn_owner", None) != ver:
            self.eng.cnn_set_weights(self.kernel_params())
            self.eng._cnn_owner = ver

    def _host_scalars(self):
        """beta and alpha as Python floats, re-read from the device only when the optimizer has changed them."""
        ver = (getattr(self.beta, "_version", 0), getattr(self.alpha, "_version", 0))
        if getattr(self, "_scalars_ver", None) != ver:
            self._scalars, self._scalars_ver = (float(self.beta), float(self.alpha)), ver
        return self._scalars

    def hash_obs(self, obs):                                                                         # :196-197
        return (obs[:, -2:, :, :].int()).view(obs.shape[0], -1).sum(dim=1)

    def compute_value(self, leaves_obs, root_obs=None):                                              # :210-218
        latent = self.extract_features(leaves_obs)
        if root_obs is None:
            return self.action_net(latent), None
        return latent, self.value_net(latent)

    @th.no_grad()
    def compute_value_kernel(self, leaves_obs, root_obs=None):
        """compute_value(leaves_obs) for the tree's width pruning (cule_bfs.py:176), on K4."""
        self._sync_kernel_weights()
        head, _ = self.eng.cnn_forward(leaves_obs.contiguous())
        return head, None

    def compute_value_with_root(self, leaves_obs, root_obs=None):                                    # :199-208
        if root_obs is None:
            return self.action_net(self.extract_features(leaves_obs)), None
        cat = self.extract_features(th.cat((root_obs.to(leaves_obs.dtype), leaves_obs)))
        return cat[1:], self.value_net(cat[:1])

    def _distribution(self, logits):
        return th.distributions.Categorical(logits=logits)

    @th.no_grad()
    def forward(self, obs, deterministic=False):                                                     # :42-104
        obs = obs.to(self.device)
        key = int(self.hash_obs(obs)[0].item())
        if key in self.obs2leaves_dict:
            leaves, rewards, first_action = self.obs2leaves_dict[key]
            del self.timestep2obs_dict[self.obs2timestep_dict[key]]
        else:
            leaves, rewards, first_action = self.cule_bfs.bfs(obs, self.cule_bfs.max_depth)
            self.obs2leaves_dict[key] = leaves, rewards, first_action
        self.obs2timestep_dict[key] = self.time_step
        self.timestep2obs_dict[self.time_step] = key
        val_coef = self.cule_bfs.gamma ** self.cule_bfs.max_depth                                     # :60
        self._sync_kernel_weights()
        A = self.action_space.n
        root_u8 = obs.round().clamp(0, 255).to(th.uint8).contiguous() if obs.dtype != th.uint8 else obs.contiguous()
        if leaves.dtype == th.uint8:                                                                 # K4: root + leaves in one pass
            head_all, val_all = self.eng.cnn_forward(th.cat([root_u8[:1], leaves]))
            head_root, val_root, head_leaf, val_leaf = head_all[:1], val_all[:1], head_all[1:], val_all[1:]
        else:                                                                                        # float_obs trees
            head_leaf, val_leaf = self.eng.cnn_forward(leaves.round().clamp(0, 255).to(th.uint8).contiguous())
            head_root, val_root = self.eng.cnn_forward(root_u8)
        beta, alpha = self._host_scalars()
        logits, _, vmax = ops.softtreemax_logits_full(                                               # K5
            self.eng, head_leaf.reshape(1, -1, A), rewards.reshape(1, -1), val_coef, beta, alpha,
            self.regularization, self.is_cumulative_mode, depth0=head_root if self.learn_alpha else None,
            values=val_leaf.reshape(1, -1) if self.use_leaves_v else None,
            group=first_action.reshape(1, -1) if first_action is not None else None)                 # :70-79
        value_root = vmax.reshape(()) if self.use_leaves_v else val_root.reshape(1, 1)               # :62-65
        dist = self._distribution(logits)
        actions = logits.argmax(dim=1) if deterministic else dist.sample()
        log_prob = dist.log_prob(actions)
        old = self.time_step - self.buffer_size                                                      # :99-103
        if old in self.timestep2obs_dict:
            del self.obs2leaves_dict[self.timestep2obs_dict[old]]
            del self.obs2timestep_dict[self.timestep2obs_dict[old]]
            del self.timestep2obs_dict[old]
        self.time_step += 1
        return actions, value_root, log_prob

    @th.no_grad()
    def forward_roots(self, obs, root_blobs, deterministic=False):
        """forward() for R roots at once (the reference evaluates one root per call: main.py:47, n_envs = 1): obs (R, 4, 84, 84)
        observation stacks, root_blobs (R, 256) emulator states of the same roots.  One stm_bfs over all roots, one K4 pass
        over the R + R*L images, one K5 launch -> (actions (R,), values (R,), log_prob (R,)).  No leaf cache is kept (a
        vectorised rollout that wants evaluate_actions later keeps the returned trees itself); full-width trees only."""
        if self.max_width != -1:
            raise NotImplementedError("forward_roots: full-width trees (max_width = -1)")
        obs = th.as_tensor(obs).to(self.device)
        root_u8 = obs.round().clamp(0, 255).to(th.uint8).contiguous() if obs.dtype != th.uint8 else obs.contiguous()
        R, A, d = root_u8.shape[0], self.action_space.n, self.cule_bfs.max_depth
        out = self.cule_bfs.bfs_batch(root_blobs.to(self.device), root_u8, d)
        val_coef = self.cule_bfs.gamma ** d
        self._sync_kernel_weights()
        head_all, val_all = self.eng.cnn_forward(th.cat([root_u8, out["stacks"]]))
        head_root, val_root, head_leaf, val_leaf = head_all[:R], val_all[:R], head_all[R:], val_all[R:]
        beta, alpha = self._host_scalars()
        logits, _, vmax = ops.softtreemax_logits_full(
            self.eng, head_leaf.reshape(R, -1, A), out["rewards"].reshape(R, -1), val_coef, beta, alpha, self.regularization,
            self.is_cumulative_mode, depth0=head_root if self.learn_alpha else None,
            values=val_leaf.reshape(R, -1) if self.use_leaves_v else None)
        values = vmax if self.use_leaves_v else val_root
        dist = self._distribution(logits)
        actions = logits.argmax(dim=1) if deterministic else dist.sample()
        return actions, values, dist.log_prob(actions), logits

    def evaluate_actions(self, obs, actions):                                                        # :106-194
        self.add_gradients_history()
        obs = obs.to(self.device)
        B, A = obs.shape[0], self.action_space.n
        keys = self.hash_obs(obs).tolist()                                                           # one D2H, not B
        leaves, rewards, groups = [], [], []
        for k in keys:
            if k not in self.obs2leaves_dict:
                # the reference prints and re-expands with the WHOLE batch as root (:129-132), a latent bug
                raise KeyError("observation not in the leaf cache: evaluate_actions must follow forward()")
            l, r, fa = self.obs2leaves_dict[k]
            leaves.append(l)
            rewards.append(r)
            groups.append(fa)
        L = leaves[0].shape[0]
        val_coef = self.cule_bfs.gamma ** self.cule_bfs.max_depth
        all_obs = th.cat([obs.round().clamp(0, 255).to(th.uint8) if obs.dtype != th.uint8 else obs] + leaves)   # :122,139
        rew = th.stack(rewards)
        if self.update_backend == "kernel":
            # K4 on the update path too: forward that keeps its activations + tcgen05 data / weight gradients (stm_cnn_backward)
            self._sync_kernel_weights()
            heads = ops.cnn_heads(self.eng, all_obs, self.kernel_param_list())                       # (B + B L, A + 1)
            values = heads[:B, A:A + 1]                                                              # :145-147
            head = heads[B:, :A].reshape(B, L, A)
            depth0 = heads[:B, :A] if self.learn_alpha else None
            leaf_values = heads[B:, A].reshape(B, L)
        else:                                                                                        # torch autograd (cuDNN), kept for A/B tests
            feats = self.extract_features(all_obs)
            values = self.value_net(feats[:B])
            head = self.action_net(feats[B:]).reshape(B, L, A)
            depth0 = self.action_net(feats[:B]) if self.learn_alpha else None
            leaf_values = self.value_net(feats[B:]).reshape(B, L) if self.use_leaves_v else None
        group = th.stack([g.reshape(-1) for g in groups]) if groups[0] is not None else None          # :163-171
        logits = ops.softtreemax_logits(self.eng, head, rew, val_coef, self.beta, self.alpha, self.regularization,
                                        self.is_cumulative_mode, depth0, group)                       # K5 fwd + bwd
        dist = self._distribution(logits)
        log_prob = dist.log_prob(actions.to(self.device).reshape(-1))
        if self.use_leaves_v:
            values = (val_coef * leaf_values + rew).max(dim=1, keepdim=True).values
        return values, log_prob, dist.entropy()

    def predict(self, obs, state=None, mask=None, deterministic=False):                              # :272-273
        # numpy out, like SB3's BasePolicy.predict (evaluate_policy feeds it to env.step)
        return self.forward(th.as_tensor(obs, dtype=th.float32, device=self.device), deterministic)[0].cpu().numpy(), None

    # ------------------------------------------------------------------------------------------
    def _get_data(self):                                                                             # :230-252
        return dict(observation_space=self.observation_space, action_space=self.action_space,
                    tree_depth=self.tree_depth, gamma=self.cule_bfs.gamma, buffer_size=self.buffer_size,
                    learn_alpha=self.learn_alpha, learn_beta=self.learn_beta,
                    is_cumulative_mode=self.is_cumulative_mode, regularization=self.regularization,
                    max_width=self.max_width, use_leaves_v=self.use_leaves_v)

    def save(self, path):                                                                            # :220-228
        th.save({"state_dict": self.state_dict(), "data": self._get_data(), "alpha": self.alpha, "beta": self.beta,
                 "time_step": self.time_step}, path)

    @classmethod
    def load(cls, path, device="cuda:0", env=None, lr_schedule=None):                                # :254-270
        saved = th.load(path, map_location=device, weights_only=False)
        model = cls(**saved["data"], lr_schedule=lr_schedule or (lambda _: 0.0), step_env=env, device=device)
        model.load_state_dict(saved["state_dict"])
        return model.to(device)


ActorCriticCnnTSPolicy = SoftTreeMaxPolicy        # the reference's class name (main.py:18)
