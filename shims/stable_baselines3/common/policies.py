"""stable_baselines3.common.policies.ActorCriticCnnPolicy as SB3 1.4.0 builds it for an image observation and a
Discrete action space — the base class of the reference's policies (policies/actor_critic_depth0.py:9,
policies/actor_critic_ts.py:15).  Restated surface:

    features_extractor (NatureCNN: conv 8x8 s4 (32) - 4x4 s2 (64) - 3x3 s1 (64) - flatten - linear 512, ReLU)
    mlp_extractor.{shared_net, policy_net, value_net, forward_actor, forward_critic}   (net_arch [] for a CNN: identities)
    action_net Linear(512, A), value_net Linear(512, 1); orthogonal init (sqrt 2 / 0.01 / 1), Adam(eps=1e-5)
    extract_features (obs / 255), forward, evaluate_actions, get_distribution, predict_values, _predict, predict
    action_dist.proba_distribution(action_logits=).{get_actions, log_prob, entropy, sample, mode}
    optimizer_class, optimizer_kwargs, optimizer, device, set_training_mode, save, load
"""
import math

import numpy as np
import torch as th
import torch.nn as nn

from .utils import get_device, obs_as_tensor


class NatureCNN(nn.Module):
    def __init__(self, observation_space, features_dim=512):
        super().__init__()
        self.features_dim = features_dim
        n_in = observation_space.shape[0]
        self.cnn = nn.Sequential(nn.Conv2d(n_in, 32, kernel_size=8, stride=4, padding=0), nn.ReLU(),
                                 nn.Conv2d(32, 64, kernel_size=4, stride=2, padding=0), nn.ReLU(),
                                 nn.Conv2d(64, 64, kernel_size=3, stride=1, padding=0), nn.ReLU(), nn.Flatten())
        with th.no_grad():
            n_flatten = self.cnn(th.zeros((1,) + tuple(observation_space.shape))).shape[1]
        self.linear = nn.Sequential(nn.Linear(n_flatten, features_dim), nn.ReLU())

    def forward(self, observations):
        return self.linear(self.cnn(observations))


class MlpExtractor(nn.Module):
    """net_arch = [] (SB3's default for ActorCriticCnnPolicy): no shared, policy or value layers."""

    def __init__(self, feature_dim):
        super().__init__()
        self.shared_net, self.policy_net, self.value_net = nn.Sequential(), nn.Sequential(), nn.Sequential()
        self.latent_dim_pi = self.latent_dim_vf = feature_dim

    def forward(self, features):
        shared = self.shared_net(features)
        return self.policy_net(shared), self.value_net(shared)

    def forward_actor(self, features):
        return self.policy_net(self.shared_net(features))

    def forward_critic(self, features):
        return self.value_net(self.shared_net(features))


class CategoricalDistribution:
    def __init__(self, action_dim):
        self.action_dim = action_dim
        self.distribution = None

    def proba_distribution_net(self, latent_dim):
        return nn.Linear(latent_dim, self.action_dim)

    def proba_distribution(self, action_logits):
        self.distribution = th.distributions.Categorical(logits=action_logits)
        return self

    def log_prob(self, actions):
        return self.distribution.log_prob(actions)

    def entropy(self):
        return self.distribution.entropy()

    def sample(self):
        return self.distribution.sample()

    def mode(self):
        return th.argmax(self.distribution.probs, dim=1)

    def get_actions(self, deterministic=False):
        return self.mode() if deterministic else self.sample()


class ActorCriticCnnPolicy(nn.Module):
    def __init__(self, observation_space, action_space, lr_schedule, net_arch=None, activation_fn=nn.Tanh, ortho_init=True,
                 use_sde=False, features_extractor_class=NatureCNN, features_extractor_kwargs=None, normalize_images=True,
                 optimizer_class=th.optim.Adam, optimizer_kwargs=None, **_unused):
        super().__init__()
        if optimizer_kwargs is None:
            optimizer_kwargs = {}
            if optimizer_class == th.optim.Adam:
                optimizer_kwargs["eps"] = 1e-5
        self.observation_space, self.action_space = observation_space, action_space
        self.normalize_images, self.ortho_init = normalize_images, ortho_init
        self.optimizer_class, self.optimizer_kwargs = optimizer_class, optimizer_kwargs
        self.features_extractor = features_extractor_class(observation_space, **(features_extractor_kwargs or {}))
        self.features_dim = self.features_extractor.features_dim
        self.mlp_extractor = MlpExtractor(self.features_dim)
        self.action_dist = CategoricalDistribution(action_space.n)
        self.action_net = self.action_dist.proba_distribution_net(self.mlp_extractor.latent_dim_pi)
        self.value_net = nn.Linear(self.mlp_extractor.latent_dim_vf, 1)
        if ortho_init:
            for module, gain in ((self.features_extractor, math.sqrt(2)), (self.mlp_extractor, math.sqrt(2)),
                                 (self.action_net, 0.01), (self.value_net, 1.0)):
                for m in module.modules():
                    if isinstance(m, (nn.Linear, nn.Conv2d)):
                        nn.init.orthogonal_(m.weight, gain=gain)
                        if m.bias is not None:
                            m.bias.data.fill_(0.0)
        self.optimizer = self.optimizer_class(self.parameters(), lr=lr_schedule(1), **self.optimizer_kwargs)

    @property
    def device(self):
        for p in self.parameters():
            return p.device
        return get_device("cpu")

    def set_training_mode(self, mode):
        self.train(mode)

    def extract_features(self, obs):
        """preprocess_obs for an image Box: float, / 255; then the features extractor."""
        x = obs.float()
        if self.normalize_images:
            x = x / 255.0
        return self.features_extractor(x)

    def _get_action_dist_from_latent(self, latent_pi):
        return self.action_dist.proba_distribution(action_logits=self.action_net(latent_pi))

    def forward(self, obs, deterministic=False):
        features = self.extract_features(obs)
        latent_pi, latent_vf = self.mlp_extractor(features)
        values = self.value_net(latent_vf)
        distribution = self._get_action_dist_from_latent(latent_pi)
        actions = distribution.get_actions(deterministic=deterministic)
        return actions, values, distribution.log_prob(actions)

    def evaluate_actions(self, obs, actions):
        features = self.extract_features(obs)
        latent_pi, latent_vf = self.mlp_extractor(features)
        distribution = self._get_action_dist_from_latent(latent_pi)
        return self.value_net(latent_vf), distribution.log_prob(actions), distribution.entropy()

    def get_distribution(self, obs):
        return self._get_action_dist_from_latent(self.mlp_extractor.forward_actor(self.extract_features(obs)))

    def predict_values(self, obs):
        return self.value_net(self.mlp_extractor.forward_critic(self.extract_features(obs)))

    def _predict(self, observation, deterministic=False):
        return self.get_distribution(observation).get_actions(deterministic=deterministic)

    def predict(self, observation, state=None, mask=None, deterministic=False):
        self.set_training_mode(False)
        obs = np.asarray(observation)
        vectorized = obs.ndim == len(self.observation_space.shape) + 1
        if not vectorized:
            obs = obs[None]
        with th.no_grad():
            actions = self._predict(obs_as_tensor(obs, self.device), deterministic=deterministic)
        actions = actions.cpu().numpy()
        return (actions if vectorized else actions[0]), state

    # --- persistence (SB3 BasePolicy.save / load: constructor data + state_dict) -----------------------------------------
    def _get_constructor_parameters(self):
        return dict(observation_space=self.observation_space, action_space=self.action_space,
                    lr_schedule=lambda _: 0.0, ortho_init=self.ortho_init, normalize_images=self.normalize_images,
                    optimizer_class=self.optimizer_class, optimizer_kwargs=self.optimizer_kwargs)

    def save(self, path):
        data = self._get_constructor_parameters()
        data.pop("lr_schedule", None)
        th.save({"state_dict": self.state_dict(), "data": data}, path)

    @classmethod
    def load(cls, path, device="auto"):
        device = get_device(device)
        saved = th.load(path, map_location=device, weights_only=False)
        model = cls(lr_schedule=lambda _: 0.0, **saved["data"])
        model.load_state_dict(saved["state_dict"])
        model.to(device)
        return model
