"""SAC actor and twin critics with the reference's module names, state-dict keys, constructor
JSON and file names (rfarl/rfarl/policy/SAC_model.py:19-355).

Actor: features -> 128 -> 128 -> (mu, log_std clamped to [-20, 2]); z ~ N(mu, sigma) by
reparameterisation, action = (2/pi) atan(z), log-probability with the reference's squashing
correction log(1 - a^2 + 1e-6) summed over the action dimensions (SAC:175-199).
Critics: the DDPG critic network, files critic_{1,2}_params.pth (SAC:333-355)."""
import copy
import json
import os

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from .AC_IQN_model import _ObservationTower, encoder
from .DDPG_model import Critic as _QCritic


class Actor(_ObservationTower):
    def __init__(self, self_dimension, object_dimension, max_object_num, self_feature_dimension,
                 object_feature_dimension, concat_feature_dimension, hidden_dimension,
                 value_ranges_of_action, device='cpu', seed=0, min_log_std=-20, max_log_std=2):
        super().__init__()
        self._init_dims(self_dimension, object_dimension, max_object_num, self_feature_dimension,
                        object_feature_dimension, concat_feature_dimension, hidden_dimension, device, seed)
        self.value_ranges_of_action = copy.deepcopy(value_ranges_of_action)
        self.action_dimension = len(self.value_ranges_of_action)
        self.atan_scale = 2.0 / np.pi
        self.min_val = 1e-6
        self.min_log_std = min_log_std
        self.max_log_std = max_log_std
        self.self_encoder = encoder(self_dimension, self_feature_dimension)
        self.object_encoder = encoder(object_dimension, object_feature_dimension)
        self.hidden_layer = nn.Linear(concat_feature_dimension, hidden_dimension)
        self.hidden_layer_2 = nn.Linear(hidden_dimension, hidden_dimension)
        self.output_layer_mu = nn.Linear(hidden_dimension, self.action_dimension)
        self.output_layer_log_std = nn.Linear(hidden_dimension, self.action_dimension)

    def observation_processor(self, x):
        return self._features(x)

    def forward(self, x, noise=None):
        """-> (actions (B, 2) in (-1, 1), log_prob (B, 1)).  ``noise`` injects the standard-normal
        draw of the reparameterisation (parity tests); default torch.randn_like."""
        h = self._trunk(x, self.hidden_layer, self.hidden_layer_2)
        mu = self.output_layer_mu(h)
        log_std = torch.clamp(self.output_layer_log_std(h), self.min_log_std, self.max_log_std)
        return self.compute_action(mu, log_std, noise)

    @torch.no_grad()
    def act(self, x, noise=None):
        """Collection: the sampled, squashed action of ``forward`` only (same draw, same arithmetic) — both heads as ONE
        128 -> 4 product and none of the log-probability terms, which acting never reads (AGENT: act of the SAC agent
        uses actions only)."""
        h = self._trunk(x, self.hidden_layer, self.hidden_layer_2)
        w = torch.cat((self.output_layer_mu.weight, self.output_layer_log_std.weight), dim=0)
        b = torch.cat((self.output_layer_mu.bias, self.output_layer_log_std.bias), dim=0)
        y = F.linear(h, w, b)
        d = self.action_dimension
        mu, log_std = y[:, :d], torch.clamp(y[:, d:], self.min_log_std, self.max_log_std)
        eps = torch.randn_like(mu) if noise is None else noise
        return self.atan_scale * torch.atan(mu + torch.exp(log_std) * eps)

    def compute_action(self, mu, log_std, noise=None):
        sigma = torch.exp(log_std)
        eps = torch.randn_like(mu) if noise is None else noise
        z = mu + sigma * eps  # Normal(mu, sigma).rsample(), SAC:191-192
        actions = self.atan_scale * torch.atan(z)
        # Normal.log_prob(z) = -(z - mu)^2 / (2 sigma^2) - log sigma - log sqrt(2 pi)
        log_prob = -0.5 * eps.pow(2) - log_std - 0.5 * float(np.log(2.0 * np.pi))
        log_prob = log_prob - torch.log(1 - actions.pow(2) + self.min_val)  # SAC:196
        return actions, log_prob.sum(dim=-1, keepdim=True)

    def get_constructor_parameters(self):
        return dict(self_dimension=self.self_dimension, object_dimension=self.object_dimension,
                    max_object_num=self.max_object_num,
                    self_feature_dimension=self.self_feature_dimension,
                    object_feature_dimension=self.object_feature_dimension,
                    concat_feature_dimension=self.concat_feature_dimension,
                    hidden_dimension=self.hidden_dimension,
                    value_ranges_of_action=self.value_ranges_of_action, seed=self.seed_id)

    def save(self, directory):
        self._dump(directory, "actor")

    @classmethod
    def load(cls, directory, device="cpu"):
        return cls._restore(directory, "actor", device)


class Critic(_QCritic):
    """Q(s, a) as in DDPG; two of them, saved as critic_<id>_params.pth (SAC:333-355)."""

    def save(self, id, directory):
        torch.save(self.state_dict(), os.path.join(directory, f"critic_{id}_params.pth"))
        with open(os.path.join(directory, f"critic_{id}_constructor_params.json"), mode="w") as f:
            json.dump(self.get_constructor_parameters(), f)

    @classmethod
    def load(cls, id, directory, device="cpu"):
        weights = torch.load(os.path.join(directory, f"critic_{id}_params.pth"), map_location=device)
        with open(os.path.join(directory, f"critic_{id}_constructor_params.json"), mode="r") as f:
            kwargs = json.load(f)
        kwargs["device"] = device
        model = cls(**kwargs)
        model.load_state_dict(weights)
        model.to(device)
        return model


class SAC_Policy():
    """Actor (seed), critic_1 (seed + 1), critic_2 (seed + 2), SAC:19-66."""

    def __init__(self, self_dimension, object_dimension, max_object_num, self_feature_dimension,
                 object_feature_dimension, concat_feature_dimension, hidden_dimension,
                 value_ranges_of_action, device='cpu', seed=0, min_log_std=-20, max_log_std=2):
        dims = (self_dimension, object_dimension, max_object_num, self_feature_dimension,
                object_feature_dimension, concat_feature_dimension, hidden_dimension)
        self.actor = Actor(*dims, value_ranges_of_action, device, seed, min_log_std, max_log_std).to(device)
        self.critic_1 = Critic(*dims, len(value_ranges_of_action), device, seed + 1).to(device)
        self.critic_2 = Critic(*dims, len(value_ranges_of_action), device, seed + 2).to(device)

    def save(self, directory):
        self.actor.save(directory)
        self.critic_1.save(1, directory)
        self.critic_2.save(2, directory)

    @classmethod
    def load(cls, directory, device="cpu"):
        actor = Actor.load(directory, device)
        a = actor
        policy = cls(a.self_dimension, a.object_dimension, a.max_object_num, a.self_feature_dimension,
                     a.object_feature_dimension, a.concat_feature_dimension, a.hidden_dimension,
                     a.value_ranges_of_action, device=device)
        policy.actor = actor
        policy.critic_1 = Critic.load(1, directory, device)
        policy.critic_2 = Critic.load(2, directory, device)
        return policy
