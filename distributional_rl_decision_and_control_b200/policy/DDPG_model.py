"""DDPG actor / critic with the reference's module names, state-dict keys, constructor JSON and
file names (rfarl/rfarl/policy/DDPG_model.py:16-317).  The actor is the AC-IQN actor's network
(same layers and squashing, DDPG:81-161); the critic is a scalar Q(s, a): features -> 128,
gated by the action encoding 2 -> 128, -> 128 -> 1 (DDPG:200-282)."""
import torch.nn as nn
import torch.nn.functional as F

from .AC_IQN_model import Actor as _SquashedActor
from .AC_IQN_model import _ObservationTower, encoder


class Actor(_SquashedActor):
    """Deterministic policy into (-1, 1)^2 by (2/pi) atan; files actor_network_params.pth /
    actor_constructor_params.json (DDPG:163-198)."""


class Critic(_ObservationTower):
    def __init__(self, self_dimension, object_dimension, max_object_num, self_feature_dimension,
                 object_feature_dimension, concat_feature_dimension, hidden_dimension,
                 action_dimension, device='cpu', seed=0):
        super().__init__()
        self._init_dims(self_dimension, object_dimension, max_object_num, self_feature_dimension,
                        object_feature_dimension, concat_feature_dimension, hidden_dimension, device, seed)
        self.action_dimension = action_dimension
        self.self_encoder = encoder(self_dimension, self_feature_dimension)
        self.object_encoder = encoder(object_dimension, object_feature_dimension)
        self.action_encoder = encoder(action_dimension, hidden_dimension)
        self.hidden_layer = nn.Linear(concat_feature_dimension, hidden_dimension)
        self.hidden_layer_2 = nn.Linear(hidden_dimension, hidden_dimension)
        self.output_layer = nn.Linear(hidden_dimension, 1)

    def observation_processor(self, x):
        return self._features(x)

    def forward(self, x, actions):
        h = F.relu(self.hidden_layer(self._features(x)))
        h = self.action_encoder(actions) * h  # DDPG:274-276
        h = F.relu(self.hidden_layer_2(h))
        return self.output_layer(h)

    def get_constructor_parameters(self):
        return dict(self_dimension=self.self_dimension, object_dimension=self.object_dimension,
                    max_object_num=self.max_object_num,
                    self_feature_dimension=self.self_feature_dimension,
                    object_feature_dimension=self.object_feature_dimension,
                    concat_feature_dimension=self.concat_feature_dimension,
                    hidden_dimension=self.hidden_dimension,
                    action_dimension=self.action_dimension, seed=self.seed_id)

    def save(self, directory):
        self._dump(directory, "critic")

    @classmethod
    def load(cls, directory, device="cpu"):
        return cls._restore(directory, "critic", device)


class DDPG_Policy():
    """Actor + critic pair; the critic is seeded with seed + 1 (DDPG:16-78)."""

    def __init__(self, self_dimension, object_dimension, max_object_num, self_feature_dimension,
                 object_feature_dimension, concat_feature_dimension, hidden_dimension,
                 value_ranges_of_action, device='cpu', seed=0):
        dims = (self_dimension, object_dimension, max_object_num, self_feature_dimension,
                object_feature_dimension, concat_feature_dimension, hidden_dimension)
        self.actor = Actor(*dims, value_ranges_of_action, device, seed).to(device)
        self.critic = Critic(*dims, len(value_ranges_of_action), device, seed + 1).to(device)

    def save(self, directory):
        self.actor.save(directory)
        self.critic.save(directory)

    @classmethod
    def load(cls, directory, device="cpu"):
        actor = Actor.load(directory, device)
        critic = Critic.load(directory, device)
        a = actor
        policy = cls(a.self_dimension, a.object_dimension, a.max_object_num, a.self_feature_dimension,
                     a.object_feature_dimension, a.concat_feature_dimension, a.hidden_dimension,
                     a.value_ranges_of_action, device=device)
        policy.actor, policy.critic = actor, critic
        return policy
