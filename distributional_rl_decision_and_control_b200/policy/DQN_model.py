"""DQN Q-network over the 25 discrete thrust-change actions with the reference's module names,
state-dict keys and file names (rfarl/rfarl/policy/DQN_model.py:14-112).  Plain dense layers on
the shared observation tower; trained from the device replay ring (kernel K2b)."""
import json
import os

import torch
import torch.nn as nn
import torch.nn.functional as F

from .AC_IQN_model import _ObservationTower, encoder


class DQN_Policy(_ObservationTower):
    def __init__(self, self_dimension, object_dimension, max_object_num, self_feature_dimension,
                 object_feature_dimension, concat_feature_dimension, hidden_dimension, action_size,
                 device='cpu', seed=0):
        super().__init__()
        self._init_dims(self_dimension, object_dimension, max_object_num, self_feature_dimension,
                        object_feature_dimension, concat_feature_dimension, hidden_dimension, device, seed)
        self.action_size = action_size
        self.self_encoder = encoder(self_dimension, self_feature_dimension)
        self.object_encoder = encoder(object_dimension, object_feature_dimension)
        self.hidden_layer = nn.Linear(concat_feature_dimension, hidden_dimension)
        self.hidden_layer_2 = nn.Linear(hidden_dimension, hidden_dimension)
        self.output_layer = nn.Linear(hidden_dimension, action_size)

    def forward(self, x):
        h = self._trunk(x, self.hidden_layer, self.hidden_layer_2)  # DQN:50-77
        return self.output_layer(h)

    def get_constructor_parameters(self):
        return dict(self_dimension=self.self_dimension, object_dimension=self.object_dimension,
                    max_object_num=self.max_object_num,
                    self_feature_dimension=self.self_feature_dimension,
                    object_feature_dimension=self.object_feature_dimension,
                    concat_feature_dimension=self.concat_feature_dimension,
                    hidden_dimension=self.hidden_dimension, action_size=self.action_size,
                    seed=self.seed_id)

    def save(self, directory):
        torch.save(self.state_dict(), os.path.join(directory, "network_params.pth"))
        with open(os.path.join(directory, "constructor_params.json"), mode="w") as f:
            json.dump(self.get_constructor_parameters(), f)

    @classmethod
    def load(cls, directory, device="cpu"):
        weights = torch.load(os.path.join(directory, "network_params.pth"), map_location=device)
        with open(os.path.join(directory, "constructor_params.json"), mode="r") as f:
            kwargs = json.load(f)
        kwargs["device"] = device
        model = cls(**kwargs)
        model.load_state_dict(weights)
        model.to(device)
        return model
