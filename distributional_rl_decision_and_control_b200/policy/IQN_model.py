"""IQN Q-network over the 25 discrete thrust-change actions, with the reference's names,
state-dict keys and file names (rfarl/rfarl/policy/IQN_model.py:14-146).  Dense layers in
PyTorch; tau sampling + cosine table is kernel K3."""
import json
import os

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from .AC_IQN_model import _ObservationTower, encoder
from .iqn_ops import TauSampler, linear_scale_relu


class IQN_Policy(_ObservationTower):
    def __init__(self, self_dimension, object_dimension, max_object_num, self_feature_dimension,
                 object_feature_dimension, concat_feature_dimension, hidden_dimension, action_size,
                 device='cpu', seed=0):
        super().__init__()
        self._init_dims(self_dimension, object_dimension, max_object_num, self_feature_dimension,
                        object_feature_dimension, concat_feature_dimension, hidden_dimension, device, seed)
        self.action_size = action_size
        self.self_encoder = encoder(self_dimension, self_feature_dimension)
        self.object_encoder = encoder(object_dimension, object_feature_dimension)
        self.K = 32  # IQN:44
        self.n = 64  # IQN:45
        self.pis = torch.FloatTensor([np.pi * i for i in range(self.n)]).view(1, 1, self.n).to(device)
        self.cos_embedding = nn.Linear(self.n, concat_feature_dimension)
        self.hidden_layer = nn.Linear(concat_feature_dimension, hidden_dimension)
        self.hidden_layer_2 = nn.Linear(hidden_dimension, hidden_dimension)
        self.output_layer = nn.Linear(hidden_dimension, action_size)
        self.tau_sampler = TauSampler(seed, self.n)
        self._ones = {}  # (B, device) -> ones (B, hidden): the unit scale of the two plain ReLU layers

    def calc_cos(self, batch_size, num_tau=8, cvar=1.0, taus=None):
        dev = self.cos_embedding.weight.device
        cos, taus = self.tau_sampler.draw(batch_size, num_tau, cvar, dev, taus)
        assert cos.shape == (batch_size, num_tau, self.n), "cos shape is incorrect"
        return cos, taus

    def forward(self, x, num_tau=8, cvar=1.0, taus=None):
        feats = self._features(x)
        B = feats.shape[0]
        cos, taus = self.calc_cos(B, num_tau, cvar, taus)
        emb = F.relu(self.cos_embedding(cos.reshape(B * num_tau, self.n)))
        h = (feats.unsqueeze(1) * emb.reshape(B, num_tau, -1)).reshape(B * num_tau, -1)
        h = F.relu(self.hidden_layer(h))
        h = F.relu(self.hidden_layer_2(h))
        return self.output_layer(h).reshape(B, num_tau, self.action_size), taus

    def forward_fused(self, x, num_tau=8, cvar=1.0, taus=None):
        """Same function as ``forward`` (IQN:74-86) with bias + ReLU (+ the feature product) of the
        three hidden layers in K5: one launch per layer forward, one backward."""
        feats = self._features(x)
        B = feats.shape[0]
        cos, taus = self.calc_cos(B, num_tau, cvar, taus)
        ones = self._ones.get((B, feats.device))
        if ones is None:
            ones = self._ones[(B, feats.device)] = torch.ones((B, self.hidden_dimension), device=feats.device)
        h = linear_scale_relu(cos.reshape(B * num_tau, self.n), self.cos_embedding.weight,
                              self.cos_embedding.bias, feats, num_tau)
        h = linear_scale_relu(h, self.hidden_layer.weight, self.hidden_layer.bias, ones, num_tau)
        h = linear_scale_relu(h, self.hidden_layer_2.weight, self.hidden_layer_2.bias, ones, num_tau)
        return self.output_layer(h).reshape(B, num_tau, self.action_size), taus

    @torch.no_grad()
    def action_values_fused(self, x, cvar=1.0):
        """(M, A) action values of IQN's act (AGENT:227-250): mean over K = 32 quantile rows of the network output.
        K6 tower, then K10 with head 1 — cosine embedding, the three hidden layers on the tcgen05 tensor cores and the
        mean over the 32 rows of the last hidden layer in one persistent launch — and the (linear) output layer applied
        to that mean.  None when the shape / dtype is not K10's."""
        from .. import _lib
        from .iqn_ops import UMMA_TRUNK
        feats = self._features(x)
        ce, h1, h2 = self.cos_embedding, self.hidden_layer, self.hidden_layer_2
        if not (UMMA_TRUNK and self.K == 32 and feats.is_cuda and feats.dtype == torch.float32
                and ce.weight.dtype == torch.float32 and tuple(ce.weight.shape) == (256, 64)
                and tuple(h1.weight.shape) == (128, 256) and tuple(h2.weight.shape) == (128, 128)):
            return None
        ts = self.tau_sampler
        ts.counter += 1
        ws = _lib.critic_workspace(id(self), feats.device)
        mean_h, _ = _lib.critic_forward(feats, None, ce.weight, ce.bias, h1.weight, h1.bias, h2.weight, h2.bias, None, None,
                                        self.K, ws, seed=ts.seed, counter=ts.counter, counter_dev=ts.counter_dev,
                                        cvar=float(cvar), mean_hidden=True)
        return torch.addmm(self.output_layer.bias, mean_h, self.output_layer.weight.t())

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
