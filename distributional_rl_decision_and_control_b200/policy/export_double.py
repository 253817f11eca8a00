"""Float64 export twins of the policies (the reference's policy/*_model_dtype_double.py, loaded by
rfarl/rfarl/scripts/save_model_to_torch_script.py:16-69 and traced for the libtorch deployment node,
virelex/src/action_planner_node.cpp:309-316).  Each class loads this package's float32 checkpoint files
(`*_network_params.pth` + constructor JSON — the reference's on-disk format), holds the weights in float64 and
has the forward signature the export script traces:

    Actor (AC-IQN / DDPG)  (self (1,7), objects (1,5,5), mask (1,5))          -> actions (1,2)
    IQN_Policy             (state, taus (1,32))                                -> quantiles (1,32,25)
    DQN_Policy             state                                               -> q (1,25)
    SAC Actor              state                                               -> (mu, log_std)
    Rainbow_Policy         state                                               -> probabilities (1,25,atoms)

CPU / module-by-module arithmetic only (tracing cannot follow a custom CUDA kernel, and float64 is not a kernel
type): the observation tower's fused path is never taken for float64 tensors.
"""
import torch
import torch.nn.functional as F

from . import AC_IQN_model, DDPG_model, DQN_model, IQN_model, Rainbow_model, SAC_model


def _to_double(model):
    model.double()
    if hasattr(model, "pis"):
        model.pis = model.pis.double()
    return model


class AC_IQN_Actor(AC_IQN_model.Actor):
    @classmethod
    def load(cls, directory, device="cpu"):
        return _to_double(super().load(directory, device))


class DDPG_Actor(DDPG_model.Actor):
    @classmethod
    def load(cls, directory, device="cpu"):
        return _to_double(super().load(directory, device))


class IQN_Policy(IQN_model.IQN_Policy):
    """forward(x, taus): the quantile fractions are an INPUT of the exported graph (IQN_dd:56-110)."""

    def forward(self, x, taus):
        feats = self._features(x)
        B, K = feats.shape[0], taus.shape[1]
        cos = torch.cos(taus.unsqueeze(-1) * self.pis.to(taus.dtype))
        emb = F.relu(self.cos_embedding(cos.reshape(B * K, self.n)))
        h = (feats.unsqueeze(1) * emb.reshape(B, K, -1)).reshape(B * K, -1)
        h = F.relu(self.hidden_layer(h))
        h = F.relu(self.hidden_layer_2(h))
        return self.output_layer(h).reshape(B, K, self.action_size)

    @classmethod
    def load(cls, directory, device="cpu"):
        return _to_double(super().load(directory, device))


class DQN_Policy(DQN_model.DQN_Policy):
    @classmethod
    def load(cls, directory, device="cpu"):
        return _to_double(super().load(directory, device))


class SAC_Actor(SAC_model.Actor):
    """forward(x) -> (mu, log_std): the deployed node takes the mean action (SAC_dd:175-187)."""

    def forward(self, x):
        h = F.relu(self.hidden_layer(self._features(x)))
        h = F.relu(self.hidden_layer_2(h))
        mu = self.output_layer_mu(h)
        return mu, torch.clamp(self.output_layer_log_std(h), self.min_log_std, self.max_log_std)

    @classmethod
    def load(cls, directory, device="cpu"):
        return _to_double(super().load(directory, device))


class Rainbow_Policy(Rainbow_model.Rainbow_Policy):
    @classmethod
    def load(cls, directory, device="cpu"):
        return _to_double(super().load(directory, device))
