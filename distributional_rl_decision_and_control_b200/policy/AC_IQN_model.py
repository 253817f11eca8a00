"""AC-IQN actor / critic with the reference's module names, state-dict keys, constructor JSON
and file names (rfarl/rfarl/policy/AC_IQN_model.py:16-515), so checkpoints written by either
side load in the other and `save_model_to_torch_script.py` keeps working.

The dense layers stay in PyTorch (the only contractions on the path); the critic's quantile
embedding input — tau sampling and cos(pi*i*tau) — is kernel K3 (`asvrl_iqn_tau_cos`).
Layers are created in the reference's order after `torch.manual_seed(seed)`, so a given seed
yields the reference's initial weights.
"""
import copy
import json
import os

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from .iqn_ops import (TauSampler, action_encoder, atan_head, critic_trunk, critic_trunk_umma, linear_scale_relu, narrow_ok,
                      observation_tower, thin_linear, umma_trunk_ok)


_UNIT = {}


def _unit_scale(B, D, device):
    """ones (B, D): the unit `scale` of K5's scale_relu when a layer is a plain Linear + ReLU."""
    key = (B, D, device)
    t = _UNIT.get(key)
    if t is None:
        if len(_UNIT) > 64:
            _UNIT.clear()
        t = _UNIT[key] = torch.ones((B, D), dtype=torch.float32, device=device)
        t._asvrl_unit = True  # lets K5's backward regroup the rows (iqn_ops._LinearScaleRelu)
    return t


def encoder(input_dimension, output_dimension):
    """Linear + ReLU as a 2-element Sequential: state-dict keys `<name>.0.weight/bias`."""
    return nn.Sequential(nn.Linear(input_dimension, output_dimension), nn.ReLU())


class _ObservationTower(nn.Module):
    """Shared front end: self 7->56, each of 5 objects 5->40 (masked), concat -> 256
    (ACIQN:284-308 / 428-451).  On CUDA float32 tensors it is kernel K6; the module-by-module
    path below it serves the float64 / CPU TorchScript export and is the parity reference."""
    fused_tower = True

    def _init_dims(self, self_dimension, object_dimension, max_object_num, self_feature_dimension,
                   object_feature_dimension, concat_feature_dimension, hidden_dimension, device, seed):
        self.self_dimension = self_dimension
        self.object_dimension = object_dimension
        self.max_object_num = max_object_num
        self.self_feature_dimension = self_feature_dimension
        self.object_feature_dimension = object_feature_dimension
        self.concat_feature_dimension = concat_feature_dimension
        self.hidden_dimension = hidden_dimension
        self.device = device
        self.seed_id = seed
        self.seed = torch.manual_seed(seed)

    def _features(self, x):
        assert len(x) == 3, "The number of elements in state must be 3!"
        self_obs, objects, mask = x
        B = self_obs.shape[0]
        if (self.fused_tower and objects is not None and self_obs.is_cuda and self_obs.dtype == torch.float32
                and self.self_encoder[0].weight.dtype == torch.float32):
            # K6: both encoders, the mask and the concat in one launch (one more for the backward)
            se, oe = self.self_encoder[0], self.object_encoder[0]
            return observation_tower(self_obs, objects.reshape(B, self.max_object_num, self.object_dimension),
                                     mask.to(torch.float32), se.weight, se.bias, oe.weight, oe.bias)
        own = self.self_encoder(self_obs)
        width = self.max_object_num * self.object_feature_dimension
        if objects is None:  # empty batch special case, RB:60-63
            others = own.new_zeros((B, width))
        else:
            enc = self.object_encoder(objects.reshape(B * self.max_object_num, self.object_dimension))
            enc = enc.reshape(B, self.max_object_num, self.object_feature_dimension)
            others = (enc * (mask >= 0.5).unsqueeze(-1).to(enc.dtype)).reshape(B, width)
        return torch.cat((own, others), dim=1)

    def _trunk(self, x, l1, l2, eff=None, image=False):
        """relu(l2(relu(l1(features(x))))).  A no-grad pass over >= 4 096 rows of the reference's fixed shape
        (7 / 5 x 5 observations, 56 + 5 x 40 features, 256 -> 128 -> 128) in fp32 mode runs tower and both layers in ONE
        persistent tcgen05 launch (K9's trunk, asvrl_policy_hidden: 42 us at 65 536 rows; K6 + 2 x K8 took 145 us).
        ``eff(layer)`` -> (weight, bias) overrides the parameters (Rainbow's noisy layers).  ``image``: return the
        tcgen05 operand image of the result (for asvrl_c51_head), or None when the one-launch path does not apply."""
        w1, b1 = eff(l1) if eff else (l1.weight, l1.bias)
        w2, b2 = eff(l2) if eff else (l2.weight, l2.bias)
        so, ob, mk = x
        if (not torch.is_grad_enabled() and self.fused_tower and ob is not None and so.is_cuda and so.dtype == torch.float32
                and so.shape[0] >= 4096 and w1.dtype == torch.float32 and not torch.backends.cuda.matmul.allow_tf32
                and (self.self_dimension, self.object_dimension, self.max_object_num) == (7, 5, 5)
                and (self.self_feature_dimension, self.object_feature_dimension) == (56, 40)
                and tuple(w1.shape) == (128, 256) and tuple(w2.shape) == (128, 128)):
            from .. import _lib
            se, oe = self.self_encoder[0], self.object_encoder[0]
            M = so.shape[0]
            return _lib.policy_hidden(so.contiguous(), ob.reshape(M, 5, 5).contiguous(), mk.to(torch.float32).contiguous(),
                                      se.weight, se.bias, oe.weight, oe.bias, w1.contiguous(), b1.contiguous(),
                                      w2.contiguous(), b2.contiguous(), image=image)
        if image:
            return None
        return self._hidden_w(self._features(x), w1, b1, w2, b2)

    def _hidden(self, feats, l1, l2):
        return self._hidden_w(feats, l1.weight, l1.bias, l2.weight, l2.bias)

    def _hidden_w(self, feats, w1, b1, w2, b2):
        """relu(relu(feats w1^T + b1) w2^T + b2).  A no-grad pass over >= 4 096 rows in fp32 mode whose shape K9's trunk
        does not cover runs the two layers through K8 (three TF32 products per step, bias + ReLU in the epilogue; 55 +
        40 us at 65 536 rows against ~250 us for two cuBLAS SIMT sgemm + bias + ReLU); everything else is F.linear."""
        if (not torch.is_grad_enabled() and feats.is_cuda and feats.dtype == torch.float32 and feats.shape[0] >= 4096
                and w1.dtype == torch.float32 and not torch.backends.cuda.matmul.allow_tf32 and self.fused_tower
                and w1.shape[1] % 32 == 0 and w2.shape[1] % 32 == 0
                and w1.shape[0] % 16 == 0 and w1.shape[0] <= 256 and w2.shape[0] % 16 == 0 and w2.shape[0] <= 256):
            from .. import _lib
            h = _lib.linear_tf32x3(feats.contiguous(), w1.contiguous(), b1.contiguous(), relu=True)
            return _lib.linear_tf32x3(h, w2.contiguous(), b2.contiguous(), relu=True)
        return F.relu(F.linear(F.relu(F.linear(feats, w1, b1)), w2, b2))

    def _dump(self, directory, stem):
        torch.save(self.state_dict(), os.path.join(directory, f"{stem}_network_params.pth"))
        with open(os.path.join(directory, f"{stem}_constructor_params.json"), mode="w") as f:
            json.dump(self.get_constructor_parameters(), f)

    @classmethod
    def _restore(cls, directory, stem, device):
        weights = torch.load(os.path.join(directory, f"{stem}_network_params.pth"), map_location=device)
        with open(os.path.join(directory, f"{stem}_constructor_params.json"), mode="r") as f:
            kwargs = json.load(f)
        kwargs["device"] = device
        model = cls(**kwargs)
        model.load_state_dict(weights)
        model.to(device)
        return model


class Actor(_ObservationTower):
    """Deterministic policy: features -> 128 -> 128 -> 2, squashed by (2/pi) atan (ACIQN:243-323)."""

    def __init__(self, self_dimension, object_dimension, max_object_num, self_feature_dimension,
                 object_feature_dimension, concat_feature_dimension, hidden_dimension,
                 value_ranges_of_action, device='cpu', seed=0):
        super().__init__()
        self._init_dims(self_dimension, object_dimension, max_object_num, self_feature_dimension,
                        object_feature_dimension, concat_feature_dimension, hidden_dimension, device, seed)
        self.value_ranges_of_action = copy.deepcopy(value_ranges_of_action)
        self.action_dimension = len(self.value_ranges_of_action)
        self.atan_scale = 2.0 / np.pi
        self.self_encoder = encoder(self_dimension, self_feature_dimension)
        self.object_encoder = encoder(object_dimension, object_feature_dimension)
        self.hidden_layer = nn.Linear(concat_feature_dimension, hidden_dimension)
        self.hidden_layer_2 = nn.Linear(hidden_dimension, hidden_dimension)
        self.output_layer = nn.Linear(hidden_dimension, self.action_dimension)

    def observation_processor(self, x):
        return self._features(x)

    def forward(self, x):
        feats = self._features(x)
        if self.fused_tower and feats.is_cuda and feats.dtype == torch.float32:
            # bias + ReLU of the two hidden layers in K5 (one launch per layer, one for its backward)
            ones = _unit_scale(feats.shape[0], self.hidden_dimension, feats.device)
            h = linear_scale_relu(feats, self.hidden_layer.weight, self.hidden_layer.bias, ones, 1)
            h = linear_scale_relu(h, self.hidden_layer_2.weight, self.hidden_layer_2.bias, ones, 1)
            if self.action_dimension == 2 and narrow_ok(h, self.hidden_dimension):
                # output layer, atan and scale in one launch (one more for the whole backward)
                return atan_head(h, self.output_layer.weight, self.output_layer.bias, self.atan_scale)
            return self.atan_scale * torch.atan(thin_linear(h, self.output_layer.weight, self.output_layer.bias))
        h = F.relu(self.hidden_layer(feats))
        h = F.relu(self.hidden_layer_2(h))
        return self.atan_scale * torch.atan(self.output_layer(h))  # into (-1, 1), ACIQN:320-321

    umma_linear = True  # collection path: dense layers through asvrl_linear_tf32x3 (False: cuBLAS + K5)
    fused_policy = os.environ.get("ASVRL_POLICY_FUSED", "1") != "0"  # K9: the whole forward in one launch

    def _fused_policy_ok(self, x):
        """K9 is compiled for the reference's fixed network shape (AGENT:20-26) in float32."""
        return (self.fused_policy and x[1] is not None and x[0].is_cuda and x[0].dtype == torch.float32
                and self.hidden_layer.weight.dtype == torch.float32 and x[2].dtype == torch.float32
                and (self.self_dimension, self.object_dimension, self.max_object_num, self.self_feature_dimension,
                     self.object_feature_dimension, self.concat_feature_dimension, self.hidden_dimension,
                     self.action_dimension) == (7, 5, 5, 56, 40, 256, 128, 2))

    @torch.no_grad()
    def act_eps_greedy(self, x, eps, seed, counter_dev):
        """Collection path: forward + epsilon-greedy for every row in one launch sequence — K6, the two
        dense layers (K8 on the tensor cores, or two GEMMs with one K5 pass between them), and the actor-head kernel that goes from the second
        hidden layer's pre-activation to the explored action (ReLU, 128 -> 2, atan squash, Philox
        exploration).  ``eps``: float or 0-d CUDA tensor; ``counter_dev``: int64 device word the caller
        advances after every call (graph-replay safe).  Two-dimensional actions only."""
        from .. import _lib
        lo = [r[0] for r in self.value_ranges_of_action]
        hi = [r[1] for r in self.value_ranges_of_action]
        if self._fused_policy_ok(x):
            # K9: tower, both dense layers (tcgen05, fp16 two-piece split), head and exploration in one
            # persistent launch; the activations never leave the SM
            se, oe = self.self_encoder[0], self.object_encoder[0]
            return _lib.policy_forward(x[0], x[1].reshape(-1, self.max_object_num, self.object_dimension), x[2],
                                       se.weight, se.bias, oe.weight, oe.bias, self.hidden_layer.weight,
                                       self.hidden_layer.bias, self.hidden_layer_2.weight, self.hidden_layer_2.bias,
                                       self.output_layer.weight, self.output_layer.bias, self.atan_scale, eps, lo, hi,
                                       seed, 0, counter_dev)
        feats = self._features(x)
        if self.umma_linear and feats.dtype == torch.float32 and not torch.backends.cuda.matmul.allow_tf32:
            # K8: both dense layers on the tcgen05 tensor cores at fp32 accuracy (3 x TF32 split), bias
            # and ReLU in the epilogue; when the caller allowed plain TF32 the one-pass cuBLAS GEMM is faster
            h = _lib.linear_tf32x3(feats, self.hidden_layer.weight, self.hidden_layer.bias, relu=True)
            pre = _lib.linear_tf32x3(h, self.hidden_layer_2.weight, None, relu=False)
        else:
            ones = _unit_scale(feats.shape[0], self.hidden_dimension, feats.device)
            h = linear_scale_relu(feats, self.hidden_layer.weight, self.hidden_layer.bias, ones, 1)
            pre = h @ self.hidden_layer_2.weight.t()
        return _lib.actor_head(pre, self.hidden_layer_2.bias, self.output_layer.weight, self.output_layer.bias,
                               self.atan_scale, eps, lo, hi, seed, 0, counter_dev)

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


class Critic(_ObservationTower):
    """IQN critic Z(s, a, tau) (ACIQN:362-480): cosine embedding 64 -> 256 (ReLU) multiplies the
    state features, 256 -> 128, multiplied by the action encoding 2 -> 128, 128 -> 128 -> 1."""

    def __init__(self, self_dimension, object_dimension, max_object_num, self_feature_dimension,
                 object_feature_dimension, concat_feature_dimension, hidden_dimension,
                 action_dimension, device='cpu', seed=0):
        super().__init__()
        self._init_dims(self_dimension, object_dimension, max_object_num, self_feature_dimension,
                        object_feature_dimension, concat_feature_dimension, hidden_dimension, device, seed)
        self.action_dimension = action_dimension
        self.K = 32  # quantiles used when acting, ACIQN:389
        self.n = 64  # cosine features, ACIQN:390
        self.self_encoder = encoder(self_dimension, self_feature_dimension)
        self.object_encoder = encoder(object_dimension, object_feature_dimension)
        self.pis = torch.FloatTensor([np.pi * i for i in range(self.n)]).view(1, 1, self.n).to(device)
        self.cos_embedding = nn.Linear(self.n, concat_feature_dimension)
        self.action_encoder = encoder(action_dimension, hidden_dimension)
        self.hidden_layer = nn.Linear(concat_feature_dimension, hidden_dimension)
        self.hidden_layer_2 = nn.Linear(hidden_dimension, hidden_dimension)
        self.output_layer = nn.Linear(hidden_dimension, 1)
        self.tau_sampler = TauSampler(seed, self.n)

    def calc_cos(self, batch_size, num_tau=8, cvar=1.0, taus=None):
        """K3: tau ~ U(0,1)*cvar from the counter RNG (or injected) and cos(tau*pi*i), i < 64."""
        dev = self.cos_embedding.weight.device
        cos, taus = self.tau_sampler.draw(batch_size, num_tau, cvar, dev, taus)
        assert cos.shape == (batch_size, num_tau, self.n), "cos shape is incorrect"
        return cos, taus

    def observation_processor(self, x, num_tau=8, cvar=1.0, taus=None):
        feats = self._features(x)
        B = feats.shape[0]
        cos, taus = self.calc_cos(B, num_tau, cvar, taus)
        emb = F.relu(self.cos_embedding(cos.reshape(B * num_tau, self.n)))
        emb = emb.reshape(B, num_tau, self.concat_feature_dimension)
        return (feats.unsqueeze(1) * emb).reshape(B * num_tau, self.concat_feature_dimension), taus

    def forward(self, x, actions, num_tau=8, cvar=1.0, taus=None):
        assert len(x) == 3, "The number of elements in state must be 3!"
        B = x[0].shape[0]
        h, taus = self.observation_processor(x, num_tau, cvar, taus)
        h = F.relu(self.hidden_layer(h)).reshape(B, num_tau, self.hidden_dimension)
        h = (self.action_encoder(actions).unsqueeze(1) * h).reshape(B * num_tau, self.hidden_dimension)
        h = F.relu(self.hidden_layer_2(h))
        return self.output_layer(h).reshape(B, num_tau), taus

    def forward_fused(self, x, actions, num_tau=8, cvar=1.0, taus=None, param_grads=True):
        """Same function as ``forward`` with the trunk's glue in K5 (one launch per layer instead
        of autograd's chain of element-wise kernels).  ``param_grads=False`` is the actor update
        (AGENT:420-427): only d z / d actions flows back, the critic's own parameter gradients —
        which the reference computes and then discards at the next zero_grad — are skipped."""
        assert len(x) == 3, "The number of elements in state must be 3!"
        ps = [self.cos_embedding.weight, self.cos_embedding.bias, self.hidden_layer.weight,
              self.hidden_layer.bias, self.hidden_layer_2.weight, self.hidden_layer_2.bias,
              self.output_layer.weight, self.output_layer.bias]
        enc = self.action_encoder[0]
        B = actions.shape[0]
        ones = _unit_scale(B, self.hidden_dimension, actions.device)
        narrow = self.action_dimension == 2 and narrow_ok(actions, self.hidden_dimension)
        if param_grads:
            feats = self._features(x)
            aenc = action_encoder(actions, enc.weight, enc.bias) if narrow else \
                linear_scale_relu(actions, enc.weight, enc.bias, ones, 1)
        else:
            with torch.no_grad():
                feats = self._features(x)
            ps = [p.detach() for p in ps]
            aenc = action_encoder(actions, enc.weight.detach(), enc.bias.detach()) if narrow else \
                linear_scale_relu(actions, enc.weight.detach(), enc.bias.detach(), ones, 1)
        need_grad = torch.is_grad_enabled() and (feats.requires_grad or aenc.requires_grad or ps[0].requires_grad)
        if umma_trunk_ok(feats, aenc, ps[0], ps[2], ps[4], num_tau, need_grad):
            # K10 (+ K11 / K12 in the backward): cosine embedding, the three dense layers and the head in one
            # persistent tcgen05 launch; the (B, N, 64) cosine table never exists
            ts = self.tau_sampler
            if taus is not None:
                tau_in, draw = taus.reshape(B, num_tau).to(torch.float32).contiguous(), (0, 0, None, 1.0)
            else:
                ts.counter += 1
                tau_in, draw = None, (ts.seed, ts.counter, ts.counter_dev, float(cvar))
            z, tau = critic_trunk_umma(feats, aenc, *ps, num_tau, tau_in, draw, id(self))
            return z, tau.unsqueeze(-1)
        cos, taus = self.calc_cos(B, num_tau, cvar, taus)
        z = critic_trunk(cos.reshape(B * num_tau, self.n), feats, aenc, *ps, num_tau)
        return z, taus

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


class AC_IQN_Policy():
    """Actor + critic pair (ACIQN:16-78); the critic is seeded with seed + 1."""

    def __init__(self, self_dimension, object_dimension, max_object_num, self_feature_dimension,
                 object_feature_dimension, concat_feature_dimension, hidden_dimension,
                 value_ranges_of_action, device='cpu', seed=0):
        self.actor = Actor(self_dimension, object_dimension, max_object_num, self_feature_dimension,
                           object_feature_dimension, concat_feature_dimension, hidden_dimension,
                           value_ranges_of_action, device, seed).to(device)
        self.critic = Critic(self_dimension, object_dimension, max_object_num,
                             self_feature_dimension, object_feature_dimension,
                             concat_feature_dimension, hidden_dimension,
                             len(value_ranges_of_action), device, seed + 1).to(device)

    def save(self, directory):
        self.actor.save(directory)
        self.critic.save(directory)

    @classmethod
    def load(cls, directory, device="cpu"):
        actor = Actor.load(directory, device)
        critic = Critic.load(directory, device)
        a = actor
        policy = cls(a.self_dimension, a.object_dimension, a.max_object_num,
                     a.self_feature_dimension, a.object_feature_dimension,
                     a.concat_feature_dimension, a.hidden_dimension, a.value_ranges_of_action,
                     device=device)
        policy.actor, policy.critic = actor, critic
        return policy
