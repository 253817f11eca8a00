"""Rainbow network (noisy, dueling, C51) with the reference's module names, parameter / buffer
names and file names (rfarl/rfarl/policy/Rainbow_model.py:17-182): factorised-noise linear
layers, value and advantage streams over the shared observation tower, 51 atoms per action."""
import json
import math
import os

import torch
import torch.nn as nn
import torch.nn.functional as F

from .AC_IQN_model import _ObservationTower, encoder

# collection: the dueling C51 head through K13 (asvrl_c51_head); ASVRL_C51_HEAD=0 keeps K8 blocks + asvrl_c51_greedy
C51_HEAD = os.environ.get("ASVRL_C51_HEAD", "1") != "0"


class NoisyLinear(nn.Module):
    """y = (W_mu + W_sigma * eps_out eps_in^T) x + (b_mu + b_sigma * eps_out) while training,
    the mean layer in eval mode; eps = sign(n) sqrt(|n|), n ~ N(0, 1) (Rainbow:17-53)."""

    def __init__(self, in_features, out_features, std_init=0.05):
        super().__init__()
        self.in_features = in_features
        self.out_features = out_features
        self.std_init = std_init
        self.weight_mu = nn.Parameter(torch.empty(out_features, in_features))
        self.weight_sigma = nn.Parameter(torch.empty(out_features, in_features))
        self.register_buffer('weight_epsilon', torch.empty(out_features, in_features))
        self.bias_mu = nn.Parameter(torch.empty(out_features))
        self.bias_sigma = nn.Parameter(torch.empty(out_features))
        self.register_buffer('bias_epsilon', torch.empty(out_features))
        self.reset_parameters()
        self.reset_noise()

    def reset_parameters(self):
        bound = 1.0 / math.sqrt(self.in_features)
        with torch.no_grad():
            self.weight_mu.uniform_(-bound, bound)
            self.weight_sigma.fill_(self.std_init / math.sqrt(self.in_features))
            self.bias_mu.uniform_(-bound, bound)
            self.bias_sigma.fill_(self.std_init / math.sqrt(self.out_features))

    def _scale_noise(self, size):
        x = torch.randn(size, device=self.weight_mu.device)
        return x.sign().mul_(x.abs().sqrt_())

    def reset_noise(self):
        eps_in = self._scale_noise(self.in_features)
        eps_out = self._scale_noise(self.out_features)
        self.weight_epsilon.copy_(torch.outer(eps_out, eps_in))
        self.bias_epsilon.copy_(eps_out)

    def forward(self, x):
        if self.training:
            return F.linear(x, self.weight_mu + self.weight_sigma * self.weight_epsilon,
                            self.bias_mu + self.bias_sigma * self.bias_epsilon)
        return F.linear(x, self.weight_mu, self.bias_mu)


class Rainbow_Policy(_ObservationTower):
    def __init__(self, self_dimension, object_dimension, max_object_num, self_feature_dimension,
                 object_feature_dimension, concat_feature_dimension, hidden_dimension, action_size,
                 atoms, device='cpu', seed=0):
        super().__init__()
        self._init_dims(self_dimension, object_dimension, max_object_num, self_feature_dimension,
                        object_feature_dimension, concat_feature_dimension, hidden_dimension, device, seed)
        self.action_size = action_size
        self.atoms = atoms
        self.self_encoder = encoder(self_dimension, self_feature_dimension)
        self.object_encoder = encoder(object_dimension, object_feature_dimension)
        self.hidden_layer_v = NoisyLinear(concat_feature_dimension, hidden_dimension)
        self.hidden_layer_a = NoisyLinear(concat_feature_dimension, hidden_dimension)
        self.hidden_layer_v_2 = NoisyLinear(hidden_dimension, hidden_dimension)
        self.hidden_layer_a_2 = NoisyLinear(hidden_dimension, hidden_dimension)
        self.output_layer_v = NoisyLinear(hidden_dimension, atoms)
        self.output_layer_a = NoisyLinear(hidden_dimension, action_size * atoms)

    def forward(self, x, log=False):
        """-> (B, action_size, atoms) probabilities (or log-probabilities), Rainbow:98-135."""
        feats = self._features(x)
        v = F.relu(self.hidden_layer_v(feats))
        v = self.output_layer_v(F.relu(self.hidden_layer_v_2(v))).view(-1, 1, self.atoms)
        a = F.relu(self.hidden_layer_a(feats))
        a = self.output_layer_a(F.relu(self.hidden_layer_a_2(a))).view(-1, self.action_size, self.atoms)
        q = v + a - a.mean(1, keepdim=True)
        return F.log_softmax(q, dim=2) if log else F.softmax(q, dim=2)

    @torch.no_grad()
    def greedy_actions(self, x, support, want_q=False):
        """Collection path (AGENT:235-246 for every row): the six noisy layers as GEMMs — the two output layers without
        their bias — and `asvrl_c51_greedy` for the dueling combination, softmax, expectation and argmax in one pass
        (the (M, A, atoms) probability tensor is never written).  int32 (M,) [and the (M, A) expected values]."""
        from .. import _lib

        def eff(layer):   # the layer's weight / bias with its current noise (Rainbow:50-53)
            if layer.training:
                return (layer.weight_mu + layer.weight_sigma * layer.weight_epsilon,
                        layer.bias_mu + layer.bias_sigma * layer.bias_epsilon)
            return layer.weight_mu, layer.bias_mu

        # the two noisy streams' hidden layers: K9's trunk (tower + both layers in one launch) per stream on the
        # reference's shape, K8 or F.linear otherwise (_ObservationTower._trunk)
        atoms, A, H = self.atoms, self.action_size, self.hidden_dimension
        v = self._trunk(x, self.hidden_layer_v, self.hidden_layer_v_2, eff)
        a_img = None
        if C51_HEAD and atoms <= 64 and A <= 32 and H == 128:
            # K13: the advantage layer, the dueling combination, softmax, expectation and argmax in one tcgen05 launch
            # from the advantage stream's hidden layer as an operand image (None: K9's trunk does not cover this shape)
            a_img = self._trunk(x, self.hidden_layer_a, self.hidden_layer_a_2, eff, image=True)
        if a_img is not None:
            wv, bv = eff(self.output_layer_v)
            wa, ba = eff(self.output_layer_a)
            wvp = torch.zeros((64, H), dtype=wv.dtype, device=wv.device)
            wvp[:atoms] = wv
            zv = _lib.linear_tf32x3(v.contiguous(), wvp)   # the value stream's product, 64-padded (K8)
            return _lib.c51_head(a_img, zv, wa.contiguous(), ba.contiguous(), bv.contiguous(), support, v.shape[0], A, want_q)
        a = self._trunk(x, self.hidden_layer_a, self.hidden_layer_a_2, eff)
        umma = v.is_cuda and v.shape[0] >= 4096 and self.hidden_dimension % 32 == 0 and self.hidden_dimension <= 256 \
            and v.dtype == torch.float32
        wv, bv = eff(self.output_layer_v)
        wa, ba = eff(self.output_layer_a)
        if umma and atoms <= 256:
            # the two output layers on the tensor cores too (K8 takes <= 256 outputs, a multiple of 16, per call): the
            # value layer padded to 64 outputs, the advantage layer `apb` actions at a time (5 x 51 = 255 -> 256) into
            # blocks that asvrl_c51_greedy reads in place; zero weight rows fill the padding
            vp = -(-atoms // 16) * 16
            wvp = torch.zeros((vp, H), dtype=wv.dtype, device=wv.device)
            wvp[:atoms] = wv
            zv = _lib.linear_tf32x3(v, wvp)
            apb = min(A, 256 // atoms)
            nb = -(-A // apb)
            ld = -(-(apb * atoms) // 16) * 16
            wap = torch.zeros((nb, ld, H), dtype=wa.dtype, device=wa.device)
            for k in range(nb):
                rows_k = min(apb, A - k * apb) * atoms
                wap[k, :rows_k] = wa[k * apb * atoms:k * apb * atoms + rows_k]
            za = torch.empty((nb, v.shape[0], ld), dtype=torch.float32, device=v.device)
            for k in range(nb):
                _lib.linear_tf32x3(a, wap[k], out=za[k])
            return _lib.c51_greedy(zv, za, bv.contiguous(), ba.contiguous(), support, A, want_q, actions_per_block=apb)
        return _lib.c51_greedy(v @ wv.t(), a @ wa.t(), bv.contiguous(), ba.contiguous(), support, A, want_q)

    def reset_noise(self):
        for name, module in self.named_children():
            if "hidden_layer" in name or "output_layer" in name:
                module.reset_noise()

    def get_constructor_parameters(self):
        return dict(self_dimension=self.self_dimension, object_dimension=self.object_dimension,
                    max_object_num=self.max_object_num,
                    self_feature_dimension=self.self_feature_dimension,
                    object_feature_dimension=self.object_feature_dimension,
                    concat_feature_dimension=self.concat_feature_dimension,
                    hidden_dimension=self.hidden_dimension, action_size=self.action_size,
                    atoms=self.atoms, seed=self.seed_id)

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
