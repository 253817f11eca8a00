"""Device time of the B = 1024 small kernels of the update (graph replays): tower_bwd, K5 backward, K4."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from distributional_rl_decision_and_control_b200 import _lib  # noqa: E402
from k9_probe import timed  # noqa: E402

dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(0)
B = 1024
r = lambda *s: torch.randn(s, device=dev, generator=g)
feats = torch.relu(r(B, 256))
dfeats, so, ob = r(B, 256), r(B, 7), r(B, 5, 5)
print("tower_bwd ms", timed(lambda: _lib.tower_bwd(dfeats, feats, so, ob, 56, 40)))
pre, bias, dout = r(B, 128), r(128), r(B, 128)
ones = torch.ones((B, 128), device=dev)
print("scale_relu_bwd N=1 ms", timed(lambda: _lib.iqn_scale_relu_bwd(dout, pre, bias, ones, 1, want_dscale=False)))
print("scale_relu_bwd grouped 32 ms", timed(lambda: _lib.iqn_scale_relu_bwd(dout, pre, bias, ones[:32], 32, want_dscale=False)))
