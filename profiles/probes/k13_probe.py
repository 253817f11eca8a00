"""K13 (asvrl_c51_head) against the float64 Rainbow head, and its time at 65 536 rows."""
import copy
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from distributional_rl_decision_and_control_b200 import _lib  # noqa: E402
from test_agents_more import build  # noqa: E402

dev = "cuda:0"
pol = build("Rainbow", 7, dev)
pol.train()
support = torch.linspace(-10.0, 10.0, 51, device=dev)
for M in (4096, 5000, 65536):
    g = torch.Generator(device=dev).manual_seed(3)
    x = (torch.randn((M, 7), device=dev, generator=g) * 5, torch.randn((M, 5, 5), device=dev, generator=g) * 3,
         (torch.rand((M, 5), device=dev, generator=g) < 0.6).float())
    x = (x[0], x[1] * x[2][..., None], x[2])
    act, q = pol.greedy_actions(x, support, want_q=True)
    torch.cuda.synchronize()
    _lib.c51_head_status()
    _lib.policy_forward_status()
    ref = copy.deepcopy(pol).double()
    ref.fused_tower = False
    ref.train()
    with torch.no_grad():
        want = (ref(tuple(t.double() for t in x)) * support.double()).sum(2)
    err = float((q.double() - want).abs().max())
    top2 = want.topk(2, dim=1).values
    clear = (top2[:, 0] - top2[:, 1]) > 1e-4
    agree = bool(torch.equal(act.long()[clear], want.argmax(1)[clear]))
    print(f"M={M}: max |q - q64| = {err:.3e} (scale {float(want.abs().max()):.3f}), greedy agree on clear rows: {agree} ({int(clear.sum())})", flush=True)
from k9_probe import timed  # noqa: E402
print("greedy_actions ms", timed(lambda: pol.greedy_actions(x, support)))
