"""ncu target: K10 / K11 / K12 alone, eager launches (B x N quantile rows; see profiles/README.md for the command)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from distributional_rl_decision_and_control_b200 import _lib  # noqa: E402
from test_gpu_critic import _critic, _inputs, _params  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
N = int(sys.argv[2]) if len(sys.argv) > 2 else 32
c = _critic()
tau, feats, aenc = _inputs(B, N, 1)
ps = _params(c)
ws = _lib.critic_workspace("probe", feats.device)
dz = torch.randn((B, N), device="cuda:0") / (B * N)
for _ in range(3):
    z, _ = _lib.critic_forward(feats, aenc, *ps, N, ws, tau_in=tau)
    if B * N <= 1 << 17:
        z, _, (cos_i, h1_i, h2_i, r3_i) = _lib.critic_forward(feats, aenc, *ps, N, ws, tau_in=tau, save=True)
        daenc, dfeats, (dp3, dp2, dp1, aux), scales = _lib.critic_backward(dz, r3_i, h2_i, h1_i, feats, aenc, ps[2], ps[4], ps[6], N,
                                                                          fmax=feats.amax().reshape(1))
        _lib.critic_wgrad(dp3, dp2, dp1, aux, cos_i, h1_i, h2_i, r3_i, scales, B, N)
        _lib.critic_backward(dz, r3_i, h2_i, None, feats, aenc, ps[2], ps[4], ps[6], N, full=False)
torch.cuda.synchronize()
_lib.critic_forward_status()
print("ok", float(z.abs().mean()))
