"""ncu target: K9 alone, eager launches at 65 536 ASVs (see profiles/README.md for the command)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from distributional_rl_decision_and_control_b200 import _lib  # noqa: E402
from distributional_rl_decision_and_control_b200.agent import Agent  # noqa: E402
from test_gpu_policy import _obs  # noqa: E402

M = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
ag = Agent(device="cuda:0", seed=100, agent_type="AC-IQN", BATCH_SIZE=64, num_tau=8, BUFFER_SIZE=256)
x = _obs(M, 1)
for _ in range(4):
    a = ag.act_batch(x, eps=0.05)
torch.cuda.synchronize()
_lib.policy_forward_status()
print("ok", float(a.abs().mean()))
