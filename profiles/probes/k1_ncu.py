"""ncu target: a few fused steps of E envs (default 148: one CTA on every fourth SM, so the samples show the CTA's serial
chain without neighbours).  ncu --set full --import-source on -k regex:env_step_kernel -s 6 -c 1 python profiles/probes/k1_ncu.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from distributional_rl_decision_and_control_b200.envs.marinenav.batched import BatchedMarineNavEnv  # noqa: E402

E = int(sys.argv[1]) if len(sys.argv) > 1 else 148
dev = torch.device("cuda:0")
gen = torch.Generator(device=dev).manual_seed(7)
env = BatchedMarineNavEnv(E, 8, 10, 8, device=dev, seed=11, fused=True, prefetch="stream")
env.reset()
pool = [torch.rand((E, 8, 2), device=dev, generator=gen) * 2 - 1 for _ in range(8)]
for i in range(10):
    env.step_fused(pool[i % 8])
torch.cuda.synchronize()
print("done")
