"""K1 (fused step) time against the number of envs: where one wave ends and what a lone CTA per SM costs (the
latency of the CTA's serial chain, which bounds the one-wave launch).  Cold L2 (256 MiB write between steps)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from distributional_rl_decision_and_control_b200.envs.marinenav.batched import BatchedMarineNavEnv  # noqa: E402

dev = torch.device("cuda:0")
gen = torch.Generator(device=dev).manual_seed(7)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
STAGGER = os.environ.get("K1_SWEEP_STAGGER", "1") == "1"   # 0: fresh episodes, (almost) no env ends inside the timed steps
for E in (148, 592, 1184, 2072, 3552, 4096, 4144, 8192):
    env = BatchedMarineNavEnv(E, 8, 10, 8, device=dev, seed=11, fused=True, prefetch="stream")
    env.reset()
    if STAGGER:
        env.state["ep_steps"].copy_(torch.randint(0, 1000, (E,), device=dev, generator=gen, dtype=torch.int32))
    pool = [torch.rand((E, 8, 2), device=dev, generator=gen) * 2 - 1 for _ in range(8)]
    for i in range(10):
        env.step_fused(pool[i % 8])
    ts = {"cold": [], "warm": []}
    for mode in ("cold",):
        for i in range(60):
            if mode == "cold":
                flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            env.step_fused(pool[i % 8])
            e1.record()
            torch.cuda.synchronize()
            ts[mode].append(e0.elapsed_time(e1) * 1e3)
    med = lambda v: sorted(v)[len(v) // 2]
    print(f"E={E:5d}  CTAs={E // 4:5d} ({E / 4 / 148:5.2f} per SM)  cold: median {med(ts['cold']):6.1f} us  min {min(ts['cold']):6.1f}  max {max(ts['cold']):6.1f}", flush=True)
