"""K9 alone: device time of the fused collection policy forward at 65 536 ASVs (the per-GPU share of
BASELINE config 4) against round 1's K6 + 2 x K8 + actor-head sequence.  CUDA events around graph replays."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from distributional_rl_decision_and_control_b200 import _lib  # noqa: E402
from distributional_rl_decision_and_control_b200.agent import Agent  # noqa: E402
from distributional_rl_decision_and_control_b200.policy.AC_IQN_model import Actor  # noqa: E402
from test_gpu_policy import _obs  # noqa: E402

DEV = "cuda:0"


def timed(fn, reps=50, per_graph=10):
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(per_graph):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (reps * per_graph)


def main():
    M = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
    ag = Agent(device=DEV, seed=100, agent_type="AC-IQN", BATCH_SIZE=64, num_tau=8, BUFFER_SIZE=256)
    x = _obs(M, 1)
    out = {"rows": M}
    Actor.fused_policy = True
    out["k9_ms"] = timed(lambda: ag.act_batch(x, eps=0.05))
    torch.cuda.synchronize()
    _lib.policy_forward_status()
    Actor.fused_policy = False
    out["k6_k8_head_ms"] = timed(lambda: ag.act_batch(x, eps=0.05))
    Actor.fused_policy = True
    out["alg_bytes"] = M * (148 + 8)
    out["alg_gbs"] = out["alg_bytes"] / (out["k9_ms"] * 1e-3) / 1e9
    out["tflops_fp32_equiv"] = 2 * M * (256 * 128 + 128 * 128 + 1392 + 256) / (out["k9_ms"] * 1e-3) / 1e12
    print(json.dumps(out))


if __name__ == "__main__":
    main()
