"""Soak: a few thousand batched steps of the graph-captured training loop per agent; parameters and the last losses
must stay finite (guards the learner shortcuts: loss gradients without autograd nodes, K13, the Rainbow append)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from distributional_rl_decision_and_control_b200.agent import Agent  # noqa: E402
from distributional_rl_decision_and_control_b200.envs.marinenav.batched import BatchedMarineNavEnv  # noqa: E402
from distributional_rl_decision_and_control_b200.vector_trainer import VectorTrainer  # noqa: E402

dev = torch.device("cuda:0")
E, R = 2048, 8
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
for kind in ("AC-IQN", "Rainbow", "SAC", "DQN"):
    agent = Agent(device=dev, seed=100, agent_type=kind, BATCH_SIZE=1024, num_tau=32, BUFFER_SIZE=1 << 20,
                  replay_lanes=E * R if kind == "Rainbow" else 1)
    env = BatchedMarineNavEnv(E, R, 10, 8, device=dev, seed=11, fused=True, prefetch="workers")
    vt = VectorTrainer(env, agent, update_every=4, learning_starts=8, cuda_graph=True)
    vt.learn(steps)
    torch.cuda.synchronize()
    nets = [m for m in (getattr(agent, "policy_local", None),) if m is not None]
    params = []
    for m in nets:
        mods = [getattr(m, n) for n in ("actor", "critic", "critic_1", "critic_2") if hasattr(m, n)] or [m]
        for mod in mods:
            params += list(mod.parameters())
    finite = all(bool(torch.isfinite(p).all()) for p in params)
    norm = float(sum(p.double().pow(2).sum() for p in params).sqrt())
    print(f"{kind}: {vt.steps} batched steps, {vt.updates} updates, parameters finite: {finite}, |theta| = {norm:.3f}", flush=True)
    assert finite
