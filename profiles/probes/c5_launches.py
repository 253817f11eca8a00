"""ncu target: a few eager batched steps + updates of one config-5 agent (launch list with durations).
    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv python profiles/probes/c5_launches.py Rainbow"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from distributional_rl_decision_and_control_b200.agent import Agent  # noqa: E402
from distributional_rl_decision_and_control_b200.envs.marinenav.batched import BatchedMarineNavEnv  # noqa: E402
from distributional_rl_decision_and_control_b200.vector_trainer import VectorTrainer  # noqa: E402

kind = sys.argv[1] if len(sys.argv) > 1 else "Rainbow"
E, R = 8192, 8
dev = torch.device("cuda:0")
agent = Agent(device=dev, seed=100, agent_type=kind, BATCH_SIZE=1024, num_tau=32, BUFFER_SIZE=1 << 20,
              replay_lanes=E * R if kind == "Rainbow" else 1)
env = BatchedMarineNavEnv(E, R, 10, 8, device=dev, seed=11, fused=True, prefetch="workers")
vt = VectorTrainer(env, agent, update_every=4, learning_starts=8, cuda_graph=False)
vt.learn(16)
torch.cuda.synchronize()
print("MARK steps", vt.steps, "updates", vt.updates, flush=True)
