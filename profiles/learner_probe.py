"""Learner probe: one AC-IQN update at C3 (B=1024, N=32) and at the reference shape (B=64, N=8), fp32 vs
TF32 GEMMs, eager vs CUDA graph (CUDA events).  Under `ncu --metrics gpu__time_duration.sum` with
LEARNER_PROBE_NCU=1 only a few eager updates run, for the launch list."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributional_rl_decision_and_control_b200.agent import Agent
from distributional_rl_decision_and_control_b200.envs.marinenav.batched import BatchedMarineNavEnv

dev = torch.device("cuda:0")
env = BatchedMarineNavEnv(4096, 8, 10, 8, device=dev, seed=1, fused=False)
env.reset()
g = torch.Generator(device=dev).manual_seed(0)
ncu = bool(os.environ.get("LEARNER_PROBE_NCU"))


def fill(agent):
    for i in range(12):
        obs = env.obs
        a = torch.rand((4096, 8, 2), device=dev, generator=g) * 2 - 1
        out = env.step_no_reset(a)
        agent.memory.push_step(obs, a, False, out)
        env.finish_episodes()


def timed(fn, n):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for B, N in ((1024, 32), (64, 8)):
    for tf32 in (False, True):
        torch.backends.cuda.matmul.allow_tf32 = tf32
        torch.backends.cudnn.allow_tf32 = tf32
        ag = Agent(device=dev, seed=100, agent_type="AC-IQN", BATCH_SIZE=B, num_tau=N, BUFFER_SIZE=1 << 19)
        fill(ag)
        if ncu:
            if tf32:
                continue
            for _ in range(3):
                ag.train_AC_IQN()
            torch.cuda.synchronize()
            continue
        eager = timed(ag.train_AC_IQN, 30)
        ag.enable_cuda_graph()
        graph = timed(ag.train_graph_replay, 100)
        print(f"B={B} N={N} tf32={tf32}: eager {eager:.3f} ms  graph {graph:.3f} ms  -> {1e3 / graph:.0f} updates/s",
              flush=True)
