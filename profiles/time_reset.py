"""Times K1b alone (CUDA events): a full reset of E envs and a sparse one (E/1000 envs), C2 shape."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributional_rl_decision_and_control_b200 import _lib
from distributional_rl_decision_and_control_b200.envs.marinenav.batched import BatchedMarineNavEnv

E = 4096
env = BatchedMarineNavEnv(E, 8, 10, 8, device="cuda:0", seed=0, fused=False)
env.reset()
sparse = torch.zeros(E, dtype=torch.uint8, device="cuda:0")
sparse[::997] = 1
for name, mask in (("full", None), ("sparse(5 envs)", sparse)):
    ts = []
    for i in range(20):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.env_reset(env.reset_params, env.state, mask, env.want_robots, env.want_cores, env.want_obs,
                       env.min_start_goal_dis, 0, 100 + i, env.OP, env.CP, 0)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    print(f"K1b {name}: median {ts[len(ts)//2]:.1f} us, min {ts[0]:.1f} us")
print("placed obstacles histogram:", torch.bincount(env.state["n_obs"], minlength=11).tolist())
print("placed cores histogram:", torch.bincount(env.state["n_cores"], minlength=9).tolist())
