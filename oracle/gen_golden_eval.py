"""Golden of the reference's own `Trainer.evaluation` (rfarl/rfarl/policy/trainer.py:266-392) — TEST
INFRASTRUCTURE, run in the build container, never imported by the product.

    python oracle/gen_golden_eval.py      ->  tests/golden/eval_reference.json

Frozen scenarios from the reference's `create_eval_configs` (TRAINER:65-81) with the perception noise switched
off in the stored configs (pos_std = vel_std = 0, r_mean_ratio = 1: the config schema of ENV:601-607 carries
them per robot), played by the UNMODIFIED reference Trainer / env / agent with the weights of the reference's
shipped trained AC-IQN policy (virelex/trained/traced_AC_IQN_model.pt; tests/golden/traced_ac_iqn.npz holds the
same weights).  Stored: the eval configs, and per scenario the reference's reward / success / time / energy,
episode length and every robot's final pose.
"""
import io
import json
import os
import sys
import warnings
from contextlib import redirect_stdout

import numpy as np
import torch

warnings.filterwarnings("ignore")
REF = "/root/reference"
sys.path.insert(0, os.path.join(REF, "rfarl"))
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "eval_reference.json")


def main():
    from rfarl.agent import Agent
    from rfarl.envs.marinenav.env import MarineNavEnv3
    from rfarl.policy.trainer import Trainer
    sched = {"num_episodes": [3, 3, 2], "num_robots": [3, 5, 5], "num_cores": [0, 0, 4], "num_obstacles": [0, 4, 4],
             "min_start_goal_dis": [30.0, 40.0, 40.0]}
    with redirect_stdout(io.StringIO()):
        train_env = MarineNavEnv3(seed=0)
        eval_env = MarineNavEnv3(seed=253, is_eval_env=True)
        agent = Agent(device="cpu", seed=100, agent_type="AC-IQN")
        shipped = torch.jit.load(os.path.join(REF, "virelex", "trained", "traced_AC_IQN_model.pt"))
        agent.policy_local.actor.load_state_dict({k: v.float() for k, v in shipped.state_dict().items()})
        tr = Trainer(train_env=train_env, eval_env=eval_env, eval_schedule=sched, rl_agent=agent)
        for cfg in tr.eval_config:  # noise-free perception
            n = len(cfg["robots"]["perception"]["pos_std"])
            cfg["robots"]["perception"]["pos_std"] = [0.0] * n
            cfg["robots"]["perception"]["vel_std"] = [0.0] * n
            cfg["robots"]["perception"]["r_mean_ratio"] = [1.0] * n
        tr.evaluation()
    res = []
    for e in range(len(tr.eval_config)):
        traj = tr.eval_trajectories[0][e]
        res.append({"reward": float(tr.eval_rewards[0][e]), "success": bool(tr.eval_successes[0][e]),
                    "time": float(tr.eval_times[0][e]), "energy": float(tr.eval_energies[0][e]),
                    "lengths": [len(t) for t in traj],
                    "final_pose": [[float(v) for v in t[-1][:3]] for t in traj],
                    "n_actions": [len(a) for a in tr.eval_actions[0][e]]})
    with open(OUT, "w") as f:
        json.dump({"gamma": agent.GAMMA, "eval_configs": tr.eval_config, "results": res}, f)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")
    for r in res:
        print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in r.items() if k != "final_pose"})


if __name__ == "__main__":
    main()
