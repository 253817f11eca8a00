"""Golden of the reference's own `Trainer.evaluation` (rfarl/rfarl/policy/trainer.py:266-392) — TEST
INFRASTRUCTURE, run in the build container, never imported by the product.

    python oracle/gen_golden_eval.py      ->  tests/golden/eval_reference.json

Frozen scenarios from the reference's `create_eval_configs` (TRAINER:65-81) with the perception noise switched
off in the stored configs (pos_std = vel_std = 0, r_mean_ratio = 1: the config schema of ENV:601-607 carries
them per robot), played by the UNMODIFIED reference Trainer / env / agent with the weights of the reference's
shipped trained AC-IQN policy (virelex/trained/traced_AC_IQN_model.pt; tests/golden/traced_ac_iqn.npz holds the
same weights).  Stored: the eval configs, and per scenario the reference's reward / success / time / energy,
episode length, every robot's final pose and the ACTIONS the reference's policy took.  The GPU tests replay those
actions (teacher forcing, SURVEY.md 8(d): "multi-step trajectories diverge chaotically; compare step-wise with
re-injection") — measured here: free-running float32 and float64 roll-outs of this trained policy agree to 1e-5
for ~8 steps and differ by metres after ~20.  Of 28 candidate scenarios the 10 whose threshold predicates (goal,
collision, detection range) stay at least 2 cm clear of their thresholds in float64 are kept: only those can be
reproduced event for event by a float32 path (the margin is stored with each scenario).
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


MIN_MARGIN, KEEP = 0.02, 10  # metres; scenarios kept (of 28 candidates)


def main():
    from rfarl.agent import Agent
    from rfarl.envs.marinenav.env import MarineNavEnv3
    from rfarl.policy.trainer import Trainer
    sched = {"num_episodes": [8, 10, 10], "num_robots": [3, 5, 5], "num_cores": [0, 0, 4], "num_obstacles": [0, 4, 4],
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
        # Threshold predicates (goal reached at 2 m, collision at contact, detection at 20 m + r) decide discrete
        # events; a float32 path can only reproduce an episode event for event while the float64 margins of those
        # predicates stay clear of zero.  Record the smallest margin of every scenario and keep the robust ones.
        margins = []
        orig_step = eval_env.step

        def step(actions, is_continuous_action=True):
            out = orig_step(actions, is_continuous_action)
            m = margins[-1]
            for i, rob in enumerate(eval_env.robots):
                if rob.deactivated:
                    continue
                m = min(m, abs(rob.dist_to_goal() - rob.goal_dis))
                for obs in eval_env.obstacles:
                    d = np.hypot(rob.x - obs.x, rob.y - obs.y)
                    m = min(m, abs(d - obs.r - rob.r), abs(d - obs.r - rob.perception.range))
                for j, other in enumerate(eval_env.robots):
                    if j != i and not other.deactivated:
                        d = np.hypot(rob.x - other.x, rob.y - other.y)
                        m = min(m, abs(d - other.r - rob.r), abs(d - other.r - rob.perception.range))
            margins[-1] = m
            return out
        eval_env.step = step
        orig_reset = eval_env.reset_with_eval_config

        def reset_cfg(cfg):
            margins.append(np.inf)
            return orig_reset(cfg)
        eval_env.reset_with_eval_config = reset_cfg
        tr.evaluation()
    keep = [e for e in np.argsort(margins)[::-1] if margins[e] >= MIN_MARGIN][:KEEP]
    keep = sorted(int(e) for e in keep)
    print("margins", np.round(margins, 4).tolist(), "kept", keep)
    res = []
    for e in keep:
        traj = tr.eval_trajectories[0][e]
        res.append({"min_margin": float(margins[e]), "reward": float(tr.eval_rewards[0][e]), "success": bool(tr.eval_successes[0][e]),
                    "time": float(tr.eval_times[0][e]), "energy": float(tr.eval_energies[0][e]),
                    "lengths": [len(t) for t in traj],
                    "final_pose": [[float(v) for v in t[-1][:3]] for t in traj],
                    # the actions the reference's policy took (teacher forcing: closed-loop roll-outs of a trained,
                    # nearly bang-bang policy amplify a 1e-7 difference to metres within ~20 steps) and a sample of the
                    # observations it took them on (the policy itself is checked on those)
                    "actions": [[[float(v) for v in a] for a in acts] for acts in tr.eval_actions[0][e]],
                    "obs_sample": [[[float(v) for v in o[0]], [[float(v) for v in ob] for ob in o[1]]]
                                   for o in tr.eval_observations[0][e][0][:8]]})
    with open(OUT, "w") as f:
        json.dump({"gamma": agent.GAMMA, "min_margin": MIN_MARGIN, "eval_configs": [tr.eval_config[e] for e in keep],
                   "results": res}, f)
    if os.environ.get("EVAL_DEBUG_TRAJ"):  # full reference trajectories / actions / observations for debugging (not committed)
        with open(os.environ["EVAL_DEBUG_TRAJ"], "w") as f:
            json.dump({"trajectories": [tr.eval_trajectories[0][e] for e in keep],
                       "actions": [tr.eval_actions[0][e] for e in keep],
                       "observations": [tr.eval_observations[0][e] for e in keep]}, f)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")
    for r in res:
        print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in r.items() if k != "final_pose"})


if __name__ == "__main__":
    main()
