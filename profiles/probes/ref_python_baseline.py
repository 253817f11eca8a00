"""The reference's own Python loop timed on THIS container's host cores: the stated
baseline, produced by a committed script instead of quoted from the survey.  /root/reference is imported
unmodified; it cannot travel to the GPU box, so the result file is committed (profiles/ref_python_baseline.json).

    python profiles/probes/ref_python_baseline.py [--seconds S] > profiles/ref_python_baseline.json

Legs (each a bounded sample):
  c1_trainer   BASELINE config 1: rfarl/config/ac_iqn.json through Trainer.learn exactly as train_RL_agents.py's
               run_trial does (TRAIN:74-109, TRAINER:85-255), -D cpu, one process, short step budget, evaluation
               reduced to one episode; env-steps/s and learner updates/s of the whole loop
  env_only     P = nproc independent MarineNavEnv3 processes (the reference's own -P fan-out, TRAIN:131-140), R=8 ASVs,
               O=10 buoys, C=8 vortex cores (BASELINE config 2's shape), random actions; summed env-steps/s
  learner      Agent.train() of AC-IQN at the reference's B=64 / N=8 (AGENT:386-432) on a filled replay buffer
"""
import argparse
import io
import json
import multiprocessing as mp
import os
import sys
import tempfile
import time
import warnings
from contextlib import redirect_stdout

REF = "/root/reference/rfarl"
warnings.filterwarnings("ignore")


def _imports():
    if REF not in sys.path:
        sys.path.insert(0, REF)
    from rfarl.agent import Agent
    from rfarl.envs.marinenav.env import MarineNavEnv3
    from rfarl.policy.trainer import Trainer
    return Agent, MarineNavEnv3, Trainer


def c1_trainer(total_timesteps):
    import torch
    Agent, MarineNavEnv3, Trainer = _imports()
    with open(os.path.join(REF, "rfarl", "config", "ac_iqn.json")) as f:
        params = json.load(f)
    sched, ev = params["training_schedule"], params["eval_schedule"]
    ev = dict(ev, num_episodes=[1, 0, 0, 0, 0, 0])
    out = {}
    for threads in (None, 1):
        if threads:
            torch.set_num_threads(threads)
        buf = io.StringIO()
        with tempfile.TemporaryDirectory() as d, redirect_stdout(buf):
            train_env = MarineNavEnv3(seed=0, schedule=sched)
            eval_env = MarineNavEnv3(seed=253, is_eval_env=True)
            agent = Agent(device="cpu", seed=100, agent_type="AC-IQN")
            n_train = [0]
            orig = agent.train

            def counted(*a, **k):
                n_train[0] += 1
                return orig(*a, **k)
            agent.train = counted
            tr = Trainer(train_env=train_env, eval_env=eval_env, eval_schedule=ev, rl_agent=agent,
                         imitation=False, il_agent=None)
            tr.save_eval_config(d)
            t0 = time.perf_counter()
            tr.learn(total_timesteps=total_timesteps, eval_freq=10 ** 9, eval_log_path=d)
            dt = time.perf_counter() - t0
        out["torch_threads_%s" % (threads or "default(%d)" % torch.get_num_threads())] = {
            "env_steps_per_sec": train_env.total_timesteps / dt, "updates_per_sec": n_train[0] / dt,
            "env_steps": train_env.total_timesteps, "updates": n_train[0], "seconds": dt}
    return out


def _env_worker(args):
    seed, seconds, R, O, C = args
    import numpy as np
    _, MarineNavEnv3, _ = _imports()
    with redirect_stdout(io.StringIO()):
        env = MarineNavEnv3(seed=seed)
        env.num_robots, env.num_obs, env.num_cores = R, O, C
        env.reset()
        rng = np.random.RandomState(seed)
        steps = asv = 0
        t0 = time.perf_counter()
        while time.perf_counter() - t0 < seconds:
            acts = [list(rng.uniform(-1, 1, 2)) if not rob.deactivated else None for rob in env.robots]
            asv += sum(a is not None for a in acts)
            _, _, dones, infos = env.step(acts, True)
            steps += 1
            for rob, info in zip(env.robots, infos):  # what Trainer.learn does (TRAINER:168-172)
                if info["state"] in ("deactivated after collision", "deactivated after reaching goal"):
                    continue
                if rob.collision or rob.reach_goal:
                    rob.deactivated = True
            if env.check_all_deactivated() or env.episode_timesteps >= 1000:
                env.reset()
        dt = time.perf_counter() - t0
    return steps / dt, asv / dt


def env_only(seconds, procs, R=8, O=10, C=8):
    with mp.get_context("fork").Pool(procs) as pool:
        res = pool.map(_env_worker, [(s, seconds, R, O, C) for s in range(procs)])
    return {"processes": procs, "robots": R, "buoys": O, "vortex_cores": C,
            "env_steps_per_sec": sum(r[0] for r in res), "active_asv_steps_per_sec": sum(r[1] for r in res),
            "env_steps_per_sec_per_core": sum(r[0] for r in res) / procs, "seconds": seconds}


def learner(seconds):
    import numpy as np
    import torch
    Agent, MarineNavEnv3, _ = _imports()
    with redirect_stdout(io.StringIO()):
        env = MarineNavEnv3(seed=1)
        agent = Agent(device="cpu", seed=100, agent_type="AC-IQN")
        obs, _, _ = env.reset()
        rng = np.random.RandomState(0)
        while agent.memory.size() < 512:
            acts = [list(rng.uniform(-1, 1, 2)) for _ in env.robots]
            nobs, rew, dones, infos = env.step(acts, True)
            for i in range(len(env.robots)):
                agent.memory.add((obs[i], acts[i], rew[i], nobs[i], dones[i]))
            obs = nobs
            if env.episode_timesteps >= 200:
                obs, _, _ = env.reset()
        n = 0
        t0 = time.perf_counter()
        while time.perf_counter() - t0 < seconds:
            agent.train()
            n += 1
        dt = time.perf_counter() - t0
    return {"batch": agent.BATCH_SIZE, "n_tau": 8, "updates_per_sec": n / dt, "updates": n,
            "torch_threads": torch.get_num_threads(), "seconds": dt}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=20.0)
    ap.add_argument("--c1-steps", type=int, default=2600, help="TRAINER learning_starts is 2000: 2600 gives ~150 updates")
    a = ap.parse_args()
    import torch
    cores = len(os.sched_getaffinity(0))
    out = {"what": "the unmodified Python reference (/root/reference) timed in the build container",
           "host": {"cores": cores, "torch": torch.__version__, "python": sys.version.split()[0]},
           "env_only": env_only(a.seconds, cores),
           "env_only_one_process": env_only(a.seconds / 2, 1),
           "learner": learner(a.seconds / 2),
           "c1_trainer": c1_trainer(a.c1_steps)}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
