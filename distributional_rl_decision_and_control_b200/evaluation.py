"""Vectorised evaluation: the frozen evaluation scenarios of ``Trainer.evaluation``
(rfarl/rfarl/policy/trainer.py:266-387) as ONE device batch.

The reference plays its 60 scenarios one after the other, one batch-1 policy forward per ASV per
step.  Here scenario e is env e of a batch: every step is one greedy policy forward over all
E x R ASVs and one K1 launch (finished scenarios are masked out), and the episode statistics the
reference keeps (discounted reward, travel time, energy, success, trajectories, actions) are
accumulated on the device.  Episode end rule, per scenario (TRAINER:344): 1000 steps, ANY collision,
or all vehicles deactivated.
"""
import numpy as np
import torch

from . import _abi, _lib
from .envs.marinenav.env import MarineNavEnv3

_CONTINUOUS = ("AC-IQN", "DDPG", "SAC")


class BatchedEvaluator:
    def __init__(self, eval_configs, device="cuda:0", seed=0, record=True, max_steps=1000):
        """``eval_configs``: list of scenario dicts in the ``episode_data`` schema (ENV:616-773), as
        written by ``Trainer.create_eval_configs`` / ``save_eval_config``."""
        _lib.lib()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("BatchedEvaluator needs a CUDA device (no CPU fallback)")
        self.configs = list(eval_configs)
        self.E = len(self.configs)
        if self.E == 0:
            raise ValueError("no evaluation scenario")
        self.RP = max(c["env"]["num_robots"] for c in self.configs)
        self.OP = max(len(c["env"]["obstacles"]["r"]) for c in self.configs)
        self.CP = max(len(c["env"]["cores"]["Gamma"]) for c in self.configs)
        self.seed = int(seed)
        self.record = bool(record)
        self.max_steps = int(max_steps)
        self.params = None
        with torch.cuda.device(self.device):
            self.frozen = _lib.new_state(self.E, self.RP, self.OP, self.CP, self.device)
            self.state = _lib.new_state(self.E, self.RP, self.OP, self.CP, self.device)
            self.out = _lib.new_out(self.E, self.RP, self.device, debug=False)
        self._load()

    def _load(self):
        """Every scenario goes through the single-env façade's loader (start pose, initial
        velocity + current at the start, tables) and its device row is copied into the batch."""
        one = MarineNavEnv3(seed=0, is_eval_env=True, device=self.device)
        for e, cfg in enumerate(self.configs):
            one.reset_with_eval_config(cfg)
            p = one._params()
            if self.params is None:
                self.params = p
            st = one._st
            r, o, c = len(one.robots), len(one.obstacles), len(one.cores)
            for k in _abi.STATE_FLOAT_FIELDS + ("flags",):
                self.frozen[k][e, :r] = st[k][0, :r]
            if o:
                self.frozen["obstacles"][e, :o] = st["obstacles"][0, :o]
            if c:
                self.frozen["cores"][e, :c] = st["cores"][0, :c]
            self.frozen["n_robots"][e], self.frozen["n_obs"][e], self.frozen["n_cores"][e] = r, o, c
        # WAMV:161-165: energy = power_coefficient * (|T_l| + |T_r|) * dt * N, the coefficient per robot of the config
        pc = torch.ones((self.E, self.RP), dtype=torch.float32)
        for e, cfg in enumerate(self.configs):
            vals = cfg["robots"].get("power_coefficient", [])
            pc[e, :len(vals)] = torch.tensor(vals, dtype=torch.float32)[:self.RP]
        self.power_coefficient = pc.to(self.device)
        self.dt_hold = float(self.params.dt * self.params.n_substeps)
        torch.cuda.synchronize(self.device)

    @torch.no_grad()
    def run(self, agent, gamma=None):
        """Play every scenario greedily with ``agent`` (eps = 0, eval-mode networks).  Returns a dict
        with the reference's per-scenario lists (rewards, successes, times, energies) and, when
        recording, ``trajectories`` [E][R] -> list of state rows and ``actions`` [E][R] -> list."""
        dev, E, RP = self.device, self.E, self.RP
        gamma = agent.GAMMA if gamma is None else gamma
        discrete = agent.agent_type not in _CONTINUOUS
        for k, v in self.frozen.items():
            self.state[k].copy_(v)
        st_ptrs, o_ptrs = _lib.state_ptrs(self.state), _lib.out_ptrs(self.out)
        mode = _abi.M_DYNAMICS | _abi.M_AUTO_DEACTIVATE | (_abi.M_DISCRETE if discrete else 0)
        _lib.env_step(self.params, self.state, None, None, self.out, 0, self.OP, self.CP, seed=self.seed,
                      step=0, st_ptrs=st_ptrs, o_ptrs=o_ptrs)  # first observation (ENV:614)
        nets = [m for m in (getattr(agent.policy_local, n, None) for n in ("actor", "critic", "critic_1", "critic_2"))
                if m is not None] or [agent.policy_local]
        for m in nets:
            m.eval()
        present = torch.arange(RP, device=dev).unsqueeze(0) < self.state["n_robots"].unsqueeze(1)
        rewards = torch.zeros((E, RP), device=dev)
        times = torch.zeros((E, RP), device=dev)
        energies = torch.zeros((E, RP), device=dev)
        running = torch.ones((E,), dtype=torch.bool, device=dev)
        lengths = torch.zeros((E,), dtype=torch.int32, device=dev)
        traj, acts = [], []
        _row = lambda: torch.stack([self.state[k] for k in ("x", "y", "theta", "velr_x", "velr_y", "velr_r", "vel_x",
                                                            "vel_y", "vel_r", "thrust_l", "thrust_r")], dim=-1).cpu()
        first_row = _row() if self.record else None  # the reset row every rob.trajectory starts with (WAMV:191-193)
        disc = 1.0
        try:
            for length in range(self.max_steps + 1):  # TRAINER:344: `length >= 1000` is tested after the step
                mask = running.to(torch.uint8)
                obs = (self.out["obs_self"].view(E * RP, 7), self.out["obs_objects"].view(E * RP, 5, 5),
                       self.out["obs_mask"].view(E * RP, 5))
                a = agent.act_batch(obs, 0.0)
                a = a.view(E, RP) if discrete else a.view(E, RP, 2)
                active = present & ((self.state["flags"] & 4) == 0) & running.unsqueeze(1)
                _lib.env_step(self.params, self.state, a, None, self.out, mode, self.OP, self.CP, seed=self.seed,
                              step=length + 1, env_mask=mask, st_ptrs=st_ptrs, o_ptrs=o_ptrs)
                act_f = active.to(torch.float32)
                rewards += disc * self.out["reward"] * act_f
                times += self.dt_hold * act_f
                energies += self.power_coefficient * (self.state["thrust_l"].abs() + self.state["thrust_r"].abs()) * self.dt_hold * act_f
                if self.record:
                    traj.append(_row())
                    acts.append((a.to(torch.float32).cpu(), active.cpu()))
                es = self.out["env_status"]
                ended = ((es & 1) != 0) | ((es & 4) != 0) | (length >= self.max_steps)
                lengths += running.to(torch.int32)
                running = running & ~ended
                disc *= gamma
                if length % 16 == 15 and not bool(running.any()):
                    break
        finally:
            for m in nets:
                m.train()
        flags = self.state["flags"]
        success = (((flags & 2) != 0) | ~present).all(dim=1)  # check_all_reach_goal, ENV:186-192
        n_rob = self.state["n_robots"].clamp(min=1).to(torch.float32)
        res = {"rewards": (rewards.sum(1) / n_rob).cpu().numpy().tolist(),
               "successes": success.cpu().numpy().tolist(),
               "times": (times.sum(1) / n_rob).cpu().numpy().tolist(),
               "energies": (energies.sum(1) / n_rob).cpu().numpy().tolist(),
               "lengths": lengths.cpu().numpy().tolist()}
        if self.record:
            res["trajectories"], res["actions"] = self._histories(first_row, traj, acts, lengths.cpu().numpy())
        return res

    def _histories(self, first_row, traj, acts, lengths):
        """[E][R] python lists shaped like the reference's rob.trajectory / rob.action_history: the reset row, then
        one row per step the robot was active in; one action per step (thruster angles, columns 9-10 of
        WAMV:191-193, are the constant 0 of training)."""
        nr = self.state["n_robots"].cpu().numpy()
        T = np.stack([t.numpy() for t in traj])                       # (steps, E, RP, 11)
        A = np.stack([a.numpy() for a, _ in acts])
        M = np.stack([m.numpy() for _, m in acts])
        out_t, out_a = [], []
        for e in range(self.E):
            te, ae = [], []
            for r in range(int(nr[e])):
                v0 = first_row[e, r].numpy()
                rows, al = [[float(x) for x in v0[:9]] + [0.0, 0.0] + [float(v0[9]), float(v0[10])]], []
                for s in range(int(lengths[e])):
                    if M[s, e, r]:
                        v = T[s, e, r]
                        rows.append([float(x) for x in v[:9]] + [0.0, 0.0] + [float(v[9]), float(v[10])])
                        al.append(A[s, e, r].tolist() if A.ndim == 4 else int(A[s, e, r]))
                te.append(rows)
                ae.append(al)
            out_t.append(te)
            out_a.append(ae)
        return out_t, out_a


def summarize(res):
    """The numbers ``Trainer.evaluation`` prints (TRAINER:366-379)."""
    succ = np.array(res["successes"], dtype=bool)
    avg_t = float(np.mean(np.array(res["times"])[succ])) if succ.any() else None
    avg_e = float(np.mean(np.array(res["energies"])[succ])) if succ.any() else None
    return {"avg_reward": float(np.mean(res["rewards"])), "success_rate": float(succ.mean()),
            "avg_time": avg_t, "avg_energy": avg_e}
