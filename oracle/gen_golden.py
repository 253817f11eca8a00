"""Generates tests/golden/*.npz by running the UNMODIFIED Python reference in this container.

    python -m oracle.gen_golden            # from the repo root; needs /root/reference

Every fixture stores the inputs (float32-representable, held as float64), the parameter
overrides, and the reference's outputs (float64 / ints).  The fixtures travel to the GPU box;
the reference does not.  TEST INFRASTRUCTURE ONLY.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from distributional_rl_decision_and_control_b200 import _abi  # noqa: E402
from oracle import ref_harness as rh  # noqa: E402
from oracle import scenes  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def fix_all_goal(st):
    """ENV:245 asserts that not every robot has reached its goal before a step."""
    E, RP = st["x"].shape
    for e in range(E):
        nr = st["n_robots"][e]
        if nr > 0 and all(st["flags"][e, i] & 2 for i in range(nr)):
            st["flags"][e, 0] = 0


def apply_overrides(params, ov):
    for k, v in ov.items():
        if isinstance(v, list):
            arr = getattr(params, k)
            for i, x in enumerate(v):
                arr[i] = x
        else:
            setattr(params, k, v)
    return params


def save(name, st0, actions, noise, mode, OP, CP, st1, out, overrides, extra=None):
    d = {"mode": np.int64(mode), "OP": np.int64(OP), "CP": np.int64(CP),
         "overrides": np.array(json.dumps(overrides))}
    for k, v in st0.items():
        d["in_" + k] = v
    for k, v in st1.items():
        d["ref_" + k] = v
    for k, v in out.items():
        d["out_" + k] = v
    if actions is not None:
        d["actions"] = actions
    d["noise"] = noise.astype(np.float32)
    if extra:
        d.update(extra)
    np.savez_compressed(os.path.join(GOLD, name), **d)
    print("wrote", name, {k: v.shape for k, v in d.items() if hasattr(v, "shape") and v.ndim > 0 and k.startswith("out_")})


def one_step_case(name, seed, E, RP, OP, CP, mode, overrides=None, **scene_kw):
    overrides = overrides or {}
    params = apply_overrides(_abi.default_params(), overrides)
    st0 = scenes.random_scene(seed, E, RP, OP, CP, **scene_kw)
    fix_all_goal(st0)
    actions = scenes.random_actions(seed, E, RP, bool(mode & _abi.M_DISCRETE)) if mode & 1 else None
    noise = scenes.random_noise(seed, E, RP, OP, params)
    st1 = scenes.copy_state(st0)
    out = rh.ref_step(params, st1, actions, noise, mode, OP,
                      auto_deactivate=bool(mode & _abi.M_AUTO_DEACTIVATE))
    save(name, st0, actions, noise, mode, OP, CP, st1, out, overrides)


def rollout_case(name, seed, E, RP, OP, CP, T):
    """T consecutive steps with the Trainer's deactivation (trainer.py:168-170) folded in."""
    params = _abi.default_params()
    mode = _abi.M_DYNAMICS | _abi.M_AUTO_DEACTIVATE
    st = scenes.random_scene(seed, E, RP, OP, CP, box=40.0, p_deact=0.0, p_timeout=0.0)
    st["ep_steps"][:] = np.array([990 + (e % 3) * 4 for e in range(E)], np.int32)
    st["flags"][:] = 0
    st0 = scenes.copy_state(st)
    acts, noises, states, outs = [], [], [], []
    alive = np.ones(E, bool)
    for t in range(T):
        a = scenes.random_actions(seed + 31 * t, E, RP, False)
        n = scenes.random_noise(seed + 17 * t, E, RP, OP, params)
        # ENV:245: stop stepping an env once every robot reached its goal (episode is over)
        for e in range(E):
            nr = st["n_robots"][e]
            if all(st["flags"][e, i] & 2 for i in range(nr)):
                alive[e] = False
        sub = {k: v[alive].copy() for k, v in st.items()}
        out = rh.ref_step(params, sub, a[alive], n[alive], mode, OP, auto_deactivate=True)
        for k in st:
            st[k][alive] = sub[k]
        full = {k: np.zeros((E,) + v.shape[1:], v.dtype) for k, v in out.items()}
        for k in out:
            full[k][alive] = out[k]
        acts.append(a), noises.append(n.astype(np.float32))
        states.append(scenes.copy_state(st)), outs.append(full)
        acts[-1] = a
        outs[-1]["alive"] = alive.copy()
    d = {"mode": np.int64(mode), "OP": np.int64(OP), "CP": np.int64(CP), "T": np.int64(T),
         "overrides": np.array(json.dumps({})), "actions": np.stack(acts), "noise": np.stack(noises)}
    for k, v in st0.items():
        d["in_" + k] = v
    for k in states[0]:
        d["ref_" + k] = np.stack([s[k] for s in states])
    for k in outs[0]:
        d["out_" + k] = np.stack([o[k] for o in outs])
    np.savez_compressed(os.path.join(GOLD, name), **d)
    print("wrote", name, "alive at end:", int(alive.sum()), "flags:", np.unique(st["flags"]))


def current_case(name, seed):
    ref_env, _ = rh.import_reference()
    import scipy.spatial
    rng = np.random.RandomState(seed)
    cores = np.zeros((6, 8, 4))
    pts = scenes.f32(55.0 * rng.rand(6, 40, 2))
    vel = np.zeros((6, 40, 2))
    ncs = [1, 2, 3, 5, 8, 8]
    for g, nc in enumerate(ncs):
        cores[g, :nc, 0:2] = scenes.f32(55.0 * rng.rand(nc, 2))
        cores[g, :nc, 2] = np.where(rng.rand(nc) < 0.5, 1.0, -1.0)
        cores[g, :nc, 3] = scenes.f32(2 * np.pi * 0.5 * (2 + 2 * rng.rand(nc)))
        # put a few query points inside a core (solid-body branch, ENV:498-499)
        pts[g, :4] = scenes.f32(cores[g, 0, 0:2] + 0.3 * (rng.rand(4, 2) - 0.5))
        env = ref_env.MarineNavEnv3(seed=0)
        env.cores = [ref_env.Core(c[0], c[1], bool(c[2] > 0), c[3]) for c in cores[g, :nc]]
        env.core_centers = scipy.spatial.KDTree(cores[g, :nc, 0:2])
        for k in range(40):
            vel[g, k] = env.get_velocity(pts[g, k, 0], pts[g, k, 1])[:2]
    np.savez_compressed(os.path.join(GOLD, name), cores=cores, n_cores=np.array(ncs), points=pts,
                        velocity=vel)
    print("wrote", name)


def iqn_case(name, seed):
    """AGENT:401-412 through the reference's own calculate_huber_loss (hard-wired to N=N'=8),
    gradient by autograd in float64; plus Critic.calc_cos (ACIQN:410-426) on injected tau."""
    rh.import_reference()
    import torch
    from rfarl.agent import calculate_huber_loss
    from rfarl.policy.AC_IQN_model import Critic
    g = torch.Generator().manual_seed(seed)
    B, N = 64, 8
    z = (3 * torch.randn(B, N, generator=g)).float().double().requires_grad_(True)
    zn = (3 * torch.randn(B, N, generator=g)).float().double()
    tau = torch.rand(B, N, generator=g).float().double()
    r = torch.randn(B, 1, generator=g).float().double()
    done = (torch.rand(B, 1, generator=g) < 0.1).double()
    gamma = 0.99
    q_t = r.unsqueeze(-1) + gamma * zn.unsqueeze(1) * (1.0 - done.unsqueeze(-1))   # AGENT:401
    td = q_t - z.unsqueeze(-1)                                                      # AGENT:403-406
    hub = calculate_huber_loss(td, 1.0)                                             # AGENT:408
    ql = abs(tau.unsqueeze(-1) - (td.detach() < 0).float()) * hub / 1.0             # AGENT:409
    lps = ql.sum(dim=1).mean(dim=1)                                                 # AGENT:411
    loss = lps.mean()                                                               # AGENT:412
    loss.backward()
    critic = Critic(7, 5, 5, 56, 40, 256, 128, 2, "cpu", 101)
    taus32 = tau.float().unsqueeze(-1)
    cos = torch.cos(taus32 * critic.pis)                                            # ACIQN:424
    np.savez_compressed(os.path.join(GOLD, name), z=z.detach().numpy(), z_next=zn.numpy(),
                        tau=tau.numpy(), r=r.numpy()[:, 0], done=done.numpy()[:, 0],
                        gamma=gamma, loss=loss.item(), loss_per_sample=lps.detach().numpy(),
                        grad_z=z.grad.numpy(), cos=cos.numpy(), pis=critic.pis.numpy().reshape(-1))
    print("wrote", name, "loss", loss.item())


def main():
    os.makedirs(GOLD, exist_ok=True)
    D, X, A = _abi.M_DYNAMICS, _abi.M_DISCRETE, _abi.M_AUTO_DEACTIVATE
    one_step_case("env_step_cont.npz", 11, 40, 8, 10, 8, D)
    one_step_case("env_step_disc.npz", 12, 16, 6, 4, 0, D | X | A)
    one_step_case("env_step_c2shape.npz", 13, 12, 8, 10, 0, D | A, ragged=False, p_timeout=0.0)
    one_step_case("env_step_hull.npz", 14, 8, 5, 3, 4, D, overrides={
        "yDotV": -30.0, "yDotR": 12.0, "nDotV": 9.0, "yR": -25.0, "yRV": -14.0, "yVR": -11.0,
        "yRR": -6.0, "nV": -18.0, "nVV": -7.0, "nRV": -5.0, "nVR": -4.0, "left_pos": 0.12,
        "right_pos": -0.07, "goal_dis": 3.0, "perc_range": 15.0, "r_mean_ratio": 0.7,
        "colregs_penalty": -0.2, "thrust_change_l": [0.0, -400.0, -800.0, 400.0, 800.0]})
    one_step_case("env_perception_only.npz", 15, 16, 8, 10, 0, 0)
    rollout_case("env_rollout.npz", 16, 8, 6, 6, 4, 24)
    current_case("current_field.npz", 17)
    iqn_case("iqn_loss.npz", 18)


if __name__ == "__main__":
    main()
