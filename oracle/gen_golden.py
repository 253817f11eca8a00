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


def reset_checks_case(name, seed, n_cases=400):
    """The reference's own acceptance predicates (ENV:358-456) on random accepted sets and
    candidates (many deliberately close to a threshold): inputs + the reference's answers."""
    ref_env, ref_wamv = rh.import_reference()
    rng = np.random.RandomState(seed)
    K = 6
    rec = {k: [] for k in ("n_r", "n_c", "n_o", "sg", "cores", "obs", "min_dis", "c_sg", "c_core",
                           "c_obs", "a_sg", "a_core", "a_obs")}
    for case in range(n_cases):
        env = ref_env.MarineNavEnv3(seed=0)
        n_r, n_c, n_o = rng.randint(0, K + 1), rng.randint(0, K + 1), rng.randint(0, K + 1)
        sg = np.zeros((K, 4)); cores = np.zeros((K, 4)); obs = np.zeros((K, 4))
        sg[:n_r] = 2.0 + 51.0 * rng.rand(n_r, 4)
        cores[:n_c, 0:2] = 55.0 * rng.rand(n_c, 2)
        cores[:n_c, 2] = np.where(rng.rand(n_c) < 0.5, 1.0, -1.0)
        cores[:n_c, 3] = 2 * np.pi * 0.5 * (2.0 + 2.0 * rng.rand(n_c))
        obs[:n_o, 0:2] = 5.0 + 45.0 * rng.rand(n_o, 2)
        obs[:n_o, 2] = 0.5 + rng.rand(n_o)
        env.robots = []
        for i in range(n_r):
            rob = ref_wamv.Robot(seed=0)
            rob.start, rob.goal = sg[i, 0:2].copy(), sg[i, 2:4].copy()
            env.robots.append(rob)
        env.cores = [ref_env.Core(c[0], c[1], bool(c[2] > 0), c[3]) for c in cores[:n_c]]
        env.obstacles = [ref_env.Obstacle(o[0], o[1], o[2]) for o in obs[:n_o]]
        env.min_start_goal_dis = float(rng.choice([20.0, 30.0, 40.0]))

        def near(points, radius):  # a candidate at ~radius from a random anchor, or uniform
            if len(points) and rng.rand() < 0.6:
                a = points[rng.randint(len(points))]
                ang = 2 * np.pi * rng.rand()
                d = radius * (1.0 + 0.05 * rng.randn())
                return a + d * np.array([np.cos(ang), np.sin(ang)])
            return -1.0 + 57.0 * rng.rand(2)

        start = near(sg[:n_r, 0:2], 10.0)
        ang = 2 * np.pi * rng.rand()
        goal = start + env.min_start_goal_dis * (1.0 + 0.1 * rng.randn()) * np.array([np.cos(ang), np.sin(ang)])
        if rng.rand() < 0.4:
            goal = near(sg[:n_r, 2:4], 10.0)
        c_sg = np.concatenate([start, goal])
        a_sg = env.check_start_and_goal(start, goal)
        anchors = np.concatenate([sg[:n_r, 0:2], sg[:n_r, 2:4], cores[:n_c, 0:2]]) if n_r + n_c else np.zeros((0, 2))
        cc = near(anchors, rng.choice([10.5, 3.0, 1.9, 3.2]))
        c_core = np.array([cc[0], cc[1], 1.0 if rng.rand() < 0.5 else -1.0, 2 * np.pi * 0.5 * (2.0 + 2.0 * rng.rand())])
        a_core = env.check_core(ref_env.Core(c_core[0], c_core[1], bool(c_core[2] > 0), c_core[3]))
        anchors = np.concatenate([anchors, obs[:n_o, 0:2]]) if len(anchors) + n_o else np.zeros((0, 2))
        oc = near(anchors, rng.choice([11.0, 1.5, 2.0]))
        c_obs = np.array([oc[0], oc[1], 0.5 + rng.rand()])
        a_obs = env.check_obstacle(ref_env.Obstacle(c_obs[0], c_obs[1], c_obs[2]))
        for k, v in (("n_r", n_r), ("n_c", n_c), ("n_o", n_o), ("sg", sg), ("cores", cores), ("obs", obs),
                     ("min_dis", env.min_start_goal_dis), ("c_sg", c_sg), ("c_core", c_core),
                     ("c_obs", c_obs), ("a_sg", a_sg), ("a_core", a_core), ("a_obs", a_obs)):
            rec[k].append(v)
    d = {k: np.array(v) for k, v in rec.items()}
    np.savez_compressed(os.path.join(GOLD, name), **d)
    print("wrote", name, "accept rates", d["a_sg"].mean(), d["a_core"].mean(), d["a_obs"].mean())


class PhiloxReplayRNG:
    """Stands in for MarineNavEnv3.rd (ENV:29) and feeds reset() the counter-indexed candidate
    sequence of include/asvrl.h (stream 2), so the UNMODIFIED reference reset() generates the
    scenario the kernel / oracle must reproduce.  Call sites: ENV:58/113 randint (robot seed,
    ignored), ENV:109-110 start / goal, ENV:126-128 core, ENV:154-155 obstacle, ENV:171 theta."""

    def __init__(self, seed, episode, env):
        from oracle import oracle as o
        self.o, self.seed, self.episode, self.env = o, seed, episode, env
        self.t = {0: 0, 1: 0, 2: 0}
        self.theta_i = 0
        self.words = None
        self.pending = None

    def _block(self, cls, t, theta=False):
        slot, block = (t, 0) if theta else (t >> 8, t & 255)
        ctr = [block | (slot << 8) | (cls << 16), self.env & 0xFFFFFFFF, self.episode & 0xFFFFFFFF,
               ((self.episode >> 32) & 0xFFFFFF) | (2 << 24)]
        return self.o.philox(ctr, [self.seed & 0xFFFFFFFF, self.seed >> 32])

    @staticmethod
    def _u(w):
        return ((int(w) >> 8) + 0.5) / 16777216.0

    def randint(self, lo, hi):
        return 0

    def binomial(self, n, p):
        return int(self.words[2]) >> 31                      # core direction, ENV:127

    def uniform(self, low=0.0, high=1.0):
        if np.ndim(low) == 0:
            if float(low) == 0.0 and abs(float(high) - 2 * np.pi) < 1e-12:   # ENV:171
                w = self._block(3, self.theta_i, theta=True)
                self.theta_i += 1
                return low + (high - low) * self._u(w[0])
            # scalar draws that complete a candidate: ENV:128 (v_edge) or ENV:155 (radius)
            idx = 3 if self.pending == 1 else 2
            return low + (high - low) * self._u(self.words[idx])
        low, high = np.asarray(low, float), np.asarray(high, float)
        if low[0] == 2.0:                                    # start, then goal: one candidate
            if self.pending == "goal":
                self.pending = None
                return low + (high - low) * np.array([self._u(self.words[2]), self._u(self.words[3])])
            self.words = self._block(0, self.t[0]); self.t[0] += 1
            self.pending = "goal"
            return low + (high - low) * np.array([self._u(self.words[0]), self._u(self.words[1])])
        cls = 1 if low[0] == 0.0 else 2                      # core centre / obstacle centre
        self.words = self._block(cls, self.t[cls]); self.t[cls] += 1
        self.pending = cls
        return low + (high - low) * np.array([self._u(self.words[0]), self._u(self.words[1])])


def reset_case(name, seed, E):
    """The reference's reset() (ENV:72-176) driven by the Philox candidate stream."""
    ref_env, _ = rh.import_reference()
    import contextlib, io
    rng = np.random.RandomState(seed)
    RP, OP, CP = 8, 12, 8
    episode, env_offset = (3 << 32) + 11, 700
    want = np.stack([rng.randint(1, RP + 1, E), rng.randint(0, CP + 1, E), rng.randint(0, OP + 1, E)], 1)
    want[0] = (8, 8, 12)    # crowded: the 500-try cap leaves it short
    want[1] = (1, 0, 0)
    min_sg = rng.choice([20.0, 30.0, 40.0], E)
    from oracle import oracle as o
    st = o.new_state(E, RP, OP, CP)
    for e in range(E):
        env = ref_env.MarineNavEnv3(seed=0)
        env.rd = PhiloxReplayRNG(0xABCDEF0123456789 + seed, episode, env_offset + e)
        env.num_robots, env.num_cores, env.num_obs = (int(v) for v in want[e])
        env.min_start_goal_dis = float(min_sg[e])
        with contextlib.redirect_stdout(io.StringIO()):
            env.reset()
        st["n_robots"][e], st["n_cores"][e], st["n_obs"][e] = len(env.robots), len(env.cores), len(env.obstacles)
        for i, rob in enumerate(env.robots):
            st["x"][e, i], st["y"][e, i], st["theta"][e, i] = rob.x, rob.y, rob.theta
            st["goal_x"][e, i], st["goal_y"][e, i] = rob.goal
            assert not rob.velocity.any() and not rob.velocity_r.any() and rob.left_thrust == 0.0
        for i, c in enumerate(env.cores):
            st["cores"][e, i] = (c.x, c.y, 1.0 if c.clockwise else -1.0, c.Gamma)
        for i, ob in enumerate(env.obstacles):
            st["obstacles"][e, i] = (ob.x, ob.y, ob.r, 0.0)
    d = {"ref_" + k: v for k, v in st.items()}
    d.update(want=want.astype(np.int32), min_sg=min_sg, seed=np.uint64(0xABCDEF0123456789 + seed),
             episode=np.uint64(episode), env_offset=np.int64(env_offset), RP=RP, OP=OP, CP=CP)
    np.savez_compressed(os.path.join(GOLD, name), **d)
    print("wrote", name, "placed", st["n_robots"].sum(), st["n_cores"].sum(), st["n_obs"].sum(), "of", want.sum(0))


def replay_case(name, seed):
    """The reference's ReplayBuffer (RB:6-69): add 40 transitions into a maxlen-32 deque, then
    state_batch + the deque's content in order (what a ring of capacity 32 must hold)."""
    rh.import_reference()
    from rfarl.utils.replay_buffer import ReplayBuffer
    rng = np.random.RandomState(seed)
    rb = ReplayBuffer(32, 8, 5, 5)
    n = 40
    n_obj = rng.randint(0, 6, n)
    n_obj2 = rng.randint(0, 6, n)
    self_s = scenes.f32(rng.randn(n, 7)); self_n = scenes.f32(rng.randn(n, 7))
    obj_s = scenes.f32(rng.randn(n, 5, 5)); obj_n = scenes.f32(rng.randn(n, 5, 5))
    act = scenes.f32(rng.rand(n, 2) * 2 - 1); rew = scenes.f32(rng.randn(n)); done = rng.rand(n) < 0.2
    for i in range(n):
        s = (list(self_s[i]), [list(o) for o in obj_s[i, :n_obj[i]]])
        s2 = (list(self_n[i]), [list(o) for o in obj_n[i, :n_obj2[i]]])
        rb.add((s, list(act[i]), float(rew[i]), s2, bool(done[i])))
    mem = list(rb.memory)
    st = rb.state_batch([m[0] for m in mem])
    nst = rb.state_batch([m[3] for m in mem])
    np.savez_compressed(os.path.join(GOLD, name), n_obj=n_obj, n_obj2=n_obj2, self_s=self_s, self_n=self_n,
                        obj_s=obj_s, obj_n=obj_n, act=act, rew=rew, done=done,
                        mem_self=np.array(st[0]), mem_obj=np.array(st[1]), mem_mask=np.array(st[2]),
                        mem_nself=np.array(nst[0]), mem_nobj=np.array(nst[1]), mem_nmask=np.array(nst[2]),
                        mem_act=np.array([m[1] for m in mem]), mem_rew=np.array([m[2] for m in mem]),
                        mem_done=np.array([m[4] for m in mem]))
    print("wrote", name, np.array(st[1]).shape)


def agent_case(name, seed):
    """The reference Agent('AC-IQN') (AGENT:19-205, 386-432): initial weights for the seed, actor and
    critic forwards, and ONE train_AC_IQN step on an injected batch with injected taus (memory.sample
    and the three calc_cos draws are patched; everything else is the reference's own code)."""
    rh.import_reference()
    import torch
    from rfarl.agent import Agent
    torch.set_num_threads(1)
    rng = np.random.RandomState(seed)
    B, N = 64, 8
    n_obj, n_obj2 = rng.randint(0, 6, B), rng.randint(0, 6, B)
    t = np.arange(5)[None, :]
    mask, nmask = (t < n_obj[:, None]).astype(np.float32), (t < n_obj2[:, None]).astype(np.float32)
    scale_s = np.array([20, 20, 2, 2, 0.5, 500, 500], np.float32)
    scale_o = np.array([15, 15, 1, 1, 1], np.float32)
    batch = {"self": (rng.randn(B, 7) * scale_s).astype(np.float32),
             "obj": ((rng.randn(B, 5, 5) * scale_o) * mask[:, :, None]).astype(np.float32), "mask": mask,
             "nself": (rng.randn(B, 7) * scale_s).astype(np.float32),
             "nobj": ((rng.randn(B, 5, 5) * scale_o) * nmask[:, :, None]).astype(np.float32), "nmask": nmask,
             "act": (rng.rand(B, 2) * 2 - 1).astype(np.float32), "rew": rng.randn(B).astype(np.float32),
             "done": (rng.rand(B) < 0.1).astype(np.float32)}
    taus = {k: rng.rand(B, N).astype(np.float32) for k in ("target", "local", "actor")}
    agent = Agent(device="cpu", seed=seed, agent_type="AC-IQN")
    init = {"init_actor." + k: v.clone().numpy() for k, v in agent.policy_local.actor.state_dict().items()}
    init.update({"init_critic." + k: v.clone().numpy() for k, v in agent.policy_local.critic.state_dict().items()})
    tt = lambda a: torch.from_numpy(a)
    states = (tt(batch["self"]), tt(batch["obj"]), tt(batch["mask"]))
    with torch.no_grad():
        actor_out = agent.policy_local.actor(states).numpy()

    queue = []

    def patched_calc_cos(self, batch_size, num_tau=8, cvar=1.0):
        tau = tt(queue.pop(0)).unsqueeze(-1)
        return torch.cos(tau * self.pis), tau

    from rfarl.policy.AC_IQN_model import Critic
    orig = Critic.calc_cos
    Critic.calc_cos = patched_calc_cos
    try:
        queue.append(taus["local"])
        with torch.no_grad():
            critic_out = agent.policy_local.critic(states, tt(batch["act"]))[0].numpy()
        as_list = lambda a: a.tolist()
        agent.memory.sample = lambda: ((as_list(batch["self"]), as_list(batch["obj"]), as_list(batch["mask"])),
                                       as_list(batch["act"]), as_list(batch["rew"]),
                                       (as_list(batch["nself"]), as_list(batch["nobj"]), as_list(batch["nmask"])),
                                       [bool(d) for d in batch["done"]])
        queue.extend([taus["target"], taus["local"], taus["actor"]])   # draw order in AGENT:399, 403, 422
        critic_loss, actor_loss = agent.train_AC_IQN()
        assert not queue
    finally:
        Critic.calc_cos = orig
    post = {"post_actor." + k: v.clone().numpy() for k, v in agent.policy_local.actor.state_dict().items()}
    post.update({"post_critic." + k: v.clone().numpy() for k, v in agent.policy_local.critic.state_dict().items()})
    d = {"seed": np.int64(seed), "actor_out": actor_out, "critic_out": critic_out,
         "critic_loss": np.float64(critic_loss), "actor_loss": np.float64(actor_loss)}
    d.update({"b_" + k: v for k, v in batch.items()})
    d.update({"tau_" + k: v for k, v in taus.items()})
    d.update(init)
    d.update(post)
    np.savez_compressed(os.path.join(GOLD, name), **d)
    print("wrote", name, "critic_loss", float(critic_loss), "actor_loss", float(actor_loss))


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
    reset_checks_case("reset_checks.npz", 19)
    reset_case("env_reset.npz", 20, 48)
    replay_case("replay.npz", 21)
    agent_case("agent_ac_iqn.npz", 100)


if __name__ == "__main__":
    main()
