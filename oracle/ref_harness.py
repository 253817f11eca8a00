"""Drives the UNMODIFIED Python reference (imported from /root/reference) on injected state
and injected sensor noise.  TEST INFRASTRUCTURE ONLY; usable only in the build container —
/root/reference does not exist on the GPU box, so nothing in `-m gpu` tests, smoke() or
bench.py imports this module.  gen_golden.py uses it to write tests/golden/*.npz.

State injector: builds rfarl.envs.marinenav.env.MarineNavEnv3 (env.py:24-67) and overwrites
`robots`, `obstacles`, `cores`, `core_centers`, `episode_timesteps` from arrays.
Noise injector: replaces every `rob.perception.rd` (wamv.py:9) by a table-replay object with
`.normal(mu, std)` / `.vonmises(mu, kappa)`, consumed in the reference's own draw order
(wamv.py:27-40, 465-510): per candidate px, py, vx, vy normals then one von Mises.
"""
import os
import sys
import warnings

import numpy as np

REF_ROOT = os.environ.get("ASVRL_REFERENCE", "/root/reference")


def import_reference():
    p = os.path.join(REF_ROOT, "rfarl")
    if not os.path.isdir(p):
        raise RuntimeError(f"reference not present at {p}")
    if p not in sys.path:
        sys.path.insert(0, p)
    warnings.filterwarnings("ignore")
    import rfarl.envs.marinenav.env as ref_env  # noqa
    import rfarl.envs.marinenav.vehicles.wamv as ref_wamv  # noqa
    return ref_env, ref_wamv


class TableRNG:
    """Replays a flat list of pre-drawn noise values; normals are stored already scaled."""

    def __init__(self):
        self.values = []
        self.pos = 0

    def load(self, values):
        self.values = list(values)
        self.pos = 0

    def _next(self):
        v = self.values[self.pos]
        self.pos += 1
        return float(v)

    def normal(self, mu, std):
        return mu + self._next()

    def vonmises(self, mu, kappa):
        return mu + self._next()


PARAM_TO_ROBOT = {  # asvrl_env_params field -> Robot attribute (wamv.py:45-126)
    "dt": "dt", "n_substeps": "N", "length": "length", "width": "width", "robot_r": "r",
    "goal_dis": "goal_dis", "min_thrust": "min_thrust", "max_thrust": "max_thrust",
    "m": "m", "Izz": "Izz", "xDotU": "xDotU", "yDotV": "yDotV", "yDotR": "yDotR", "nDotR": "nDotR",
    "nDotV": "nDotV", "xU": "xU", "xUU": "xUU", "yV": "yV", "yVV": "yVV", "yR": "yR", "yRV": "yRV",
    "yVR": "yVR", "yRR": "yRR", "nR": "nR", "nRR": "nRR", "nV": "nV", "nVV": "nVV", "nRV": "nRV",
    "nVR": "nVR",
}


def build_env(params, st, e, OP, is_eval_env=False):
    """Reference env holding env `e` of the float64 state dict `st` (layout of oracle.new_state)."""
    ref_env, ref_wamv = import_reference()
    env = ref_env.MarineNavEnv3(seed=0, is_eval_env=is_eval_env)
    env.r = params.core_r
    env.timestep_penalty = params.timestep_penalty
    env.COLREGs_penalty = params.colregs_penalty
    env.collision_penalty = params.collision_penalty
    env.goal_reward = params.goal_reward
    nr, no, nc = int(st["n_robots"][e]), int(st["n_obs"][e]), int(st["n_cores"][e])
    env.robots = []
    for i in range(nr):
        rob = ref_wamv.Robot(seed=0)
        for pf, rf in PARAM_TO_ROBOT.items():
            setattr(rob, rf, getattr(params, pf))
        rob.detect_r = rob.r
        rob.compute_constant_matrices()
        rob.left_thrust_change = np.array(list(params.thrust_change_l))
        rob.right_thrust_change = np.array(list(params.thrust_change_r))
        rob.compute_actions()
        rob.perception.range = params.perc_range
        rob.perception.angle = params.perc_angle
        rob.perception.pos_std = params.pos_std
        rob.perception.vel_std = params.vel_std
        rob.perception.r_kappa = params.r_kappa
        rob.perception.r_mean_ratio = params.r_mean_ratio
        rob.perception.rd = TableRNG()
        rob.x, rob.y, rob.theta = float(st["x"][e, i]), float(st["y"][e, i]), float(st["theta"][e, i])
        rob.velocity_r = np.array([st["velr_x"][e, i], st["velr_y"][e, i], st["velr_r"][e, i]], float)
        rob.velocity = np.array([st["vel_x"][e, i], st["vel_y"][e, i], st["vel_r"][e, i]], float)
        rob.left_pos, rob.right_pos = params.left_pos, params.right_pos
        rob.left_thrust, rob.right_thrust = float(st["thrust_l"][e, i]), float(st["thrust_r"][e, i])
        rob.goal = np.array([st["goal_x"][e, i], st["goal_y"][e, i]], float)
        rob.start = np.array([rob.x, rob.y])
        f = int(st["flags"][e, i])
        rob.collision, rob.reach_goal, rob.deactivated = bool(f & 1), bool(f & 2), bool(f & 4)
        env.robots.append(rob)
    env.num_robots = nr
    env.obstacles = [ref_env.Obstacle(*map(float, st["obstacles"][e, k, :3])) for k in range(no)]
    env.cores = [ref_env.Core(float(c[0]), float(c[1]), bool(c[2] > 0), float(c[3]))
                 for c in st["cores"][e, :nc]]
    if nc > 0:
        import scipy.spatial
        env.core_centers = scipy.spatial.KDTree(np.array([[c.x, c.y] for c in env.cores]))
    env.episode_timesteps = int(st["ep_steps"][e])
    return env


def load_noise(env, noise_e, OP):
    """noise_e [RP][OP+RP][5] -> each robot's TableRNG in the reference's draw order."""
    for i, rob in enumerate(env.robots):
        seq = []
        if not rob.deactivated:
            for k in range(len(env.obstacles)):
                seq.extend(noise_e[i, k])
            for j, other in enumerate(env.robots):
                if other is rob or other.deactivated:
                    continue
                seq.extend(noise_e[i, OP + j])
        rob.perception.rd.load(seq)


INFO_CODE = {"normal": 0, "collision": 1, "reach goal": 2, "too long episode": 3,
             "deactivated after collision": 4, "deactivated after reaching goal": 5}


def write_back(env, st, e):
    for i, rob in enumerate(env.robots):
        st["x"][e, i], st["y"][e, i], st["theta"][e, i] = rob.x, rob.y, rob.theta
        st["velr_x"][e, i], st["velr_y"][e, i], st["velr_r"][e, i] = rob.velocity_r
        st["vel_x"][e, i], st["vel_y"][e, i], st["vel_r"][e, i] = rob.velocity
        st["thrust_l"][e, i], st["thrust_r"][e, i] = rob.left_thrust, rob.right_thrust
        st["flags"][e, i] = (1 if rob.collision else 0) | (2 if rob.reach_goal else 0) | \
                            (4 if rob.deactivated else 0)
    st["ep_steps"][e] = env.episode_timesteps


def collect(env, st, out, e, OP, observations, rewards=None, dones=None, infos=None,
            was_deactivated=None):
    """Fill row `e` of the out dict (layout of oracle.new_out) from the reference objects."""
    RP = st["x"].shape[1]
    nr = len(env.robots)
    for i in range(RP):
        if i >= nr:
            out["status"][e, i] = 7 << 8
            out["top_ids"][e, i] = -1
            continue
        rob = env.robots[i]
        status = 0
        if rob.collision:
            status |= 2
        if rob.reach_goal:
            status |= 4
        wd = was_deactivated[i] if was_deactivated is not None else rob.deactivated
        out["top_ids"][e, i] = -1
        if wd:
            status |= 8
            if dones is None:  # perception-only call: the batched API still reports done + why
                status |= 1 | ((4 if rob.collision else 5) << 8)
        else:
            self_obs, objs = observations[i]
            out["obs_self"][e, i] = self_obs
            for k, o in enumerate(objs):
                out["obs_objects"][e, i, k] = o
                out["obs_mask"][e, i, k] = 1.0
            slots = list(rob.perception.observed_obs) + [OP + j for j in rob.perception.observed_objs]
            det = 0
            for s in slots:
                det |= 1 << s
            out["detected"][e, i] = det
            full = rob.perception.observation["objects"]
            assert len(full) == len(slots)
            keys = [rob.compute_distance(o[0], o[1], o[4], True) for o in full]
            order = sorted(range(len(full)), key=lambda k: keys[k])  # stable, like heapq.nsmallest
            for k, idx in enumerate(order[:5]):
                out["top_ids"][e, i, k] = slots[idx]
                assert list(full[idx]) == list(objs[k])
            if rob.apply_COLREGs:
                status |= 16
                out["phi"][e, i] = rob.phi
        if dones is not None:
            if dones[i]:
                status |= 1
            status |= INFO_CODE[infos[i]["state"]] << 8
            out["reward"][e, i] = rewards[i]
        out["status"][e, i] = status
    es = 0
    if env.check_all_deactivated():
        es |= 1
    if env.check_any_collision():
        es |= 4
    if env.check_all_reach_goal():
        es |= 8
    out["env_status"][e] = es


def ref_step(params, st, actions, noise, mode, OP, auto_deactivate=False):
    """Reference equivalent of oracle.env_step over every env of `st` (in place)."""
    from . import oracle as _o  # only for new_out's layout
    E, RP = st["x"].shape
    out = _o.new_out(E, RP)
    out.pop("margin_perc"), out.pop("margin_colregs")
    discrete = bool(mode & 2)
    for e in range(E):
        env = build_env(params, st, e, OP)
        load_noise(env, noise[e], OP)
        was = [rob.deactivated for rob in env.robots]
        ep_before = env.episode_timesteps
        if mode & 1:
            acts = []
            for i, rob in enumerate(env.robots):
                if rob.deactivated:
                    acts.append(None)
                elif discrete:
                    acts.append(int(actions[e, i]))
                else:
                    acts.append([float(actions[e, i, 0]), float(actions[e, i, 1])])
            obs, rewards, dones, infos = env.step(acts, not discrete)
            if auto_deactivate:  # trainer.py:168-170
                for rob in env.robots:
                    if rob.collision or rob.reach_goal:
                        rob.deactivated = True
            collect(env, st, out, e, OP, obs, rewards, dones, infos, was)
        else:
            obs, _, _ = env.get_observations()
            collect(env, st, out, e, OP, obs, None, None, None, was)
        if ep_before >= params.max_episode_steps:
            out["env_status"][e] |= 2
        for rob in env.robots:  # every injected draw must have been consumed, and no more
            assert rob.perception.rd.pos == len(rob.perception.rd.values)
        write_back(env, st, e)
    return out
