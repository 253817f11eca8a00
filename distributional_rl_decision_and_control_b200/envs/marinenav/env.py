"""``MarineNavEnv3`` with the reference's interface (rfarl/rfarl/envs/marinenav/env.py:24-778)
on top of the batched CUDA kernels, for ONE environment: the drop-in used by the unchanged
``Trainer`` / ``train_RL_agents.py`` / ``run_vrx_experiments.py`` call sites.

reset() / step() / get_observations() launch K1b / K1 (include/asvrl.h) on a 1-env state and
copy the small result back; ``robots[i]`` are host mirrors of the device state with the
attributes the callers read (``collision``, ``reach_goal``, ``x`` …) and the one they write
(``deactivated``, TRAINER:168-170), which is pushed to the device before the next launch.
Thousands of environments per launch are the business of ``batched.BatchedMarineNavEnv``.

Differences by construction: scenario generation and sensor noise draw from the counter-based
Philox streams of include/asvrl.h instead of numpy's Mersenne Twister (same distributions, same
acceptance rules), and state is float32.
"""
import copy
import json

import numpy as np
import torch

from ... import _abi, _lib
from .vehicles.wamv import Robot


class Core:
    """Vortex core (ENV:7-14)."""

    def __init__(self, x: float, y: float, clockwise: bool, Gamma: float):
        self.x, self.y, self.clockwise, self.Gamma = x, y, clockwise, Gamma


class Obstacle:
    """Static circular obstacle (ENV:16-22)."""

    def __init__(self, x: float, y: float, r: float):
        self.x, self.y, self.r = x, y, r


class MarineNavEnv3():

    def __init__(self, seed: int = 0, schedule: dict = None, is_eval_env: bool = False,
                 device="cuda:0"):
        _lib.lib()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("MarineNavEnv3 runs on a CUDA device (no CPU fallback)")
        self.seed = seed
        self.is_eval_env = is_eval_env
        # ENV:33-54
        self.width = 55
        self.height = 55
        self.r = 0.5
        self.v_rel_max = 1.0
        self.p = 0.8
        self.v_range = [3, 3]
        self.obs_r_range = [1, 1]
        self.clear_r = 10.0
        self.angular_speed_max = np.pi / 2
        self.angular_speed_penalty = -1.0
        self.steering_reward_angle_max = np.pi / 3
        self.steering_reward_speed_min = 1.0
        self.steering_reward = 0.5
        self.timestep_penalty = -0.1
        self.COLREGs_penalty = -0.1
        self.collision_penalty = -5.0
        self.goal_reward = 10.0
        self.num_cores = 8
        self.num_obs = 8
        self.min_start_goal_dis = 30.0
        self.num_robots = 6

        self.robots = [Robot(seed=0) for _ in range(self.num_robots)]
        self.cores = []
        self.obstacles = []

        self.schedule = schedule
        self.episode_timesteps = 0
        self.total_timesteps = 0
        self.observation_in_robot_frame = True

        self._episodes = 0      # reset() calls: the scenario-stream counter
        self._obs_calls = 0     # perception launches: the sensor-noise counter
        self._shape = None
        self._alloc(8, 8, 8)

    # ------------------------------------------------------------------ device plumbing
    def _alloc(self, RP, OP, CP):
        """(Re)allocate the 1-env device state for at most RP robots, OP buoys, CP cores."""
        RP, OP, CP = max(RP, 1), max(OP, 0), max(CP, 0)
        if self._shape is not None and RP <= self._shape[0] and OP <= self._shape[1] and CP <= self._shape[2]:
            return
        if self._shape is not None:
            RP, OP, CP = max(RP, self._shape[0]), max(OP, self._shape[1]), max(CP, self._shape[2])
        if RP > 32 or OP + RP > _abi.MAX_SLOTS:
            raise RuntimeError("at most 32 robots and 64 robots + obstacles per environment")
        self._shape = (RP, OP, CP)
        with torch.cuda.device(self.device):
            self._st = _lib.new_state(1, RP, OP, CP, self.device)
            self._out = _lib.new_out(1, RP, self.device, debug=True)
            self._actions_f = torch.zeros((1, RP, 2), dtype=torch.float32, device=self.device)
            self._actions_i = torch.zeros((1, RP), dtype=torch.int32, device=self.device)

    def _params(self):
        """asvrl_env_params from the env's and robot 0's attributes (one parameter set per
        launch: heterogeneous robots are rejected)."""
        p = _abi.default_params()
        p.timestep_penalty, p.colregs_penalty = self.timestep_penalty, self.COLREGs_penalty
        p.collision_penalty, p.goal_reward = self.collision_penalty, self.goal_reward
        p.core_r = self.r
        if self.robots:
            first = self.robots[0].param_tuple()
            for rob in self.robots[1:]:
                if rob.param_tuple() != first:
                    raise RuntimeError("robots with different hull / perception parameters in one "
                                       "environment are not supported by the batched kernels")
            self.robots[0].fill_params(p)
        return p

    def _reset_params(self):
        rp = _abi.default_reset_params()
        rp.map_w, rp.map_h = float(self.width), float(self.height)
        rp.clear_r, rp.core_r = float(self.clear_r), float(self.r)
        rp.v_rel_max, rp.p = float(self.v_rel_max), float(self.p)
        rp.v_min, rp.v_max = float(self.v_range[0]), float(self.v_range[1])
        rp.obs_r_min, rp.obs_r_max = float(self.obs_r_range[0]), float(self.obs_r_range[1])
        return rp

    def _pull_state(self):
        """device state -> host mirrors (robots, cores, obstacles)."""
        st = {k: v.cpu().numpy() for k, v in self._st.items()}
        nr, no, nc = int(st["n_robots"][0]), int(st["n_obs"][0]), int(st["n_cores"][0])
        return st, nr, no, nc

    def _push_flags(self):
        """host ``deactivated`` / ``collision`` / ``reach_goal`` -> device flags word."""
        RP = self._shape[0]
        f = np.zeros((1, RP), np.int32)
        for i, rob in enumerate(self.robots):
            f[0, i] = ((_abi.F_COLLISION if rob.collision else 0) | (_abi.F_REACH_GOAL if rob.reach_goal else 0) |
                       (_abi.F_DEACTIVATED if rob.deactivated else 0))
        self._st["flags"].copy_(torch.from_numpy(f))

    def get_action_space_dimension(self):
        return self.robots[0].compute_actions_dimension()

    # ------------------------------------------------------------------ reset (ENV:72-164)
    def reset(self):
        if self.schedule is not None:
            steps = self.schedule["timesteps"]
            diffs = np.array(steps) - self.total_timesteps
            idx = len(diffs[diffs <= 0]) - 1  # interval the current timestep falls into
            self.num_robots = self.schedule["num_robots"][idx]
            assert self.num_robots > 0, "Number of robots is 0!"
            self.num_cores = self.schedule["num_cores"][idx]
            self.num_obs = self.schedule["num_obstacles"][idx]
            self.min_start_goal_dis = self.schedule["min_start_goal_dis"][idx]
            print("\n======== training schedule ========")
            print("num of robots: ", self.num_robots)
            print("num of cores: ", self.num_cores)
            print("num of obstacles: ", self.num_obs)
            print("min start goal dis: ", self.min_start_goal_dis)
            print("======== training schedule ========\n")

        self.episode_timesteps = 0
        self._alloc(self.num_robots, self.num_obs, self.num_cores)
        RP, OP, CP = self._shape
        self._episodes += 1
        dev = self.device
        with torch.cuda.device(dev):
            i32 = lambda v: torch.tensor([int(v)], dtype=torch.int32, device=dev)
            _lib.env_reset(self._reset_params(), self._st, None, i32(self.num_robots), i32(self.num_cores),
                           i32(self.num_obs),
                           torch.tensor([float(self.min_start_goal_dis)], dtype=torch.float32, device=dev),
                           int(self.seed), self._episodes, OP, CP)
        st, nr, no, nc = self._pull_state()
        self.cores = [Core(float(c[0]), float(c[1]), bool(c[2] > 0), float(c[3])) for c in st["cores"][0, :nc]]
        self.obstacles = [Obstacle(float(o[0]), float(o[1]), float(o[2])) for o in st["obstacles"][0, :no]]
        self.robots = []
        for i in range(nr):
            rob = Robot(seed=0)
            rob.start = np.array([float(st["x"][0, i]), float(st["y"][0, i])])
            rob.goal = np.array([float(st["goal_x"][0, i]), float(st["goal_y"][0, i])])
            rob.init_theta = float(st["theta"][0, i])
            rob.reset_state()  # no cores exist yet when the reference resets a robot (ENV:175)
            self.robots.append(rob)
        return self.get_observations()

    def check_all_deactivated(self):
        return all(rob.deactivated for rob in self.robots)

    def check_all_reach_goal(self):
        return all(rob.reach_goal for rob in self.robots)

    def check_any_collision(self):
        return any(rob.collision for rob in self.robots)

    def compute_COLREGs_penalty(self, rob):
        return self.COLREGs_penalty * rob.phi if rob.apply_COLREGs else 0.0  # ENV:232-238

    # ------------------------------------------------------------------ kernel launch + unpack
    def _launch(self, mode, actions=None):
        RP, OP, CP = self._shape
        self._obs_calls += 1
        with torch.cuda.device(self.device):
            self._push_flags()
            _lib.env_step(self._params(), self._st, actions, None, self._out, mode, OP, CP,
                          seed=int(self.seed) ^ 0x5851F42D4C957F2D, step=self._obs_calls)
            out = {k: v.cpu().numpy() for k, v in self._out.items()}
            st = {k: self._st[k].cpu().numpy() for k in _abi.STATE_FLOAT_FIELDS + ("flags",)}
        return out, st

    def _unpack(self, out, st):
        """Refresh the host mirrors and build the reference's observation lists
        (WAMV:436-529: (None, None) for a deactivated robot, <=5 object rows otherwise)."""
        observations, collisions, reach_goals = [], [], []
        OP = self._shape[1]
        for i, rob in enumerate(self.robots):
            status = int(out["status"][0, i])
            rob.x, rob.y, rob.theta = (float(st[k][0, i]) for k in ("x", "y", "theta"))
            rob.velocity_r = np.array([st["velr_x"][0, i], st["velr_y"][0, i], st["velr_r"][0, i]], dtype=float)
            rob.velocity = np.array([st["vel_x"][0, i], st["vel_y"][0, i], st["vel_r"][0, i]], dtype=float)
            rob.left_thrust, rob.right_thrust = float(st["thrust_l"][0, i]), float(st["thrust_r"][0, i])
            rob.collision = bool(status & _abi.S_COLLISION)
            rob.reach_goal = bool(status & _abi.S_REACH_GOAL)
            if rob.deactivated:
                observation = (None, None)
            else:
                rob.apply_COLREGs = bool(status & _abi.S_COLREGS)
                if rob.apply_COLREGs:
                    rob.phi = float(out["phi"][0, i])
                n = int(out["obs_mask"][0, i].sum())
                det = int(out["detected"][0, i].view(np.uint64)) if "detected" in out else 0
                rob.perception.observed_obs = [s for s in range(OP) if (det >> s) & 1]
                rob.perception.observed_objs = [s - OP for s in range(OP, 64) if (det >> s) & 1]
                observation = (out["obs_self"][0, i].tolist(), out["obs_objects"][0, i, :n].tolist())
                rob.perception.observation = dict(self=observation[0], objects=observation[1])
                if self.is_eval_env:
                    rob.observation_history.append([observation[0], observation[1]])
            observations.append(observation)
            collisions.append(rob.collision)
            reach_goals.append(rob.reach_goal)
        return observations, collisions, reach_goals

    def get_observations(self):
        """ENV:341-356: perception only (K1 with mode 0)."""
        out, st = self._launch(0)
        return self._unpack(out, st)

    # ------------------------------------------------------------------ step (ENV:240-333)
    def step(self, actions, is_continuous_action=True):
        assert len(actions) == len(self.robots), "Number of actions not equal number of robots!"
        assert self.check_all_reach_goal() is not True, "All robots reach goals, not actions are available!"
        for rob in self.robots:
            if rob.deactivated and not (rob.collision or rob.reach_goal):
                raise RuntimeError("Robot being deactived can only be caused by collsion or reaching goal!")
        mode = _abi.M_DYNAMICS
        if is_continuous_action:
            a = np.zeros((1, self._shape[0], 2), np.float32)
            for i, act in enumerate(actions):
                if act is not None and not self.robots[i].deactivated:
                    a[0, i] = act
            self._actions_f.copy_(torch.from_numpy(a))
            dev_actions = self._actions_f
        else:
            mode |= _abi.M_DISCRETE
            a = np.zeros((1, self._shape[0]), np.int32)
            for i, act in enumerate(actions):
                if act is not None and not self.robots[i].deactivated:
                    a[0, i] = int(act)
            self._actions_i.copy_(torch.from_numpy(a))
            dev_actions = self._actions_i
        # the kernel reads ep_steps from the device; keep it equal to the host counter
        self._st["ep_steps"].fill_(int(self.episode_timesteps))
        out, st = self._launch(mode, dev_actions)
        was_active = [not rob.deactivated for rob in self.robots]
        observations, _, _ = self._unpack(out, st)
        rewards, dones, infos = [], [], []
        for i, rob in enumerate(self.robots):
            status = int(out["status"][0, i])
            rewards.append(float(out["reward"][0, i]) if was_active[i] else 0)
            dones.append(bool(status & _abi.S_DONE))
            infos.append({"state": _abi.INFO_STRINGS[(status >> _abi.S_INFO_SHIFT) & 15]})
            if self.is_eval_env and was_active[i]:
                rob.action_history.append(actions[i])
                rob.trajectory.append(rob.state_row())
        self.episode_timesteps += 1
        self.total_timesteps += 1
        return observations, rewards, dones, infos

    # ------------------------------------------------------------------ current field (ENV:458-501)
    def get_velocity(self, x: float, y: float):
        """Current at (x, y) = sum over all vortex cores, evaluated with device tensor ops
        (used only to initialise ``velocity`` when a stored scenario is loaded, ENV:612)."""
        if len(self.cores) == 0:
            return np.zeros(3)
        dev = self.device
        c = torch.tensor([[k.x, k.y, 1.0 if k.clockwise else -1.0, k.Gamma] for k in self.cores],
                         dtype=torch.float64, device=dev)
        rx, ry = c[:, 0] - float(x), c[:, 1] - float(y)
        d = torch.sqrt(rx * rx + ry * ry)
        speed = torch.where(d <= self.r, c[:, 3] / (2 * np.pi * self.r * self.r) * d, c[:, 3] / (2 * np.pi * d))
        vx = (-(ry / d) * c[:, 2] * speed).sum()
        vy = ((rx / d) * c[:, 2] * speed).sum()
        return np.array([float(vx), float(vy), 0.0])

    # ------------------------------------------------------------------ stored scenarios
    _ENV_KEYS = ("width", "height", "r", "v_rel_max", "p", "clear_r", "angular_speed_max",
                 "angular_speed_penalty", "timestep_penalty", "COLREGs_penalty", "collision_penalty",
                 "goal_reward")
    _ROB_KEYS = ("dt", "N", "length", "width", "detect_r", "r", "goal_dis", "power_coefficient",
                 "min_thrust", "max_thrust", "m", "Izz", "xDotU", "yDotV", "yDotR", "nDotR", "nDotV",
                 "xU", "xUU", "yV", "yVV", "yR", "yRV", "yVR", "yRR", "nR", "nRR", "nV", "nVV", "nRV",
                 "nVR", "init_theta", "init_left_pos", "init_right_pos", "init_left_thrust",
                 "init_right_thrust")
    _PERC_KEYS = ("range", "angle", "max_obj_num", "pos_std", "vel_std", "r_kappa", "r_mean_ratio")

    def reset_with_eval_config(self, eval_config):
        """Load a scenario written by ``episode_data`` (ENV:503-614) and return its first
        observation."""
        self.episode_timesteps = 0
        env_cfg, rob_cfg = eval_config["env"], eval_config["robots"]
        self.seed = env_cfg["seed"]
        self.is_eval_env = env_cfg["is_eval_env"]
        for k in self._ENV_KEYS:
            setattr(self, k, env_cfg[k])
        self.v_range = copy.deepcopy(env_cfg["v_range"])
        self.obs_r_range = copy.deepcopy(env_cfg["obs_r_range"])
        self.cores = [Core(pos[0], pos[1], cw, g) for pos, cw, g in
                      zip(env_cfg["cores"]["positions"], env_cfg["cores"]["clockwise"], env_cfg["cores"]["Gamma"])]
        self.obstacles = [Obstacle(pos[0], pos[1], r) for pos, r in
                          zip(env_cfg["obstacles"]["positions"], env_cfg["obstacles"]["r"])]
        self.robots = []
        for i in range(env_cfg["num_robots"]):
            rob = Robot(seed=0)
            for k in self._ROB_KEYS:
                setattr(rob, k, rob_cfg[k][i])
            rob.left_thrust_change = np.array(rob_cfg["left_thrust_change"][i])
            rob.right_thrust_change = np.array(rob_cfg["right_thrust_change"][i])
            rob.compute_actions()
            rob.start = np.array(rob_cfg["start"][i])
            rob.goal = np.array(rob_cfg["goal"][i])
            rob.init_velocity_r = np.array(rob_cfg["init_velocity_r"][i])
            for k in self._PERC_KEYS:
                setattr(rob.perception, k, rob_cfg["perception"][k][i])
            rob.reset_state(current_velocity=self.get_velocity(rob.start[0], rob.start[1]))
            self.robots.append(rob)
        self._upload_scene()
        return self.get_observations()

    def _upload_scene(self):
        """host mirrors (robots, cores, obstacles) -> device state of the single env."""
        self._alloc(len(self.robots), len(self.obstacles), len(self.cores))
        RP, OP, CP = self._shape
        host = {k: np.zeros((1, RP), np.float32) for k in _abi.STATE_FLOAT_FIELDS}
        for i, rob in enumerate(self.robots):
            vals = dict(x=rob.x, y=rob.y, theta=rob.theta, velr_x=rob.velocity_r[0], velr_y=rob.velocity_r[1],
                        velr_r=rob.velocity_r[2], vel_x=rob.velocity[0], vel_y=rob.velocity[1],
                        vel_r=rob.velocity[2], thrust_l=rob.left_thrust, thrust_r=rob.right_thrust,
                        goal_x=rob.goal[0], goal_y=rob.goal[1])
            for k, v in vals.items():
                host[k][0, i] = v
        obst = np.zeros((1, max(OP, 1), 4), np.float32)
        for i, o in enumerate(self.obstacles):
            obst[0, i, :3] = (o.x, o.y, o.r)
        cores = np.zeros((1, max(CP, 1), 4), np.float32)
        for i, c in enumerate(self.cores):
            cores[0, i] = (c.x, c.y, 1.0 if c.clockwise else -1.0, c.Gamma)
        with torch.cuda.device(self.device):
            for k, v in host.items():
                self._st[k].copy_(torch.from_numpy(v))
            self._st["obstacles"].copy_(torch.from_numpy(obst))
            self._st["cores"].copy_(torch.from_numpy(cores))
            self._st["n_robots"].fill_(len(self.robots))
            self._st["n_obs"].fill_(len(self.obstacles))
            self._st["n_cores"].fill_(len(self.cores))
            self._st["ep_steps"].fill_(0)

    def episode_data(self):
        """Scenario + histories in the JSON schema of the reference (ENV:616-773)."""
        env = {"seed": self.seed, "is_eval_env": self.is_eval_env}
        for k in self._ENV_KEYS:
            env[k] = getattr(self, k)
        env["v_range"] = copy.deepcopy(self.v_range)
        env["obs_r_range"] = copy.deepcopy(self.obs_r_range)
        env["num_robots"] = self.num_robots
        env["cores"] = {"positions": [[c.x, c.y] for c in self.cores],
                        "clockwise": [c.clockwise for c in self.cores],
                        "Gamma": [c.Gamma for c in self.cores]}
        env["obstacles"] = {"positions": [[o.x, o.y] for o in self.obstacles],
                            "r": [o.r for o in self.obstacles]}
        robots = {k: [getattr(rob, k) for rob in self.robots] for k in self._ROB_KEYS}
        robots["left_thrust_change"] = [list(map(float, rob.left_thrust_change)) for rob in self.robots]
        robots["right_thrust_change"] = [list(map(float, rob.right_thrust_change)) for rob in self.robots]
        robots["start"] = [list(map(float, rob.start)) for rob in self.robots]
        robots["goal"] = [list(map(float, rob.goal)) for rob in self.robots]
        robots["init_velocity_r"] = [list(map(float, rob.init_velocity_r)) for rob in self.robots]
        robots["perception"] = {k: [getattr(rob.perception, k) for rob in self.robots] for k in self._PERC_KEYS}
        robots["observation_history"] = [copy.deepcopy(rob.observation_history) for rob in self.robots]
        robots["action_history"] = [copy.deepcopy(rob.action_history) for rob in self.robots]
        robots["trajectory"] = [copy.deepcopy(rob.trajectory) for rob in self.robots]
        return {"env": env, "robots": robots}

    def save_episode(self, filename):
        with open(filename, "w") as file:
            json.dump(self.episode_data(), file)
