"""Host mirror of one ASV with the attribute surface of the reference ``Robot`` and
``Perception`` (rfarl/rfarl/envs/marinenav/vehicles/wamv.py:5-193).

The dynamics, perception and COLREGs arithmetic of the reference class live in kernel K1
(csrc/env_step.cu); this object only carries parameters, the per-episode bookkeeping the
Trainer reads and writes (``deactivated``, ``collision``, ``reach_goal``, histories), and a copy
of the device state refreshed after every launch.
"""
import numpy as np

# hull / hydrodynamic attributes that map one-to-one onto asvrl_env_params (WAMV:46-125)
_DIRECT = ("dt", "length", "width", "goal_dis", "min_thrust", "max_thrust", "m", "Izz", "xDotU",
           "yDotV", "yDotR", "nDotR", "nDotV", "xU", "xUU", "yV", "yVV", "yR", "yRV", "yVR", "yRR",
           "nR", "nRR", "nV", "nVV", "nRV", "nVR")


class Perception:
    """Sensor model parameters (WAMV:5-25); the noise itself is drawn inside K1."""

    def __init__(self, seed: int = 0):
        self.seed = seed
        self.range = 20.0
        self.angle = 2 * np.pi
        self.max_obj_num = 5
        self.observation = dict(self=[], objects=[])
        self.observed_obs = []
        self.observed_objs = []
        self.pos_std = 0.05
        self.vel_std = 0.05
        self.r_kappa = 1.0
        self.r_mean_ratio = 0.8


class Robot:

    def __init__(self, seed: int = 0):
        self.dt = 0.05
        self.N = 10
        self.perception = Perception(seed)
        self.length = 5.0
        self.width = 2.5
        self.detect_r = 0.5 * np.sqrt(self.length ** 2 + self.width ** 2)
        self.r = self.detect_r
        # COLREGs zones (WAMV:73-76)
        self.head_on_zone_x_dim = 17.0
        self.head_on_zone_y_dim = 9.0
        self.left_crossing_zone_x_dim = np.array([-9.0, 12.0])
        self.left_crossing_zone_y_dim_front = np.array([-17.0, -7.0])
        self.safe_dis = 10.0
        self.goal_dis = 2.0
        self.power_coefficient = 1.0
        self.min_thrust = -500.0
        self.max_thrust = 1000.0
        self.left_thrust_change = np.array([0.0, -500.0, -1000.0, 500.0, 1000.0])
        self.right_thrust_change = np.array([0.0, -500.0, -1000.0, 500.0, 1000.0])
        self.compute_actions()
        self.x = self.y = self.theta = None
        self.velocity_r = None
        self.velocity = None
        self.left_pos = self.right_pos = None
        self.left_thrust = self.right_thrust = None
        self.m, self.Izz = 400, 450
        self.xDotU, self.yDotV, self.yDotR, self.nDotR, self.nDotV = 20, 0, 0, -980, 0
        self.xU, self.xUU, self.yV, self.yVV = -100, -150, -100, -150
        self.yR = self.yRV = self.yVR = self.yRR = 0
        self.nR, self.nRR = -980, -950
        self.nV = self.nVV = self.nRV = self.nVR = 0
        self.start = None
        self.goal = None
        self.collision = False
        self.reach_goal = False
        self.deactivated = False
        self.apply_COLREGs = False
        self.phi = 0.0
        self.init_theta = 0.0
        self.init_velocity_r = np.array([0.0, 0.0, 0.0])
        self.init_left_pos = self.init_right_pos = 0.0
        self.init_left_thrust = self.init_right_thrust = 0.0
        self.observation_history = []
        self.action_history = []
        self.trajectory = []

    # ------------------------------------------------------------------ reference helpers
    def compute_actions(self):
        """5 x 5 thrust-change table, left-major (WAMV:146-147)."""
        self.actions = [(l, r) for l in self.left_thrust_change for r in self.right_thrust_change]

    def compute_actions_dimension(self):
        return len(self.actions)

    def compute_constant_matrices(self):
        """The reference caches M_RB / M_A / D here (WAMV:152-159); K1 derives them per launch."""

    def compute_step_energy_cost(self):
        """WAMV:161-165."""
        l = self.power_coefficient * np.abs(self.left_thrust) * self.dt * self.N
        r = self.power_coefficient * np.abs(self.right_thrust) * self.dt * self.N
        return l + r

    def dist_to_goal(self):
        return np.linalg.norm(np.asarray(self.goal) - np.array([self.x, self.y]))

    def state_row(self):
        """trajectory record (ENV:266-268, WAMV:190-192)."""
        return [self.x, self.y, self.theta, self.velocity_r[0], self.velocity_r[1], self.velocity_r[2],
                self.velocity[0], self.velocity[1], self.velocity[2], self.left_pos, self.right_pos,
                self.left_thrust, self.right_thrust]

    def reset_state(self, current_velocity=np.zeros(3)):
        """WAMV:177-193."""
        self.observation_history.clear()
        self.action_history.clear()
        self.trajectory.clear()
        self.x, self.y = float(self.start[0]), float(self.start[1])
        self.theta = self.init_theta
        self.velocity_r = np.array(self.init_velocity_r, dtype=float)
        self.velocity = self.velocity_r + np.asarray(current_velocity, dtype=float)
        self.left_pos, self.right_pos = self.init_left_pos, self.init_right_pos
        self.left_thrust, self.right_thrust = self.init_left_thrust, self.init_right_thrust
        self.trajectory.append(self.state_row())

    # ------------------------------------------------------------------ kernel parameter block
    def param_tuple(self):
        pc = self.perception
        return (tuple(float(getattr(self, k)) for k in _DIRECT) + (int(self.N), float(self.r)) +
                tuple(float(v) for v in self.left_thrust_change) +
                tuple(float(v) for v in self.right_thrust_change) +
                (float(self.init_left_pos), float(self.init_right_pos), float(pc.range), float(pc.angle),
                 float(pc.pos_std), float(pc.vel_std), float(pc.r_kappa), float(pc.r_mean_ratio),
                 float(self.head_on_zone_x_dim), float(self.head_on_zone_y_dim)) +
                tuple(float(v) for v in self.left_crossing_zone_x_dim) +
                tuple(float(v) for v in self.left_crossing_zone_y_dim_front))

    def fill_params(self, p):
        """Write this robot's parameters into an asvrl_env_params block."""
        for k in _DIRECT:
            setattr(p, k, float(getattr(self, k)))
        p.n_substeps = int(self.N)
        p.robot_r = float(self.r)
        if len(self.left_thrust_change) != 5 or len(self.right_thrust_change) != 5:
            raise RuntimeError("the discrete action table is 5 x 5 (WAMV:86-87)")
        for i in range(5):
            p.thrust_change_l[i] = float(self.left_thrust_change[i])
            p.thrust_change_r[i] = float(self.right_thrust_change[i])
        p.left_pos, p.right_pos = float(self.init_left_pos), float(self.init_right_pos)
        pc = self.perception
        if pc.max_obj_num != 5:
            raise RuntimeError("the observation layout is fixed at 5 objects (WAMV:14)")
        p.perc_range, p.perc_angle = float(pc.range), float(pc.angle)
        p.pos_std, p.vel_std = float(pc.pos_std), float(pc.vel_std)
        p.r_kappa, p.r_mean_ratio = float(pc.r_kappa), float(pc.r_mean_ratio)
        p.head_on_x, p.head_on_y = float(self.head_on_zone_x_dim), float(self.head_on_zone_y_dim)
        p.lc_x0, p.lc_x1 = (float(v) for v in self.left_crossing_zone_x_dim)
        p.lc_y0, p.lc_y1 = (float(v) for v in self.left_crossing_zone_y_dim_front)
        return p
