"""ctypes mirror of include/asvrl.h (struct layouts, constants).

Pure declarations: importing this module loads no native code.  The field order of
``EnvParams`` must match ``asvrl_env_params`` exactly; tests/test_abi.py checks the size and
the defaults against ``asvrl_env_default_params`` of the built library.
Reference defaults: rfarl/rfarl/envs/marinenav/vehicles/wamv.py:45-126 and
rfarl/rfarl/envs/marinenav/env.py:33-54.
"""
import ctypes as C
import math

VERSION = 200
MAX_OBJ = 5
SELF_DIM = 7
OBJ_DIM = 5
NOISE_DIM = 5
MAX_SLOTS = 64

E_BADARG, E_LIMIT, E_NODEV = -1, -2, -3

F_COLLISION, F_REACH_GOAL, F_DEACTIVATED = 1, 2, 4

S_DONE, S_COLLISION, S_REACH_GOAL, S_WAS_DEACTIVATED, S_COLREGS = 1, 2, 4, 8, 16
S_INFO_SHIFT = 8
INFO_NORMAL, INFO_COLLISION, INFO_REACH_GOAL, INFO_TOO_LONG = 0, 1, 2, 3
INFO_DEACT_COLLISION, INFO_DEACT_GOAL, INFO_DEACT_INVALID, INFO_ABSENT = 4, 5, 6, 7
# ENV:291-328
INFO_STRINGS = {
    INFO_NORMAL: "normal",
    INFO_COLLISION: "collision",
    INFO_REACH_GOAL: "reach goal",
    INFO_TOO_LONG: "too long episode",
    INFO_DEACT_COLLISION: "deactivated after collision",
    INFO_DEACT_GOAL: "deactivated after reaching goal",
}

M_DYNAMICS, M_DISCRETE, M_AUTO_DEACTIVATE = 1, 2, 4

ES_ALL_DEACTIVATED, ES_TIMEOUT, ES_ANY_COLLISION, ES_ALL_REACH_GOAL = 1, 2, 4, 8

ROW_SELF, ROW_OBJ, ROW_NOBJ, ROW_ACT, ROW_REW = 0, 7, 32, 33, 35
ROW_NSELF, ROW_NOBJ_F, ROW_NNOBJ, ROW_DONE, ROW = 36, 43, 68, 69, 70


class EnvParams(C.Structure):
    _fields_ = [
        ("dt", C.c_double),
        ("n_substeps", C.c_int32),
        ("max_episode_steps", C.c_int32),
        ("length", C.c_double), ("width", C.c_double), ("robot_r", C.c_double),
        ("goal_dis", C.c_double),
        ("min_thrust", C.c_double), ("max_thrust", C.c_double),
        ("thrust_rate", C.c_double),
        ("thrust_change_l", C.c_double * 5),
        ("thrust_change_r", C.c_double * 5),
        ("left_pos", C.c_double), ("right_pos", C.c_double),
        ("m", C.c_double), ("Izz", C.c_double),
        ("xDotU", C.c_double), ("yDotV", C.c_double), ("yDotR", C.c_double),
        ("nDotR", C.c_double), ("nDotV", C.c_double),
        ("xU", C.c_double), ("xUU", C.c_double), ("yV", C.c_double), ("yVV", C.c_double),
        ("yR", C.c_double), ("yRV", C.c_double), ("yVR", C.c_double), ("yRR", C.c_double),
        ("nR", C.c_double), ("nRR", C.c_double), ("nV", C.c_double), ("nVV", C.c_double),
        ("nRV", C.c_double), ("nVR", C.c_double),
        ("perc_range", C.c_double), ("perc_angle", C.c_double),
        ("pos_std", C.c_double), ("vel_std", C.c_double), ("r_kappa", C.c_double),
        ("r_mean_ratio", C.c_double),
        ("head_on_x", C.c_double), ("head_on_y", C.c_double),
        ("lc_x0", C.c_double), ("lc_x1", C.c_double), ("lc_y0", C.c_double),
        ("lc_y1", C.c_double),
        ("colregs_min_speed", C.c_double),
        ("timestep_penalty", C.c_double), ("colregs_penalty", C.c_double),
        ("collision_penalty", C.c_double), ("goal_reward", C.c_double),
        ("core_r", C.c_double),
        ("drift_vx", C.c_double), ("drift_vy", C.c_double),
    ]

    def copy(self):
        q = EnvParams()
        C.memmove(C.byref(q), C.byref(self), C.sizeof(EnvParams))
        return q


def default_params() -> EnvParams:
    """Reference defaults (WAMV:45-126, ENV:33-54)."""
    p = EnvParams()
    p.dt, p.n_substeps, p.max_episode_steps = 0.05, 10, 1000
    p.length, p.width = 5.0, 2.5
    p.robot_r = 0.5 * math.sqrt(p.length ** 2 + p.width ** 2)
    p.goal_dis = 2.0
    p.min_thrust, p.max_thrust, p.thrust_rate = -500.0, 1000.0, 1000.0
    for i, v in enumerate([0.0, -500.0, -1000.0, 500.0, 1000.0]):
        p.thrust_change_l[i] = v
        p.thrust_change_r[i] = v
    p.left_pos = p.right_pos = 0.0
    p.m, p.Izz = 400.0, 450.0
    p.xDotU, p.yDotV, p.yDotR, p.nDotR, p.nDotV = 20.0, 0.0, 0.0, -980.0, 0.0
    p.xU, p.xUU, p.yV, p.yVV = -100.0, -150.0, -100.0, -150.0
    p.yR = p.yRV = p.yVR = p.yRR = 0.0
    p.nR, p.nRR = -980.0, -950.0
    p.nV = p.nVV = p.nRV = p.nVR = 0.0
    p.perc_range, p.perc_angle = 20.0, 2.0 * math.pi
    p.pos_std, p.vel_std, p.r_kappa, p.r_mean_ratio = 0.05, 0.05, 1.0, 0.8
    p.head_on_x, p.head_on_y = 17.0, 9.0
    p.lc_x0, p.lc_x1, p.lc_y0, p.lc_y1 = -9.0, 12.0, -17.0, -7.0
    p.colregs_min_speed = 0.5
    p.timestep_penalty, p.colregs_penalty = -0.1, -0.1
    p.collision_penalty, p.goal_reward = -5.0, 10.0
    p.core_r = 0.5
    p.drift_vx = p.drift_vy = 0.0
    return p


STATE_FLOAT_FIELDS = ("x", "y", "theta", "velr_x", "velr_y", "velr_r", "vel_x", "vel_y", "vel_r",
                      "thrust_l", "thrust_r", "goal_x", "goal_y")
STATE_ENV_INT_FIELDS = ("n_robots", "n_obs", "n_cores", "ep_steps")


class EnvStatePtrs(C.Structure):
    """asvrl_env_state: raw device pointers (void* here; typed in the header)."""
    _fields_ = ([(n, C.c_void_p) for n in STATE_FLOAT_FIELDS] + [("flags", C.c_void_p)] +
                [(n, C.c_void_p) for n in STATE_ENV_INT_FIELDS] +
                [("obstacles", C.c_void_p), ("cores", C.c_void_p)])


OUT_FIELDS = ("obs_self", "obs_objects", "obs_mask", "reward", "status", "phi", "top_ids",
              "detected", "env_status", "ended")


class EnvOutPtrs(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in OUT_FIELDS]


class ResetParams(C.Structure):
    """asvrl_reset_params (ENV:33-41, 107-155)."""
    _fields_ = [(n, C.c_double) for n in ("map_w", "map_h", "start_margin", "obs_margin", "clear_r",
                                          "core_r", "v_rel_max", "p", "v_min", "v_max", "obs_r_min",
                                          "obs_r_max")] + [("max_tries", C.c_int32),
                                                           ("reserved", C.c_int32)]


def default_reset_params() -> ResetParams:
    """Reference defaults (ENV:33-41; 500 tries ENV:107,124,152)."""
    p = ResetParams()
    p.map_w = p.map_h = 55.0
    p.start_margin, p.obs_margin, p.clear_r, p.core_r = 2.0, 5.0, 10.0, 0.5
    p.v_rel_max, p.p, p.v_min, p.v_max = 1.0, 0.8, 3.0, 3.0
    p.obs_r_min = p.obs_r_max = 1.0
    p.max_tries, p.reserved = 500, 0
    return p


class EnvAutoReset(C.Structure):
    """asvrl_env_autoreset: arguments of the fused step + auto-reset launch (include/asvrl.h)."""
    _fields_ = [("reset_params", C.POINTER(ResetParams)), ("want_robots", C.c_void_p),
                ("want_cores", C.c_void_p), ("want_obs", C.c_void_p),
                ("min_start_goal_dis", C.c_void_p), ("episode_idx", C.c_void_p),
                ("observe_seed", C.c_uint64), ("term_self", C.c_void_p),
                ("term_objects", C.c_void_p), ("term_mask", C.c_void_p),
                ("stage", C.POINTER(EnvStatePtrs)), ("stage_tag", C.c_void_p),
                ("stage_self", C.c_void_p), ("stage_objects", C.c_void_p), ("stage_mask", C.c_void_p),
                ("refill_in_launch", C.c_int32), ("reserved", C.c_int32)]


OPT_MAX_TENSORS = 64
OPT_SCRATCH_FLOATS = 320


class AdamTensors(C.Structure):
    """asvrl_adam_tensors (include/asvrl.h)."""
    _fields_ = [("n", C.c_int),
                ("param", C.c_void_p * OPT_MAX_TENSORS), ("grad", C.c_void_p * OPT_MAX_TENSORS),
                ("exp_avg", C.c_void_p * OPT_MAX_TENSORS), ("exp_avg_sq", C.c_void_p * OPT_MAX_TENSORS),
                ("step", C.c_void_p * OPT_MAX_TENSORS),
                ("numel", C.c_longlong * OPT_MAX_TENSORS)]


MAX_PEERS = 8


class PeerBuckets(C.Structure):
    """asvrl_peer_buckets (include/asvrl.h)."""
    _fields_ = [("world", C.c_int), ("rank", C.c_int), ("bucket", C.c_void_p * MAX_PEERS),
                ("half_floats", C.c_longlong)]
