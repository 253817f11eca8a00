"""Seeded synthetic scenes for the parity tests and the bench (TEST/BENCH INFRASTRUCTURE).

Dense on purpose: ASVs, buoys and goals are packed into a small box so that detections,
collisions, goal arrivals, top-5 truncation and COLREGs encounters all occur often.  Every
input value is float32-representable (feed the float64 oracle and the
fp32 kernel the same numbers).
"""
import math

import numpy as np

from . import oracle as _o


def f32(a):
    return np.asarray(a, np.float32).astype(np.float64)


def random_scene(seed, E, RP, OP, CP, box=30.0, p_deact=0.1, p_timeout=0.05, ragged=True,
                 speed=3.0):
    rng = np.random.RandomState(seed)
    st = _o.new_state(E, RP, OP, CP)
    st["n_robots"][:] = rng.randint(1, RP + 1, E) if ragged else RP
    st["n_obs"][:] = rng.randint(0, OP + 1, E) if (ragged and OP > 0) else OP
    st["n_cores"][:] = rng.randint(0, CP + 1, E) if (ragged and CP > 0) else CP
    off = 0.5 * (55.0 - box)
    st["x"][:] = f32(off + box * rng.rand(E, RP))
    st["y"][:] = f32(off + box * rng.rand(E, RP))
    st["theta"][:] = f32(np.minimum(2 * math.pi * rng.rand(E, RP), 6.2831))
    for k in ("velr_x", "velr_y"):
        st[k][:] = f32(speed * (rng.rand(E, RP) - 0.5))
    st["velr_r"][:] = f32(1.0 * (rng.rand(E, RP) - 0.5))
    # velocity = stale velocity_r + a small current-like offset (WAMV:201-202)
    st["vel_x"][:] = f32(st["velr_x"] + 0.3 * (rng.rand(E, RP) - 0.5))
    st["vel_y"][:] = f32(st["velr_y"] + 0.3 * (rng.rand(E, RP) - 0.5))
    st["vel_r"][:] = st["velr_r"]
    st["thrust_l"][:] = f32(-500 + 1500 * rng.rand(E, RP))
    st["thrust_r"][:] = f32(-500 + 1500 * rng.rand(E, RP))
    # a quarter of the goals sit within a few metres of the ASV so goal arrivals happen
    near = rng.rand(E, RP) < 0.25
    gx = off + box * rng.rand(E, RP)
    gy = off + box * rng.rand(E, RP)
    gx = np.where(near, st["x"] + 4.5 * (rng.rand(E, RP) - 0.5), gx)
    gy = np.where(near, st["y"] + 4.5 * (rng.rand(E, RP) - 0.5), gy)
    st["goal_x"][:] = f32(gx)
    st["goal_y"][:] = f32(gy)
    deact = rng.rand(E, RP) < p_deact
    why = rng.rand(E, RP) < 0.5
    sticky = rng.rand(E, RP) < 0.05  # collided/arrived last step but not yet deactivated
    fl = np.where(deact, np.where(why, 1 | 4, 2 | 4), np.where(sticky, np.where(why, 1, 2), 0))
    st["flags"][:] = fl.astype(np.uint32)
    st["ep_steps"][:] = np.where(rng.rand(E) < p_timeout, 1000 + rng.randint(0, 3, E),
                                 rng.randint(0, 1000, E)).astype(np.int32)
    if OP > 0:
        st["obstacles"][:, :, 0] = f32(off + box * rng.rand(E, OP))
        st["obstacles"][:, :, 1] = f32(off + box * rng.rand(E, OP))
        st["obstacles"][:, :, 2] = f32(0.5 + 1.5 * rng.rand(E, OP))
    if CP > 0:
        st["cores"][:, :, 0] = f32(55.0 * rng.rand(E, CP))
        st["cores"][:, :, 1] = f32(55.0 * rng.rand(E, CP))
        st["cores"][:, :, 2] = np.where(rng.rand(E, CP) < 0.5, 1.0, -1.0)
        st["cores"][:, :, 3] = f32(2 * math.pi * 0.5 * (2.0 + 2.0 * rng.rand(E, CP)))
    # zero the padding so fixtures are canonical
    rmask = np.arange(RP)[None, :] < st["n_robots"][:, None]
    for k in _o._abi.STATE_FLOAT_FIELDS:
        st[k][~rmask] = 0.0
    st["flags"][~rmask] = 0
    if OP > 0:
        st["obstacles"][~(np.arange(OP)[None, :] < st["n_obs"][:, None])] = 0.0
    if CP > 0:
        st["cores"][~(np.arange(CP)[None, :] < st["n_cores"][:, None])] = 0.0
    return st


def random_actions(seed, E, RP, discrete):
    rng = np.random.RandomState(seed + 7919)
    if discrete:
        return rng.randint(0, 25, (E, RP)).astype(np.int32)
    return f32(2.0 * rng.rand(E, RP, 2) - 1.0)


def random_noise(seed, E, RP, OP, params):
    """Sensor noise with the reference's distributions (WAMV:22-40), float32-representable."""
    rng = np.random.RandomState(seed + 104729)
    S = OP + RP
    n = np.zeros((E, RP, S, 5))
    n[..., 0:2] = rng.normal(0.0, params.pos_std, (E, RP, S, 2))
    n[..., 2:4] = rng.normal(0.0, params.vel_std, (E, RP, S, 2))
    n[..., 4] = rng.vonmises(0.0, params.r_kappa, (E, RP, S))
    return f32(n)


def copy_state(st):
    return {k: v.copy() for k, v in st.items()}
