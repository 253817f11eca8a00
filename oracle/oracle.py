"""ctypes binding of oracle/liboracle.so — the float64 CPU oracle.  TEST INFRASTRUCTURE ONLY.

Only tests/, ``__graft_entry__.smoke()`` and bench.py's ``cpu_baseline`` / ``--impl reference``
legs may import this module; the product package never does.  See asv_oracle.h for the parity
status (pinned by tests/golden/, generated from the Python reference by gen_golden.py).
"""
import ctypes as C
import os
import subprocess

import numpy as np

from distributional_rl_decision_and_control_b200 import _abi

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = [os.path.join(_HERE, f) for f in ("asv_oracle.c", "asv_oracle.h")]
    stale = (not os.path.exists(so)) or any(
        os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(so) for s in src)
    if force or stale:
        cmd = ["gcc", "-O2", "-fPIC", "-std=c11", "-ffp-contract=off", "-shared", "-o", so,
               os.path.join(_HERE, "asv_oracle.c"), "-lm", "-lpthread"]
        subprocess.run(cmd, check=True, cwd=_HERE)
    return so


class _State(C.Structure):
    _fields_ = ([(n, C.c_void_p) for n in _abi.STATE_FLOAT_FIELDS] + [("flags", C.c_void_p)] +
                [(n, C.c_void_p) for n in _abi.STATE_ENV_INT_FIELDS] +
                [("obstacles", C.c_void_p), ("cores", C.c_void_p)])


class _Out(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in _abi.OUT_FIELDS if n != "ended"] + [("margin_perc", C.c_void_p),
                                                              ("margin_colregs", C.c_void_p)]


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.oracle_env_step.restype = C.c_int
        _LIB.oracle_env_step.argtypes = [C.POINTER(_abi.EnvParams), C.POINTER(_State), C.c_void_p,
                                         C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(_Out),
                                         C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_uint32,
                                         C.c_int]
        _LIB.oracle_env_noise.restype = C.c_int
        _LIB.oracle_env_noise.argtypes = [C.POINTER(_abi.EnvParams), C.c_uint64, C.c_uint64,
                                          C.c_int, C.c_int, C.c_int, C.c_int64, C.c_void_p]
        _LIB.oracle_current.restype = None
        _LIB.oracle_current.argtypes = [C.POINTER(_abi.EnvParams), C.c_void_p, C.c_int, C.c_double,
                                        C.c_double, C.c_void_p]
        _LIB.oracle_qhuber.restype = C.c_int
        _LIB.oracle_qhuber.argtypes = [C.c_void_p] * 5 + [C.c_double, C.c_double, C.c_int, C.c_int,
                                                          C.c_int, C.c_void_p, C.c_void_p,
                                                          C.c_void_p]
        _LIB.oracle_tau_cos.restype = C.c_int
        _LIB.oracle_tau_cos.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        dp = C.c_void_p
        RP_ = C.POINTER(_abi.ResetParams)
        _LIB.oracle_check_start_goal.restype = C.c_int
        _LIB.oracle_check_start_goal.argtypes = [RP_, dp, C.c_int, dp, C.c_double, dp]
        _LIB.oracle_check_core.restype = C.c_int
        _LIB.oracle_check_core.argtypes = [RP_, dp, C.c_int, dp, C.c_int, dp, dp]
        _LIB.oracle_check_obstacle.restype = C.c_int
        _LIB.oracle_check_obstacle.argtypes = [RP_, dp, C.c_int, dp, C.c_int, dp, C.c_int, dp, dp]
        _LIB.oracle_env_reset.restype = C.c_int
        _LIB.oracle_env_reset.argtypes = [RP_, C.POINTER(_State), dp, dp, dp, dp, dp, C.c_uint64,
                                          C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, dp]
        _LIB.oracle_philox4x32_10.restype = None
        _LIB.oracle_philox4x32_10.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def new_state(E, RP, OP, CP):
    """Zeroed float64 state arrays with the layout of oracle_env_state."""
    st = {n: np.zeros((E, RP), np.float64) for n in _abi.STATE_FLOAT_FIELDS}
    st["flags"] = np.zeros((E, RP), np.uint32)
    for n in _abi.STATE_ENV_INT_FIELDS:
        st[n] = np.zeros((E,), np.int32)
    st["obstacles"] = np.zeros((E, max(OP, 1), 4), np.float64)
    st["cores"] = np.zeros((E, max(CP, 1), 4), np.float64)
    return st


def new_out(E, RP):
    return {
        "obs_self": np.zeros((E, RP, 7)), "obs_objects": np.zeros((E, RP, 5, 5)),
        "obs_mask": np.zeros((E, RP, 5)), "reward": np.zeros((E, RP)),
        "status": np.zeros((E, RP), np.uint32), "phi": np.zeros((E, RP)),
        "top_ids": np.zeros((E, RP, 5), np.int32), "detected": np.zeros((E, RP), np.uint64),
        "env_status": np.zeros((E,), np.uint32),
        "margin_perc": np.zeros((E, RP)), "margin_colregs": np.zeros((E, RP)),
    }


def env_step(params, st, actions, noise, mode, max_obs, max_cores, seed=0, step=0, env_offset=0,
             n_threads=1, out=None, margins=True):
    """One batched step IN PLACE on the float64 state dict ``st``; returns the out dict."""
    E, RP = st["x"].shape
    for k, v in st.items():
        assert v.flags["C_CONTIGUOUS"], k
    if out is None:
        out = new_out(E, RP)
    cs = _State(**{k: _p(v) for k, v in st.items()})
    co = _Out(**{k: _p(v) for k, v in out.items()})
    if not margins:
        co.margin_perc = None
        co.margin_colregs = None
    if actions is not None:
        actions = np.ascontiguousarray(
            actions, np.int32 if (mode & _abi.M_DISCRETE) else np.float64)
    if noise is not None:
        noise = np.ascontiguousarray(noise, np.float64)
        assert noise.shape == (E, RP, max_obs + RP, 5), noise.shape
    rc = lib().oracle_env_step(C.byref(params), C.byref(cs), _p(actions), _p(noise), seed, step,
                               C.byref(co), E, RP, max_obs, max_cores, env_offset, mode,
                               n_threads)
    if rc != 0:
        raise RuntimeError(f"oracle_env_step failed: {rc}")
    return out


def env_noise(params, seed, step, E, RP, OP, env_offset=0):
    out = np.zeros((E, RP, OP + RP, 5), np.float64)
    lib().oracle_env_noise(C.byref(params), seed, step, E, RP, OP, env_offset, _p(out))
    return out


def current(params, cores, x, y):
    cores = np.ascontiguousarray(cores, np.float64).reshape(-1, 4)
    out = np.zeros(2)
    lib().oracle_current(C.byref(params), _p(cores), cores.shape[0], float(x), float(y), _p(out))
    return out


def qhuber(z, tau, z_next, r, done, gamma=0.99, kappa=1.0):
    z, tau, z_next, r, done = (np.ascontiguousarray(a, np.float64) for a in (z, tau, z_next, r, done))
    B, N = z.shape
    Np = z_next.shape[1]
    lps, grad, loss = np.zeros(B), np.zeros((B, N)), np.zeros(1)
    lib().oracle_qhuber(_p(z), _p(tau), _p(z_next), _p(r.reshape(-1)), _p(done.reshape(-1)), gamma,
                        kappa, B, N, Np, _p(lps), _p(grad), _p(loss))
    return float(loss[0]), lps, grad


def tau_cos(tau, n_cos=64):
    tau = np.ascontiguousarray(tau, np.float64)
    B, N = tau.shape
    out = np.zeros((B, N, n_cos))
    lib().oracle_tau_cos(_p(tau), B, N, n_cos, _p(out))
    return out


def philox(ctr, key):
    ctr = np.ascontiguousarray(ctr, np.uint32)
    key = np.ascontiguousarray(key, np.uint32)
    out = np.zeros(4, np.uint32)
    lib().oracle_philox4x32_10(_p(ctr), _p(key), _p(out))
    return out


def _arr(a, n=4):
    a = np.ascontiguousarray(a, np.float64).reshape(-1, n) if a is not None and len(a) else np.zeros((0, n))
    return a


def check_start_goal(rp, sg, cand, min_dis):
    """ENV:358-376 -> (accept, margin)."""
    sg, cand = _arr(sg), np.ascontiguousarray(cand, np.float64)
    m = np.array([np.inf])
    ok = lib().oracle_check_start_goal(C.byref(rp), _p(sg), len(sg), _p(cand), float(min_dis), _p(m))
    return bool(ok), float(m[0])


def check_core(rp, sg, cores, cand):
    """ENV:378-418 -> (accept, margin)."""
    sg, cores, cand = _arr(sg), _arr(cores), np.ascontiguousarray(cand, np.float64)
    m = np.array([np.inf])
    ok = lib().oracle_check_core(C.byref(rp), _p(sg), len(sg), _p(cores), len(cores), _p(cand), _p(m))
    return bool(ok), float(m[0])


def check_obstacle(rp, sg, cores, obs, cand):
    """ENV:420-456 -> (accept, margin)."""
    sg, cores, obs = _arr(sg), _arr(cores), _arr(obs)
    cand = np.ascontiguousarray(cand, np.float64)
    m = np.array([np.inf])
    ok = lib().oracle_check_obstacle(C.byref(rp), _p(sg), len(sg), _p(cores), len(cores), _p(obs),
                                     len(obs), _p(cand), _p(m))
    return bool(ok), float(m[0])


def env_reset(rp, st, reset_mask, want_robots, want_cores, want_obs, min_sg, seed, episode,
              max_obs, max_cores, env_offset=0):
    """ENV:72-176 over the Philox candidate sequence, IN PLACE on the float64 state dict;
    returns the per-env smallest predicate margin."""
    E, RP = st["x"].shape
    cs = _State(**{k: _p(v) for k, v in st.items()})
    i32 = lambda a: np.ascontiguousarray(a, np.int32)
    wr, wc, wo = i32(want_robots), i32(want_cores), i32(want_obs)
    ms = np.ascontiguousarray(min_sg, np.float64)
    mask = None if reset_mask is None else np.ascontiguousarray(reset_mask, np.uint8)
    margin = np.full((E,), np.inf)
    rc = lib().oracle_env_reset(C.byref(rp), C.byref(cs), _p(mask), _p(wr), _p(wc), _p(wo), _p(ms),
                                seed, episode, E, RP, max_obs, max_cores, env_offset, _p(margin))
    if rc != 0:
        raise RuntimeError("oracle_env_reset failed")
    return margin
