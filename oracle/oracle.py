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
