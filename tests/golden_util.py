"""Shared helpers of the parity tests: fixture loading and the comparison rules.

Tolerances (BASELINE.md section 5 / SURVEY.md 8d):
  * integer / flag / index outputs: bit-exact, except where the float64 oracle reports that a
    threshold predicate was decided by a margin below EPS_* (then the case is counted, not
    compared) — for the float64 oracle against the float64 reference no margin rule is used.
  * floating point: |a-b| <= rtol * max(|b|, scale_of_the_vector) + atol.
"""
import json
import os

import numpy as np

from distributional_rl_decision_and_control_b200 import _abi

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

STATE_KEYS = _abi.STATE_FLOAT_FIELDS + ("flags",) + _abi.STATE_ENV_INT_FIELDS + ("obstacles", "cores")


def load(name):
    z = np.load(os.path.join(GOLD, name), allow_pickle=False)
    d = {k: z[k] for k in z.files}
    d["overrides"] = json.loads(str(d["overrides"])) if "overrides" in d else {}
    return d


def params_from(overrides):
    p = _abi.default_params()
    for k, v in overrides.items():
        if isinstance(v, list):
            arr = getattr(p, k)
            for i, x in enumerate(v):
                arr[i] = x
        else:
            setattr(p, k, v)
    return p


def state_in(d, t=None):
    return {k: np.ascontiguousarray(d["in_" + k]).copy() for k in STATE_KEYS}


def vec_close(a, b, rtol, atol, axis=-1):
    """|a-b| <= rtol*max(|b|, ||b|| along axis) + atol; returns (ok, worst relative error)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    scale = np.abs(b)
    if axis is not None and b.ndim > 0:
        scale = np.maximum(scale, np.linalg.norm(b, axis=axis, keepdims=True))
    err = np.abs(a - b)
    bound = rtol * scale + atol
    worst = float(np.max(err / np.maximum(scale, atol / max(rtol, 1e-300)))) if err.size else 0.0
    return bool(np.all(err <= bound)), worst
