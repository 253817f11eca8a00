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
    ok = bool(np.all(err <= bound))
    if not ok:
        i = np.unravel_index(np.argmax(err - bound), err.shape)
        worst = (worst, "at", tuple(int(x) for x in i), "got", float(a[i]), "want", float(b[i]), "row", np.round(b[i[:-1]], 6).tolist() if b.ndim > 1 else None)
    return ok, worst


# ----------------------------------------------------------------------------------------
# fp32 CUDA path against a float64 truth (the reference's golden outputs or the oracle)
# ----------------------------------------------------------------------------------------
RTOL = 1e-5          # north_star: "within a stated fp32 tolerance (e.g. 1e-5 relative)"
ATOL = 1e-6          # absolute floor, as BASELINE.md section 5 states it
EPS_PERC = 1e-5      # metres: detection / collision / goal / top-5 order decided by less -> not compared
EPS_COLREGS = 2e-4   # COLREGs predicates are evaluated on fp32 observations

VEC_GROUPS_SELF = ((0, 2), (2, 4))           # goal xy, velocity xy are 2-vectors; rest scalars
VEC_GROUPS_OBJ = ((0, 2), (2, 4))


def _grouped_close(a, b, groups, rtol=RTOL, atol=ATOL):
    """last axis holds fields; `groups` are slices compared relative to their vector norm."""
    ok_all, worst_all = True, 0.0
    covered = set()
    for lo, hi in groups:
        ok, worst = vec_close(a[..., lo:hi], b[..., lo:hi], rtol, atol, -1)
        if not ok:
            return False, ("group", lo, hi, worst)
        worst_all = max(worst_all, worst)
        covered.update(range(lo, hi))
    rest = [i for i in range(a.shape[-1]) if i not in covered]
    if rest:
        ok, worst = vec_close(a[..., rest], b[..., rest], rtol, atol, None)
        if not ok:
            return False, ("scalars", rest, worst)
        worst_all = max(worst_all, worst)
    return ok_all, worst_all


def compare_env(got_out, got_st, ref_out, ref_st, margin_perc, margin_colregs, rows=None):
    """Asserts parity of one batched step.  `rows` optionally restricts to env indices.
    Returns a dict of statistics (worst relative errors, number of ASVs skipped by the margin
    rule)."""
    def sel(v):
        return np.asarray(v) if rows is None else np.asarray(v)[rows]

    stats = {}
    status_r = sel(ref_out["status"]).astype(np.int64)
    active = ((status_r >> 8) & 15) != 7
    active &= (status_r & 8) == 0
    mp, mc = sel(margin_perc), sel(margin_colregs)
    sure_p = (mp > EPS_PERC) | ~active
    sure_c = (mc > EPS_COLREGS) | ~active
    sure = sure_p & sure_c
    stats["asv"] = int(active.sum())
    stats["skipped_perc"] = int((~sure_p).sum())
    stats["skipped_colregs"] = int((sure_p & ~sure_c).sum())
    # --- bit-exact: indices and flags ---
    for k in ("top_ids", "detected"):
        if k in got_out:
            g, r = sel(got_out[k]), sel(ref_out[k])
            assert np.array_equal(g[sure_p], r[sure_p]), k
    g, r = sel(got_out["status"]).astype(np.int64), status_r
    assert np.array_equal(g[sure], r[sure]), "status"
    no_colregs = ~np.int64(16)
    assert np.array_equal(g[sure_p] & no_colregs & 0xFF, r[sure_p] & no_colregs & 0xFF), "status w/o COLREGs"
    assert np.array_equal(sel(got_out["obs_mask"])[sure_p], sel(ref_out["obs_mask"])[sure_p])
    gf, rf = sel(got_st["flags"]).astype(np.int64), sel(ref_st["flags"]).astype(np.int64)
    assert np.array_equal(gf[sure_p], rf[sure_p]), "flags"
    assert np.array_equal(sel(got_st["ep_steps"]), sel(ref_st["ep_steps"]))
    env_sure = sure_p.all(axis=1)
    assert np.array_equal(sel(got_out["env_status"])[env_sure], sel(ref_out["env_status"])[env_sure])
    # --- floating point ---
    ok, w = _grouped_close(sel(got_out["obs_self"])[sure_p], sel(ref_out["obs_self"])[sure_p], VEC_GROUPS_SELF)
    assert ok, ("obs_self", w)
    stats["obs_self"] = w
    ok, w = _grouped_close(sel(got_out["obs_objects"])[sure_p], sel(ref_out["obs_objects"])[sure_p], VEC_GROUPS_OBJ)
    assert ok, ("obs_objects", w)
    stats["obs_objects"] = w
    ok, w = vec_close(sel(got_out["reward"])[sure], sel(ref_out["reward"])[sure], RTOL, ATOL, None)
    assert ok, ("reward", w)
    stats["reward"] = w
    ok, w = vec_close(sel(got_out["phi"])[sure], sel(ref_out["phi"])[sure], RTOL, 2e-6, None)
    assert ok, ("phi", w)
    stats["phi"] = w
    pos_g = np.stack([sel(got_st["x"]), sel(got_st["y"])], -1)
    pos_r = np.stack([sel(ref_st["x"]), sel(ref_st["y"])], -1)
    ok, w = vec_close(pos_g, pos_r, RTOL, ATOL, -1)
    assert ok, ("position", w)
    stats["position"] = w
    for a, b in (("velr_x", "velr_y"), ("vel_x", "vel_y")):
        vg = np.stack([sel(got_st[a]), sel(got_st[b])], -1)
        vr = np.stack([sel(ref_st[a]), sel(ref_st[b])], -1)
        ok, w = vec_close(vg, vr, RTOL, ATOL, -1)
        assert ok, (a, w)
        stats[a[:-2]] = w
    for k in ("theta", "velr_r", "vel_r", "thrust_l", "thrust_r"):
        ok, w = vec_close(sel(got_st[k]), sel(ref_st[k]), RTOL, ATOL, None)
        assert ok, (k, w)
        stats[k] = w
    return stats


def to_device_state(st, device):
    """float64 numpy state dict -> float32 CUDA tensors (layout of _lib.new_state)."""
    import torch
    out = {}
    for k, v in st.items():
        if k == "flags":
            out[k] = torch.from_numpy(v.astype(np.int32)).to(device)
        elif v.dtype == np.float64:
            out[k] = torch.from_numpy(v.astype(np.float32)).to(device).contiguous()
        else:
            out[k] = torch.from_numpy(v.copy()).to(device)
    return out


def to_host(d):
    out = {}
    for k, v in d.items():
        a = v.detach().cpu().numpy()
        if k in ("flags", "status", "env_status"):
            a = a.astype(np.uint32) if a.dtype == np.int32 else a
        if k == "detected":
            a = a.view(np.uint64)
        out[k] = a
    return out
