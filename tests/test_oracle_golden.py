"""Pins the C oracle (oracle/asv_oracle.c) against the Python reference's own outputs
(tests/golden/*.npz, written by oracle/gen_golden.py).  Float64 against float64: flags and
indices bit-exact, floats to 1e-9 relative.  CPU only."""
import numpy as np
import pytest

from distributional_rl_decision_and_control_b200 import _abi
import golden_util as gu

RT, AT = 1e-9, 1e-10
FLOAT_STATE = _abi.STATE_FLOAT_FIELDS


def check_out(out, d, t=None, sel=None):
    def ref(k):
        v = d["out_" + k]
        v = v if t is None else v[t]
        return v if sel is None else v[sel]

    def got(k):
        return out[k] if sel is None else out[k][sel]

    for k in ("status", "top_ids", "detected", "env_status"):
        assert np.array_equal(got(k), ref(k)), k
    assert np.array_equal(got("obs_mask"), ref("obs_mask"))
    for k, axis in (("obs_self", None), ("obs_objects", None), ("reward", None), ("phi", None)):
        ok, worst = gu.vec_close(got(k), ref(k), RT, AT, axis)
        assert ok, (k, worst)


def check_state(st, d, t=None, sel=None):
    for k in FLOAT_STATE + ("flags", "ep_steps"):
        r = d["ref_" + k] if t is None else d["ref_" + k][t]
        a = st[k]
        if sel is not None:
            r, a = r[sel], a[sel]
        if k in ("flags", "ep_steps"):
            assert np.array_equal(a, r), k
        else:
            ok, worst = gu.vec_close(a, r, RT, AT, None)
            assert ok, (k, worst)


@pytest.mark.parametrize("name", ["env_step_cont.npz", "env_step_disc.npz", "env_step_c2shape.npz",
                                  "env_step_hull.npz", "env_perception_only.npz"])
def test_one_step(oracle, name):
    d = gu.load(name)
    p = gu.params_from(d["overrides"])
    st = gu.state_in(d)
    out = oracle.env_step(p, st, d.get("actions"), d["noise"].astype(np.float64), int(d["mode"]),
                          int(d["OP"]), int(d["CP"]))
    check_out(out, d)
    check_state(st, d)


def test_one_step_threads_equal_serial(oracle):
    d = gu.load("env_step_cont.npz")
    p = gu.params_from(d["overrides"])
    st = gu.state_in(d)
    out = oracle.env_step(p, st, d["actions"], d["noise"].astype(np.float64), int(d["mode"]),
                          int(d["OP"]), int(d["CP"]), n_threads=4)
    check_out(out, d)
    check_state(st, d)


def test_rollout(oracle):
    """24 consecutive steps from the oracle's OWN state track the reference's trajectory,
    including the Trainer's deactivation folded into the step."""
    d = gu.load("env_rollout.npz")
    p = gu.params_from(d["overrides"])
    st = gu.state_in(d)
    for t in range(int(d["T"])):
        alive = d["out_alive"][t]
        sub = {k: np.ascontiguousarray(v[alive]) for k, v in st.items()}
        out = oracle.env_step(p, sub, d["actions"][t][alive], d["noise"][t][alive].astype(np.float64),
                              int(d["mode"]), int(d["OP"]), int(d["CP"]))
        for k in st:
            st[k][alive] = sub[k]
        full = {k: np.zeros((alive.size,) + v.shape[1:], v.dtype) for k, v in out.items()}
        for k in out:
            full[k][alive] = out[k]
        check_out(full, d, t, alive)
        check_state(st, d, t, alive)


def test_current_field(oracle):
    d = gu.load("current_field.npz")
    p = _abi.default_params()
    for g in range(d["cores"].shape[0]):
        nc = int(d["n_cores"][g])
        for k in range(d["points"].shape[1]):
            v = oracle.current(p, d["cores"][g, :nc], d["points"][g, k, 0], d["points"][g, k, 1])
            ok, worst = gu.vec_close(v, d["velocity"][g, k], 1e-12, 1e-13)
            assert ok, (g, k, worst)


def test_iqn_loss(oracle):
    d = gu.load("iqn_loss.npz")
    loss, lps, grad = oracle.qhuber(d["z"], d["tau"], d["z_next"], d["r"], d["done"], float(d["gamma"]), 1.0)
    assert abs(loss - float(d["loss"])) <= 1e-12 * abs(float(d["loss"]))
    assert np.allclose(lps, d["loss_per_sample"], rtol=1e-12, atol=1e-14)
    assert np.allclose(grad, d["grad_z"], rtol=1e-12, atol=1e-16)


def test_tau_cos(oracle):
    d = gu.load("iqn_loss.npz")
    cos = oracle.tau_cos(d["tau"], 64)
    # the reference evaluates cosf in float32 (torch CPU): agree to float32 rounding of cos
    assert np.max(np.abs(cos - d["cos"])) <= 2.5e-7
    assert np.array_equal(d["pis"], (np.pi * np.arange(64)).astype(np.float32))


def test_philox_known_answers(oracle):
    """Random123 kat_vectors for philox4x32-10."""
    kat = [
        ((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
        ((0xffffffff,) * 4, (0xffffffff, 0xffffffff), (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
        ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
         (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
    ]
    for ctr, key, want in kat:
        assert tuple(int(x) for x in oracle.philox(ctr, key)) == want
