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


def test_reset_predicates_against_reference(oracle):
    """check_start_and_goal / check_core / check_obstacle (ENV:358-456): the reference's own
    answers on 400 random (accepted set, candidate) cases, many near a threshold."""
    d = gu.load("reset_checks.npz")
    rp = _abi.default_reset_params()
    for i in range(len(d["n_r"])):
        nr, nc, no = int(d["n_r"][i]), int(d["n_c"][i]), int(d["n_o"][i])
        sg, cores, obs = d["sg"][i, :nr], d["cores"][i, :nc], d["obs"][i, :no]
        assert oracle.check_start_goal(rp, sg, d["c_sg"][i], float(d["min_dis"][i]))[0] == bool(d["a_sg"][i]), i
        assert oracle.check_core(rp, sg, cores, d["c_core"][i])[0] == bool(d["a_core"][i]), i
        assert oracle.check_obstacle(rp, sg, cores, obs, d["c_obs"][i])[0] == bool(d["a_obs"][i]), i


def test_reset_against_reference_reset(oracle):
    """The unmodified reference reset() (ENV:72-176), fed the Philox candidate stream through a
    stand-in for its RandomState, against oracle_env_reset: same counts, same scenario."""
    d = gu.load("env_reset.npz")
    rp = _abi.default_reset_params()
    E = len(d["min_sg"])
    RP, OP, CP = int(d["RP"]), int(d["OP"]), int(d["CP"])
    st = oracle.new_state(E, RP, OP, CP)
    for v in st.values():
        v[...] = 7  # every field of a reset env must be overwritten
    oracle.env_reset(rp, st, None, d["want"][:, 0], d["want"][:, 1], d["want"][:, 2], d["min_sg"],
                     int(d["seed"]), int(d["episode"]), OP, CP, env_offset=int(d["env_offset"]))
    for k in ("n_robots", "n_cores", "n_obs"):
        assert np.array_equal(st[k], d["ref_" + k]), k
    assert (st["n_obs"] < d["want"][:, 2]).any()  # the 500-try cap is exercised
    for k in ("x", "y", "theta", "goal_x", "goal_y", "obstacles", "cores"):
        assert np.allclose(st[k], d["ref_" + k], rtol=1e-13, atol=1e-13), k
    for k in ("velr_x", "velr_y", "velr_r", "vel_x", "vel_y", "vel_r", "thrust_l", "thrust_r", "flags", "ep_steps"):
        assert not st[k].any(), k


def test_reset_mask_and_offset(oracle):
    rp = _abi.default_reset_params()
    E, RP, OP, CP = 12, 4, 5, 3
    want = (np.full(E, 4), np.full(E, 3), np.full(E, 5), np.full(E, 30.0))
    a = oracle.new_state(E, RP, OP, CP)
    oracle.env_reset(rp, a, None, *want, 5, 9, OP, CP, env_offset=100)
    b = oracle.new_state(4, RP, OP, CP)
    oracle.env_reset(rp, b, None, *(w[:4] for w in want), 5, 9, OP, CP, env_offset=104)
    for k in a:
        assert np.array_equal(a[k][4:8], b[k]), k  # streams follow the GLOBAL env index
    c = {k: v.copy() for k, v in a.items()}
    mask = (np.arange(E) % 2).astype(np.uint8)
    oracle.env_reset(rp, c, mask, *want, 5, 10, OP, CP, env_offset=100)
    assert np.array_equal(c["x"][mask == 0], a["x"][mask == 0])
    assert not np.array_equal(c["x"][mask == 1], a["x"][mask == 1])


def test_replay_ring_against_reference_buffer():
    """ReplayBuffer (RB:6-69): deque(maxlen) content order, state_batch padding and masks."""
    from oracle import replay_oracle as ro
    d = gu.load("replay.npz")
    n = len(d["rew"])
    t = np.arange(5)[None, :]
    mask = (t < d["n_obj"][:, None]).astype(np.float32)
    nmask = (t < d["n_obj2"][:, None]).astype(np.float32)
    obj = d["obj_s"] * mask[:, :, None]     # the env emits zero rows beyond the kept objects
    nobj = d["obj_n"] * nmask[:, :, None]
    status = np.where(d["done"], _abi.S_DONE, 0).astype(np.uint32)
    ring = ro.Ring(32)
    for lo in range(0, n, 8):  # five "batched steps" of 8 ASVs each
        sl = slice(lo, lo + 8)
        rows = ro.make_rows(d["self_s"][sl], obj[sl], mask[sl], d["act"][sl], False, d["rew"][sl],
                            d["self_n"][sl], nobj[sl], nmask[sl], status[sl])
        ring.push(rows, ro.valid_rows(status[sl]))
    assert ring.size == 32 and ring.total == n
    order = [(ring.total + i) % 32 for i in range(32)]  # oldest -> newest, like list(deque)
    got = ring.gather(order)
    for k, ref in (("self", "mem_self"), ("obj", "mem_obj"), ("mask", "mem_mask"), ("nself", "mem_nself"),
                   ("nobj", "mem_nobj"), ("nmask", "mem_nmask"), ("act", "mem_act"), ("rew", "mem_rew")):
        assert np.array_equal(got[k], d[ref].astype(np.float32)), k
    assert np.array_equal(got["done"], d["mem_done"].astype(np.float32))
    s, o, m = ro.state_batch_rows([(list(d["self_s"][i]), [list(x) for x in d["obj_s"][i, :d["n_obj"][i]]]) for i in range(8, 40)])
    assert np.array_equal(s, d["mem_self"].astype(np.float32)) and np.array_equal(m, d["mem_mask"].astype(np.float32))


def _learner_case(dtype):
    import os
    import torch
    from oracle import learner_oracle as lo
    z = np.load(os.path.join(gu.GOLD, "agent_ac_iqn.npz"))
    tt = lambda a: torch.from_numpy(np.asarray(a)).to(dtype)
    actor = {k[len("init_actor."):]: tt(z[k]) for k in z.files if k.startswith("init_actor.")}
    critic = {k[len("init_critic."):]: tt(z[k]) for k in z.files if k.startswith("init_critic.")}
    batch = {k[2:]: tt(z[k]) for k in z.files if k.startswith("b_")}
    taus = {k[4:]: tt(z[k]) for k in z.files if k.startswith("tau_")}
    return lo, z, actor, critic, batch, taus


def test_learner_restatement_forward_against_reference():
    """oracle/learner_oracle.py: Actor / Critic forward == the reference modules' outputs recorded in
    the fixture (same weights, states, actions and tau table)."""
    import torch
    lo, z, actor, critic, batch, taus = _learner_case(torch.float32)
    s = (batch["self"], batch["obj"], batch["mask"])
    with torch.no_grad():
        a = lo.actor_forward(actor, s).numpy()
        q = lo.critic_forward(critic, s, batch["act"], taus["local"]).numpy()
    assert np.allclose(a, z["actor_out"], rtol=2e-5, atol=2e-6)
    assert np.allclose(q, z["critic_out"], rtol=2e-5, atol=2e-5)


@pytest.mark.parametrize("dtype", ["float32", "float64"])
def test_learner_restatement_update_against_reference_train(dtype):
    """One train_AC_IQN step (AGENT:386-432) from the recorded weights, batch and taus: both losses
    and every weight after the two clipped Adam steps equal the reference's."""
    import torch
    lo, z, actor, critic, batch, taus = _learner_case(getattr(torch, dtype))
    ta, tc = {k: v.clone() for k, v in actor.items()}, {k: v.clone() for k, v in critic.items()}
    cl, al = lo.train_ac_iqn(actor, critic, ta, tc, batch, taus, lo.new_adam_state(actor), lo.new_adam_state(critic))
    assert abs(cl - float(z["critic_loss"])) <= 2e-5 * abs(float(z["critic_loss"])) + 1e-6
    assert abs(al - float(z["actor_loss"])) <= 2e-5 * abs(float(z["actor_loss"])) + 1e-6
    for name, w in (("actor", actor), ("critic", critic)):
        for k, v in w.items():
            ref, init = z[f"post_{name}.{k}"], z[f"init_{name}.{k}"]
            # the first Adam step is lr * g / (|g| + eps): ~ +-1e-4 for every touched weight, and
            # unstable where |g| ~ eps (a sign flip from summation order moves the weight by 2 lr).
            # Compare the step itself: every element within 2 lr, 99.5 % within 2e-6.
            err = np.abs((v.numpy() - init) - (ref - init))
            assert err.max() <= 2.1e-4 and np.mean(err <= 2e-6) >= 0.995, (name, k, err.max(), np.mean(err <= 2e-6))
