"""K1b (scenario re-generation) and K2 (device replay ring) parity on the B200, through the
C-ABI, against the CPU oracles (pinned to the reference by tests/test_oracle_golden.py) and the
reference-generated fixtures tests/golden/env_reset.npz and replay.npz."""
import numpy as np
import pytest
import torch

from distributional_rl_decision_and_control_b200 import _abi, _lib
from oracle import replay_oracle as ro
import golden_util as gu

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
EPS = 1e-4  # metres: acceptance decisions with a float64 margin below this are not compared


def gpu_reset(E, RP, OP, CP, want, min_sg, seed, episode, off=0, mask=None, st=None):
    rp = _abi.default_reset_params()
    st = _lib.new_state(E, RP, OP, CP, DEV) if st is None else st
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a, dt)).to(DEV)
    _lib.env_reset(rp, st, None if mask is None else t(mask, np.uint8), t(want[0], np.int32),
                   t(want[1], np.int32), t(want[2], np.int32), t(min_sg, np.float32), seed, episode,
                   OP, CP, env_offset=off)
    torch.cuda.synchronize()
    return st


def compare_reset(got, want, margin, sel=None):
    """Envs whose every acceptance decision had a float64 margin > EPS must agree exactly in
    counts and to fp32 rounding in coordinates."""
    E = len(margin)
    ok = margin > EPS
    if sel is not None:
        ok &= sel
    for k in ("n_robots", "n_cores", "n_obs", "ep_steps"):
        assert np.array_equal(got[k][ok], want[k][ok]), k
    assert np.array_equal(got["flags"][ok], want["flags"][ok])
    for k in _abi.STATE_FLOAT_FIELDS + ("obstacles", "cores"):
        assert np.allclose(got[k][ok], want[k][ok], rtol=1e-6, atol=5e-6), k
    return int(E - ok.sum())


def test_reset_against_reference_generated_fixture():
    d = gu.load("env_reset.npz")
    E, RP, OP, CP = len(d["min_sg"]), int(d["RP"]), int(d["OP"]), int(d["CP"])
    st = gpu_reset(E, RP, OP, CP, (d["want"][:, 0], d["want"][:, 1], d["want"][:, 2]), d["min_sg"],
                   int(d["seed"]), int(d["episode"]), off=int(d["env_offset"]))
    got = gu.to_host(st)
    want = {k[4:]: v for k, v in d.items() if k.startswith("ref_")}
    for k in ("n_robots", "n_cores", "n_obs"):
        assert np.array_equal(got[k], want[k]), k
    for k in ("x", "y", "theta", "goal_x", "goal_y", "obstacles", "cores"):
        assert np.allclose(got[k], want[k], rtol=1e-6, atol=5e-6), k


@pytest.mark.parametrize("E,RP,OP,CP,wr,wc,wo", [(4096, 8, 10, 8, 8, 8, 10), (257, 3, 0, 0, 3, 0, 0),
                                                 (100, 6, 6, 4, 5, 4, 4), (64, 32, 32, 16, 20, 16, 32),
                                                 (33, 8, 10, 8, 0, 0, 0)])
def test_reset_oracle(oracle, E, RP, OP, CP, wr, wc, wo):
    rp = _abi.default_reset_params()
    rng = np.random.RandomState(E)
    want = (np.full(E, wr), np.full(E, wc), np.full(E, wo))
    min_sg = rng.choice([30.0, 35.0, 40.0], E)
    st_o = oracle.new_state(E, RP, OP, CP)
    margin = oracle.env_reset(rp, st_o, None, *want, min_sg, 77, 5, OP, CP, env_offset=12345)
    got = gu.to_host(gpu_reset(E, RP, OP, CP, want, min_sg, 77, 5, off=12345))
    skipped = compare_reset(got, st_o, margin)
    # every env makes up to 3 x 500 threshold decisions; P(|margin| < EPS) ~ 1e-5 each
    assert skipped <= max(2, E // 10)
    # reference invariants on the generated scenarios (ENV:358-456)
    nr = got["n_robots"]
    for e in range(min(E, 200)):
        s = np.stack([got["x"][e, :nr[e]], got["y"][e, :nr[e]]], 1)
        g = np.stack([got["goal_x"][e, :nr[e]], got["goal_y"][e, :nr[e]]], 1)
        assert (np.linalg.norm(s - g, axis=1) >= min_sg[e] - 1e-4).all()
        for i in range(nr[e]):
            for j in range(i):
                assert np.linalg.norm(s[i] - s[j]) > 10.0 - 1e-4 and np.linalg.norm(g[i] - g[j]) > 10.0 - 1e-4
        ob = got["obstacles"][e, :got["n_obs"][e]]
        for o in ob:
            assert (np.linalg.norm(s - o[:2], axis=1) >= o[2] + 10.0 - 1e-4).all()
            assert (np.linalg.norm(g - o[:2], axis=1) >= o[2] + 10.0 - 1e-4).all()


def test_reset_mask_leaves_other_envs_untouched(oracle):
    E, RP, OP, CP = 300, 8, 10, 8
    want = (np.full(E, 8), np.full(E, 8), np.full(E, 10))
    min_sg = np.full(E, 30.0)
    st = gpu_reset(E, RP, OP, CP, want, min_sg, 1, 1)
    before = gu.to_host(st)
    mask = (np.arange(E) % 5 == 0).astype(np.uint8)
    gpu_reset(E, RP, OP, CP, want, min_sg, 1, 2, mask=mask, st=st)
    after = gu.to_host(st)
    off = mask == 0
    for k in before:
        assert np.array_equal(before[k][off], after[k][off]), k
    assert not np.array_equal(before["x"][~off], after["x"][~off])
    # the masked re-generation equals the oracle's
    st_o = {k: v.astype(np.float64) if v.dtype == np.float32 else v.copy() for k, v in before.items()}
    rp = _abi.default_reset_params()
    margin = oracle.env_reset(rp, st_o, mask, *want, min_sg, 1, 2, OP, CP)
    compare_reset(after, st_o, margin, sel=~off)


# ------------------------------------------------------------------------------------ K2
def _t(a, dt=np.float32):
    return torch.from_numpy(np.ascontiguousarray(a, dt)).to(DEV)


def random_step_outputs(rng, n):
    n_obj, n_obj2 = rng.randint(0, 6, n), rng.randint(0, 6, n)
    t = np.arange(5)[None, :]
    mask, nmask = (t < n_obj[:, None]).astype(np.float32), (t < n_obj2[:, None]).astype(np.float32)
    d = {"self": rng.randn(n, 7).astype(np.float32), "nself": rng.randn(n, 7).astype(np.float32),
         "obj": rng.randn(n, 5, 5).astype(np.float32) * mask[:, :, None],
         "nobj": rng.randn(n, 5, 5).astype(np.float32) * nmask[:, :, None], "mask": mask, "nmask": nmask,
         "act": (rng.rand(n, 2) * 2 - 1).astype(np.float32), "act_id": rng.randint(0, 25, n).astype(np.int32),
         "rew": rng.randn(n).astype(np.float32)}
    kind = rng.rand(n)
    status = np.zeros(n, np.uint32)
    status[kind < 0.15] = _abi.S_DONE | _abi.S_WAS_DEACTIVATED | (_abi.INFO_DEACT_GOAL << _abi.S_INFO_SHIFT)
    status[(kind >= 0.15) & (kind < 0.25)] = _abi.INFO_ABSENT << _abi.S_INFO_SHIFT
    status[(kind >= 0.25) & (kind < 0.4)] = _abi.S_DONE | (_abi.INFO_COLLISION << _abi.S_INFO_SHIFT)
    d["status"] = status
    return d


def push_both(rb, ring, d, discrete=False):
    act = d["act_id"] if discrete else d["act"]
    _lib.replay_push(rb, _t(d["self"]), _t(d["obj"]), _t(d["mask"]), _t(act, np.int32 if discrete else np.float32),
                     discrete, _t(d["rew"]), _t(d["nself"]), _t(d["nobj"]), _t(d["nmask"]),
                     _t(d["status"].astype(np.int32), np.int32))
    rows = ro.make_rows(d["self"], d["obj"], d["mask"], act, discrete, d["rew"], d["nself"], d["nobj"],
                        d["nmask"], d["status"])
    ring.push(rows, ro.valid_rows(d["status"]))


@pytest.mark.parametrize("capacity,n,steps,discrete", [(1000, 300, 9, False), (5000, 1, 40, False),
                                                        (70000, 32768, 3, False), (513, 513, 4, True)])
def test_push_matches_ring_oracle(capacity, n, steps, discrete):
    rng = np.random.RandomState(capacity)
    rb = _lib.new_replay(capacity, n, DEV)
    ring = ro.Ring(capacity)
    for s in range(steps):
        push_both(rb, ring, random_step_outputs(rng, n), discrete)
    torch.cuda.synchronize()
    hs = rb["head_size"].cpu().numpy()
    assert int(hs[0]) == ring.total and int(hs[1]) == ring.size
    assert np.array_equal(rb["rows"].cpu().numpy(), ring.rows)  # bit-exact, incl. wrap-around
    assert int(rb["workspace"][0].item()) == 0  # ticket left clean


def test_push_reference_fixture_then_gather():
    """The reference ReplayBuffer's deque content (tests/golden/replay.npz) through push+sample."""
    d = gu.load("replay.npz")
    n = len(d["rew"])
    t = np.arange(5)[None, :]
    mask = (t < d["n_obj"][:, None]).astype(np.float32)
    nmask = (t < d["n_obj2"][:, None]).astype(np.float32)
    status = np.where(d["done"], _abi.S_DONE, 0).astype(np.int32)
    rb = _lib.new_replay(32, 8, DEV)
    for lo in range(0, n, 8):
        sl = slice(lo, lo + 8)
        _lib.replay_push(rb, _t(d["self_s"][sl]), _t(d["obj_s"][sl] * mask[sl, :, None]), _t(mask[sl]),
                         _t(d["act"][sl]), False, _t(d["rew"][sl]), _t(d["self_n"][sl]),
                         _t(d["obj_n"][sl] * nmask[sl, :, None]), _t(nmask[sl]), _t(status[sl], np.int32))
    order = torch.tensor([(n + i) % 32 for i in range(32)], dtype=torch.int64, device=DEV)
    b = _lib.replay_sample(rb, _lib.new_batch(32, DEV), indices=order)
    torch.cuda.synchronize()
    for k, ref in (("self", "mem_self"), ("obj", "mem_obj"), ("mask", "mem_mask"), ("nself", "mem_nself"),
                   ("nobj", "mem_nobj"), ("nmask", "mem_nmask"), ("act", "mem_act"), ("rew", "mem_rew"),
                   ("done", "mem_done")):
        assert np.array_equal(b[k].cpu().numpy(), d[ref].astype(np.float32)), k


@pytest.mark.parametrize("capacity,filled,B", [(100000, 100000, 1024), (5000, 777, 64), (64, 64, 64), (10, 3, 8)])
def test_sample_distinct_uniform_and_gather(capacity, filled, B):
    rng = np.random.RandomState(B)
    rb = _lib.new_replay(capacity, filled, DEV)
    ring = ro.Ring(capacity)
    d = random_step_outputs(rng, filled)
    d["status"][:] = 0
    push_both(rb, ring, d)
    batch = _lib.new_batch(B, DEV)
    seen = []
    for c in range(20):
        _lib.replay_sample(rb, batch, seed=9, counter=c)
        idx = batch["indices"].cpu().numpy()
        assert idx.min() >= 0 and idx.max() < ring.size
        if B <= ring.size:
            assert len(np.unique(idx)) == B  # random.sample draws WITHOUT replacement (RB:28)
        want = ring.gather(idx)
        for k in want:
            assert np.array_equal(batch[k].cpu().numpy(), want[k]), k
        seen.append(idx.copy())
    again = _lib.replay_sample(rb, batch, seed=9, counter=3)["indices"].cpu().numpy()
    assert np.array_equal(again, seen[3])
    if ring.size >= 64:
        assert not np.array_equal(seen[3], seen[4])
    if ring.size >= 50000:
        from scipy import stats
        allidx = np.concatenate(seen).astype(np.float64) / ring.size
        assert stats.kstest(allidx, "uniform").pvalue > 1e-3


def test_push_rejects_bad_arguments():
    rb = _lib.new_replay(4, 8, DEV)
    z = torch.zeros(8, 7, device=DEV)
    with pytest.raises(RuntimeError, match="exceeds capacity"):
        _lib.replay_push(rb, z, torch.zeros(8, 5, 5, device=DEV), torch.zeros(8, 5, device=DEV),
                         torch.zeros(8, 2, device=DEV), False, torch.zeros(8, device=DEV), z,
                         torch.zeros(8, 5, 5, device=DEV), torch.zeros(8, 5, device=DEV),
                         torch.zeros(8, dtype=torch.int32, device=DEV))
