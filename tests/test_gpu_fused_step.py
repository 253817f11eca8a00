"""The fused step (asvrl_env_step_autoreset: K1 + K1b + first observation in one launch, with or
without pre-generated next episodes) against the call sequence include/asvrl.h defines it by
(asvrl_env_step, then per ended env asvrl_env_reset + masked asvrl_env_step with the env's own
episode number), which is itself parity-checked against the oracle in test_gpu_env_step.py /
test_gpu_reset_replay.py.  Everything is compared BIT-exactly: same arithmetic, same Philox counters."""
import numpy as np
import pytest
import torch

from distributional_rl_decision_and_control_b200 import _abi, _lib
from distributional_rl_decision_and_control_b200.envs.marinenav.batched import BatchedMarineNavEnv
from distributional_rl_decision_and_control_b200.utils.replay_buffer import ReplayBuffer

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


class CallSequence:
    """The definition of the fused step, spelled out with the separate entry points."""

    def __init__(self, E, R, O, C, seed):
        self.env = BatchedMarineNavEnv(E, R, O, C, device=DEV, seed=seed, fused=False)
        self.env.reset()
        self.idx = np.full((E,), self.env.episode_counter, dtype=np.int64)

    def step(self, act, discrete=False):
        env = self.env
        out = env.step_no_reset(act, is_continuous_action=not discrete)
        trans = {k: out[k].clone() for k in ("reward", "status", "phi", "env_status", "ended",
                                             "obs_self", "obs_objects", "obs_mask")}
        for e in np.nonzero(out["ended"].cpu().numpy())[0]:
            self.idx[e] += 1
            mask = torch.zeros((env.E,), dtype=torch.uint8, device=DEV)
            mask[e] = 1
            n = int(self.idx[e])
            _lib.env_reset(env.reset_params, env.state, mask, env.want_robots, env.want_cores, env.want_obs,
                           env.min_start_goal_dis, env.seed, n, env.OP, env.CP, env.env_offset)
            _lib.env_step(env.params, env.state, None, None, out, 0, env.OP, env.CP,
                          seed=env.seed ^ env.OBSERVE_SALT, step=n, env_offset=env.env_offset, env_mask=mask)
        return trans, out


def make_pair(E, R, O, C, seed, prefetch="stream", spread=13):
    """A fused env and the call-sequence emulation in the same state; episode clocks are set so
    that timeouts happen in every one of the next `spread` steps."""
    a = BatchedMarineNavEnv(E, R, O, C, device=DEV, seed=seed, fused=True, prefetch=prefetch)
    a.reset()
    b = CallSequence(E, R, O, C, seed)
    clocks = 1000 - (torch.arange(E, device=DEV, dtype=torch.int32) % spread)
    clocks[1::3] = 0
    for env in (a, b.env):
        env.state["ep_steps"].copy_(clocks)
    return a, b


def same(x, y):
    return torch.equal(x, y)


@pytest.mark.parametrize("prefetch", [False, "workers", "stream"])
@pytest.mark.parametrize("E,R,O,C,discrete", [(512, 8, 10, 8, False), (130, 3, 0, 0, False),
                                              (96, 5, 6, 2, True), (33, 1, 4, 0, False)])
def test_fused_equals_call_sequence(E, R, O, C, discrete, prefetch):
    a, b = make_pair(E, R, O, C, seed=11, prefetch=prefetch)
    g = torch.Generator(device=DEV).manual_seed(5)
    n_ended = n_staged = 0
    for t in range(16):
        if discrete:
            act = torch.randint(0, 25, (E, R), device=DEV, generator=g, dtype=torch.int32)
        else:
            act = torch.rand((E, R, 2), device=DEV, generator=g) * 2 - 1
        if prefetch:
            tag_before = a.stage_tag.clone()
            idx_before = a.episode_idx.clone()
        oa = a.step(act, is_continuous_action=not discrete)
        trans, ob = b.step(act, discrete)
        torch.cuda.synchronize()
        ended = trans["ended"].bool()
        n_ended += int(ended.sum())
        if prefetch:  # envs whose stage was ready when the launch began certainly took it
            n_staged += int((ended & (tag_before == idx_before + 1)).sum())
        for k in ("reward", "status", "phi", "env_status", "ended"):  # the transition's scalars are kept
            assert same(oa[k], trans[k]), (t, k)
        for k in ("obs_self", "obs_objects", "obs_mask"):  # what to act on next / the terminal rows
            assert same(oa[k], ob[k]), (t, k)
            assert same(a.term[k][ended], trans[k][ended]), (t, "term " + k)
        for k, v in a.state.items():
            assert same(v, b.env.state[k]), (t, "state " + k)
        assert np.array_equal(a.episode_idx.cpu().numpy(), b.idx), t
    assert n_ended > 0
    if prefetch:
        assert n_staged > 0
        # the stage of every env that did not just end holds its next episode
        torch.cuda.synchronize()
        ok = a.stage_tag == a.episode_idx + 1
        assert int(ok.sum()) >= E // 2


def test_schedule_change_invalidates_the_stage():
    E, R = 64, 8
    a, b = make_pair(E, R, 10, 4, seed=7, spread=5)
    g = torch.Generator(device=DEV).manual_seed(2)
    for t in range(10):
        if t == 3:  # curriculum stage switch (ENV:75-87): fewer robots and buoys from now on
            a.set_schedule(num_robots=3, num_obs=4, num_cores=2, min_start_goal_dis=20.0)
            b.env.set_schedule(num_robots=3, num_obs=4, num_cores=2, min_start_goal_dis=20.0)
        act = torch.rand((E, R, 2), device=DEV, generator=g) * 2 - 1
        a.step(act)
        b.step(act)
    torch.cuda.synchronize()
    for k, v in a.state.items():
        assert same(v, b.env.state[k]), k
    assert int((a.state["n_robots"] == 3).sum()) > 0


def test_fused_cuda_graph_equals_eager():
    E, R = 256, 8
    a, _ = make_pair(E, R, 10, 4, seed=3)
    c, _ = make_pair(E, R, 10, 4, seed=3)
    a.enable_cuda_graph(True)
    for _ in range(2):  # the two warm-up steps of enable_cuda_graph act on zeros
        c.step(torch.zeros((E, R, 2), device=DEV))
    g = torch.Generator(device=DEV).manual_seed(9)
    for t in range(12):
        act = torch.rand((E, R, 2), device=DEV, generator=g) * 2 - 1
        a.step_graph(act)
        c.step(act)
    torch.cuda.synchronize()
    for k, v in a.state.items():
        assert same(v, c.state[k]), k
    for k in ("obs_self", "obs_objects", "obs_mask", "reward", "status"):
        assert same(a.out[k], c.out[k]), k


def test_replay_push_reads_terminal_rows():
    """Ring contents after fused steps == after the call sequence with the push in between."""
    E, R = 200, 8
    a, b = make_pair(E, R, 10, 0, seed=21)
    ra = ReplayBuffer(50000, 64, 5, 5, device=DEV)
    rb = ReplayBuffer(50000, 64, 5, 5, device=DEV)
    g = torch.Generator(device=DEV).manual_seed(1)
    for t in range(8):
        act = torch.rand((E, R, 2), device=DEV, generator=g) * 2 - 1
        obs_a = tuple(x.clone() for x in a.obs)
        obs_b = tuple(x.clone() for x in b.env.obs)
        oa = a.step_fused(act)
        ra.push_step(obs_a, act, False, oa, term=a.term)
        trans, ob = b.step(act)
        rb.push_step(obs_b, act, False, trans)  # the clones taken before the re-generation
    torch.cuda.synchronize()
    assert same(ra.ring["head_size"], rb.ring["head_size"])
    n = int(ra.ring["head_size"][1])
    assert n > 0
    assert same(ra.ring["rows"][:n], rb.ring["rows"][:n])


def test_autoreset_rejects_bad_arguments():
    env = BatchedMarineNavEnv(8, 8, 10, 0, device=DEV, seed=0)
    env.reset()
    act = torch.zeros((8, 8, 2), device=DEV)
    with pytest.raises(RuntimeError):  # needs ASVRL_M_DYNAMICS
        _lib.env_step_autoreset(env.params, env.reset_params, env.state, act, env.out, 0, env.OP, env.CP,
                                env.want_robots, env.want_cores, env.want_obs, env.min_start_goal_dis,
                                env.episode_idx)
    with pytest.raises(RuntimeError):  # term without ended
        _lib.replay_push(_lib.new_replay(100, 64, DEV), *env.obs, act, False, env.out["reward"], *env.obs,
                         env.out["status"], term=env.term)


def test_host_pipelined_env_matches_plain_shards():
    """send / recv over two shards returns, per shard, exactly what BatchedMarineNavEnv.step produces."""
    from distributional_rl_decision_and_control_b200.envs.marinenav.host_pipeline import HostPipelinedEnv
    E, R, G = 64, 8, 2
    h = HostPipelinedEnv(E, R, 10, 4, groups=G, device=DEV, seed=5)
    plain = [BatchedMarineNavEnv(E // G, R, 10, 4, device=DEV, seed=5, env_offset=g * (E // G)) for g in range(G)]
    for p in plain:
        p.reset()
    for env in plain + h.shards:
        env.state["ep_steps"][::3] = 999
    torch.cuda.synchronize()
    gen = torch.Generator().manual_seed(3)
    for t in range(6):
        acts = [(torch.rand((E // G, R, 2), generator=gen) * 2 - 1).pin_memory() for _ in range(G)]
        for g in range(G):
            h.send(g, acts[g])
        for g in range(G):
            got = h.recv(g)
            want = plain[g].step(acts[g].to(DEV))
            torch.cuda.synchronize()
            for k in ("obs_self", "obs_objects", "reward", "status"):
                assert torch.equal(got[k], want[k].cpu()), (t, g, k)
            mask = (got["obs_objects"][..., 4] != 0).float()  # the mask the caller derives
            assert torch.equal(mask, want["obs_mask"].cpu()), (t, g)


@pytest.mark.parametrize("zero_copy", [False, True])
def test_host_pipelined_env_cuda_graph_matches_plain_shards(zero_copy):
    """The captured [H2D, fused step, D2H] graph of every shard returns what BatchedMarineNavEnv.step
    produces on the same actions (the two warm-up steps of the capture are replayed on the plain envs)."""
    from distributional_rl_decision_and_control_b200.envs.marinenav.host_pipeline import HostPipelinedEnv
    E, R, G = 64, 8, 2
    h = HostPipelinedEnv(E, R, 10, 4, groups=G, device=DEV, seed=5)
    plain = [BatchedMarineNavEnv(E // G, R, 10, 4, device=DEV, seed=5, env_offset=g * (E // G)) for g in range(G)]
    for p in plain:
        p.reset()
    for env in plain + h.shards:
        env.state["ep_steps"][::3] = 997
    torch.cuda.synchronize()
    h.enable_cuda_graph(zero_copy=zero_copy)  # zero_copy: the kernel stores straight into pinned host memory
    zero = torch.zeros((E // G, R, 2), device=DEV)
    for p in plain:  # enable_cuda_graph stepped every shard twice with zero actions
        p.step(zero)
        p.step(zero)
    gen = torch.Generator().manual_seed(3)
    for t in range(7):
        acts = [(torch.rand((E // G, R, 2), generator=gen) * 2 - 1).pin_memory() for _ in range(G)]
        for g in range(G):
            if t % 2:  # in place in the staging buffer, or copied into it by send()
                h.action_buffer(g).copy_(acts[g])
                h.send(g, h.action_buffer(g))
            else:
                h.send(g, acts[g])
        for g in range(G):
            got = h.recv(g)
            want = plain[g].step(acts[g].to(DEV))
            torch.cuda.synchronize()
            for k in ("obs_self", "obs_objects", "reward", "status"):
                assert torch.equal(got[k], want[k].cpu()), (t, g, k)


_SHAPE_PROBE = r"""
import hashlib, sys, torch
sys.path.insert(0, sys.argv[1])
from distributional_rl_decision_and_control_b200.envs.marinenav.batched import BatchedMarineNavEnv
C = int(sys.argv[2])
env = BatchedMarineNavEnv(64, 8, 10, C, device="cuda:0", seed=77, fused=True, prefetch="workers")
env.reset()
env.state["ep_steps"].copy_(1000 - (torch.arange(64, device="cuda:0", dtype=torch.int32) % 7))
g = torch.Generator(device="cuda:0").manual_seed(5)
h = hashlib.sha256()
for i in range(12):
    out = env.step_fused(torch.rand((64, 8, 2), device="cuda:0", generator=g) * 2 - 1)
    for k in ("obs_self", "obs_objects", "obs_mask", "reward", "status", "ended"):
        h.update(out[k].cpu().numpy().tobytes())
for k in sorted(env.state):
    if torch.is_tensor(env.state[k]):
        h.update(env.state[k].cpu().numpy().tobytes())
print("HASH", h.hexdigest())
"""


@pytest.mark.parametrize("cores", [8, 0])
def test_shape_specialised_kernel_equals_generic(cores):
    """env_step_kernel<.., SHAPE> for the BASELINE config-2 table sizes (8 ASVs, 10 buoys, 8 or 0 vortex
    cores) is the same arithmetic as the run-time-sized instantiation: 12 fused steps with resets,
    every output and the final state hashed, with and without ASVRL_K1_GENERIC=1."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    got = []
    for generic in ("0", "1"):
        env = dict(os.environ, ASVRL_K1_GENERIC=generic)
        r = subprocess.run([sys.executable, "-c", _SHAPE_PROBE, root, str(cores)], env=env, capture_output=True,
                           text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        got.append([l for l in r.stdout.splitlines() if l.startswith("HASH")][0])
    assert got[0] == got[1]
