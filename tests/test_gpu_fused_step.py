"""The fused step (asvrl_env_step_autoreset: K1 + K1b + first observation in one launch) against
the three-call sequence it replaces (asvrl_env_step, asvrl_env_reset, masked asvrl_env_step), which
is itself parity-checked against the oracle in test_gpu_env_step.py / test_gpu_reset_replay.py.
Everything is compared BIT-exactly: same kernels' arithmetic, same Philox counters."""
import pytest
import torch

from distributional_rl_decision_and_control_b200 import _abi, _lib
from distributional_rl_decision_and_control_b200.envs.marinenav.batched import BatchedMarineNavEnv
from distributional_rl_decision_and_control_b200.utils.replay_buffer import ReplayBuffer

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def make_pair(E, R, O, C, seed, near_timeout=True):
    """Two envs in the same state, one fused, one not; some envs one step before the timeout."""
    a = BatchedMarineNavEnv(E, R, O, C, device=DEV, seed=seed, fused=True)
    b = BatchedMarineNavEnv(E, R, O, C, device=DEV, seed=seed, fused=False)
    a.reset()
    b.reset()
    if near_timeout:
        for env in (a, b):
            env.state["ep_steps"][::7] = 1000   # ENV:312: the next step ends these episodes
            env.state["ep_steps"][3::11] = 999  # and these one step later
    return a, b


def same(x, y):
    return torch.equal(x, y)


@pytest.mark.parametrize("E,R,O,C,discrete", [(512, 8, 10, 8, False), (130, 3, 0, 0, False),
                                              (96, 5, 6, 2, True), (33, 1, 4, 0, False)])
def test_fused_equals_three_calls(E, R, O, C, discrete):
    a, b = make_pair(E, R, O, C, seed=11)
    g = torch.Generator(device=DEV).manual_seed(5)
    n_ended = 0
    for t in range(6):
        if discrete:
            act = torch.randint(0, 25, (E, R), device=DEV, generator=g, dtype=torch.int32)
        else:
            act = torch.rand((E, R, 2), device=DEV, generator=g) * 2 - 1
        oa = a.step(act, is_continuous_action=not discrete)
        ob = b.step_no_reset(act, is_continuous_action=not discrete)
        trans = {k: ob[k].clone() for k in ("reward", "status", "phi", "env_status", "ended",
                                            "obs_self", "obs_objects", "obs_mask")}
        b.finish_episodes()
        torch.cuda.synchronize()
        ended = trans["ended"].bool()
        n_ended += int(ended.sum())
        # the transition's scalars are kept for every env (the three-call path overwrites them)
        for k in ("reward", "status", "phi", "env_status", "ended"):
            assert same(oa[k], trans[k]), (t, k)
        # the observation to act on next: fresh for ended envs, the step's own otherwise
        for k in ("obs_self", "obs_objects", "obs_mask"):
            assert same(oa[k], b.out[k]), (t, k)
            assert same(a.term[k][ended], trans[k][ended]), (t, "term " + k)
        for k, v in a.state.items():
            assert same(v, b.state[k]), (t, "state " + k)
    assert n_ended > 0


def test_fused_keeps_untouched_term_rows_and_cuda_graph():
    E, R = 256, 8
    a, b = make_pair(E, R, 10, 4, seed=3)
    a.enable_cuda_graph(True)
    b.enable_cuda_graph(True)
    g = torch.Generator(device=DEV).manual_seed(9)
    for t in range(8):
        act = torch.rand((E, R, 2), device=DEV, generator=g) * 2 - 1
        a.step_graph(act)
        b.step_graph(act)
    torch.cuda.synchronize()
    for k, v in a.state.items():
        assert same(v, b.state[k]), k
    for k in ("obs_self", "obs_objects", "obs_mask"):
        assert same(a.out[k], b.out[k]), k


def test_replay_push_reads_terminal_rows():
    """Ring contents after fused steps == after the three-call sequence with the push in between."""
    E, R = 200, 8
    a, b = make_pair(E, R, 10, 0, seed=21)
    ra = ReplayBuffer(50000, 64, 5, 5, device=DEV)
    rb = ReplayBuffer(50000, 64, 5, 5, device=DEV)
    g = torch.Generator(device=DEV).manual_seed(1)
    for t in range(5):
        act = torch.rand((E, R, 2), device=DEV, generator=g) * 2 - 1
        obs_a = tuple(x.clone() for x in a.obs)
        obs_b = tuple(x.clone() for x in b.obs)
        oa = a.step_fused(act)
        ra.push_step(obs_a, act, False, oa, term=a.term)
        ob = b.step_no_reset(act)
        rb.push_step(obs_b, act, False, ob)
        b.finish_episodes()
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
                                env.want_robots, env.want_cores, env.want_obs, env.min_start_goal_dis)
    with pytest.raises(RuntimeError):  # term without ended
        _lib.replay_push(_lib.new_replay(100, 64, DEV), *env.obs, act, False, env.out["reward"], *env.obs,
                         env.out["status"], term=env.term)
