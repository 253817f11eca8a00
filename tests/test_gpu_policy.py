"""K9 (`asvrl_policy_forward`): the whole collection policy forward in one persistent tcgen05 launch,
through the C-ABI, against (1) the float64 module-by-module actor (the reference's arithmetic,
ACIQN:284-323) and (2) round 1's K6 + K8 + actor-head path for the exploration draws (same Philox
stream, so the explored rows and their uniform actions are bit-identical)."""
import copy

import pytest
import torch

from distributional_rl_decision_and_control_b200 import _lib
from distributional_rl_decision_and_control_b200.agent import Agent

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ATOL = 1e-5  # actions live in (-1, 1): BASELINE.md's 1e-5 relative bar on a unit-scale quantity


def _obs(M, seed):
    """Observation batch with the reference's value ranges (thrust up to 1000 N, positions in metres)."""
    g = torch.Generator(device=DEV).manual_seed(seed)
    scale_s = torch.tensor([20.0, 20.0, 2.0, 2.0, 0.5, 500.0, 500.0], device=DEV)
    scale_o = torch.tensor([15.0, 15.0, 1.0, 1.0, 1.0], device=DEV)
    s = torch.randn((M, 7), device=DEV, generator=g) * scale_s
    cnt = torch.randint(0, 6, (M,), device=DEV, generator=g)
    mask = (torch.arange(5, device=DEV)[None, :] < cnt[:, None]).float()
    o = torch.randn((M, 5, 5), device=DEV, generator=g) * scale_o * mask[..., None]
    return s.contiguous(), o.contiguous(), mask.contiguous()


def _actor64(actor):
    a = copy.deepcopy(actor).double()
    a.fused_tower = False
    return a


def _agent(kind="AC-IQN", seed=100):
    return Agent(device=DEV, seed=seed, agent_type=kind, BATCH_SIZE=64, num_tau=8, BUFFER_SIZE=256)


@pytest.mark.parametrize("M", [65536, 20000, 128, 129, 1, 31, 148 * 128 * 2 + 77])
def test_policy_forward_matches_float64_actor(M):
    ag = _agent()
    actor = ag.policy_local.actor
    assert actor._fused_policy_ok(_obs(4, 0)), "K9 must be the path that runs"
    x = _obs(M, 1000 + M)
    want = _actor64(actor)(tuple(t.double() for t in x))
    got = ag.act_batch(x, eps=0.0)
    torch.cuda.synchronize()
    _lib.policy_forward_status()
    assert got.shape == (M, 2) and got.dtype == torch.float32
    err = (got.double() - want).abs().max().item()
    assert err <= ATOL, f"greedy actions: max abs err {err:.3e}"
    # larger, non-uniform scales in every layer (pre-activations ~50 x larger): the yardstick is the error of the same
    # network evaluated in plain fp32 by torch / cuBLAS sgemm; 22-bit operands (two fp16 pieces) against fp32's 24
    # bits is up to 4 x the representation error per product (measured: <= 5.5 x the fp32 modules' error)
    g = torch.Generator(device=DEV).manual_seed(7)
    with torch.no_grad():
        for p in actor.parameters():
            p.mul_(1.0 + 2.0 * torch.rand(p.shape, device=DEV, generator=g))
    want = _actor64(actor)(tuple(t.double() for t in x))
    got = ag.act_batch(x, eps=0.0)
    plain = copy.deepcopy(actor)
    plain.fused_tower = False
    err32 = (plain(x).double() - want).abs().max().item()
    err = (got.double() - want).abs().max().item()
    assert err <= max(ATOL, 8.0 * err32), f"greedy actions, rescaled weights: max abs err {err:.3e} (fp32 modules: {err32:.3e})"
    _lib.policy_forward_status()


def test_policy_forward_exploration_is_the_actor_head_draw():
    """eps-greedy: the same rows explore, with the same uniform actions, as asvrl_actor_head (Philox
    stream 5 keyed on seed, call counter, row); about eps of the rows; fresh noise per call."""
    ag = _agent()
    actor = ag.policy_local.actor
    M = 50000
    x = _obs(M, 5)
    greedy = ag.act_batch(x, eps=0.0)
    a1 = ag.act_batch(x, eps=torch.tensor(0.3, device=DEV))
    ctr = ag._explore_counter.clone()
    a2 = ag.act_batch(x, eps=0.3)
    explored = (a1 - greedy).abs().max(dim=1).values > 1e-6
    frac = explored.float().mean().item()
    assert 0.29 < frac < 0.31, frac
    rnd = a1[explored]
    assert rnd.min().item() >= -1.0 and rnd.max().item() <= 1.0
    assert abs(rnd.mean().item()) < 0.02 and abs(rnd.var().item() - 1.0 / 3.0) < 0.02
    assert torch.equal(a1[~explored], greedy[~explored])
    assert not torch.equal(a1, a2)
    # the unfused path with the counter rewound draws the same exploration
    type(actor).fused_policy = False
    try:
        greedy_b = ag.act_batch(x, eps=0.0)
        ag._explore_counter.copy_(ctr)
        b2 = ag.act_batch(x, eps=0.3)
    finally:
        type(actor).fused_policy = True
    ex_a = (a2 != greedy).any(dim=1)
    ex_b = (b2 != greedy_b).any(dim=1)
    assert torch.equal(ex_a, ex_b)
    assert torch.equal(a2[ex_a], b2[ex_b])
    assert (a2 - b2).abs().max().item() <= ATOL  # greedy rows: fp16-split K9 against tf32-split K8


def test_policy_forward_unaligned_inputs_and_graph_replay():
    ag = _agent(seed=3)
    actor = ag.policy_local.actor
    M = 4096 + 50
    x = _obs(M, 11)
    want = ag.act_batch(x, eps=0.0)
    # views whose base pointer is 4-byte aligned only: the plain-load input path
    bufs = [torch.empty((t.numel() + 1,), device=DEV) for t in x]
    y = []
    for b, t in zip(bufs, x):
        v = b[1:].view(t.shape)
        v.copy_(t)
        assert v.data_ptr() % 16 != 0
        y.append(v)
    got = ag.act_batch(tuple(y), eps=0.0)
    assert torch.equal(got, want)
    # captured in a CUDA graph with a device-side eps and call counter; weights updated between replays
    eps = torch.tensor(0.0, device=DEV)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        ag.act_batch(x, eps=eps)
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        out = ag.act_batch(x, eps=eps)
    graph.replay()
    assert torch.equal(out, want)
    with torch.no_grad():
        actor.hidden_layer.weight.mul_(1.5)
    graph.replay()  # the pack launch is part of the graph: the new weights are used
    want2 = _actor64(actor)(tuple(t.double() for t in x))
    assert (out.double() - want2).abs().max().item() <= ATOL
    assert (out - want).abs().max().item() > 1e-3
    torch.cuda.synchronize()
    _lib.policy_forward_status()


def test_policy_forward_ddpg_actor():
    ag = _agent("DDPG", seed=9)
    x = _obs(3000, 21)
    want = _actor64(ag.policy_local.actor)(tuple(t.double() for t in x))
    got = ag.act_batch(x, eps=0.0)
    assert (got.double() - want).abs().max().item() <= ATOL


@pytest.mark.parametrize("M", [65536, 5000, 129, 1])
def test_policy_hidden_matches_float64_trunk(M):
    """asvrl_policy_hidden (K9's trunk with the second hidden layer written out: the front end DQN, SAC's actor and
    Rainbow's streams share) against the float64 modules, on the reference's observation ranges."""
    actor = _agent().policy_local.actor
    x = _obs(M, 2000 + M)
    se, oe = actor.self_encoder[0], actor.object_encoder[0]
    l1, l2 = actor.hidden_layer, actor.hidden_layer_2
    got = _lib.policy_hidden(x[0], x[1], x[2], se.weight, se.bias, oe.weight, oe.bias, l1.weight, l1.bias, l2.weight, l2.bias)
    torch.cuda.synchronize()
    _lib.policy_forward_status()
    a64 = _actor64(actor)
    f64 = a64._features(tuple(t.double() for t in x))
    want = torch.relu(a64.hidden_layer_2(torch.relu(a64.hidden_layer(f64))))
    assert got.shape == (M, 128)
    scale = float(want.abs().max())
    err = float((got.double() - want).abs().max())
    assert err <= 1e-5 * scale + 1e-6, f"hidden layer: max abs err {err:.3e} at scale {scale:.3e}"
