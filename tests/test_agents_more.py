"""DDPG / DQN / SAC / Rainbow against the reference's own outputs (tests/golden/agents_more.npz, written
by oracle/gen_golden_agents.py from the unmodified reference): initial weights for the seed and forward
passes on CPU; one full train step from an injected batch on the GPU (loss 1e-5 relative, updated
weights by digest)."""
import os

import numpy as np
import pytest
import torch

from distributional_rl_decision_and_control_b200 import _lib
from distributional_rl_decision_and_control_b200.policy.DDPG_model import DDPG_Policy
from distributional_rl_decision_and_control_b200.policy.DQN_model import DQN_Policy
from distributional_rl_decision_and_control_b200.policy.Rainbow_model import Rainbow_Policy
from distributional_rl_decision_and_control_b200.policy.SAC_model import SAC_Policy

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "agents_more.npz")
DIMS = (7, 5, 5, 56, 40, 256, 128)
RANGES = [[-1.0, 1.0], [-1.0, 1.0]]
PARTS = {"DDPG": ("actor", "critic"), "SAC": ("actor", "critic_1", "critic_2")}


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(GOLD))


def build(kind, seed, device="cpu"):
    if kind == "DDPG":
        return DDPG_Policy(*DIMS, RANGES, device, seed)
    if kind == "SAC":
        return SAC_Policy(*DIMS, RANGES, device, seed)
    if kind == "DQN":
        return DQN_Policy(*DIMS, 25, device, seed).to(device)
    return Rainbow_Policy(*DIMS, 25, 51, device, seed).to(device)


def parts_of(kind, policy):
    names = PARTS.get(kind)
    return {n: getattr(policy, n) for n in names} if names else {"net": policy}


def check_digest(gold, prefix, modules, rtol):
    n = 0
    for name, m in modules.items():
        for k, v in m.state_dict().items():
            want = gold[f"{prefix}.{name}.{k}"]
            a = v.detach().double().cpu().numpy().ravel()
            got = np.concatenate([[a.sum(), np.abs(a).sum()], a[:6]])
            scale = max(want[1], 1e-3)
            assert np.allclose(got[:2], want[:2], rtol=0, atol=rtol * scale), (prefix, name, k, got[:2], want[:2])
            assert np.allclose(got[2:], want[2:len(got)], rtol=0, atol=rtol * max(1.0, np.abs(want[2:]).max())), (prefix, name, k)
            n += 1
    assert n > 0


def states_of(gold, kind, device="cpu", nxt=False):
    p = "n" if nxt else ""
    t = lambda k: torch.from_numpy(gold[f"{kind}.b_{p}{k}"]).to(device)
    return (t("self"), t("obj"), t("mask"))


@pytest.mark.parametrize("kind", ["DDPG", "DQN", "SAC", "Rainbow"])
def test_initial_weights_and_forward(gold, kind):
    seed = int(gold["seed"])
    pol = build(kind, seed)
    check_digest(gold, f"{kind}.init", parts_of(kind, pol), 1e-6)
    x = states_of(gold, kind)
    tol = dict(rtol=1e-5, atol=1e-6)
    with torch.no_grad():
        if kind == "DDPG":
            assert np.allclose(pol.actor(x).numpy(), gold["DDPG.actor_out"], **tol)
            act = torch.from_numpy(gold["DDPG.b_act"])
            assert np.allclose(pol.critic(x, act).numpy(), gold["DDPG.critic_out"], **tol)
        elif kind == "DQN":
            assert np.allclose(pol(x).numpy(), gold["DQN.q_out"], rtol=1e-5, atol=1e-5)
        elif kind == "SAC":
            a, lp = pol.actor(x, torch.from_numpy(gold["SAC.eps_fwd"]))
            assert np.allclose(a.numpy(), gold["SAC.actor_out"], **tol)
            assert np.allclose(lp.numpy(), gold["SAC.log_prob"], rtol=1e-5, atol=1e-5)
        else:
            assert np.allclose(pol(x).numpy(), gold["Rainbow.p_out"], rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("kind", ["DDPG", "DQN", "SAC", "Rainbow"])
def test_checkpoint_round_trip(tmp_path, kind):
    pol = build(kind, 3)
    pol.save(str(tmp_path))
    back = type(pol).load(str(tmp_path))
    for (n, a), (_, b) in zip(parts_of(kind, pol).items(), parts_of(kind, back).items()):
        for (k, v), (k2, v2) in zip(a.state_dict().items(), b.state_dict().items()):
            assert k == k2 and torch.equal(v, v2), (n, k)
    want = {"DDPG": {"actor_network_params.pth", "actor_constructor_params.json", "critic_network_params.pth",
                     "critic_constructor_params.json"},
            "SAC": {"actor_network_params.pth", "actor_constructor_params.json", "critic_1_params.pth",
                    "critic_1_constructor_params.json", "critic_2_params.pth", "critic_2_constructor_params.json"}
            }.get(kind, {"network_params.pth", "constructor_params.json"})
    assert set(os.listdir(tmp_path)) == want


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["DDPG", "DQN", "SAC", "Rainbow"])
def test_train_step_matches_reference(gold, kind):
    from distributional_rl_decision_and_control_b200.agent import Agent
    dev = "cuda:0"
    B = 64
    agent = Agent(device=dev, seed=int(gold["seed"]), agent_type=kind, BUFFER_SIZE=4096)
    s, ns = states_of(gold, kind, dev), states_of(gold, kind, dev, nxt=True)
    t = lambda k: torch.from_numpy(gold[f"{kind}.{k}"]).to(dev)
    if kind == "Rainbow":
        sample = (torch.arange(B, device=dev), s, t("b_act").long(), t("returns"), ns, t("nonterminals"), t("weights"))
        agent.memory.update_priorities = lambda i, p: None

        def reset_noise_like_the_cpu_reference():  # same draws as Rainbow:44-53 on the CPU generator
            torch.manual_seed(777)
            scale = lambda x: x.sign().mul_(x.abs().sqrt_())
            for name, mod in agent.policy_target.named_children():
                if "hidden_layer" in name or "output_layer" in name:
                    e_in, e_out = scale(torch.randn(mod.in_features)), scale(torch.randn(mod.out_features))
                    mod.weight_epsilon.copy_(torch.outer(e_out, e_in))
                    mod.bias_epsilon.copy_(e_out)
        agent.policy_target.reset_noise = reset_noise_like_the_cpu_reference
        loss = agent.train_Rainbow(sample)
        got = [float(loss.mean())]
        assert np.allclose(loss.cpu().numpy(), gold["Rainbow.loss_per_sample"], rtol=2e-5, atol=1e-5)
    else:
        act = t("b_act")
        if kind == "DQN":
            act = torch.stack([act, torch.zeros_like(act)], dim=1)
        batch = {"self": s[0], "obj": s[1], "mask": s[2], "nself": ns[0], "nobj": ns[1], "nmask": ns[2],
                 "act": act.contiguous(), "rew": t("b_rew"), "done": t("b_done")}
        if kind == "SAC":
            out = agent.train_SAC(batch, noise={"target": t("eps_target"), "actor": t("eps_actor")})
        else:
            out = agent.train_DDPG(batch) if kind == "DDPG" else (agent.train_DQN(batch),)
        got = [float(x) for x in out]
    assert np.allclose(got, gold[f"{kind}.losses"], rtol=2e-5, atol=1e-6), (got, gold[f"{kind}.losses"])
    check_digest(gold, f"{kind}.post", parts_of(kind, agent.policy_local), 2e-5)


@pytest.mark.gpu
def test_rainbow_memory_n_step_window():
    """One lane = the reference's cyclic buffer: n-step return, n-th next state and the blanking of
    windows that cross an episode start (REPLAY:131-179), on a hand-made stream."""
    from distributional_rl_decision_and_control_b200.utils.replay_memory_rainbow import ReplayMemory
    m = ReplayMemory("cuda:0", 64)
    # two episodes: lengths 5 (terminal at its last step) and 7; reward = global index / 10
    terms = [False] * 4 + [True] + [False] * 7
    for i, term in enumerate(terms):
        m.append(([float(i)] * 7, [[float(i)] * 5] * (i % 3)), i % 25, i / 10.0, term)
    assert m.transitions.num_elements() == len(terms)
    idxs = torch.tensor([0, 2, 3, 4, 5, 8], device="cuda:0")
    s, a, R, ns, nonterm = m._gather(idxs)
    g = [1.0, 0.99, 0.99 ** 2]
    r = lambda i: i / 10.0
    want_R = [r(0) * g[0] + r(1) * g[1] + r(2) * g[2],       # inside episode 1
              r(2) * g[0] + r(3) * g[1] + r(4) * g[2],       # reaches the terminal step: all three rewards count
              r(3) * g[0] + r(4) * g[1],                     # position 2 is the next episode's first step: blank
              r(4) * g[0],                                   # the terminal step itself
              r(5) * g[0] + r(6) * g[1] + r(7) * g[2],
              r(8) * g[0] + r(9) * g[1] + r(10) * g[2]]
    assert np.allclose(R.cpu().numpy(), want_R, rtol=1e-6)
    assert s[0][:, 0].cpu().tolist() == [0.0, 2.0, 3.0, 4.0, 5.0, 8.0]
    assert ns[0][:, 0].cpu().tolist() == [3.0, 0.0, 0.0, 0.0, 8.0, 11.0]     # blank next states are zeros
    assert nonterm[:, 0].cpu().tolist() == [1.0, 0.0, 0.0, 0.0, 1.0, 1.0]
    assert a.cpu().tolist() == [0, 2, 3, 4, 5, 8]
    assert s[2].sum(1).cpu().tolist() == [float(i % 3) for i in (0, 2, 3, 4, 5, 8)]  # masks from the counts
    # sampling never returns the n newest slots, and priorities follow updates
    idx, *_ = m.sample(8)
    assert int(idx.max()) < len(terms) - m.n
    m.update_priorities(idx[:2], torch.tensor([4.0, 9.0], device="cuda:0"))
    assert m.transitions.max == 3.0


@pytest.mark.gpu
@pytest.mark.parametrize("M,head", [(3000, True), (5000, True), (5000, False), (65536 + 77, True)])
def test_rainbow_greedy_actions_match_probability_path(M, head, monkeypatch):
    """asvrl_c51_greedy (dueling combination, softmax, expectation, argmax in one pass) against the module's own
    probability tensor in float64 (Rainbow:127-135, AGENT:235-246), noisy layers in training mode.  M = 5000 takes the
    tensor-core paths: K9's trunk per stream, then K13 (asvrl_c51_head: advantage layer + dueling C51 in one tcgen05
    launch) or, with head = False, K8 blocks + asvrl_c51_greedy."""
    from distributional_rl_decision_and_control_b200.policy import Rainbow_model
    monkeypatch.setattr(Rainbow_model, "C51_HEAD", head)
    import copy
    dev = "cuda:0"
    pol = build("Rainbow", 7, dev)
    pol.train()
    g = torch.Generator(device=dev).manual_seed(3)
    x = (torch.randn((M, 7), device=dev, generator=g) * 5, torch.randn((M, 5, 5), device=dev, generator=g) * 3,
         (torch.rand((M, 5), device=dev, generator=g) < 0.6).float())
    x = (x[0], x[1] * x[2][..., None], x[2])
    support = torch.linspace(-10.0, 10.0, 51, device=dev)
    act, q = pol.greedy_actions(x, support, want_q=True)
    torch.cuda.synchronize()
    _lib.c51_head_status()
    _lib.policy_forward_status()
    ref = copy.deepcopy(pol).double()
    ref.fused_tower = False
    ref.train()
    want = (ref(tuple(t.double() for t in x)) * support.double()).sum(2)
    assert float((q.double() - want).abs().max()) <= 1e-5 * float(want.abs().max()) + 1e-6
    top2 = want.topk(2, dim=1).values
    clear = (top2[:, 0] - top2[:, 1]) > 1e-4
    assert torch.equal(act.long()[clear], want.argmax(1)[clear])
    assert int(clear.sum()) > M // 2


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["DQN", "SAC"])
def test_collection_hidden_layers_on_tensor_cores_match_float64(kind):
    """A no-grad forward over >= 4 096 rows takes K8 for the two hidden layers (`_ObservationTower._hidden`); the
    action values / the actor's mean stay within 1e-5 of the float64 module."""
    import copy
    dev = "cuda:0"
    pol = build(kind, 11, dev)
    net = pol if kind == "DQN" else pol.actor
    g = torch.Generator(device=dev).manual_seed(5)
    M = 5000
    x = (torch.randn((M, 7), device=dev, generator=g) * 5, torch.randn((M, 5, 5), device=dev, generator=g) * 3,
         (torch.rand((M, 5), device=dev, generator=g) < 0.6).float())
    x = (x[0], x[1] * x[2][..., None], x[2])
    ref = copy.deepcopy(net).double()
    ref.fused_tower = False
    x64 = tuple(t.double() for t in x)
    with torch.no_grad():
        if kind == "DQN":
            got, want = net(x), ref(x64)
        else:
            z = torch.zeros((M, 2), device=dev)
            got, want = net(x, noise=z)[0], ref(x64, noise=z.double())[0]
    torch.cuda.synchronize()
    _lib.linear_tf32x3_status()
    assert float((got.double() - want).abs().max()) <= 1e-5 * float(want.abs().max()) + 1e-6


@pytest.mark.gpu
def test_rainbow_batched_append_kernel_matches_torch_statements():
    """asvrl_rainbow_append (one launch per batched step) against the torch statements it replaces, over a ring that
    wraps, with deactivated / absent / finishing ASVs."""
    from distributional_rl_decision_and_control_b200.utils.replay_memory_rainbow import ReplayMemory
    dev = "cuda:0"
    E, R = 37, 8
    L = E * R
    mems = [ReplayMemory(dev, 8 * L, lanes=L), ReplayMemory(dev, 8 * L, lanes=L)]
    mems[1].fused_append = False
    g = torch.Generator(device=dev).manual_seed(9)
    for step in range(13):
        obs = (torch.randn((E, R, 7), device=dev, generator=g), torch.randn((E, R, 5, 5), device=dev, generator=g),
               (torch.rand((E, R, 5), device=dev, generator=g) < 0.5).float())
        actions = torch.randint(0, 25, (E, R), device=dev, generator=g, dtype=torch.int32)
        u = torch.rand((E, R), device=dev, generator=g)
        status = torch.zeros((E, R), dtype=torch.int32, device=dev)
        status |= (u < 0.15).int() * 1                    # done
        status |= ((u > 0.8) & (u < 0.9)).int() * 8       # was deactivated
        status |= (u >= 0.9).int() * (7 << 8)             # absent
        status |= ((u >= 0.3) & (u < 0.8)).int() * (1 << 8)
        out = {"status": status, "reward": torch.randn((E, R), device=dev, generator=g)}
        for m in mems:
            m.max_priority.fill_(1.0 + 0.25 * step)
            m.append_step(obs, actions, out)
    torch.cuda.synchronize()
    a, b = mems
    for name in ("rows", "tstep", "prio", "lane_t", "count", "head_dev", "filled_dev"):
        assert torch.equal(getattr(a, name), getattr(b, name)), name
    assert a.head == b.head and a.filled == b.filled and int(a._append_scratch.abs().sum()) == 0


@pytest.mark.gpu
def test_sac_act_equals_forward_actions():
    """Actor.act (one 128 -> 4 product, no log-probability terms) returns the actions of Actor.forward for the same draw."""
    dev = "cuda:0"
    actor = build("SAC", 13, dev).actor
    g = torch.Generator(device=dev).manual_seed(8)
    for M in (300, 5000):
        x = (torch.randn((M, 7), device=dev, generator=g) * 5, torch.randn((M, 5, 5), device=dev, generator=g) * 3,
             (torch.rand((M, 5), device=dev, generator=g) < 0.6).float())
        x = (x[0], x[1] * x[2][..., None], x[2])
        z = torch.randn((M, 2), device=dev, generator=g)
        with torch.no_grad():
            want = actor(x, noise=z)[0]
        got = actor.act(x, noise=z)
        assert float((got - want).abs().max()) <= 1e-5   # two cuBLAS products of different shapes: summation order only
