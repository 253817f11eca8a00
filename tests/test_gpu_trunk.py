"""K5 (critic trunk glue) on the B200 through the C-ABI: each kernel against the same op written
in stock PyTorch (float64 on the GPU as the yardstick), and the fused critic path against the
module-by-module autograd path of the same Critic — forward quantiles, every parameter gradient,
and d z / d actions of the actor update."""
import pytest
import torch

from distributional_rl_decision_and_control_b200 import _lib
from distributional_rl_decision_and_control_b200.agent import Agent
from distributional_rl_decision_and_control_b200.policy.AC_IQN_model import Critic

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
RTOL = 1e-5


def close(got, want, what, rtol=RTOL):
    """1e-5 of the tensor's scale (sums of up to 32768 fp32 terms are compared to float64)."""
    want = want.to(torch.float64)
    err = (got.to(torch.float64) - want).abs().max().item()
    scale = max(want.abs().max().item(), 1e-6)
    assert err <= rtol * scale, f"{what}: err {err:.3e} vs scale {scale:.3e}"


@pytest.mark.parametrize("B,N,D", [(1024, 32, 256), (64, 8, 128), (5, 3, 256), (1, 1, 128), (130, 7, 64)])
def test_scale_relu_fwd_bwd(B, N, D):
    g = torch.Generator(device=DEV).manual_seed(B + N + D)
    M = B * N
    pre = torch.randn((M, D), device=DEV, generator=g)
    bias = torch.randn((D,), device=DEV, generator=g)
    scale = torch.relu(torch.randn((B, D), device=DEV, generator=g))  # a ReLU feature: zeros included
    dout = torch.randn((M, D), device=DEV, generator=g)
    out = _lib.iqn_scale_relu_fwd(pre, bias, scale, N)
    p64, b64, s64 = (t.double().requires_grad_(True) for t in (pre, bias, scale))
    ref = (s64.unsqueeze(1) * torch.relu(p64 + b64).view(B, N, D)).view(M, D)
    close(out, ref.detach(), "out")
    ref.backward(dout.double())
    dpre, dscale, dbias = _lib.iqn_scale_relu_bwd(dout, pre, bias, scale, N)
    close(dpre, p64.grad, "dpre")
    close(dscale, s64.grad, "dscale")
    close(dbias, b64.grad, "dbias")
    # partial requests: the actor update asks for dscale alone
    a, b, c = _lib.iqn_scale_relu_bwd(dout, pre, bias, scale, N, want_dpre=False, want_dbias=False)
    assert a is None and c is None
    assert torch.equal(b, dscale)
    # deterministic reductions, and the counter is left clean for the next call
    again = _lib.iqn_scale_relu_bwd(dout, pre, bias, scale, N)
    assert all(torch.equal(x, y) for x, y in zip(again, (dpre, dscale, dbias)))


@pytest.mark.parametrize("M,D", [(32768, 128), (512, 128), (15, 256), (1, 64), (1000, 32)])
def test_head_fwd_bwd(M, D):
    g = torch.Generator(device=DEV).manual_seed(M + D)
    pre = torch.randn((M, D), device=DEV, generator=g)
    bias = torch.randn((D,), device=DEV, generator=g)
    w = torch.randn((D,), device=DEV, generator=g)
    bo = torch.randn((1,), device=DEV, generator=g)
    dz = torch.randn((M,), device=DEV, generator=g)
    z = _lib.iqn_head_fwd(pre, bias, w, bo)
    p64, b64, w64, o64 = (t.double().requires_grad_(True) for t in (pre, bias, w, bo))
    ref = torch.relu(p64 + b64) @ w64 + o64
    close(z, ref.detach(), "z")
    ref.backward(dz.double())
    dpre, dw, dbias, dbo = _lib.iqn_head_bwd(dz, pre, bias, w)
    close(dpre, p64.grad, "dpre")
    close(dw, w64.grad, "dw")
    close(dbias, b64.grad, "dbias")
    close(dbo, o64.grad, "db_out")
    only, n1, n2, n3 = _lib.iqn_head_bwd(dz, pre, bias, w, want_params=False)
    assert n1 is None and n2 is None and n3 is None and torch.equal(only, dpre)


def _critic_inputs(B, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    own = torch.randn((B, 7), device=DEV, generator=g) * torch.tensor([20, 20, 2, 2, 0.5, 500, 500.0], device=DEV)
    obj = torch.randn((B, 5, 5), device=DEV, generator=g) * torch.tensor([15, 15, 1, 1, 1.0], device=DEV)
    cnt = torch.randint(0, 6, (B,), device=DEV, generator=g)
    mask = (torch.arange(5, device=DEV)[None, :] < cnt[:, None]).float()
    obj = obj * mask[..., None]
    act = torch.rand((B, 2), device=DEV, generator=g) * 2 - 1
    return (own, obj, mask), act


def close_most(got, want, what, rtol, frac=0.99, rtol_max=2e-2):
    """Per-sample quantities on the tensor-core path: a ReLU gate inside the rounding margin that flips against the
    float32 module path moves ONE sample's gradient by a few per cent (1 of its 32 quantile rows), so nearly all
    samples must agree to `rtol` of the tensor's scale and none may be off by more than a flipped gate's worth."""
    want = want.to(torch.float64)
    err = (got.to(torch.float64) - want).abs()
    scale = max(want.abs().max().item(), 1e-6)
    ok = (err <= rtol * scale).double().mean().item()
    assert ok >= frac and err.max().item() <= rtol_max * scale, \
        f"{what}: {ok:.4f} of the elements within {rtol:g}, max err {err.max().item():.3e} vs scale {scale:.3e}"


@pytest.mark.parametrize("umma", [False, True])
@pytest.mark.parametrize("B,N", [(64, 8), (1024, 32), (3, 5)])
def test_fused_critic_matches_module_path(B, N, umma, monkeypatch):
    """umma = False: K3 + cuBLAS + K5 (the same sgemm as the module path: agreement to rounding).  umma = True: K10 / K11 /
    K12 where they apply (>= 4096 quantile rows): an independent evaluation order, so ReLU gates of pre-activations
    inside the rounding margin differ from the torch float32 path's (as that path's differ from float64), and one such
    gate moves a weight gradient by ~1 / sqrt(B N) of its scale — the gradient arithmetic itself is pinned to 1e-5 with
    matched gates in tests/test_gpu_critic.py; here the bar on gradients is 2e-4 (measured against float64 on these
    inputs: 1e-6 .. 4e-6, profiles/probes/critic_grad_err.py; torch's float32 path: up to 1e-4 on cos_embedding)."""
    from distributional_rl_decision_and_control_b200.policy import iqn_ops
    monkeypatch.setattr(iqn_ops, "UMMA_TRUNK", umma)
    gtol = 2e-4 if umma and B * N >= iqn_ops.UMMA_TRUNK_MIN_ROWS else 2e-5
    crit = Critic(7, 5, 5, 56, 40, 256, 128, 2, device=DEV, seed=101).to(DEV)
    x, act = _critic_inputs(B, 7 + B)
    taus = torch.rand((B, N), device=DEV, generator=torch.Generator(device=DEV).manual_seed(1))
    wz = torch.randn((B, N), device=DEV, generator=torch.Generator(device=DEV).manual_seed(2))

    def run(fn, **kw):
        for p in crit.parameters():
            p.grad = None
        a = act.clone().requires_grad_(True)
        z, _ = fn(x, a, N, taus=taus, **kw)
        (z * wz).sum().backward()
        return z.detach(), a.grad, {n: (None if p.grad is None else p.grad.clone()) for n, p in crit.named_parameters()}

    crit.fused_tower = False       # module-by-module autograd path: the torch fp32 reference
    z0, da0, g0 = run(crit.forward)
    crit.fused_tower = True        # K6 tower + K5 trunk
    z1, da1, g1 = run(crit.forward_fused)
    close(z1, z0, "quantiles")
    tensor_path = gtol > 2e-5
    (close_most if tensor_path else close)(da1, da0, "d/d actions", rtol=1e-4 if tensor_path else gtol / 2)
    for n in g0:
        close(g1[n], g0[n], f"grad {n}", rtol=gtol)
    # the actor update's variant: the same d/d actions, no parameter gradients at all
    z2, da2, g2 = run(crit.forward_fused, param_grads=False)
    assert torch.equal(z2, z1)
    (close_most if tensor_path else close)(da2, da0, "d/d actions (actor pass)", rtol=1e-4 if tensor_path else gtol / 2)
    assert torch.equal(da2, da1) or not tensor_path   # K11's two forms are bit-identical
    assert all(v is None for v in g2.values())


def test_train_step_fused_equals_unfused():
    """One AC-IQN update with injected batch and taus: K5 path vs the autograd path."""
    B, N = 64, 8
    x, act = _critic_inputs(B, 11)
    nx, _ = _critic_inputs(B, 12)
    g = torch.Generator(device=DEV).manual_seed(5)
    batch = {"self": x[0], "obj": x[1], "mask": x[2], "act": act, "rew": torch.randn((B,), device=DEV, generator=g),
             "nself": nx[0], "nobj": nx[1], "nmask": nx[2],
             "done": (torch.rand((B,), device=DEV, generator=g) < 0.1).float()}
    taus = {k: torch.rand((B, N), device=DEV, generator=g) for k in ("target", "local", "actor")}
    res = []
    for fused in (False, True):
        ag = Agent(device=DEV, seed=100, agent_type="AC-IQN", BATCH_SIZE=B, num_tau=N, BUFFER_SIZE=1024,
                   fused_trunk=fused)
        before = [p.detach().clone() for p in ag.policy_local.critic.parameters()]
        cl, al = ag.train_AC_IQN(batch=batch, taus=taus)
        res.append((cl, al, [p.grad.clone() for p in ag.policy_local.actor.parameters()],
                    torch.cat([(p.detach() - q).reshape(-1) for p, q in
                               zip(ag.policy_local.critic.parameters(), before)])))
    close(res[1][0], res[0][0], "critic loss")
    close(res[1][1], res[0][1], "actor loss")
    for a, b in zip(res[1][2], res[0][2]):
        close(a, b, "clipped actor gradient", rtol=2e-5)
    # Adam's first step is lr * g / (|g| + eps): elements with |g| near the fp32 noise of the two
    # summation orders may differ, all others must agree
    diff = (res[1][3] - res[0][3]).abs()
    assert (diff > 1e-6).float().mean().item() < 0.01
    assert diff.max().item() <= 2.1e-4


@pytest.mark.parametrize("B", [1, 64, 1024, 32768, 77])
def test_tower_fwd_bwd(B):
    """K6 against the encoders written module by module (float64 on the GPU)."""
    from distributional_rl_decision_and_control_b200.policy.AC_IQN_model import Actor
    actor = Actor(7, 5, 5, 56, 40, 256, 128, [[-1.0, 1.0], [-1.0, 1.0]], device=DEV, seed=3).to(DEV)
    x, _ = _critic_inputs(B, B)
    wf = torch.randn((B, 256), device=DEV, generator=torch.Generator(device=DEV).manual_seed(4))
    ref = Actor(7, 5, 5, 56, 40, 256, 128, [[-1.0, 1.0], [-1.0, 1.0]], device=DEV, seed=3).to(DEV).double()
    ref.fused_tower = False
    f_ref = ref._features(tuple(t.double() for t in x))
    (f_ref * wf.double()).sum().backward()
    f = actor._features(x)
    (f * wf).sum().backward()
    close(f.detach(), f_ref.detach(), "features")
    assert torch.equal(f.detach()[:, 56:].reshape(B, 5, 40).abs().sum(-1) > 0,
                       (f_ref.detach()[:, 56:].reshape(B, 5, 40).abs().sum(-1) > 0))  # masked rows are exactly 0
    for name in ("self_encoder.0.weight", "self_encoder.0.bias", "object_encoder.0.weight", "object_encoder.0.bias"):
        close(dict(actor.named_parameters())[name].grad, dict(ref.named_parameters())[name].grad, name, rtol=2e-5)
    # deterministic
    g1 = actor.self_encoder[0].weight.grad.clone()
    actor.zero_grad(set_to_none=True)
    (actor._features(x) * wf).sum().backward()
    assert torch.equal(actor.self_encoder[0].weight.grad, g1)


@pytest.mark.parametrize("gscale", [1.0, 40.0])
def test_clip_adam_matches_torch(gscale):
    """K7 against torch.nn.utils.clip_grad_norm_(0.5) + torch.optim.Adam over several steps, with
    gradients below (gscale=1: some steps unclipped) and above the clipping threshold."""
    from distributional_rl_decision_and_control_b200.utils.clip_adam import ClipAdam
    g = torch.Generator(device=DEV).manual_seed(9)
    shapes = [(56, 7), (56,), (40, 5), (40,), (128, 256), (128,), (1, 128), (1,), (3, 5, 7)]
    mine = [torch.nn.Parameter(torch.randn(s, device=DEV, generator=g)) for s in shapes]
    ref = [torch.nn.Parameter(p.detach().clone()) for p in mine]
    opt = ClipAdam(mine, lr=1e-4, max_norm=0.5)
    topt = torch.optim.Adam(ref, lr=1e-4)
    for it in range(5):
        scale = gscale * (0.002 if it % 2 == 0 else 1.0)
        for p, q in zip(mine, ref):
            p.grad = torch.randn(p.shape, device=DEV, generator=g) * scale
            q.grad = p.grad.clone()
        if it == 3:  # a parameter without gradient is skipped by both
            mine[2].grad = None
            ref[2].grad = None
        want_norm = torch.nn.utils.clip_grad_norm_(ref, 0.5)
        topt.step()
        opt.step()
        close(opt.grad_norm, want_norm.reshape(1), "total norm", rtol=1e-6)
        for p, q in zip(mine, ref):
            assert (p.detach() - q.detach()).abs().max().item() <= 2e-7 * max(1.0, q.abs().max().item())
    assert opt.step_count.tolist() == [5.0, 5.0, 4.0] + [5.0] * 6
    assert int(opt._counter.item()) == 0


@pytest.mark.parametrize("B,N", [(64, 8), (256, 32), (3, 5)])
def test_fused_iqn_network_matches_module_path(B, N):
    """IQN_Policy.forward_fused (K5 + K6) against the module-by-module autograd path."""
    from distributional_rl_decision_and_control_b200.policy.IQN_model import IQN_Policy
    net = IQN_Policy(7, 5, 5, 56, 40, 256, 128, 25, device=DEV, seed=7).to(DEV)
    x, _ = _critic_inputs(B, 100 + B)
    taus = torch.rand((B, N), device=DEV, generator=torch.Generator(device=DEV).manual_seed(1))
    wz = torch.randn((B, N, 25), device=DEV, generator=torch.Generator(device=DEV).manual_seed(2))

    def run(fn):
        net.zero_grad(set_to_none=True)
        z, _ = fn(x, N, taus=taus)
        (z * wz).sum().backward()
        return z.detach(), {n: p.grad.clone() for n, p in net.named_parameters()}

    net.fused_tower = False
    z0, g0 = run(net.forward)
    net.fused_tower = True
    z1, g1 = run(net.forward_fused)
    close(z1, z0, "quantiles")
    for n in g0:
        close(g1[n], g0[n], f"grad {n}", rtol=2e-5)


def test_clip_adam_skips_non_finite_steps_and_reports_them():
    from distributional_rl_decision_and_control_b200.utils.clip_adam import ClipAdam
    p = torch.nn.Parameter(torch.ones(300, device=DEV))
    opt = ClipAdam([p], lr=1e-2)
    p.grad = torch.full_like(p, float("nan"))
    opt.step()
    torch.cuda.synchronize()
    assert torch.equal(p.detach(), torch.ones(300, device=DEV))  # untouched
    assert float(opt.step_count[0]) == 0.0 and float(opt.exp_avg[0].abs().max()) == 0.0  # nor the moments / step count
    p.grad = torch.ones_like(p)
    opt.step()  # a good step AFTER the skipped one: the skip is still reported (counted on the device, not "last norm")
    with pytest.raises(RuntimeError, match="1 optimiser step"):
        opt.check()
    assert (p.detach() < 1.0).all() and float(opt.step_count[0]) == 1.0
    opt.step()
    opt.check()  # reported once; nothing new was skipped


def test_actor_head_eps_greedy_matches_actor_forward():
    """The collection path's fused forward: eps = 0 reproduces Actor.forward; eps = 0.3 explores about
    30 % of the rows uniformly in range, leaves the others greedy, and draws fresh noise per call."""
    ag = Agent(device=DEV, seed=100, agent_type="AC-IQN", BATCH_SIZE=64, num_tau=8, BUFFER_SIZE=256)
    ag.policy_local.actor.fused_policy = False  # this test covers round 1's launch sequence (K9: tests/test_gpu_policy.py)
    B = 20000
    x, _ = _critic_inputs(B, 77)
    greedy = ag.policy_local.actor(x)
    a0 = ag.act_batch(x, eps=0.0)
    close(a0, greedy, "greedy actions", rtol=1e-5)  # K8 (3 x TF32 on tcgen05) against cuBLAS sgemm
    ag.policy_local.actor.umma_linear = False
    close(ag.act_batch(x, eps=0.0), greedy, "greedy actions, cuBLAS + K5 path", rtol=2e-6)
    ag.policy_local.actor.umma_linear = True
    a1 = ag.act_batch(x, eps=torch.tensor(0.3, device=DEV))
    a2 = ag.act_batch(x, eps=0.3)
    explored = (a1 - greedy).abs().max(dim=1).values > 1e-5
    frac = explored.float().mean().item()
    assert 0.28 < frac < 0.32, frac
    rnd = a1[explored]
    assert rnd.min().item() >= -1.0 and rnd.max().item() <= 1.0
    assert abs(rnd.mean().item()) < 0.02 and abs(rnd.var().item() - 1.0 / 3.0) < 0.02  # U(-1, 1)
    assert not torch.equal(a1, a2)  # the device counter advanced


@pytest.mark.parametrize("M,K,N,bias,relu", [(65536, 256, 128, True, True), (1000, 128, 128, False, False),
                                             (300, 64, 16, True, False), (257, 32, 256, True, True),
                                             (128, 256, 64, False, True), (1, 96, 48, True, True)])
def test_linear_tf32x3_matches_float64(M, K, N, bias, relu):
    """K8: the tcgen05 dense layer (three-pass TF32 split) against the float64 product.  The bound is
    the split's own: ~2^-21 of sum_k |x||w| per output — fp32-GEMM level, 1000 x below one TF32 pass."""
    g = torch.Generator(device=DEV).manual_seed(M + K + N)
    x = torch.randn((M, K), device=DEV, generator=g) * 3.0
    w = (torch.rand((N, K), device=DEV, generator=g) * 2 - 1) / K ** 0.5
    b = torch.randn((N,), device=DEV, generator=g) if bias else None
    got = _lib.linear_tf32x3(x, w, b, relu=relu)
    want = x.double() @ w.double().t()
    if bias:
        want = want + b.double()
    if relu:
        want = want.clamp_min(0.0)
    mag = x.double().abs() @ w.double().abs().t()
    err = (got.double() - want).abs()
    assert float((err / mag.clamp_min(1e-30)).max()) <= 2e-6, float((err / mag).max())
    # one TF32 pass would be ~2^-11 of mag: the split really ran
    assert float(err.max()) <= 1e-5 * float(want.abs().max().clamp_min(1e-6))
    torch.cuda.synchronize()


def test_linear_tf32x3_in_a_cuda_graph_and_rejects_bad_shapes():
    x = torch.randn((4096, 256), device=DEV)
    w = torch.randn((128, 256), device=DEV) / 16
    out = torch.empty((4096, 128), device=DEV)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        _lib.linear_tf32x3(x, w, None, out=out)
    torch.cuda.current_stream().wait_stream(s)
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        _lib.linear_tf32x3(x, w, None, out=out)
    x.mul_(2.0)
    gr.replay()
    torch.cuda.synchronize()
    close(out, x.double() @ w.double().t(), "graph replay", rtol=1e-5)
    for bad in ((4096, 250), (4096, 16)):
        with pytest.raises(RuntimeError):
            _lib.linear_tf32x3(torch.zeros(bad, device=DEV), torch.zeros((128, bad[1]), device=DEV))
    with pytest.raises(RuntimeError):
        _lib.linear_tf32x3(torch.zeros((64, 64), device=DEV), torch.zeros((100, 64), device=DEV))


@pytest.mark.parametrize("B", [1024, 64, 1, 77, 4096])
def test_narrow_layers_match_torch(B):
    """The action encoder and the actor's atan head (one launch each way) against the same layers in float64 autograd."""
    from distributional_rl_decision_and_control_b200.policy.iqn_ops import action_encoder, atan_head
    g = torch.Generator(device=DEV).manual_seed(B)
    r = lambda *s: torch.randn(s, device=DEV, generator=g)
    D = 128
    a, wa, ba, up = (r(B, 2).clamp(-1, 1).requires_grad_(True), (r(D, 2) * 0.7).requires_grad_(True),
                     (r(D) * 0.7).requires_grad_(True), r(B, D))
    out = action_encoder(a, wa, ba)
    out.backward(up)
    a64, wa64, ba64 = (t.detach().double().requires_grad_(True) for t in (a, wa, ba))
    ref = torch.relu(a64 @ wa64.t() + ba64)
    ref.backward(up.double())
    close(out.detach(), ref.detach(), "aenc")
    for name, got, want in (("d actions", a.grad, a64.grad), ("d wa", wa.grad, wa64.grad), ("d ba", ba.grad, ba64.grad)):
        close(got, want, name)
    h, w3, b3, upa = r(B, D).relu().requires_grad_(True), (r(2, D) * 0.1).requires_grad_(True), r(2).requires_grad_(True), r(B, 2)
    s = 2.0 / 3.141592653589793
    act = atan_head(h, w3, b3, s)
    act.backward(upa)
    h64, w64, b64 = (t.detach().double().requires_grad_(True) for t in (h, w3, b3))
    ref = s * torch.atan(h64 @ w64.t() + b64)
    ref.backward(upa.double())
    close(act.detach(), ref.detach(), "actions")
    for name, got, want in (("d h", h.grad, h64.grad), ("d w3", w3.grad, w64.grad), ("d b3", b3.grad, b64.grad)):
        close(got, want, name)
    act2 = atan_head(h.detach().requires_grad_(True), w3, b3, s)
    assert torch.equal(act2.detach(), act.detach())   # deterministic
