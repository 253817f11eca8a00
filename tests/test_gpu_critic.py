"""K10 (`asvrl_critic_forward`): the IQN critic's quantile trunk in one persistent tcgen05 launch, through the
C-ABI, against the float64 restatement of ACIQN:410-426 / 453-480 (the reference's arithmetic with injected
taus), the tau stream of K3, and — SAVE mode — the operand images against the activations they must hold."""
import numpy as np
import pytest
import torch

from distributional_rl_decision_and_control_b200 import _lib
from distributional_rl_decision_and_control_b200.agent import Agent

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
RTOL = 1e-5  # quantiles to 1e-5 relative (of the vector's scale), abs floor 1e-6


def _critic(seed=100):
    ag = Agent(device=DEV, seed=seed, agent_type="AC-IQN", BATCH_SIZE=64, num_tau=8, BUFFER_SIZE=256)
    return ag.policy_local.critic


def _inputs(B, N, seed, feat_scale=1.0):
    g = torch.Generator(device=DEV).manual_seed(seed)
    tau = torch.rand((B, N), device=DEV, generator=g)
    # features / action encodings are ReLU outputs: non-negative, a third of them exactly zero
    feats = torch.relu(torch.randn((B, 256), device=DEV, generator=g) + 0.4) * feat_scale
    aenc = torch.relu(torch.randn((B, 128), device=DEV, generator=g) + 0.4)
    return tau.contiguous(), feats.contiguous(), aenc.contiguous()


def _params(c):
    return (c.cos_embedding.weight, c.cos_embedding.bias, c.hidden_layer.weight, c.hidden_layer.bias,
            c.hidden_layer_2.weight, c.hidden_layer_2.bias, c.output_layer.weight.reshape(-1), c.output_layer.bias)


def _trunk64(tau, feats, aenc, ps):
    """float64 trunk; the cosine argument is the reference's float32 product tau * fl32(pi i) (ACIQN:397, 424)."""
    Wc, bc, Wh, bh, Wh2, bh2, wo, bo = [p.detach().double() for p in ps]
    B, N = tau.shape
    pis = torch.tensor([np.pi * i for i in range(64)], dtype=torch.float32, device=tau.device)
    cos = torch.cos((tau.reshape(B, N, 1) * pis.view(1, 1, 64)).double())
    R1 = torch.relu(cos @ Wc.t() + bc)
    H1 = feats.double().unsqueeze(1) * R1
    R2 = torch.relu(H1 @ Wh.t() + bh)
    H2 = (aenc.double().unsqueeze(1) if aenc is not None else 1.0) * R2
    R3 = torch.relu(H2 @ Wh2.t() + bh2)
    z = R3 @ wo + bo
    return z, cos, H1, H2, R3


def _run(c, tau, feats, aenc, save=False, **kw):
    ws = _lib.critic_workspace("test", feats.device)
    out = _lib.critic_forward(feats, aenc, *_params(c), tau.shape[1], ws, tau_in=tau, save=save, **kw)
    torch.cuda.synchronize()
    _lib.critic_forward_status()
    return out


def _check(z, want, what):
    scale = want.abs().max().item()
    err = (z.double() - want).abs().max().item()
    assert err <= RTOL * scale + 1e-6, f"{what}: max abs err {err:.3e} at scale {scale:.3e}"


@pytest.mark.parametrize("B,N", [(1024, 32), (64, 8), (37, 8), (100, 5), (4096, 32), (1, 1), (3, 64), (148 * 8 + 3, 32)])
def test_quantiles_match_float64_trunk(B, N):
    c = _critic()
    tau, feats, aenc = _inputs(B, N, 10 * B + N)
    z, tau_out = _run(c, tau, feats, aenc)
    assert z.shape == (B, N) and tau_out is tau
    want = _trunk64(tau, feats, aenc, _params(c))[0]
    _check(z, want, "z")
    # without the action encoding (the IQN network's trunk, IQN:74-86)
    z2, _ = _run(c, tau, feats, None)
    _check(z2, _trunk64(tau, feats, None, _params(c))[0], "z (unit scale)")


def test_larger_scales():
    """Observation features of the reference's range (thrust up to 1000 N gives features in the hundreds)."""
    c = _critic(7)
    tau, feats, aenc = _inputs(512, 32, 5, feat_scale=150.0)
    z, _ = _run(c, tau, feats, aenc)
    _check(z, _trunk64(tau, feats, aenc, _params(c))[0], "z at feature scale 150")


def test_tau_stream_is_k3s():
    c = _critic()
    B, N = 300, 32
    _, feats, aenc = _inputs(B, N, 3)
    ws = _lib.critic_workspace("test", feats.device)
    cdev = torch.tensor([5], dtype=torch.int64, device=DEV)
    z, tau = _lib.critic_forward(feats, aenc, *_params(c), N, ws, seed=1234, counter=7, counter_dev=cdev, cvar=0.8)
    tau3, _ = _lib.iqn_tau_cos(B, N, 64, DEV, seed=1234, counter=7, cvar=0.8, counter_dev=cdev)
    torch.cuda.synchronize()
    _lib.critic_forward_status()
    assert torch.equal(tau, tau3.reshape(B, N))
    _check(z, _trunk64(tau, feats, aenc, _params(c))[0], "z with drawn taus")


def _decode(img):
    """(tiles, chunks, 32 KB) uint8 operand image -> (tiles * 128, chunks * 64) float64 of hi + lo."""
    tiles, chunks, _ = img.shape
    h = img.reshape(tiles, chunks, 2, 128 * 64 * 2).view(torch.float16).reshape(tiles, chunks, 2, 128 * 64)
    r = torch.arange(128, device=img.device).view(128, 1)
    k = torch.arange(64, device=img.device).view(1, 64)
    idx = ((r >> 3) * 1024 + (r & 7) * 128 + (((k >> 3) ^ (r & 7)) << 4) + (k & 7) * 2) // 2   # half index in the tile
    v = h[..., idx.reshape(-1)].reshape(tiles, chunks, 2, 128, 64).double()
    v = v[:, :, 0] + v[:, :, 1]
    return v.permute(0, 2, 1, 3).reshape(tiles * 128, chunks * 64)


@pytest.mark.parametrize("B,N", [(1024, 32), (70, 8)])
def test_saved_operand_images(B, N):
    c = _critic()
    tau, feats, aenc = _inputs(B, N, 77 + B)
    z, _, (cos_i, h1_i, h2_i, r3_i) = _run(c, tau, feats, aenc, save=True)
    want_z, cos, H1, H2, R3 = _trunk64(tau, feats, aenc, _params(c))
    _check(z, want_z, "z (save mode)")
    z0, _ = _run(c, tau, feats, aenc)
    assert torch.equal(z, z0), "save mode must not change the arithmetic"
    M = B * N
    for name, img, want, width in (("cos", cos_i, cos, 64), ("H1", h1_i, H1, 256), ("H2", h2_i, H2, 128), ("R3", r3_i, R3, 128)):
        got = _decode(img)[:M]
        want = want.reshape(M, width)
        scale = want.abs().max().item()
        err = (got - want).abs().max().item()
        assert err <= 5e-6 * scale + 1e-7, f"{name} image: max abs err {err:.3e} at scale {scale:.3e}"


def _encode(x):
    """(M, C) float -> (tiles, C / 64, 32 KB) uint8 operand image (fp16 hi | lo, 128-byte swizzle); inverse of _decode."""
    M, Cc = x.shape
    tiles, chunks = (M + 127) // 128, Cc // 64
    xp = torch.zeros((tiles * 128, Cc), dtype=torch.float32, device=x.device)
    xp[:M] = x.float()
    hi = xp.half()
    lo = (xp - hi.float()).half()
    v = torch.stack((hi, lo), 0).reshape(2, tiles, 128, chunks, 64).permute(1, 3, 0, 2, 4)   # tile, chunk, piece, r, k
    r = torch.arange(128, device=x.device).view(128, 1)
    k = torch.arange(64, device=x.device).view(1, 64)
    idx = (((r >> 3) * 1024 + (r & 7) * 128 + (((k >> 3) ^ (r & 7)) << 4) + (k & 7) * 2) // 2).reshape(-1)
    out = torch.zeros((tiles, chunks, 2, 128 * 64), dtype=torch.float16, device=x.device)
    out[..., idx] = v.reshape(tiles, chunks, 2, 128 * 64)
    return out.view(torch.uint8).reshape(tiles, chunks, 32768).contiguous()


def test_encode_decode_round_trip():
    g = torch.Generator(device=DEV).manual_seed(0)
    x = torch.randn((300, 128), device=DEV, generator=g)
    got = _decode(_encode(x))[:300]
    assert (got - x.double()).abs().max().item() <= 2e-7 * x.abs().max().item()


@pytest.mark.parametrize("M", [32768, 512, 8192 + 128])
def test_wgrad_from_images(M):
    """K12 alone on synthetic images: every product against float64 on the decoded (hi + lo) operands."""
    g = torch.Generator(device=DEV).manual_seed(M)
    rnd = lambda *s: torch.randn(s, device=DEV, generator=g)
    S, S1 = 256.0, 2.0
    dp3, dp2, dp1 = rnd(M, 128) * 8, rnd(M, 128) * 4, rnd(M, 256) * 2
    h2, h1, cos, r3 = torch.relu(rnd(M, 128)) * 3, torch.relu(rnd(M, 256)) * 20, torch.cos(rnd(M, 64) * 50), torch.relu(rnd(M, 128))
    dz = rnd(M) * 100
    aux = torch.zeros((M, 64), device=DEV)
    aux[:, 0], aux[:, 1] = dz, 1.0
    imgs = [_encode(t) for t in (dp3, dp2, dp1, aux, cos, h1, h2, r3)]
    scales = torch.tensor([1.0 / S, 1.0 / S1], device=DEV)
    B, N = M // 32, 32
    got = _lib.critic_wgrad(*imgs, scales, B, N)
    torch.cuda.synchronize()
    _lib.critic_wgrad_status()
    d = [_decode(i)[:M] for i in imgs]
    dp3, dp2, dp1, aux, cos, h1, h2, r3 = d
    want = ((dp1.t() @ cos) / S1, dp1.sum(0) / S1, (dp2.t() @ h1) / S, dp2.sum(0) / S, (dp3.t() @ h2) / S, dp3.sum(0) / S,
            (r3.t() @ aux[:, 0]) / S)
    for name, a, w in zip(("d_wc", "d_bc", "d_wh", "d_bh", "d_wh2", "d_bh2", "d_wo"), got, want):
        scale = w.abs().max().item()
        err = (a.double() - w).abs().max().item()
        assert err <= 3e-6 * scale, f"{name}: max abs err {err:.3e} at scale {scale:.3e}"


def _autograd64(tau, feats, aenc, ps, dz, gates=None):
    """float64 autograd of the trunk: gradients of feats, aenc and the eight parameters for upstream dz.
    ``gates``: (g1, g2, g3) boolean ReLU gates to use instead of the float64 path's own (see the test below);
    also returns the float64 pre-activations."""
    f64 = feats.double().requires_grad_(True)
    a64 = aenc.double().requires_grad_(True)
    p64 = [p.detach().double().requires_grad_(True) for p in ps]
    Wc, bc, Wh, bh, Wh2, bh2, wo, bo = p64
    B, N = tau.shape
    pis = torch.tensor([np.pi * i for i in range(64)], dtype=torch.float32, device=tau.device)
    cos = torch.cos((tau.reshape(B, N, 1) * pis.view(1, 1, 64)).double())
    act = (lambda P, i: torch.relu(P)) if gates is None else (lambda P, i: P * gates[i].reshape(P.shape))
    P1 = cos @ Wc.t() + bc
    H1 = f64.unsqueeze(1) * act(P1, 0)
    P2 = H1 @ Wh.t() + bh
    H2 = a64.unsqueeze(1) * act(P2, 1)
    P3 = H2 @ Wh2.t() + bh2
    z = act(P3, 2) @ wo + bo
    z.backward(dz.double())
    return f64.grad, a64.grad, [p.grad for p in p64], [P.detach() for P in (P1, P2, P3)]


def _grad_check(name, got, want, tol=1e-5):
    scale = want.abs().max().item()
    err = (got.double() - want).abs().max().item()
    assert err <= tol * scale + 1e-12, f"{name}: max abs err {err:.3e} at scale {scale:.3e} ({err / max(scale, 1e-300):.2e} rel)"


@pytest.mark.parametrize("B,N,fs", [(1024, 32, 1.0), (64, 8, 1.0), (37, 8, 1.0), (512, 32, 150.0), (2048, 16, 1.0), (300, 1, 1.0)])
def test_backward_chain_and_wgrad_match_float64_autograd(B, N, fs):
    """K10 (SAVE) -> K11 -> K12 against float64 autograd of the same trunk, gradients to 1e-5 of each tensor's scale.

    ReLU gates of pre-activations within rounding error of zero differ between ANY float32-level evaluation and
    float64 (torch's own float32 autograd included), and one flipped gate moves a weight gradient by ~1 / sqrt(B N) of
    its scale.  So, as for the env step's flags, the gates are compared separately — they must agree with float64
    wherever |pre-activation| exceeds a 2e-5 margin of the layer's scale — and the gradient arithmetic is checked with
    the float64 path using the product's gates."""
    c = _critic()
    tau, feats, aenc = _inputs(B, N, 5 * B + N, feat_scale=fs)
    ps = _params(c)
    g = torch.Generator(device=DEV).manual_seed(B)
    dz = (torch.randn((B, N), device=DEV, generator=g) * (1.0 / (B * N))).contiguous()   # the scale of a mean loss
    _, _, (cos_i, h1_i, h2_i, r3_i) = _run(c, tau, feats, aenc, save=True)
    daenc, dfeats, (dp3, dp2, dp1, aux), scales = _lib.critic_backward(dz, r3_i, h2_i, h1_i, feats, aenc, ps[2], ps[4], ps[6], N,
                                                                      fmax=feats.amax().reshape(1))
    grads = _lib.critic_wgrad(dp3, dp2, dp1, aux, cos_i, h1_i, h2_i, r3_i, scales, B, N)
    daenc_t = _lib.critic_backward(dz, r3_i, h2_i, None, feats, aenc, ps[2], ps[4], ps[6], N, full=False)
    torch.cuda.synchronize()
    _lib.critic_backward_status()
    _lib.critic_wgrad_status()
    M = B * N
    Ps = _autograd64(tau, feats, aenc, ps, dz)[3]
    fb = (feats > 0).unsqueeze(1).expand(B, N, 256).reshape(M, 256)
    ab = (aenc > 0).unsqueeze(1).expand(B, N, 128).reshape(M, 128)
    gates, flips = [], 0
    fe = feats.double().unsqueeze(1).expand(B, N, 256).reshape(M, 256)
    ae = aenc.double().unsqueeze(1).expand(B, N, 128).reshape(M, 128)
    for img, P, known, fac in ((h1_i, Ps[0].reshape(M, 256), fb, fe), (h2_i, Ps[1].reshape(M, 128), ab, ae),
                               (r3_i, Ps[2].reshape(M, 128), None, None)):
        gate = _decode(img)[:M] > 0
        if known is not None:
            # the saved product factor * relu(P) says nothing where the factor is zero, or where the product is below the
            # two-piece fp16 image's resolution (2^-25); there dP = dH . factor . gate is zero / negligible either way
            # (an image entry > 0 always is the product's open gate; a pre-activation inside the rounding margin whose
            # product vanished is a closed gate of the product — one of the flips counted below)
            gate = gate | (~gate & (fac * P.abs() <= 1e-7) & (P > 2e-5 * P.abs().max()))
        diff = gate != (P > 0)
        flips += int(diff.sum())
        assert float((P.abs() * diff).max()) <= 2e-5 * float(P.abs().max()), "a gate differs outside the rounding margin"
        gates.append(gate.double())
    assert flips <= 1e-4 * 4 * M * 128 + 8, f"{flips} flipped gates"   # 4 M 128 gates; ~1e-4 of them lie inside the margin
    wf, wa, wp, _ = _autograd64(tau, feats, aenc, ps, dz, gates)
    # feats and aenc are ReLU outputs: where they are exactly zero the producing gate is closed and the gradient is
    # discarded downstream (ACIQN:9-13, 441-451); K11 returns 0 there (it divides the saved product by the factor)
    wf, wa = wf * (feats > 0), wa * (aenc > 0)
    # K11 recovers relu(P + b) as H / factor from the two-piece fp16 image of H (absolute resolution 2^-25): entries
    # whose factor is below ~1e-2 of the tensor's scale lose relative accuracy in proportion; what flows on into the
    # parameter gradients is factor-weighted, so that is the second yardstick
    for name, got, want, fac in (("dfeats", dfeats, wf, feats), ("daenc", daenc, wa, aenc), ("daenc (actor pass)", daenc_t, wa, aenc)):
        big = fac >= 1e-2 * fac.max()
        _grad_check(name + " where the factor is >= 1e-2 of its scale", got * big, want * big)
        _grad_check(name + " x factor", got * fac, want * fac)
        _grad_check(name + " overall", got, want, tol=2e-2)
    d_wc, d_bc, d_wh, d_bh, d_wh2, d_bh2, d_wo = grads
    for name, got, want in (("d_wc", d_wc, wp[0]), ("d_bc", d_bc, wp[1]), ("d_wh", d_wh, wp[2]), ("d_bh", d_bh, wp[3]),
                            ("d_wh2", d_wh2, wp[4]), ("d_bh2", d_bh2, wp[5]), ("d_wo", d_wo, wp[6]), ("d_bo", scales[2:3], wp[7])):
        _grad_check(name, got, want)


@pytest.mark.parametrize("M", [4096, 300, 1])
def test_iqn_action_values_match_float64_network(M):
    """IQN's act (AGENT:227-250) through K6 + K10 (head 1): mean over the 32 quantile rows, against the float64 module."""
    import copy
    from test_gpu_policy import _obs
    ag = Agent(device=DEV, seed=5, agent_type="IQN", BATCH_SIZE=64, num_tau=8, BUFFER_SIZE=256)
    net = ag.policy_local
    x = _obs(M, 3 + M)
    ts = net.tau_sampler
    q = net.action_values_fused(x, cvar=0.9)
    torch.cuda.synchronize()
    _lib.critic_forward_status()
    assert q is not None and q.shape == (M, net.action_size)
    taus, _ = _lib.iqn_tau_cos(M, 32, 64, DEV, seed=ts.seed, counter=ts.counter, cvar=0.9, counter_dev=ts.counter_dev)
    ref = copy.deepcopy(net).double()
    ref.fused_tower = False
    feats64 = ref._features(tuple(t.double() for t in x))
    ps = (net.cos_embedding.weight, net.cos_embedding.bias, net.hidden_layer.weight, net.hidden_layer.bias,
          net.hidden_layer_2.weight, net.hidden_layer_2.bias, torch.zeros(128, device=DEV), torch.zeros(1, device=DEV))
    R3 = _trunk64(taus, feats64, None, ps)[4]                                 # (M, 32, 128), IQN:74-85
    want = (R3 @ ref.output_layer.weight.t() + ref.output_layer.bias).mean(dim=1)   # IQN:86, AGENT:239
    scale = want.abs().max().item()
    err = (q.double() - want).abs().max().item()
    assert err <= RTOL * scale + 1e-6, f"action values: max abs err {err:.3e} at scale {scale:.3e}"
    # the greedy action agrees wherever the two best values are further apart than the error bar
    top2 = want.topk(2, dim=1).values
    clear = (top2[:, 0] - top2[:, 1]) > 4 * RTOL * scale
    assert torch.equal(q.argmax(1)[clear], want.argmax(1)[clear])


@pytest.mark.parametrize("B,N,fs", [(1024, 32, 150.0), (64, 8, 1.0), (37, 8, 9.0)])
def test_feats_maximum_left_by_the_pack_launch(B, N, fs):
    """K10's weight-pack launch leaves max feats behind K11's image: K11 with fmax = NULL / repack = 0 on that workspace is
    bit-identical to K11 with the caller's feats.amax() (the scale S1 is a power of two chosen from the same maximum)."""
    c = _critic()
    tau, feats, aenc = _inputs(B, N, 7 * B + N, feat_scale=fs)
    ps = _params(c)
    g = torch.Generator(device=DEV).manual_seed(B + 1)
    dz = (torch.randn((B, N), device=DEV, generator=g) * (1.0 / (B * N))).contiguous()
    bws = _lib.critic_backward_workspace(DEV, shared=False)
    bws.fill_(0xFF)   # NaN words where nothing is written
    _, _, (cos_i, h1_i, h2_i, r3_i) = _run(c, tau, feats, aenc, save=True, pack_backward=bws)
    tail = bws[-256:].view(torch.float32)
    assert float(tail.max()) == float(feats.amax()) and not bool(torch.isnan(tail).any())
    a = _lib.critic_backward(dz, r3_i, h2_i, h1_i, feats, aenc, ps[2], ps[4], ps[6], N, fmax=None, repack=False, workspace=bws)
    b = _lib.critic_backward(dz, r3_i, h2_i, h1_i, feats, aenc, ps[2], ps[4], ps[6], N, fmax=feats.amax().reshape(1))
    torch.cuda.synchronize()
    _lib.critic_backward_status()
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]) and torch.equal(a[3], b[3])
    # the dP images (their padding is never written, so compare what K12 makes of them)
    ga = _lib.critic_wgrad(*a[2], cos_i, h1_i, h2_i, r3_i, a[3], B, N)
    gb = _lib.critic_wgrad(*b[2], cos_i, h1_i, h2_i, r3_i, b[3], B, N)
    for x, y in zip(ga, gb):
        assert torch.equal(x, y)


@pytest.mark.parametrize("n", [1, 37, 512, 32768, 32768 + 5, 300000])
def test_neg_mean(n):
    g = torch.Generator(device=DEV).manual_seed(n)
    x = torch.randn((n + 1,), device=DEV, generator=g) * 3 + 0.5
    for v in (x[:n], x[1:]):   # 16-byte aligned and not
        got = _lib.neg_mean(v)
        want = -v.double().mean()
        assert got.shape == () and abs(float(got) - float(want)) <= 1e-6 * max(1.0, float(v.abs().mean()))
