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
