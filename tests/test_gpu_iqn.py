"""K3 / K4 parity on the B200 through the C-ABI."""
import numpy as np
import pytest
import torch

from distributional_rl_decision_and_control_b200 import _lib
import golden_util as gu

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
RTOL = 1e-5


def gpu_qhuber(z, tau, zn, r, done, gamma=0.99, kappa=1.0):
    ws = torch.zeros(1, dtype=torch.int32, device=DEV)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(DEV)
    loss, lps, grad = _lib.qhuber_fwd_bwd(t(z), t(tau), t(zn), t(r), t(done), gamma, kappa, ws)
    torch.cuda.synchronize()
    assert int(ws.item()) == 0  # workspace left clean
    return float(loss.item()), lps.cpu().numpy(), grad.cpu().numpy()


def test_qhuber_golden_reference():
    """AGENT:401-412 evaluated by the reference's own calculate_huber_loss + autograd."""
    d = gu.load("iqn_loss.npz")
    loss, lps, grad = gpu_qhuber(d["z"], d["tau"], d["z_next"], d["r"], d["done"], float(d["gamma"]))
    assert abs(loss - float(d["loss"])) <= RTOL * abs(float(d["loss"]))
    assert np.allclose(lps, d["loss_per_sample"], rtol=RTOL, atol=1e-6)
    assert np.allclose(grad, d["grad_z"], rtol=RTOL, atol=1e-9)


@pytest.mark.parametrize("B,N,Np,kappa", [(1024, 32, 32, 1.0), (64, 8, 8, 1.0), (7, 5, 13, 0.5),
                                          (300, 40, 70, 2.0), (1, 1, 1, 1.0), (4096, 8, 32, 1.0)])
def test_qhuber_oracle(oracle, B, N, Np, kappa):
    rng = np.random.RandomState(B + N)
    f = lambda a: np.asarray(a, np.float32).astype(np.float64)
    z, zn = f(3 * rng.randn(B, N)), f(3 * rng.randn(B, Np))
    tau, r = f(rng.rand(B, N)), f(rng.randn(B))
    done = (rng.rand(B) < 0.1).astype(np.float64)
    want_loss, want_lps, want_grad = oracle.qhuber(z, tau, zn, r, done, 0.99, kappa)
    loss, lps, grad = gpu_qhuber(z, tau, zn, r, done, 0.99, kappa)
    assert abs(loss - want_loss) <= RTOL * abs(want_loss)
    assert np.allclose(lps, want_lps, rtol=RTOL, atol=1e-6)
    # gradient entries are sums of N' signed terms of size <= 1/(B N'): relative to that scale
    assert np.max(np.abs(grad - want_grad)) <= RTOL / B


def test_qhuber_matches_torch_autograd_fp32():
    """The same loss written as the reference writes it, in stock PyTorch fp32 on the GPU."""
    B, N, Np = 512, 32, 32
    g = torch.Generator(device="cpu").manual_seed(3)
    z = (3 * torch.randn(B, N, generator=g)).to(DEV).requires_grad_(True)
    zn = (3 * torch.randn(B, Np, generator=g)).to(DEV)
    tau = torch.rand(B, N, generator=g).to(DEV)
    r = torch.randn(B, generator=g).to(DEV)
    done = (torch.rand(B, generator=g) < 0.1).float().to(DEV)
    q_t = r[:, None, None] + 0.99 * zn[:, None, :] * (1.0 - done[:, None, None])
    td = q_t - z[:, :, None]
    hub = torch.where(td.abs() <= 1.0, 0.5 * td.pow(2), 1.0 * (td.abs() - 0.5))
    ql = (tau[:, :, None] - (td.detach() < 0).float()).abs() * hub
    loss = ql.sum(1).mean(1).mean()
    loss.backward()
    ws = torch.zeros(1, dtype=torch.int32, device=DEV)
    l2, _, g2 = _lib.qhuber_fwd_bwd(z.detach(), tau, zn, r, done, 0.99, 1.0, ws)
    assert abs(l2.item() - loss.item()) <= RTOL * abs(loss.item())
    assert torch.max((g2 - z.grad).abs()).item() <= RTOL / B


def test_qhuber_deterministic():
    rng = np.random.RandomState(0)
    args = (rng.randn(1024, 32), rng.rand(1024, 32), rng.randn(1024, 32), rng.randn(1024),
            (rng.rand(1024) < 0.1).astype(np.float64))
    a = gpu_qhuber(*args)
    b = gpu_qhuber(*args)
    assert a[0] == b[0] and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])


def test_tau_cos_golden_reference():
    """cos table of the reference Critic.calc_cos (torch CPU float32) on injected tau."""
    d = gu.load("iqn_loss.npz")
    tau = torch.from_numpy(d["tau"].astype(np.float32)).to(DEV)
    tau_out, cos = _lib.iqn_tau_cos(tau.shape[0], tau.shape[1], 64, DEV, tau_in=tau)
    torch.cuda.synchronize()
    assert torch.equal(tau_out, tau)
    assert np.max(np.abs(cos.cpu().numpy() - d["cos"])) <= 2.5e-7  # 2 ulp of cosf near 1


def test_tau_cos_oracle_c3_and_ragged(oracle):
    for B, N, n_cos in ((1024, 32, 64), (5, 7, 64), (3, 1, 16), (33, 8, 100), (4, 3, 30), (9, 2, 7)):
        rng = np.random.RandomState(B)
        tau = rng.rand(B, N).astype(np.float32)
        _, cos = _lib.iqn_tau_cos(B, N, n_cos, DEV, tau_in=torch.from_numpy(tau).to(DEV))
        want = oracle.tau_cos(tau.astype(np.float64), n_cos)
        assert np.max(np.abs(cos.cpu().numpy() - want)) <= 2.5e-7


def test_tau_rng_uniform_and_reproducible():
    from scipy import stats
    t1, c1 = _lib.iqn_tau_cos(1024, 32, 64, DEV, seed=5, counter=9)
    t2, _ = _lib.iqn_tau_cos(1024, 32, 64, DEV, seed=5, counter=9)
    t3, _ = _lib.iqn_tau_cos(1024, 32, 64, DEV, seed=5, counter=10)
    assert torch.equal(t1, t2) and not torch.equal(t1, t3)
    x = t1.cpu().numpy().ravel().astype(np.float64)
    assert x.min() > 0.0 and x.max() < 1.0
    assert stats.kstest(x, "uniform").pvalue > 1e-3
    th, _ = _lib.iqn_tau_cos(64, 8, 64, DEV, seed=5, counter=9, cvar=0.25)
    assert th.max().item() < 0.25
    pis = torch.from_numpy((np.pi * np.arange(64)).astype(np.float32)).to(DEV)
    assert torch.max((c1 - torch.cos(t1[..., None] * pis)).abs()).item() <= 2.5e-7
