"""torch-facing wrappers of K3 (tau + cosine embedding) and K4 (fused quantile-Huber loss with
its backward).  Both are hand-written sm_100a kernels behind the C-ABI (include/asvrl.h); they
need CUDA tensors and raise otherwise — there is no CPU fallback.

Reference: rfarl/rfarl/policy/AC_IQN_model.py:410-426 (calc_cos), rfarl/rfarl/agent.py:401-412
and 701-707 (loss core).
"""
import torch

from .. import _lib


class TauSampler:
    """Counter-based tau stream of one network (replaces torch.rand at ACIQN:419 / IQN:65).

    ``counter`` advances by one per call on the host; ``counter_dev`` (optional int64 device
    word) is added on the device so a captured CUDA graph draws fresh taus on every replay."""

    def __init__(self, seed, n_cos=64):
        self.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        self.n_cos = int(n_cos)
        self.counter = 0
        self.counter_dev = None

    def draw(self, batch_size, num_tau, cvar, device, taus=None):
        """-> cos (B, N, n_cos), taus (B, N, 1).  ``taus`` injects the samples (parity tests)."""
        if taus is not None:
            tau_in = taus.reshape(batch_size, num_tau).to(torch.float32).contiguous()
            tau, cos = _lib.iqn_tau_cos(batch_size, num_tau, self.n_cos, device, tau_in=tau_in)
        else:
            self.counter += 1
            tau, cos = _lib.iqn_tau_cos(batch_size, num_tau, self.n_cos, device, seed=self.seed,
                                        counter=self.counter, cvar=float(cvar),
                                        counter_dev=self.counter_dev)
        return cos, tau.unsqueeze(-1)


class _QuantileHuber(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, tau, z_next, r, done, gamma, kappa, workspace):
        loss, lps, grad = _lib.qhuber_fwd_bwd(z.contiguous(), tau.contiguous(), z_next.contiguous(),
                                              r.contiguous(), done.contiguous(), float(gamma),
                                              float(kappa), workspace)
        ctx.save_for_backward(grad)
        ctx.mark_non_differentiable(lps)
        return loss.reshape(()), lps

    @staticmethod
    def backward(ctx, g_loss, _g_lps):
        (grad,) = ctx.saved_tensors
        return grad * g_loss, None, None, None, None, None, None, None


_WS = {}


def quantile_huber_loss(z, tau, z_next, r, done, gamma=0.99, kappa=1.0):
    """AGENT:401-412 in one launch: z (B,N) expected quantiles (differentiable), tau (B,N) their
    fractions, z_next (B,N') target-network quantiles, r (B), done (B).  Returns (scalar loss,
    per-sample loss); d loss / d z comes from the same launch (targets and the indicator are
    constants, AGENT:400, 409)."""
    if not z.is_cuda:
        raise RuntimeError("quantile_huber_loss runs on CUDA only (no CPU fallback)")
    key = (z.device.index, torch.cuda.current_stream(z.device).cuda_stream)
    ws = _WS.get(key)
    if ws is None:
        ws = _WS[key] = torch.zeros(1, dtype=torch.int32, device=z.device)
    return _QuantileHuber.apply(z, tau.reshape(z.shape), z_next, r.reshape(-1), done.reshape(-1),
                                gamma, kappa, ws)
