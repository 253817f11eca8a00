"""torch-facing wrappers of K3 (tau + cosine embedding) and K4 (fused quantile-Huber loss with
its backward).  Both are hand-written sm_100a kernels behind the C-ABI (include/asvrl.h); they
need CUDA tensors and raise otherwise — there is no CPU fallback.

Reference: rfarl/rfarl/policy/AC_IQN_model.py:410-426 (calc_cos), rfarl/rfarl/agent.py:401-412
and 701-707 (loss core).
"""
import os

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


def quantile_huber_loss_grad(z, tau, z_next, r, done, gamma=0.99, kappa=1.0):
    """The same launch without an autograd node: -> (scalar loss, per-sample loss, d loss / d z).  For a caller that
    continues with ``z.backward(grad)`` itself — ``loss.backward()`` would only multiply ``grad`` by a freshly filled
    one and zero-fill a gradient for the per-sample output (three small launches per update)."""
    if not z.is_cuda:
        raise RuntimeError("quantile_huber_loss_grad runs on CUDA only (no CPU fallback)")
    key = (z.device.index, torch.cuda.current_stream(z.device).cuda_stream)
    ws = _WS.get(key)
    if ws is None:
        ws = _WS[key] = torch.zeros(1, dtype=torch.int32, device=z.device)
    with torch.no_grad():
        loss, lps, grad = _lib.qhuber_fwd_bwd(z.detach().contiguous(), tau.reshape(z.shape).contiguous(),
                                              z_next.contiguous(), r.reshape(-1).contiguous(),
                                              done.reshape(-1).contiguous(), float(gamma), float(kappa), ws)
    return loss.reshape(()), lps, grad


def _wgrad(d_out, x, splits=64):
    """d_out^T x for (M, a) and (M, b) with M = B*N in the tens of thousands: cuBLAS runs the plain
    product (K = M, a 128 x 256 output) on a handful of CTAs (~110 us at C3), so the contraction is
    split into `splits` batched GEMMs and the partial products are summed (~35-50 us)."""
    M = d_out.shape[0]
    if (x.shape[1] <= 8 or d_out.shape[1] <= 8) and M >= 256 and M % 16 == 0:
        # a 2-wide factor (the action encoder's input, the actor's output): cuBLAS runs the thin product on one CTA
        # (~27 us at M = 1024); 16 batched slices and a sum take ~6 us
        part = torch.bmm(d_out.view(16, M // 16, -1).transpose(1, 2), x.view(16, M // 16, -1))
        return part.sum(0)
    if M >= 4096 and M % splits == 0:
        part = torch.bmm(d_out.view(splits, M // splits, -1).transpose(1, 2), x.view(splits, M // splits, -1))
        return part.sum(0)
    return d_out.t() @ x


# forward GEMMs of the fused trunks through K8 (tcgen05, 3 x TF32) in fp32 mode; ASVRL_UMMA_TRUNK=0/1 overrides
UMMA_FORWARD = os.environ.get("ASVRL_UMMA_TRUNK", "0") == "1"
UMMA_MIN_ROWS = 4096  # below this a 128-row-tile kernel leaves most SMs idle: cuBLAS


def _gemm_nt(x, W):
    """x @ W.T without epilogue.  When the caller asked for fp32 arithmetic (TF32 not allowed) and the
    shape fits, K8 computes it on the tensor cores as a three-pass TF32 split at fp32-GEMM accuracy
    (asvrl_linear_tf32x3); otherwise cuBLAS (fp32 SIMT, or one-pass TF32 when allowed)."""
    M, K = x.shape
    N = W.shape[0]
    if (UMMA_FORWARD and M >= UMMA_MIN_ROWS and x.dtype == torch.float32 and K % 32 == 0 and N % 16 == 0
            and N <= 256 and not torch.backends.cuda.matmul.allow_tf32 and W.is_contiguous()):
        return _lib.linear_tf32x3(x.contiguous(), W.detach(), None, relu=False)
    return x @ W.t()


class _CriticTrunk(torch.autograd.Function):
    """The critic's quantile trunk (ACIQN:453-480) with K5 between the GEMMs:

        P1 = cos Wc^T          H1 = feats[b] * relu(P1 + bc)
        P2 = H1 Wh^T           H2 = aenc[b]  * relu(P2 + bh)
        P3 = H2 Wh2^T          z  = relu(P3 + bh2) . wo + bo

    The three contractions are plain torch.mm calls (cuBLAS); everything else is one launch per
    layer forward and one backward.  The backward computes only what ``needs_input_grad`` asks
    for: the actor update needs d z / d aenc alone (one GEMM and two launches)."""

    @staticmethod
    def forward(ctx, cos, feats, aenc, Wc, bc, Wh, bh, Wh2, bh2, wo, bo, n_tau):
        cos, feats, aenc = cos.contiguous(), feats.contiguous(), aenc.contiguous()
        P1 = _gemm_nt(cos, Wc)
        H1 = _lib.iqn_scale_relu_fwd(P1, bc, feats, n_tau)
        P2 = _gemm_nt(H1, Wh)
        H2 = _lib.iqn_scale_relu_fwd(P2, bh, aenc, n_tau)
        P3 = _gemm_nt(H2, Wh2)
        z = _lib.iqn_head_fwd(P3, bh2, wo.reshape(-1), bo)
        ctx.save_for_backward(cos, feats, aenc, Wc, bc, Wh, bh, Wh2, bh2, wo, P1, H1, P2, H2, P3)
        ctx.n_tau = n_tau
        return z.view(feats.shape[0], n_tau)

    @staticmethod
    def backward(ctx, dz):
        cos, feats, aenc, Wc, bc, Wh, bh, Wh2, bh2, wo, P1, H1, P2, H2, P3 = ctx.saved_tensors
        n_tau = ctx.n_tau
        need = ctx.needs_input_grad
        g = [None] * 12
        head_params = need[7] or need[8] or need[9] or need[10]
        dP3, dwo, dbh2, dbo = _lib.iqn_head_bwd(dz.contiguous().view(-1), P3, bh2, wo.reshape(-1), head_params)
        if head_params:
            g[8], g[9], g[10] = dbh2, dwo.view_as(wo), dbo
            if need[7]:
                g[7] = _wgrad(dP3, H2)
        layer1 = need[1] or need[3] or need[4]
        layer2 = layer1 or need[5] or need[6]
        if not (layer2 or need[2]):
            return tuple(g)
        dH2 = dP3 @ Wh2
        dP2, daenc, dbh = _lib.iqn_scale_relu_bwd(dH2, P2, bh, aenc, n_tau, want_dpre=layer2,
                                                  want_dscale=need[2], want_dbias=need[6])
        g[2], g[6] = daenc, dbh
        if not layer2:
            return tuple(g)
        if need[5]:
            g[5] = _wgrad(dP2, H1)
        if layer1:
            dH1 = dP2 @ Wh
            dP1, dfeats, dbc = _lib.iqn_scale_relu_bwd(dH1, P1, bc, feats, n_tau, want_dpre=need[3],
                                                       want_dscale=need[1], want_dbias=need[4])
            g[1], g[4] = dfeats, dbc
            if need[3]:
                g[3] = _wgrad(dP1, cos)
        return tuple(g)


# K10 / K11 / K12: the whole trunk, forward and backward, on the tcgen05 tensor cores (ASVRL_CRITIC_UMMA=0: K3 + cuBLAS + K5)
UMMA_TRUNK = os.environ.get("ASVRL_CRITIC_UMMA", "1") != "0"
UMMA_TRUNK_MIN_ROWS = int(os.environ.get("ASVRL_CRITIC_UMMA_MIN_ROWS", "4096"))  # fewer 128-row tiles leave most SMs idle


def umma_trunk_ok(feats, aenc, Wc, Wh, Wh2, n_tau, need_grad):
    """K10 - K12 are compiled for the reference's fixed trunk shape (64 -> 256 -> 128 -> 128 -> 1, AGENT:20-26) in
    float32; the backward's per-sample sums need a power-of-two quantile count <= 32 (the reference uses 8 and 32)."""
    return (UMMA_TRUNK and feats.shape[0] * n_tau >= UMMA_TRUNK_MIN_ROWS
            and feats.is_cuda and feats.dtype == torch.float32 and Wc.dtype == torch.float32
            and tuple(Wc.shape) == (256, 64) and tuple(Wh.shape) == (128, 256) and tuple(Wh2.shape) == (128, 128)
            and feats.shape[1] == 256 and (aenc is None or aenc.shape[1] == 128)
            and (not need_grad or (aenc is not None and n_tau in (1, 2, 4, 8, 16, 32))))


class _CriticTrunkUmma(torch.autograd.Function):
    """The critic's quantile trunk (ACIQN:410-480) as three tcgen05 kernels: K10 forward (cosine embedding generated in
    the kernel; in a differentiable pass it also leaves the operand images), K11 the backward chain (d feats, d aenc and
    the dP images), K12 every weight / bias gradient from the images.  When only d z / d aenc is asked for — the actor
    update, AGENT:420-427 — K11 runs its short form and K12 not at all."""

    @staticmethod
    def forward(ctx, feats, aenc, Wc, bc, Wh, bh, Wh2, bh2, wo, bo, n_tau, tau_in, draw, ws_key):
        feats, aenc = feats.contiguous(), aenc.contiguous()
        need = ctx.needs_input_grad
        save = any(need[:10])
        ws = _lib.critic_workspace(ws_key, feats.device)
        seed, counter, counter_dev, cvar = draw
        # the backward of THIS pass reads the same weights: its (transposed) image is packed by the same launch
        bws = _lib.critic_backward_workspace(feats.device, shared=False) if save else None
        out = _lib.critic_forward(feats, aenc, Wc, bc, Wh, bh, Wh2, bh2, wo.reshape(-1), bo, n_tau, ws, tau_in=tau_in,
                                  seed=seed, counter=counter, counter_dev=counter_dev, cvar=cvar, save=save,
                                  pack_backward=bws)
        z, tau = out[0], out[1]
        if save:
            # max feats (the scale of K11's dP1 image) is left behind the backward image by the same pack launch
            ctx.save_for_backward(feats, aenc, Wh, Wh2, wo, bws, *out[2])
        ctx.n_tau = n_tau
        ctx.mark_non_differentiable(tau)
        ctx.set_materialize_grads(False)   # no (B, N) zero fill for d tau
        return z, tau

    @staticmethod
    def backward(ctx, dz, _dtau):
        g = [None] * 14
        if dz is None:
            return tuple(g)
        feats, aenc, Wh, Wh2, wo, bws, cos_i, h1_i, h2_i, r3_i = ctx.saved_tensors
        need = ctx.needs_input_grad
        n_tau = ctx.n_tau
        dz = dz.contiguous()
        if need[0] or any(need[2:10]):
            daenc, dfeats, (dp3, dp2, dp1, aux), scales = _lib.critic_backward(dz, r3_i, h2_i, h1_i, feats, aenc, Wh, Wh2,
                                                                              wo.reshape(-1), n_tau, fmax=None, repack=False,
                                                                              workspace=bws)
            d_wc, d_bc, d_wh, d_bh, d_wh2, d_bh2, d_wo = _lib.critic_wgrad(dp3, dp2, dp1, aux, cos_i, h1_i, h2_i, r3_i,
                                                                           scales, feats.shape[0], n_tau)
            g[0], g[1] = dfeats, daenc
            g[2], g[3], g[4], g[5], g[6], g[7], g[8] = d_wc, d_bc, d_wh, d_bh, d_wh2, d_bh2, d_wo.view_as(wo)
            g[9] = scales[2:3]   # sum dz, reduced in K11's prologue
        else:
            g[1] = _lib.critic_backward(dz, r3_i, h2_i, None, feats, aenc, Wh, Wh2, wo.reshape(-1), n_tau, full=False,
                                        repack=False, workspace=bws)
        return tuple(g)


def critic_trunk_umma(feats, aenc, Wc, bc, Wh, bh, Wh2, bh2, wo, bo, n_tau, tau_in, draw, ws_key):
    """-> z (B, N), tau (B, N).  ``tau_in`` (B, N) injects the fractions, else ``draw`` = (seed, counter, counter_dev,
    cvar) of the network's Philox tau stream (the same draw as K3)."""
    if not feats.is_cuda:
        raise RuntimeError("critic_trunk_umma runs on CUDA only (no CPU fallback)")
    if torch.is_grad_enabled() and (feats.requires_grad or aenc.requires_grad or Wc.requires_grad):
        return _CriticTrunkUmma.apply(feats, aenc, Wc, bc, Wh, bh, Wh2, bh2, wo, bo, int(n_tau), tau_in, draw, ws_key)
    ws = _lib.critic_workspace(ws_key, feats.device)
    seed, counter, counter_dev, cvar = draw
    return _lib.critic_forward(feats.contiguous(), aenc.contiguous(), Wc, bc, Wh, bh, Wh2, bh2, wo.reshape(-1), bo,
                               int(n_tau), ws, tau_in=tau_in, seed=seed, counter=counter, counter_dev=counter_dev,
                               cvar=cvar)


def critic_trunk(cos, feats, aenc, Wc, bc, Wh, bh, Wh2, bh2, wo, bo, n_tau):
    """cos (B*N, n_cos), feats (B, D1), aenc (B, D2) and the four layers' parameters -> z (B, N)."""
    if not cos.is_cuda:
        raise RuntimeError("critic_trunk runs on CUDA only (no CPU fallback)")
    return _CriticTrunk.apply(cos, feats, aenc, Wc, bc, Wh, bh, Wh2, bh2, wo, bo, int(n_tau))


class _Tower(torch.autograd.Function):
    """K6: self / object encoders, mask and concat (ACIQN:284-308) in one launch; the backward is one
    launch for the four parameter gradients (observations carry no gradient)."""

    @staticmethod
    def forward(ctx, self_obs, objects, mask, Ws, bs, Wo, bo):
        self_obs, objects, mask = self_obs.contiguous(), objects.contiguous(), mask.contiguous()
        feats = _lib.tower_fwd(self_obs, objects, mask, Ws, bs, Wo, bo)
        ctx.save_for_backward(feats, self_obs, objects)
        ctx.dims = (Ws.shape[0], Wo.shape[0])
        return feats

    @staticmethod
    def backward(ctx, dfeats):
        feats, self_obs, objects = ctx.saved_tensors
        dws, dbs, dwo, dbo = _lib.tower_bwd(dfeats.contiguous(), feats, self_obs, objects, *ctx.dims)
        return None, None, None, dws, dbs, dwo, dbo


def observation_tower(self_obs, objects, mask, Ws, bs, Wo, bo):
    """-> (B, SF + K*OF) features; CUDA float32 only (no CPU fallback)."""
    if not self_obs.is_cuda:
        raise RuntimeError("observation_tower runs on CUDA only (no CPU fallback)")
    return _Tower.apply(self_obs, objects, mask, Ws, bs, Wo, bo)


class _LinearScaleRelu(torch.autograd.Function):
    """out[m] = scale[m // n_tau] * relu(x[m] W^T + bias): a cuBLAS GEMM without epilogue, then K5's
    scale_relu pass; the backward is K5's one pass (dpre, dscale, dbias) plus the two GEMMs.
    The IQN network's trunk (IQN:74-86) is three of these (scale = the state features, then ones)."""

    @staticmethod
    def forward(ctx, x, W, bias, scale, n_tau):
        x, scale = x.contiguous(), scale.contiguous()
        pre = _gemm_nt(x, W)
        out = _lib.iqn_scale_relu_fwd(pre, bias, scale, n_tau)
        ctx.save_for_backward(x, W, bias, scale, pre)
        ctx.n_tau = n_tau
        return out

    @staticmethod
    def backward(ctx, dout):
        x, W, bias, scale, pre = ctx.saved_tensors
        need = ctx.needs_input_grad
        want_dpre = need[0] or need[1]
        n_tau, B = ctx.n_tau, scale.shape[0]
        if n_tau == 1 and not need[3] and getattr(scale, "_asvrl_unit", False) and B % 32 == 0:
            # a plain Linear + ReLU over B rows (unit scale, no d scale): K5 walks them as B / 32 groups of 32 rows, so the
            # bias gradient is 16 x fewer per-CTA partials for the last CTA to add
            scale, n_tau = scale[:B // 32], 32
        dpre, dscale, dbias = _lib.iqn_scale_relu_bwd(dout.contiguous(), pre, bias, scale, n_tau,
                                                      want_dpre=want_dpre, want_dscale=need[3], want_dbias=need[2])
        dx = dpre @ W if need[0] else None
        dW = _wgrad(dpre, x) if need[1] else None
        return dx, dW, dbias, dscale, None


class _ThinLinear(torch.autograd.Function):
    """y = x W^T + b for a layer with a handful of outputs (the actor's 128 -> 2 head, ACIQN:318): the forward and
    d x are plain cuBLAS calls; the weight gradient goes through `_wgrad`'s sliced product."""

    @staticmethod
    def forward(ctx, x, W, b):
        ctx.save_for_backward(x, W)
        return torch.addmm(b, x, W.t())

    @staticmethod
    def backward(ctx, dy):
        x, W = ctx.saved_tensors
        need = ctx.needs_input_grad
        dy = dy.contiguous()
        return (dy @ W if need[0] else None, _wgrad(dy, x) if need[1] else None, dy.sum(0) if need[2] else None)


def thin_linear(x, W, b):
    return _ThinLinear.apply(x, W, b)


NARROW_MAX_ROWS = 8192   # the backward kernels of the two narrow layers use at most 64 CTAs


def narrow_ok(x, D):
    return (x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.shape[0] <= NARROW_MAX_ROWS
            and 32 <= D <= 256 and 1024 % D == 0)


class _ActionEncoder(torch.autograd.Function):
    """aenc = relu(Wa a + ba) (ACIQN:399, 472) for the B sampled states: one launch forward, one backward (d Wa, d ba and,
    for the actor update, d a) instead of sgemm + K5 and K5 backward + two thin products."""

    @staticmethod
    def forward(ctx, actions, wa, ba):
        actions = actions.contiguous()
        aenc = _lib.action_encoder_fwd(actions, wa, ba)
        ctx.save_for_backward(actions, wa, aenc)
        return aenc

    @staticmethod
    def backward(ctx, daenc):
        actions, wa, aenc = ctx.saved_tensors
        need = ctx.needs_input_grad
        dwa, dba, dact = _lib.action_encoder_bwd(daenc.contiguous(), aenc, actions, wa, want_dact=need[0])
        return dact, dwa if need[1] else None, dba if need[2] else None


def action_encoder(actions, wa, ba):
    return _ActionEncoder.apply(actions, wa, ba)


class _AtanHead(torch.autograd.Function):
    """act = s atan(W3 h + b3) (ACIQN:318-321): one launch forward, one backward."""

    @staticmethod
    def forward(ctx, h, w3, b3, scale):
        h = h.contiguous()
        y, act = _lib.atan_head_fwd(h, w3, b3, scale)
        ctx.save_for_backward(h, w3, y)
        ctx.scale = float(scale)
        return act

    @staticmethod
    def backward(ctx, dact):
        h, w3, y = ctx.saved_tensors
        need = ctx.needs_input_grad
        dh, dw3, db3 = _lib.atan_head_bwd(dact.contiguous(), y, h, w3, ctx.scale, want_dh=need[0])
        return dh, dw3 if need[1] else None, db3 if need[2] else None, None


def atan_head(h, w3, b3, scale):
    return _AtanHead.apply(h, w3, b3, scale)


def linear_scale_relu(x, W, bias, scale, n_tau):
    if not x.is_cuda:
        raise RuntimeError("linear_scale_relu runs on CUDA only (no CPU fallback)")
    return _LinearScaleRelu.apply(x, W, bias, scale, int(n_tau))
