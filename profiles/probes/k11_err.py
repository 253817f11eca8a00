import sys, os
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import torch
import numpy as np
from distributional_rl_decision_and_control_b200 import _lib
DEV = "cuda:0"
from test_gpu_critic import _critic, _inputs, _params, _run, _autograd64
for B, N, fs in [(1024, 32, 1.0), (64, 8, 1.0), (512, 32, 150.0), (2048, 16, 1.0)]:
    c = _critic()
    tau, feats, aenc = _inputs(B, N, 5 * B + N, feat_scale=fs)
    ps = _params(c)
    g = torch.Generator(device=DEV).manual_seed(B)
    dz = (torch.randn((B, N), device=DEV, generator=g) * (1.0 / (B * N))).contiguous()
    _, _, (cos_i, h1_i, h2_i, r3_i) = _run(c, tau, feats, aenc, save=True)
    daenc, dfeats, (dp3, dp2, dp1, aux), scales = _lib.critic_backward(dz, r3_i, h2_i, h1_i, feats, aenc, ps[2], ps[4], ps[6], N)
    grads = _lib.critic_wgrad(dp3, dp2, dp1, aux, cos_i, h1_i, h2_i, r3_i, scales, B, N)
    torch.cuda.synchronize()
    wf, wa, wp = _autograd64(tau, feats, aenc, ps, dz)
    wf, wa = wf * (feats > 0), wa * (aenc > 0)
    def rel(got, want): return ((got.double() - want).abs().max() / want.abs().max()).item()
    print(B, N, fs, "scales", scales.tolist())
    print("  dfeats %.2e daenc %.2e" % (rel(dfeats, wf), rel(daenc, wa)))
    big = feats > 1e-2 * fs
    print("  dfeats (feats>1e-2) %.2e  daenc (aenc>1e-2) %.2e" % (rel(dfeats * big, wf * big), rel(daenc * (aenc > 1e-2), wa * (aenc > 1e-2))))
    print("  weighted dfeats*feats %.2e  daenc*aenc %.2e" % (rel(dfeats * feats, wf * feats), rel(daenc * aenc, wa * aenc)))
    names = ("d_wc", "d_bc", "d_wh", "d_bh", "d_wh2", "d_bh2", "d_wo")
    print("  " + " ".join("%s %.2e" % (n, rel(gq, w)) for n, gq, w in zip(names, grads, wp[:7])))
    # the fp32 torch path for comparison: autograd in fp32
    f32 = feats.clone().requires_grad_(True); a32 = aenc.clone().requires_grad_(True)
    p32 = [p.detach().clone().requires_grad_(True) for p in ps]
    pis = torch.tensor([np.pi * i for i in range(64)], dtype=torch.float32, device=DEV)
    cos = torch.cos(tau.reshape(B, N, 1) * pis.view(1, 1, 64))
    H1 = f32.unsqueeze(1) * torch.relu(cos @ p32[0].t() + p32[1])
    H2 = a32.unsqueeze(1) * torch.relu(H1 @ p32[2].t() + p32[3])
    z = torch.relu(H2 @ p32[4].t() + p32[5]) @ p32[6] + p32[7]
    z.backward(dz)
    print("  torch fp32: dfeats %.2e daenc %.2e " % (rel(f32.grad * (feats > 0), wf), rel(a32.grad * (aenc > 0), wa)) + " ".join("%s %.2e" % (n, rel(p.grad, w)) for n, p, w in zip(names, p32[:7], wp[:7])))
