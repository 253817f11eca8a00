import sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np, torch
from distributional_rl_decision_and_control_b200 import _lib
from test_gpu_critic import _critic, _inputs, _params, _run, _autograd64, _decode
DEV = "cuda:0"
B, N, fs = 1024, 32, 1.0
c = _critic()
tau, feats, aenc = _inputs(B, N, 5 * B + N, feat_scale=fs)
ps = _params(c)
g = torch.Generator(device=DEV).manual_seed(B)
dz = (torch.randn((B, N), device=DEV, generator=g) * (1.0 / (B * N))).contiguous()
_, _, (cos_i, h1_i, h2_i, r3_i) = _run(c, tau, feats, aenc, save=True)
daenc, dfeats, (dp3, dp2, dp1, aux), scales = _lib.critic_backward(dz, r3_i, h2_i, h1_i, feats, aenc, ps[2], ps[4], ps[6], N)
torch.cuda.synchronize()
M = B * N
Ps = _autograd64(tau, feats, aenc, ps, dz)[3]
P1, P2, P3 = Ps[0].reshape(M, 256), Ps[1].reshape(M, 128), Ps[2].reshape(M, 128)
fe = feats.double().unsqueeze(1).expand(B, N, 256).reshape(M, 256)
ae = aenc.double().unsqueeze(1).expand(B, N, 128).reshape(M, 128)
H1, H2, R3 = _decode(h1_i)[:M], _decode(h2_i)[:M], _decode(r3_i)[:M]
def mk(img, P, fac):
    gate = img > 0
    if fac is not None:
        gate = gate | (~gate & (fac * P.abs() <= 1e-7) & (P > 0))
    return gate
g1, g2, g3 = mk(H1, P1, fe), mk(H2, P2, ae), mk(R3, P3, None)
for nm, gg, P in (("g1", g1, P1), ("g2", g2, P2), ("g3", g3, P3)):
    d = gg != (P > 0)
    print(nm, "flips", int(d.sum()), "max |P| at flips", float((P.abs() * d).max()))
wf, wa, wp, _ = _autograd64(tau, feats, aenc, ps, dz, [g1.double(), g2.double(), g3.double()])
wf = wf * (feats > 0)
err = (dfeats.double() - wf).abs() * (feats >= 1e-2 * feats.max())
idx = int(err.argmax()); b, k = idx // 256, idx % 256
print("worst", b, k, "got", dfeats[b, k].item(), "want", wf[b, k].item(), "feats", feats[b, k].item())
S = 1.0 / scales[0].item(); S1 = 1.0 / scales[1].item()
rows = slice(b * N, (b + 1) * N)
DP1 = _decode(dp1)[:M]
dH1_ours = DP1[rows, k] / (feats[b, k].double() * S1)
# float64 dH1 with forced gates
Wc, bc, Wh, bh, Wh2, bh2, wo, bo = [p.detach().double() for p in ps]
dP3 = dz.reshape(M, 1).double() * wo.view(1, 128) * g3
dH2 = dP3 @ Wh2
dP2 = dH2 * ae * g2
dH1 = dP2 @ Wh
print("dH1 ours vs ref (rows of the sample):")
print(torch.stack((dH1_ours, dH1[rows, k] * g1[rows, k], H1[rows, k], (fe * P1 * g1)[rows, k]), 1))
DP2 = _decode(dp2)[:M] / S
print("dP2 row err max", float((DP2[rows] - dP2[rows]).abs().max()), "scale", float(dP2.abs().max()))
DP3 = _decode(dp3)[:M] / S
print("dP3 row err max", float((DP3[rows] - dP3[rows]).abs().max()), "scale", float(dP3.abs().max()))
print("dP3 all err", float((DP3 - dP3).abs().max()), "dP2 all err", float((DP2 - dP2).abs().max()), "dP1 all err",
      float((DP1 / S1 - (dH1 * fe * g1)).abs().max()), float((dH1 * fe * g1).abs().max()))
