import sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import torch
from distributional_rl_decision_and_control_b200 import _lib
from test_gpu_critic import _critic, _inputs, _params, _run
for B, N, fs in [(1024, 32, 1.0), (1024, 32, 150.0), (4096, 32, 1.0), (512, 8, 1.0)]:
    c = _critic(); ps = _params(c)
    tau, feats, aenc = _inputs(B, N, 3, feat_scale=fs)
    dz = torch.randn((B, N), device="cuda:0").contiguous()
    _, _, (ci, h1i, h2i, r3i) = _run(c, tau, feats, aenc, save=True)
    fm = feats.amax().reshape(1)
    outs = []
    for rep in range(6):
        da, df, imgs, sc = _lib.critic_backward(dz, r3i, h2i, h1i, feats, aenc, ps[2], ps[4], ps[6], N, fmax=fm)
        dat = _lib.critic_backward(dz, r3i, h2i, None, feats, aenc, ps[2], ps[4], ps[6], N, full=False)
        g = _lib.critic_wgrad(*imgs, ci, h1i, h2i, r3i, sc, B, N)
        torch.cuda.synchronize()
        outs.append((da.clone(), df.clone(), dat.clone(), [x.clone() for x in g]))
    ok_rep = all(torch.equal(outs[0][0], o[0]) and torch.equal(outs[0][1], o[1]) and torch.equal(outs[0][2], o[2])
                 and all(torch.equal(a, b) for a, b in zip(outs[0][3], o[3])) for o in outs[1:])
    print(B, N, fs, "repeatable", ok_rep, "full==trunc daenc", torch.equal(outs[0][0], outs[0][2]),
          "max diff", float((outs[0][0] - outs[0][2]).abs().max()), "scale", float(outs[0][0].abs().max()))
