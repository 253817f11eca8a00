"""Critic gradients on realistic observation ranges (the inputs of tests/test_gpu_trunk.py): torch's float32 module
path and the tensor-core path (K6 + K10 / K11 / K12), each against the float64 module path."""
import copy
import sys

sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import torch
from distributional_rl_decision_and_control_b200.policy import iqn_ops
from distributional_rl_decision_and_control_b200.policy.AC_IQN_model import Critic
from test_gpu_trunk import _critic_inputs

DEV = "cuda:0"
B, N = 1024, 32
crit = Critic(7, 5, 5, 56, 40, 256, 128, 2, device=DEV, seed=101).to(DEV)
x, act = _critic_inputs(B, 7 + B)
taus = torch.rand((B, N), device=DEV, generator=torch.Generator(device=DEV).manual_seed(1))
wz = torch.randn((B, N), device=DEV, generator=torch.Generator(device=DEV).manual_seed(2))


def run(c, fn, xs, a0, tz, w):
    for p in c.parameters():
        p.grad = None
    a = a0.clone().requires_grad_(True)
    z, _ = fn(xs, a, N, taus=tz)
    (z * w).sum().backward()
    return z.detach(), a.grad, {n: p.grad.clone() for n, p in c.named_parameters()}


c64 = copy.deepcopy(crit).double()
c64.fused_tower = False
pis32 = torch.tensor([3.141592653589793 * i for i in range(64)], dtype=torch.float32, device=DEV)


def cos64(batch_size, num_tau=8, cvar=1.0, taus=None):   # the reference's float32 argument, cosine in float64
    t = taus.reshape(batch_size, num_tau, 1).float()
    return torch.cos((t * pis32.view(1, 1, 64)).double()), t.double()


c64.calc_cos = cos64
z64, da64, g64 = run(c64, c64.forward, tuple(t.double() for t in x), act.double(), taus.double(), wz.double())
crit.fused_tower = False
z0, da0, g0 = run(crit, crit.forward, x, act, taus, wz)
crit.fused_tower = True
iqn_ops.UMMA_TRUNK = True
z1, da1, g1 = run(crit, crit.forward_fused, x, act, taus, wz)
iqn_ops.UMMA_TRUNK = False
z2, da2, g2 = run(crit, crit.forward_fused, x, act, taus, wz)
rel = lambda a, b: float((a.double() - b).abs().max() / b.abs().max())
print("z           torch32 %.2e  K5 %.2e  tensor %.2e" % (rel(z0, z64), rel(z2, z64), rel(z1, z64)))
print("d/d actions torch32 %.2e  K5 %.2e  tensor %.2e" % (rel(da0, da64), rel(da2, da64), rel(da1, da64)))
for n in g64:
    print("%-28s torch32 %.2e  K5 %.2e  tensor %.2e" % (n, rel(g0[n], g64[n]), rel(g2[n], g64[n]), rel(g1[n], g64[n])))
