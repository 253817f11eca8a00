"""Timeline of CTA 0 of K11 (debug build: ASVRL_NVCC_FLAGS=-DCF_TRACE python -m ...csrc.build --force).
writer 0 = MMA thread (16 s + 1 / 2: Sa start / issued, + 3 / 4: Sb), 1 / 2 = slot 0 / 1 epilogue (1 tile start, 2-3 dP3
chunks published, 4 dH2 ready, 5-6 dP2 published, 7 B1 done, 8 dH1 ready, 9-12 dP1 published), 3 = kernel start / prologue end."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from distributional_rl_decision_and_control_b200 import _lib  # noqa: E402
from test_gpu_critic import _critic, _inputs, _params  # noqa: E402

B, N = int(sys.argv[1]), int(sys.argv[2])
c = _critic()
tau, feats, aenc = _inputs(B, N, 1)
ps = _params(c)
ws = _lib.critic_workspace("probe", feats.device)
dz = (torch.randn((B, N), device="cuda:0") / (B * N)).contiguous()
_, _, (ci, h1i, h2i, r3i) = _lib.critic_forward(feats, aenc, *ps, N, ws, tau_in=tau, save=True)
fmx = feats.amax().reshape(1)
so = C.CDLL(_lib.SO_PATH)
ev = np.zeros((4, 1024), dtype=np.uint64)
cnt = np.zeros(4, dtype=np.int32)
for rep in range(3):
    _lib.critic_backward(dz, r3i, h2i, h1i, feats, aenc, ps[2], ps[4], ps[6], N, fmax=fmx)
    torch.cuda.synchronize()
    so.asvrl_critic_backward_trace(ev.ctypes.data_as(C.c_void_p), cnt.ctypes.data_as(C.c_void_p))
t0 = int(ev[3, 0]) >> 8
for w in range(4):
    print("writer", w, "n", cnt[w])
    print(" ".join(f"{int(e) & 255}@{(int(e) >> 8) - t0}" for e in ev[w, :min(cnt[w], 120)]))
