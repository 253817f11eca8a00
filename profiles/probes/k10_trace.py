"""Timeline of CTA 0 of K10 (debug build: ASVRL_NVCC_FLAGS=-DCF_TRACE python -m ...csrc.build --force).
Prints, per writer (0 = MMA thread, 1 / 2 = slot 0 / 1 epilogue, 3 / 4 = the W producers), the events in ns."""
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
so = C.CDLL(_lib.SO_PATH)
ev = np.zeros((8, 2048), dtype=np.uint64)
cnt = np.zeros(8, dtype=np.int32)
for rep in range(3):
    _lib.critic_forward(feats, aenc, *ps, N, ws, tau_in=tau)
    torch.cuda.synchronize()
    so.asvrl_critic_forward_trace(ev.ctypes.data_as(C.c_void_p), cnt.ctypes.data_as(C.c_void_p))
t0 = min(int(ev[w, 0]) >> 8 for w in range(5) if cnt[w] > 0)
for w in range(5):
    print("writer", w, "n", cnt[w])
    print(" ".join(f"{int(e) & 255}@{(int(e) >> 8) - t0}" for e in ev[w, :min(cnt[w], 200)]))
