"""Timeline of the first step CTA of K1 (debug build: ASVRL_NVCC_FLAGS=-DK1_TRACE python -m ...csrc.build --force).
SM clocks at the phase boundaries, thread 0 (warp 0, the integration) and thread 32 (warp 1, compaction / noise)."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from distributional_rl_decision_and_control_b200 import _lib  # noqa: E402
from distributional_rl_decision_and_control_b200.envs.marinenav.batched import BatchedMarineNavEnv  # noqa: E402

E = int(sys.argv[1]) if len(sys.argv) > 1 else 592
dev = torch.device("cuda:0")
gen = torch.Generator(device=dev).manual_seed(7)
env = BatchedMarineNavEnv(E, 8, 10, 8, device=dev, seed=11, fused=True, prefetch="stream")
env.reset()
pool = [torch.rand((E, 8, 2), device=dev, generator=gen) * 2 - 1 for _ in range(8)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
so = C.CDLL(_lib.SO_PATH)
names = {0: "entry", 1: "phase-0 loads consumed", 2: "after first barrier (w0)", 3: "after first barrier (w1)",
         4: "w1: pair compaction done", 5: "w0: integration loop starts", 6: "w0: integration loop ends",
         7: "w0: past bar 1", 8: "w0: noise tickets done", 9: "w1: noise tickets done", 10: "phase B starts",
         11: "phase B2 starts", 12: "phase C1 starts", 13: "phase C2 starts", 14: "phase D starts", 15: "phase D done"}
for cold in (True, False):
    acc = []
    for rep in range(8):
        if cold:
            flush.zero_()
        env.step_fused(pool[rep % 8])
        torch.cuda.synchronize()
        buf = np.zeros(32, dtype=np.uint64)
        so.asvrl_env_step_trace(buf.ctypes.data_as(C.c_void_p))
        acc.append(buf.astype(np.int64))
    t = np.median(np.stack(acc[2:]), axis=0)
    print(f"--- E={E} {'cold' if cold else 'warm'} L2 (cycles from entry; / 1965 = us)")
    for i in range(16):
        print(f"  {i:2d} {names[i]:32s} {int(t[i] - t[0]):7d}  {float(t[i] - t[0]) / 1965:6.2f} us")
