"""Timeline of CTA 0 of K13 (debug build: ASVRL_NVCC_FLAGS=-DCH_TRACE python -m ...csrc.build --force).
writer 0 = weight producer (tag = stage index within the tile), 1 = MMA issuer (2u: unit u starts, 2u + 1: committed),
2 / 3 = epilogue warps e = 0 / 1 of quadrant 2 (2u: unit u's accumulator seen, 2u + 1: unit done)."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from distributional_rl_decision_and_control_b200 import _lib  # noqa: E402
from test_agents_more import build  # noqa: E402

dev = "cuda:0"
pol = build("Rainbow", 7, dev)
pol.train()
support = torch.linspace(-10.0, 10.0, 51, device=dev)
M = 65536
g = torch.Generator(device=dev).manual_seed(3)
x = (torch.randn((M, 7), device=dev, generator=g) * 5, torch.randn((M, 5, 5), device=dev, generator=g) * 3,
     (torch.rand((M, 5), device=dev, generator=g) < 0.6).float())
so = C.CDLL(_lib.SO_PATH)
ev = np.zeros((4, 512), dtype=np.uint64)
cnt = np.zeros(4, dtype=np.int32)
for rep in range(3):
    pol.greedy_actions(x, support)
    torch.cuda.synchronize()
    so.asvrl_c51_head_trace(ev.ctypes.data_as(C.c_void_p), cnt.ctypes.data_as(C.c_void_p))
t0 = min(int(ev[w, 0] >> 8) for w in range(4) if cnt[w])
names = ["producer", "mma", "epi e=0", "epi e=1"]
for w in range(4):
    n = int(cnt[w])
    seq = [((int(ev[w, i]) >> 8) - t0, int(ev[w, i]) & 255) for i in range(n)]
    print(f"--- {names[w]}: {n} events; first tile (cycles from the first event, tag):")
    print("  " + "  ".join(f"{t}:{tag}" for t, tag in seq[:60 if w else 28]))
    if n:
        print(f"  last event at {seq[-1][0]} cycles = {seq[-1][0] / 1965:.1f} us")

ns = np.zeros((256, 2), dtype=np.uint64)
so.asvrl_c51_head_cta_times(ns.ctypes.data_as(C.c_void_p))
ns = ns[:148].astype(np.int64)
k0 = ns[:, 0].min()
start, end = (ns[:, 0] - k0) / 1e3, (ns[:, 1] - k0) / 1e3
print(f"--- per CTA (us from the first CTA's start): start min {start.min():.1f} max {start.max():.1f}; end min {end.min():.1f} "
      f"median {np.median(end):.1f} max {end.max():.1f}; duration min {(end - start).min():.1f} max {(end - start).max():.1f}")
order = np.argsort(end)
print("  slowest CTAs:", [(int(i), round(float(start[i]), 1), round(float(end[i]), 1)) for i in order[-6:]])
print("  fastest CTAs:", [(int(i), round(float(start[i]), 1), round(float(end[i]), 1)) for i in order[:6]])
