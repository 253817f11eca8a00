import torch, sys
dev = "cuda:0"
M = 32768
def t(fn, n=50):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
for tf32 in (False, True):
    torch.backends.cuda.matmul.allow_tf32 = tf32
    for (a, b) in ((128, 128), (128, 256), (256, 64)):
        dP = torch.randn(M, a, device=dev); H = torch.randn(M, b, device=dev)
        base = t(lambda: dP.t() @ H)
        alt = t(lambda: (H.t() @ dP).t())
        line = f"tf32={tf32} {a}x{b}: mm {base:.1f}us  swapped {alt:.1f}us "
        ref = dP.double().t() @ H.double()
        for S in (16, 32, 64, 128, 256):
            f = lambda: torch.bmm(dP.view(S, M // S, a).transpose(1, 2), H.view(S, M // S, b)).sum(0)
            us = t(f)
            err = (f().double() - ref).abs().max().item() / ref.abs().max().item()
            line += f" S{S}:{us:.1f}us({err:.0e})"
        print(line, flush=True)
    # forward / dgrad shapes for context
    X = torch.randn(M, 256, device=dev); W = torch.randn(128, 256, device=dev)
    print(f"tf32={tf32} fwd 256->128: {t(lambda: X @ W.t()):.1f}us  dgrad 128->256: {t(lambda: (X @ W.t()) @ W):.1f}us (both)")
