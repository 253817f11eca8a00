import torch, time
dev = "cuda:0"
for mb in (2.2, 8, 64):
    n = int(mb * 1e6)
    d = torch.empty(n, dtype=torch.uint8, device=dev); h = torch.empty(n, dtype=torch.uint8).pin_memory()
    for _ in range(3): h.copy_(d, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): h.copy_(d, non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    t = e0.elapsed_time(e1) / 20 * 1e-3
    e0.record()
    for _ in range(20): d.copy_(h, non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    t2 = e0.elapsed_time(e1) / 20 * 1e-3
    print(f"{mb} MB: D2H {n / t / 1e9:.1f} GB/s ({t * 1e6:.0f} us)  H2D {n / t2 / 1e9:.1f} GB/s")
# two streams concurrently D2H + H2D
import subprocess
print(subprocess.run("nvidia-smi --query-gpu=pcie.link.gen.current,pcie.link.width.current,pcie.link.gen.max --format=csv,noheader", shell=True, capture_output=True, text=True).stdout)
