"""K7's peer-memory all-reduce (asvrl_clip_adam_step_p2p) under torchrun with 2 ranks on one box:
same parameters as the NCCL all-reduce path, replicas bit-identical, capturable in a CUDA graph.
Skipped on a single-GPU box (the world_size-2 gloo test of the exchange step runs on CPU)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_SCRIPT = r'''
import os, sys
sys.path.insert(0, {root!r})
import torch, torch.distributed as dist
from distributional_rl_decision_and_control_b200.utils.clip_adam import ClipAdam
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dev = torch.device("cuda", rank)
dist.init_process_group("nccl", init_method="tcp://127.0.0.1:{port}", rank=rank, world_size=world, device_id=dev)
shapes = [(56, 7), (56,), (256, 64), (128, 256), (1, 128), (1,), (3, 5, 7)]
def make(p2p):
    os.environ["ASVRL_P2P_ALLREDUCE"] = "1" if p2p else "0"
    g = torch.Generator(device=dev).manual_seed(1)
    ps = [torch.nn.Parameter(torch.randn(s, device=dev, generator=g)) for s in shapes]
    return ps, ClipAdam(ps, lr=1e-4, max_norm=0.5, process_group=dist.group.WORLD)
pa, oa = make(True)
pb, ob = make(False)
assert oa._peers is not None, "peer-memory path did not come up"
assert ob._peers is None
g = torch.Generator(device=dev).manual_seed(100 + rank)
def grads(scale):
    for p, q in zip(pa, pb):
        p.grad = torch.randn(p.shape, device=dev, generator=g) * scale
        q.grad = p.grad.clone()
for it in range(5):
    grads(0.002 if it % 2 == 0 else 1.0)
    oa.step(); ob.step()
    torch.cuda.synchronize()
    assert abs(float(oa.grad_norm) - float(ob.grad_norm)) <= 1e-6 * float(ob.grad_norm)
    for p, q in zip(pa, pb):
        assert (p.detach() - q.detach()).abs().max().item() <= 2e-7 * max(1.0, q.abs().max().item())
# captured: the epoch word advances on the device, so replays keep the ranks in lock step
grads(0.01)
side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side):
    oa.step()
torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize()
gr = torch.cuda.CUDAGraph()
with torch.cuda.graph(gr):
    oa.step()
for _ in range(7):
    gr.replay()
torch.cuda.synchronize()
assert torch.isfinite(oa.grad_norm).all()
flat = torch.cat([p.detach().reshape(-1) for p in pa])
got = [torch.empty_like(flat) for _ in range(world)]
dist.all_gather(got, flat)
assert all(torch.equal(got[0], x) for x in got), "replicas diverged"
dist.barrier(); torch.cuda.synchronize()
print("rank", rank, "ok", flush=True)
os._exit(0)
'''


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs on one box")
def test_clip_adam_peer_memory_allreduce_two_ranks(tmp_path):
    port = 29600 + os.getpid() % 2000
    script = tmp_path / "p2p.py"
    script.write_text(_SCRIPT.format(root=ROOT, port=port))
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(os.environ, RANK=str(r), WORLD_SIZE="2"),
                              stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for r in range(2)]
    for p in procs:
        try:
            out, err = p.communicate(timeout=180)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        assert p.returncode == 0 and "ok" in out, err[-2000:]
