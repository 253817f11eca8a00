import os, sys, time
sys.path.insert(0, os.getcwd())
import torch, torch.distributed as dist
from distributional_rl_decision_and_control_b200.utils.clip_adam import ClipAdam
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
pg = dist.group.WORLD
shapes = [(56, 7), (56,), (40, 5), (40,), (256, 64), (256,), (128, 2), (128,), (128, 256), (128,), (128, 128), (128,), (1, 128), (1,)]
def make(p2p):
    os.environ["ASVRL_P2P_ALLREDUCE"] = "1" if p2p else "0"
    g = torch.Generator(device=dev).manual_seed(1)
    ps = [torch.nn.Parameter(torch.randn(s, device=dev, generator=g)) for s in shapes]
    return ps, ClipAdam(ps, lr=1e-4, max_norm=0.5, process_group=pg)
pa, oa = make(True)
pb, ob = make(False)
print(rank, "p2p active:", oa._peers is not None, flush=True)
g = torch.Generator(device=dev).manual_seed(100 + rank)
for it in range(6):
    sc = 0.002 if it % 2 == 0 else 1.0
    for p, q in zip(pa, pb):
        p.grad = torch.randn(p.shape, device=dev, generator=g) * sc
        q.grad = p.grad.clone()
    oa.step(); ob.step()
    torch.cuda.synchronize()
    err = max((p.detach() - q.detach()).abs().max().item() for p, q in zip(pa, pb))
    print(rank, it, "norm", float(oa.grad_norm), float(ob.grad_norm), "max param diff vs NCCL", err, flush=True)
flat = torch.cat([p.detach().reshape(-1) for p in pa])
gathered = [torch.empty_like(flat) for _ in range(world)]
dist.all_gather(gathered, flat)
print(rank, "replicas bit-identical:", all(torch.equal(gathered[0], x) for x in gathered), flush=True)
def timeit(opt, ps, n=300):
    for p in ps: p.grad = torch.randn(p.shape, device=dev, generator=g) * 0.01
    for _ in range(5): opt.step()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): opt.step()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
ta, tb = timeit(oa, pa), timeit(ob, pb)
print(rank, f"eager us/step: p2p {ta:.1f}  nccl {tb:.1f}", flush=True)
# graphs
def graph_time(opt, ps, n=300):
    side = torch.cuda.Stream(); side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3): opt.step()
    torch.cuda.current_stream().wait_stream(side); torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        opt.step()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): gr.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
ga, gb = graph_time(oa, pa), graph_time(ob, pb)
print(rank, f"graph us/step: p2p {ga:.1f}  nccl {gb:.1f}", flush=True)
dist.barrier(); torch.cuda.synchronize()
os._exit(0)
