"""ClipAdam: `clip_grad_norm_(params, max_norm)` + `Adam.step()` of the reference's updates
(rfarl/rfarl/agent.py:414-416, 424-427 and the other agents' twins) as kernel K7 — two launches
per optimiser step instead of torch's ~45 (foreach norm / stack / clamp / mul, then the capturable
foreach Adam whose scalar-tensor pow falls back to one launch per parameter).

Same arithmetic and hyper-parameter defaults as torch.optim.Adam(lr, betas=(0.9, 0.999), eps=1e-8,
weight_decay=0, amsgrad=False); the step counts live in device floats (one per tensor, like torch), so the update is CUDA-graph
capturable.  With a process group the gradients of all ranks are summed in ONE flat bucket and the
average is folded into the kernel's gradient scale — every rank clips the same global gradient and
the replicas stay bit-identical.  The sum is computed by K7 itself over NVLink peer
memory (a one-shot all-reduce fused into the norm kernel: pack + flag, wait + sum + norm, Adam)
when the ranks can map each other's buckets, else by one NCCL all-reduce.  CUDA only.
"""
import os
import warnings

import torch

from .. import _abi, _lib


class ClipAdam:
    def __init__(self, params, lr=1e-3, max_norm=0.5, betas=(0.9, 0.999), eps=1e-8, process_group=None):
        self.params = [p for p in params]
        if not self.params:
            raise ValueError("ClipAdam got an empty parameter list")
        dev = self.params[0].device
        if dev.type != "cuda":
            raise RuntimeError("ClipAdam runs on CUDA parameters only (no CPU fallback)")
        if any(p.dtype != torch.float32 or not p.is_contiguous() for p in self.params):
            raise RuntimeError("ClipAdam needs contiguous float32 parameters")
        self.lr, self.max_norm, self.betas, self.eps = float(lr), float(max_norm), tuple(betas), float(eps)
        self.process_group = process_group
        sizes = [p.numel() for p in self.params]
        self._state = torch.zeros((2, sum(sizes)), dtype=torch.float32, device=dev)
        self.exp_avg, self.exp_avg_sq = [], []
        off = 0
        for n in sizes:
            self.exp_avg.append(self._state[0, off:off + n])
            self.exp_avg_sq.append(self._state[1, off:off + n])
            off += n
        self.step_count = torch.zeros((len(sizes),), dtype=torch.float32, device=dev)  # t per tensor, on the device
        self.grad_norm = torch.zeros((1,), dtype=torch.float32, device=dev)   # total norm of the last step
        self._scratch = torch.zeros((_abi.OPT_SCRATCH_FLOATS,), dtype=torch.float32, device=dev)
        self._counter = torch.zeros((1,), dtype=torch.int32, device=dev)
        # N > 1: the exchange step runs inside K7 over NVLink peer memory when the ranks can map each
        # other's buckets (torch symmetric memory); otherwise one NCCL all-reduce of the flat bucket
        self._peers = None
        if process_group is not None and os.environ.get("ASVRL_P2P_ALLREDUCE", "1") != "0":
            try:
                self._setup_p2p(sum(sizes), dev)
            except Exception as e:  # no peer access / no symmetric memory on this stack: NCCL path
                warnings.warn(f"ClipAdam: peer-memory all-reduce unavailable ({type(e).__name__}: {e}); using NCCL")
                self._peers = None
        # torch.optim-shaped view for code that inspects the learning rate
        self.param_groups = [{"params": self.params, "lr": self.lr, "betas": self.betas, "eps": self.eps}]

    def _setup_p2p(self, total, dev):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        world = dist.get_world_size(self.process_group)
        if world > _abi.MAX_PEERS:
            raise RuntimeError(f"more than {_abi.MAX_PEERS} ranks")
        half = (total + 63) // 64 * 64
        bucket = symm_mem.empty(2 * half + 64, dtype=torch.float32, device=dev)  # [half 0 | half 1 | flags]
        bucket.zero_()
        torch.cuda.synchronize(dev)
        hdl = symm_mem.rendezvous(bucket, group=self.process_group)
        peers = _abi.PeerBuckets()
        peers.world, peers.rank, peers.half_floats = hdl.world_size, hdl.rank, half
        ptrs = list(hdl.buffer_ptrs)
        for r in range(hdl.world_size):
            peers.bucket[r] = int(ptrs[r])
        self._bucket, self._hdl = bucket, hdl          # keep the mapping alive
        self._grad_sum = torch.zeros((total,), dtype=torch.float32, device=dev)
        self._epoch = torch.zeros((1,), dtype=torch.int32, device=dev)
        dist.barrier(group=self.process_group)          # every bucket is zeroed before anyone signals
        torch.cuda.synchronize(dev)
        self._peers = peers

    def check(self):
        """Host-side health check (one device read): no step since construction may have been skipped.  K7
        skips a step (parameters, moments and step counts untouched) on a non-finite gradient norm or — on
        the peer-memory path — when a rank never arrived at the exchange; it counts skipped steps in the last
        scratch float, and a missed exchange raises a sticky fault word on every rank (include/asvrl.h), so
        a fault in the middle of an interval is still seen here."""
        bad = ~torch.isfinite(self.grad_norm).all() | (self._scratch[-1] != 0)
        if self._peers is not None:  # sticky fault word of the peer-memory exchange (a rank timed out)
            bad = bad | (self._bucket[2 * self._peers.half_floats + _abi.MAX_PEERS] != 0)
        if bool(bad):
            n = int(self._scratch[-1].item())
            self._scratch[-1] = 0  # reported once: a caller that catches this and carries on is told about NEW skips only
            self.grad_norm.zero_()  # (the peer-memory fault word stays raised: that exchange cannot be resumed)
            raise RuntimeError("ClipAdam: %d optimiser step(s) were skipped (non-finite gradient norm: diverged "
                               "gradients, or a rank missed the gradient exchange — all ranks must step in lock "
                               "step; after a missed exchange every rank skips every later step)"
                               % n)

    def zero_grad(self, set_to_none=True):
        for p in self.params:
            if set_to_none:
                p.grad = None
            elif p.grad is not None:
                p.grad.zero_()

    def step(self):
        """Clip to ``max_norm`` and apply Adam to every parameter that has a gradient."""
        self.lr = float(self.param_groups[0]["lr"])
        live = [(p, p.grad, m, v, self.step_count[i:i + 1]) for i, (p, m, v) in
                enumerate(zip(self.params, self.exp_avg, self.exp_avg_sq)) if p.grad is not None]
        if not live:
            return
        grads = [g if g.is_contiguous() else g.contiguous() for _, g, _, _, _ in live]
        scale = 1.0
        if self._peers is not None:  # the exchange step happens inside K7 (peer-memory one-shot all-reduce)
            if len(live) != len(self.params):
                raise RuntimeError("ClipAdam (peer-memory path): every parameter needs a gradient on every rank")
            _lib.clip_adam_step([x[0].data for x in live], grads, [x[2] for x in live], [x[3] for x in live],
                                [x[4] for x in live], self.max_norm, self.lr, self.betas[0], self.betas[1],
                                self.eps, 1.0, self.grad_norm, self._scratch, self._counter,
                                peers=self._peers, grad_sum=self._grad_sum, epoch=self._epoch)
            return
        if self.process_group is not None:  # the one exchange step: a flat fp32 bucket, summed
            import torch.distributed as dist
            flat = torch.cat([g.reshape(-1) for g in grads])
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.process_group)
            scale = 1.0 / dist.get_world_size(self.process_group)
            out, off = [], 0
            for g in grads:
                out.append(flat[off:off + g.numel()])
                off += g.numel()
            grads = out
        M = _abi.OPT_MAX_TENSORS
        if len(live) > M:
            raise RuntimeError(f"ClipAdam: more than {M} parameter tensors in one optimiser")
        _lib.clip_adam_step([x[0].data for x in live], grads, [x[2] for x in live], [x[3] for x in live],
                            [x[4] for x in live], self.max_norm, self.lr, self.betas[0], self.betas[1],
                            self.eps, scale, self.grad_norm, self._scratch, self._counter)
