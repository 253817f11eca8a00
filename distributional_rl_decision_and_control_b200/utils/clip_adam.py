"""ClipAdam: `clip_grad_norm_(params, max_norm)` + `Adam.step()` of the reference's updates
(rfarl/rfarl/agent.py:414-416, 424-427 and the other agents' twins) as kernel K7 — two launches
per optimiser step instead of torch's ~45 (foreach norm / stack / clamp / mul, then the capturable
foreach Adam whose scalar-tensor pow falls back to one launch per parameter).

Same arithmetic and hyper-parameter defaults as torch.optim.Adam(lr, betas=(0.9, 0.999), eps=1e-8,
weight_decay=0, amsgrad=False); the step counts live in device floats (one per tensor, like torch), so the update is CUDA-graph
capturable.  With a process group the gradients are summed over the ranks in ONE flat bucket
(NCCL all-reduce) and the average is folded into the kernel's gradient scale — every rank clips
the same global gradient and the replicas stay bit-identical.  CUDA only.
"""
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
        # torch.optim-shaped view for code that inspects the learning rate
        self.param_groups = [{"params": self.params, "lr": self.lr, "betas": self.betas, "eps": self.eps}]

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
