"""GPU-resident uniform replay buffer behind the reference's ``ReplayBuffer`` interface
(rfarl/rfarl/utils/replay_buffer.py:6-69): ``add(item)``, ``sample()``, ``size()``,
``state_batch(states)``.

Storage is the device ring of dense float32[70] rows of include/asvrl.h; sampling and the
pad-to-5 / mask construction run in kernel K2b (``asvrl_replay_sample``), whole batched env
steps are appended by kernel K2a (``asvrl_replay_push``).  ``add`` of single Python tuples
(the reference Trainer's path, TRAINER:164-166) is staged in pinned host memory and moved into
the ring with one copy before the next sample.  No CPU fallback: a CUDA device is required.
"""
import random

import numpy as np
import torch

from .. import _abi, _lib


def state_batch(states, obj_len=_abi.OBJ_DIM, max_obj_num=_abi.MAX_OBJ):
    """RB:51-69: list of (self, objects) -> (self batch, padded objects, masks); objects and
    masks are empty lists when no state has any object."""
    self_b = [s[0] for s in states]
    obj_lists = [s[1] for s in states]
    if max(len(o) for o in obj_lists) == 0:
        return (self_b, [], [])
    pad = [0.] * obj_len
    objs = [list(o) + [pad] * (max_obj_num - len(o)) for o in obj_lists]
    masks = [[1.] * len(o) + [0.] * (max_obj_num - len(o)) for o in obj_lists]
    return (self_b, objs, masks)


class ReplayBuffer:
    """Fixed-size buffer of (state, action, reward, next_state, done) transitions."""

    STAGE = 4096  # host-staged rows before a forced flush

    def __init__(self, buffer_size, batch_size, obj_len, max_object_num, seed=249, device="cuda:0",
                 max_push=1 << 20):
        if obj_len != _abi.OBJ_DIM or max_object_num != _abi.MAX_OBJ:
            raise RuntimeError("the device row layout is fixed at 5 objects x 5 features (include/asvrl.h)")
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("ReplayBuffer needs a CUDA device (no CPU fallback)")
        _lib.lib()
        self.buffer_size = int(buffer_size)
        self.batch_size = int(batch_size)
        self.obj_len = obj_len
        self.max_obj_num = max_object_num
        self.seed = random.seed(seed)  # RB:20 (kept: the Agent's epsilon draws share this stream)
        self.rng_seed = int(seed)
        self.draws = 0
        self.counter_dev = None  # optional int64 device word added to the draw counter (CUDA graphs)
        with torch.cuda.device(self.device):
            self.ring = _lib.new_replay(self.buffer_size, min(int(max_push), self.buffer_size), self.device)
            self._stage = torch.zeros((self.STAGE, _abi.ROW), dtype=torch.float32).pin_memory()
            self._batch = _lib.new_batch(self.batch_size, self.device)
        self._stage_np = self._stage.numpy()
        self._staged = 0

    # ------------------------------------------------------------------ reference interface
    def add(self, item):
        """Add one experience tuple (RB:22-24).  state = (self[7], objects[<=5][5])."""
        state, action, reward, next_state, done = item
        row = self._stage_np[self._staged]
        row[:] = 0.0
        self._pack_state(row, _abi.ROW_SELF, _abi.ROW_OBJ, _abi.ROW_NOBJ, state)
        if np.ndim(action) == 0:
            row[_abi.ROW_ACT] = float(action)
        else:
            row[_abi.ROW_ACT:_abi.ROW_ACT + len(action)] = action
        row[_abi.ROW_REW] = reward
        self._pack_state(row, _abi.ROW_NSELF, _abi.ROW_NOBJ_F, _abi.ROW_NNOBJ, next_state)
        row[_abi.ROW_DONE] = 1.0 if done else 0.0
        self._staged += 1
        if self._staged == self.STAGE:
            self.flush()

    def _pack_state(self, row, o_self, o_obj, o_n, state):
        s, objs = state
        row[o_self:o_self + _abi.SELF_DIM] = s
        n = min(len(objs), self.max_obj_num)
        for t in range(n):
            row[o_obj + t * self.obj_len:o_obj + (t + 1) * self.obj_len] = objs[t]
        row[o_n] = n

    def flush(self):
        """Move host-staged rows into the device ring (ring order = insertion order, RB:14)."""
        n = self._staged
        if n == 0:
            return
        cap = self.buffer_size
        with torch.cuda.device(self.device):
            hs = self.ring["head_size"]
            total = int(hs[0].item())
            pos = total % cap
            src = self._stage[:n].to(self.device, non_blocking=True)
            skip = max(0, n - cap)  # more staged rows than the ring holds: keep the newest `cap`, as deque(maxlen) does
            pos = (pos + skip) % cap
            keep = n - skip
            first = min(keep, cap - pos)
            self.ring["rows"][pos:pos + first].copy_(src[skip:skip + first])
            if first < keep:
                self.ring["rows"][0:keep - first].copy_(src[skip + first:n])
            hs[0] = total + n
            hs[1] = min(total + n, cap)
            torch.cuda.current_stream().synchronize()  # the pinned stage is reused right away
        self._staged = 0

    def size(self):
        """Current number of stored transitions (RB:47-49)."""
        self.flush()
        return int(self.ring["head_size"][1].item())

    def sample(self):
        """Reference-shaped sample (RB:26-45): python lists on the host.  The learner itself
        uses ``sample_device`` and never crosses PCIe."""
        b = self.sample_device()
        host = {k: v.cpu().numpy() for k, v in b.items()}
        states = (host["self"].tolist(), host["obj"].tolist(), host["mask"].tolist())
        next_states = (host["nself"].tolist(), host["nobj"].tolist(), host["nmask"].tolist())
        return states, host["act"].tolist(), host["rew"].tolist(), next_states, (host["done"] > 0.5).tolist()

    def state_batch(self, states):
        return state_batch(states, self.obj_len, self.max_obj_num)

    # ------------------------------------------------------------------ device-side interface
    def sample_device(self, batch=None, indices=None):
        """K2b: ``batch_size`` distinct uniform rows gathered into dense CUDA tensors
        {self, obj, mask, act, rew, nself, nobj, nmask, done, indices}."""
        self.flush()
        batch = self._batch if batch is None else batch
        if self.counter_dev is None and indices is None:
            # eager path (one host read, like size()): random.sample raises when the population is smaller than
            # the sample (RB:28); under a captured graph the caller checks readiness before capturing
            have = int(self.ring["head_size"][1].item())
            want = batch if isinstance(batch, int) else int(batch["self"].shape[0])  # a dict of preallocated outputs
            if want > have:
                raise ValueError(f"Sample larger than population: batch {want} > {have} stored transitions")
        self.draws += 1
        with torch.cuda.device(self.device):
            return _lib.replay_sample(self.ring, batch, seed=self.rng_seed, counter=self.draws,
                                      indices=indices, counter_dev=self.counter_dev)

    def push_step(self, obs, actions, discrete, out, term=None):
        """K2a: append the transitions of one batched env step.  ``obs`` = (self, objects, mask)
        the policy acted on, ``out`` = the step's output dict (next observation, reward, status).
        ``term``: the terminal-observation buffers of a fused auto-reset step (``env.term``)."""
        self.flush()
        with torch.cuda.device(self.device):
            _lib.replay_push(self.ring, obs[0], obs[1], obs[2], actions, discrete, out["reward"],
                             out["obs_self"], out["obs_objects"], out["obs_mask"], out["status"],
                             term=term, ended=out["ended"] if term is not None else None)
