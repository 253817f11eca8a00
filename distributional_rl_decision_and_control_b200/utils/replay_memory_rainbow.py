"""GPU-resident prioritised n-step replay for Rainbow behind the reference's ``ReplayMemory``
interface (rfarl/rfarl/policy/replay_memory_rainbow.py:99-199): ``append``, ``sample``,
``update_priorities``, ``transitions.num_elements()``.

Layout: a ring of T time slots x L lanes.  The host path (the reference Trainer, one env) uses one
lane, which is exactly the reference's cyclic buffer; the batched path appends one row per ASV per
step (lane = env * R + robot), so the n steps that follow a sampled transition belong to the SAME
vehicle.  Priorities live in a flat device vector; a batch is drawn by stratified sampling over
their running sum (``cumsum`` + ``searchsorted``: the sum-tree walk of REPLAY:14-93 as two dense
passes), the n-step window is gathered with the reference's blanking rule (a transition whose
episode timestep is 0 cuts the window, REPLAY:131-141).

The ring position, the per-lane episode clocks, the element count and the running maximum priority are
device words updated IN PLACE, so append_step / sample / update_priorities replay inside CUDA graphs
(``head`` / ``filled`` are host mirrors, exact outside graph replay)."""
import numpy as np
import torch

W_SELF, W_OBJ = 7, 25
C_NOBJ, C_ACT, C_REW, C_NONTERM = 32, 33, 34, 35
WIDTH = 36


class _Transitions:
    """What the reference exposes as ``memory.transitions`` (a SegmentTree): only the element
    count is read from outside (TRAINER:184)."""

    def __init__(self, owner):
        self._o = owner

    def num_elements(self):
        return int(self._o.count.item()) if self._o.lanes > 1 else self._o._host_count

    @property
    def max(self):
        return float(self._o.max_priority.item())


class ReplayMemory:
    def __init__(self, device, capacity, lanes=1, n=3):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("ReplayMemory needs a CUDA device (no CPU fallback)")
        self.lanes = int(lanes)
        self.slots = max(int(capacity) // self.lanes, 8)
        self.capacity = self.slots * self.lanes
        self.history = 1
        self.discount = 0.99
        self.n = int(n)
        self.priority_weight = 0.4   # beta, annealed towards 1 by the caller
        self.priority_exponent = 0.5
        dev = self.device
        self.rows = torch.zeros((self.slots, self.lanes, WIDTH), dtype=torch.float32, device=dev)
        self.tstep = torch.zeros((self.slots, self.lanes), dtype=torch.int32, device=dev)
        self.prio = torch.zeros((self.slots * self.lanes,), dtype=torch.float32, device=dev)
        self.lane_t = torch.zeros((self.lanes,), dtype=torch.int32, device=dev)  # REPLAY:109 `self.t`
        self.max_priority = torch.ones((), dtype=torch.float32, device=dev)       # SegmentTree.max
        self.count = torch.zeros((), dtype=torch.int64, device=dev)
        self.head = 0          # next time slot to write (host mirror of head_dev)
        self.filled = 0        # time slots written so far (<= slots; host mirror)
        self.head_dev = torch.zeros((1,), dtype=torch.int64, device=dev)
        self.filled_dev = torch.zeros((1,), dtype=torch.int64, device=dev)
        self._host_count = 0
        self.n_step_scaling = torch.tensor([self.discount ** i for i in range(self.n)],
                                           dtype=torch.float32, device=dev)
        self.transitions = _Transitions(self)
        self.fused_append = True   # append_step through asvrl_rainbow_append (False: the torch statements, its reference)
        self._append_scratch = torch.zeros((2,), dtype=torch.int32, device=dev)
        self._gen = torch.Generator(device=dev)
        self._gen.manual_seed(0x5EED)

    # ------------------------------------------------------------------ reference interface (one lane)
    def state_batch(self, states):
        from .replay_buffer import state_batch
        return state_batch(states)

    def append(self, state, action, reward, terminal):
        """Adds state and action at time t, reward and terminal at time t + 1 (REPLAY:125-131)."""
        if self.lanes != 1:
            raise RuntimeError("append() is the single-lane host path; use append_step() for a batch")
        row = np.zeros((WIDTH,), dtype=np.float32)
        row[:W_SELF] = state[0]
        k = min(len(state[1]), 5)
        for i in range(k):
            row[W_SELF + 5 * i:W_SELF + 5 * (i + 1)] = state[1][i]
        row[C_NOBJ], row[C_ACT], row[C_REW] = k, float(action), float(reward)
        row[C_NONTERM] = 0.0 if terminal else 1.0
        t = self.head
        self.rows[t, 0].copy_(torch.from_numpy(row))
        self.tstep[t, 0] = self.lane_t[0]
        self.prio[t] = self.max_priority
        self.lane_t[0] = 0 if terminal else self.lane_t[0] + 1
        self._advance()
        self._host_count = min(self._host_count + 1, self.capacity)

    def _advance(self):
        self.head = (self.head + 1) % self.slots
        self.filled = min(self.filled + 1, self.slots)
        self.head_dev.add_(1).remainder_(self.slots)
        self.filled_dev.add_(1).clamp_(max=self.slots)

    def slots_filled(self):
        """Time slots written so far, read from the device word (exact under graph replay; one host sync)."""
        return int(self.filled_dev.item())

    # ------------------------------------------------------------------ batched path
    def append_step(self, obs, actions, out):
        """One row per ASV of a batched env step: ``obs`` = (self, objects, mask) acted on,
        ``actions`` int (E, R), ``out`` = the step's output dict.  Lanes whose ASV was already
        deactivated (or is padding) get priority 0 and cut every window that reaches them."""
        L = self.lanes
        o_self, o_obj, o_mask = obs
        if self.fused_append and o_self.dtype == torch.float32 and out["status"].dtype in (torch.int32, torch.uint32):
            # the whole append in one launch (asvrl_rainbow_append); the torch statements below are its specification
            from .. import _lib
            f, i32 = torch.float32, torch.int32
            acts = actions.reshape(L)
            acts = acts if acts.dtype == i32 else acts.to(i32)
            _lib.check(_lib.lib().asvrl_rainbow_append(
                _lib.ptr(o_self.reshape(L, W_SELF).contiguous(), f), _lib.ptr(o_obj.reshape(L, W_OBJ).contiguous(), f),
                _lib.ptr(o_mask.reshape(L, 5).to(f).contiguous(), f), _lib.ptr(acts.contiguous(), i32),
                _lib.ptr(out["reward"].reshape(L).contiguous(), f), out["status"].reshape(L).contiguous().data_ptr(),
                L, self.slots, self.capacity, _lib.ptr(self.rows, f), _lib.ptr(self.tstep, i32), _lib.ptr(self.prio, f),
                _lib.ptr(self.lane_t, i32), _lib.ptr(self.max_priority, f), _lib.ptr(self.head_dev, torch.int64),
                _lib.ptr(self.filled_dev, torch.int64), _lib.ptr(self.count, torch.int64),
                self._append_scratch.data_ptr(), _lib.stream_ptr()), "asvrl_rainbow_append")
            self.head = (self.head + 1) % self.slots
            self.filled = min(self.filled + 1, self.slots)
            return
        status = out["status"].reshape(L)
        info = (status >> 8) & 15
        active = ((status & 8) == 0) & (info != 7)
        done = (status & 1) != 0
        t = self.head_dev                                            # (1,) device index: graph-replay safe
        row = torch.cat((o_self.reshape(L, W_SELF), o_obj.reshape(L, W_OBJ), o_mask.reshape(L, 5).sum(dim=1, keepdim=True),
                         actions.reshape(L, 1).to(torch.float32), out["reward"].reshape(L, 1),
                         (~done).to(torch.float32).unsqueeze(1)), dim=1)
        self.rows.index_copy_(0, t, row.unsqueeze(0))
        zero_t = torch.zeros_like(self.lane_t)
        self.tstep.index_copy_(0, t, torch.where(active, self.lane_t, zero_t).unsqueeze(0))
        self.prio.view(self.slots, L).index_copy_(
            0, t, torch.where(active, self.max_priority, torch.zeros((), device=self.device)).unsqueeze(0))
        self.lane_t.copy_(torch.where(active & ~done, self.lane_t + 1, zero_t))
        self.count.copy_(torch.clamp(self.count + active.sum(), max=self.capacity))
        self._advance()

    # ------------------------------------------------------------------ sampling
    def _gather(self, idxs):
        """The reference's _get_transitions + the tensors built from it (REPLAY:133-179) for flat
        indices ``idxs`` (slot * lanes + lane)."""
        L, T, n = self.lanes, self.slots, self.n
        slot, lane = idxs // L, idxs % L
        offs = torch.arange(0, n + 1, device=self.device)
        slots = (slot.unsqueeze(1) + offs) % T                      # (B, n+1)
        lanes = lane.unsqueeze(1).expand_as(slots)
        rows = self.rows[slots, lanes]                              # (B, n+1, WIDTH)
        first = self.tstep[slots, lanes] == 0
        first[:, 0] = False                                         # only FUTURE episode starts cut
        blank = torch.cumsum(first.to(torch.int32), dim=1) > 0      # REPLAY:139-140
        rows = rows * (~blank).unsqueeze(-1).to(rows.dtype)         # blank_trans: zeros, nonterminal False
        def state(k):
            r = rows[:, k]
            nobj = r[:, C_NOBJ]
            mask = (torch.arange(5, device=self.device).unsqueeze(0) < nobj.unsqueeze(1)).to(torch.float32)
            return (r[:, :W_SELF].contiguous(), r[:, W_SELF:W_SELF + W_OBJ].reshape(-1, 5, 5).contiguous(), mask)
        actions = rows[:, 0, C_ACT].to(torch.int64)
        R = rows[:, :n, C_REW] @ self.n_step_scaling                # REPLAY:172-173
        nonterminals = rows[:, n, C_NONTERM].unsqueeze(1)           # REPLAY:175
        return state(0), actions, R, state(n), nonterminals

    def sample(self, batch_size):
        """-> (idxs, states, actions, returns, next_states, nonterminals, weights), REPLAY:181-188.
        The n newest time slots cannot be drawn (their n-step window is not written yet)."""
        L, T = self.lanes, self.slots
        p = self.prio.clone().view(T, L)
        recent = (self.head_dev - 1 - torch.arange(self.n, device=self.device)) % T
        p.index_fill_(0, recent, 0.0)
        p = p.view(-1)
        csum = torch.cumsum(p.to(torch.float64), dim=0)
        total = csum[-1]
        seg = total / batch_size                                     # REPLAY:144-149
        gen = None if torch.cuda.is_current_stream_capturing() else self._gen   # the default generator is graph-safe
        u = (torch.rand((batch_size,), device=self.device, generator=gen, dtype=torch.float64) +
             torch.arange(batch_size, device=self.device, dtype=torch.float64)) * seg
        idxs = torch.searchsorted(csum, u, right=True).clamp_(max=p.numel() - 1)
        probs = p[idxs] / total.to(torch.float32)
        size = torch.clamp(self.count, min=1).to(torch.float32) if L > 1 else \
            torch.tensor(float(max(self._host_count, 1)), device=self.device)
        weights = (size * probs).pow(-self.priority_weight)          # REPLAY:185-187
        weights = weights / weights.max()
        states, actions, R, next_states, nonterminals = self._gather(idxs)
        return idxs, states, actions, R, next_states, nonterminals, weights

    def update_priorities(self, idxs, priorities):
        """priorities ** 0.5 into the sampled entries; tracks the running maximum (REPLAY:190-192)."""
        if not torch.is_tensor(priorities):
            priorities = torch.as_tensor(np.asarray(priorities), dtype=torch.float32, device=self.device)
        if not torch.is_tensor(idxs):
            idxs = torch.as_tensor(np.asarray(idxs), dtype=torch.int64, device=self.device)
        pr = priorities.detach().to(torch.float32).pow(self.priority_exponent)
        self.prio[idxs] = pr
        self.max_priority.copy_(torch.maximum(self.max_priority, pr.max()))
