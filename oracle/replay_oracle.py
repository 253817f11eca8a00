"""numpy restatement of the replay path.  TEST INFRASTRUCTURE ONLY (see asv_oracle.h).

Reference (paths under rfarl/rfarl/): utils/replay_buffer.py:22-24 `add`, :26-69 `sample` +
`state_batch` (pad to 5 objects, float mask), agent.py:360-368 `state_to_tensor`,
policy/trainer.py:157-166 (only robots that were active when the step started are stored).
Parity pin: tests/test_oracle_golden.py checks Ring + gather against the reference's own
ReplayBuffer (deque content, state_batch padding and masks) stored in tests/golden/replay.npz.
"""
import numpy as np

from distributional_rl_decision_and_control_b200 import _abi

ROW = _abi.ROW


def valid_rows(status):
    """TRAINER:157-159: skip robots deactivated before the step; padding slots never exist."""
    status = np.asarray(status).astype(np.uint32).reshape(-1)
    info = (status >> _abi.S_INFO_SHIFT) & 15
    return ((status & _abi.S_WAS_DEACTIVATED) == 0) & (info != _abi.INFO_ABSENT)


def make_rows(obs_self, obs_obj, obs_mask, actions, discrete, reward, nself, nobj, nmask, status):
    """Dense float32[70] rows (layout of include/asvrl.h ASVRL_ROW_*) for every ASV."""
    n = np.asarray(reward).size
    rows = np.zeros((n, ROW), np.float32)
    rows[:, _abi.ROW_SELF:_abi.ROW_OBJ] = np.asarray(obs_self, np.float32).reshape(n, 7)
    rows[:, _abi.ROW_OBJ:_abi.ROW_NOBJ] = np.asarray(obs_obj, np.float32).reshape(n, 25)
    rows[:, _abi.ROW_NOBJ] = np.asarray(obs_mask, np.float32).reshape(n, 5).sum(1)
    if discrete:
        rows[:, _abi.ROW_ACT] = np.asarray(actions).reshape(n).astype(np.float32)
    else:
        rows[:, _abi.ROW_ACT:_abi.ROW_REW] = np.asarray(actions, np.float32).reshape(n, 2)
    rows[:, _abi.ROW_REW] = np.asarray(reward, np.float32).reshape(n)
    rows[:, _abi.ROW_NSELF:_abi.ROW_NOBJ_F] = np.asarray(nself, np.float32).reshape(n, 7)
    rows[:, _abi.ROW_NOBJ_F:_abi.ROW_NNOBJ] = np.asarray(nobj, np.float32).reshape(n, 25)
    rows[:, _abi.ROW_NNOBJ] = np.asarray(nmask, np.float32).reshape(n, 5).sum(1)
    st = np.asarray(status).astype(np.uint32).reshape(n)
    rows[:, _abi.ROW_DONE] = ((st & _abi.S_DONE) != 0).astype(np.float32)
    return rows


class Ring:
    """deque(maxlen=capacity) of rows (RB:14, 22-24) as a ring with explicit counters."""

    def __init__(self, capacity):
        self.capacity = int(capacity)
        self.rows = np.zeros((self.capacity, ROW), np.float32)
        self.total = 0

    @property
    def size(self):
        return min(self.total, self.capacity)

    def push(self, rows, valid):
        for r in rows[valid]:
            self.rows[self.total % self.capacity] = r
            self.total += 1

    def gather(self, indices):
        """RB:30-69 + AGENT:387-392 for the given row indices."""
        idx = np.asarray(indices, np.int64) % max(self.size, 1)
        r = self.rows[idx]
        B = len(idx)
        t = np.arange(5, dtype=np.float32)[None, :]
        return {
            "self": r[:, _abi.ROW_SELF:_abi.ROW_OBJ].copy(),
            "obj": r[:, _abi.ROW_OBJ:_abi.ROW_NOBJ].reshape(B, 5, 5).copy(),
            "mask": (t < r[:, _abi.ROW_NOBJ:_abi.ROW_NOBJ + 1]).astype(np.float32),   # RB:65
            "act": r[:, _abi.ROW_ACT:_abi.ROW_REW].copy(),
            "rew": r[:, _abi.ROW_REW].copy(),
            "nself": r[:, _abi.ROW_NSELF:_abi.ROW_NOBJ_F].copy(),
            "nobj": r[:, _abi.ROW_NOBJ_F:_abi.ROW_NNOBJ].reshape(B, 5, 5).copy(),
            "nmask": (t < r[:, _abi.ROW_NNOBJ:_abi.ROW_NNOBJ + 1]).astype(np.float32),
            "done": r[:, _abi.ROW_DONE].copy(),
        }


def state_batch_rows(states):
    """RB:41-69 `state_batch` restated: list of (self[7], objects[<=5][5]) -> padded arrays."""
    B = len(states)
    self_b = np.zeros((B, 7), np.float32)
    obj_b = np.zeros((B, 5, 5), np.float32)
    mask_b = np.zeros((B, 5), np.float32)
    for b, (s, objs) in enumerate(states):
        self_b[b] = s
        for t, o in enumerate(objs[:5]):
            obj_b[b, t] = o
            mask_b[b, t] = 1.0
    return self_b, obj_b, mask_b
