"""Host-facing, pipelined stepping of the batched environment (send / recv, like EnvPool's async mode).

A CPU-side caller of the reference's ``step()`` hands in actions and reads back observations, rewards
and done flags; with the environments on the GPU that is one small H2D copy, the fused step launch
and one large D2H copy per call, and the D2H copy (34 words per ASV) is what bounds the rate.  The
environments are therefore split into ``groups`` independent shards, each with its own stream and
pinned buffers: while the caller works on the results of one shard, the copy of the next one and the
kernel of the one after are in flight.  Every shard is stepped synchronously with respect to itself
(``recv(g)`` returns the results of the actions passed to the last ``send(g, ...)``), so the semantics
per environment are those of ``BatchedMarineNavEnv.step``.
"""
import torch

from ... import _lib
from .batched import BatchedMarineNavEnv


class HostPipelinedEnv:
    def __init__(self, n_envs, num_robots=8, num_obs=10, num_cores=0, groups=2, device="cuda:0", seed=0,
                 env_offset=0, with_mask=False, **kw):
        if n_envs % groups:
            raise ValueError("n_envs must be a multiple of groups")
        self.device = torch.device(device)
        self.groups = int(groups)
        self.E = n_envs // groups
        self.with_mask = bool(with_mask)
        self.shards, self.streams, self.host_out, self._act_dev, self._done = [], [], [], [], []
        for g in range(groups):
            st = torch.cuda.Stream(device=self.device)
            with torch.cuda.stream(st):
                env = BatchedMarineNavEnv(self.E, num_robots, num_obs, num_cores, device=self.device, seed=seed,
                                          env_offset=env_offset + g * self.E, **kw)
                env.reset()
            self.shards.append(env)
            self.streams.append(st)
            self.host_out.append(_lib.new_out(self.E, env.RP, None, pin=True))
            self._act_dev.append(None)
            self._done.append(torch.cuda.Event())
        self.RP = self.shards[0].RP
        self.words = (sum(w for _, w in _lib.HOST_BLOCK_FIELDS) if with_mask else _lib.HOST_BLOCK_NO_MASK)
        self.d2h_bytes = self.E * self.RP * self.words * 4
        torch.cuda.synchronize(self.device)

    def send(self, g, actions_pinned, is_continuous_action=True):
        """Enqueue on shard g's stream: actions H2D, the fused step, results D2H.  Returns at once."""
        env, st = self.shards[g], self.streams[g]
        with torch.cuda.stream(st):
            if self._act_dev[g] is None:
                self._act_dev[g] = torch.empty(actions_pinned.shape, dtype=actions_pinned.dtype, device=self.device)
            self._act_dev[g].copy_(actions_pinned, non_blocking=True)
            out = env.step(self._act_dev[g], is_continuous_action)
            n = self.E * self.RP * self.words
            self.host_out[g]["host_block"][:n].copy_(out["host_block"][:n], non_blocking=True)
            self._done[g].record(st)

    def recv(self, g):
        """Block until shard g's last send has landed in pinned memory; returns the host views
        (obs_self, obs_objects, reward, status[, obs_mask])."""
        self._done[g].synchronize()
        return self.host_out[g]
