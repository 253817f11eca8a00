"""Host-facing, pipelined stepping of the batched environment (send / recv, like EnvPool's async mode).

A CPU-side caller of the reference's ``step()`` hands in actions and reads back observations, rewards
and done flags; with the environments on the GPU that is one small H2D copy, the fused step launch
and one large D2H copy per call (34 words per ASV).  The
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
        self._graphs = None      # per shard: {phase: CUDAGraph} once enable_cuda_graph() has run
        self._act_host = [None] * groups
        self.words = (sum(w for _, w in _lib.HOST_BLOCK_FIELDS) if with_mask else _lib.HOST_BLOCK_NO_MASK)
        self.d2h_bytes = self.E * self.RP * self.words * 4
        torch.cuda.synchronize(self.device)

    def enable_cuda_graph(self, is_continuous_action=True, zero_copy=False):
        """Capture, per shard and ping-pong phase, ONE graph of [actions H2D from the shard's pinned
        staging buffer, the fused step, results D2H into the shard's pinned block].  A send() is then
        a host memcpy into the staging buffer (none when the caller filled ``action_buffer(g)`` in
        place) and one graph launch: the per-step host work drops from four torch / ctypes calls to
        one, which is what bounded the rate (the D2H copy itself runs at ~52 GB/s on this box).
        RNG counters advance through the env's device word, as in BatchedMarineNavEnv.enable_cuda_graph."""
        dev = self.device
        self._graphs = []
        self.zero_copy = bool(zero_copy)
        for g, (env, st) in enumerate(zip(self.shards, self.streams)):
            if zero_copy:
                # the step kernel stores observations / rewards / status straight into the shard's pinned
                # host block (device-accessible under unified addressing): the D2H transfer rides on the
                # kernel's own coalesced row stores and the separate copy — a third of a shard's
                # H2D -> kernel -> D2H chain — disappears.  The mask stays on the device.
                ho = self.host_out[g]
                for op in env._o_ptrs:
                    op.obs_self, op.obs_objects = ho["obs_self"].data_ptr(), ho["obs_objects"].data_ptr()
                    op.reward, op.status = ho["reward"].data_ptr(), ho["status"].data_ptr()
            if env.prefetch == "stream":
                env.prefetch = "workers"  # a captured step refills its stage inside the launch
            torch.cuda.synchronize(dev)
            shape, dt = ((self.E, self.RP, 2), torch.float32) if is_continuous_action else ((self.E, self.RP), torch.int32)
            self._act_host[g] = torch.zeros(shape, dtype=dt).pin_memory()
            n = self.E * self.RP * self.words
            graphs = {}
            with torch.cuda.device(dev), torch.cuda.stream(st):
                env.counter_dev = torch.zeros((1,), dtype=torch.int64, device=dev)
                self._act_dev[g] = torch.zeros(shape, dtype=dt, device=dev)
                for _ in range(2):  # warm-up on the capture stream
                    env.step(self._act_dev[g], is_continuous_action)
                st.synchronize()
                for i in range(2):
                    phase = env._cur
                    gr = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(gr, stream=st):
                        self._act_dev[g].copy_(self._act_host[g], non_blocking=True)
                        out = env.step(self._act_dev[g], is_continuous_action)  # flips env._cur
                        if not zero_copy:
                            self.host_out[g]["host_block"][:n].copy_(out["host_block"][:n], non_blocking=True)
                        if i == 1:
                            env.counter_dev.add_(2)
                    graphs[phase] = gr
            self._graphs.append(graphs)
        torch.cuda.synchronize(dev)

    def action_buffer(self, g):
        """Pinned staging buffer of shard g (after enable_cuda_graph): write actions here and pass it
        to send() to skip the host copy."""
        return self._act_host[g]

    def send(self, g, actions_pinned, is_continuous_action=True):
        """Enqueue on shard g's stream: actions H2D, the fused step, results D2H.  Returns at once."""
        env, st = self.shards[g], self.streams[g]
        if self._graphs is not None:
            buf = self._act_host[g]
            if actions_pinned is not buf and actions_pinned.data_ptr() != buf.data_ptr():
                buf.copy_(actions_pinned)
            home = torch.cuda.current_stream(self.device)
            torch.cuda.set_stream(st)   # (the stream context manager costs several microseconds per call)
            self._graphs[g][env._cur].replay()
            self._done[g].record(st)
            torch.cuda.set_stream(home)
            env._cur = 1 - env._cur
            env.launches += 1 if env.fused else 3
            return
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
