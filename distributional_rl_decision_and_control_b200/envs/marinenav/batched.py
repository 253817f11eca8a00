"""Batched, GPU-resident marine-navigation environment (the vectorised face of K1 / K1b).

Thousands of independent ``MarineNavEnv3`` instances (reference:
rfarl/rfarl/envs/marinenav/env.py:24-333) advanced by ONE kernel launch per step, with the
Trainer's bookkeeping folded in (rfarl/rfarl/policy/trainer.py:157-172, 240-245):
deactivation of ASVs that collided or arrived, end-of-episode detection (all deactivated or
1000 steps) and re-generation of finished scenarios.  State lives in HBM as structure-of-arrays
``[field][env][robot]``; nothing crosses PCIe unless the caller asks for host copies.
"""
import ctypes as C

import torch

from ... import _abi, _lib


class BatchedMarineNavEnv:
    """``n_envs`` environments x ``max_robots`` ASVs on one GPU.

    Observation buffers are ping-ponged: ``obs`` (what the policy acts on, post-reset for
    finished episodes) and ``next_obs`` views are valid until the next ``step``.
    """

    def __init__(self, n_envs, num_robots=8, num_obs=10, num_cores=0, min_start_goal_dis=30.0,
                 device="cuda:0", seed=0, env_offset=0, max_robots=None, max_obs=None,
                 max_cores=None, params=None, reset_params=None, debug_outputs=False, fused=True,
                 prefetch="stream", prefetch_every=4):
        _lib.lib()  # fail loudly if the CUDA library is missing
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("BatchedMarineNavEnv needs a CUDA device (no CPU fallback)")
        self.E = int(n_envs)
        self.RP = int(max_robots or num_robots)
        self.OP = int(max_obs if max_obs is not None else num_obs)
        self.CP = int(max_cores if max_cores is not None else num_cores)
        self.params = params if params is not None else _abi.default_params()
        self.reset_params = reset_params if reset_params is not None else _lib.default_reset_params()
        self.seed = int(seed)
        self.env_offset = int(env_offset)
        self.total_steps = 0     # batched steps taken (drives the sensor-noise counter)
        self.episode_counter = 0  # reset launches (drives the scenario counter)
        with torch.cuda.device(self.device):
            self.state = _lib.new_state(self.E, self.RP, self.OP, self.CP, self.device)
            self._outs = [_lib.new_out(self.E, self.RP, self.device, debug=debug_outputs) for _ in range(2)]
            self._st_ptrs = _lib.state_ptrs(self.state)
            self._o_ptrs = [_lib.out_ptrs(o) for o in self._outs]
            self._cur = 0
            i32 = dict(dtype=torch.int32, device=self.device)
            self.want_robots = torch.full((self.E,), int(num_robots), **i32)
            self.want_obs = torch.full((self.E,), int(num_obs), **i32)
            self.want_cores = torch.full((self.E,), int(num_cores), **i32)
            self.min_start_goal_dis = torch.full((self.E,), float(min_start_goal_dis),
                                                 dtype=torch.float32, device=self.device)
            # terminal observation rows of envs whose episode ended in the last fused step
            self.term = {"obs_self": torch.zeros((self.E, self.RP, 7), dtype=torch.float32, device=self.device),
                         "obs_objects": torch.zeros((self.E, self.RP, 5, 5), dtype=torch.float32, device=self.device),
                         "obs_mask": torch.zeros((self.E, self.RP, 5), dtype=torch.float32, device=self.device)}
            # per-env episode counters: the scenario RNG counter of env e's next episode is idx[e] + 1
            self.episode_idx = torch.zeros((self.E,), dtype=torch.int32, device=self.device)
            # pre-generated next episodes (a second state block + the episode number each holds)
            #   prefetch="stream":  asvrl_env_stage_refill on a low-priority side stream every
            #                       `prefetch_every` steps: scenario AND first observation staged, the
            #                       step never waits for a generation that is ready in time
            #   prefetch="workers": extra CTAs of the fused launch itself stage the scenario (no second
            #                       stream: what a captured CUDA graph of the step uses)
            #   prefetch=False:     ended envs are always re-generated in place by their CTA
            if prefetch is True:
                prefetch = "stream"
            if prefetch not in (False, None, "stream", "workers"):
                raise ValueError("prefetch must be False, 'stream' or 'workers'")
            self.prefetch = prefetch if fused else False
            self.prefetch_every = max(1, int(prefetch_every))
            self.stage = self.stage_tag = self._stage_ptrs = self.stage_out = None
            if self.prefetch:
                self.stage = _lib.new_state(self.E, self.RP, self.OP, self.CP, self.device)
                self._stage_ptrs = _lib.state_ptrs(self.stage)
                self.stage_tag = torch.zeros((self.E,), dtype=torch.int32, device=self.device)
                self.stage_out = _lib.new_out(self.E, self.RP, self.device, debug=False)
                self._stage_out_ptrs = _lib.out_ptrs(self.stage_out)
                self._need_ws = torch.zeros((self.E,), dtype=torch.uint8, device=self.device)
                self._ep_ws = torch.zeros((self.E,), dtype=torch.int32, device=self.device)
                self._side = torch.cuda.Stream(device=self.device, priority=0)  # lowest priority
                self._since_refill = 0
        self.fused = bool(fused)  # step(auto_reset=True) = ONE launch (asvrl_env_step_autoreset)
        self.launches = 0
        # optional int64 device word added to both RNG counters on the device: a captured CUDA
        # graph of step()/finish_episodes() bumps it and so draws fresh streams on every replay
        self.counter_dev = None

    # ------------------------------------------------------------------ schedule (ENV:75-87)
    def set_schedule(self, num_robots=None, num_obs=None, num_cores=None, min_start_goal_dis=None):
        """Curriculum stage applied to episodes generated from now on."""
        if num_robots is not None:
            assert 0 < num_robots <= self.RP, "Number of robots is 0!"
            self.want_robots.fill_(int(num_robots))
        if num_obs is not None:
            self.want_obs.fill_(min(int(num_obs), self.OP))
        if num_cores is not None:
            self.want_cores.fill_(min(int(num_cores), self.CP))
        if min_start_goal_dis is not None:
            self.min_start_goal_dis.fill_(float(min_start_goal_dis))
        if self.stage_tag is not None:
            # staged episodes were generated for the previous stage; a refill still in flight on the
            # side stream must not publish after the tags are cleared
            if self.prefetch == "stream":
                torch.cuda.current_stream(self.device).wait_stream(self._side)
            self.stage_tag.zero_()

    # ------------------------------------------------------------------ views
    @property
    def out(self):
        return self._outs[self._cur]

    @property
    def obs(self):
        o = self._outs[self._cur]
        return o["obs_self"], o["obs_objects"], o["obs_mask"]

    # ------------------------------------------------------------------ reset / step
    def _reset_and_observe(self, mask, out_idx):
        """K1b on the masked envs, then a perception-only K1 pass on the same envs (ENV:164)."""
        with torch.cuda.device(self.device):
            self.episode_counter += 1
            _lib.env_reset(self.reset_params, self.state, mask, self.want_robots, self.want_cores,
                           self.want_obs, self.min_start_goal_dis, self.seed, self.episode_counter,
                           self.OP, self.CP, self.env_offset, st_ptrs=self._st_ptrs,
                           counter_dev=self.counter_dev)
            _lib.env_step(self.params, self.state, None, None, self._outs[out_idx], 0, self.OP,
                          self.CP, seed=self.seed ^ self.OBSERVE_SALT, step=self.episode_counter,
                          env_offset=self.env_offset, env_mask=mask, st_ptrs=self._st_ptrs,
                          o_ptrs=self._o_ptrs[out_idx], counter_dev=self.counter_dev)
            self.launches += 2

    def reset(self):
        """Fresh scenario in every env; returns the observation tensors (device views)."""
        self._reset_and_observe(None, self._cur)
        # every env now holds episode number `episode_counter`; fused steps count on from there
        # (episode_counter also grows by one per fused step, an upper bound of every idx)
        self.episode_idx.fill_(self.episode_counter)
        if self.prefetch == "stream":
            self.refill_stage(wait=True)
        return self.obs

    def refill_stage(self, wait=False):
        """Enqueue asvrl_env_stage_refill on the side stream: every env whose stage does not hold
        its next episode gets scenario + first observation generated beside the steps."""
        with torch.cuda.device(self.device):
            main = torch.cuda.current_stream(self.device)
            self._side.wait_stream(main)  # counters / schedule written so far are visible
            _lib.env_stage_refill(self.params, self.reset_params, self._stage_ptrs, self._stage_out_ptrs,
                                  self.want_robots, self.want_cores, self.want_obs, self.min_start_goal_dis,
                                  self.episode_idx, self.stage_tag, self._need_ws, self._ep_ws, self.seed,
                                  self.seed ^ self.OBSERVE_SALT, self.E, self.RP, self.OP, self.CP,
                                  self.env_offset, stream=self._side)
            self.launches += 4
            if wait:
                main.wait_stream(self._side)
        self._since_refill = 0

    def step(self, actions, is_continuous_action=True, auto_reset=True):
        """One action for every ASV of every env.

        actions: CUDA float32 [E, RP, 2] in (-1, 1) (continuous, WAMV:220-223) or int32 [E, RP]
        ids into the 5x5 thrust-change table (WAMV:146-147).  Returns the out dict of THIS
        step (next observation, reward, status, env_status, ended); with ``auto_reset`` the
        finished envs are then re-generated and re-observed in place, so ``self.obs`` is what
        the policy should act on next while the returned dict's reward/status still describe the
        transition (use ``step_no_reset`` + ``finish_episodes`` to store next_obs in between).
        """
        if auto_reset and self.fused:
            return self.step_fused(actions, is_continuous_action)
        out = self.step_no_reset(actions, is_continuous_action)
        if auto_reset:
            self.finish_episodes()
        return out

    OBSERVE_SALT = 0x5DEECE66D

    def step_fused(self, actions, is_continuous_action=True):
        """K1, K1b and the first observation of re-generated envs in ONE launch.  Returns the out
        dict of this step: obs_* = what the policy acts on next (fresh for ended envs), reward /
        status / env_status / ended = the transition; ``self.term`` holds the terminal observation
        rows of the ended envs (the replay row's next state)."""
        mode = _abi.M_DYNAMICS | _abi.M_AUTO_DEACTIVATE
        if not is_continuous_action:
            mode |= _abi.M_DISCRETE
        nxt = 1 - self._cur
        with torch.cuda.device(self.device):
            self.total_steps += 1
            self.episode_counter += 1
            _lib.env_step_autoreset(self.params, self.reset_params, self.state, actions, self._outs[nxt],
                                    mode, self.OP, self.CP, self.want_robots, self.want_cores,
                                    self.want_obs, self.min_start_goal_dis, self.episode_idx,
                                    seed=self.seed, step=self.total_steps,
                                    observe_seed=self.seed ^ self.OBSERVE_SALT, term=self.term,
                                    stage_ptrs=self._stage_ptrs, stage_tag=self.stage_tag,
                                    stage_obs=self.stage_out if self.prefetch == "stream" else None,
                                    refill_in_launch=self.prefetch == "workers",
                                    env_offset=self.env_offset, st_ptrs=self._st_ptrs,
                                    o_ptrs=self._o_ptrs[nxt], counter_dev=self.counter_dev)
            self.launches += 1
        self._cur = nxt
        if self.prefetch == "stream" and not torch.cuda.is_current_stream_capturing():
            self._since_refill += 1
            if self._since_refill >= self.prefetch_every:
                self.refill_stage()
        return self._outs[nxt]

    def step_no_reset(self, actions, is_continuous_action=True):
        mode = _abi.M_DYNAMICS | _abi.M_AUTO_DEACTIVATE
        if not is_continuous_action:
            mode |= _abi.M_DISCRETE
        nxt = 1 - self._cur
        with torch.cuda.device(self.device):
            self.total_steps += 1
            _lib.env_step(self.params, self.state, actions, None, self._outs[nxt], mode, self.OP,
                          self.CP, seed=self.seed, step=self.total_steps, env_offset=self.env_offset,
                          st_ptrs=self._st_ptrs, o_ptrs=self._o_ptrs[nxt], counter_dev=self.counter_dev)
            self.launches += 1
        self._cur = nxt
        return self._outs[nxt]

    def finish_episodes(self):
        """Re-generate every env whose episode ended in the last step (TRAINER:172, 240-245) and
        overwrite its observation rows with the fresh episode's first observation."""
        self._reset_and_observe(self._outs[self._cur]["ended"], self._cur)

    # ------------------------------------------------------------------ CUDA graph of one step
    def enable_cuda_graph(self, is_continuous_action=True):
        """Capture step() (the fused launch, or K1, K1b, masked K1) once per ping-pong phase.  The RNG counters baked
        into the two graphs differ by one and a device word advances them by two per pair of
        replays, so every replay draws fresh noise and fresh scenarios."""
        dev = self.device
        if self.prefetch == "stream":
            self.prefetch = "workers"  # a captured step cannot fork to the side stream: refill in the launch
            torch.cuda.synchronize(dev)
        with torch.cuda.device(dev):
            self.counter_dev = torch.zeros((1,), dtype=torch.int64, device=dev)
            shape, dt = ((self.E, self.RP, 2), torch.float32) if is_continuous_action else ((self.E, self.RP), torch.int32)
            self.graph_actions = torch.zeros(shape, dtype=dt, device=dev)
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                for _ in range(2):
                    self.step(self.graph_actions, is_continuous_action)
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            self._graphs = {}
            for i in range(2):
                phase = self._cur
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    self.step(self.graph_actions, is_continuous_action)  # flips self._cur
                    if i == 1:
                        self.counter_dev.add_(2)
                self._graphs[phase] = g
        return self._graphs

    def step_graph(self, actions=None):
        """Replay the captured step for the current phase; ``actions`` (device tensor) is copied
        into the graph's static action buffer first.  Returns the out dict of this step."""
        if actions is not None:
            self.graph_actions.copy_(actions)
        self._graphs[self._cur].replay()
        self._cur = 1 - self._cur
        self.launches += 1 if self.fused else 3
        return self._outs[self._cur]

    # ------------------------------------------------------------------ host-facing convenience
    def new_host_out(self):
        """Pinned host mirror of the step outputs a CPU-side caller reads (one flat block)."""
        return _lib.new_out(self.E, self.RP, None, pin=True)

    def step_host(self, actions_pinned, host_out, is_continuous_action=True):
        """End-to-end call with HOST buffers: pinned actions in; observations (post auto-reset),
        rewards and status out (the call shape a CPU-side caller of the reference's step() makes).
        One H2D copy, the kernels, one D2H copy of the flat output block; the caller synchronises
        the stream before reading ``host_out``."""
        if not hasattr(self, "_act_dev"):
            self._act_dev = torch.empty(actions_pinned.shape, dtype=actions_pinned.dtype, device=self.device)
        if getattr(self, "_graphs", None):
            self.graph_actions.copy_(actions_pinned, non_blocking=True)
            self.step_graph()
        else:
            self._act_dev.copy_(actions_pinned, non_blocking=True)
            self.step(self._act_dev, is_continuous_action)
        host_out["host_block"].copy_(self.out["host_block"], non_blocking=True)
        return host_out
