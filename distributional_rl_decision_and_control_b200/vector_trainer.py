"""Many-environment training loop: the reference's ``Trainer.learn`` schedule
(rfarl/rfarl/policy/trainer.py:85-255) with every per-robot Python loop replaced by one launch
over all environments of this GPU.

Per batched step:
  act     ``Agent.act_batch``           one policy forward for E x R ASVs        (TRAINER:110-135)
  step    K1 ``asvrl_env_step``         dynamics, perception, reward, deactivate (TRAINER:140, 168-170)
  store   K2a ``asvrl_replay_push``     transitions of ASVs active at step start (TRAINER:157-166)
  reset   K1b + masked K1               finished episodes re-generated           (TRAINER:172, 240-245)
and every ``update_every`` batched steps ``updates_per_round`` learner updates (K2b, K3, K4 +
the torch MLPs), a hard target sync every ``target_update_interval`` updates.

Nothing crosses PCIe inside the loop.  With ``cuda_graph=True`` two consecutive collect steps
(the observation buffers ping-pong) are captured in one CUDA graph and the learner update in
another; RNG counters live in device words that the graphs advance themselves.
Under torchrun each rank owns an env shard and a replay shard; the only collective is the
gradient all-reduce inside ``Agent.train``.
"""
import time

import torch

from .envs.marinenav.batched import BatchedMarineNavEnv


class VectorTrainer:
    def __init__(self, env: BatchedMarineNavEnv, agent, update_every=4, updates_per_round=1,
                 learning_starts=2000, target_update_interval=10000, exploration_fraction=0.25,
                 initial_eps=0.6, final_eps=0.05, schedule=None, cuda_graph=False):
        self.env = env
        self.agent = agent
        self.update_every = int(update_every)
        self.updates_per_round = int(updates_per_round)
        self.learning_starts = int(learning_starts)          # in batched steps
        self.target_update_interval = int(target_update_interval)  # in learner updates
        self.exploration_fraction = exploration_fraction
        self.initial_eps = initial_eps
        self.final_eps = final_eps
        self.schedule = schedule          # same dict as the reference's training_schedule (ENV:75-87)
        self.discrete = agent.agent_type in ("IQN", "DQN", "Rainbow")
        self.rainbow = agent.agent_type == "Rainbow"
        if self.rainbow and agent.memory.lanes != env.E * env.RP:
            raise RuntimeError("Rainbow on a batched env needs Agent(replay_lanes=n_envs * max_robots)")
        self.steps = 0                    # batched steps
        self.updates = 0
        self.env_timesteps = 0            # env-steps = batched steps x envs (the reference's unit)
        self.cuda_graph = bool(cuda_graph)
        self._collect_graph = None
        self._eps_dev = torch.zeros((), dtype=torch.float32, device=env.device)
        self._stage = -1
        self._buffer_ready = False

    # ------------------------------------------------------------------ schedules
    def linear_eps(self, total_steps):
        """TRAINER:257-264 over batched steps."""
        progress = self.steps / max(total_steps, 1)
        if progress < self.exploration_fraction:
            return self.initial_eps + progress / self.exploration_fraction * (self.final_eps - self.initial_eps)
        return self.final_eps

    def _apply_schedule(self):
        """Curriculum stage from the env-step count (ENV:75-87); applies to episodes generated
        from now on, exactly like the reference's lookup at reset()."""
        if self.schedule is None:
            return
        idx = sum(1 for t in self.schedule["timesteps"] if t <= self.env_timesteps) - 1
        if idx != self._stage:
            self._stage = idx
            s = self.schedule
            self.env.set_schedule(s["num_robots"][idx], s["num_obstacles"][idx], s["num_cores"][idx],
                                  s["min_start_goal_dis"][idx])

    # ------------------------------------------------------------------ one batched step
    def _collect_eager(self):
        env, agent = self.env, self.agent
        E, RP = env.E, env.RP
        o_self, o_obj, o_mask = env.obs
        flat = (o_self.view(E * RP, 7), o_obj.view(E * RP, 5, 5), o_mask.view(E * RP, 5))
        actions = agent.act_batch(flat, self._eps_dev)
        actions = actions.view(E, RP) if self.discrete else actions.view(E, RP, 2)
        if self.rainbow:  # the n-step memory stores (s_t, a_t, r_t+1, terminal), no next state (TRAINER:163-164)
            out = env.step_fused(actions, is_continuous_action=False) if env.fused else \
                env.step(actions, is_continuous_action=False)
            agent.memory.append_step((o_self, o_obj, o_mask), actions, out)
            return out
        if env.fused:  # one launch: step + re-generation + first observation; terminal rows in env.term
            out = env.step_fused(actions, is_continuous_action=not self.discrete)
            agent.memory.push_step((o_self, o_obj, o_mask), actions, self.discrete, out, term=env.term)
            return out
        out = env.step_no_reset(actions, is_continuous_action=not self.discrete)
        agent.memory.push_step((o_self, o_obj, o_mask), actions, self.discrete, out)
        env.finish_episodes()
        return out

    def _capture_collect(self):
        env = self.env
        dev = env.device
        if getattr(env, "prefetch", None) == "stream":
            env.prefetch = "workers"  # a captured step refills the stage inside its own launch
            torch.cuda.synchronize(dev)
        with torch.cuda.device(dev):
            env.counter_dev = torch.zeros((1,), dtype=torch.int64, device=dev)
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                for _ in range(2):
                    self._collect_eager()
                env.counter_dev.add_(2)
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for _ in range(2):  # two steps: the ping-pong observation buffers return to phase
                    self._collect_eager()
                env.counter_dev.add_(2)  # the host counters baked into the graph differ by 1
            self._collect_graph = g
        self.steps += 2  # the warm-up steps really ran; the captured ones did not
        self.env_timesteps += 2 * env.E

    def collect(self, eps):
        """Advance every environment by one action (two when graphed); returns batched steps."""
        self._eps_dev.fill_(float(eps))
        if not self.cuda_graph:
            self._collect_eager()
            n = 1
        else:
            if self._collect_graph is None:
                self._capture_collect()
            self._collect_graph.replay()
            n = 2
        self.steps += n
        self.env_timesteps += n * self.env.E
        return n

    # ------------------------------------------------------------------ loop
    def learn(self, total_steps, log_every=0):
        """Run ``total_steps`` batched steps.  Returns a dict of throughput counters."""
        env, agent = self.env, self.agent
        self._apply_schedule()
        env.reset()
        t0 = time.perf_counter()
        last_losses = None
        since_update = 0
        while self.steps < total_steps:
            self._apply_schedule()
            n = self.collect(self.linear_eps(total_steps))
            since_update += n
            if self.steps >= self.learning_starts and agent.training and since_update >= self.update_every:
                since_update = 0
                if not self._buffer_ready:  # one host sync until the ring holds a batch (TRAINER:186)
                    have = agent.memory.transitions.num_elements() if self.rainbow else agent.memory.size()
                    ready = have > agent.BATCH_SIZE and (not self.rainbow or agent.memory.slots_filled() > agent.memory.n + 1)
                    pg = getattr(agent, "process_group", None)
                    if pg is not None:
                        # every rank enters the gradient exchange in the same round: ready = MIN over ranks
                        import torch.distributed as dist
                        flag = torch.tensor([1 if ready else 0], dtype=torch.int32, device=env.device)
                        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=pg)
                        ready = bool(flag.item())
                    self._buffer_ready = ready
                if self._buffer_ready and self.cuda_graph and agent._graph is None:
                    agent.enable_cuda_graph()
                if self._buffer_ready:
                    for _ in range(self.updates_per_round):
                        last_losses = agent.train_graph_replay() if agent._graph is not None else \
                            getattr(agent, agent._TRAINERS[agent.agent_type])()
                        self.updates += 1
                        if self.updates % self.target_update_interval == 0:
                            agent.soft_update()
            if log_every and self.steps % log_every < n:
                torch.cuda.synchronize(env.device)
                dt = time.perf_counter() - t0
                print(f"steps {self.steps} env-steps {self.env_timesteps} updates {self.updates} "
                      f"env-steps/s {self.env_timesteps / dt:.3e}")
        torch.cuda.synchronize(env.device)
        dt = time.perf_counter() - t0
        if last_losses is not None:
            last_losses = tuple(float(x) for x in last_losses) if isinstance(last_losses, tuple) \
                else float(last_losses.mean())
        return {"steps": self.steps, "env_steps": self.env_timesteps, "updates": self.updates,
                "seconds": dt, "env_steps_per_sec": self.env_timesteps / dt,
                "updates_per_sec": self.updates / dt, "last_losses": last_losses}
