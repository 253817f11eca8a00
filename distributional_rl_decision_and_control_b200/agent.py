"""Agent façade with the reference's constructor, ``act_*`` / ``train`` / ``soft_update`` /
``save_latest_model`` / ``load_model`` surface (rfarl/rfarl/agent.py:18-707) for the two
distributional agents on the hot path, AC-IQN and IQN.

What changed underneath (same results, different machinery):
  * replay sampling, padding and masks come from the device ring (kernel K2b), not from python
    lists re-uploaded on every call (AGENT:387-392);
  * tau sampling + the cosine table are kernel K3, the Bellman target + pairwise TD matrix +
    quantile-Huber loss + its gradient are ONE launch of kernel K4 (AGENT:401-412);
  * ``act_batch`` evaluates the policy for every ASV of every environment in one forward
    (the reference runs one batch-1 forward per ASV per step, AGENT:207-225);
  * ``enable_cuda_graph()`` captures a whole update (sample -> critic step -> actor step) in a
    CUDA graph: at B=64 the update is ~100 tiny launches and is launch-latency bound;
  * under ``torch.distributed`` the flat gradient of each optimiser step is averaged with one
    NCCL all-reduce BEFORE clipping, so all ranks clip and step identically.
The other four agent types of the reference (DDPG, DQN, SAC, Rainbow; AGENT:478-641) run on the
same device env batch and device replay: their networks are plain dense layers in PyTorch, their
updates read the K2b batch tensors (Rainbow: the device prioritised n-step memory of
utils/replay_memory_rainbow.py) and share the gradient exchange / clip / Adam tail.
"""
import copy
import random

import numpy as np
import torch
import torch.optim as optim

from .policy.AC_IQN_model import AC_IQN_Policy
from .policy.DDPG_model import DDPG_Policy
from .policy.DQN_model import DQN_Policy
from .policy.IQN_model import IQN_Policy
from .policy.Rainbow_model import Rainbow_Policy
from .policy.SAC_model import SAC_Policy
from . import _lib
from .policy.iqn_ops import quantile_huber_loss, quantile_huber_loss_grad
from .utils.replay_buffer import ReplayBuffer, state_batch

_TARGET_SALT = 0x9E3779B97F4A7C15


class Agent():
    def __init__(self,
                 self_dimension=7,
                 object_dimension=5,
                 max_object_num=5,
                 self_feature_dimension=56,
                 object_feature_dimension=40,
                 concat_feature_dimension=256,
                 hidden_dimension=128,
                 value_ranges_of_action=[[-1.0, 1.0], [-1.0, 1.0]],  # AC-IQN
                 action_size=25,  # IQN
                 multi_steps=3,
                 BATCH_SIZE=64,
                 BUFFER_SIZE=1_000_000,
                 LR=1e-4,
                 TAU=1.0,
                 GAMMA=0.99,
                 device="cpu",
                 seed=0,
                 training=True,
                 agent_type=None,
                 num_tau=8,
                 process_group=None,
                 rank=0,
                 matmul_precision="fp32",
                 replay_lanes=1,
                 fused_trunk=True,
                 fused_optimizer=True):
        self.device = torch.device(device) if not isinstance(device, torch.device) else device
        self.LR = LR
        self.TAU = TAU
        self.GAMMA = GAMMA
        self.BUFFER_SIZE = BUFFER_SIZE
        self.BATCH_SIZE = BATCH_SIZE
        self.training = training
        self.value_ranges_of_action = copy.deepcopy(value_ranges_of_action)
        self.action_size = action_size
        self.agent_type = agent_type
        self.num_tau = int(num_tau)  # N = N' (8 in the reference: ACIQN:462, AGENT:407)
        self.process_group = process_group
        self.rank = int(rank)
        # "fp32": the reference's arithmetic (parity mode, 1e-5 on quantiles and losses);
        # "tf32": the Linear layers of train() run on the tensor cores (throughput mode, ~1e-3)
        if matmul_precision not in ("fp32", "tf32"):
            raise ValueError("matmul_precision must be 'fp32' or 'tf32'")
        self.matmul_precision = matmul_precision
        # AC-IQN: the critic's quantile trunk runs through K5 (policy/iqn_ops.critic_trunk); False
        # keeps the module-by-module autograd path (the torch fp32 reference of the parity tests)
        self.fused_trunk = bool(fused_trunk)
        self._actor_dz = None   # the constant gradient of the actor loss, (B, N) of -1 / (B N)
        self._graph = None
        self._dims = (self_dimension, object_dimension, max_object_num, self_feature_dimension,
                      object_feature_dimension, concat_feature_dimension, hidden_dimension)

        if training:
            if self.device.type != "cuda":
                raise RuntimeError("training runs on a CUDA device only: the replay ring and the IQN "
                                   "kernels have no CPU fallback")
            capt = dict(lr=self.LR, capturable=True)
            if fused_optimizer:  # K7: clip_grad_norm_(0.5) + Adam in two launches (utils/clip_adam.py)
                from .utils.clip_adam import ClipAdam
                _Adam = lambda ps, **kw: ClipAdam(ps, lr=self.LR, max_norm=0.5, process_group=process_group)
            else:
                _Adam = optim.Adam
            if agent_type == "AC-IQN":
                self.policy_local = AC_IQN_Policy(*self._dims, value_ranges_of_action, self.device, seed)
                self.policy_target = AC_IQN_Policy(*self._dims, value_ranges_of_action, self.device, seed)
                self.actor_optimizer = _Adam(self.policy_local.actor.parameters(), **capt)
                self.critic_optimizer = _Adam(self.policy_local.critic.parameters(), **capt)
                samplers = [(self.policy_local.critic.tau_sampler, 0),
                            (self.policy_target.critic.tau_sampler, _TARGET_SALT)]
            elif agent_type == "IQN":
                self.policy_local = IQN_Policy(*self._dims, action_size, self.device, seed).to(self.device)
                self.policy_target = IQN_Policy(*self._dims, action_size, self.device, seed).to(self.device)
                self.optimizer = _Adam(self.policy_local.parameters(), **capt)
                samplers = [(self.policy_local.tau_sampler, 0),
                            (self.policy_target.tau_sampler, _TARGET_SALT)]
            elif agent_type == "DDPG":  # AGENT:96-120
                self.policy_local = DDPG_Policy(*self._dims, value_ranges_of_action, self.device, seed)
                self.policy_target = DDPG_Policy(*self._dims, value_ranges_of_action, self.device, seed)
                self.actor_optimizer = _Adam(self.policy_local.actor.parameters(), **capt)
                self.critic_optimizer = _Adam(self.policy_local.critic.parameters(), **capt)
                samplers = []
            elif agent_type == "DQN":  # AGENT:121-143
                self.policy_local = DQN_Policy(*self._dims, action_size, self.device, seed).to(self.device)
                self.policy_target = DQN_Policy(*self._dims, action_size, self.device, seed).to(self.device)
                self.optimizer = _Adam(self.policy_local.parameters(), **capt)
                samplers = []
            elif agent_type == "SAC":  # AGENT:144-168
                self.alpha = 0.078
                self.policy_local = SAC_Policy(*self._dims, value_ranges_of_action, self.device, seed)
                self.policy_target = SAC_Policy(*self._dims, value_ranges_of_action, self.device, seed)
                self.actor_optimizer = _Adam(self.policy_local.actor.parameters(), **capt)
                self.critic_1_optimizer = _Adam(self.policy_local.critic_1.parameters(), **capt)
                self.critic_2_optimizer = _Adam(self.policy_local.critic_2.parameters(), **capt)
                samplers = []
            elif agent_type == "Rainbow":  # AGENT:169-198
                self.atoms = 51
                self.Vmin = -1.0
                self.Vmax = 1.0
                self.support = torch.linspace(self.Vmin, self.Vmax, self.atoms).to(self.device)
                self.delta_z = (self.Vmax - self.Vmin) / (self.atoms - 1)
                self.n = multi_steps
                self.policy_local = Rainbow_Policy(*self._dims, action_size, self.atoms, self.device, seed).to(self.device)
                self.policy_target = Rainbow_Policy(*self._dims, action_size, self.atoms, self.device, seed).to(self.device)
                self.optimizer = _Adam(self.policy_local.parameters(), **capt)
                samplers = []
            else:
                raise RuntimeError("Agent type not implemented!")
            if not self.fused_trunk:  # the torch reference path end to end: no K5, no K6
                for pol in (self.policy_local, self.policy_target):
                    mods = [pol] if isinstance(pol, torch.nn.Module) else \
                        [m for m in vars(pol).values() if isinstance(m, torch.nn.Module)]
                    for m in mods:
                        m.fused_tower = False
            # local and target nets must not share a tau stream, and neither may two ranks
            for s, salt in samplers:
                s.seed = (s.seed ^ salt ^ (self.rank * 0xD1B54A32D192ED03)) & 0xFFFFFFFFFFFFFFFF
            if agent_type == "Rainbow":  # prioritised n-step memory (AGENT:202-203); ``replay_lanes`` =
                # ASVs per batched step when fed by append_step (1 = the reference's host path)
                from .utils.replay_memory_rainbow import ReplayMemory
                self.memory = ReplayMemory(self.device, BUFFER_SIZE, lanes=replay_lanes, n=multi_steps)
            else:
                self.memory = ReplayBuffer(BUFFER_SIZE, BATCH_SIZE, object_dimension, max_object_num,
                                           seed=249 + 7919 * self.rank, device=self.device)

    # ------------------------------------------------------------------ acting
    def state_to_tensor(self, states):
        """AGENT:360-368: host lists -> float32 tensors on the agent's device; the object tensors
        are None when the batch has no object at all."""
        self_b, obj_b, mask_b = states
        dev = self.device
        self_t = torch.tensor(self_b).float().to(dev)
        if len(obj_b) == 0:
            return (self_t, None, None)
        return (self_t, torch.tensor(obj_b).float().to(dev), torch.tensor(mask_b).float().to(dev))

    def act_ac_iqn(self, state, eps=0.0, cvar=1.0, use_eval=True):
        """epsilon-greedy action of ONE ASV (AGENT:207-225)."""
        if random.random() > eps:
            x = self.state_to_tensor(state_batch([state]))
            actor = self.policy_local.actor
            actor.eval() if use_eval else actor.train()
            with torch.no_grad():
                action = actor(x).cpu().data.numpy()[0].tolist()
            actor.train()
            return action
        return [np.random.uniform(low=lo, high=hi) for lo, hi in self.value_ranges_of_action]

    def act_iqn(self, state, eps=0.0, cvar=1.0, use_eval=True):
        """(action id, quantiles, taus) of ONE ASV (AGENT:227-250)."""
        x = self.state_to_tensor(state_batch([state]))
        net = self.policy_local
        net.eval() if use_eval else net.train()
        with torch.no_grad():
            quantiles, taus = net(x, net.K, cvar)
            action_values = quantiles.mean(dim=1)
        net.train()
        if random.random() > eps:
            action = np.argmax(action_values.cpu().data.numpy())
        else:
            action = random.choice(np.arange(self.action_size))
        return action, quantiles.cpu().data.numpy(), taus.cpu().data.numpy()

    def _greedy_continuous(self, state, eps, use_eval, sac=False):
        if random.random() > eps:
            x = self.state_to_tensor(state_batch([state]))
            actor = self.policy_local.actor
            actor.eval() if use_eval else actor.train()
            with torch.no_grad():
                out = actor(x)
                action = (out[0] if sac else out).cpu().data.numpy()[0].tolist()
            actor.train()
            return action
        return [np.random.uniform(low=lo, high=hi) for lo, hi in self.value_ranges_of_action]

    def act_ddpg(self, state, eps=0.0, use_eval=True):
        """AGENT:252-269."""
        return self._greedy_continuous(state, eps, use_eval)

    def act_sac(self, state, eps=0.0, use_eval=True):
        """AGENT:289-306: the sampled (squashed) action of the stochastic actor."""
        return self._greedy_continuous(state, eps, use_eval, sac=True)

    def act_dqn(self, state, eps=0.0, use_eval=True):
        """AGENT:271-287."""
        x = self.state_to_tensor(state_batch([state]))
        net = self.policy_local
        net.eval() if use_eval else net.train()
        with torch.no_grad():
            action_values = net(x)
        net.train()
        if random.random() > eps:
            return np.argmax(action_values.cpu().data.numpy())
        return random.choice(np.arange(self.action_size))

    def act_rainbow(self, state, eps=0.0, use_eval=True):
        """AGENT:308-324."""
        x = self.state_to_tensor(state_batch([state]))
        net = self.policy_local
        net.eval() if use_eval else net.train()
        with torch.no_grad():
            probs = net(x)
        net.train()
        if random.random() > eps:
            return (probs * self.support).sum(2).argmax(1).item()
        return random.choice(np.arange(self.action_size))

    @torch.no_grad()
    def act_batch(self, obs, eps=0.0, cvar=1.0):
        """``_act_batch`` under this agent's GEMM arithmetic (fp32 like the reference, or TF32 with
        ``matmul_precision="tf32"``)."""
        with self._precision():
            return self._act_batch(obs, eps, cvar)

    def _act_batch(self, obs, eps=0.0, cvar=1.0):
        """Vectorised epsilon-greedy for M ASVs at once: ``obs`` = (self (M,7), objects (M,5,5),
        mask (M,5)) CUDA tensors.  Returns float32 (M,2) actions (AC-IQN) or int32 (M,) action ids
        (IQN), still on the device.  ``eps`` may be a 0-d CUDA tensor (so a captured CUDA graph
        follows the exploration schedule); exploration noise comes from torch's default CUDA
        generator, which is graph-safe."""
        M = obs[0].shape[0]
        actor = getattr(self.policy_local, "actor", None)
        if (self.agent_type in ("AC-IQN", "DDPG") and self.fused_trunk and obs[0].is_cuda
                and hasattr(actor, "act_eps_greedy") and actor.action_dimension == 2):
            # forward and exploration of all M ASVs without the ~10 element-wise launches between them
            if not hasattr(self, "_explore_counter"):
                self._explore_counter = torch.zeros((1,), dtype=torch.int64, device=self.device)
                self._explore_seed = (0xA5A5A5A5 ^ (self.rank * 0x9E3779B97F4A7C15) ^ int(actor.seed_id)) & 0xFFFFFFFFFFFFFFFF
            actions = actor.act_eps_greedy(obs, eps, self._explore_seed, self._explore_counter)
            self._explore_counter.add_(1)  # on the stream (and inside a captured graph): fresh noise every call
            return actions
        explore = torch.rand((M,), device=self.device) < eps
        if self.agent_type in ("AC-IQN", "DDPG", "SAC"):
            if self.agent_type == "SAC":
                greedy = self.policy_local.actor.act(obs)   # the sampled action without the log-probability terms
            else:
                greedy = self.policy_local.actor(obs)
            if not hasattr(self, "_act_lo"):
                self._act_lo = torch.tensor([r[0] for r in self.value_ranges_of_action], device=self.device)
                self._act_hi = torch.tensor([r[1] for r in self.value_ranges_of_action], device=self.device)
            lo, hi = self._act_lo, self._act_hi
            rnd = lo + (hi - lo) * torch.rand((M, lo.numel()), device=self.device)
            return torch.where(explore.unsqueeze(-1), rnd, greedy).contiguous()
        net = self.policy_local
        if self.agent_type == "IQN":
            q = net.action_values_fused(obs, cvar) if self.fused_trunk else None   # K6 + K10 (head 1) + one small GEMM
            if q is None:
                quantiles, _ = net(obs, net.K, cvar)
                q = quantiles.mean(dim=1)
            greedy = q.argmax(dim=1)
        elif self.agent_type == "DQN":
            greedy = net(obs).argmax(dim=1)
        elif self.agent_type == "Rainbow":  # noisy layers stay in training mode while collecting (TRAINER:131)
            if self.fused_trunk and obs[0].is_cuda and obs[0].dtype == torch.float32 and self.atoms <= 64:
                greedy = net.greedy_actions(obs, self.support)   # one pass over the output layers' products
            else:
                greedy = (net(obs) * self.support).sum(2).argmax(1)
        else:
            raise RuntimeError("Agent type not implemented!")
        rnd = torch.randint(0, self.action_size, (M,), device=self.device)
        return torch.where(explore, rnd, greedy).to(torch.int32).contiguous()

    # ------------------------------------------------------------------ learning
    def train(self):
        """One learner update (AGENT:370-384).  AC-IQN returns (critic_loss, actor_loss), IQN the
        loss, as numpy scalars like the reference (AGENT:432, 476)."""
        if self._graph is not None:
            out = self.train_graph_replay()
        elif self.agent_type in self._TRAINERS:
            out = getattr(self, self._TRAINERS[self.agent_type])()
        else:
            raise RuntimeError("Agent type not implemented!")
        if isinstance(out, tuple):
            return tuple(o.detach().cpu().numpy() for o in out)
        return out.detach().cpu().numpy()

    _TRAINERS = {"AC-IQN": "train_AC_IQN", "IQN": "train_IQN", "DDPG": "train_DDPG", "DQN": "train_DQN",
                 "SAC": "train_SAC", "Rainbow": "train_Rainbow"}

    def _sync_grads(self, params):
        """Data-parallel exchange step: average ONE flat fp32 bucket over the ranks."""
        if self.process_group is None:
            return
        import torch.distributed as dist
        grads = [p.grad for p in params if p.grad is not None]
        flat = torch.cat([g.reshape(-1) for g in grads])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.process_group)
        flat.div_(dist.get_world_size(self.process_group))
        off = 0
        for g in grads:
            n = g.numel()
            g.copy_(flat[off:off + n].view_as(g))
            off += n

    def _clip_and_step(self, optimizer, params):
        """all-reduce (N > 1) -> clip_grad_norm_(0.5) -> Adam, AGENT:414-416 and twins."""
        if hasattr(optimizer, "max_norm"):  # ClipAdam: exchange, clip and update inside K7's two launches
            optimizer.step()
            return
        self._sync_grads(params)
        torch.nn.utils.clip_grad_norm_(params, 0.5)
        optimizer.step()

    def _batch(self, batch):
        b = self.memory.sample_device() if batch is None else batch
        return ((b["self"], b["obj"], b["mask"]), b["act"], b["rew"], (b["nself"], b["nobj"], b["nmask"]),
                b["done"])

    def _precision(self):
        """Context that selects the GEMM arithmetic of a learner update (and of its graph capture)."""
        import contextlib

        @contextlib.contextmanager
        def ctx():
            old = torch.backends.cuda.matmul.allow_tf32
            torch.backends.cuda.matmul.allow_tf32 = self.matmul_precision == "tf32"
            try:
                yield
            finally:
                torch.backends.cuda.matmul.allow_tf32 = old
        return ctx()

    def train_AC_IQN(self, batch=None, taus=None):
        with self._precision():
            return self._train_AC_IQN(batch, taus)

    def train_IQN(self, batch=None, taus=None):
        with self._precision():
            return self._train_IQN(batch, taus)

    def _train_AC_IQN(self, batch=None, taus=None):
        """AGENT:386-432 on device tensors.  Returns (critic_loss, actor_loss) as 0-d CUDA tensors.
        ``batch`` injects the sample; ``taus`` = {"target", "local", "actor"} -> (B, N) tensors injects
        the three tau draws in the reference's order (AGENT:399, 403, 422) for the parity tests."""
        states, actions, rewards, next_states, dones = self._batch(batch)
        N = self.num_tau
        taus = taus or {}
        local, target = self.policy_local, self.policy_target

        # critic step ------------------------------------------------------------------
        self.critic_optimizer.zero_grad(set_to_none=True)
        fused = self.fused_trunk
        with torch.no_grad():
            if fused and target.actor._fused_policy_ok(next_states):
                # K9 with eps = 0: tower, both dense layers, head and atan of the target actor in one tcgen05 launch
                next_actions = target.actor.act_eps_greedy(next_states, 0.0, 0, None)   # AGENT:396
            else:
                next_actions = target.actor(next_states)                 # AGENT:396
            crit_t = target.critic.forward_fused if fused else target.critic
            z_next, _ = crit_t(next_states, next_actions, N, taus=taus.get("target"))  # AGENT:399
        crit = local.critic.forward_fused if fused else local.critic
        z, tau = crit(states, actions, N, taus=taus.get("local"))        # AGENT:403
        if fused:
            # K4 = AGENT:401-412; its launch already holds d loss / d z, so the backward starts at z
            critic_loss, _, dz = quantile_huber_loss_grad(z, tau.squeeze(-1), z_next, rewards, dones, self.GAMMA, 1.0)
            z.backward(dz)
        else:
            critic_loss, _ = quantile_huber_loss(z, tau.squeeze(-1), z_next, rewards, dones,
                                                 self.GAMMA, 1.0)        # K4 = AGENT:401-412
            critic_loss.backward()
        self._clip_and_step(self.critic_optimizer, list(local.critic.parameters()))  # AGENT:414-416

        # actor step -------------------------------------------------------------------
        self.actor_optimizer.zero_grad(set_to_none=True)
        if fused:
            q, _ = local.critic.forward_fused(states, local.actor(states), N, taus=taus.get("actor"),
                                              param_grads=False)         # AGENT:420-422
        else:
            q, _ = local.critic(states, local.actor(states), N, taus=taus.get("actor"))
        if fused:
            # AGENT:423-425: the loss in one launch; its gradient is the constant -1 / (B N), kept between updates
            actor_loss = _lib.neg_mean(q.detach())
            if self._actor_dz is None or self._actor_dz.shape != q.shape:
                self._actor_dz = torch.full_like(q, -1.0 / q.numel(), requires_grad=False)
            q.backward(self._actor_dz)
        else:
            actor_loss = -q.mean()
            actor_loss.backward()
        self._clip_and_step(self.actor_optimizer, list(local.actor.parameters()))  # AGENT:424-427
        return critic_loss.detach(), actor_loss.detach()

    def _train_IQN(self, batch=None, taus=None):
        """AGENT:434-476 on device tensors.  Returns the loss as a 0-d CUDA tensor."""
        states, actions, rewards, next_states, dones = self._batch(batch)
        N = self.num_tau
        B = rewards.shape[0]
        taus = taus or {}
        self.optimizer.zero_grad(set_to_none=True)
        with torch.no_grad():
            net_t = self.policy_target.forward_fused if self.fused_trunk else self.policy_target
            zn_all, _ = net_t(next_states, N, taus=taus.get("target"))  # (B, N, A)
            z_next = zn_all.max(dim=2)[0]                                # AGENT:451-452
        net = self.policy_local.forward_fused if self.fused_trunk else self.policy_local
        z_all, tau = net(states, N, taus=taus.get("local"))
        a = actions[:, 0].long().view(B, 1, 1).expand(B, N, 1)           # AGENT:457-458
        z = z_all.gather(2, a).squeeze(-1)
        loss, _ = quantile_huber_loss(z, tau.squeeze(-1), z_next, rewards, dones, self.GAMMA, 1.0)
        loss.backward()
        self._clip_and_step(self.optimizer, list(self.policy_local.parameters()))  # AGENT:470-473
        return loss.detach()

    # ------------------------------------------------------------------ the four non-distributional agents
    def _step(self, optimizer, loss, params):
        """zero_grad / backward / all-reduce / clip / Adam, as every update of the reference does
        (AGENT:414-416 and twins).  zero_grad first: losses of SAC share nothing across optimisers."""
        optimizer.zero_grad(set_to_none=True)
        loss.backward()
        self._clip_and_step(optimizer, params)

    def train_DDPG(self, batch=None):
        """AGENT:478-516 on device tensors.  Returns (critic_loss, actor_loss) as 0-d CUDA tensors."""
        with self._precision():
            states, actions, rewards, next_states, dones = self._batch(batch)
            rewards, dones = rewards.unsqueeze(-1), dones.unsqueeze(-1)
            local, target = self.policy_local, self.policy_target
            with torch.no_grad():
                q_next = target.critic(next_states, target.actor(next_states))
                q_target = rewards + self.GAMMA * q_next * (1. - dones)
            critic_loss = torch.nn.functional.smooth_l1_loss(local.critic(states, actions), q_target)
            self._step(self.critic_optimizer, critic_loss, list(local.critic.parameters()))
            actor_loss = -local.critic(states, local.actor(states)).mean()
            self._step(self.actor_optimizer, actor_loss, list(local.actor.parameters()))
            return critic_loss.detach(), actor_loss.detach()

    def train_DQN(self, batch=None):
        """AGENT:518-545 on device tensors.  Returns the loss as a 0-d CUDA tensor."""
        with self._precision():
            states, actions, rewards, next_states, dones = self._batch(batch)
            rewards, dones = rewards.unsqueeze(-1), dones.unsqueeze(-1)
            with torch.no_grad():
                q_next = self.policy_target(next_states).max(dim=1, keepdim=True)[0]
                q_target = rewards + (1 - dones) * self.GAMMA * q_next
            q = self.policy_local(states).gather(1, actions[:, :1].long())
            loss = torch.nn.functional.smooth_l1_loss(q, q_target)
            self._step(self.optimizer, loss, list(self.policy_local.parameters()))
            return loss.detach()

    def train_SAC(self, batch=None, noise=None):
        """AGENT:547-595 on device tensors.  Returns (critic_1_loss, critic_2_loss, actor_loss).
        ``noise`` = {"target", "actor"} -> (B, 2) standard-normal draws (parity tests)."""
        with self._precision():
            states, actions, rewards, next_states, dones = self._batch(batch)
            rewards, dones = rewards.unsqueeze(-1), dones.unsqueeze(-1)
            noise = noise or {}
            local, target = self.policy_local, self.policy_target
            mse = torch.nn.functional.mse_loss
            with torch.no_grad():
                next_actions, next_log_probs = target.actor(next_states, noise.get("target"))
                tv = torch.min(target.critic_1(next_states, next_actions),
                               target.critic_2(next_states, next_actions)) - self.alpha * next_log_probs
                tv = rewards + self.GAMMA * tv * (1. - dones)
            loss_1 = mse(local.critic_1(states, actions), tv)
            loss_2 = mse(local.critic_2(states, actions), tv)
            self._step(self.critic_1_optimizer, loss_1, list(local.critic_1.parameters()))
            self._step(self.critic_2_optimizer, loss_2, list(local.critic_2.parameters()))
            new_actions, log_probs = local.actor(states, noise.get("actor"))
            values_new = torch.min(local.critic_1(states, new_actions), local.critic_2(states, new_actions))
            actor_loss = (self.alpha * log_probs - values_new).mean()
            self._step(self.actor_optimizer, actor_loss, list(local.actor.parameters()))
            return loss_1.detach(), loss_2.detach(), actor_loss.detach()

    def train_Rainbow(self, sample=None):
        """AGENT:597-641 (categorical double-Q with n-step returns and prioritised replay) on device
        tensors; priorities are updated on the device.  Returns the per-sample loss (B,)."""
        with self._precision():
            B, A = self.BATCH_SIZE, self.atoms
            idxs, states, actions, returns, next_states, nonterminals, weights = \
                self.memory.sample(B) if sample is None else sample
            rows = torch.arange(B, device=self.device)
            log_ps_a = self.policy_local(states, log=True)[rows, actions]          # log p(s_t, a_t)
            with torch.no_grad():
                pns = self.policy_local(next_states)                              # online net picks a*
                a_star = (self.support.expand_as(pns) * pns).sum(2).argmax(1)
                self.policy_target.reset_noise()
                pns_a = self.policy_target(next_states)[rows, a_star]             # target net evaluates it
                # Tz = R^n + gamma^n z, clamped, projected onto the fixed support (AGENT:616-632)
                Tz = returns.unsqueeze(1) + nonterminals * (self.GAMMA ** self.n) * self.support.unsqueeze(0)
                Tz = Tz.clamp(min=self.Vmin, max=self.Vmax)
                b = (Tz - self.Vmin) / self.delta_z
                lo, up = b.floor().to(torch.int64), b.ceil().to(torch.int64)
                lo = torch.where((up > 0) & (lo == up), lo - 1, lo)                # b integral: keep the mass
                up = torch.where((lo < (A - 1)) & (lo == up), up + 1, up)
                m = torch.zeros((B, A), dtype=torch.float32, device=self.device)
                m.scatter_add_(1, lo, pns_a * (up.float() - b))
                m.scatter_add_(1, up, pns_a * (b - lo.float()))
            loss = -torch.sum(m * log_ps_a, 1)                                    # cross entropy
            self._step(self.optimizer, (weights * loss).mean(), list(self.policy_local.parameters()))
            self.memory.update_priorities(idxs, loss.detach())
            return loss.detach()

    # ------------------------------------------------------------------ CUDA graph of one update
    def enable_cuda_graph(self, warmup=3):
        """Capture sample -> critic step -> actor step as one CUDA graph.  RNG counters of the
        sampler and of the tau streams are read from a device word that the graph itself
        advances, so every replay draws a fresh batch and fresh taus."""
        have = self.memory.transitions.num_elements() if self.agent_type == "Rainbow" else self.memory.size()
        if have < self.BATCH_SIZE:
            raise RuntimeError("fill the replay buffer with at least one batch before capturing")
        step = getattr(self, self._TRAINERS[self.agent_type])
        dev = self.device
        import gc
        gc.collect()  # no finaliser of an unrelated CUDA tensor may run in the middle of the capture
        with torch.cuda.device(dev):
            self._ctr = torch.zeros((1,), dtype=torch.int64, device=dev)
            self.memory.counter_dev = self._ctr
            nets = [self.policy_local.critic, self.policy_target.critic] if self.agent_type == "AC-IQN" \
                else [self.policy_local, self.policy_target] if self.agent_type == "IQN" else []
            for n in nets:
                n.tau_sampler.counter_dev = self._ctr
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                for _ in range(warmup):
                    step()
                    self._ctr.add_(16)
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                out = step()
                self._ctr.add_(16)
            self._graph, self._graph_out = graph, out
        return graph

    def train_graph_replay(self):
        self._graph.replay()
        return self._graph_out

    def disable_cuda_graph(self):
        self._graph = None

    # ------------------------------------------------------------------ target sync, checkpoints
    def soft_update(self):
        """theta_target = TAU * theta_local + (1 - TAU) * theta_target (AGENT:643-656; TAU = 1.0 in
        the reference, i.e. a hard copy every target_update_interval steps)."""
        for name in ("optimizer", "actor_optimizer", "critic_optimizer", "critic_1_optimizer", "critic_2_optimizer"):
            opt = getattr(self, name, None)
            if hasattr(opt, "check"):
                opt.check()  # one device read per target sync: K7 skips non-finite steps, say so loudly
        if self.agent_type in ("AC-IQN", "DDPG"):
            pairs = [(self.policy_target.actor, self.policy_local.actor),
                     (self.policy_target.critic, self.policy_local.critic)]
        elif self.agent_type == "SAC":
            pairs = [(self.policy_target.actor, self.policy_local.actor),
                     (self.policy_target.critic_1, self.policy_local.critic_1),
                     (self.policy_target.critic_2, self.policy_local.critic_2)]
        elif self.agent_type in ("IQN", "DQN", "Rainbow"):
            pairs = [(self.policy_target, self.policy_local)]
        else:
            raise RuntimeError("Agent type not implemented!")
        with torch.no_grad():
            for tgt, src in pairs:
                tp, sp = list(tgt.parameters()), list(src.parameters())
                if self.TAU == 1.0:
                    torch._foreach_copy_(tp, sp)
                else:
                    torch._foreach_mul_(tp, 1.0 - self.TAU)
                    torch._foreach_add_(tp, sp, alpha=self.TAU)

    def save_latest_model(self, directory):
        self.policy_local.save(directory)

    def load_model(self, path, device="cpu"):
        """Restores policy_local only, like the reference (AGENT:684-698).  Parameters are copied
        in place so optimisers and a captured graph stay valid."""
        cls = {"AC-IQN": AC_IQN_Policy, "IQN": IQN_Policy, "DDPG": DDPG_Policy, "DQN": DQN_Policy,
               "SAC": SAC_Policy, "Rainbow": Rainbow_Policy}.get(self.agent_type)
        if cls is None:
            raise RuntimeError("Agent type not implemented!")
        loaded = cls.load(path, device)
        if not hasattr(self, "policy_local"):
            self.policy_local = loaded
            return
        parts = {"AC-IQN": ("actor", "critic"), "DDPG": ("actor", "critic"),
                 "SAC": ("actor", "critic_1", "critic_2")}.get(self.agent_type)
        if parts:
            for name in parts:
                getattr(self.policy_local, name).load_state_dict(getattr(loaded, name).state_dict())
        else:
            self.policy_local.load_state_dict(loaded.state_dict())


def calculate_huber_loss(td_errors, k=1.0):
    """Element-wise Huber loss (AGENT:701-707) — kept for callers of the reference helper; the
    training path uses the fused kernel (policy/iqn_ops.quantile_huber_loss) instead."""
    a = td_errors.abs()
    return torch.where(a <= k, 0.5 * td_errors.pow(2), k * (a - 0.5 * k))
