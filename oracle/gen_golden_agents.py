"""Generates tests/golden/agents_more.npz: the reference's DDPG / DQN / SAC / Rainbow agents
(rfarl/rfarl/agent.py:96-198, 478-641) run UNMODIFIED in this container — initial weights for a
seed, forwards on a fixed batch, and ONE train step each on an injected batch (memory.sample is
patched; SAC's reparameterisation noise and Rainbow's target-net noise seed are recorded).
Weights are stored as per-tensor digests (sum, sum of |.|, first values).  TEST INFRASTRUCTURE ONLY.

    python -m oracle.gen_golden_agents
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_harness as rh  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def digest(sd, prefix):
    out = {}
    for k, v in sd.items():
        a = v.detach().double().numpy().ravel()
        out[f"{prefix}.{k}"] = np.concatenate([[a.sum(), np.abs(a).sum()], a[:6]])
    return out


def make_batch(rng, B, discrete):
    n_obj, n_obj2 = rng.randint(0, 6, B), rng.randint(0, 6, B)
    t = np.arange(5)[None, :]
    mask, nmask = (t < n_obj[:, None]).astype(np.float32), (t < n_obj2[:, None]).astype(np.float32)
    scale_s = np.array([20, 20, 2, 2, 0.5, 500, 500], np.float32)
    scale_o = np.array([15, 15, 1, 1, 1], np.float32)
    b = {"self": (rng.randn(B, 7) * scale_s).astype(np.float32),
         "obj": ((rng.randn(B, 5, 5) * scale_o) * mask[:, :, None]).astype(np.float32), "mask": mask,
         "nself": (rng.randn(B, 7) * scale_s).astype(np.float32),
         "nobj": ((rng.randn(B, 5, 5) * scale_o) * nmask[:, :, None]).astype(np.float32), "nmask": nmask,
         "rew": rng.randn(B).astype(np.float32), "done": (rng.rand(B) < 0.1).astype(np.float32)}
    b["act"] = rng.randint(0, 25, B).astype(np.float32) if discrete else (rng.rand(B, 2) * 2 - 1).astype(np.float32)
    return b


def main():
    rh.import_reference()
    import torch
    from rfarl.agent import Agent
    torch.set_num_threads(1)
    seed, B = 100, 64
    d = {"seed": np.int64(seed)}
    tt = torch.from_numpy
    for kind in ("DDPG", "DQN", "SAC", "Rainbow"):
        discrete = kind in ("DQN", "Rainbow")
        rng = np.random.RandomState(seed + len(kind))
        batch = make_batch(rng, B, discrete)
        d.update({f"{kind}.b_{k}": v for k, v in batch.items()})
        agent = Agent(device="cpu", seed=seed, agent_type=kind)
        nets = {"DDPG": ("actor", "critic"), "SAC": ("actor", "critic_1", "critic_2")}.get(kind)
        parts = {n: getattr(agent.policy_local, n) for n in nets} if nets else {"net": agent.policy_local}
        for n, m in parts.items():
            d.update(digest(m.state_dict(), f"{kind}.init.{n}"))
        states = (tt(batch["self"]), tt(batch["obj"]), tt(batch["mask"]))
        nstates = (tt(batch["nself"]), tt(batch["nobj"]), tt(batch["nmask"]))
        lists = lambda s: tuple(x.tolist() for x in s)
        dones = [bool(x) for x in batch["done"]]
        if kind != "Rainbow":
            acts = [int(a) for a in batch["act"]] if discrete else batch["act"].tolist()
            agent.memory.sample = lambda: (lists(states), acts, batch["rew"].tolist(), lists(nstates), dones)
        with torch.no_grad():
            if kind == "DDPG":
                a = agent.policy_local.actor(states)
                d[f"{kind}.actor_out"] = a.numpy()
                d[f"{kind}.critic_out"] = agent.policy_local.critic(states, tt(batch["act"])).numpy()
            elif kind == "DQN":
                d[f"{kind}.q_out"] = agent.policy_local(states).numpy()
            elif kind == "Rainbow":
                d[f"{kind}.p_out"] = agent.policy_local(states).numpy()
        if kind == "SAC":
            from torch.distributions import Normal
            eps = {k: rng.randn(B, 2).astype(np.float32) for k in ("fwd", "target", "actor")}
            d.update({f"{kind}.eps_{k}": v for k, v in eps.items()})
            queue = [eps["fwd"]]
            orig = Normal.rsample
            Normal.rsample = lambda self, sample_shape=torch.Size(): self.loc + self.scale * tt(queue.pop(0))
            try:
                with torch.no_grad():
                    a, lp = agent.policy_local.actor(states)
                d[f"{kind}.actor_out"], d[f"{kind}.log_prob"] = a.numpy(), lp.numpy()
                queue.extend([eps["target"], eps["actor"]])  # draw order AGENT:557, 581
                losses = agent.train_SAC()
                assert not queue
            finally:
                Normal.rsample = orig
        elif kind == "Rainbow":
            idxs = np.arange(B)
            R = rng.randn(B).astype(np.float32) * 0.3
            nonterm = (rng.rand(B, 1) > 0.1).astype(np.float32)
            w = rng.rand(B).astype(np.float32) * 0.5 + 0.5
            d.update({f"{kind}.returns": R, f"{kind}.nonterminals": nonterm, f"{kind}.weights": w})
            agent.memory.sample = lambda bs: (idxs, states, tt(batch["act"]).long(), tt(R), nstates, tt(nonterm), tt(w))
            agent.memory.update_priorities = lambda i, p: None
            torch.manual_seed(777)  # the target net's reset_noise inside train_Rainbow
            losses = (agent.train_Rainbow(),)
        else:
            losses = agent.train_DDPG() if kind == "DDPG" else (agent.train_DQN(),)
        d[f"{kind}.losses"] = np.array([np.asarray(x, np.float64).mean() for x in losses])
        if kind == "Rainbow":
            d[f"{kind}.loss_per_sample"] = np.asarray(losses[0], np.float64)
        for n, m in parts.items():
            d.update(digest(m.state_dict(), f"{kind}.post.{n}"))
        print(kind, "losses", d[f"{kind}.losses"])
    np.savez_compressed(os.path.join(GOLD, "agents_more.npz"), **d)
    print("wrote agents_more.npz")


if __name__ == "__main__":
    main()
