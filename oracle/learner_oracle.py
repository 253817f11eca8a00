"""CPU restatement of the reference's AC-IQN learner update.  TEST INFRASTRUCTURE, not product code:
only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import it.

Plain torch on the host, written as functions over weight dictionaries that carry the reference's
state-dict keys (so a golden fixture's `init_*` arrays drop straight in).  Parity is pinned:
tests/test_oracle_golden.py runs one update from tests/golden/agent_ac_iqn.npz (weights, batch and the
three tau tables recorded from the reference itself, oracle/gen_golden.py:agent_case) and compares
outputs, losses and every updated weight.

Reference lines restated (rfarl/rfarl/...):
  policy/AC_IQN_model.py:9-13     encoder = Linear + ReLU
  policy/AC_IQN_model.py:284-323  Actor.observation_processor / forward
  policy/AC_IQN_model.py:410-480  Critic.calc_cos / observation_processor / forward
  agent.py:386-432                train_AC_IQN
  agent.py:701-707                calculate_huber_loss
  torch.optim.Adam (lr 1e-4, betas (0.9, 0.999), eps 1e-8) and clip_grad_norm_(0.5), agent.py:36-40, 414-416
"""
import math

import torch

ACTOR_KEYS = ("self_encoder.0", "object_encoder.0", "hidden_layer", "hidden_layer_2", "output_layer")
CRITIC_KEYS = ("self_encoder.0", "object_encoder.0", "cos_embedding", "action_encoder.0", "hidden_layer",
               "hidden_layer_2", "output_layer")


def _lin(w, name, x):
    return x @ w[name + ".weight"].t() + w[name + ".bias"]


def features(w, x):
    """observation_processor without the quantile part (ACIQN:284-308): [self 56 | 5 x object 40]."""
    x_self, x_obj, x_mask = x
    B = x_self.shape[0]
    f_self = torch.relu(_lin(w, "self_encoder.0", x_self))
    if x_obj is None:  # RB:60-63 / ACIQN:293-294
        f_obj = torch.zeros((B, 200), dtype=f_self.dtype)
    else:
        f_obj = torch.relu(_lin(w, "object_encoder.0", x_obj.reshape(B * 5, 5))).reshape(B, 5, -1)
        f_obj = f_obj.masked_fill(x_mask.unsqueeze(-1) < 0.5, 0.0).reshape(B, -1)
    return torch.cat((f_self, f_obj), 1)


def actor_forward(w, x):
    """ACIQN:310-323: two ReLU layers, 128 -> 2, (2/pi) atan."""
    h = torch.relu(_lin(w, "hidden_layer", features(w, x)))
    h = torch.relu(_lin(w, "hidden_layer_2", h))
    return (2.0 / math.pi) * torch.atan(_lin(w, "output_layer", h))


def critic_forward(w, x, actions, tau):
    """ACIQN:410-480 with the tau table given (B, N) instead of drawn: returns quantiles (B, N)."""
    B, N = tau.shape
    pis = math.pi * torch.arange(64, dtype=tau.dtype).view(1, 1, 64)
    cos = torch.cos(tau.unsqueeze(-1) * pis).reshape(B * N, 64)
    cos_f = torch.relu(_lin(w, "cos_embedding", cos)).reshape(B, N, -1)
    f = (features(w, x).unsqueeze(1) * cos_f).reshape(B * N, -1)
    h = torch.relu(_lin(w, "hidden_layer", f)).reshape(B, N, -1)
    a = torch.relu(_lin(w, "action_encoder.0", actions))
    h = (a.unsqueeze(1) * h).reshape(B * N, -1)
    h = torch.relu(_lin(w, "hidden_layer_2", h))
    return _lin(w, "output_layer", h).reshape(B, N)


def quantile_huber(q_expected, tau, q_next, rew, done, gamma=0.99, kappa=1.0):
    """AGENT:399-412 + 701-705 for any (N, N'): mean_b mean_j sum_i |tau_i - 1{delta<0}| * Huber(delta)."""
    target = rew.view(-1, 1, 1) + gamma * q_next.detach().unsqueeze(1) * (1.0 - done.view(-1, 1, 1))
    delta = target - q_expected.unsqueeze(-1)
    huber = torch.where(delta.abs() <= kappa, 0.5 * delta.pow(2), kappa * (delta.abs() - 0.5 * kappa))
    rho = (tau.unsqueeze(-1) - (delta.detach() < 0).to(delta.dtype)).abs() * huber / kappa
    return rho.sum(dim=1).mean(dim=1).mean()


def new_adam_state(w):
    return {"step": 0, "m": {k: torch.zeros_like(v) for k, v in w.items()},
            "v": {k: torch.zeros_like(v) for k, v in w.items()}}


def clip_adam(w, grads, state, lr=1e-4, max_norm=0.5, b1=0.9, b2=0.999, eps=1e-8):
    """clip_grad_norm_(max_norm) then one torch.optim.Adam step, in place on `w`."""
    norm = torch.sqrt(sum((g.double() ** 2).sum() for g in grads.values())).to(next(iter(grads.values())).dtype)
    coef = torch.clamp(max_norm / (norm + 1e-6), max=1.0)
    state["step"] += 1
    t = state["step"]
    c1, c2 = 1.0 - b1 ** t, 1.0 - b2 ** t
    for k, g in grads.items():
        g = g * coef
        m, v = state["m"][k], state["v"][k]
        m.mul_(b1).add_(g, alpha=1.0 - b1)
        v.mul_(b2).addcmul_(g, g, value=1.0 - b2)
        w[k].sub_((lr / c1) * m / (v.sqrt() / math.sqrt(c2) + eps))
    return float(norm)


def _grads(loss, w):
    keys = list(w)
    gs = torch.autograd.grad(loss, [w[k] for k in keys], allow_unused=True)
    return {k: g for k, g in zip(keys, gs) if g is not None}


def train_ac_iqn(actor, critic, t_actor, t_critic, batch, taus, opt_actor, opt_critic, gamma=0.99):
    """One AGENT:386-432 update, in place on `actor` / `critic`.  batch: dict with self, obj, mask, nself,
    nobj, nmask, act, rew, done; taus: dict with target (B, N'), local (B, N), actor (B, N) in the
    reference's draw order (AGENT:399, 403, 422).  Returns (critic_loss, actor_loss)."""
    s = (batch["self"], batch["obj"], batch["mask"])
    ns = (batch["nself"], batch["nobj"], batch["nmask"])
    with torch.no_grad():
        q_next = critic_forward(t_critic, ns, actor_forward(t_actor, ns), taus["target"])
    cw = {k: v.requires_grad_(True) for k, v in critic.items()}
    q = critic_forward(cw, s, batch["act"], taus["local"])
    critic_loss = quantile_huber(q, taus["local"], q_next, batch["rew"], batch["done"], gamma)
    g = _grads(critic_loss, cw)
    with torch.no_grad():
        for v in critic.values():
            v.requires_grad_(False)
        clip_adam(critic, g, opt_critic)
    aw = {k: v.requires_grad_(True) for k, v in actor.items()}
    actor_loss = -critic_forward(critic, s, actor_forward(aw, s), taus["actor"]).mean()
    g = _grads(actor_loss, aw)
    with torch.no_grad():
        for v in actor.values():
            v.requires_grad_(False)
        clip_adam(actor, g, opt_actor)
    return float(critic_loss.detach()), float(actor_loss.detach())


def init_weights(seed, dtype=torch.float32):
    """Random weights of the reference's shapes (nn.Linear's default U(-1/sqrt(fan_in), ..) ranges);
    for timing and randomised parity runs, not the reference's seeded init."""
    g = torch.Generator().manual_seed(seed)
    shapes = {"self_encoder.0": (56, 7), "object_encoder.0": (40, 5), "cos_embedding": (256, 64),
              "action_encoder.0": (128, 2), "hidden_layer": (128, 256), "hidden_layer_2": (128, 128)}

    def make(keys, out):
        w = {}
        for k in keys:
            o, i = out if k == "output_layer" else shapes[k]
            b = 1.0 / math.sqrt(i)
            w[k + ".weight"] = ((torch.rand((o, i), generator=g, dtype=torch.float64) * 2 - 1) * b).to(dtype)
            w[k + ".bias"] = ((torch.rand((o,), generator=g, dtype=torch.float64) * 2 - 1) * b).to(dtype)
        return w
    return make(ACTOR_KEYS, (2, 128)), make(CRITIC_KEYS, (1, 128))


def synthetic_batch(B, N, seed, dtype=torch.float32):
    """SURVEY 8(d) config-3 inputs: scaled normal states, 0..5 objects, U(-1,1) actions, 2 % done."""
    g = torch.Generator().manual_seed(seed)
    rn = lambda *s: torch.randn(s, generator=g, dtype=torch.float64)
    ss = torch.tensor([20, 20, 2, 2, 0.5, 500, 500], dtype=torch.float64)
    so = torch.tensor([15, 15, 1, 1, 1], dtype=torch.float64)
    t = torch.arange(5).view(1, 5)
    out = {}
    for p in ("", "n"):
        cnt = torch.randint(0, 6, (B, 1), generator=g)
        mask = (t < cnt).to(torch.float64)
        out[p + "self"] = rn(B, 7) * ss
        out[p + "obj"] = rn(B, 5, 5) * so * mask.unsqueeze(-1)
        out[p + "mask"] = mask
    out["act"] = torch.rand((B, 2), generator=g, dtype=torch.float64) * 2 - 1
    out["rew"] = rn(B)
    out["done"] = (torch.rand((B,), generator=g, dtype=torch.float64) < 0.02).to(torch.float64)
    taus = {k: torch.rand((B, N), generator=g, dtype=torch.float64).to(dtype) for k in ("target", "local", "actor")}
    return {k: v.to(dtype) for k, v in out.items()}, taus


def time_updates(B, N, seconds=4.0, seed=0):
    """updates/s of train_ac_iqn on the host (torch's intra-op threads), for bench.py's CPU legs."""
    import time
    actor, critic = init_weights(seed)
    ta, tc = ({k: v.clone() for k, v in actor.items()}, {k: v.clone() for k, v in critic.items()})
    oa, oc = new_adam_state(actor), new_adam_state(critic)
    batch, taus = synthetic_batch(B, N, seed + 1)
    train_ac_iqn(actor, critic, ta, tc, batch, taus, oa, oc)  # warm-up
    n, t0 = 0, time.perf_counter()
    while True:
        train_ac_iqn(actor, critic, ta, tc, batch, taus, oa, oc)
        n += 1
        dt = time.perf_counter() - t0
        if dt >= seconds:
            return n / dt, n, dt
