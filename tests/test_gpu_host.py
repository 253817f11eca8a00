"""Host layer on the B200: the rfarl-shaped classes over the C-ABI kernels.
 * critic forward and ONE full Agent.train_AC_IQN step against the reference's own outputs
   (tests/golden/agent_ac_iqn.npz: injected batch and taus, reference weights for the seed);
 * MarineNavEnv3.reset()/step() surface against the float64 oracle on the same state and the
   same Philox noise stream;
 * Trainer.learn / VectorTrainer.learn end to end, eager and CUDA-graph."""
import json
import os

import numpy as np
import pytest
import torch

from distributional_rl_decision_and_control_b200 import _abi
from distributional_rl_decision_and_control_b200.agent import Agent
from distributional_rl_decision_and_control_b200.envs.marinenav.batched import BatchedMarineNavEnv
from distributional_rl_decision_and_control_b200.envs.marinenav.env import MarineNavEnv3
from distributional_rl_decision_and_control_b200.policy.trainer import Trainer
from distributional_rl_decision_and_control_b200.vector_trainer import VectorTrainer
import golden_util as gu

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def fx():
    z = np.load(os.path.join(gu.GOLD, "agent_ac_iqn.npz"))
    return {k: z[k] for k in z.files}


def dev_batch(d):
    t = lambda k: torch.from_numpy(d["b_" + k]).to(DEV)
    return {k: t(k) for k in ("self", "obj", "mask", "act", "rew", "nself", "nobj", "nmask", "done")}


def test_critic_forward_matches_reference(fx):
    agent = Agent(device=DEV, seed=int(fx["seed"]), agent_type="AC-IQN", BUFFER_SIZE=1024)
    b = dev_batch(fx)
    with torch.no_grad():
        z, tau = agent.policy_local.critic((b["self"], b["obj"], b["mask"]), b["act"],
                                           taus=torch.from_numpy(fx["tau_local"]).to(DEV))
    assert tau.shape == (64, 8, 1)
    assert np.allclose(z.cpu().numpy(), fx["critic_out"], rtol=1e-5, atol=1e-5)


def test_train_step_matches_reference(fx):
    """AGENT:386-432 end to end: losses to 1e-5 relative; after one Adam step the weights agree
    (an Adam step is lr * g / (|g| + 1e-8): only gradients within ~1e-7 of zero may differ)."""
    agent = Agent(device=DEV, seed=int(fx["seed"]), agent_type="AC-IQN", BUFFER_SIZE=1024)
    taus = {k: torch.from_numpy(fx["tau_" + k]).to(DEV) for k in ("target", "local", "actor")}
    critic_loss, actor_loss = agent.train_AC_IQN(batch=dev_batch(fx), taus=taus)
    assert abs(critic_loss.item() - float(fx["critic_loss"])) <= 1e-5 * abs(float(fx["critic_loss"]))
    assert abs(actor_loss.item() - float(fx["actor_loss"])) <= 1e-5 * abs(float(fx["actor_loss"]))
    for name, mod in (("actor", agent.policy_local.actor), ("critic", agent.policy_local.critic)):
        for k, v in mod.state_dict().items():
            got, want, init = v.cpu().numpy(), fx[f"post_{name}.{k}"], fx[f"init_{name}.{k}"]
            assert not np.array_equal(want, init), k          # the step moved the reference too
            close = np.abs(got - want) <= 2e-6
            assert close.mean() >= 0.995, (name, k, close.mean())
            assert np.abs(got - want).max() <= 2.1e-4, (name, k)  # at most one flipped Adam sign


@pytest.mark.parametrize("fused", [True, False])
def test_train_steps_at_config3_match_float64_restatement(fused):
    """BASELINE config 3 (B=1024, N=N'=32), three consecutive updates from random weights, batches and
    tau tables: the product update (K3, K4, K5, K6, K7 + the GEMMs) against the float64 restatement of
    AGENT:386-432 (oracle/learner_oracle.py, pinned to the reference by tests/test_oracle_golden.py)."""
    from oracle import learner_oracle as lo
    B, N = 1024, 32
    agent = Agent(device=DEV, seed=3, agent_type="AC-IQN", BATCH_SIZE=B, num_tau=N, BUFFER_SIZE=2048,
                  fused_trunk=fused, fused_optimizer=fused)
    actor32, critic32 = lo.init_weights(41)
    for pol in (agent.policy_local, agent.policy_target):
        pol.actor.load_state_dict({k: v.to(DEV) for k, v in actor32.items()})
        pol.critic.load_state_dict({k: v.to(DEV) for k, v in critic32.items()})
    actor, critic = ({k: v.double() for k, v in actor32.items()}, {k: v.double() for k, v in critic32.items()})
    ta, tc = {k: v.clone() for k, v in actor.items()}, {k: v.clone() for k, v in critic.items()}
    oa, oc = lo.new_adam_state(actor), lo.new_adam_state(critic)
    for step in range(3):
        b32, t32 = lo.synthetic_batch(B, N, 100 + step)
        want = lo.train_ac_iqn(actor, critic, ta, tc, {k: v.double() for k, v in b32.items()},
                               {k: v.double() for k, v in t32.items()}, oa, oc)
        got = agent.train_AC_IQN(batch={k: v.to(DEV) for k, v in b32.items()},
                                 taus={k: v.to(DEV) for k, v in t32.items()})
        for g, w in zip(got, want):  # float32 product against float64 restatement
            assert abs(g.item() - w) <= 2e-5 * abs(w) + 2e-6, (step, g.item(), w)
        for name, mod, ref in (("actor", agent.policy_local.actor, actor), ("critic", agent.policy_local.critic, critic)):
            for k, v in mod.state_dict().items():
                err = (v.double().cpu() - ref[k]).abs().numpy()
                # Adam steps are +-lr-sized and sign-unstable where |g| ~ eps: nearly all elements agree
                # closely, none is off by more than one flipped step per update
                assert np.mean(err <= 5e-6) >= 0.99 and err.max() <= 2.1e-4 * (step + 1), (step, name, k, err.max())


def test_agent_train_from_device_replay_and_graph():
    agent = Agent(device=DEV, seed=1, agent_type="AC-IQN", BUFFER_SIZE=4096)
    rng = np.random.RandomState(0)
    for i in range(300):
        n1, n2 = rng.randint(0, 6), rng.randint(0, 6)
        s = (rng.randn(7).tolist(), rng.randn(n1, 5).tolist())
        s2 = (rng.randn(7).tolist(), rng.randn(n2, 5).tolist())
        agent.memory.add((s, (rng.rand(2) * 2 - 1).tolist(), float(rng.randn()), s2, bool(rng.rand() < 0.1)))
    assert agent.memory.size() == 300
    st, ac, rw, ns, dn = agent.memory.sample()
    assert len(st[0]) == 64 and len(st[1][0]) == 5 and len(ac[0]) == 2 and isinstance(dn[0], bool)
    c0, a0 = agent.train()
    assert np.isfinite(c0) and np.isfinite(a0) and c0.shape == ()
    w0 = agent.policy_local.critic.output_layer.weight.detach().clone()
    agent.enable_cuda_graph()
    losses, idx = [], []
    for _ in range(4):
        c, a = agent.train()
        losses.append(float(c))
        idx.append(agent.memory._batch["indices"].clone())
    assert len(set(losses)) == 4                       # fresh batch + fresh taus on every replay
    assert not torch.equal(idx[0], idx[1])
    assert not torch.equal(w0, agent.policy_local.critic.output_layer.weight)
    agent.soft_update()
    for a, b in zip(agent.policy_local.critic.parameters(), agent.policy_target.critic.parameters()):
        assert torch.equal(a, b)


def test_iqn_agent_trains():
    agent = Agent(device=DEV, seed=2, agent_type="IQN", BUFFER_SIZE=2048)
    rng = np.random.RandomState(1)
    for i in range(200):
        s = (rng.randn(7).tolist(), rng.randn(rng.randint(0, 6), 5).tolist())
        s2 = (rng.randn(7).tolist(), rng.randn(rng.randint(0, 6), 5).tolist())
        agent.memory.add((s, int(rng.randint(0, 25)), float(rng.randn()), s2, False))
    loss = agent.train()
    assert np.isfinite(loss)
    a, q, t = agent.act_iqn(([0.0] * 7, []), eps=0.0)
    assert 0 <= int(a) < 25 and q.shape == (1, 32, 25) and t.shape == (1, 32, 1)


# ------------------------------------------------------------------------------------ env façade
def test_env_facade_surface_and_oracle_parity(oracle, capsys):
    sched = {"timesteps": [0, 100], "num_robots": [5, 6], "num_cores": [4, 4], "num_obstacles": [6, 6],
             "min_start_goal_dis": [30.0, 30.0]}
    env = MarineNavEnv3(seed=11, schedule=sched, device=DEV)
    obs, collisions, reach_goals = env.reset()
    assert "training schedule" in capsys.readouterr().out
    n = len(env.robots)
    assert n == 5 and len(obs) == n and collisions == [False] * n and reach_goals == [False] * n
    assert len(obs[0][0]) == 7 and all(len(o) == 5 for o in obs[0][1]) and len(obs[0][1]) <= 5
    assert len(env.obstacles) <= 6 and len(env.cores) <= 4 and env.get_action_space_dimension() == 25
    rng = np.random.RandomState(3)
    for step in range(6):
        # oracle on the very same device state, same Philox stream the facade will use
        st_host = gu.to_host(env._st)
        st_o = {k: (v.astype(np.float64) if v.dtype == np.float32 else v.copy()) for k, v in st_host.items()}
        for i, rob in enumerate(env.robots):
            st_o["flags"][0, i] = (1 if rob.collision else 0) | (2 if rob.reach_goal else 0) | (4 if rob.deactivated else 0)
        st_o["ep_steps"][:] = env.episode_timesteps
        RP, OP, CP = env._shape
        acts = [None if rob.deactivated else (rng.rand(2) * 2 - 1).astype(np.float32).tolist() for rob in env.robots]
        a_arr = np.zeros((1, RP, 2))
        for i, a in enumerate(acts):
            if a is not None:
                a_arr[0, i] = a
        want = oracle.env_step(env._params(), st_o, a_arr, None, _abi.M_DYNAMICS, OP, CP,
                               seed=11 ^ 0x5851F42D4C957F2D, step=env._obs_calls + 1)
        obs, rewards, dones, infos = env.step(acts, is_continuous_action=True)
        assert len(obs) == len(rewards) == len(dones) == len(infos) == n
        for i, rob in enumerate(env.robots):
            code = (int(want["status"][0, i]) >> _abi.S_INFO_SHIFT) & 15
            assert infos[i]["state"] == _abi.INFO_STRINGS[code]
            assert dones[i] == bool(want["status"][0, i] & _abi.S_DONE)
            if rob.deactivated:
                assert obs[i] == (None, None) and rewards[i] == 0
                continue
            assert abs(rewards[i] - want["reward"][0, i]) <= 1e-4 + 1e-5 * abs(want["reward"][0, i])
            k = int(want["obs_mask"][0, i].sum())
            assert len(obs[i][1]) == k
            assert np.allclose(obs[i][0], want["obs_self"][0, i], rtol=1e-5, atol=2e-4)
            assert np.allclose(np.array(obs[i][1]).reshape(k, 5), want["obs_objects"][0, i, :k], rtol=1e-5, atol=2e-4)
            assert abs(rob.x - st_o["x"][0, i]) < 1e-4 and abs(rob.theta - st_o["theta"][0, i]) < 1e-5
            if step == 2 and i == 0:  # the Trainer owns `deactivated` (TRAINER:168-170)
                rob.collision, rob.deactivated = True, True
    assert env.episode_timesteps == 6 and env.total_timesteps == 6
    with pytest.raises(AssertionError, match="Number of actions"):
        env.step([None], True)


def test_episode_data_round_trip_and_eval_histories(tmp_path):
    env = MarineNavEnv3(seed=253, is_eval_env=True, device=DEV)
    env.num_robots, env.num_cores, env.num_obs, env.min_start_goal_dis = 4, 3, 5, 30.0
    env.reset()
    cfg = json.loads(json.dumps(env.episode_data()))   # must be JSON-serialisable (TRAINER:80-83)
    assert set(cfg) == {"env", "robots"} and cfg["env"]["num_robots"] == 4
    assert len(cfg["robots"]["start"]) == len(env.robots) and "perception" in cfg["robots"]
    env2 = MarineNavEnv3(seed=0, is_eval_env=True, device=DEV)
    obs, _, _ = env2.reset_with_eval_config(cfg)
    assert len(obs) == len(env.robots)
    for a, b in zip(env.robots, env2.robots):
        assert np.allclose(a.start, b.start) and np.allclose(a.goal, b.goal) and a.init_theta == b.init_theta
        assert np.allclose(b.velocity - b.velocity_r, env2.get_velocity(b.x, b.y), atol=1e-6)
    assert [(o.x, o.y, o.r) for o in env.obstacles] == [(o.x, o.y, o.r) for o in env2.obstacles]
    env2.step([[0.5, 0.5]] * len(env2.robots), True)
    rob = env2.robots[0]
    assert len(rob.trajectory) == 2 and len(rob.trajectory[1]) == 13 and rob.action_history == [[0.5, 0.5]]
    assert len(rob.observation_history) == 2
    env2.save_episode(str(tmp_path / "ep.json"))
    assert json.load(open(tmp_path / "ep.json"))["robots"]["trajectory"][0][1][11] == 250.0


def test_trainer_learn_end_to_end(tmp_path, capsys):
    """train_RL_agents.py's run_trial body (TRAIN:74-109) on a short budget."""
    sched = {"timesteps": [0], "num_robots": [3], "num_cores": [0], "num_obstacles": [0], "min_start_goal_dis": [30.0]}
    ev = {"num_episodes": [1], "num_robots": [3], "num_cores": [0], "num_obstacles": [0], "min_start_goal_dis": [30.0]}
    train_env = MarineNavEnv3(seed=0, schedule=sched, device=DEV)
    eval_env = MarineNavEnv3(seed=253, is_eval_env=True, device=DEV)
    agent = Agent(device=DEV, seed=100, agent_type="AC-IQN", BUFFER_SIZE=10000)
    trainer = Trainer(train_env=train_env, eval_env=eval_env, eval_schedule=ev, rl_agent=agent,
                      learning_starts=40, imitation=False, il_agent=None)
    trainer.save_eval_config(str(tmp_path))
    trainer.learn(total_timesteps=120, eval_freq=80, eval_log_path=str(tmp_path), verbose=False)
    capsys.readouterr()
    files = set(os.listdir(tmp_path))
    assert {"eval_configs.json", "evaluations.npz", "actor_network_params.pth", "critic_network_params.pth"} <= files
    ev_data = np.load(tmp_path / "evaluations.npz", allow_pickle=True)
    assert set(ev_data.files) == {"timesteps", "observations", "actions", "trajectories", "rewards",
                                  "successes", "times", "energies", "relations"}
    assert list(ev_data["timesteps"]) == [40, 80]
    assert agent.memory.size() > 64 and trainer.current_timestep == 121


@pytest.mark.parametrize("graph", [False, True])
@pytest.mark.parametrize("kind", ["AC-IQN", "IQN"])
def test_vector_trainer(kind, graph):
    env = BatchedMarineNavEnv(128, num_robots=5, num_obs=4, num_cores=2, device=DEV, seed=5,
                              max_robots=8, max_obs=10, max_cores=8)
    agent = Agent(device=DEV, seed=7, agent_type=kind, BUFFER_SIZE=50000)
    sched = {"timesteps": [0, 128 * 20], "num_robots": [5, 8], "num_cores": [2, 4], "num_obstacles": [4, 6],
             "min_start_goal_dis": [30.0, 30.0]}
    vt = VectorTrainer(env, agent, update_every=4, learning_starts=8, target_update_interval=3,
                       schedule=sched, cuda_graph=graph)
    w0 = [p.detach().clone() for p in (agent.policy_local.critic if kind == "AC-IQN" else agent.policy_local).parameters()]
    res = vt.learn(60)
    assert res["steps"] >= 60 and res["updates"] >= 10 and res["env_steps"] == res["steps"] * 128
    losses = res["last_losses"]
    assert np.all(np.isfinite(losses))
    w1 = list((agent.policy_local.critic if kind == "AC-IQN" else agent.policy_local).parameters())
    assert any(not torch.equal(a, b) for a, b in zip(w0, w1))
    hs = agent.memory.ring["head_size"].cpu().numpy()
    assert 0 < hs[1] <= 50000 and hs[0] >= hs[1]
    rows = agent.memory.ring["rows"][: int(hs[1])]
    assert torch.isfinite(rows).all()
    assert int(env.want_robots[0].item()) == 8            # the curriculum advanced (ENV:75-87)
    assert int(env.state["n_robots"].max().item()) in (5, 8)  # 8 once an episode ended after the switch


@pytest.mark.parametrize("kind,graph", [("DDPG", False), ("DDPG", True), ("DQN", True), ("SAC", False),
                                        ("SAC", True), ("Rainbow", False)])
def test_vector_trainer_other_agents(kind, graph):
    """The four non-distributional agents of the reference (AGENT:478-641) on the shared device env
    batch: collection, device replay (Rainbow: prioritised n-step memory), updates, target sync."""
    E, RP = 96, 8
    env = BatchedMarineNavEnv(E, num_robots=6, num_obs=4, num_cores=2, device=DEV, seed=9,
                              max_robots=RP, max_obs=10, max_cores=8)
    env.reset()
    env.state["ep_steps"][::5] = 990  # some episodes end inside the run
    agent = Agent(device=DEV, seed=7, agent_type=kind, BUFFER_SIZE=40000,
                  replay_lanes=E * RP if kind == "Rainbow" else 1)
    vt = VectorTrainer(env, agent, update_every=4, learning_starts=8, target_update_interval=3, cuda_graph=graph)
    nets = {"DDPG": lambda p: p.critic, "SAC": lambda p: p.critic_1}.get(kind, lambda p: p)
    w0 = [p.detach().clone() for p in nets(agent.policy_local).parameters()]
    # learn() resets the env: keep the shortened clocks by patching after the first reset
    res = vt.learn(48)
    assert res["steps"] >= 48 and res["updates"] >= 8
    assert np.all(np.isfinite(res["last_losses"]))
    assert any(not torch.equal(a, b) for a, b in zip(w0, nets(agent.policy_local).parameters()))
    if kind == "Rainbow":
        m = agent.memory
        assert m.transitions.num_elements() > 0 and torch.isfinite(m.rows).all()
        idxs, s, a, R, ns, nonterm, w = m.sample(64)
        assert s[0].shape == (64, 7) and ns[1].shape == (64, 5, 5) and a.dtype == torch.int64
        assert torch.isfinite(R).all() and float(w.max()) == 1.0 and (m.prio[idxs] > 0).all()
    else:
        hs = agent.memory.ring["head_size"].cpu().numpy()
        assert 0 < hs[1] <= 40000


def test_batched_evaluation_matches_sequential(tmp_path, capsys):
    """Trainer.evaluation over the frozen scenarios as one device batch == one scenario after the
    other (TRAINER:266-387).  Sensor noise is switched off in the stored scenarios so that both
    paths see the same observations (their noise streams are keyed by the env index)."""
    from distributional_rl_decision_and_control_b200.evaluation import BatchedEvaluator, summarize
    ev = {"num_episodes": [3, 3], "num_robots": [3, 5], "num_cores": [0, 4], "num_obstacles": [0, 6],
          "min_start_goal_dis": [30.0, 35.0]}
    sched = {"timesteps": [0], "num_robots": [3], "num_cores": [0], "num_obstacles": [0], "min_start_goal_dis": [30.0]}
    agent = Agent(device=DEV, seed=100, agent_type="AC-IQN", BUFFER_SIZE=1000)
    tr = Trainer(train_env=MarineNavEnv3(seed=0, schedule=sched, device=DEV),
                 eval_env=MarineNavEnv3(seed=253, is_eval_env=True, device=DEV), eval_schedule=ev, rl_agent=agent)
    for cfg in tr.eval_config:
        n = cfg["env"]["num_robots"]
        cfg["robots"]["perception"]["pos_std"] = [0.0] * n
        cfg["robots"]["perception"]["vel_std"] = [0.0] * n
        cfg["robots"]["perception"]["r_mean_ratio"] = [1.0] * n
    tr.evaluation()
    capsys.readouterr()
    seq = {k: getattr(tr, "eval_" + k)[-1] for k in ("rewards", "successes", "times", "energies", "trajectories")}
    res = BatchedEvaluator(tr.eval_config, device=DEV, seed=253).run(agent)
    assert len(res["rewards"]) == 6 and res["successes"] == [bool(x) for x in seq["successes"]]
    lengths_seq = [max(len(t) for t in ep) - 1 for ep in seq["trajectories"]]
    assert res["lengths"] == lengths_seq
    assert np.allclose(res["times"], seq["times"], rtol=1e-5)
    assert np.allclose(res["rewards"], seq["rewards"], rtol=5e-3, atol=5e-3)
    assert np.allclose(res["energies"], seq["energies"], rtol=5e-3)
    # recorded histories have the reference's row layout (13 columns, WAMV:191-193)
    t0 = res["trajectories"][0][0]
    assert len(t0[0]) == 13 and len(res["actions"][0][0]) == len(t0) - 1  # the reset row has no action
    assert np.allclose(t0[-1][:3], seq["trajectories"][0][0][-1][:3], atol=2e-2)
    s = summarize(res)
    assert 0.0 <= s["success_rate"] <= 1.0 and np.isfinite(s["avg_reward"])
    # the Trainer switch fills the same containers
    tr2 = Trainer(train_env=tr.train_env, eval_env=tr.eval_env, eval_schedule=ev, rl_agent=agent, batched_evaluation=True)
    tr2.eval_config = tr.eval_config
    tr2.evaluation()
    capsys.readouterr()
    assert np.allclose(tr2.eval_rewards[-1], res["rewards"], rtol=1e-6)
    tr2.save_evaluation(str(tmp_path))
    assert "evaluations.npz" in os.listdir(tmp_path)


def _reference_eval_golden():
    import json
    with open(os.path.join(gu.GOLD, "eval_reference.json")) as f:
        g = json.load(f)
    w = np.load(os.path.join(gu.GOLD, "traced_ac_iqn.npz"))
    return g, {k[2:]: torch.from_numpy(w[k]) for k in w.files if k.startswith("w.")}


class _ReplayPolicy:
    """Stands in for the Agent during evaluation and hands out the actions the REFERENCE's policy took (teacher
    forcing).  Serves both call styles: act_batch (the vectorised evaluator; one (E, R, 2) tensor per step) and
    act_ac_iqn (the sequential Trainer: robots in index order, deactivated robots skipped)."""
    agent_type, GAMMA, training = "AC-IQN", 0.99, False

    def __init__(self, results, gamma, RP):
        self.GAMMA = gamma
        self.policy_local = torch.nn.Module()
        self.results, self.RP, self.step = results, RP, 0
        self.flat = []
        for w in results:  # sequential order: scenario, step, robot (only robots that still act at that step)
            acts = w["actions"]
            for s in range(max(len(a) for a in acts)):
                self.flat.extend(a[s] for a in acts if s < len(a))
        self.flat_pos = 0

    def act_batch(self, obs, eps=0.0):
        out = torch.zeros((len(self.results), self.RP, 2), dtype=torch.float32)
        for e, w in enumerate(self.results):
            for r, a in enumerate(w["actions"]):
                if self.step < len(a):
                    out[e, r] = torch.tensor(a[self.step])
        self.step += 1
        return out.to(obs[0].device).view(-1, 2)

    def act_ac_iqn(self, state, eps=0.0, use_eval=True):
        a = self.flat[self.flat_pos]
        self.flat_pos += 1
        return list(a)


@pytest.mark.parametrize("batched", [True, False])
def test_evaluation_matches_the_reference_trainer(batched, capsys):
    """`Trainer.evaluation` of the REFERENCE (TRAINER:266-392; golden made by oracle/gen_golden_eval.py from the
    unmodified reference playing its shipped trained AC-IQN policy on noise-free frozen scenarios).  The recorded
    actions are replayed (teacher forcing — a closed loop around a trained, nearly bang-bang policy amplifies a
    1e-7 difference to metres within ~20 steps, see the generator), so what is compared is everything the
    evaluation itself computes: end rule and success flags and every robot's episode length exact, travel time
    exact, discounted reward and energy to 1e-4, final poses to 1e-3 m — for the vectorised evaluator (one device
    batch) and for the sequential Trainer over the facade."""
    from distributional_rl_decision_and_control_b200.evaluation import BatchedEvaluator
    g, _ = _reference_eval_golden()
    want = g["results"]
    RP = max(len(w["actions"]) for w in want)
    agent = _ReplayPolicy(want, g["gamma"], RP)
    if batched:
        res = BatchedEvaluator(g["eval_configs"], device=DEV, seed=253).run(agent)
        traj = res["trajectories"]
    else:
        sched = {"timesteps": [0], "num_robots": [3], "num_cores": [0], "num_obstacles": [0], "min_start_goal_dis": [30.0]}
        ev = {"num_episodes": [0], "num_robots": [3], "num_cores": [0], "num_obstacles": [0], "min_start_goal_dis": [30.0]}
        tr = Trainer(train_env=MarineNavEnv3(seed=0, schedule=sched, device=DEV),
                     eval_env=MarineNavEnv3(seed=253, is_eval_env=True, device=DEV), eval_schedule=ev, rl_agent=agent)
        tr.eval_config = g["eval_configs"]
        tr.evaluation()
        capsys.readouterr()
        assert agent.flat_pos == len(agent.flat)  # every recorded action was asked for, in the reference's order
        res = {k: getattr(tr, "eval_" + k)[-1] for k in ("rewards", "successes", "times", "energies")}
        traj = tr.eval_trajectories[-1]
    assert [bool(x) for x in res["successes"]] == [w["success"] for w in want]
    assert [[len(t) for t in ep] for ep in traj] == [w["lengths"] for w in want]
    assert np.allclose(res["times"], [w["time"] for w in want], rtol=1e-6)
    assert np.allclose(res["rewards"], [w["reward"] for w in want], rtol=1e-4, atol=1e-4)
    assert np.allclose(res["energies"], [w["energy"] for w in want], rtol=1e-4)
    for ep, w in zip(traj, want):
        for t, fp in zip(ep, w["final_pose"]):
            assert np.allclose(t[-1][:2], fp[:2], atol=1e-3), (t[-1][:3], fp)


def test_policy_on_the_reference_evaluation_observations():
    """The other half of the evaluation parity: the observations the reference's robots saw (first robot of every
    golden scenario, first steps) -> the actions its policy took, through this package's Agent (K9 for the batch
    call, the batch-1 act_ac_iqn of the sequential path), to 5e-5 (trained weights: float32 modules are ~2e-5 away)."""
    g, weights = _reference_eval_golden()
    agent = Agent(device=DEV, seed=100, agent_type="AC-IQN", BUFFER_SIZE=1000)
    agent.policy_local.actor.load_state_dict(weights)
    states, want = [], []
    for w in g["results"]:
        for o, a in zip(w["obs_sample"], w["actions"][0]):
            states.append((o[0], o[1]))
            want.append(a)
    got_seq = np.array([agent.act_ac_iqn(s) for s in states])
    assert np.abs(got_seq - np.array(want)).max() <= 5e-5
    self_b = torch.tensor([s[0] for s in states], device=DEV)
    obj_b = torch.zeros((len(states), 5, 5), device=DEV)
    mask_b = torch.zeros((len(states), 5), device=DEV)
    for i, s in enumerate(states):
        for k, ob in enumerate(s[1][:5]):
            obj_b[i, k] = torch.tensor(ob, device=DEV)
            mask_b[i, k] = 1.0
    got_batch = agent.act_batch((self_b, obj_b, mask_b), 0.0).cpu().numpy()
    assert np.abs(got_batch - np.array(want)).max() <= 5e-5
