"""CPU-only checks of the host layer: module surface mirrored from the reference, the actor
(pure torch) against the reference's own weights / outputs, schedules, the rfarl import shim,
and the data-parallel gradient exchange over gloo with world_size 2.  No kernel is launched."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

import golden_util as gu
from distributional_rl_decision_and_control_b200.policy.AC_IQN_model import AC_IQN_Policy, Actor, Critic
from distributional_rl_decision_and_control_b200.policy.IQN_model import IQN_Policy
from distributional_rl_decision_and_control_b200.policy.trainer import Trainer
from distributional_rl_decision_and_control_b200.utils.replay_buffer import state_batch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DIMS = (7, 5, 5, 56, 40, 256, 128)


def load_agent_fixture():
    z = np.load(os.path.join(gu.GOLD, "agent_ac_iqn.npz"))
    return {k: z[k] for k in z.files}


def test_state_dict_keys_and_shapes_match_the_reference():
    """on-disk formats.  Keys / shapes of the reference's own state dicts."""
    d = load_agent_fixture()
    pol = AC_IQN_Policy(*DIMS, [[-1.0, 1.0], [-1.0, 1.0]], "cpu", 100)
    for name, mod in (("actor", pol.actor), ("critic", pol.critic)):
        want = {k[len("init_" + name) + 1:]: v.shape for k, v in d.items() if k.startswith("init_" + name + ".")}
        got = {k: tuple(v.shape) for k, v in mod.state_dict().items()}
        assert got == {k: tuple(s) for k, s in want.items()}, name
    assert sum(p.numel() for p in pol.actor.parameters()) == 50354
    assert sum(p.numel() for p in pol.critic.parameters()) == 67249
    assert sum(p.numel() for p in IQN_Policy(*DIMS, 25, "cpu", 0).parameters()) == 69961


def test_same_seed_gives_the_reference_initial_weights():
    d = load_agent_fixture()
    pol = AC_IQN_Policy(*DIMS, [[-1.0, 1.0], [-1.0, 1.0]], "cpu", int(d["seed"]))
    for k, v in pol.actor.state_dict().items():
        assert np.array_equal(v.numpy(), d["init_actor." + k]), k
    for k, v in pol.critic.state_dict().items():
        assert np.array_equal(v.numpy(), d["init_critic." + k]), k


def test_actor_forward_matches_reference_output():
    d = load_agent_fixture()
    actor = Actor(*DIMS, [[-1.0, 1.0], [-1.0, 1.0]], "cpu", int(d["seed"]))
    t = torch.from_numpy
    with torch.no_grad():
        out = actor((t(d["b_self"]), t(d["b_obj"]), t(d["b_mask"]))).numpy()
    assert np.allclose(out, d["actor_out"], rtol=1e-6, atol=1e-7)
    assert np.abs(out).max() < 1.0
    with torch.no_grad():  # no object in the whole batch: RB:60-63 / ACIQN:295-296
        none = actor((t(d["b_self"]), None, None)).numpy()
        zero = actor((t(d["b_self"]), torch.zeros(64, 5, 5), torch.zeros(64, 5))).numpy()
    assert np.array_equal(none, zero)


def test_checkpoint_files_round_trip(tmp_path):
    pol = AC_IQN_Policy(*DIMS, [[-1.0, 1.0], [-1.0, 1.0]], "cpu", 3)
    pol.save(str(tmp_path))
    assert sorted(os.listdir(tmp_path)) == ["actor_constructor_params.json", "actor_network_params.pth",
                                            "critic_constructor_params.json", "critic_network_params.pth"]
    import json
    keys = set(json.load(open(tmp_path / "actor_constructor_params.json")))
    assert keys == {"self_dimension", "object_dimension", "max_object_num", "self_feature_dimension",
                    "object_feature_dimension", "concat_feature_dimension", "hidden_dimension",
                    "value_ranges_of_action", "seed"}
    assert "action_dimension" in json.load(open(tmp_path / "critic_constructor_params.json"))
    back = AC_IQN_Policy.load(str(tmp_path), "cpu")
    for a, b in zip(pol.actor.parameters(), back.actor.parameters()):
        assert torch.equal(a, b)
    q = IQN_Policy(*DIMS, 25, "cpu", 4)
    q.save(str(tmp_path))
    assert os.path.exists(tmp_path / "network_params.pth") and os.path.exists(tmp_path / "constructor_params.json")
    q2 = IQN_Policy.load(str(tmp_path), "cpu")
    assert all(torch.equal(a, b) for a, b in zip(q.parameters(), q2.parameters()))


def test_actor_traces_to_torchscript_float64(tmp_path):
    """The export path (save_model_to_torch_script.py:16-24): float64 actor, example inputs
    ((1,7),(1,5,5),(1,5)) -> (1,2)."""
    actor = Actor(*DIMS, [[-1.0, 1.0], [-1.0, 1.0]], "cpu", 5).double()
    ex = ((torch.rand(1, 7).double(), torch.rand(1, 5, 5).double(), torch.ones(1, 5).double()),)
    traced = torch.jit.trace(actor, ex)
    path = str(tmp_path / "traced_AC_IQN_model.pt")
    traced.save(path)
    loaded = torch.jit.load(path)
    out = loaded(ex[0])
    assert out.shape == (1, 2) and out.dtype == torch.float64
    assert torch.allclose(out, actor(ex[0]))


def test_critic_needs_cuda():
    critic = Critic(*DIMS, 2, "cpu", 1)
    with pytest.raises(RuntimeError):
        critic((torch.zeros(2, 7), None, None), torch.zeros(2, 2))


def test_state_batch_padding():
    s = [([1.0] * 7, [[1.0] * 5, [2.0] * 5]), ([2.0] * 7, [])]
    self_b, obj_b, mask_b = state_batch(s)
    assert len(obj_b) == 2 and len(obj_b[0]) == 5 and obj_b[1] == [[0.0] * 5] * 5
    assert mask_b == [[1.0, 1.0, 0.0, 0.0, 0.0], [0.0] * 5]
    assert state_batch([([0.0] * 7, [])]) == ([[0.0] * 7], [], [])


def test_linear_eps_schedule():
    t = Trainer.__new__(Trainer)
    t.exploration_fraction, t.initial_eps, t.final_eps = 0.25, 0.6, 0.05
    t.current_timestep = 0
    assert t.linear_eps(1000) == 0.6
    t.current_timestep = 125
    assert abs(t.linear_eps(1000) - 0.325) < 1e-12
    t.current_timestep = 250
    assert t.linear_eps(1000) == 0.05


def test_rfarl_import_shim():
    env = dict(os.environ, PYTHONPATH=os.path.join(ROOT, "distributional_rl_decision_and_control_b200", "compat"))
    code = ("from rfarl.envs.marinenav.env import MarineNavEnv3; from rfarl.agent import Agent; "
            "from rfarl.policy.trainer import Trainer; from rfarl.utils.replay_buffer import ReplayBuffer; "
            "from rfarl.policy.AC_IQN_model import Actor; print('ok')")
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, cwd="/tmp")
    assert r.returncode == 0 and "ok" in r.stdout, r.stderr


def test_training_agent_refuses_cpu():
    from distributional_rl_decision_and_control_b200.agent import Agent
    with pytest.raises(RuntimeError, match="CUDA"):
        Agent(device="cpu", agent_type="AC-IQN")


_DDP = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, {root!r})
from distributional_rl_decision_and_control_b200.agent import Agent
rank = int(os.environ["RANK"])
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:{port}", rank=rank, world_size=2)
agent = Agent(device="cpu", agent_type="AC-IQN", training=False, process_group=dist.group.WORLD, rank=rank)
torch.manual_seed(0)
params = [torch.nn.Parameter(torch.zeros(5, 3)), torch.nn.Parameter(torch.zeros(7))]
for i, p in enumerate(params):
    p.grad = torch.full_like(p, float(rank + 1) * (i + 1))
agent._sync_grads(params)
assert torch.allclose(params[0].grad, torch.full((5, 3), 1.5)), params[0].grad
assert torch.allclose(params[1].grad, torch.full((7,), 3.0)), params[1].grad
dist.destroy_process_group()
print("rank", rank, "ok")
"""


def test_gradient_exchange_gloo_world_size_2(tmp_path):
    """the one exchange step of the path is an all-reduce (average) of the flat
    gradient bucket; checked on CPU with gloo."""
    port = 29500 + os.getpid() % 2000
    script = tmp_path / "ddp.py"
    script.write_text(_DDP.format(root=ROOT, port=port))
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(os.environ, RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for r in range(2)]
    for p in procs:
        out, err = p.communicate(timeout=120)
        assert p.returncode == 0 and "ok" in out, err


def test_bench_reference_arm_contract():
    """`bench.py --impl reference`: one JSON line with the contract's keys, timed on the CPU port alone
    (no GPU needed); under torchrun every rank but 0 exits without work or output."""
    import json
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "3", "--warmup", "1",
           "--envs", "64"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=dict(os.environ, RANK="0", WORLD_SIZE="1"))
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "env_steps_per_sec" and d["unit"] == "env-steps/s"
    assert d["higher_is_better"] is True and d["steps"] == 3 and d["value"] > 0 and d["dtype"] == "f64"
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"]
    # the workload description is the one the b200 arm prints (same function): the driver's same_config check
    sys.path.insert(0, root)
    import bench
    assert d["config"] == bench.workload_config(64, 8, 1, True)
    assert d["config"]["auto_deactivate"] and d["config"]["auto_reset"]
    assert all(len(v) <= 128 for v in d["config"].values() if isinstance(v, str))  # the driver's record truncates longer strings
    assert cb["learner"]["c3_B1024_N32"]["updates_per_sec"] > 0 and cb["learner"]["ref_B64_N8"]["updates_per_sec"] > 0
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=60, env=dict(os.environ, RANK="1", WORLD_SIZE="2"))
    assert r.returncode == 0 and r.stdout.strip() == ""
