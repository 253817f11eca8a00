"""Export / deployment format: checkpoints written by this package load in the REFERENCE's
float64 export classes and trace as save_model_to_torch_script.py does (EXPORT:16-24), and the weights of the
reference's SHIPPED TorchScript policy (virelex/trained/traced_AC_IQN_model.pt, golden: tests/golden/traced_ac_iqn.npz,
made by oracle/gen_golden_export.py) give the shipped model's outputs in this package's actors — float64 twin, live
float32 module and, on the GPU, the fused K9 kernel."""
import importlib
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

import golden_util as gu
from distributional_rl_decision_and_control_b200.policy.AC_IQN_model import AC_IQN_Policy, Actor
from distributional_rl_decision_and_control_b200.policy.AC_IQN_model_dtype_double import Actor as ActorDouble

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/rfarl/rfarl"
DIMS = (7, 5, 5, 56, 40, 256, 128)
needs_ref = pytest.mark.skipif(not os.path.isdir(REF), reason="the reference tree is only in the build container")


def _golden():
    z = np.load(os.path.join(gu.GOLD, "traced_ac_iqn.npz"))
    return {k: z[k] for k in z.files}


def _logit(a):
    return np.tan(np.asarray(a, np.float64) * (np.pi / 2))


def _unit_inputs(n, seed=5):
    g = torch.Generator().manual_seed(seed)
    mask = (torch.rand((n, 5), generator=g) < 0.6).double()
    return (torch.rand((n, 7), dtype=torch.float64, generator=g),
            torch.rand((n, 5, 5), dtype=torch.float64, generator=g) * mask.unsqueeze(-1), mask)


def _shipped_actor(dtype):
    d = _golden()
    actor = Actor(*DIMS, [[-1.0, 1.0], [-1.0, 1.0]], "cpu", 0)
    actor.load_state_dict({k[2:]: torch.from_numpy(v) for k, v in d.items() if k.startswith("w.")})
    return actor.to(dtype), d


def test_shipped_policy_weights_reproduce_the_shipped_outputs():
    """The reference's trained, exported AC-IQN policy: same weights -> same actions, in float64 to 1e-12 and in
    the live float32 actor to 1e-6, compared on the logits so saturation cannot hide errors."""
    a64, d = _shipped_actor(torch.float64)
    x = tuple(torch.from_numpy(d[k]) for k in ("self", "objects", "mask"))
    with torch.no_grad():
        got64 = a64(x).numpy()
        a32 = a64.float()
        got32 = a32(tuple(t.float() for t in x)).numpy()
    assert np.abs(got64 - d["actions"]).max() <= 1e-12
    # float32 against the float64 export: 1e-6 holds for freshly initialised weights (next test); the TRAINED
    # policy has logits in the hundreds, where float32 itself (torch on the CPU) is ~2e-5 away
    assert np.abs(got32 - d["actions"]).max() <= 5e-5
    unsat = np.abs(d["actions"]) < 0.9
    assert unsat.mean() > 0.1
    rel = np.abs(_logit(got32) - _logit(d["actions"]))[unsat] / np.maximum(np.abs(_logit(d["actions"]))[unsat], 1.0)
    assert rel.max() <= 1e-4


def test_double_twin_loads_this_package_checkpoint_and_traces(tmp_path):
    pol = AC_IQN_Policy(*DIMS, [[-1.0, 1.0], [-1.0, 1.0]], "cpu", 11)
    pol.save(str(tmp_path))
    twin = ActorDouble.load(str(tmp_path), "cpu")
    assert all(p.dtype == torch.float64 for p in twin.parameters())
    g = torch.Generator().manual_seed(3)
    ex = ((torch.rand((1, 7), dtype=torch.float64, generator=g), torch.rand((1, 5, 5), dtype=torch.float64, generator=g),
           torch.rand((1, 5), dtype=torch.float64, generator=g)),)
    traced = torch.jit.trace(twin, ex)
    path = str(tmp_path / "traced_AC_IQN_model.pt")
    traced.save(path)
    loaded = torch.jit.load(path)
    d = _golden()
    x = tuple(torch.from_numpy(d[k]) for k in ("self", "objects", "mask"))
    u = _unit_inputs(64)
    with torch.no_grad():
        # traced float64 against the live float32 actor: 1e-6 on inputs of the export script's own kind
        # (torch.rand, EXPORT:19-22); full-range observations (thrust in the hundreds) cost float32 a few 1e-6
        assert (loaded(u) - pol.actor(tuple(t.float() for t in u)).double()).abs().max().item() <= 1e-6
        live32 = pol.actor(tuple(t.float() for t in x)).double()
        assert (loaded(x) - live32).abs().max().item() <= 5e-6
        assert (loaded(x) - twin(x)).abs().max().item() <= 1e-14


@needs_ref
def test_reference_export_classes_load_this_package_checkpoint(tmp_path):
    """EXPORT:16-24 with the reference's OWN class: Actor.load of rfarl/policy/AC_IQN_model_dtype_double.py on a
    checkpoint this package wrote, torch.jit.trace with the script's example input, outputs against the live
    float32 actor at 1e-6."""
    pol = AC_IQN_Policy(*DIMS, [[-1.0, 1.0], [-1.0, 1.0]], "cpu", 12)
    pol.save(str(tmp_path))
    sys.path.insert(0, REF)
    try:
        for m in [m for m in sys.modules if m == "policy" or m.startswith("policy.")]:
            del sys.modules[m]
        ref = importlib.import_module("policy.AC_IQN_model_dtype_double")
        model = ref.Actor.load(str(tmp_path), "cpu")
    finally:
        sys.path.remove(REF)
        for m in [m for m in sys.modules if m == "policy" or m.startswith("policy.")]:
            del sys.modules[m]
    example_input = ((torch.rand((1, 7), dtype=torch.float64), torch.rand((1, 5, 5), dtype=torch.float64),
                      torch.rand((1, 5), dtype=torch.float64)),)
    traced = torch.jit.trace(model, example_input)
    d = _golden()
    x = tuple(torch.from_numpy(d[k]) for k in ("self", "objects", "mask"))
    u = _unit_inputs(64)
    with torch.no_grad():
        want = traced(x)
        live32 = pol.actor(tuple(t.float() for t in x)).double()
        twin = ActorDouble.load(str(tmp_path), "cpu")(x)
        assert (traced(u) - pol.actor(tuple(t.float() for t in u)).double()).abs().max().item() <= 1e-6
    assert (want - live32).abs().max().item() <= 5e-6
    assert (want - twin).abs().max().item() <= 1e-13  # this package's float64 twin = the reference's, same weights


@needs_ref
def test_reference_export_script_runs_on_the_shim(tmp_path):
    """save_model_to_torch_script.py itself, unmodified except for its three path constants, with the `policy`
    package it imports resolved to this repo's shim: writes traced_AC_IQN_model.pt that the deployment loads."""
    pol = AC_IQN_Policy(*DIMS, [[-1.0, 1.0], [-1.0, 1.0]], "cpu", 13)
    pol.save(str(tmp_path))
    src = open(os.path.join(REF, "scripts", "save_model_to_torch_script.py")).read()
    src = (src.replace('model_type = "RL/agent/type"', 'model_type = "AC_IQN"')
              .replace('load_dir = "directory/to/your/trained/RL/model"', f'load_dir = {str(tmp_path)!r}')
              .replace('save_dir = "torch/script/output/directory"', f'save_dir = {str(tmp_path)!r}'))
    script = tmp_path / "export.py"
    script.write_text(src)
    shim = os.path.join(ROOT, "distributional_rl_decision_and_control_b200", "compat", "rfarl")
    r = subprocess.run([sys.executable, "-W", "ignore", str(script)], capture_output=True, text=True, cwd=str(tmp_path),
                       env=dict(os.environ, PYTHONPATH=shim + os.pathsep + ROOT))
    assert r.returncode == 0, r.stderr[-1500:]
    loaded = torch.jit.load(str(tmp_path / "traced_AC_IQN_model.pt"))
    x = (torch.rand((1, 7), dtype=torch.float64), torch.rand((1, 5, 5), dtype=torch.float64), torch.ones((1, 5), dtype=torch.float64))
    with torch.no_grad():
        out = loaded(x)
        assert out.shape == (1, 2) and out.dtype == torch.float64
        assert (out - pol.actor(tuple(t.float() for t in x)).double()).abs().max().item() <= 1e-6


def test_other_export_twins_have_the_traced_signatures(tmp_path):
    """IQN (state, taus) -> (1, 32, 25); DQN -> (1, 25); SAC -> (mu, log_std); DDPG -> (1, 2); Rainbow -> (1, 25, atoms):
    the call signatures EXPORT:26-69 traces, from float32 checkpoints of this package."""
    from distributional_rl_decision_and_control_b200.policy import (DDPG_model, DDPG_model_dtype_double, DQN_model,
                                                                    DQN_model_dtype_double, IQN_model,
                                                                    IQN_model_dtype_double, Rainbow_model,
                                                                    Rainbow_model_dtype_double, SAC_model,
                                                                    SAC_model_dtype_double)
    x = (torch.rand((1, 7), dtype=torch.float64), torch.rand((1, 5, 5), dtype=torch.float64), torch.ones((1, 5), dtype=torch.float64))
    x32 = tuple(t.float() for t in x)
    rng = [[-1.0, 1.0], [-1.0, 1.0]]
    d = tmp_path
    q = IQN_model.IQN_Policy(*DIMS, 25, "cpu", 1); q.save(str(d))
    iq = IQN_model_dtype_double.IQN_Policy.load(str(d), "cpu")
    taus = torch.rand((1, 32), dtype=torch.float64)
    tr = torch.jit.trace(iq, (x, taus))
    assert tr(x, taus).shape == (1, 32, 25) and tr(x, taus).dtype == torch.float64
    dq = DQN_model.DQN_Policy(*DIMS, 25, "cpu", 2); dq.save(str(d))
    dd = DQN_model_dtype_double.DQN_Policy.load(str(d), "cpu")
    with torch.no_grad():
        assert (torch.jit.trace(dd, (x,))(x) - dq(x32).double()).abs().max().item() <= 1e-5
    sac = SAC_model.SAC_Policy(*DIMS, rng, "cpu", 3); sac.save(str(d))
    sa = SAC_model_dtype_double.Actor.load(str(d), "cpu")
    mu, log_std = torch.jit.trace(sa, (x,))(x)
    assert mu.shape == (1, 2) and log_std.shape == (1, 2) and mu.dtype == torch.float64
    dp = DDPG_model.DDPG_Policy(*DIMS, rng, "cpu", 4); dp.save(str(d))
    da = DDPG_model_dtype_double.Actor.load(str(d), "cpu")
    with torch.no_grad():
        assert (torch.jit.trace(da, (x,))(x) - dp.actor(x32).double()).abs().max().item() <= 1e-6
    rb = Rainbow_model.Rainbow_Policy(*DIMS, 25, 51, "cpu", 5); rb.save(str(d))
    rd = Rainbow_model_dtype_double.Rainbow_Policy.load(str(d), "cpu")
    with torch.no_grad():
        out = rd(x)
    assert out.shape == (1, 25, 51) and out.dtype == torch.float64 and abs(float(out.sum()) - 25.0) < 1e-9


@pytest.mark.gpu
def test_k9_with_the_shipped_trained_policy():
    """The fused collection forward (K9, tcgen05) evaluated with the reference's TRAINED policy weights against
    the shipped float64 TorchScript outputs: actions to 1e-5, logits of the unsaturated rows to 1e-5 relative."""
    from distributional_rl_decision_and_control_b200 import _lib
    from distributional_rl_decision_and_control_b200.agent import Agent
    ag = Agent(device="cuda:0", seed=100, agent_type="AC-IQN", BATCH_SIZE=64, num_tau=8, BUFFER_SIZE=256)
    d = _golden()
    ag.policy_local.actor.load_state_dict({k[2:]: torch.from_numpy(v) for k, v in d.items() if k.startswith("w.")})
    reps = 40  # 3840 rows: several tiles per CTA
    x = tuple(torch.from_numpy(np.tile(d[k], (reps,) + (1,) * (d[k].ndim - 1))).float().cuda().contiguous()
              for k in ("self", "objects", "mask"))
    assert ag.policy_local.actor._fused_policy_ok(x)
    got = ag.act_batch(x, eps=0.0).cpu().numpy()
    torch.cuda.synchronize()
    _lib.policy_forward_status()
    want = np.tile(d["actions"], (reps, 1))
    # yardstick: the same trained network in plain float32 modules (torch, CPU) is ~2e-5 from the float64 export;
    # K9's operands carry 22 bits (two fp16 pieces) against float32's 24
    a32, _ = _shipped_actor(torch.float32)
    with torch.no_grad():
        e32 = np.abs(a32(tuple(torch.from_numpy(d[k]).float() for k in ("self", "objects", "mask"))).numpy() - d["actions"]).max()
    err = np.abs(got - want).max()
    assert err <= max(1e-5, 4.0 * e32), (err, e32)
    unsat = np.abs(want) < 0.9
    rel = np.abs(_logit(got) - _logit(want))[unsat] / np.maximum(np.abs(_logit(want))[unsat], 1.0)
    assert rel.max() <= 4e-4, rel.max()
