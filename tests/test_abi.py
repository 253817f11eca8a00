"""CPU-only: the C-ABI library loads, exports every symbol include/asvrl.h declares, and its
host-side entry points agree with the ctypes mirror.  No kernel is launched."""
import ctypes as C
import os
import re

import pytest

from distributional_rl_decision_and_control_b200 import _abi, _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def so():
    if not os.path.exists(_lib.SO_PATH):
        from distributional_rl_decision_and_control_b200.csrc import build
        build.build()
    return _lib.lib()


def test_exports_every_declared_symbol(so):
    hdr = open(os.path.join(ROOT, "include", "asvrl.h")).read()
    declared = set(re.findall(r"ASVRL_API\s+[\w\s\*]+?\b(asvrl_\w+)\s*\(", hdr))
    assert len(declared) >= 11
    assert declared == set(_lib.EXPORTED)
    for name in declared:
        assert getattr(so, name) is not None


def test_version(so):
    assert so.asvrl_version() == _abi.VERSION


def test_default_params_match_mirror(so):
    p = _abi.EnvParams()
    so.asvrl_env_default_params(C.byref(p))
    q = _abi.default_params()
    assert bytes(p) == bytes(q)
    assert abs(p.robot_r - 2.7950849718747373) < 1e-15


def test_reset_defaults(so):
    r = _lib.default_reset_params()
    assert (r.map_w, r.map_h, r.clear_r, r.core_r, r.max_tries) == (55.0, 55.0, 10.0, 0.5, 500)
    assert (r.v_min, r.v_max, r.obs_r_min, r.obs_r_max, r.p, r.v_rel_max) == (3.0, 3.0, 1.0, 1.0, 0.8, 1.0)


def test_no_cpu_fallback(so):
    """Compute entry points refuse host tensors instead of silently computing on the CPU."""
    import torch
    with pytest.raises(RuntimeError, match="CUDA tensors only"):
        _lib.ptr(torch.zeros(4))
    z = torch.zeros(2, 8)
    with pytest.raises(RuntimeError):
        _lib.qhuber_fwd_bwd(z, z, z, torch.zeros(2), torch.zeros(2), 0.99, 1.0, torch.zeros(1, dtype=torch.int32))


def test_bad_argument_codes_without_gpu(so):
    rc = so.asvrl_qhuber_fwd_bwd(None, None, None, None, None, 0.99, 1.0, 1, 1, 1, None, None, None, None, None)
    assert rc == _abi.E_BADARG
    assert b"null pointer" in so.asvrl_last_error_string()


def test_bad_argument_codes_of_k5_k6_k7_without_gpu(so):
    """Argument validation happens on the host before any launch: checkable on a CPU-only box."""
    import ctypes as C
    fake = 0x1000  # never dereferenced: the calls below fail validation first
    assert so.asvrl_iqn_scale_relu_fwd(None, None, None, 1, 1, 128, None, None) == _abi.E_BADARG
    assert so.asvrl_iqn_scale_relu_fwd(fake, fake, fake, 4, 8, 130, fake, None) == _abi.E_BADARG  # dim % 4
    assert so.asvrl_iqn_scale_relu_bwd(fake, fake, fake, fake, 4, 8, 128, None, None, None, None, None, None) == _abi.E_BADARG
    assert so.asvrl_iqn_scale_relu_bwd(fake, fake, fake, fake, 4, 8, 128, None, None, fake, None, None, None) == _abi.E_BADARG
    assert so.asvrl_iqn_head_fwd(None, None, None, None, 4, 128, None, None) == _abi.E_BADARG
    assert so.asvrl_iqn_head_bwd(fake, fake, fake, fake, 4, 128, fake, fake, None, None, None, None, None) == _abi.E_BADARG
    assert so.asvrl_tower_fwd(None, None, None, None, None, None, None, 1, 7, 5, 5, 56, 40, None, None) == _abi.E_BADARG
    assert so.asvrl_tower_fwd(fake, fake, fake, fake, fake, fake, fake, 1, 9, 5, 5, 56, 40, fake, None) == _abi.E_LIMIT
    assert so.asvrl_iqn_scale_relu_bwd_partials(1024, 256) == 512 * 256
    assert so.asvrl_tower_bwd_partials(1024, 5, 56, 40) == 64 * 96 * 9
    t = _abi.AdamTensors()
    t.n = _abi.OPT_MAX_TENSORS + 1
    assert so.asvrl_clip_adam_step(C.byref(t), 0.5, 1e-4, 0.9, 0.999, 1e-8, 1.0, None, fake, fake, None) == _abi.E_LIMIT
    t.n = 1
    assert so.asvrl_clip_adam_step(C.byref(t), 0.5, 1e-4, 0.9, 0.999, 1e-8, 1.0, None, fake, fake, None) == _abi.E_BADARG
    assert so.asvrl_clip_adam_step(C.byref(t), -1.0, 1e-4, 0.9, 0.999, 1e-8, 1.0, None, fake, fake, None) == _abi.E_BADARG
    p = _abi.PeerBuckets()
    p.world, p.rank = 9, 0
    assert so.asvrl_clip_adam_step_p2p(C.byref(t), C.byref(p), fake, fake, 0.5, 1e-4, 0.9, 0.999, 1e-8, None,
                                       fake, fake, None) == _abi.E_BADARG
    assert so.asvrl_actor_head(None, None, None, None, 0.6366, None, 0.1, None, None, 1, 0, None, 8, 128, None, None) == _abi.E_BADARG
    assert so.asvrl_linear_tf32x3(None, None, None, 1, 128, 256, 128, None, None) == _abi.E_BADARG
    assert ctypes_sizes_match()


def ctypes_sizes_match():
    """The ctypes mirrors of the by-value tensor / peer lists have the C layout (LP64)."""
    import ctypes as C
    n = _abi.OPT_MAX_TENSORS
    return C.sizeof(_abi.AdamTensors) == 8 + 5 * 8 * n + 8 * n and C.sizeof(_abi.PeerBuckets) == 8 + 8 * _abi.MAX_PEERS + 8
