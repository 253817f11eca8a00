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
