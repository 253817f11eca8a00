"""asvrl_tower_bwd_static — the allocation-free variant of K6's backward (scratch in a file-scope device array of
libasvrl.so; no counterpart in the reference) against asvrl_tower_bwd,
through the C-ABI, on two calls in a row.  Collected last on purpose."""
import ctypes as C

import pytest
import torch

from distributional_rl_decision_and_control_b200 import _lib

pytestmark = pytest.mark.gpu
SD, OD, K = 7, 5, 5


def _case(B, SF, OF, seed, dev):
    g = torch.Generator(device=dev).manual_seed(seed)
    r = lambda *s: torch.randn(s, device=dev, generator=g)
    W = SF + K * OF
    feats = torch.relu(r(B, W)).contiguous()      # zeros close ReLU gates / stand for masked objects
    return r(B, W).contiguous(), feats, r(B, SD).contiguous(), r(B, K, OD).contiguous()


def _static(dfeats, feats, so, ob, SF, OF):
    """Straight through the C-ABI: the caller owns inputs and outputs, the library owns the scratch."""
    dev = dfeats.device
    B = so.shape[0]
    new = lambda *s: torch.full(s, float("nan"), dtype=torch.float32, device=dev)
    dws, dbs, dwo, dbo = new(SF, SD), new(SF), new(OF, OD), new(OF)
    p = lambda t: C.c_void_p(t.data_ptr())
    rc = _lib.lib().asvrl_tower_bwd_static(p(dfeats), p(feats), p(so), p(ob), B, SD, OD, K, SF, OF, p(dws), p(dbs),
                                           p(dwo), p(dbo), C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    return rc, (dws, dbs, dwo, dbo)


def test_static_scratch_matches_caller_scratch_twice_in_a_row():
    dev = torch.device("cuda:0")
    # two calls in a row with different data and geometry: the second one meets the first one's scratch contents
    for B, SF, OF, seed in ((1024, 56, 40, 1), (333, 56, 40, 2), (64, 88, 40, 3), (1024, 56, 40, 1)):
        dfeats, feats, so, ob = _case(B, SF, OF, seed, dev)
        want = _lib.tower_bwd(dfeats, feats, so, ob, SF, OF)
        rc, got = _static(dfeats, feats, so, ob, SF, OF)
        assert rc == 0, _lib.lib().asvrl_last_error_string()
        torch.cuda.synchronize()
        for name, a, b in zip(("dw_self", "db_self", "dw_obj", "db_obj"), got, want):
            assert not bool(torch.isnan(a).any()), name
            scale = float(b.abs().max())
            assert float((a - b).abs().max()) <= 1e-6 * scale + 1e-12, f"{name} at B={B}"


def test_static_scratch_rejects_sizes_beyond_its_maximum():
    dev = torch.device("cuda:0")
    SF, OF = 100, 40   # 140 columns > 128
    dfeats, feats, so, ob = _case(32, SF, OF, 4, dev)
    rc, _ = _static(dfeats, feats, so, ob, SF, OF)
    assert rc == -2   # ASVRL_E_LIMIT
    msg = _lib.lib().asvrl_last_error_string().decode()
    assert "exceeds the static scratch" in msg and "128" in msg
    # the limit itself is accepted
    dfeats, feats, so, ob = _case(32, 88, 40, 5, dev)
    rc, got = _static(dfeats, feats, so, ob, 88, 40)
    assert rc == 0
    want = _lib.tower_bwd(dfeats, feats, so, ob, 88, 40)
    torch.cuda.synchronize()
    for a, b in zip(got, want):
        assert float((a - b).abs().max()) <= 1e-6 * float(b.abs().max()) + 1e-12
