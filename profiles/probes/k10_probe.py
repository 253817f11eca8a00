"""K10 alone: device time of the fused critic trunk forward (B x N quantile rows) in inference and SAVE mode against the
K3 + 3 x (cuBLAS sgemm + K5) sequence it replaces.  CUDA events around graph replays."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from distributional_rl_decision_and_control_b200 import _lib  # noqa: E402
from distributional_rl_decision_and_control_b200.policy.iqn_ops import critic_trunk  # noqa: E402
from test_gpu_critic import _critic, _inputs, _params  # noqa: E402
from k9_probe import timed  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    N = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    c = _critic()
    tau, feats, aenc = _inputs(B, N, 1)
    ps = _params(c)
    ws = _lib.critic_workspace("probe", feats.device)
    out = {"B": B, "N": N}
    out["k10_ms"] = timed(lambda: _lib.critic_forward(feats, aenc, *ps, N, ws, tau_in=tau))
    out["k10_nopack_ms"] = timed(lambda: _lib.critic_forward(feats, aenc, *ps, N, ws, tau_in=tau, repack=False))
    out["k10_save_ms"] = timed(lambda: _lib.critic_forward(feats, aenc, *ps, N, ws, tau_in=tau, save=True))
    torch.cuda.synchronize()
    _lib.critic_forward_status()

    def old():
        with torch.no_grad():
            _, cos = _lib.iqn_tau_cos(B, N, 64, feats.device, tau_in=tau)
            return critic_trunk(cos.reshape(B * N, 64), feats, aenc, ps[0], ps[1], ps[2], ps[3], ps[4], ps[5],
                                c.output_layer.weight, ps[7], N)
    out["k3_gemm_k5_ms"] = timed(old)
    M = B * N
    out["gflop_fp32_equiv"] = 2 * M * (64 * 256 + 256 * 128 + 128 * 128 + 128) / 1e9
    out["tflops_fp32_equiv"] = out["gflop_fp32_equiv"] / out["k10_ms"]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
