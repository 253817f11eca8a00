"""Golden vectors of the export / deployment format, generated in the build container from
the reference's own artefacts — TEST INFRASTRUCTURE, never imported by the product.

    python oracle/gen_golden_export.py        ->  tests/golden/traced_ac_iqn.npz

  * the weights of the SHIPPED float64 TorchScript policy virelex/trained/traced_AC_IQN_model.pt (they are
    float32-representable: the float32 checkpoint cast by save_model_to_torch_script.py), stored as float32;
  * seeded, realistic observations (goal / velocities / thrusts / up to 5 objects, some rows without objects) and
    the shipped model's float64 outputs on them, batched and one row at a time (the (1,7),(1,5,5),(1,5) call of
    virelex/src/action_planner_node.cpp:309-313).
"""
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "traced_ac_iqn.npz")


def observations(n, seed):
    rng = np.random.RandomState(seed)
    f32 = lambda a: np.asarray(a, np.float32).astype(np.float64)
    # a third of the rows at full range (the trained policy is nearly bang-bang there), the rest scaled down so
    # that the outputs leave saturation and the comparison sees the logits
    scale = np.where(np.arange(n) % 3 == 0, 1.0, np.where(np.arange(n) % 3 == 1, 0.1, 0.03))[:, None]
    s = rng.randn(n, 7) * np.array([18.0, 18.0, 1.5, 0.4, 0.3, 450.0, 450.0]) * scale
    s[:, 5:] = np.clip(s[:, 5:] + 250.0 * scale, -500.0, 1000.0)
    cnt = rng.randint(0, 6, n)
    cnt[:4] = (0, 5, 1, 3)
    mask = (np.arange(5)[None, :] < cnt[:, None]).astype(np.float64)
    o = rng.randn(n, 5, 5) * np.array([12.0, 12.0, 1.2, 1.2, 0.0]) + np.array([0, 0, 0, 0, 1.0])
    o[..., 4] = np.where(rng.rand(n, 5) < 0.5, 1.0, 2.79508497)
    return f32(s), f32(o * mask[..., None]), mask


def main():
    m = torch.jit.load(os.path.join(REF, "virelex", "trained", "traced_AC_IQN_model.pt"))
    sd = {k: v.detach().numpy() for k, v in m.state_dict().items()}
    out = {}
    for k, v in sd.items():
        assert np.array_equal(v.astype(np.float32).astype(np.float64), v), k
        out["w." + k] = v.astype(np.float32)
    s, o, mask = observations(96, 20261104)
    t = lambda a: torch.from_numpy(a)
    with torch.no_grad():
        batched = m((t(s), t(o), t(mask))).numpy()
        single = np.concatenate([m((t(s[i:i + 1]), t(o[i:i + 1]), t(mask[i:i + 1]))).numpy() for i in range(len(s))])
    assert np.abs(batched - single).max() < 1e-13
    print("unsaturated share", (np.abs(batched) < 0.9).mean())
    assert (np.abs(batched) < 0.9).mean() > 0.1, "a part of the rows must not saturate the policy"
    out.update(self=s, objects=o, mask=mask, actions=single)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, {k: v.shape for k, v in out.items()}, "|a| mean", np.abs(batched).mean())


if __name__ == "__main__":
    sys.exit(main())
