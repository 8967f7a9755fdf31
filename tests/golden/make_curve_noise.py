"""Add 100-step training curves of the LIVE reference (fp32 and fp64) to tests/golden/golden.json.

    python tests/golden/make_curve_noise.py          # build container only (needs /root/reference)

The fp64 run is the yardstick for how far ANY correct fp32 implementation may drift from the reference's fp32
curve: Adam's sign-like updates amplify rounding-level differences, so `|ref32 - ref64|` per step is the
reference's own numerical noise and the parity test (tests/test_gpu_model.py) bounds ours by a multiple of it.
A third fixture is the per-step maximum deviation of six fp32 runs whose weights were perturbed by 1e-6
(relative): the spread a rounding-level change produces in the reference itself.  golden.npz is not touched.
"""
from __future__ import annotations

import json
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as G  # noqa: E402
from s2lc_b200 import synthetic  # noqa: E402

STEPS = 100


def curve(dtype: torch.dtype, shapes, perturb_seed: int | None = None, steps: int = STEPS) -> list[float]:
    from modules.efficientnet_unet import EfficientNetConfig, EfficientnetUnet
    from src.losses import FocalLoss
    cfg = EfficientNetConfig(version="b0", in_channels=6, num_classes=4, drop_connect_rate=None,
                             class_distribution=list(synthetic.CLASS_DISTRIBUTION))
    model = EfficientnetUnet(cfg)
    model.load_state_dict(synthetic.state_dict_like({k: tuple(s) for k, s in shapes["b0"].items()}, seed=0))
    if perturb_seed is not None:   # rounding-level (1e-6 relative) perturbation of every weight
        g = torch.Generator().manual_seed(perturb_seed)
        with torch.no_grad():
            for p in model.parameters():
                p.mul_(1 + 1e-6 * torch.randn(p.shape, generator=g))
    model = model.to(dtype).train()
    alpha = torch.tensor(synthetic.CLASS_DISTRIBUTION)
    alpha[1:] = 1 - alpha[1:]
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, weight_decay=0.05)
    lf = FocalLoss(alpha=alpha.to(dtype), gamma=2.0, label_smoothing=0.0, ignore_index=0)
    x, y = synthetic.tiles(2, 224).to(dtype), synthetic.labels(2, 224)
    out = []
    for _ in range(steps):
        opt.zero_grad()
        val = lf(model(x), y)
        val.backward()
        opt.step()
        out.append(float(val.detach()))
    return out


def main() -> None:
    G.install_reference()
    with open(os.path.join(HERE, "state_shapes.json")) as f:
        shapes = json.load(f)
    with open(os.path.join(HERE, "golden.json")) as f:
        meta = json.load(f)
    c32, c64 = curve(torch.float32, shapes), curve(torch.float64, shapes)
    assert all(abs(a - b) < 1e-6 for a, b in zip(c32[:20], meta["curve_b0_224_adam1e-3"])), "fp32 curve changed"
    meta["curve100_b0_224_adam1e-3_f32"] = c32
    meta["curve100_b0_224_adam1e-3_f64"] = c64
    # envelope of the reference's OWN sensitivity: fp32 runs whose weights were perturbed by 1e-6 (relative)
    pert = [curve(torch.float32, shapes, perturb_seed=100 + i, steps=50) for i in range(6)]
    meta["curve50_b0_224_adam1e-3_perturbed_maxdev"] = [
        max(abs(c[s] - c32[s]) / abs(c32[s]) for c in pert) for s in range(50)]
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(meta, f)
    print("perturbed max dev per step:", [round(v, 4) for v in meta["curve50_b0_224_adam1e-3_perturbed_maxdev"]])
    dev = [abs(a - b) / abs(b) for a, b in zip(c32, c64)]
    print("max rel |f32-f64| first 20:", max(dev[:20]), "all:", max(dev))


if __name__ == "__main__":
    main()
