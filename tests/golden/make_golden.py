"""Generate the golden fixtures under tests/golden/ from the LIVE, UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

The reference is imported from a writable copy (importing it writes log files — SURVEY.md F2) with
a 4-line ``timm`` stub (F3).  Inputs/weights are the deterministic synthetic ones from
``s2lc_b200.synthetic``; outputs are stored as sub-sampled tensors plus whole-tensor moments
(sum, L2 norm, fixed projection) so the fixtures stay small.  Nothing at test/bench time reads
/root/reference.
"""
from __future__ import annotations

import json
import os
import shutil
import sys
import tempfile
import types
import zlib

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import s2lc_b200  # noqa: E402,F401
from s2lc_b200 import synthetic  # noqa: E402


def install_reference() -> str:
    ref = tempfile.mkdtemp(prefix="refcopy_")
    shutil.copytree("/root/reference", ref, dirs_exist_ok=True)
    sys.path[:0] = [ref, ref + "/src"]
    timm = types.ModuleType("timm")
    models = types.ModuleType("timm.models")
    layers = types.ModuleType("timm.models.layers")
    vt = types.ModuleType("timm.models.vision_transformer")
    layers.to_2tuple = lambda x: (x, x)
    vt.Block = type("Block", (torch.nn.Module,), {})
    sys.modules.update({"timm": timm, "timm.models": models, "timm.models.layers": layers,
                        "timm.models.vision_transformer": vt})
    return ref


def moments(t: torch.Tensor, key: str) -> list[float]:
    """[sum, l2, projection] in float64; the projection vector is closed-form (no RNG)."""
    v = t.detach().double().flatten()
    h = zlib.crc32(key.encode())
    i = torch.arange(v.numel(), dtype=torch.float64)
    r = torch.sin(i * (0.001 * (1 + h % 97)) + (h % 13))
    return [float(v.sum()), float(v.norm()), float((v * r).sum())]


def generalised_forward(model, x):
    """Tier B: the reference modules with the skip rule 'drop the map whose HxW equals
    the head's' instead of the literal (7, 7) of efficientnet_unet.py:259."""
    enc = model.encoder
    identity = x
    h = enc.stem(x)
    head_hw = (-(-x.shape[-2] // 32), -(-x.shape[-1] // 32))
    maps = []
    dcr = None
    for i, blk in enumerate(enc.blocks):
        if enc.drop_connect_rate is not None:
            dcr = enc.drop_connect_rate * (i / len(enc.blocks))
        h = blk(h, drop_connect_rate=dcr)
        if h.shape[-2:] not in [m.shape[-2:] for m in maps] and tuple(h.shape[-2:]) != head_hw:
            maps.insert(0, h)
    h = enc.conv_head(h)
    for up, dc, fm in zip(model.up_convs, model.double_convs, maps):
        h = dc(torch.cat([up(h), fm], dim=1))
    h = model.input_double_conv(torch.cat([model.input_up_conv(h), identity], dim=1))
    return model.out_conv1x1(h)


def main() -> None:
    install_reference()
    from modules.efficientnet_unet import EfficientNetConfig, EfficientnetUnet  # reference
    from src.losses import FocalLoss  # reference
    import modules.efficientnet_unet as ref_mod

    torch.set_num_threads(8)
    meta: dict = {"torch": torch.__version__, "cases": {}}
    arrays: dict[str, np.ndarray] = {}

    # ---- 1. state-dict key/shape contract for b0..b7
    shapes = {}
    for v in ["b0", "b1", "b2", "b3", "b4", "b5", "b6", "b7"]:
        m = EfficientnetUnet(EfficientNetConfig(version=v, in_channels=6, num_classes=4,
                                                class_distribution=list(synthetic.CLASS_DISTRIBUTION)))
        shapes[v] = {k: list(t.shape) for k, t in m.state_dict().items()}
    with open(os.path.join(HERE, "state_shapes.json"), "w") as f:
        json.dump(shapes, f)

    alpha = torch.tensor(synthetic.CLASS_DISTRIBUTION)
    alpha[1:] = 1 - alpha[1:]   # get_loss losses.py:27-29 with masked_loss=True

    def run_case(name, version, size, batch, tier, dc_rate=None, loss="focal"):
        cfg = EfficientNetConfig(version=version, in_channels=6, num_classes=4, drop_connect_rate=dc_rate,
                                 class_distribution=list(synthetic.CLASS_DISTRIBUTION))
        model = EfficientnetUnet(cfg)
        sd = synthetic.state_dict_like({k: tuple(s) for k, s in shapes[version].items()}, seed=0)
        model.load_state_dict(sd)
        model.train()
        x = synthetic.tiles(batch, size)
        y = synthetic.labels(batch, size)
        masks = None
        real_rand = torch.rand
        if dc_rate is not None:
            nb = len(model.encoder.blocks)
            rs = np.random.RandomState(77)
            u = rs.random_sample((nb, batch))
            queue = [torch.from_numpy(u[i].astype(np.float32)).view(batch, 1, 1, 1)
                     for i, b in enumerate(model.encoder.blocks)
                     if b.skip_connection and b.stride == 1 and b.input_filters == b.output_filters
                     and dc_rate * (i / nb)]
            ref_mod.torch.rand = lambda *a, **k: queue.pop(0)
        try:
            if tier == "A":
                logits = model(x)
                if dc_rate is None:
                    assert torch.equal(generalised_forward(model_clone(model, cfg, sd), x), logits), "tier B != tier A at 224"
            else:
                logits = generalised_forward(model, x)
        finally:
            ref_mod.torch.rand = real_rand
        if loss == "focal":
            lf = FocalLoss(alpha=alpha, gamma=2.0, label_smoothing=0.0, ignore_index=0)
        else:
            lf = torch.nn.CrossEntropyLoss(weight=alpha, label_smoothing=0.0, ignore_index=0)
        val = lf(logits, y)
        val.backward()
        case = {"version": version, "size": size, "batch": batch, "tier": tier, "dc_rate": dc_rate,
                "loss_type": loss, "loss": float(val), "logits_moments": moments(logits, "logits"),
                "grad_moments": {}, "buffer_moments": {}}
        arrays[f"{name}/logits_sub"] = logits.detach()[:, :, ::8, ::8].numpy().copy()
        for k, p in model.named_parameters():
            if p.grad is not None:
                case["grad_moments"][k] = moments(p.grad, k)
        for k, b in model.named_buffers():
            case["buffer_moments"][k] = moments(b, k)
        meta["cases"][name] = case
        print(name, "loss", float(val), flush=True)

    def model_clone(model, cfg, sd):
        c = EfficientnetUnet(cfg)
        c.load_state_dict(sd)
        c.train()
        return c

    run_case("b0_224_A", "b0", 224, 2, "A")
    run_case("b0_224_A_ce", "b0", 224, 2, "A", loss="ce")
    run_case("b0_224_A_dc", "b0", 224, 2, "A", dc_rate=0.2)
    run_case("b5_224_A", "b5", 224, 1, "A")
    run_case("b0_256_B", "b0", 256, 1, "B")
    run_case("b5_512_B", "b5", 512, 1, "B")
    run_case("b7_256_B", "b7", 256, 1, "B")

    # ---- loss kernels: full tensors on small random logits (incl. ignored pixels, bf16-range logits)
    rs = np.random.RandomState(5)
    lg = torch.from_numpy((3 * rs.standard_normal((2, 4, 32, 32))).astype(np.float32)).requires_grad_(True)
    yy = synthetic.labels(2, 32, seed=9, block=4)
    for nm, fn in [("focal_g2", FocalLoss(alpha=alpha, gamma=2.0, label_smoothing=0.0, ignore_index=0)),
                   ("focal_g0.5_noignore", FocalLoss(alpha=torch.ones(4), gamma=0.5, label_smoothing=0.0)),
                   ("focal_g2_sum", FocalLoss(alpha=alpha, gamma=2.0, label_smoothing=0.0, ignore_index=0, reduce_type="sum")),
                   ("ce_weighted", torch.nn.CrossEntropyLoss(weight=alpha, ignore_index=0)),
                   ("ce_plain", torch.nn.CrossEntropyLoss(ignore_index=-100))]:
        lg.grad = None
        v = fn(lg, yy)
        v.backward()
        arrays[f"loss/{nm}/value"] = np.asarray(float(v))
        arrays[f"loss/{nm}/dlogits"] = lg.grad.numpy().copy()
    arrays["loss/logits"] = lg.detach().numpy().copy()
    arrays["loss/y"] = yy.numpy().copy()

    # ---- short training curve: plain-PyTorch restatement of Lightning automatic optimisation around
    #      the reference modules (train_segmentation.py:109-114,129-147), fp32, Adam(lr=1e-3, wd=0.05)
    cfg = EfficientNetConfig(version="b0", in_channels=6, num_classes=4, drop_connect_rate=None,
                             class_distribution=list(synthetic.CLASS_DISTRIBUTION))
    model = EfficientnetUnet(cfg)
    model.load_state_dict(synthetic.state_dict_like({k: tuple(s) for k, s in shapes["b0"].items()}, seed=0))
    model.train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, weight_decay=0.05)
    lf = FocalLoss(alpha=alpha, gamma=2.0, label_smoothing=0.0, ignore_index=0)
    x, y = synthetic.tiles(2, 224), synthetic.labels(2, 224)
    curve = []
    for step in range(20):
        opt.zero_grad()
        val = lf(model(x), y)
        val.backward()
        opt.step()
        curve.append(float(val))
    meta["curve_b0_224_adam1e-3"] = curve
    print("curve", curve[:3], curve[-1])

    np.savez_compressed(os.path.join(HERE, "golden.npz"), **arrays)
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(meta, f)


if __name__ == "__main__":
    main()
