"""Well-conditioned fixture from the LIVE, UNMODIFIED reference (build container only; needs /root/reference):

    python tests/golden/make_trained.py

Runs the recipe of tests/_trained.py (b0, 224x224, bs 8, Adam lr 1e-3, no decay, focal loss, 300 steps on
label-carrying synthetic tiles) on the reference's own ``EfficientnetUnet`` + ``FocalLoss`` and commits
``tests/golden/trained.json``: the loss curve, moments of every tensor of the trained state, and — at the trained
state, on a fresh batch — loss, sub-sampled logits, per-tensor gradient moments and the reference's OWN fp32-vs-fp64
gradient error (so a test can tell which tensors the 1e-4 bar is meaningful for).  Nothing at test time reads
/root/reference.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), HERE]
import s2lc_b200  # noqa: E402,F401
import _trained as T  # noqa: E402
from make_golden import install_reference, moments  # noqa: E402
from s2lc_b200 import synthetic  # noqa: E402


def main() -> None:
    install_reference()
    from modules.efficientnet_unet import EfficientNetConfig, EfficientnetUnet  # reference
    from src.losses import FocalLoss  # reference

    torch.set_num_threads(8)
    cfg = EfficientNetConfig(version=T.VERSION, in_channels=6, num_classes=4, drop_connect_rate=None,
                             class_distribution=list(synthetic.CLASS_DISTRIBUTION))
    model = EfficientnetUnet(cfg)
    model.load_state_dict(T.initial_state())
    model.train()
    # the reference's optimizer (train_segmentation.py:110-114) hands ALL parameters to Adam; encoder.fc.* get no
    # gradient and are skipped by torch, so the update set equals the oracle's used_param_keys
    opt = torch.optim.Adam(model.parameters(), lr=T.LR, weight_decay=T.WD)
    lf = FocalLoss(alpha=T.alpha(), gamma=2.0, label_smoothing=0.0, ignore_index=0)
    x, y = T.train_batch()
    curve = []
    for step in range(T.STEPS):
        opt.zero_grad()
        val = lf(model(x), y)
        val.backward()
        opt.step()
        curve.append(float(val.detach()))
        if step % 20 == 0:
            print(step, curve[-1], flush=True)
    meta = {"torch": torch.__version__, "recipe": {"version": T.VERSION, "size": T.SIZE, "batch": T.BATCH,
                                                   "steps": T.STEPS, "lr": T.LR, "weight_decay": T.WD},
            "curve": curve, "state_moments": {k: moments(v, k) for k, v in model.state_dict().items()
                                              if v.is_floating_point()}}
    # ---- evaluation at the trained state on a fresh batch (training-mode BatchNorm, as the step runs it)
    xe, ye = T.eval_batch()
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}

    def grads_of(dtype):
        m = EfficientnetUnet(cfg)
        m.load_state_dict(sd)
        m = m.to(dtype).train()
        l2 = FocalLoss(alpha=T.alpha().to(dtype), gamma=2.0, label_smoothing=0.0, ignore_index=0)
        logits = m(xe.to(dtype))
        v = l2(logits, ye)
        v.backward()
        return logits.detach(), float(v.detach()), {k: p.grad.detach() for k, p in m.named_parameters() if p.grad is not None}

    lg32, loss32, g32 = grads_of(torch.float32)
    lg64, loss64, g64 = grads_of(torch.float64)
    ev = {"loss": loss32, "loss_f64": loss64, "logits_moments": moments(lg32, "logits"), "grad_moments": {},
          "ref_f32_vs_f64": {}}
    mx = max(float(g.norm()) for g in g64.values())
    for k in g32:
        ev["grad_moments"][k] = moments(g32[k], k)
        ev["ref_f32_vs_f64"][k] = float((g32[k].double() - g64[k]).norm() / max(float(g64[k].norm()), 1e-6 * mx))
    ev["logits_f32_vs_f64"] = float((lg32.double() - lg64).norm() / lg64.norm())
    am = lg32.argmax(1)
    ev["argmax_accuracy_vs_labels"] = float((am == ye)[ye != 0].float().mean())
    top2 = lg32.topk(2, dim=1).values
    ev["median_top2_margin"] = float((top2[:, 0] - top2[:, 1]).median())
    meta["eval"] = ev
    np.savez_compressed(os.path.join(HERE, "trained_eval.npz"), logits_sub=lg32[:, :, ::8, ::8].numpy().copy())
    with open(os.path.join(HERE, "trained.json"), "w") as f:
        json.dump(meta, f)
    errs = sorted(ev["ref_f32_vs_f64"].values())
    print("ref f32-vs-f64: share <= 1e-5", sum(e <= 1e-5 for e in errs) / len(errs), "<= 1e-4",
          sum(e <= 1e-4 for e in errs) / len(errs), "<= 1e-3", sum(e <= 1e-3 for e in errs) / len(errs),
          "zero-norm grads", sum(m[1] == 0 for m in ev["grad_moments"].values()))
    print("final loss", curve[-1], "eval loss", loss32, "acc", ev["argmax_accuracy_vs_labels"], "margin",
          ev["median_top2_margin"], "ref f32-vs-f64 grad err median", errs[len(errs) // 2], "max", errs[-1])


if __name__ == "__main__":
    main()
