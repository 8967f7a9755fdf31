"""Small fixtures that pin host-side rows of SURVEY.md §8 to the LIVE reference (build container only):

    python tests/golden/make_pins.py        # -> tests/golden/pins.json

  * CNES label remap: `_cnes_transform` (src/configs/cnes_labell_mappings.py:85-95) applied to every uint8 value for
    the reference's four simplified label maps, plus the `CNES_TO_SIMPLIFIED` table itself (row f3);
  * `get_loss(config)` (src/losses.py:24-63): class of the returned loss, class weights, ignore_index, gamma, label
    smoothing for every combination of weighted / masked / loss type (row a12);
  * `initialize_classification_layer_bias` (src/utils.py:174-188): multi-class log-prior and the binary log-odds
    branch, and the exceptions it raises on bad distributions (row a11).
"""
from __future__ import annotations

import itertools
import json
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, HERE]
from make_golden import install_reference  # noqa: E402


def main() -> None:
    install_reference()
    import configs.cnes_labell_mappings as cn
    from src.losses import FocalLoss, LossType, get_loss
    from src.utils import initialize_classification_layer_bias

    pins: dict = {}
    maps = {"multiclass": cn.CNES_SIMPLIFIED_MULTICLASS, "binary_impervious": cn.CNES_SIMPLIFIED_BINARY_IMPERVIOUS,
            "binary_nature": cn.CNES_SIMPLIFIED_BINARY_NATURE, "binary_agriculture": cn.CNES_SIMPLIFIED_BINARY_AGRICULTURE}
    allv = np.arange(256)
    pins["cnes"] = {"to_simplified": {str(k): v for k, v in cn.CNES_TO_SIMPLIFIED.items()},
                    "maps": {name: {"classes": list(m.keys()),
                                    "remap_0_255": cn._cnes_transform(allv, label_map=m).tolist()} for name, m in maps.items()}}

    dist4, dist2 = [0.0, 0.4737, 0.3684, 0.1579], [0.7, 0.3]
    cases = []
    for weighted, masked, lt, dist in itertools.product([True, False], [True, False], ["ce", "focal"], [dist4, dist2]):
        cfg = types.SimpleNamespace(num_classes=len(dist), train=types.SimpleNamespace(
            weighted_loss=weighted, class_distribution=list(dist), masked_loss=masked, loss_type=LossType(lt),
            label_smoothing=0.05, focal_loss_gamma=1.5))
        fn = get_loss(cfg)
        if isinstance(fn, FocalLoss):
            rec = {"cls": "FocalLoss", "weights": fn.alpha.tolist(), "gamma": fn.gamma,
                   "label_smoothing": fn.label_smoothing, "ignore_index": fn.ignore_index, "reduce_type": fn.reduce_type}
        else:
            rec = {"cls": type(fn).__name__, "weights": None if fn.weight is None else fn.weight.tolist(),
                   "label_smoothing": fn.label_smoothing, "ignore_index": fn.ignore_index, "reduction": fn.reduction}
        cases.append({"weighted": weighted, "masked": masked, "loss_type": lt, "dist": dist, "out": rec})
    pins["get_loss"] = cases

    bias = {}
    for name, dist in [("multi", dist4), ("binary", dist2), ("multi_uniform", [0.25] * 4)]:
        layer = torch.nn.Conv2d(32, len(dist), 1)
        initialize_classification_layer_bias(layer, dist)
        bias[name] = {"dist": dist, "bias": layer.bias.detach().flatten().tolist(), "shape": list(layer.bias.shape)}
    errs = {}
    for name, dist in [("none", None), ("bad_sum", [0.5, 0.2, 0.2, 0.2]), ("single", [1.0])]:
        try:
            initialize_classification_layer_bias(torch.nn.Conv2d(32, 4, 1), dist)
            errs[name] = None
        except Exception as e:   # noqa: BLE001
            errs[name] = type(e).__name__
    pins["bias"] = {"ok": bias, "raises": errs}
    with open(os.path.join(HERE, "pins.json"), "w") as f:
        json.dump(pins, f)
    print(json.dumps(pins["bias"], indent=1)[:800])
    print(pins["get_loss"][0])


if __name__ == "__main__":
    main()
