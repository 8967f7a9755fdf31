"""Host-side rows pinned to fixtures made from the LIVE reference (tests/golden/make_pins.py): the CNES label remap
(row f3), ``get_loss(config)`` (row a12) and the class-prior bias of ``out_conv1x1`` incl. the binary branch and the
reference's error behaviour (row a11).  CPU only."""
import json
import os
import types

import numpy as np
import pytest
import torch

from conftest import GOLDEN_DIR
from s2lc_b200.efficientnet_unet import EfficientNetConfig, EfficientnetUnet
from s2lc_b200.input_pipeline import build_label_lut
from s2lc_b200.losses import CrossEntropyLoss, FocalLoss, LossType, get_loss


@pytest.fixture(scope="module")
def pins():
    with open(os.path.join(GOLDEN_DIR, "pins.json")) as f:
        return json.load(f)


def test_label_lut_matches_reference_cnes_transform(pins):
    table = {int(k): v for k, v in pins["cnes"]["to_simplified"].items()}
    assert len(table) == 23
    for name, m in pins["cnes"]["maps"].items():
        lut = build_label_lut(table, m["classes"]).numpy()
        assert np.array_equal(lut, np.asarray(m["remap_0_255"])), name
        assert lut.max() == len(m["classes"]) - 1


def test_get_loss_matches_reference(pins):
    for case in pins["get_loss"]:
        cfg = types.SimpleNamespace(num_classes=len(case["dist"]), train=types.SimpleNamespace(
            weighted_loss=case["weighted"], class_distribution=list(case["dist"]), masked_loss=case["masked"],
            loss_type=LossType(case["loss_type"]), label_smoothing=0.05, focal_loss_gamma=1.5))
        fn, want = get_loss(cfg), case["out"]
        if want["cls"] == "FocalLoss":
            assert isinstance(fn, FocalLoss)
            assert fn.gamma == want["gamma"] and fn.reduce_type == want["reduce_type"]
            w = fn.alpha
        else:
            assert isinstance(fn, CrossEntropyLoss) and want["cls"] == "CrossEntropyLoss"
            assert fn.reduction == want["reduction"]
            w = fn.weight
        assert fn.ignore_index == want["ignore_index"] and fn.label_smoothing == want["label_smoothing"]
        if want["weights"] is None:
            assert w is None
        else:
            assert torch.allclose(w, torch.tensor(want["weights"]), rtol=0, atol=0), case
    # string loss types are accepted like the enum (LossType is a str enum), unknown ones raise like the reference
    cfg.train.loss_type = "focal"
    assert isinstance(get_loss(cfg), FocalLoss)
    cfg.train.loss_type = "nope"
    with pytest.raises(ValueError):
        get_loss(cfg)
    cfg.train.loss_type, cfg.num_classes, cfg.train.weighted_loss = "ce", 7, True
    with pytest.raises(AssertionError):
        get_loss(cfg)


def test_class_prior_bias_matches_reference(pins):
    for name, rec in pins["bias"]["ok"].items():
        m = EfficientnetUnet(EfficientNetConfig(version="b0", in_channels=6, num_classes=len(rec["dist"]),
                                                class_distribution=list(rec["dist"])))
        b = m.out_conv1x1.bias.detach()
        assert list(b.shape) == rec["shape"], name
        assert torch.equal(b, torch.tensor(rec["bias"])), name
    bad = {"none": None, "bad_sum": [0.5, 0.2, 0.2, 0.2], "single": [1.0]}
    for name, exc in pins["bias"]["raises"].items():
        with pytest.raises({"TypeError": TypeError, "AssertionError": AssertionError}[exc]):
            EfficientnetUnet(EfficientNetConfig(version="b0", in_channels=6, num_classes=4, class_distribution=bad[name]))


def test_concat_input_false_is_declared_unbuilt():
    with pytest.raises(NotImplementedError):
        EfficientnetUnet(EfficientNetConfig(version="b0", in_channels=6, num_classes=4,
                                            class_distribution=[0.25] * 4), concat_input=False)
