"""CPU: the metrics oracle against hand-computed values, and the host-side ratio code against the oracle."""
import importlib.util
import os

import torch

from oracle.metrics_oracle import confusion_counts, reference_metrics

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_oracle_known_answer():
    # 1 x 3 classes x 2 x 2: predictions [[0, 1], [2, 2]], labels [[0, 2], [2, 1]]
    logits = torch.tensor([[[[5., 0.], [0., 0.]], [[0., 5.], [0., 0.]], [[0., 0.], [5., 5.]]]])
    labels = torch.tensor([[[0, 2], [2, 1]]])
    cm = confusion_counts(logits, labels, 3)
    assert cm.tolist() == [[1, 0, 0], [0, 0, 1], [0, 1, 1]]
    m = reference_metrics(logits, labels, 3, ignore_index=0)
    assert m["confusion_matrix"].tolist() == [[0, 0, 0], [0, 0, 1], [0, 0.5, 0.5]]
    assert abs(float(m["accuracy"]) - 0.5) < 1e-7
    # IoU per class: 1/1, 0/2, 1/3 -> macro 4/9
    assert abs(float(m["iou"]) - 4 / 9) < 1e-6


def test_host_ratios_match_oracle():
    import s2lc_b200  # noqa: F401
    from s2lc_b200.metrics import metrics_from_counts
    g = torch.Generator().manual_seed(0)
    logits = torch.randn(2, 5, 16, 16, generator=g)
    labels = torch.randint(0, 4, (2, 16, 16), generator=g)       # class 4 never a target
    counts = confusion_counts(logits, labels, 5)
    for ign in (None, 0, 3):
        got, want = metrics_from_counts(counts, ign), reference_metrics(logits, labels, 5, ign)
        for k in want:
            assert torch.allclose(got[k], want[k], atol=1e-6), (k, ign)


def test_input_oracle_known_answer():
    import numpy as np
    from oracle.input_oracle import prepare_sample, remap_labels
    tile = np.arange(2 * 3 * 4, dtype=np.int16).reshape(2, 3, 4) * 10
    lab = np.array([[0, 1, 2, 3], [4, 5, 6, 7], [8, 9, 10, 11]], dtype=np.uint8)
    names = {1: "a", 2: "b", 5: "a", 6: "zz"}
    rm = remap_labels(lab, names, ["other", "a", "b"])
    assert rm.tolist() == [[0, 1, 2, 0], [0, 1, 0, 0], [0, 0, 0, 0]]
    x, y = prepare_sample(tile, rm, 1, 1, 2, True, False, np.array([0.0, 0.1]), np.array([1.0, 2.0]), 10.0)
    # crop rows 1..2, cols 1..2 then horizontal flip
    assert y.tolist() == [[0, 1], [0, 0]]
    assert np.allclose(x[0], np.array([[60, 50], [100, 90]]) / 10.0)
    assert np.allclose(x[1], (np.array([[180, 170], [220, 210]]) - 1.0) / 20.0)


def test_label_lut_equals_vectorized_remap():
    import numpy as np
    import s2lc_b200  # noqa: F401
    from oracle.input_oracle import remap_labels
    from s2lc_b200.input_pipeline import build_label_lut
    names = {1: "urban", 2: "urban", 3: "crops", 4: "water", 7: "forest", 200: "crops"}
    new = ["other", "urban", "crops", "forest"]
    lut = build_label_lut(names, new).numpy()
    allv = np.arange(256, dtype=np.uint8)
    assert np.array_equal(lut[allv], remap_labels(allv, names, new))
