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
