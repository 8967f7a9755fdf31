"""ORACLE (test infrastructure only — never imported by the package): per-step segmentation metrics.

Restates what the reference computes every step (src/train_segmentation.py:53-67, :145-151):
``predictions = logits.argmax(dim=1)`` fed to torchmetrics ``MulticlassConfusionMatrix(num_classes, ignore_index,
normalize="true")``, ``JaccardIndex`` (macro), ``Accuracy`` (micro) and ``F1Score`` (micro).  torchmetrics
(requirements.txt: ``torchmetrics``, no pinned version) is NOT installed in this image and is not vendored under
/root/reference, so this oracle restates its published definitions with plain torch ops; the integer part (argmax +
confusion counts) is exact by construction, the derived ratios are "parity unpinned" against torchmetrics itself.
"""
from __future__ import annotations

import torch


def confusion_counts(logits: torch.Tensor, labels: torch.Tensor, num_classes: int) -> torch.Tensor:
    """counts[t, p] = #pixels with label t (in range) and argmax p.  logits [B, C, H, W], labels int64 [B, H, W]."""
    pred = logits.float().argmax(dim=1).reshape(-1)
    tgt = labels.reshape(-1)
    ok = (tgt >= 0) & (tgt < num_classes)
    idx = tgt[ok] * num_classes + pred[ok]
    return torch.bincount(idx, minlength=num_classes * num_classes).reshape(num_classes, num_classes)


def reference_metrics(logits: torch.Tensor, labels: torch.Tensor, num_classes: int, ignore_index: int | None):
    counts = confusion_counts(logits, labels, num_classes).double()
    cmi = counts.clone()
    if ignore_index is not None:
        cmi[ignore_index] = 0                       # torchmetrics drops samples whose target is ignore_index
    row = cmi.sum(1, keepdim=True)
    conf = torch.nan_to_num(cmi / row, nan=0.0)     # normalize="true"
    tp = counts.diag()
    union = counts.sum(0) + counts.sum(1) - tp
    present = union > 0
    iou = (tp[present] / union[present]).sum() / max(int(present.sum()), 1)   # macro over classes that occur
    acc = tp.sum() / counts.sum().clamp_min(1)      # micro accuracy == micro F1 for single-label multiclass
    return {"confusion_matrix": conf.float(), "iou": iou.float(), "accuracy": acc.float(), "f1": acc.float()}
