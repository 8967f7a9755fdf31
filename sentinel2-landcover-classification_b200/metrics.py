"""Per-step segmentation metrics on the device and uint8 class maps (row f4).

The reference updates four torchmetrics objects every step with ``logits.argmax(dim=1)`` and the labels
(src/train_segmentation.py:53-67, :145-151): ``MulticlassConfusionMatrix(num_classes, ignore_index=0 if masked_loss
else None, normalize="true")``, ``JaccardIndex`` ("IoU"), ``Accuracy`` and ``F1Score`` (multiclass task, library
defaults: IoU macro-averaged, accuracy/F1 micro-averaged; none of the last three is given an ignore_index).  All of
them are functions of the C x C confusion matrix, so here ONE libs2lc kernel pass over the logits accumulates
``counts[target, prediction]`` on the device (``s2lc_argmax_confusion``) and the four values are derived from it on
the host at ``compute()`` time.  torchmetrics is not installed in this image: the formulas below restate its
documented definitions and are pinned by ``oracle/metrics_oracle.py`` (plain torch ``argmax`` + ``bincount``).
"""
from __future__ import annotations

import torch

from . import _lib as L


def _nhwc_rows(logits: torch.Tensor) -> tuple[torch.Tensor, int, int]:
    """[B, C, H, W] logits (any strides) -> a [rows, C] channels-innermost view/copy, its row stride, dtype code."""
    if logits.dim() != 4:
        raise ValueError(f"expected [B, C, H, W] logits, got {tuple(logits.shape)}")
    if not logits.is_cuda:
        raise RuntimeError("s2lc metrics run on the GPU: logits must be a CUDA tensor (there is no CPU fallback)")
    if logits.dtype not in (torch.float32, torch.bfloat16):
        logits = logits.float()
    rows = logits.permute(0, 2, 3, 1)
    if not rows.is_contiguous():
        rows = rows.contiguous()
    return rows, logits.shape[1], (L.F32 if logits.dtype == torch.float32 else L.BF16)


def argmax_uint8(logits: torch.Tensor) -> torch.Tensor:
    """``logits.argmax(dim=1)`` as a uint8 [B, H, W] class map (the GeoTIFF-ready buffer of a mosaic writer, replacing
    the reference's ``torch.save`` of whole logit tensors in experiments/inference_demo.py:14-41)."""
    rows, c, dt = _nhwc_rows(logits)
    b, h, w = logits.shape[0], logits.shape[2], logits.shape[3]
    pred = torch.empty(b, h, w, dtype=torch.uint8, device=logits.device)
    L.check(L.load().s2lc_argmax_confusion(rows.data_ptr(), c, None, None, pred.data_ptr(), b * h * w, c, dt,
                                           torch.cuda.current_stream().cuda_stream))
    return pred


class SegmentationMetrics:
    """The reference's per-mode metric set (train_segmentation.py:53-67) over one accumulated confusion matrix."""

    def __init__(self, num_classes: int, ignore_index: int | None = None, device: torch.device | str = "cuda") -> None:
        if not 2 <= num_classes <= 16:
            raise ValueError("num_classes must be in 2..16")
        self.num_classes, self.ignore_index = num_classes, ignore_index
        self.counts = torch.zeros(num_classes, num_classes, dtype=torch.int64, device=device)
        self.lib = L.load()

    def reset(self) -> None:
        self.counts.zero_()

    def update(self, logits: torch.Tensor, labels: torch.Tensor, pred_out: torch.Tensor | None = None) -> None:
        """``metric.update(logits.argmax(dim=1), labels)`` for all four metrics at once (one kernel launch)."""
        rows, c, dt = _nhwc_rows(logits)
        if c != self.num_classes:
            raise ValueError(f"logits have {c} classes, metrics were built for {self.num_classes}")
        n = logits.shape[0] * logits.shape[2] * logits.shape[3]
        if labels.dtype != torch.int64 or labels.numel() != n or not labels.is_cuda:
            raise ValueError("labels must be a CUDA int64 [B, H, W] tensor matching the logits")
        labels = labels.contiguous()
        L.check(self.lib.s2lc_argmax_confusion(rows.data_ptr(), c, labels.data_ptr(), self.counts.data_ptr(),
                                               pred_out.data_ptr() if pred_out is not None else None, n, c, dt,
                                               torch.cuda.current_stream().cuda_stream))

    # ------------------------------------------------------------------ compute (host side, C x C numbers)
    def compute(self) -> dict[str, torch.Tensor]:
        return metrics_from_counts(self.counts, self.ignore_index)


def metrics_from_counts(counts: torch.Tensor, ignore_index: int | None) -> dict[str, torch.Tensor]:
    """counts[target, prediction] (int64) -> the reference's four metrics.

    * confusion_matrix: samples whose target is ``ignore_index`` are dropped, rows normalised to sum 1
      (``normalize="true"``; an empty row stays 0 as torchmetrics maps its NaN to 0);
    * iou: multiclass Jaccard, macro average over classes; a class that appears neither as target nor as prediction
      carries weight 0 (torchmetrics' macro rule), no ignore_index (the reference passes none);
    * accuracy, f1: micro average = trace / total (identical for single-label multiclass), no ignore_index."""
    cm = counts.to(torch.float64)
    cmi = cm.clone()
    if ignore_index is not None and 0 <= ignore_index < cm.shape[0]:
        cmi[ignore_index, :] = 0
    row = cmi.sum(1, keepdim=True)
    conf = torch.where(row > 0, cmi / row.clamp_min(1), torch.zeros_like(cmi))
    tp = cm.diag()
    fp = cm.sum(0) - tp
    fn = cm.sum(1) - tp
    denom = tp + fp + fn
    present = denom > 0
    iou_c = torch.where(present, tp / denom.clamp_min(1), torch.zeros_like(tp))
    iou = iou_c.sum() / present.sum().clamp_min(1)
    total = cm.sum()
    acc = tp.sum() / total.clamp_min(1)
    return {"confusion_matrix": conf.float(), "iou": iou.float(), "accuracy": acc.float(), "f1": acc.float()}
