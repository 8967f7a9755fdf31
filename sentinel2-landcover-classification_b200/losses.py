"""Focal / class-weighted cross-entropy losses with the reference's call signatures, computed by ONE fused
CUDA kernel (forward value + d loss / d logits in the same pass over the logits).

Mirrors ``src/losses.py`` of the reference: ``Loss`` (:13), ``LossType`` (:16-20), ``get_loss(config)`` (:24-63),
``FocalLoss`` dataclass (:69-89) and the weighted ``nn.CrossEntropyLoss`` (:35-40).  ``DiceLoss`` /
``CombinedLoss`` are out of scope (the reference's DiceLoss raises on any real shape — SURVEY.md §2).
"""
from __future__ import annotations

import enum
import typing
from dataclasses import dataclass, field

import torch

from . import _lib as L

Loss = typing.Callable[[torch.Tensor, torch.Tensor], torch.Tensor]
ReduceType = typing.Literal["mean", "sum"]


class LossType(str, enum.Enum):
    CE = "ce"
    FOCAL = "focal"
    DICE = "dice"
    DICE_FOCAL = "dice_focal"


class _FusedLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits: torch.Tensor, y: torch.Tensor, weights: torch.Tensor | None, gamma: float, ls: float,
                ignore_index: int, kind: int, reduce_mean: bool):
        if not logits.is_cuda:
            raise RuntimeError("s2lc losses run on sm_100a CUDA devices only (no CPU fallback)")
        lib = L.load()
        n, c, h, w = logits.shape
        if logits.dtype not in (torch.float32, torch.bfloat16):
            logits = logits.float()
        nhwc = logits.permute(0, 2, 3, 1)
        if not nhwc.is_contiguous():
            nhwc = nhwc.contiguous()
        y = y.contiguous()
        if y.dtype != torch.int64:
            y = y.long()
        dl = torch.empty_like(nhwc)
        out = torch.empty(2, dtype=torch.float32, device=logits.device)   # [loss, weight-sum scratch]
        st = torch.cuda.current_stream().cuda_stream
        dt = L.F32 if logits.dtype == torch.float32 else L.BF16
        L.check(lib.s2lc_loss_fwd_bwd(nhwc.data_ptr(), c, y.data_ptr(), L.ptr(weights), float(gamma), float(ls),
                                      int(ignore_index), kind, int(reduce_mean), out.data_ptr(),
                                      out.data_ptr() + 4, dl.data_ptr(), c, n * h * w, c, dt, st))
        ctx.dl, ctx.dt, ctx.shape = dl, dt, (n, c, h, w)
        return out[0]

    @staticmethod
    def backward(ctx, g: torch.Tensor):
        dl = ctx.dl
        n, c, h, w = ctx.shape
        lib = L.load()
        g32 = g.detach().to(torch.float32).contiguous()
        st = torch.cuda.current_stream().cuda_stream
        L.check(lib.s2lc_scale_inplace(dl.data_ptr(), c, g32.data_ptr(), n * h * w, c, ctx.dt, st))
        return dl.permute(0, 3, 1, 2), None, None, None, None, None, None, None


def _dev_weights(w: torch.Tensor | None, device: torch.device, cache: dict) -> torch.Tensor | None:
    if w is None:
        return None
    key = (device, w.data_ptr(), w._version)
    if key not in cache:
        cache.clear()
        cache[key] = w.detach().to(device=device, dtype=torch.float32).contiguous()
    return cache[key]


@dataclass
class FocalLoss:
    """alpha_y (1 - p_y)^gamma CE, mean over ALL pixels (ignored ones count in the denominator) — reference :69-89."""
    alpha: torch.Tensor  # (C,)
    gamma: float
    label_smoothing: float
    ignore_index: int = -100
    reduce_type: ReduceType = "mean"
    _cache: dict = field(default_factory=dict, repr=False, compare=False)

    def __call__(self, y_hat: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        if self.gamma is None:
            raise TypeError("focal_loss_gamma must be set (the reference default None raises too: SURVEY.md F4)")
        if self.reduce_type not in ("mean", "sum"):
            raise ValueError(f"Invalid reduction: {self.reduce_type}.")
        a = _dev_weights(self.alpha, y_hat.device, self._cache)
        return _FusedLoss.apply(y_hat, y, a, self.gamma, self.label_smoothing, self.ignore_index, L.LOSS_FOCAL,
                                self.reduce_type == "mean")


class CrossEntropyLoss:
    """``nn.CrossEntropyLoss(weight, label_smoothing, ignore_index)`` of reference :36-40 (weight-normalised mean)."""

    def __init__(self, weight: torch.Tensor | None = None, label_smoothing: float = 0.0, ignore_index: int = -100,
                 reduction: str = "mean") -> None:
        if reduction not in ("mean", "sum"):
            raise ValueError(f"Invalid reduction: {reduction}.")
        self.weight, self.label_smoothing, self.ignore_index, self.reduction = weight, label_smoothing, ignore_index, reduction
        self._cache: dict = {}

    def __call__(self, y_hat: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        w = _dev_weights(self.weight, y_hat.device, self._cache)
        return _FusedLoss.apply(y_hat, y, w, 0.0, self.label_smoothing, self.ignore_index, L.LOSS_CE,
                                self.reduction == "mean")


def get_loss(config) -> Loss:
    """Same factory as reference :24-63 over a config with ``.num_classes`` and ``.train.{weighted_loss,
    class_distribution, masked_loss, loss_type, label_smoothing, focal_loss_gamma}``."""
    class_weights = None
    if config.train.weighted_loss:
        class_weights = torch.tensor(config.train.class_distribution, dtype=torch.float32)
        first = int(config.train.masked_loss)
        class_weights[first:] = 1 - class_weights[first:]
    if class_weights is not None and len(class_weights) != config.num_classes:
        raise AssertionError(f"{len(class_weights)}!={config.num_classes}")
    ignore = 0 if config.train.masked_loss else -100
    lt = LossType(config.train.loss_type)
    if lt == LossType.CE:
        return CrossEntropyLoss(weight=class_weights, label_smoothing=config.train.label_smoothing, ignore_index=ignore)
    if lt == LossType.FOCAL:
        alpha = class_weights if class_weights is not None else torch.ones(config.num_classes)
        return FocalLoss(alpha=alpha, gamma=config.train.focal_loss_gamma,
                         label_smoothing=config.train.label_smoothing, ignore_index=ignore)
    if lt in (LossType.DICE, LossType.DICE_FOCAL):
        raise NotImplementedError(f"{lt.value}: out of scope — the reference's DiceLoss (src/losses.py:92-110) raises "
                                  "on every real input shape, so there is no behaviour to match")
    raise ValueError(f"Unknown loss type: {config.train.loss_type}.\nValid options: {list(LossType)}.")
