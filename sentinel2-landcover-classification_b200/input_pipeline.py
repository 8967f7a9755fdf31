"""The input pipeline on the device.

The reference prepares every sample on the CPU inside DataLoader workers: read the int16 tile and the first label band
(src/data/s2osm_dataset.py:51-56), remap CNES labels with ``np.vectorize`` (src/configs/cnes_labell_mappings.py:85-95),
albumentations ``RandomCrop | CenterCrop -> HorizontalFlip -> VerticalFlip -> Normalize(mean, std)``
(src/data/s2osm_datamodule.py:75-87), ``.float()`` / ``.long()`` (:65-70).  Here the raw rasters are uploaded once
(int16 + uint8: 13 B per source pixel instead of 32 B of fp32 + int64 per output pixel) and ONE libs2lc kernel does
crop + flips + normalisation + label remap straight into the fp32 NCHW / int64 tensors the network and the loss take.
The random draws stay on the host (``draw_augmentation``) so the caller's RNG policy is reproduced exactly.
rasterio / albumentations are not installed here: file IO is out of scope, and the arithmetic is pinned by
``oracle/input_oracle.py`` (numpy restatement of albumentations' documented behaviour — "parity unpinned" against
albumentations itself).
"""
from __future__ import annotations

import torch

from . import _lib as L


def build_label_lut(source_to_name: dict[int, str], new_class_labels: list[str]) -> torch.Tensor:
    """256-entry table equivalent to the reference's ``map_func`` (cnes_labell_mappings.py:89-93): label 0 and labels
    whose simplified name is not in ``new_class_labels`` map to 0, the rest to the index of their name."""
    lut = torch.zeros(256, dtype=torch.int64)
    for label in range(1, 256):
        target = source_to_name.get(label, "_")
        if target in new_class_labels:
            lut[label] = new_class_labels.index(target)
    return lut


def draw_augmentation(n: int, src_hw: tuple[int, int], crop: int, augment: bool, p_hflip: float = 0.5,
                      p_vflip: float = 0.5, generator: torch.Generator | None = None):
    """Per-sample crop offsets and flip flags: random crop + flips when ``augment`` (train), centre crop and no flips
    otherwise (val/test) — the two transform lists of s2osm_datamodule.py:75-87."""
    h, w = src_hw
    if augment:
        top = torch.randint(0, h - crop + 1, (n,), generator=generator)
        left = torch.randint(0, w - crop + 1, (n,), generator=generator)
        flips = torch.stack([torch.rand(n, generator=generator) < p_hflip,
                             torch.rand(n, generator=generator) < p_vflip], 1).to(torch.uint8)
    else:
        top = torch.full((n,), (h - crop) // 2)
        left = torch.full((n,), (w - crop) // 2)
        flips = torch.zeros(n, 2, dtype=torch.uint8)
    return torch.stack([top, left], 1).to(torch.int32), flips


class DeviceBatchPreparer:
    """crop -> flips -> Normalize -> label remap for a whole batch in one kernel launch."""

    def __init__(self, mean, std, crop: int, label_lut: torch.Tensor | None = None, max_pixel_value: float = 255.0,
                 device: torch.device | str = "cuda") -> None:
        self.mean = torch.as_tensor(mean, dtype=torch.float32).to(device).contiguous()
        self.std = torch.as_tensor(std, dtype=torch.float32).to(device).contiguous()
        self.crop, self.mpv, self.device = int(crop), float(max_pixel_value), torch.device(device)
        self.lut = None
        if label_lut is not None:
            if label_lut.numel() != 256:
                raise ValueError("label_lut must have 256 entries (uint8 label rasters)")
            self.lut = label_lut.to(torch.int64).to(device).contiguous()
        self.lib = L.load()

    def __call__(self, raw: torch.Tensor, labels: torch.Tensor, crop_yx: torch.Tensor, flips: torch.Tensor | None = None):
        if raw.dtype != torch.int16 or raw.dim() != 4 or not raw.is_cuda:
            raise ValueError("raw tiles must be a CUDA int16 [N, C, H, W] tensor")
        n, c, hs, ws = raw.shape
        if labels.dtype != torch.uint8 or tuple(labels.shape) != (n, hs, ws) or not labels.is_cuda:
            raise ValueError("labels must be a CUDA uint8 [N, H, W] tensor matching the tiles")
        if c != self.mean.numel() or c != self.std.numel():
            raise ValueError(f"{c} bands but {self.mean.numel()} means / {self.std.numel()} stds")
        cy = crop_yx.to(torch.int32)
        if tuple(cy.shape) != (n, 2) or int(cy.min()) < 0 or int(cy[:, 0].max()) > hs - self.crop \
                or int(cy[:, 1].max()) > ws - self.crop:
            raise ValueError("crop offsets out of range")
        cy = cy.to(self.device).contiguous()
        fl = flips.to(torch.uint8).to(self.device).contiguous() if flips is not None else None
        raw, labels = raw.contiguous(), labels.contiguous()
        x = torch.empty(n, c, self.crop, self.crop, dtype=torch.float32, device=self.device)
        y = torch.empty(n, self.crop, self.crop, dtype=torch.int64, device=self.device)
        L.check(self.lib.s2lc_prepare_batch(raw.data_ptr(), labels.data_ptr(), cy.data_ptr(), L.ptr(fl),
                                            self.mean.data_ptr(), self.std.data_ptr(), L.ptr(self.lut), x.data_ptr(),
                                            y.data_ptr(), n, c, hs, ws, self.crop, self.mpv,
                                            torch.cuda.current_stream().cuda_stream))
        return x, y
