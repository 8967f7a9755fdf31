"""ORACLE (test infrastructure only — never imported by the package): the reference's per-sample input pipeline.

numpy restatement of src/data/s2osm_dataset.py:51-71 + src/data/s2osm_datamodule.py:75-87 +
src/configs/cnes_labell_mappings.py:85-95 for one sample with GIVEN random draws.  albumentations (requirements.txt:
``albumentations``, unpinned) is not installed here and is not vendored: crop / flip are index operations (exact), and
``A.Normalize`` is restated from its documented definition ``img = (img - mean * max_pixel_value) / (std *
max_pixel_value)`` as albumentations implements it in fp32 (subtract, then multiply by the fp32 reciprocal) —
"parity unpinned" against the library itself.
"""
from __future__ import annotations

import numpy as np


def remap_labels(labels: np.ndarray, source_to_name: dict, new_class_labels: list) -> np.ndarray:
    """cnes_labell_mappings.py:85-95, literally (np.vectorize of map_func)."""
    def map_func(label: int) -> int:
        target = source_to_name.get(label, "_")
        if label == 0 or target not in new_class_labels:
            return 0
        return new_class_labels.index(source_to_name[label])
    return np.vectorize(map_func)(labels).astype(int)


def prepare_sample(tile: np.ndarray, labels: np.ndarray, top: int, left: int, crop: int, hflip: bool, vflip: bool,
                   mean: np.ndarray, std: np.ndarray, max_pixel_value: float = 255.0):
    """tile int16 [C, H, W], labels [H, W] (already remapped) -> (x float32 [C, crop, crop], y int64 [crop, crop])."""
    img = np.transpose(tile, (1, 2, 0))                                  # "c h w -> h w c" (s2osm_dataset.py:60)
    img = img[top:top + crop, left:left + crop]                           # RandomCrop / CenterCrop
    msk = labels[top:top + crop, left:left + crop]
    if hflip:
        img, msk = img[:, ::-1], msk[:, ::-1]                             # HorizontalFlip
    if vflip:
        img, msk = img[::-1], msk[::-1]                                   # VerticalFlip
    m = (np.asarray(mean, dtype=np.float32) * np.float32(max_pixel_value)).astype(np.float32)
    d = np.reciprocal(np.asarray(std, dtype=np.float32) * np.float32(max_pixel_value), dtype=np.float32)
    out = img.astype(np.float32)
    out = (out - m) * d                                                   # A.Normalize (fp32)
    return np.ascontiguousarray(np.transpose(out, (2, 0, 1))), np.ascontiguousarray(msk).astype(np.int64)
