"""ORACLE (test infrastructure only — never imported by the package): tiled prediction of a scene into a class mosaic.

numpy restatement of what the reference's predict path computes per pixel — ``predict_step`` returns ``self.net(x)``
(src/train_segmentation.py:106-107, :129-136) on Normalize'd tiles (src/data/s2osm_datamodule.py:75-87) and the class
map is ``argmax`` over the class axis — applied tile by tile with overlapping context.  The reference has no mosaic
writer of its own (experiments/inference_demo.py:14-41 saves per-batch logits), so the tiling convention (edge
replication, keep the tile centre) is this repo's; the per-pixel arithmetic is the reference's.  "parity unpinned"
for the tiling convention; the network itself is pinned by oracle/unet_oracle.py.
"""
from __future__ import annotations

import numpy as np


def tile_origins(height: int, width: int, tile: int, margin: int) -> np.ndarray:
    keep = tile - 2 * margin
    return np.array([(t - margin, l - margin) for t in range(0, height, keep) for l in range(0, width, keep)],
                    dtype=np.int32).reshape(-1, 2)


def gather_tile(scene: np.ndarray, top: int, left: int, tile: int, mean, std, max_pixel_value: float = 255.0):
    """int16 [C, H, W] -> float32 [C, tile, tile]; rows / columns outside the scene replicate the nearest edge."""
    _, h, w = scene.shape
    ys = np.clip(np.arange(top, top + tile), 0, h - 1)
    xs = np.clip(np.arange(left, left + tile), 0, w - 1)
    img = scene[:, ys][:, :, xs].astype(np.float32)
    m = (np.asarray(mean, dtype=np.float32) * np.float32(max_pixel_value)).astype(np.float32)
    d = np.reciprocal(np.asarray(std, dtype=np.float32) * np.float32(max_pixel_value), dtype=np.float32)
    return ((img - m[:, None, None]) * d[:, None, None]).astype(np.float32)


def write_tile(mosaic: np.ndarray, logits: np.ndarray, top: int, left: int, margin: int) -> None:
    """logits float [classes, tile, tile] -> argmax of the tile centre into the uint8 [H, W] mosaic (in place)."""
    tile = logits.shape[1]
    cls = np.argmax(logits, axis=0).astype(np.uint8)                     # first maximum wins, as torch.argmax
    h, w = mosaic.shape
    for ky in range(margin, tile - margin):
        my = top + ky
        if not 0 <= my < h:
            continue
        x0, x1 = max(left + margin, 0), min(left + tile - margin, w)
        if x1 > x0:
            mosaic[my, x0:x1] = cls[ky, x0 - left:x1 - left]


def predict_mosaic(scene: np.ndarray, tile: int, margin: int, mean, std, logits_fn, max_pixel_value: float = 255.0):
    """``logits_fn(x float32 [1, C, tile, tile]) -> [1, classes, tile, tile]`` is the network (oracle or otherwise)."""
    _, h, w = scene.shape
    mosaic = np.zeros((h, w), dtype=np.uint8)
    for top, left in tile_origins(h, w, tile, margin):
        x = gather_tile(scene, int(top), int(left), tile, mean, std, max_pixel_value)
        write_tile(mosaic, np.asarray(logits_fn(x[None]))[0], int(top), int(left), margin)
    return mosaic
