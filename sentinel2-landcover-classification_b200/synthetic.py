"""Deterministic synthetic Sentinel-2 tiles, labels and weights.

Everything is drawn from ``numpy.random.RandomState`` (a frozen, version-stable stream) so that the
golden fixtures made in the build container, the oracle and the CUDA path all see identical data on
any machine.  Nothing here touches the GPU path; it is input plumbing only.
"""
from __future__ import annotations

import zlib

import numpy as np
import torch

# class prior of the synthetic label maps and what get_class_probabilities(ignore_zero_label=True)
# (reference src/utils.py:152-171) would derive from it
LABEL_P = (0.05, 0.45, 0.35, 0.15)
CLASS_DISTRIBUTION = (0.0, 0.4737, 0.3684, 0.1579)


def tiles(batch: int, size: int, channels: int = 6, seed: int = 1234) -> torch.Tensor:
    """fp32 [B, C, H, W] i.i.d. N(0,1): the post-normalisation domain the net sees."""
    rs = np.random.RandomState(seed)
    return torch.from_numpy(rs.standard_normal((batch, channels, size, size)).astype(np.float32))


def labels(batch: int, size: int, num_classes: int = 4, seed: int = 1235, block: int = 16) -> torch.Tensor:
    """int64 [B, H, W]; one class per ``block``x``block`` patch (blocky like land cover), contains class 0."""
    rs = np.random.RandomState(seed)
    p = np.asarray(LABEL_P[:num_classes], dtype=np.float64)
    p = p / p.sum()
    nb = (size + block - 1) // block
    coarse = rs.choice(num_classes, size=(batch, nb, nb), p=p)
    full = np.repeat(np.repeat(coarse, block, axis=1), block, axis=2)[:, :size, :size]
    return torch.from_numpy(np.ascontiguousarray(full).astype(np.int64))


def state_dict_like(shapes: dict[str, tuple[int, ...]], seed: int = 0) -> dict[str, torch.Tensor]:
    """A full, non-trivial state dict for the given ``{key: shape}`` (reference key naming).

    Conv / ConvT weights ~ N(0, 2/fan_out) (same law as the reference's kaiming init,
    efficientnet_unet.py:401-405); every BatchNorm affine, bias and running buffer is perturbed away
    from its 1/0 default so that all gradient paths and the running-stat updates are exercised.
    """
    out: dict[str, torch.Tensor] = {}
    for key, shape in shapes.items():
        rs = np.random.RandomState((seed * 1000003 + zlib.crc32(key.encode())) & 0x7FFFFFFF)
        shape = tuple(shape)
        if key.endswith("num_batches_tracked"):
            out[key] = torch.zeros((), dtype=torch.int64)
            continue
        if key.endswith("running_mean"):
            v = 0.1 * rs.standard_normal(shape)
        elif key.endswith("running_var"):
            v = 1.0 + 0.2 * rs.random_sample(shape)
        elif len(shape) == 4:
            if "up_conv" in key:          # ConvTranspose2d weight [Cin, Cout, 2, 2]: fan_out = Cin*k*k (torch convention)
                fan_out = shape[0] * shape[2] * shape[3]
            else:
                fan_out = shape[0] * shape[2] * shape[3]
            v = rs.standard_normal(shape) * np.sqrt(2.0 / fan_out)
        elif len(shape) == 2:
            v = rs.standard_normal(shape) * np.sqrt(2.0 / (shape[0] + shape[1]))
        elif key.endswith(".weight"):     # BatchNorm gamma
            v = 1.0 + 0.1 * rs.standard_normal(shape)
        else:                             # biases / BatchNorm beta
            v = 0.1 * rs.standard_normal(shape)
        out[key] = torch.from_numpy(np.asarray(v, dtype=np.float32))
    return out


def drop_connect_masks(n_blocks: int, batch: int, rate: float | None, seed: int = 77):
    """(binary [n_blocks, B] fp32 in {0,1}, keep [n_blocks] fp32) with keep_i = 1 - rate * i / n_blocks and
    binary = floor(keep + U) (reference efficientnet_unet.py:257, :390-398).  Block 0 has rate 0 -> keep 1.
    Returns None when ``rate`` is None (drop-connect disabled, a legal config value)."""
    if rate is None:
        return None
    rs = np.random.RandomState(seed)
    u = rs.random_sample((n_blocks, batch)).astype(np.float32)
    keep = (1.0 - rate * (np.arange(n_blocks) / n_blocks)).astype(np.float32)
    return torch.from_numpy(np.floor(keep[:, None] + u)), torch.from_numpy(keep)


# per-class, per-band spectral signatures of the "learnable" synthetic tiles (rows: classes 0..3, columns: 6 bands)
_SIGNATURES = ((0.0, 0.0, 0.0, 0.0, 0.0, 0.0), (0.9, -0.6, 0.4, -0.8, 0.2, 0.7), (-0.7, 0.8, -0.5, 0.6, -0.9, 0.1),
               (0.3, 0.5, 0.9, -0.2, 0.8, -0.8))


def signal_tiles(lab: torch.Tensor, channels: int = 6, seed: int = 4321, noise: float = 0.6) -> torch.Tensor:
    """fp32 [B, C, H, W] tiles that CARRY the label: per-class band signature + N(0, noise^2).  Real Sentinel-2 tiles
    correlate with land cover; i.i.d. noise tiles do not, so a net trained on them can only memorise.  Used by the
    well-conditioned ("trained") parity fixture: after a few hundred Adam steps the logits separate."""
    rs = np.random.RandomState(seed)
    sig = torch.tensor(_SIGNATURES, dtype=torch.float32)[:, :channels]          # [classes, C]
    base = sig[lab].permute(0, 3, 1, 2).contiguous()                            # [B, C, H, W]
    return base + noise * torch.from_numpy(rs.standard_normal(tuple(base.shape)).astype(np.float32))
