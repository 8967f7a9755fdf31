"""Mosaic inference writer.

The reference predicts through Lightning: ``trainer.predict`` calls ``predict_step`` (src/train_segmentation.py:106-107,
``_model_step(mode="test")`` -> raw logits, :129-136) on the validation DataLoader and a ``BasePredictionWriter``
``torch.save``s every batch of fp32 logits (src/experiments/inference_demo.py:14-41); the class map is taken later, on
the CPU, when plotting.  For an AOI that is C x 4 B written and re-read per pixel to end up with one byte.

Here the whole scene stays resident in HBM as the int16 raster it was read as, and a prediction is three launches per
batch of tiles, none of which leaves the device:

  ``s2lc_gather_tiles``   cut B overlapping tiles out of the scene (edge-replicated outside), Normalize -> fp32 NCHW
  network forward          eval mode: running BN statistics, no drop-connect (the engine's inference workspace)
  ``s2lc_argmax_mosaic``  argmax over classes of each tile's CENTRE written as uint8 straight into the scene mosaic

Tiles are strided over ranks (one process per GPU, no exchange while predicting); the per-rank mosaics are disjoint,
so one uint8 SUM all-reduce at the end assembles the scene on every rank.  GeoTIFF encoding (rasterio) is file IO and
out of scope: the product is the headless uint8 [H, W] buffer (``write_npy`` dumps it for inspection).
"""
from __future__ import annotations

import torch

from . import _lib as L
from .metrics import _nhwc_rows


def plan_mosaic(height: int, width: int, tile: int, margin: int = 0) -> torch.Tensor:
    """Tile origins (top, left) int32 [T, 2], row-major.  Each tile keeps its centre of ``tile - 2*margin`` px; the
    kept squares tile [0, height) x [0, width) exactly once, so the first origin is (-margin, -margin) and the last
    row / column of tiles may overhang the scene (the writer drops what falls outside)."""
    if tile <= 0 or tile % 32:
        raise ValueError("tile must be a positive multiple of 32 (the network has 5 stride-2 stages)")
    if margin < 0 or 2 * margin >= tile:
        raise ValueError("margin must satisfy 0 <= 2*margin < tile")
    if height <= 0 or width <= 0:
        raise ValueError("empty scene")
    keep = tile - 2 * margin
    tops = torch.arange(0, height, keep, dtype=torch.int32) - margin
    lefts = torch.arange(0, width, keep, dtype=torch.int32) - margin
    return torch.stack(torch.meshgrid(tops, lefts, indexing="ij"), -1).reshape(-1, 2).contiguous()


def rank_batches(n_tiles: int, batch: int, rank: int, world_size: int) -> list[tuple[int, int]]:
    """[start, stop) tile ranges of this rank: batches of ``batch`` consecutive tiles dealt round-robin to ranks."""
    if batch <= 0 or world_size <= 0 or not 0 <= rank < world_size:
        raise ValueError("bad batch / rank / world_size")
    starts = range(0, n_tiles, batch)
    return [(s, min(s + batch, n_tiles)) for i, s in enumerate(starts) if i % world_size == rank]


class MosaicPredictor:
    """Tiled predict of one scene into a uint8 class mosaic.

    ``model`` is an ``EfficientnetUnet`` (or any module mapping fp32 [B, C, t, t] CUDA tensors to [B, classes, t, t]
    logits); ``mean`` / ``std`` / ``max_pixel_value`` are the datamodule's Normalize constants
    (s2osm_datamodule.py:75-87)."""

    def __init__(self, model: torch.nn.Module, mean, std, tile: int = 512, margin: int = 0, batch: int = 8,
                 max_pixel_value: float = 255.0, rank: int | None = None, world_size: int | None = None,
                 device: torch.device | str = "cuda", engine_batch: int = 64) -> None:
        plan_mosaic(tile, tile, tile, margin)                     # validates tile / margin
        self.model, self.tile, self.margin, self.batch = model, int(tile), int(margin), int(batch)
        self.device = torch.device(device)
        self.mean = torch.as_tensor(mean, dtype=torch.float32).to(self.device).contiguous()
        self.std = torch.as_tensor(std, dtype=torch.float32).to(self.device).contiguous()
        self.mpv = float(max_pixel_value)
        dist_on = torch.distributed.is_available() and torch.distributed.is_initialized()
        self.rank = rank if rank is not None else (torch.distributed.get_rank() if dist_on else 0)
        self.world_size = world_size if world_size is not None else (torch.distributed.get_world_size() if dist_on else 1)
        self.lib = L.load()
        # tiles per network launch list: a predict call of ``batch`` tiles (bs = 256 in BASELINE.json configs[4]) runs
        # as batch / engine_batch forward passes.  Eval mode has no cross-sample arithmetic (running BatchNorm
        # statistics, per-sample squeeze-excite), so the split is exact; it keeps every activation tensor below
        # 2^31 elements (256 x 256 x 256 x 144 = 2.4e9 at bs = 256) and the workspace at ~16 GB instead of ~64 GB.
        self.engine_batch = max(1, min(int(engine_batch), self.batch))
        while self.batch % self.engine_batch:
            self.engine_batch -= 1

    # ------------------------------------------------------------------ the two device stages around the network
    def gather(self, scene: torch.Tensor, origins: torch.Tensor) -> torch.Tensor:
        """int16 [C, H, W] scene + int32 [B, 2] origins (both on the device) -> normalised fp32 [B, C, tile, tile]."""
        c, h, w = scene.shape
        x = torch.empty(origins.shape[0], c, self.tile, self.tile, dtype=torch.float32, device=scene.device)
        L.check(self.lib.s2lc_gather_tiles(scene.data_ptr(), origins.data_ptr(), self.mean.data_ptr(),
                                           self.std.data_ptr(), x.data_ptr(), origins.shape[0], c, h, w, self.tile,
                                           self.mpv, torch.cuda.current_stream().cuda_stream))
        return x

    def scatter_argmax(self, logits: torch.Tensor, origins: torch.Tensor, mosaic: torch.Tensor) -> None:
        """argmax(dim=1) of [B, classes, tile, tile] logits -> the tile centres of the uint8 [H, W] mosaic."""
        rows, c, dt = _nhwc_rows(logits)
        if logits.shape[0] != origins.shape[0] or logits.shape[2] != self.tile or logits.shape[3] != self.tile:
            raise ValueError(f"logits {tuple(logits.shape)} do not match {origins.shape[0]} tiles of {self.tile} px")
        L.check(self.lib.s2lc_argmax_mosaic(rows.data_ptr(), c, origins.data_ptr(), mosaic.data_ptr(), mosaic.stride(0),
                                            origins.shape[0], self.tile, self.margin, mosaic.shape[0], mosaic.shape[1],
                                            c, dt, torch.cuda.current_stream().cuda_stream))

    # ------------------------------------------------------------------ whole scene
    @torch.no_grad()
    def predict(self, scene: torch.Tensor, reduce: bool = True) -> torch.Tensor:
        """scene: int16 [C, H, W] digital numbers (host or device) -> uint8 [H, W] class mosaic on the device."""
        if scene.dim() != 3 or scene.dtype != torch.int16:
            raise ValueError("scene must be an int16 [C, H, W] raster")
        scene = scene.to(self.device, non_blocking=True).contiguous()
        if scene.shape[0] != self.mean.numel():
            raise ValueError(f"scene has {scene.shape[0]} bands, Normalize constants {self.mean.numel()}")
        _, h, w = scene.shape
        origins = plan_mosaic(h, w, self.tile, self.margin)
        mine = rank_batches(origins.shape[0], self.batch, self.rank, self.world_size)
        origins_d = origins.to(self.device)
        mosaic = torch.zeros(h, w, dtype=torch.uint8, device=self.device)
        was_training = self.model.training
        self.model.eval()
        try:
            for s, e in mine:
                o = origins_d[s:e]
                if e - s < self.batch and origins.shape[0] > self.batch:
                    # ragged last batch: pad with tiles parked entirely outside the scene (the writer drops every
                    # pixel of them) instead of building a second workspace shape for one launch
                    park = origins_d.new_full((self.batch - (e - s), 2), -self.tile)
                    o = torch.cat([o, park])
                o = o.contiguous()
                for c0 in range(0, o.shape[0], self.engine_batch):
                    oc = o[c0:c0 + self.engine_batch]
                    self.scatter_argmax(self.model(self.gather(scene, oc)), oc, mosaic)
        finally:
            self.model.train(was_training)
        return self.assemble(mosaic) if reduce else mosaic

    def assemble(self, mosaic: torch.Tensor) -> torch.Tensor:
        """Union of the per-rank mosaics on every rank: they are disjoint and zero elsewhere, so SUM == union (uint8 is
        a native NCCL / gloo type).  The only exchange of the predict path."""
        if self.world_size > 1:
            torch.distributed.all_reduce(mosaic, op=torch.distributed.ReduceOp.SUM)
        return mosaic

    @staticmethod
    def write_npy(mosaic: torch.Tensor, path: str) -> None:
        import numpy as np
        np.save(path, mosaic.cpu().numpy())
