"""The training step around the hot path — a plain-PyTorch/NCCL restatement of what Lightning does for the
reference's ``SegmentationModule`` (src/train_segmentation.py:43-147), without Lightning (not installed here).

  * ``SegmentationStep``      = ``_model_step`` (:129-147) + ``configure_optimizers`` (:109-114): logits = net(x),
    loss = loss_fn(logits, y), loss.backward(), Adam(lr, weight_decay — L2, not AdamW) step.
  * ``FusedAdam``             = torch.optim.Adam semantics as ONE kernel over the flat parameter buffer
    (SURVEY.md §8 f1); parameters are re-pointed at views of one flat fp32 buffer that has the same layout as the
    engine's flat gradient buffer.
  * ``DataParallelTrainer``   = what ``pl.Trainer(devices=N)`` would do with DDP (the reference never exercises
    it — SURVEY.md F5): per-rank BatchNorm, mean of per-rank fp32 gradients.  The gradient buffer is laid out in
    reverse execution order and all-reduced in ~25 MB buckets on a side stream, each bucket launched from inside
    the backward launch list the moment its last gradient kernel has been enqueued, so NCCL over NVLink overlaps
    the rest of backward.  ``encoder.fc.3.*`` never receives a gradient and is not in any bucket.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import _lib as L
from .efficientnet_unet import EfficientnetUnet


class FusedAdam:
    """``torch.optim.Adam(params, lr, betas, eps, weight_decay)`` (L2 decay folded into the gradient) as a single
    libs2lc kernel over flat buffers.  Only parameters that receive gradients are stepped (as torch skips
    ``grad is None``)."""

    def __init__(self, model: EfficientnetUnet, lr: float = 1.5e-6, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.05) -> None:
        self.model, self.lr, self.betas, self.eps, self.wd = model, lr, betas, eps, weight_decay
        eng = model.engine()
        self.eng = eng
        self.flat_param = torch.empty_like(eng.flat_grad)
        self.flat_param.zero_()
        with torch.no_grad():
            for k, p in eng.used:
                o = eng.grad_offsets[k]
                view = self.flat_param[o:o + p.numel()].view_as(p)
                view.copy_(p)
                p.data = view                      # parameters now live in the flat buffer
        eng._param_ptrs = None                     # force the engine to rebuild its pointer tables
        self.exp_avg = torch.zeros_like(self.flat_param)
        self.exp_avg_sq = torch.zeros_like(self.flat_param)
        self.step_count = 0
        self.lib = L.load()
        self._table, self._n_entries, self._table_version, self._table_dtypes, self._table_complete = None, 0, -1, [], False

    def _entry_table(self):
        """Device table of every packed tensor-core weight operand (all dtypes), sorted by flat parameter offset;
        rebuilt when the engine has created new packed weights."""
        eng = self.eng
        if self._table_version != eng.packs_version:
            base = self.flat_param.data_ptr()
            end = base + 4 * self.flat_param.numel()
            ents = [pw.entry(base) for d in eng.packs.values() for pw in d.values()
                    if base <= pw.src.data_ptr() < end]
            ents.sort(key=lambda e: e.start)
            self._n_entries = len(ents)
            if ents:
                arr = (L.PackEntry * len(ents))(*ents)
                raw = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
                self._table = raw.to(self.flat_param.device)
            else:
                self._table = None
            self._table_dtypes = list(eng.packs.keys())
            self._table_complete = all(base <= pw.src.data_ptr() < end for d in eng.packs.values() for pw in d.values())
            self._table_version = eng.packs_version
        return self._table, self._n_entries

    def step(self, grad_scale: float = 1.0) -> None:
        """One launch: Adam over the flat buffer AND the re-pack of every tensor-core weight operand made from it."""
        self.step_count += 1
        st = torch.cuda.current_stream().cuda_stream
        eng = self.eng
        if eng._ptr_signature() != eng._param_ptrs:        # parameters moved since the last forward: no packs to refresh
            table, n = None, 0
        else:
            table, n = self._entry_table()
        L.check(self.lib.s2lc_adam_pack_step(self.flat_param.data_ptr(), eng.flat_grad.data_ptr(),
                                             self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(),
                                             self.flat_param.numel(), self.lr, self.betas[0], self.betas[1], self.eps,
                                             self.wd, self.step_count, grad_scale, L.ptr(table), n, st))
        eng.weight_gen += 1                                # parameters changed behind autograd's back
        if table is not None and self._table_complete and self._table_version == eng.packs_version:
            sig = eng.weight_signature()                   # ... and every packed copy was refreshed in the same pass
            for dt in self._table_dtypes:
                eng._pack_sig[dt] = sig


class SegmentationStep:
    """One optimisation step of the reference loop on one device."""

    def __init__(self, net: EfficientnetUnet, loss_fn, lr: float = 1.5e-6, weight_decay: float = 0.05,
                 fused_optimizer: bool = True, metrics=None) -> None:
        self.net, self.loss_fn = net, loss_fn
        self.fused = fused_optimizer
        # optional ``metrics.SegmentationMetrics``: the reference's ``update_metrics(logits.argmax(1), y)`` (:145),
        # one extra kernel pass over the logits per step instead of four torchmetrics updates
        self.metrics = metrics
        if fused_optimizer:
            self.optimizer = FusedAdam(net, lr=lr, weight_decay=weight_decay)
            net.engine().direct_grad = True
        else:
            self.optimizer = torch.optim.Adam(net.parameters(), lr=lr, weight_decay=weight_decay)

    def forward_backward(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        if not self.fused:
            self.optimizer.zero_grad(set_to_none=True)
        logits = self.net(x)
        loss = self.loss_fn(logits, y)
        if self.metrics is not None:
            self.metrics.update(logits.detach(), y)
        loss.backward()
        return loss.detach()

    def __call__(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        loss = self.forward_backward(x, y)
        self.optimizer.step()
        return loss


class BucketReducer:
    """Averages contiguous buckets of one flat fp32 gradient buffer across the ranks of a process group.
    NCCL reduces with AVG in one collective; gloo (CPU tests) has no AVG, so it sums and divides."""

    def __init__(self, flat_grad: torch.Tensor, buckets, process_group=None) -> None:
        self.flat_grad, self.buckets, self.pg = flat_grad, buckets, process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.avg_native = dist.is_initialized() and dist.get_backend(process_group) == "nccl"

    def reduce(self, bucket_index: int) -> None:
        if self.world == 1:
            return
        a, b = self.buckets[bucket_index][:2]
        view = self.flat_grad[a:b]
        if self.avg_native:
            dist.all_reduce(view, op=dist.ReduceOp.AVG, group=self.pg)
        else:
            dist.all_reduce(view, op=dist.ReduceOp.SUM, group=self.pg)
            view.div_(self.world)


class DataParallelTrainer(SegmentationStep):
    """N ranks, one process per GPU; gradients averaged with bucketed NCCL all-reduce overlapped with backward."""

    def __init__(self, net: EfficientnetUnet, loss_fn, lr: float = 1.5e-6, weight_decay: float = 0.05,
                 bucket_mb: float = 25.0, process_group=None) -> None:
        super().__init__(net, loss_fn, lr, weight_decay, fused_optimizer=True)
        self.pg = process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.eng = net.engine()
        self.bucket_bytes = int(bucket_mb * 2 ** 20)
        self.comm_stream = torch.cuda.Stream() if net_is_cuda(net) else None
        self.buckets = make_buckets(self.eng.grad_offsets, {k: p.numel() for k, p in self.eng.used},
                                    self.eng.flat_grad.numel(), self.bucket_bytes // 4)
        self.reducer = BucketReducer(self.eng.flat_grad, self.buckets, self.pg)
        self.eng.bucket_hook = self._bucket_ready if self.world > 1 else None
        self.eng.buckets = self.buckets
        if self.world > 1:   # identical initial weights on every rank (DDP broadcasts rank 0's at construction)
            dist.broadcast(self.optimizer.flat_param, src=0, group=self.pg)

    snapshot: torch.Tensor | None = None   # bench.py dp_check: receives every bucket's LOCAL values at fire time
    snapshot_only: bool = False            # ... and skips the all-reduce (the pass then ends with the local gradient)

    def _bucket_ready(self, bucket_index: int) -> None:
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self.comm_stream.wait_event(ev)
        with torch.cuda.stream(self.comm_stream):
            if self.snapshot is not None:
                a, b = self.buckets[bucket_index][:2]
                self.snapshot[a:b].copy_(self.eng.flat_grad[a:b])
            if not self.snapshot_only:
                self.reducer.reduce(bucket_index)

    def __call__(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        loss = self.forward_backward(x, y)
        if self.world > 1:
            torch.cuda.current_stream().wait_stream(self.comm_stream)
        self.optimizer.step()
        return loss


def net_is_cuda(net: torch.nn.Module) -> bool:
    return next(net.parameters()).is_cuda


def make_buckets(offsets: dict[str, int], numels: dict[str, int], total: int, bucket_elems: int):
    """Split the flat gradient buffer [0, total) into contiguous buckets of >= ``bucket_elems`` elements on
    parameter boundaries.  Returns [(start, end, [keys])]; offsets ascend in reverse execution order, so bucket 0
    is complete first during backward."""
    items = sorted(offsets.items(), key=lambda kv: kv[1])
    buckets, start, keys = [], 0, []
    for i, (k, o) in enumerate(items):
        keys.append(k)
        end = items[i + 1][1] if i + 1 < len(items) else total
        if end - start >= bucket_elems or i + 1 == len(items):
            buckets.append((start, end, keys))
            start, keys = end, []
    return buckets
