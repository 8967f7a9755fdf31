"""bench.py — EffNet-UNet-b5 training throughput (tiles/s) on N B200s, the reference CPU path beside it.

    python bench.py --gpus 1 --steps 10 --warmup 3                  # our arm (torchrun for N > 1)
    python bench.py --impl reference --gpus 1 --steps 3 --warmup 1  # the reference's CPU path (oracle port)

A "step" = forward + focal loss + backward + gradient all-reduce + Adam over one batch of synthetic 6-band
512x512 tiles (BASELINE.json configs[1]: b5, bs=32/GPU, bf16).  One JSON line is printed by rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# stdout must carry exactly ONE JSON line, but libraries write there too (NCCL prints its "NCCL version ..." banner to
# fd 1 at every NCCL_DEBUG level >= VERSION, which includes WARN): keep the real stdout aside for the result line and
# point fd 1 at stderr for everything else.
_RESULT_FD = os.dup(1)
os.dup2(2, 1)


def emit(result: dict) -> None:
    os.write(_RESULT_FD, (json.dumps(result) + "\n").encode())


import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--version", default="b5")
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--mode", default="train", choices=["train", "infer"],
                    help="train: BASELINE.json configs[1-3] (the headline); infer: configs[4], forward-only mosaic predict")
    ap.add_argument("--ref-batch", type=int, default=1,
                    help="--impl reference: tiles per CPU step (BASELINE.md section 4 rows; the driver's arm uses 1)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-baseline", action="store_true")
    ap.add_argument("--no-breakdown", action="store_true")
    a = ap.parse_args()
    if a.mode == "infer":
        if a.batch == 32:
            a.batch = 256          # configs[4]: bs = 256 tiles per predict call
        if a.steps == 10:
            a.steps = 0            # 0 = every strip of the 12.4k-tile mosaic that falls on this rank
    return a


def peaks() -> tuple[dict, str]:
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index: int) -> None:
        self.index, self.rows, self.proc = index, [], None

    def start(self) -> None:
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self) -> None:
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        rows = [r for r in self.rows if len(r) >= 7 and r[0].isdigit()]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        mhz = sorted(int(r[0]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": mhz[len(mhz) // 2], "sm_max_mhz": int(rows[0][1]), "reasons": reasons, "samples": len(rows)}


def cpu_reference_step_fn(version: str, size: int, batch: int):
    """The reference's CPU path for this hot path (oracle port of its PyTorch modules), fp32, fwd+loss+bwd+Adam."""
    import s2lc_b200  # noqa: F401
    from oracle import unet_oracle as O
    from s2lc_b200 import synthetic
    from s2lc_b200.efficientnet_unet import EfficientNetConfig, EfficientnetUnet
    cfg = EfficientNetConfig(version=version, in_channels=6, num_classes=4, drop_connect_rate=None,
                             class_distribution=list(synthetic.CLASS_DISTRIBUTION))
    shapes = {k: tuple(v.shape) for k, v in EfficientnetUnet(cfg).state_dict().items()}
    sd = synthetic.state_dict_like(shapes, seed=0)
    keys = O.used_param_keys(sd)
    state = dict(sd)
    for k in keys:
        state[k] = sd[k].clone().requires_grad_(True)
    opt = torch.optim.Adam([state[k] for k in keys], lr=1.5e-6, weight_decay=0.05)
    alpha = O.loss_class_weights(list(synthetic.CLASS_DISTRIBUTION), masked_loss=True)
    x, y = synthetic.tiles(batch, size), synthetic.labels(batch, size)

    def step() -> float:
        opt.zero_grad(set_to_none=True)
        logits = O.unet_forward(state, x, version, training=True, update_buffers=True)
        loss = O.focal_loss(logits, y, alpha, 2.0, 0.0, 0)
        loss.backward()
        opt.step()
        return float(loss)

    return step


def run_reference(a) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    bs = max(int(a.ref_batch), 1)   # bounded sample: one tile per step of the same workload unless --ref-batch says more
    # all the host cores this process may use (torchrun exports OMP_NUM_THREADS=1 for its workers)
    try:
        ncpu = len(os.sched_getaffinity(0))
    except AttributeError:
        ncpu = os.cpu_count() or 1
    torch.set_num_threads(max(ncpu, 1))
    infer = a.mode == "infer"
    if infer and a.steps <= 0:
        a.steps = 4
    step = cpu_eval_step_fn(a.version, a.size, bs) if infer else cpu_reference_step_fn(a.version, a.size, bs)
    for _ in range(max(a.warmup, 1)):
        step()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        step()
    dt = (time.perf_counter() - t0) / a.steps
    v = bs / dt
    cores = torch.get_num_threads()
    what = "eval forward + argmax" if infer else "fwd+focal+bwd+Adam"
    sample = f"{a.steps} steps of {bs} tile(s) {a.size}x{a.size} fp32 {what} (oracle port of the reference modules)"
    emit({
        "impl": "reference", "metric": "infer_tiles_per_s" if infer else "train_tiles_per_s", "value": v, "unit": "tiles/s", "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"efficientnet-unet-{a.version} {what}, focal loss, {a.size}x{a.size} 6-band tiles",
                   "tier": "B (generalised skip rule, SURVEY.md F1)", "batch_per_step": bs, "host_threads": cores,
                   "os_cpu_count": os.cpu_count(), "torch": torch.__version__},
        "cpu_baseline": {"value": v, "unit": "tiles/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "tiles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })



def gpu_eager_step_fn(version: str, size: int, batch: int, train: bool = True):
    """The reference's own GPU path on THIS box: its PyTorch modules (oracle port, same ATen /
    cuDNN calls) in channels_last under torch.autocast(bf16) — what Lightning's precision="bf16" runs
    (src/configs/segmentation.py:153) — fwd + focal loss + bwd + torch.optim.Adam (train) or eval forward (infer)."""
    from oracle import unet_oracle as O
    from s2lc_b200 import synthetic
    from s2lc_b200.efficientnet_unet import EfficientNetConfig, EfficientnetUnet
    torch.backends.cudnn.benchmark = True
    cfg = EfficientNetConfig(version=version, in_channels=6, num_classes=4, drop_connect_rate=None,
                             class_distribution=list(synthetic.CLASS_DISTRIBUTION))
    shapes = {k: tuple(v.shape) for k, v in EfficientnetUnet(cfg).state_dict().items()}
    sd = synthetic.state_dict_like(shapes, seed=0)
    state = {}
    for k, v in sd.items():
        v = v.cuda()
        state[k] = v.contiguous(memory_format=torch.channels_last) if v.dim() == 4 else v
    keys = O.used_param_keys(state)
    x = synthetic.tiles(batch, size).cuda().contiguous(memory_format=torch.channels_last)
    y = synthetic.labels(batch, size).cuda()
    alpha = O.loss_class_weights(list(synthetic.CLASS_DISTRIBUTION), masked_loss=True).cuda()
    nb = len(O.blocks_for(version)[1])
    if not train:
        def step():
            with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
                return O.unet_forward(state, x, version, training=False).argmax(1)
        return step
    for k in keys:
        state[k] = state[k].clone().requires_grad_(True)
    opt = torch.optim.Adam([state[k] for k in keys], lr=1.5e-6, weight_decay=0.05)
    g = torch.Generator(device="cuda").manual_seed(1)

    def step():
        opt.zero_grad(set_to_none=True)
        # drop-connect as the reference draws it (efficientnet_unet.py:257, :390-398): rate 0.2*i/n per block
        keep = 1.0 - 0.2 * torch.arange(nb, device="cuda", dtype=torch.float32) / nb
        binary = torch.floor(keep[:, None] + torch.rand(nb, batch, device="cuda", generator=g))
        with torch.autocast("cuda", dtype=torch.bfloat16):
            logits = O.unet_forward(state, x, version, training=True, update_buffers=True, dc_masks=(binary, keep))
            loss = O.focal_loss(logits, y, alpha, 2.0, 0.0, 0)
        loss.backward()
        opt.step()
        return loss

    return step


def gpu_eager_baseline(a, train: bool = True, batch: int | None = None) -> dict:
    """Timed in the same process AFTER the main arm (whose memory has been released); CUDA events, 3 warm-up + 5."""
    batch = batch or a.batch
    torch.cuda.empty_cache()
    try:
        step = gpu_eager_step_fn(a.version, a.size, batch, train)
        for _ in range(3):
            step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 5
        e0.record()
        for _ in range(n):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        peak_gb = torch.cuda.max_memory_allocated() / 2 ** 30
    except torch.OutOfMemoryError as e:        # noqa: F841
        torch.cuda.empty_cache()
        if batch > 8:
            return gpu_eager_baseline(a, train, batch // 2)
        return {"unavailable": "out of memory"}
    finally:
        torch.backends.cudnn.benchmark = False
    return {"value": batch / (ms / 1e3), "unit": "tiles/s", "ms_per_step": ms, "batch": batch,
            "what": ("reference modules (oracle port: the same F.conv2d / F.batch_norm / F.silu ... calls) eager on this "
                     "GPU: cuDNN/ATen, channels_last, torch.autocast(bf16), cudnn.benchmark, "
                     + ("fwd + focal loss + bwd + torch.optim.Adam, drop-connect 0.2" if train else "eval forward + argmax")),
            "torch": torch.__version__, "peak_mem_gb": round(peak_gb, 1)}


def dp_check(trainer, model, x, y, world: int) -> dict:
    """SURVEY.md §8(e) on NCCL.  (1) every bucket fires only after its last gradient kernel: an overlapped pass whose
    hook snapshots each bucket on the communication stream at fire time but does NOT reduce it must end with a local
    gradient that is BIT-IDENTICAL to the snapshots (a bucket fired early would miss a kernel's contribution);
    (2) the bucket-reduced buffer of a real overlapped step equals the mean of the ranks' fire-time local gradients
    (<= 1e-6); (3) the parameters are bit-identical on all ranks after the timed steps."""
    eng = model.engine()
    fg = eng.flat_grad
    snap = torch.zeros_like(fg)
    trainer.snapshot, trainer.snapshot_only = snap, True
    trainer.forward_backward(x, y)
    torch.cuda.current_stream().wait_stream(trainer.comm_stream)
    torch.cuda.synchronize()
    fire_time_complete = bool(torch.equal(snap, fg))
    trainer.snapshot_only = False
    snap.zero_()
    trainer.forward_backward(x, y)
    torch.cuda.current_stream().wait_stream(trainer.comm_stream)
    torch.cuda.synchronize()
    trainer.snapshot = None
    gathered = [torch.empty_like(snap) for _ in range(world)]
    dist.all_gather(gathered, snap)
    mean = torch.stack(gathered).double().mean(0)
    err_mean = float((fg.double() - mean).norm() / mean.norm().clamp_min(1e-30))
    t = torch.tensor([err_mean, 0.0 if fire_time_complete else 1.0], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    # (3) parameter hash after the timed steps: exact integer sum of the raw bits
    bits = trainer.optimizer.flat_param.view(torch.int32).to(torch.int64)
    h = torch.stack([bits.sum(), (bits * (torch.arange(bits.numel(), device="cuda") % 8191 + 1)).sum()])
    hs = [torch.empty_like(h) for _ in range(world)]
    dist.all_gather(hs, h)
    same = all(torch.equal(hs[0], o) for o in hs)
    return {"ranks": world, "grad_vs_mean_of_rank_grads_rel_err": float(t[0]), "tolerance": 1e-6,
            "buckets_complete_at_fire_time_all_ranks": bool(float(t[1]) == 0.0),
            "param_hash_equal_all_ranks": bool(same), "buckets": len(trainer.buckets),
            "ok": bool(float(t[0]) <= 1e-6 and float(t[1]) == 0.0 and same)}


def run_infer(a) -> None:
    """BASELINE.json configs[4]: b5 forward-only, bs = 256 tiles per predict call, over a synthetic country-scale
    mosaic: 49 scene strips of 2 x 128 tiles of 512 x 512 px (12 544 tiles >= AOI `fr`, src/configs/data_config.py:65)
    dealt round-robin over the ranks.  A step = one strip through MosaicPredictor.predict: int16 scene strip ->
    tiles (gather + Normalize) -> eval forward -> argmax -> uint8 mosaic strip
    (src/train_segmentation.py:106-107,135-136; src/experiments/inference_demo.py:14-41)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import numpy as np
    import s2lc_b200  # noqa: F401
    from s2lc_b200 import synthetic
    from s2lc_b200.efficientnet_unet import EfficientNetConfig, EfficientnetUnet
    from s2lc_b200.inference import MosaicPredictor
    cfg = EfficientNetConfig(version=a.version, in_channels=6, num_classes=4,
                             class_distribution=list(synthetic.CLASS_DISTRIBUTION))
    model = EfficientnetUnet(cfg)
    model.load_state_dict(synthetic.state_dict_like({k: tuple(v.shape) for k, v in model.state_dict().items()}, seed=0))
    model = model.cuda().eval().set_precision(a.precision)
    tile, cols = a.size, 128
    rows_per_strip = a.batch // cols if a.batch % cols == 0 and a.batch >= cols else 1
    if rows_per_strip == 1:
        cols = a.batch
    total_strips = -(-12400 // a.batch)
    mine = [s for s in range(total_strips) if s % world == rank]
    steps = a.steps if a.steps > 0 else len(mine)
    mean = np.array([5.1, 4.2, 3.9, 9.8, 7.7, 6.0], dtype=np.float32)
    std = np.array([2.0, 1.7, 1.9, 3.1, 2.6, 2.2], dtype=np.float32)
    mp = MosaicPredictor(model, mean, std, tile=tile, margin=0, batch=a.batch, rank=0, world_size=1)
    rs = np.random.RandomState(99 + rank)
    hs, ws_ = rows_per_strip * tile, cols * tile
    strip_h = torch.from_numpy(rs.randint(0, 2600, size=(6, hs, ws_), dtype=np.int16)).pin_memory()
    out_h = torch.empty(hs, ws_, dtype=torch.uint8).pin_memory()
    strip_d = strip_h.cuda()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(a.warmup, 1)):
        m = mp.predict(strip_d, reduce=False)
    torch.cuda.synchronize()
    eng = model.engine()
    ws = next(w for pool in eng.workspaces.values() for w in pool)
    agg, top = None, None
    if not a.no_breakdown:
        ev = []
        with torch.no_grad():
            ws.run_forward(mp.gather(strip_d, mp_origins(mp, hs, ws_)[:ws.N].cuda()), events=ev)
        torch.cuda.synchronize()
        agg = {}
        for e0, e1, i in ev:
            label, nbytes, flops, _ = ws.fwd_meta[i]
            r = agg.setdefault(label, [0, 0.0, 0.0, 0.0])
            r[0] += 1; r[1] += e0.elapsed_time(e1); r[2] += nbytes; r[3] += flops
        top = max(agg.items(), key=lambda kv: kv[1][1])[0]
        eng.timing_hook = {"fwd": [], "bwd": [], "only": {top}}
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        m = mp.predict(strip_d, reduce=False)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1) / steps
    clocks = sampler.stop() if rank == 0 else None
    hook, eng.timing_hook = eng.timing_hook, None
    # ---- end to end: host int16 strip in, host uint8 mosaic strip out, every step (next strip's copy overlaps)
    copy_stream = torch.cuda.Stream()
    bufs = [strip_d, torch.empty_like(strip_d)]
    barrier()
    t0 = time.perf_counter()
    with torch.cuda.stream(copy_stream):
        bufs[0].copy_(strip_h, non_blocking=True)
    for i in range(steps):
        cur, nxt = bufs[i % 2], bufs[(i + 1) % 2]
        torch.cuda.current_stream().wait_stream(copy_stream)
        with torch.cuda.stream(copy_stream):
            nxt.copy_(strip_h, non_blocking=True)
        m = mp.predict(cur, reduce=False)
        out_h.copy_(m, non_blocking=True)
        torch.cuda.synchronize()
    ms_e2e = (time.perf_counter() - t0) * 1e3 / steps
    barrier()
    t = torch.tensor([ms, ms_e2e], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = float(t[0]), float(t[1])
    out = None
    if rank == 0:
        pk, pk_kind = peaks()
        out = {
            "metric": "infer_tiles_per_s", "value": a.batch * world / (ms / 1e3), "unit": "tiles/s", "n_gpus": world,
            "steps": steps, "warmup": max(a.warmup, 1), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": a.precision.replace("fp32", "f32"), "data": "synthetic",
            "config": {"workload": f"efficientnet-unet-{a.version} forward-only mosaic predict (gather+Normalize -> eval "
                                   f"forward -> argmax -> uint8 mosaic), bs={a.batch} tiles/call, {tile}x{tile} 6-band "
                                   f"tiles, {total_strips * a.batch} tiles = {total_strips} strips dealt over {world} rank(s)",
                       "tier": "B (generalised skip rule, SURVEY.md F1)", "engine_chunk": mp.engine_batch,
                       "parallelism": f"dp{world} (tiles strided over ranks, no exchange)",
                       "l2_policy": "working set >> L2 (805 MB scene strip + GBs of activations per call)"},
            "e2e": {"value": a.batch * world / (ms_e2e / 1e3), "unit": "tiles/s",
                    "h2d_bytes_per_step": strip_h.numel() * 2, "d2h_bytes_per_step": out_h.numel()},
            "gpu_launches": (len(ws.fwd_list) * (a.batch // ws.N) + 2 * (a.batch // ws.N)) * steps,
            "clocks": clocks, "simt_fallbacks": int(eng.lib.s2lc_simt_fallbacks(0)),
        }
        if agg is not None:
            out["roofline"] = roofline_of(top, ws, hook["fwd"], [], steps * (a.batch // ws.N), pk, pk_kind, ms / (a.batch // ws.N))
            out["kernel_breakdown"] = breakdown_table(agg)
    if world == 1 and rank == 0:
        del mp, model, eng, ws
        if not a.no_gpu_baseline:
            out["gpu_eager_baseline"] = gpu_eager_baseline(a, train=False, batch=min(a.batch, 64))
        if not a.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(a, train=False)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        emit(out)


def mp_origins(mp, h, w):
    from s2lc_b200.inference import plan_mosaic
    return plan_mosaic(h, w, mp.tile, mp.margin)

def main() -> None:
    a = parse()
    if a.impl == "reference":
        run_reference(a)
        return
    if a.mode == "infer":
        run_infer(a)
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import s2lc_b200  # noqa: F401
    from s2lc_b200 import synthetic
    from s2lc_b200.efficientnet_unet import EfficientNetConfig, EfficientnetUnet
    from s2lc_b200.losses import FocalLoss
    from s2lc_b200.train_segmentation import DataParallelTrainer

    cfg = EfficientNetConfig(version=a.version, in_channels=6, num_classes=4,
                             class_distribution=list(synthetic.CLASS_DISTRIBUTION))
    model = EfficientnetUnet(cfg)
    model.load_state_dict(synthetic.state_dict_like({k: tuple(v.shape) for k, v in model.state_dict().items()}, seed=0))
    model = model.cuda().train().set_precision(a.precision)
    alpha = torch.tensor(synthetic.CLASS_DISTRIBUTION)
    alpha[1:] = 1 - alpha[1:]
    loss_fn = FocalLoss(alpha, 2.0, 0.0, ignore_index=0)
    trainer = DataParallelTrainer(model, loss_fn, lr=1.5e-6, weight_decay=0.05)

    xh = synthetic.tiles(a.batch, a.size, seed=1234 + rank).pin_memory()
    yh = synthetic.labels(a.batch, a.size, seed=1235 + rank).pin_memory()
    x, y = xh.cuda(non_blocking=True), yh.cuda(non_blocking=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(a.warmup):
        trainer(x, y)
    eng = model.engine()
    ws = next(w for pool in eng.workspaces.values() for w in pool if w.need_grad)
    agg, top = None, None
    if not a.no_breakdown:   # extra instrumented warm-up step: per-class times, picks the dominant kernel class
        agg = kernel_breakdown(trainer, model, ws, x, y)
        top = max(agg.items(), key=lambda kv: kv[1][1])[0]
        eng.timing_hook = {"fwd": [], "bwd": [], "only": {top}}
    # ---- device-resident throughput
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        loss = trainer(x, y)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1) / a.steps
    clocks = sampler.stop() if rank == 0 else None
    hook, eng.timing_hook = eng.timing_hook, None
    # ---- end to end: host buffers in, loss out, every step
    # Every step copies ONE batch from pinned host memory and reads its loss back.  The copy of batch i+1 runs on a
    # side stream while step i computes (double-buffered device inputs), as a prefetching loader would do it; the
    # priming copy of batch 0 is inside the timed region too (steps + 1 copies for steps steps).
    copy_stream = torch.cuda.Stream()
    bufs = [(x, y), (torch.empty_like(x), torch.empty_like(y))]
    barrier()
    t0 = time.perf_counter()
    with torch.cuda.stream(copy_stream):
        bufs[0][0].copy_(xh, non_blocking=True)
        bufs[0][1].copy_(yh, non_blocking=True)
    for i in range(a.steps):
        cur, nxt = bufs[i % 2], bufs[(i + 1) % 2]
        torch.cuda.current_stream().wait_stream(copy_stream)      # batch i has landed
        with torch.cuda.stream(copy_stream):                      # batch i+1 (its buffers were last used by step i-1,
            nxt[0].copy_(xh, non_blocking=True)                   # which the loss read-back below has synchronised)
            nxt[1].copy_(yh, non_blocking=True)
        loss_host = float(trainer(cur[0], cur[1]))
    torch.cuda.synchronize()
    ms_e2e = (time.perf_counter() - t0) * 1e3 / a.steps
    barrier()
    t = torch.tensor([ms, ms_e2e], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = float(t[0]), float(t[1])

    dpc = dp_check(trainer, model, x, y, world) if world > 1 else None
    out = None
    if rank == 0:
        pk, pk_kind = peaks()
        launches = len(ws.fwd_list) + len(ws.bwd_list) + 4    # + loss, loss-scale, 2 memsets-as-kernels, Adam+re-pack
        out = {
            "metric": "train_tiles_per_s", "value": a.batch * world / (ms / 1e3), "unit": "tiles/s", "n_gpus": world,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": a.precision.replace("fp32", "f32"), "data": "synthetic",
            "config": {"workload": f"efficientnet-unet-{a.version} train step (fwd + focal loss + bwd + grad all-reduce + "
                                   f"Adam), bs={a.batch}/GPU, {a.size}x{a.size} 6-band tiles",
                       "tier": "B (generalised skip rule, SURVEY.md F1)", "global_batch": a.batch * world,
                       "parallelism": f"dp{world}", "l2_policy": "working set >> L2 (activations ~10 GB/step)",
                       "final_loss": loss_host},
            "e2e": {"value": a.batch * world / (ms_e2e / 1e3), "unit": "tiles/s",
                    "h2d_bytes_per_step": xh.numel() * 4 + yh.numel() * 8, "d2h_bytes_per_step": 4},
            "gpu_launches": launches * a.steps,
            "clocks": clocks, "simt_fallbacks": int(eng.lib.s2lc_simt_fallbacks(0)),
        }
        if dpc is not None:
            out["dp_check"] = dpc
        if agg is not None:
            out["roofline"] = roofline_of(top, ws, hook["fwd"], hook["bwd"], a.steps, pk, pk_kind, ms)
            out["kernel_breakdown"] = breakdown_table(agg)
            bt = getattr(kernel_breakdown, "best_tensor", None)
            if bt is not None:   # the GEMM side of BASELINE.json's metric: the best tensor-core launch of the step
                tpk = pk.get("bf16_tflops_sustained", pk["bf16_tflops"])
                out["tensor_roofline"] = {
                    "bound": "tensor", "achieved": round(bt[0], 1), "peak": tpk, "unit": "TFLOP/s",
                    "frac": round(bt[0] / tpk, 4), "kernel": bt[1], "layer": bt[2], "launch_ms": round(bt[3], 4),
                    "algorithmic_flops": bt[4], "peak_source": pk_kind,
                    "timing": "CUDA events around this one launch in the instrumented warm-up step (every other launch "
                              "of the step is in kernel_breakdown); ncu: profiles/r02e_tconv3x3_32_ncu.txt"}
    if world == 1 and rank == 0:
        del trainer, model, eng, ws, bufs, x, y          # release the engine's HBM before the eager comparator
        if not a.no_gpu_baseline:
            out["gpu_eager_baseline"] = gpu_eager_baseline(a, train=True)
            if "value" in out["gpu_eager_baseline"]:
                out["gpu_eager_baseline"]["ours_over_it"] = round(out["value"] / out["gpu_eager_baseline"]["value"], 3)
        if not a.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(a)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        emit(out)


def kernel_breakdown(trainer, model, ws, x, y) -> dict:
    """One instrumented warm-up step: CUDA events around every launch of the forward/backward launch lists (on the
    launching stream), aggregated per kernel class.  Returns {label: [launches, ms, algorithmic bytes, flops]}."""
    ev_f, ev_b = [], []
    eng = model.engine()
    hook, eng.bucket_hook = getattr(eng, "bucket_hook", None), None
    logits = ws.run_forward(x.float().contiguous(), events=ev_f)
    lg = logits.clone().requires_grad_(True)
    loss = trainer.loss_fn(lg, y)
    loss.backward()
    d = lg.grad.permute(0, 2, 3, 1).contiguous()
    ws.run_backward(d, events=ev_b)
    eng.bucket_hook = hook
    torch.cuda.synchronize()
    agg: dict[str, list] = {}
    best = None     # fastest tensor-core launch of the step (TFLOP/s), for the `tensor_roofline` key
    for metas, evs in ((ws.fwd_meta, ev_f), (ws.bwd_meta, ev_b)):
        for e0, e1, i in evs:
            label, nbytes, flops, _desc = metas[i]
            t = e0.elapsed_time(e1)
            r = agg.setdefault(label, [0, 0.0, 0.0, 0.0])
            r[0] += 1
            r[1] += t
            r[2] += nbytes
            r[3] += flops
            if flops > 1e10 and t > 0 and label.startswith(("conv3x3", "conv1x1", "convT")):
                tf = flops / t / 1e9
                if best is None or tf > best[0]:
                    best = (tf, label, _desc, t, flops)
    kernel_breakdown.best_tensor = best
    return agg


def breakdown_table(agg: dict) -> list:
    total_ms = sum(r[1] for r in agg.values())
    rows = sorted(agg.items(), key=lambda kv: -kv[1][1])
    return [{"kernel": k, "launches": r[0], "ms": round(r[1], 3), "share": round(r[1] / total_ms, 4),
             "GBps": round(r[2] / r[1] / 1e6, 1) if r[1] > 0 else None,
             "TFLOPs": round(r[3] / r[1] / 1e9, 2) if r[1] > 0 else None} for k, r in rows[:12]]


# kernel-class label -> CUDA kernel behind it (for the ncu cross-reference in profiles/)
KERNEL_OF = {"conv1x1_wgrad": "twgrad_kernel (TMA-fed, blocks <= 128 px) / swgrad_kernel<1,false> (streaming, shallow)",
             "conv3x3_wgrad": "twgrad_kernel (TMA-fed, <= 64 px and packed taps at 256 px) / swgrad_kernel<9,false> (streaming)",
             "convT_wgrad": "twgrad_kernel (TMA-fed, children packed along N for <= 64 output channels)",
             "conv3x3": "tconv_kernel (TMA-fed, <= 128 px rows and 88->64 @256) / sconv_kernel<9,*> (streaming, 256 / 512 px)",
             "conv3x3_dgrad": "tconv_kernel (TMA-fed, <= 128 px rows) / sconv_kernel<9,*> (streaming, 256 / 512 px)",
             "conv1x1": "tconv_kernel (TMA-fed, every plain operand; TMA-store epilogue) / sconv_kernel<1,*> (lazy shallow operands)",
             "conv1x1_dgrad": "tconv_kernel (TMA-fed, blocks <= 128 px) / sconv_kernel<1,*> (lazy shallow operands)",
             "dwconv_fwd": "dw2::dw2_fwd_kernel",
             "dwconv_bwd": "dw2::dw2_bwd_kernel (fused data + weight gradient)"}


def roofline_of(label: str, ws, events_f: list, events_b: list, steps: int, pk: dict, pk_kind: str, step_ms: float) -> dict:
    """Roofline of the dominant kernel class from the events recorded around ITS launches inside the timed region."""
    n, ms, nbytes, flops = 0, 0.0, 0.0, 0.0
    for metas, evs in ((ws.fwd_meta, events_f), (ws.bwd_meta, events_b)):
        for e0, e1, i in evs:
            n += 1
            ms += e0.elapsed_time(e1)
            nbytes += metas[i][1]
            flops += metas[i][2]
    gbps, tf = nbytes / ms / 1e6, flops / ms / 1e9
    tpeak = pk.get("bf16_tflops_sustained", pk["bf16_tflops"])     # timed inside a long step: sustained figure
    frac_hbm, frac_tc = gbps / pk["hbm_gbs"], tf / tpeak
    if frac_tc > frac_hbm:
        roof = {"bound": "tensor", "achieved": tf, "peak": tpeak, "unit": "TFLOP/s", "frac": frac_tc}
    else:
        roof = {"bound": "hbm", "achieved": gbps, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": frac_hbm}
    traffic = None
    prof = os.path.join(ROOT, "profiles", "top_kernel_traffic.json")
    if os.path.exists(prof):
        with open(prof) as f:
            t = json.load(f)
        if t.get("kernel_class") == label:      # ncu capture of ONE launch of this class: scale its DRAM/algorithmic
            traffic = t.get("dram_bytes_per_launch")   # ratio to the class-average launch the roofline is quoted on
            if traffic is None and t.get("dram_over_algorithmic"):
                traffic = t["dram_over_algorithmic"] * nbytes / max(n, 1)
    roof.update({"kernel": label, "cuda_kernel": KERNEL_OF.get(label, label), "launches_per_step": n // max(steps, 1),
                 "avg_launch_ms": ms / max(n, 1), "algorithmic_bytes_per_launch": nbytes / max(n, 1),
                 "algorithmic_flops_per_launch": flops / max(n, 1), "share_of_step": ms / max(steps, 1) / step_ms,
                 "peak_source": pk_kind, "traffic": traffic,
                 "timing": "CUDA events around each launch of this class on the launching stream, inside the timed region"})
    return roof


def cpu_eval_step_fn(version: str, size: int, batch: int):
    """The reference's CPU predict path (predict_step -> logits -> argmax), oracle port, fp32, eval mode."""
    import s2lc_b200  # noqa: F401
    from oracle import unet_oracle as O
    from s2lc_b200 import synthetic
    from s2lc_b200.efficientnet_unet import EfficientNetConfig, EfficientnetUnet
    cfg = EfficientNetConfig(version=version, in_channels=6, num_classes=4, drop_connect_rate=None,
                             class_distribution=list(synthetic.CLASS_DISTRIBUTION))
    shapes = {k: tuple(v.shape) for k, v in EfficientnetUnet(cfg).state_dict().items()}
    sd = synthetic.state_dict_like(shapes, seed=0)
    x = synthetic.tiles(batch, size)

    def step():
        with torch.no_grad():
            return O.unet_forward(sd, x, version, training=False).argmax(1)

    return step


def cpu_baseline(a, train: bool = True) -> dict:
    step = cpu_reference_step_fn(a.version, a.size, 1) if train else cpu_eval_step_fn(a.version, a.size, 1)
    step()
    t0 = time.perf_counter()
    n = 2 if train else 4
    for _ in range(n):
        step()
    dt = (time.perf_counter() - t0) / n
    what = "fp32, fwd+focal+bwd+Adam" if train else "fp32, eval forward + argmax"
    return {"value": 1.0 / dt, "unit": "tiles/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{n} steps of 1 tile {a.size}x{a.size}, {what} on the host cores "
                      f"(os.cpu_count()={os.cpu_count()})"}


if __name__ == "__main__":
    main()
