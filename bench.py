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
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-breakdown", action="store_true")
    return ap.parse_args()


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
    bs = 1   # bounded sample: one 512x512 tile per step of the same workload
    # all the host cores this process may use (torchrun exports OMP_NUM_THREADS=1 for its workers)
    try:
        ncpu = len(os.sched_getaffinity(0))
    except AttributeError:
        ncpu = os.cpu_count() or 1
    torch.set_num_threads(max(ncpu, 1))
    step = cpu_reference_step_fn(a.version, a.size, bs)
    for _ in range(max(a.warmup, 1)):
        step()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        step()
    dt = (time.perf_counter() - t0) / a.steps
    v = bs / dt
    cores = torch.get_num_threads()
    sample = f"{a.steps} steps of 1 tile {a.size}x{a.size} fp32 fwd+focal+bwd+Adam (oracle port of the reference modules)"
    emit({
        "impl": "reference", "metric": "train_tiles_per_s", "value": v, "unit": "tiles/s", "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"efficientnet-unet-{a.version} fwd+bwd+Adam, focal loss, {a.size}x{a.size} 6-band tiles",
                   "tier": "B (generalised skip rule, SURVEY.md F1)", "batch_per_step": bs, "host_threads": cores,
                   "os_cpu_count": os.cpu_count(), "torch": torch.__version__},
        "cpu_baseline": {"value": v, "unit": "tiles/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "tiles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


def main() -> None:
    a = parse()
    if a.impl == "reference":
        run_reference(a)
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

    out = None
    if rank == 0:
        pk, pk_kind = peaks()
        launches = len(ws.fwd_list) + len(ws.bwd_list) + len(eng.packs.get(ws.dtype, {})) + 3
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
            "clocks": clocks,
        }
        if agg is not None:
            out["roofline"] = roofline_of(top, ws, hook["fwd"], hook["bwd"], a.steps, pk, pk_kind, ms)
            out["kernel_breakdown"] = breakdown_table(agg)
        if world == 1 and not a.no_cpu_baseline:
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
    for metas, evs in ((ws.fwd_meta, ev_f), (ws.bwd_meta, ev_b)):
        for e0, e1, i in evs:
            label, nbytes, flops, _desc = metas[i]
            r = agg.setdefault(label, [0, 0.0, 0.0, 0.0])
            r[0] += 1
            r[1] += e0.elapsed_time(e1)
            r[2] += nbytes
            r[3] += flops
    return agg


def breakdown_table(agg: dict) -> list:
    total_ms = sum(r[1] for r in agg.values())
    rows = sorted(agg.items(), key=lambda kv: -kv[1][1])
    return [{"kernel": k, "launches": r[0], "ms": round(r[1], 3), "share": round(r[1] / total_ms, 4),
             "GBps": round(r[2] / r[1] / 1e6, 1) if r[1] > 0 else None,
             "TFLOPs": round(r[3] / r[1] / 1e9, 2) if r[1] > 0 else None} for k, r in rows[:12]]


# kernel-class label -> CUDA kernel behind it (for the ncu cross-reference in profiles/)
KERNEL_OF = {"conv1x1_wgrad": "swgrad_kernel<1,false> (streaming; wgrad_tc_kernel for the shapes it declines)",
             "conv3x3_wgrad": "swgrad_kernel<9,false> (streaming, every decoder level)",
             "convT_wgrad": "swgrad_kernel<1,true> (streaming, four children)",
             "conv3x3": "sconv_kernel<9,*> (W % 128 == 0) / conv_tc_kernel<2>",
             "conv3x3_dgrad": "sconv_kernel<9,*> (W % 128 == 0) / conv_tc_kernel<2>",
             "conv1x1": "sconv_kernel<1,*> (shallow) / conv_tc_kernel<2> (deep)",
             "conv1x1_dgrad": "sconv_kernel<1,*> (shallow) / conv_tc_kernel<2> (deep)",
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


def cpu_baseline(a) -> dict:
    step = cpu_reference_step_fn(a.version, a.size, 1)
    step()
    t0 = time.perf_counter()
    n = 2
    for _ in range(n):
        step()
    dt = (time.perf_counter() - t0) / n
    return {"value": 1.0 / dt, "unit": "tiles/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{n} steps of 1 tile {a.size}x{a.size}, fp32, fwd+focal+bwd+Adam on the host cores "
                      f"(os.cpu_count()={os.cpu_count()})"}


if __name__ == "__main__":
    main()
