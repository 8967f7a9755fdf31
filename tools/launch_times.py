"""Per-launch device times of one training step (CUDA events around every libs2lc call of the forward and backward
launch lists), slowest first, with the algorithmic bytes / flops of each launch -> achieved GB/s and TFLOP/s.

    python tools/launch_times.py [--version b5] [--batch 32] [--size 512] [--top 60] [--out gpurun_out/launch_times.txt]

This is a development aid (which launch to optimise next); the judged numbers come from bench.py and ncu.
"""
from __future__ import annotations

import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import s2lc_b200  # noqa: E402,F401
from s2lc_b200 import synthetic  # noqa: E402
from s2lc_b200.efficientnet_unet import EfficientNetConfig, EfficientnetUnet  # noqa: E402
from s2lc_b200.losses import FocalLoss  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--version", default="b5")
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--top", type=int, default=60)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "launch_times.txt"))
    a = ap.parse_args()
    cfg = EfficientNetConfig(version=a.version, in_channels=6, num_classes=4,
                             class_distribution=list(synthetic.CLASS_DISTRIBUTION))
    model = EfficientnetUnet(cfg)
    model.load_state_dict(synthetic.state_dict_like({k: tuple(v.shape) for k, v in model.state_dict().items()}, seed=0))
    model = model.cuda().train().set_precision("bf16")
    alpha = torch.tensor(synthetic.CLASS_DISTRIBUTION)
    alpha[1:] = 1 - alpha[1:]
    loss_fn = FocalLoss(alpha, 2.0, 0.0, ignore_index=0)
    x, y = synthetic.tiles(a.batch, a.size).cuda(), synthetic.labels(a.batch, a.size).cuda()
    for _ in range(2):
        model.zero_grad(set_to_none=True)
        loss_fn(model(x), y).backward()
    eng = model.engine()
    ws = next(w for pool in eng.workspaces.values() for w in pool if w.need_grad)
    acc: dict[tuple, list] = {}
    for _ in range(a.reps):
        ev_f, ev_b = [], []
        logits = ws.run_forward(x.float().contiguous(), events=ev_f)
        lg = logits.clone().requires_grad_(True)
        loss_fn(lg, y).backward()
        ws.run_backward(lg.grad.permute(0, 2, 3, 1).contiguous(), events=ev_b)
        torch.cuda.synchronize()
        for tag, metas, evs in (("F", ws.fwd_meta, ev_f), ("B", ws.bwd_meta, ev_b)):
            for e0, e1, i in evs:
                acc.setdefault((tag, i), [metas[i], []])[1].append(e0.elapsed_time(e1))
    rows = []
    for (tag, i), (m, ts) in acc.items():
        ms = sorted(ts)[len(ts) // 2]
        rows.append((ms, tag, i, m))
    total = sum(r[0] for r in rows)
    rows.sort(key=lambda r: -r[0])
    lines = [f"# {a.version} bs={a.batch} {a.size}x{a.size} bf16: {len(rows)} launches, sum of launch times {total:.2f} ms "
             f"(median of {a.reps} instrumented steps)",
             f"{'ms':>8} {'share':>6} {'GB/s':>8} {'TF/s':>8}  pass idx  kernel  shape"]
    for ms, tag, i, m in rows[:a.top]:
        gb = m[1] / ms / 1e6 if ms > 0 else 0
        tf = m[2] / ms / 1e9 if ms > 0 else 0
        lines.append(f"{ms:8.3f} {ms / total:6.3f} {gb:8.1f} {tf:8.1f}  {tag} {i:4d}  {m[0]}  {m[3]}")
    agg: dict[str, list] = {}
    for ms, tag, i, m in rows:
        r = agg.setdefault(m[0], [0, 0.0, 0.0, 0.0])
        r[0] += 1; r[1] += ms; r[2] += m[1]; r[3] += m[2]
    lines.append("")
    lines.append(f"{'ms':>8} {'share':>6} {'n':>4} {'GB/s':>8} {'TF/s':>8}  kernel class")
    for k, r in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        lines.append(f"{r[1]:8.3f} {r[1] / total:6.3f} {r[0]:4d} {r[2] / r[1] / 1e6 if r[1] else 0:8.1f} "
                     f"{r[3] / r[1] / 1e9 if r[1] else 0:8.1f}  {k}")
    txt = "\n".join(lines)
    print(txt)
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    with open(a.out, "w") as f:
        f.write(txt + "\n")


if __name__ == "__main__":
    main()
