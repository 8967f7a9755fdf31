"""Verbose parity report: CUDA path vs oracle on one config (run on the GPU box).
    python tools/diag_parity.py --version b0 --size 224 --batch 2 --precision fp32 [--dc 0.2] [--loss ce]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import s2lc_b200  # noqa
import torch  # noqa
from _parity import grad_report, rel, run_pair  # noqa

ap = argparse.ArgumentParser()
ap.add_argument("--version", default="b0")
ap.add_argument("--size", type=int, default=224)
ap.add_argument("--batch", type=int, default=2)
ap.add_argument("--precision", default="fp32")
ap.add_argument("--loss", default="focal")
ap.add_argument("--dc", type=float, default=None)
ap.add_argument("--oracle-device", default="cpu")
ap.add_argument("--oracle-bf16", action="store_true")
ap.add_argument("--top", type=int, default=25)
ap.add_argument("--all", action="store_true")
ap.add_argument("--oracle-f64", action="store_true")
a = ap.parse_args()
res = run_pair(a.version, a.size, a.batch, a.precision, a.loss, a.dc, a.oracle_device, a.oracle_bf16,
               torch.float64 if a.oracle_f64 else torch.float32)
print(f"== {a.version} {a.size}^2 bs={a.batch} {a.precision} loss={a.loss} dc={a.dc}")
print(f"loss ours {res['loss']:.7f} oracle {res['o_loss']:.7f}")
print(f"logits rel-err {rel(res['logits'], res['o_logits']):.3e}  finite={bool(torch.isfinite(res['logits']).all())}")
agree = (res['logits'].argmax(1) == res['o_logits'].argmax(1)).float().mean()
print(f"argmax agreement {float(agree) * 100:.4f}%")
rows = grad_report(res)
bad4 = sum(r[0] > 1e-4 for r in rows)
bad2 = sum(r[0] > 2e-2 for r in rows)
print(f"grads: {len(rows)} tensors, max rel-err {rows[0][0]:.3e}, >1e-4: {bad4}, >2e-2: {bad2}")
for r in rows[:a.top]:
    print(f"  {r[0]:.3e}  {r[1]:55s} ours|g|={r[2]:.3e} oracle|g|={r[3]:.3e}")
if a.all:
    print("  --- all, in parameter (execution) order:")
    order = {k: i for i, k in enumerate(res["o_grads"].keys())}
    for r in sorted(rows, key=lambda r: order[r[1]]):
        print(f"  {r[0]:.3e}  {r[1]:55s} ours|g|={r[2]:.3e} oracle|g|={r[3]:.3e}")
print("  ... best:")
for r in rows[-5:]:
    print(f"  {r[0]:.3e}  {r[1]:55s} ours|g|={r[2]:.3e} oracle|g|={r[3]:.3e}")
