"""Debug aid: compare decoder intermediates (raw conv outputs and their gradients) of the CUDA path with an fp64
oracle, level by level.   python tools/diag_inter.py --version b0 --size 224 --batch 2"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import s2lc_b200  # noqa
import torch  # noqa
from oracle import unet_oracle as O  # noqa
from s2lc_b200 import synthetic  # noqa
from s2lc_b200.losses import FocalLoss  # noqa
from _parity import alpha_weights, build_model, rel  # noqa

ap = argparse.ArgumentParser()
ap.add_argument("--version", default="b0")
ap.add_argument("--size", type=int, default=224)
ap.add_argument("--batch", type=int, default=2)
ap.add_argument("--precision", default="fp32")
a = ap.parse_args()
model, sd = build_model(a.version, a.precision)
x, y = synthetic.tiles(a.batch, a.size), synthetic.labels(a.batch, a.size)
al = alpha_weights()
dt = torch.float64
leaves = {k: (v.to(dt).requires_grad_("running" not in k) if v.is_floating_point() else v.clone()) for k, v in sd.items()}
probe = {}
lg = O.unet_forward(leaves, x.to(dt), a.version, training=True, probe=probe)
loss = O.focal_loss(lg, y, al.to(dt), 2.0, 0.0, 0)
loss.backward()

logits = model(x.cuda())
val = FocalLoss(al, 2.0, 0.0, ignore_index=0)(logits, y.cuda())
val.backward()
torch.cuda.synchronize()
ws = [w for pool in model.engine().workspaces.values() for w in pool if w.need_grad][0]
N = a.batch


def nchw(t, h, w, c=None):
    t = t.float().view(N, h, w, -1).permute(0, 3, 1, 2)
    return t if c is None else t[:, :c]


for j, lv in reversed(list(enumerate(ws.levels))):
    p = lv.kdc
    print(f"--- level {j} ({lv.h}x{lv.w}, rows={lv.rows})")
    print(f"  fwd  cat   {rel(nchw(lv.cat, lv.h, lv.w, lv.c_cat), probe[p + '/cat']):.2e}   z1 {rel(nchw(lv.zc1, lv.h, lv.w), probe[p + '/z1']):.2e}"
          f"   z2 {rel(nchw(lv.zc2, lv.h, lv.w), probe[p + '/z2']):.2e}")
    for nm, dy, z, bn in (("z2", lv.dyB, lv.zc2, lv.bnB), ("z1", lv.dyA, lv.zc1, lv.bnA)):
        g = bn.p.view(1, -1) * dy.float() + bn.q.view(1, -1) * z.float() + bn.t.view(1, -1)
        want = probe[p + "/" + nm].grad
        got = nchw(g, lv.h, lv.w)
        # decompose the error: per-channel least-squares scale and offset of (got - want)
        d = (got.double().cpu() - want)
        ch_off = d.mean((0, 2, 3))
        print(f"  bwd  g({nm}) rel {rel(got, want):.2e}   |mean offset per ch|/rms(want) {float(ch_off.abs().mean() / want.pow(2).mean().sqrt()):.2e}"
              f"   sum-check ours {float(got.double().sum((0, 2, 3)).abs().max()):.2e}")
    want = probe[p + "/cat"].grad[:, :lv.c_dcat if lv.last else lv.c_cat]
    got = nchw(lv.dcat, lv.h, lv.w, want.shape[1])
    print(f"  bwd  dcat  rel {rel(got, want):.2e}")
