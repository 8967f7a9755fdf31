"""Achieved parity numbers at the TRAINED (well-conditioned) state of tests/_trained.py — what tests/test_gpu_trained.py
asserts.  python tools/diag_trained.py"""
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch  # noqa: E402

import s2lc_b200  # noqa: E402,F401
import _trained as T  # noqa: E402
from oracle import unet_oracle as O  # noqa: E402
from s2lc_b200 import synthetic  # noqa: E402
from s2lc_b200.efficientnet_unet import EfficientNetConfig, EfficientnetUnet  # noqa: E402
from s2lc_b200.losses import FocalLoss  # noqa: E402

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def rel(a, b):
    a, b = a.detach().double().flatten().cpu(), b.detach().double().flatten().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


sd, curve = T.train_oracle("cuda")
print("curve", curve[:3], "...", curve[-3:])
xe, ye = T.eval_batch()
xe, ye = xe.cuda(), ye.cuda()
a = T.alpha().cuda()
sd32 = {k: v.clone() for k, v in sd.items()}
lg32, l32, g32 = O.loss_and_grads(sd32, xe, ye, T.VERSION, alpha=a, gamma=2.0, ignore_index=0)
sd64 = {k: (v.double() if v.is_floating_point() else v) for k, v in sd.items()}
lg64, l64, g64 = O.loss_and_grads(sd64, xe.double(), ye, T.VERSION, alpha=a.double(), gamma=2.0, ignore_index=0)
lgbf, lbf, gbf = O.loss_and_grads(sd32, xe, ye, T.VERSION, alpha=a, gamma=2.0, ignore_index=0, autocast_bf16=True)
cfg = EfficientNetConfig(version=T.VERSION, in_channels=6, num_classes=4, drop_connect_rate=None,
                         class_distribution=list(synthetic.CLASS_DISTRIBUTION))
out = {}
for prec in ("fp32", "bf16"):
    m = EfficientnetUnet(cfg)
    m.load_state_dict({k: v.cpu() for k, v in sd.items()})
    m = m.cuda().train().set_precision(prec)
    fn = FocalLoss(T.alpha(), 2.0, 0.0, ignore_index=0)
    logits = m(xe)
    val = fn(logits, ye)
    val.backward()
    torch.cuda.synchronize()
    out[prec] = (logits.detach().float(), float(val), {k: p.grad.detach().float() for k, p in m.named_parameters() if p.grad is not None})
print("loss f64 %.6e f32 %.6e ours32 %.6e oursbf %.6e refbf %.6e" % (float(l64), float(l32), out["fp32"][1], out["bf16"][1], float(lbf)))
print("logits rel-err vs fp64: ref32 %.2e ours32 %.2e | vs fp32 oracle: oursbf %.2e refbf %.2e" % (
    rel(lg32, lg64), rel(out["fp32"][0], lg64), rel(out["bf16"][0], lg32), rel(lgbf.float(), lg32)))
am = lg32.argmax(1)
print("argmax agreement with fp32 oracle: ours bf16 %.5f, reference autocast %.5f; ours fp32 %.6f" % (
    float((out["bf16"][0].argmax(1) == am).float().mean()), float((lgbf.argmax(1) == am).float().mean()),
    float((out["fp32"][0].argmax(1) == am).float().mean())))
mx = max(float(v.norm()) for v in g64.values())
rows = []
for k, g in g64.items():
    den = max(float(g.norm()), 1e-6 * mx)
    rows.append((k, float((g32[k].double() - g).norm()) / den, float((out["fp32"][2][k].double() - g).norm()) / den,
                 float((out["bf16"][2][k].double() - g).norm()) / den, float((gbf[k].double() - g).norm()) / den))
for name, idx in (("ref32", 1), ("ours32", 2), ("oursbf", 3), ("refbf", 4)):
    e = sorted(r[idx] for r in rows)
    print(f"grad rel-err vs fp64 {name}: median {e[len(e)//2]:.2e} p90 {e[int(len(e)*0.9)]:.2e} max {e[-1]:.2e}; share <= 1e-4: {sum(x <= 1e-4 for x in e)/len(e):.3f}, <= 2e-2: {sum(x <= 2e-2 for x in e)/len(e):.3f}")
well = [r for r in rows if r[1] <= 1e-5]
print("tensors with ref32-vs-64 <= 1e-5:", len(well), "of", len(rows), "; of those ours32 <= 1e-4:", sum(r[2] <= 1e-4 for r in well),
      "worst ours32 on them %.2e" % max((r[2] for r in well), default=0.0))
for r in sorted(well, key=lambda r: -r[2]):
    print(f"  well {r[0]:45s} ref32 {r[1]:.1e} ours32 {r[2]:.1e}")
print("decoder tensors (ref32 / ours32):")
for r in rows:
    if not r[0].startswith("encoder."):
        print(f"  {r[0]:45s} {r[1]:.1e} {r[2]:.1e}  |g| {float(g64[r[0]].norm()):.2e}")
