"""Debug aid: after k fused Adam + re-pack steps, which packed operands differ from a fresh s2lc_weight_pack of the
stepped parameters, and where (row, tap, channel)?    python tools/diag_pack.py [steps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch  # noqa: E402

import s2lc_b200  # noqa: E402,F401
from _parity import alpha_weights, build_model  # noqa: E402
from s2lc_b200 import _lib, synthetic  # noqa: E402
from s2lc_b200.losses import FocalLoss  # noqa: E402
from s2lc_b200.train_segmentation import SegmentationStep  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 1
ma, _ = build_model("b0", "bf16", None)
fn = FocalLoss(alpha_weights(), 2.0, 0.0, ignore_index=0)
sa = SegmentationStep(ma, fn, lr=1e-3, weight_decay=0.05, fused_optimizer=True)
x, y = synthetic.tiles(4, 128).cuda(), synthetic.labels(4, 128).cuda()
for _ in range(steps):
    sa(x, y)
torch.cuda.synchronize()
eng = ma.engine()
lib = _lib.load()
st = torch.cuda.current_stream().cuda_stream
opt = sa.optimizer
base = opt.flat_param.data_ptr()
print("table entries", opt._n_entries, "complete", opt._table_complete, "packs", sum(len(d) for d in eng.packs.values()))
bad = 0
for key, pw in eng.packs[torch.bfloat16].items():
    have = pw.buf.clone()
    _lib.check(lib.s2lc_weight_pack(*pw.args, st))
    torch.cuda.synchronize()
    if not torch.equal(have, pw.buf):
        bad += 1
        d = (have != pw.buf).nonzero()
        n1, n2, taps, c, s1, s2, stp, sc, flip = pw.geom
        inside = base <= pw.src.data_ptr() < base + 4 * opt.flat_param.numel()
        print(f"{key}: {len(d)} of {have.numel()} differ; geom {pw.geom} cpad {pw.cpad} kpad {pw.kpad} src.shape "
              f"{tuple(pw.src.shape)} start {(pw.src.data_ptr() - base) // 4} in-flat {inside} args-src-ok "
              f"{pw.args[0] == pw.src.data_ptr()}")
        if taps == 9 and flip:
            w = pw.src.detach().float()                      # [o, i, 3, 3]
            exp = torch.zeros_like(have, dtype=torch.float32)
            ex = w.flatten(2).flip(2).permute(1, 2, 0)       # [i, 9 (flipped), o]
            exp.view(exp.shape[0], -1)[: n2, : 9 * pw.cpad].view(n2, 9, pw.cpad)[:, :, :c] = ex
            exp = exp.to(have.dtype)
            print("    fused == torch:", torch.equal(have, exp), " fresh == torch:", torch.equal(pw.buf, exp))
            rows = sorted(set(d[:, 0].tolist())); ks = sorted(set((d[:, 1] // pw.cpad).tolist())); cs = sorted(set((d[:, 1] % pw.cpad).tolist()))
            print("    rows", rows[:20], len(rows), "taps", ks, "channels", cs[:20], len(cs))
        for n, k in d[:6].tolist():
            print(f"    row {n} k {k} (tap {k // pw.cpad}, c {k % pw.cpad}): fused {float(have[n, k]):.5f} fresh {float(pw.buf[n, k]):.5f}")
print("mismatching packs:", bad)
