"""Diagnostic: 3x3 streaming conv vs tile engine vs torch on one small shape, under descriptor variants."""
import math, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, torch.nn.functional as F
import s2lc_b200  # noqa
from test_gpu_kernels import L, lazy, mk, nhwc, nchw, pack, run_conv
from s2lc_b200 import _lib
lib = _lib.load()
BF = torch.bfloat16
def rel(a, b): return float((a - b).norm() / b.norm())
for (cin, cout, n, h, w, only_tap) in [(32, 32, 1, 8, 128, None), (32, 32, 1, 8, 128, 4), (32, 32, 1, 8, 128, 3), (32, 32, 1, 8, 128, 5), (32, 32, 1, 8, 128, 1), (64, 64, 1, 8, 256, None)]:
    rows = n * h * w
    x = mk((n, cin, h, w), 1)
    xr = nhwc(x, BF)
    wt = mk((cout, cin, 3, 3), 2, 1 / math.sqrt(9 * cin))
    if only_tap is not None:
        m = torch.zeros(9, device="cuda"); m[only_tap] = 1
        wt = wt * m.view(1, 1, 3, 3)
    ref = F.conv2d(nchw(xr, n, h, w), wt.to(BF).float(), padding=1)
    pw = pack(lib, wt, BF, 1, 1, cout, 9, cin, 0, cin * 9, 1, 9, cpad=(cin + 63) // 64 * 64)
    lz = lazy(xr, cin)
    res = {}
    for name, flag in [("tile", 1), ("stream", 0)]:
        lib.s2lc_set_debug_flags(flag)
        out = torch.zeros(rows, cout, dtype=BF, device="cuda")
        run_conv(lib, 1, 1, lz, n, h, w, cin, pw, out, cout)
        torch.cuda.synchronize()
        got = nchw(out, n, h, w)
        res[name] = rel(got, ref)
        if name != "tile" and only_tap is None and cin == 32:
            # per-row / per-column error map to see a pattern
            e = (got - ref).abs().amax(dim=(0, 1))
            print(name, "err by h:", [round(float(v), 3) for v in e.amax(1)])
            print(name, "err by w (first 12, last 4):", [round(float(v), 3) for v in e.amax(0)[:12]], [round(float(v), 3) for v in e.amax(0)[-4:]])
    print(cin, cout, n, h, w, "tap", only_tap, res)
