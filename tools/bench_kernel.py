"""Time (CUDA events) ONE libs2lc kernel on a production-sized shape through the C ABI — the unit ncu captures.

    python tools/bench_kernel.py wgrad3x3 --n 32 --h 512 --w 512 --cin 32 --cout 32
    python tools/bench_kernel.py wgrad1x1 --n 32 --h 256 --w 256 --cin 24 --cout 144
    python tools/bench_kernel.py conv3x3  --n 32 --h 512 --w 512 --cin 32 --cout 32
    python tools/bench_kernel.py dwfwd    --n 32 --h 256 --w 256 --cin 144 --k 3 --s 2

Inputs are random bf16 tensors larger than L2 or, for small ones, preceded by an L2 flush (a 256 MB write).
Development aid; judged numbers come from bench.py.
"""
from __future__ import annotations

import argparse
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import s2lc_b200  # noqa: E402,F401
from s2lc_b200 import _lib as L  # noqa: E402
from s2lc_b200._engine import PackedWeight  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["wgrad3x3", "wgrad1x1", "wgradT", "conv3x3", "conv1x1", "dwfwd", "dwbwd", "dwwgrad", "dwbwdf"])
    ap.add_argument("--n", type=int, default=32)
    ap.add_argument("--h", type=int, default=512)
    ap.add_argument("--w", type=int, default=512)
    ap.add_argument("--cin", type=int, default=32)
    ap.add_argument("--cout", type=int, default=32)
    ap.add_argument("--k", type=int, default=3)
    ap.add_argument("--s", type=int, default=1)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--plain", action="store_true", help="no BatchNorm/activation transforms on the operands")
    a = ap.parse_args()
    lib = L.load()
    dev = torch.device("cuda")
    st = torch.cuda.current_stream().cuda_stream
    n, h, w, cin, cout = a.n, a.h, a.w, a.cin, a.cout
    rows = n * h * w
    bf = torch.bfloat16
    g = torch.Generator(device=dev).manual_seed(0)

    def rnd(*shape, dtype=bf):
        return (torch.randn(*shape, generator=g, device=dev) * 0.5).to(dtype)

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    keep = []

    def lazy(t, ld, c, act=1):
        if a.plain:
            s = L.Lazy(t.data_ptr(), ld, None, None, None, 0, 0)
        else:
            sc, sh = 1 + 0.1 * rnd(c, dtype=torch.float32), 0.1 * rnd(c, dtype=torch.float32)
            keep.extend([sc, sh])
            s = L.Lazy(t.data_ptr(), ld, sc.data_ptr(), sh.data_ptr(), None, act, 0)
        keep.append(s)
        return s

    def lazygrad(dy, ld, c, z=None):
        if a.plain or z is None:
            s = L.LazyGrad(dy.data_ptr(), ld, None, 0, None, None, None, None)
        else:
            p, q, t = (1 + 0.1 * rnd(c, dtype=torch.float32), 0.05 * rnd(c, dtype=torch.float32),
                       0.01 * rnd(c, dtype=torch.float32))
            keep.extend([p, q, t])
            s = L.LazyGrad(dy.data_ptr(), ld, z.data_ptr(), ld, p.data_ptr(), q.data_ptr(), t.data_ptr(), None)
        keep.append(s)
        return s

    cb = (cin + 7) // 8 * 8
    if a.what.startswith("wgrad"):
        mode = {"wgrad1x1": L.CONV_1X1, "wgrad3x3": L.CONV_3X3, "wgradT": L.CONVT_FWD}[a.what]
        up = 2 if mode == L.CONVT_FWD else 1
        x = rnd(rows, cb)
        dy, z = rnd(rows * up * up, cout), rnd(rows * up * up, cout)
        taps = {L.CONV_1X1: 1, L.CONV_3X3: 9, L.CONVT_FWD: 4}[mode]
        dw = torch.zeros(cout * cin * taps, dtype=torch.float32, device=dev)
        db = torch.zeros(cout, dtype=torch.float32, device=dev) if mode != L.CONV_1X1 else None
        la = lazy(x, cb, cin, act=2)
        lg = lazygrad(dy, cout, cout, None if mode == L.CONVT_FWD else z)
        p = L.WgradParams(mode, L.BF16, n, h, w, cin, cout, 0, C.pointer(la), C.pointer(lg), dw.data_ptr(), L.ptr(db))
        keep.append(p)
        nbytes = rows * cb * 2 + rows * up * up * cout * 2 * (1 if (a.plain or mode == L.CONVT_FWD) else 2)
        flops = 2.0 * rows * taps * cin * cout

        def run():
            L.check(lib.s2lc_conv_wgrad(C.byref(p), st))
    elif a.what in ("conv3x3", "conv1x1"):
        mode = L.CONV_3X3 if a.what == "conv3x3" else L.CONV_1X1
        taps = 9 if mode == L.CONV_3X3 else 1
        x = rnd(rows, cb)
        wt = torch.randn(cout, cin, 3 if taps == 9 else 1, 3 if taps == 9 else 1, device=dev) * 0.05
        if taps == 9:
            pw = PackedWeight(wt, bf, 1, cout, 9, cin, 0, cin * 9, 1, 9, cpad=((cb + 63) // 64 * 64 if (w % 128 == 0 and cb <= 128) else cb))
        else:
            pw = PackedWeight(wt, bf, 1, cout, 1, cin, 0, cin, 0, 1)
        L.check(lib.s2lc_weight_pack(*pw.args, st))
        out = torch.empty(rows, cout, dtype=bf, device=dev)
        stats = torch.zeros(2 * cout, dtype=torch.float64, device=dev)
        la = lazy(x, cb, cb, act=2)
        cp = L.ConvParams()
        cp.mode, cp.dtype, cp.N, cp.H, cp.W, cp.C = mode, L.BF16, n, h, w, cb
        cp.Cpad, cp.Kpad, cp.Nout, cp.Npad = pw.cpad, pw.kpad, cout, pw.npad
        cp.a_fwd = C.pointer(la)
        cp.w_packed, cp.out, cp.ld_out = pw.buf.data_ptr(), out.data_ptr(), cout
        cp.stats, cp.stats_kind = stats.data_ptr(), L.STATS_MOMENTS
        keep.extend([cp, pw])
        nbytes = rows * cb * 2 + rows * cout * 2
        flops = 2.0 * rows * taps * cin * cout

        def run():
            L.check(lib.s2lc_conv_fwd(C.byref(cp), st))
    else:
        k, s = a.k, a.s
        ho, wo = (h + s - 1) // s, (w + s - 1) // s
        rows_o = n * ho * wo
        c = cin
        x = rnd(rows, c)
        wt = torch.randn(c, 1, k, k, device=dev) * 0.2
        la = lazy(x, c, c, act=1)
        flops = 2.0 * rows_o * c * k * k
        if a.what == "dwfwd":
            out = torch.empty(rows_o, c, dtype=bf, device=dev)
            stats = torch.zeros(2 * c, dtype=torch.float64, device=dev)
            nbytes = (rows + rows_o) * c * 2

            def run():
                L.check(lib.s2lc_dwconv_fwd(C.byref(la), wt.data_ptr(), out.data_ptr(), stats.data_ptr(), n, h, w, c, k, s,
                                            L.BF16, st))
        elif a.what == "dwbwd":
            dy, z = rnd(rows_o, c), rnd(rows_o, c)
            lg = lazygrad(dy, c, c, z)
            out = torch.empty(rows, c, dtype=bf, device=dev)
            stats = torch.zeros(2 * c, dtype=torch.float64, device=dev)
            nbytes = (2 * rows_o + 2 * rows) * c * 2

            def run():
                L.check(lib.s2lc_dwconv_bwd_data(C.byref(lg), wt.data_ptr(), C.byref(la), None, 0, out.data_ptr(), c,
                                                 stats.data_ptr(), n, h, w, c, k, s, L.BF16, st))
        elif a.what == "dwbwdf":
            dy, z = rnd(rows_o, c), rnd(rows_o, c)
            lg = lazygrad(dy, c, c, z)
            out = torch.empty(rows, c, dtype=bf, device=dev)
            stats = torch.zeros(2 * c, dtype=torch.float64, device=dev)
            dw = torch.zeros(c * k * k, dtype=torch.float32, device=dev)
            nbytes = (2 * rows_o + 2 * rows) * c * 2
            flops *= 2

            def run():
                L.check(lib.s2lc_dwconv_bwd(C.byref(lg), wt.data_ptr(), C.byref(la), 1, None, 0, out.data_ptr(), c,
                                            stats.data_ptr(), dw.data_ptr(), n, h, w, c, k, s, L.BF16, st))
        else:
            dy, z = rnd(rows_o, c), rnd(rows_o, c)
            lg = lazygrad(dy, c, c, z)
            dw = torch.zeros(c * k * k, dtype=torch.float32, device=dev)
            nbytes = (2 * rows_o + rows) * c * 2

            def run():
                L.check(lib.s2lc_dwconv_bwd_weight(C.byref(lg), C.byref(la), dw.data_ptr(), n, h, w, c, k, s, L.BF16, st))

    run()
    torch.cuda.synchronize()
    ts = []
    for _ in range(a.reps):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    print(f"{a.what} n={n} {h}x{w} cin={cin} cout={cout} k={a.k} s={a.s} plain={a.plain}: {ms:.3f} ms  "
          f"{nbytes / ms / 1e6:.0f} GB/s (algorithmic)  {flops / ms / 1e9:.1f} TFLOP/s")


if __name__ == "__main__":
    main()
