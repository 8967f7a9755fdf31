"""One-process sweep of the depthwise kernels over the b5 / bs 32 / 512^2 stage shapes (the launches of a bench step):
forward and fused backward, production dispatch next to the generic kernels (debug flag 128), CUDA events, L2 flushed.

    python tools/dw_sweep.py [--reps 5] [--only bwd]

Development aid; judged numbers come from bench.py."""
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

# (H = W, C, k, s, launches per step)
SHAPES = [(256, 24, 3, 1, 2), (256, 48, 3, 1, 1), (256, 144, 3, 2, 1), (128, 240, 3, 1, 4), (128, 240, 5, 2, 1),
          (64, 384, 5, 1, 4), (64, 384, 3, 2, 1), (32, 768, 3, 1, 6), (32, 768, 5, 1, 1), (32, 1056, 5, 1, 6),
          (32, 1056, 5, 2, 1), (16, 1824, 5, 1, 8), (16, 1824, 3, 1, 1), (16, 3072, 3, 1, 2)]


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--n", type=int, default=32)
    ap.add_argument("--only", choices=["fwd", "bwd"], default=None)
    a = ap.parse_args()
    lib = L.load()
    dev = torch.device("cuda")
    st = torch.cuda.current_stream().cuda_stream
    bf = torch.bfloat16
    g = torch.Generator(device=dev).manual_seed(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def rnd(*shape, dtype=bf):
        return (torch.randn(*shape, generator=g, device=dev) * 0.5).to(dtype)

    def timed(run):
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
        return sorted(ts)[len(ts) // 2]

    tot = {}
    n = a.n
    for (hw, c, k, s, cnt) in SHAPES:
        h = w = hw
        ho, wo = (h + s - 1) // s, (w + s - 1) // s
        rows, rows_o = n * h * w, n * ho * wo
        x = rnd(rows, c)
        wt = torch.randn(c, 1, k, k, device=dev) * 0.2
        sc, sh = 1 + 0.1 * rnd(c, dtype=torch.float32), 0.1 * rnd(c, dtype=torch.float32)
        la = L.Lazy(x.data_ptr(), c, sc.data_ptr(), sh.data_ptr(), None, 1, 0)
        dy, z = rnd(rows_o, c), rnd(rows_o, c)
        p, q, t = (1 + 0.1 * rnd(c, dtype=torch.float32), 0.05 * rnd(c, dtype=torch.float32),
                   0.01 * rnd(c, dtype=torch.float32))
        lg = L.LazyGrad(dy.data_ptr(), c, z.data_ptr(), c, p.data_ptr(), q.data_ptr(), t.data_ptr(), None)
        out_f = torch.empty(rows_o, c, dtype=bf, device=dev)
        out_b = torch.empty(rows, c, dtype=bf, device=dev)
        stats = torch.zeros(2 * c, dtype=torch.float64, device=dev)
        dw = torch.zeros(c * k * k, dtype=torch.float32, device=dev)

        def fwd():
            L.check(lib.s2lc_dwconv_fwd(C.byref(la), wt.data_ptr(), out_f.data_ptr(), stats.data_ptr(), n, h, w, c, k,
                                        s, L.BF16, st))

        def bwd():
            L.check(lib.s2lc_dwconv_bwd(C.byref(lg), wt.data_ptr(), C.byref(la), 1, None, 0, out_b.data_ptr(), c,
                                        stats.data_ptr(), dw.data_ptr(), n, h, w, c, k, s, L.BF16, st))

        for name, fn, nbytes in (("fwd", fwd, (rows + rows_o) * c * 2), ("bwd", bwd, (2 * rows_o + 2 * rows) * c * 2)):
            if a.only and a.only != name:
                continue
            ms = timed(fn)
            old = lib.s2lc_set_debug_flags(0)
            lib.s2lc_set_debug_flags(old | 128)
            ms_g = timed(fn)
            lib.s2lc_set_debug_flags(old)
            tot[name] = tot.get(name, 0.0) + cnt * ms
            tot[name + "_generic"] = tot.get(name + "_generic", 0.0) + cnt * ms_g
            print(f"{name} {hw:4d}^2 C{c:<5d} k{k}s{s} x{cnt}: {ms:.3f} ms {nbytes / ms / 1e6:6.0f} GB/s | generic "
                  f"{ms_g:.3f} ms {nbytes / ms_g / 1e6:6.0f} GB/s | x{ms_g / ms:.2f}", flush=True)
        del x, dy, z, out_f, out_b
    print("per step (ms):", {k: round(v, 3) for k, v in tot.items()})


if __name__ == "__main__":
    main()
