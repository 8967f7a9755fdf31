"""One-process timing of the streaming (element-wise / reduction) kernels on b5 / bs 32 stage shapes: bn_apply with a
squeeze-excite gate (the materialised project-conv operand), lazygrad_apply, se_pool, se_bwd_reduce, se_bwd_apply.
CUDA events, L2 flushed before every launch.

    python tools/stream_sweep.py [--reps 5] [--only bn_apply] [--shape 32 1056]

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

SHAPES = [(256, 144), (128, 240), (64, 384), (32, 768), (32, 1056), (16, 1824), (16, 3072)]   # (H = W, C) after the dw conv


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--n", type=int, default=32)
    ap.add_argument("--only", default=None)
    ap.add_argument("--shape", type=int, nargs=2, default=None)
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

    n = a.n
    tot = {}
    for (hw, c) in ([tuple(a.shape)] if a.shape else SHAPES):
        HW = hw * hw
        rows = n * HW
        z, da, out = rnd(rows, c), rnd(rows, c), torch.empty(rows, c, dtype=bf, device=dev)
        sc, sh = 1 + 0.1 * rnd(c, dtype=torch.float32), 0.1 * rnd(c, dtype=torch.float32)
        gate = torch.rand(n, c, device=dev)
        pooled, dpool = torch.empty(n, c, device=dev), torch.randn(n, c, device=dev)
        p, q, t = (1 + 0.1 * rnd(c, dtype=torch.float32), 0.05 * rnd(c, dtype=torch.float32),
                   0.01 * rnd(c, dtype=torch.float32))
        stats = torch.zeros(2 * c, dtype=torch.float64, device=dev)
        lz = L.Lazy(z.data_ptr(), c, sc.data_ptr(), sh.data_ptr(), None, 1, 0)
        lzg = L.Lazy(z.data_ptr(), c, sc.data_ptr(), sh.data_ptr(), gate.data_ptr(), 1, 0)
        lg = L.LazyGrad(da.data_ptr(), c, z.data_ptr(), c, p.data_ptr(), q.data_ptr(), t.data_ptr(), None)
        es = 2
        cases = {
            "bn_apply": (lambda: L.check(lib.s2lc_bn_apply(C.byref(lzg), None, 0, None, out.data_ptr(), c, rows, c, HW,
                                                           L.BF16, st)), 2 * rows * c * es),
            "lazygrad_apply": (lambda: L.check(lib.s2lc_lazygrad_apply(C.byref(lg), out.data_ptr(), c, rows, c, HW,
                                                                       L.BF16, st)), 3 * rows * c * es),
            "se_pool": (lambda: L.check(lib.s2lc_se_pool(C.byref(lz), pooled.data_ptr(), n, HW, c, L.BF16, st)),
                        rows * c * es),
            "se_bwd_reduce": (lambda: L.check(lib.s2lc_se_bwd_reduce(da.data_ptr(), c, C.byref(lz), pooled.data_ptr(), n,
                                                                     HW, c, L.BF16, st)), 2 * rows * c * es),
            "se_bwd_apply": (lambda: L.check(lib.s2lc_se_bwd_apply(da.data_ptr(), c, C.byref(lzg), dpool.data_ptr(),
                                                                   stats.data_ptr(), n, HW, c, L.BF16, st)),
                             3 * rows * c * es),
        }
        for name, (fn, nbytes) in cases.items():
            if a.only and a.only != name:
                continue
            ms = timed(fn)
            tot[name] = tot.get(name, 0.0) + ms
            print(f"{name:15s} {hw:4d}^2 C{c:<5d}: {ms:.4f} ms {nbytes / ms / 1e6:6.0f} GB/s", flush=True)
        del z, da, out
    print("sum over shapes (ms):", {k: round(v, 3) for k, v in tot.items()})


if __name__ == "__main__":
    main()
