"""The streaming tcgen05 conv engine (csrc/sconv_tc.cu) against the tile-at-a-time engine (csrc/igemm_tc.cu) and
the CUDA-core fp32-accumulate engine on identical bf16 operands: forward with lazy BatchNorm+activation(+gate)
operands, bias and moment statistics; data gradient with the BatchNorm-backward affine operand, act'(prev),
residual and BNBWD statistics.  Shapes are chosen so the streaming path is eligible (row width % 128 == 0 for
3x3, H*W % 128 == 0 and >= 296 tiles for 1x1) and cover N splits, two K chunks, ragged row units and pad columns."""
import ctypes as C
import math
import os

import pytest
import torch

from conftest import rel_err
from test_gpu_kernels import L, lazy, lazygrad, lib, mk, nhwc, pack, run_conv, stream  # noqa: F401

pytestmark = pytest.mark.gpu
BF = torch.bfloat16


def _both(fn):
    """Run fn() with the streaming engine disabled (reference) and enabled; return both results."""
    from s2lc_b200 import _lib
    old = _lib.load().s2lc_set_debug_flags(1)      # bit 0: streaming engines off
    try:
        ref = fn()
    finally:
        _lib.load().s2lc_set_debug_flags(old)
    got = fn()
    torch.cuda.synchronize()
    return ref, got


@pytest.mark.parametrize("cin,cbuf,cout,n,h,w", [(38, 40, 32, 2, 12, 128), (32, 32, 32, 1, 70, 256), (88, 88, 64, 2, 10, 128),
                                                 (64, 64, 88, 1, 9, 128), (128, 128, 128, 1, 8, 128), (64, 64, 64, 3, 5, 256)])
def test_stream_conv3x3(lib, cin, cbuf, cout, n, h, w):
    rows = n * h * w
    x = mk((rows, cbuf), 101).to(BF)
    x[:, cin:] = 0
    sc, sh = 1 + 0.1 * mk((cbuf,), 102), 0.1 * mk((cbuf,), 103)
    wt = mk((cout, cin, 3, 3), 104, 1 / math.sqrt(9 * cin))
    bias = 0.1 * mk((cout,), 105)
    cpad = (cbuf + 63) // 64 * 64
    pw = pack(lib, wt, BF, 1, 1, cout, 9, cin, 0, cin * 9, 1, 9, cpad=cpad)
    lz = lazy(x, cbuf, sc, sh, act=2)

    def fwd():
        out = torch.zeros(rows, cout, dtype=BF, device="cuda")
        stats = torch.zeros(2 * cout, dtype=torch.float64, device="cuda")
        run_conv(lib, 1, 1, lz, n, h, w, cbuf, pw, out, cout, bias=bias, stats=stats, kind=1)
        return out.float(), stats

    (ro, rs), (go, gs) = _both(fwd)
    assert rel_err(go, ro) < 2e-3
    assert rel_err(gs, rs) < 1e-3
    # data gradient: A = p*dy + q*z + t, epilogue act'(prev) + BNBWD statistics
    dy, z = mk((rows, cout), 106).to(BF), mk((rows, cout), 107).to(BF)
    p_, q_, t_ = 1 + 0.1 * mk((cout,), 108), 0.05 * mk((cout,), 109), 0.01 * mk((cout,), 110)
    lg = lazygrad(dy, cout, z, p_, q_, t_)
    ncol = (cin + 7) // 8 * 8
    cpad_d = (cout + 63) // 64 * 64
    pwd = pack(lib, wt, BF, 1, 1, cin, 9, cout, 0, 9, 1, cin * 9, flip=True, cpad=cpad_d)
    prev = lazy(x, cbuf, sc, sh, act=2)

    def bwd():
        dx = torch.zeros(rows, ncol, dtype=BF, device="cuda")
        stats = torch.zeros(2 * ncol, dtype=torch.float64, device="cuda")
        run_conv(lib, 1, 1, lg, n, h, w, cout, pwd, dx, ncol, nout=ncol, prev=prev, stats=stats, kind=2)
        return dx.float(), stats

    (ro, rs), (go, gs) = _both(bwd)
    assert rel_err(go, ro) < 2e-3
    assert rel_err(gs, rs) < 2e-3


@pytest.mark.parametrize("cin,cout,n,h,w,gated", [(24, 144, 2, 144, 144, False), (144, 40, 2, 144, 144, True),
                                                  (240, 64, 3, 128, 112, True), (64, 384, 2, 160, 128, False),
                                                  (384, 128, 2, 144, 144, True)])
def test_stream_conv1x1(lib, cin, cout, n, h, w, gated):
    rows = n * h * w
    x = mk((rows, cin), 111).to(BF)
    sc, sh = 1 + 0.1 * mk((cin,), 112), 0.1 * mk((cin,), 113)
    gate = torch.sigmoid(mk((n, cin), 114)) if gated else None
    wt = mk((cout, cin, 1, 1), 115, 1 / math.sqrt(cin))
    pw = pack(lib, wt, BF, 1, 1, cout, 1, cin, 0, cin, 0, 1)
    lz = lazy(x, cin, sc, sh, gate, act=1)

    def fwd():
        out = torch.zeros(rows, cout, dtype=BF, device="cuda")
        stats = torch.zeros(2 * cout, dtype=torch.float64, device="cuda")
        run_conv(lib, 1, 0, lz, n, h, w, cin, pw, out, cout, stats=stats, kind=1)
        return out.float(), stats

    (ro, rs), (go, gs) = _both(fwd)
    assert rel_err(go, ro) < 2e-3
    assert rel_err(gs, rs) < 1e-3
    # data gradient cout -> cin with drop-connect row scale, residual, act'(prev), BNBWD statistics
    dy, z = mk((rows, cout), 116).to(BF), mk((rows, cout), 117).to(BF)
    p_, q_, t_ = 1 + 0.1 * mk((cout,), 118), 0.05 * mk((cout,), 119), 0.01 * mk((cout,), 120)
    rsc = 0.5 + torch.sigmoid(torch.randn(n, generator=torch.Generator(device="cuda").manual_seed(97), device="cuda"))
    lg = L().LazyGrad(dy.data_ptr(), cout, z.data_ptr(), cout, p_.data_ptr(), q_.data_ptr(), t_.data_ptr(), rsc.data_ptr())
    pwd = pack(lib, wt, BF, 1, 1, cin, 1, cout, 0, 1, 0, cin)
    res = mk((rows, cin), 121).to(BF)
    prev = lazy(x, cin, sc, sh, act=1)

    def bwd():
        dx = torch.zeros(rows, cin, dtype=BF, device="cuda")
        stats = torch.zeros(2 * cin, dtype=torch.float64, device="cuda")
        run_conv(lib, 1, 0, lg, n, h, w, cout, pwd, dx, cin, prev=prev, stats=stats, kind=2, res1=res, ld_res1=cin)
        return dx.float(), stats

    (ro, rs), (go, gs) = _both(bwd)
    assert rel_err(go, ro) < 2e-3
    assert rel_err(gs, rs) < 2e-3


def _run_wgrad(lib, mode, lz, lg, n, h, w, cin, cout, shape, bias):
    from test_gpu_kernels import run_wgrad
    dw = torch.zeros(shape, device="cuda")
    db = torch.zeros(cout, device="cuda") if bias else None
    run_wgrad(lib, 1, mode, lz, lg, n, h, w, cin, cout, dw, db)
    return dw, (db if bias else torch.zeros(1, device="cuda"))


@pytest.mark.parametrize("cin,cbuf,cout,n,h,w", [(38, 40, 32, 2, 12, 128), (32, 32, 32, 1, 70, 256), (88, 88, 64, 2, 10, 128),
                                                 (64, 64, 64, 3, 5, 256), (128, 128, 64, 1, 9, 128),
                                                 # deep decoder levels: strips of 64 / 32 px, G slices of 128 channels
                                                 (168, 168, 128, 2, 24, 128), (128, 128, 128, 2, 128, 128),
                                                 (320, 320, 256, 2, 64, 64), (256, 256, 256, 1, 40, 64),
                                                 (640, 640, 512, 2, 32, 32), (512, 512, 512, 3, 32, 32),
                                                 (64, 64, 64, 2, 32, 64), (96, 96, 88, 1, 17, 32)])
def test_stream_wgrad3x3(lib, cin, cbuf, cout, n, h, w):
    """3x3 weight gradient, both operands lazy (BatchNorm+ReLU on A, BatchNorm-backward affine on G) + bias gradient."""
    rows = n * h * w
    x = mk((rows, cbuf), 131).to(BF)
    x[:, cin:] = 0
    plain_a = cbuf != cin                      # the concat buffer is consumed untransformed
    sc, sh = 1 + 0.1 * mk((cbuf,), 132), 0.1 * mk((cbuf,), 133)
    lz = lazy(x, cbuf) if plain_a else lazy(x, cbuf, sc, sh, act=2)
    dy, z = mk((rows, cout), 134).to(BF), mk((rows, cout), 135).to(BF)
    p_, q_, t_ = 1 + 0.1 * mk((cout,), 136), 0.05 * mk((cout,), 137), 0.01 * mk((cout,), 138)
    lg = lazygrad(dy, cout, z, p_, q_, t_)
    (rw, rb), (gw, gb) = _both(lambda: _run_wgrad(lib, 1, lz, lg, n, h, w, cin, cout, (cout, cin, 3, 3), True))
    assert rel_err(gw, rw) < 1e-3
    assert rel_err(gb, rb) < 1e-3


@pytest.mark.parametrize("cin,cout,n,h,w,gated", [(24, 144, 2, 144, 144, False), (144, 40, 2, 144, 144, True),
                                                  (240, 64, 3, 128, 112, True), (64, 384, 2, 160, 128, False),
                                                  (384, 128, 2, 144, 144, True), (48, 24, 2, 144, 144, True),
                                                  (1056, 176, 8, 32, 32, True), (304, 1824, 16, 16, 16, False),
                                                  (3072, 512, 16, 16, 16, True), (128, 768, 8, 32, 32, False)])
def test_stream_wgrad1x1(lib, cin, cout, n, h, w, gated):
    rows = n * h * w
    x = mk((rows, cin), 141).to(BF)
    sc, sh = 1 + 0.1 * mk((cin,), 142), 0.1 * mk((cin,), 143)
    gate = torch.sigmoid(mk((n, cin), 144)) if gated else None
    lz = lazy(x, cin, sc, sh, gate, act=1)
    dy, z = mk((rows, cout), 145).to(BF), mk((rows, cout), 146).to(BF)
    p_, q_, t_ = 1 + 0.1 * mk((cout,), 147), 0.05 * mk((cout,), 148), 0.01 * mk((cout,), 149)
    rsc = 0.5 + torch.sigmoid(torch.randn(n, generator=torch.Generator(device="cuda").manual_seed(97), device="cuda"))
    lg = L().LazyGrad(dy.data_ptr(), cout, z.data_ptr(), cout, p_.data_ptr(), q_.data_ptr(), t_.data_ptr(), rsc.data_ptr())
    (rw, _), (gw, _) = _both(lambda: _run_wgrad(lib, 0, lz, lg, n, h, w, cin, cout, (cout, cin, 1, 1), False))
    assert rel_err(gw, rw) < 1e-3


@pytest.mark.parametrize("cin,cout,n,h,w,bnz", [(64, 32, 2, 256, 256, True), (128, 64, 2, 128, 128, True),
                                                (256, 128, 2, 64, 64, True), (512, 256, 3, 32, 32, False),
                                                (2048, 512, 4, 16, 16, True)])
def test_stream_wgradT(lib, cin, cout, n, h, w, bnz):
    """2x2 stride-2 transposed-conv weight + bias gradient: input pixel (n, h, w) pairs with the gradient rows of its
    four children (n, 2h+qy, 2w+qx); lazy BatchNorm+ReLU on A, BatchNorm-backward affine on G."""
    rows = n * h * w
    x = mk((rows, cin), 151).to(BF)
    sc, sh = 1 + 0.1 * mk((cin,), 152), 0.1 * mk((cin,), 153)
    lz = lazy(x, cin, sc, sh, act=2)
    dy, z = mk((rows * 4, cout), 154).to(BF), mk((rows * 4, cout), 155).to(BF)
    p_, q_, t_ = 1 + 0.1 * mk((cout,), 156), 0.05 * mk((cout,), 157), 0.01 * mk((cout,), 158)
    lg = lazygrad(dy, cout, z, p_, q_, t_) if bnz else lazygrad(dy, cout)
    (rw, rb), (gw, gb) = _both(lambda: _run_wgrad(lib, 2, lz, lg, n, h, w, cin, cout, (cin, cout, 2, 2), True))
    assert rel_err(gw, rw) < 1e-3
    assert rel_err(gb, rb) < 1e-3
