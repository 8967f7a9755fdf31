"""The PRODUCTION tcgen05 engines (streaming sconv / swgrad, and whatever the dispatcher picks for a shape) against
plain torch fp32 ops on the same bf16-rounded operands — one hop to ground truth (VERDICT r01 "what's weak" 3;
tests/test_gpu_stream.py compares the same shapes with the sibling tile/slab engines).

Reference arithmetic: the lazy operand transform is evaluated in fp32 and rounded to bf16 exactly where the engines
round it (the tensor-core operand), weights are rounded to bf16, the contraction runs in true fp32 (TF32 off, see
conftest).  What remains is fp32 accumulation order and the bf16 rounding of the stored output (2^-9 per element,
~1.2e-3 in L2), so the bars are 4e-3 for bf16 outputs and 1e-3 for fp32 weight gradients."""
import math

import pytest
import torch
import torch.nn.functional as F

from conftest import rel_err
from test_gpu_kernels import L, lazy, lazygrad, lib, mk, nchw, pack, run_conv, run_wgrad  # noqa: F401

pytestmark = pytest.mark.gpu
BF = torch.bfloat16
OUT_TOL, DW_TOL = 4e-3, 1e-3


def bfr(t):
    return t.to(BF).float()


def act_fwd(u, act):
    return F.silu(u) if act == 1 else (F.relu(u) if act == 2 else u)


def act_bwd(u, act):
    if act == 1:
        s = torch.sigmoid(u)
        return s * (1 + u * (1 - s))
    if act == 2:
        return (u > 0).float()
    return torch.ones_like(u)


def operand(x_rows, n, h, w, sc=None, sh=None, act=0, gate=None):
    """bf16-rounded tensor-core operand [n, C, h, w] of a lazy activation stored as bf16 rows."""
    u = nchw(x_rows, n, h, w)
    if sc is not None:
        u = u * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1)
    a = act_fwd(u, act)
    if gate is not None:
        a = a * gate.view(n, -1, 1, 1)
    return bfr(a)


def db_err(db, dy_rows, z_rows, n, h, w, p, q, t):
    """The bias gradient is the per-channel sum of the gradient; an engine may sum it before or after the bf16
    rounding of the tensor-core operand (the two differ by a random walk of roundings, ~2^-9 relative to a sum that is
    itself a random walk), so it is held against the nearer of the two."""
    g32 = nchw(dy_rows, n, h, w)
    if z_rows is not None:
        g32 = p.view(1, -1, 1, 1) * g32 + q.view(1, -1, 1, 1) * nchw(z_rows, n, h, w) + t.view(1, -1, 1, 1)
    return min(rel_err(db, g32.sum((0, 2, 3))), rel_err(db, bfr(g32).sum((0, 2, 3))))


def grad_operand(dy_rows, z_rows, n, h, w, p, q, t, rs=None):
    g = nchw(dy_rows, n, h, w)
    if rs is not None:
        g = g * rs.view(n, 1, 1, 1)
    if z_rows is not None:
        g = p.view(1, -1, 1, 1) * g + q.view(1, -1, 1, 1) * nchw(z_rows, n, h, w) + t.view(1, -1, 1, 1)
    return bfr(g)


CONV3 = [(38, 40, 32, 2, 12, 128), (32, 32, 32, 1, 70, 256), (88, 88, 64, 2, 10, 128), (64, 64, 88, 1, 9, 128),
         (128, 128, 128, 1, 8, 128), (64, 64, 64, 3, 5, 256),
         # deep decoder levels (rows narrower than 128 px, wide K): whatever engine the dispatcher picks
         (168, 168, 128, 2, 24, 128), (320, 320, 256, 2, 64, 64), (256, 256, 256, 1, 40, 64), (640, 640, 512, 2, 32, 32),
         (512, 512, 512, 1, 32, 32)]


@pytest.mark.parametrize("cin,cbuf,cout,n,h,w", CONV3)
def test_conv3x3_vs_torch(lib, cin, cbuf, cout, n, h, w):
    rows = n * h * w
    x = mk((rows, cbuf), 201).to(BF)
    x[:, cin:] = 0
    sc, sh = 1 + 0.1 * mk((cbuf,), 202), 0.1 * mk((cbuf,), 203)
    wt = mk((cout, cin, 3, 3), 204, 1 / math.sqrt(9 * cin))
    bias = 0.1 * mk((cout,), 205)
    width_rule = w % 128 == 0 and cbuf <= 128          # _engine.cpad3: per-tap K stride of the packed weight
    cpad = (cbuf + 63) // 64 * 64 if width_rule else cbuf
    pw = pack(lib, wt, BF, 1, 1, cout, 9, cin, 0, cin * 9, 1, 9, cpad=cpad)
    lz = lazy(x, cbuf, sc, sh, act=2)
    out = torch.zeros(rows, cout, dtype=BF, device="cuda")
    stats = torch.zeros(2 * cout, dtype=torch.float64, device="cuda")
    run_conv(lib, 1, 1, lz, n, h, w, cbuf, pw, out, cout, bias=bias, stats=stats, kind=1)
    a = operand(x, n, h, w, sc, sh, 2)[:, :cin]
    ref = F.conv2d(a, bfr(wt), bias, padding=1)
    got = nchw(out, n, h, w)
    assert rel_err(got, ref) < OUT_TOL
    assert rel_err(stats[:cout], got.double().sum((0, 2, 3))) < 1e-5
    assert rel_err(stats[cout:], (got.double() ** 2).sum((0, 2, 3))) < 1e-5
    # data gradient: A = p*dy + q*z + t, epilogue act'(prev), BNBWD statistics
    dy, z = mk((rows, cout), 206).to(BF), mk((rows, cout), 207).to(BF)
    p_, q_, t_ = 1 + 0.1 * mk((cout,), 208), 0.05 * mk((cout,), 209), 0.01 * mk((cout,), 210)
    lg = lazygrad(dy, cout, z, p_, q_, t_)
    ncol = (cin + 7) // 8 * 8
    cpad_d = (cout + 63) // 64 * 64 if (w % 128 == 0 and cout <= 128) else cout
    pwd = pack(lib, wt, BF, 1, 1, cin, 9, cout, 0, 9, 1, cin * 9, flip=True, cpad=cpad_d)
    dx = torch.zeros(rows, ncol, dtype=BF, device="cuda")
    st2 = torch.zeros(2 * ncol, dtype=torch.float64, device="cuda")
    run_conv(lib, 1, 1, lg, n, h, w, cout, pwd, dx, ncol, nout=ncol, prev=lazy(x, cbuf, sc, sh, act=2), stats=st2, kind=2)
    g = grad_operand(dy, z, n, h, w, p_, q_, t_)
    u = nchw(x, n, h, w) * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1)
    refx = F.conv_transpose2d(g, bfr(wt), padding=1) * act_bwd(u[:, :cin], 2)
    gotx = nchw(dx, n, h, w)[:, :cin]
    assert rel_err(gotx, refx) < OUT_TOL
    assert rel_err(st2[:cin], gotx.double().sum((0, 2, 3))) < 1e-5
    assert rel_err(st2[ncol:ncol + cin], (gotx.double() * nchw(x, n, h, w)[:, :cin].double()).sum((0, 2, 3))) < 1e-5


CONV1 = [(24, 144, 2, 144, 144, False), (144, 40, 2, 144, 144, True), (240, 64, 3, 128, 112, True),
         (64, 384, 2, 160, 128, False), (384, 128, 2, 144, 144, True),
         # deep layers (tile engine / split-K territory)
         (1056, 176, 8, 32, 32, True), (304, 1824, 16, 16, 16, False), (3072, 512, 16, 16, 16, True),
         (512, 2048, 16, 16, 16, False)]


@pytest.mark.parametrize("cin,cout,n,h,w,gated", CONV1)
def test_conv1x1_vs_torch(lib, cin, cout, n, h, w, gated):
    rows = n * h * w
    x = mk((rows, cin), 211).to(BF)
    sc, sh = 1 + 0.1 * mk((cin,), 212), 0.1 * mk((cin,), 213)
    gate = torch.sigmoid(mk((n, cin), 214)) if gated else None
    wt = mk((cout, cin, 1, 1), 215, 1 / math.sqrt(cin))
    pw = pack(lib, wt, BF, 1, 1, cout, 1, cin, 0, cin, 0, 1)
    out = torch.zeros(rows, cout, dtype=BF, device="cuda")
    stats = torch.zeros(2 * cout, dtype=torch.float64, device="cuda")
    run_conv(lib, 1, 0, lazy(x, cin, sc, sh, gate, act=1), n, h, w, cin, pw, out, cout, stats=stats, kind=1)
    ref = F.conv2d(operand(x, n, h, w, sc, sh, 1, gate), bfr(wt))
    got = nchw(out, n, h, w)
    assert rel_err(got, ref) < OUT_TOL
    assert rel_err(stats[:cout], got.double().sum((0, 2, 3))) < 1e-5
    assert rel_err(stats[cout:], (got.double() ** 2).sum((0, 2, 3))) < 1e-5
    # data gradient cout -> cin with drop-connect row scale, residual, act'(prev), BNBWD statistics
    dy, z = mk((rows, cout), 216).to(BF), mk((rows, cout), 217).to(BF)
    p_, q_, t_ = 1 + 0.1 * mk((cout,), 218), 0.05 * mk((cout,), 219), 0.01 * mk((cout,), 220)
    rsc = 0.5 + torch.sigmoid(torch.randn(n, generator=torch.Generator(device="cuda").manual_seed(97), device="cuda"))
    lg = L().LazyGrad(dy.data_ptr(), cout, z.data_ptr(), cout, p_.data_ptr(), q_.data_ptr(), t_.data_ptr(), rsc.data_ptr())
    pwd = pack(lib, wt, BF, 1, 1, cin, 1, cout, 0, 1, 0, cin)
    res = mk((rows, cin), 221).to(BF)
    dx = torch.zeros(rows, cin, dtype=BF, device="cuda")
    st2 = torch.zeros(2 * cin, dtype=torch.float64, device="cuda")
    run_conv(lib, 1, 0, lg, n, h, w, cout, pwd, dx, cin, prev=lazy(x, cin, sc, sh, act=1), stats=st2, kind=2, res1=res,
             ld_res1=cin)
    g = grad_operand(dy, z, n, h, w, p_, q_, t_, rsc)
    u = nchw(x, n, h, w) * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1)
    refx = (F.conv_transpose2d(g, bfr(wt)) + nchw(res, n, h, w)) * act_bwd(u, 1)
    gotx = nchw(dx, n, h, w)
    assert rel_err(gotx, refx) < OUT_TOL
    assert rel_err(st2[:cin], gotx.double().sum((0, 2, 3))) < 1e-5
    assert rel_err(st2[cin:], (gotx.double() * nchw(x, n, h, w).double()).sum((0, 2, 3))) < 1e-5


WG3 = [(38, 40, 32, 2, 12, 128), (32, 32, 32, 1, 70, 256), (88, 88, 64, 2, 10, 128), (64, 64, 64, 3, 5, 256),
       (168, 168, 128, 2, 24, 128), (128, 128, 128, 2, 128, 128), (320, 320, 256, 2, 64, 64),
       (256, 256, 256, 1, 40, 64), (640, 640, 512, 2, 32, 32), (512, 512, 512, 3, 32, 32)]


@pytest.mark.parametrize("cin,cbuf,cout,n,h,w", WG3)
def test_wgrad3x3_vs_torch(lib, cin, cbuf, cout, n, h, w):
    rows = n * h * w
    x = mk((rows, cbuf), 231).to(BF)
    x[:, cin:] = 0
    plain_a = cbuf != cin
    sc, sh = 1 + 0.1 * mk((cbuf,), 232), 0.1 * mk((cbuf,), 233)
    lz = lazy(x, cbuf) if plain_a else lazy(x, cbuf, sc, sh, act=2)
    dy, z = mk((rows, cout), 234).to(BF), mk((rows, cout), 235).to(BF)
    p_, q_, t_ = 1 + 0.1 * mk((cout,), 236), 0.05 * mk((cout,), 237), 0.01 * mk((cout,), 238)
    dw = torch.zeros(cout, cin, 3, 3, device="cuda")
    db = torch.zeros(cout, device="cuda")
    run_wgrad(lib, 1, 1, lz, lazygrad(dy, cout, z, p_, q_, t_), n, h, w, cin, cout, dw, db)
    a = (operand(x, n, h, w) if plain_a else operand(x, n, h, w, sc, sh, 2))[:, :cin]
    g = grad_operand(dy, z, n, h, w, p_, q_, t_)
    ref = torch.nn.grad.conv2d_weight(a, (cout, cin, 3, 3), g, padding=1)
    assert rel_err(dw, ref) < DW_TOL
    assert db_err(db, dy, z, n, h, w, p_, q_, t_) < DW_TOL


WG1 = [(24, 144, 2, 144, 144, False), (144, 40, 2, 144, 144, True), (240, 64, 3, 128, 112, True),
       (64, 384, 2, 160, 128, False), (384, 128, 2, 144, 144, True), (48, 24, 2, 144, 144, True),
       (1056, 176, 8, 32, 32, True), (304, 1824, 16, 16, 16, False), (3072, 512, 16, 16, 16, True)]


@pytest.mark.parametrize("cin,cout,n,h,w,gated", WG1)
def test_wgrad1x1_vs_torch(lib, cin, cout, n, h, w, gated):
    rows = n * h * w
    x = mk((rows, cin), 241).to(BF)
    sc, sh = 1 + 0.1 * mk((cin,), 242), 0.1 * mk((cin,), 243)
    gate = torch.sigmoid(mk((n, cin), 244)) if gated else None
    dy, z = mk((rows, cout), 245).to(BF), mk((rows, cout), 246).to(BF)
    p_, q_, t_ = 1 + 0.1 * mk((cout,), 247), 0.05 * mk((cout,), 248), 0.01 * mk((cout,), 249)
    rsc = 0.5 + torch.sigmoid(torch.randn(n, generator=torch.Generator(device="cuda").manual_seed(97), device="cuda"))
    lg = L().LazyGrad(dy.data_ptr(), cout, z.data_ptr(), cout, p_.data_ptr(), q_.data_ptr(), t_.data_ptr(), rsc.data_ptr())
    dw = torch.zeros(cout, cin, 1, 1, device="cuda")
    run_wgrad(lib, 1, 0, lazy(x, cin, sc, sh, gate, act=1), lg, n, h, w, cin, cout, dw)
    a = operand(x, n, h, w, sc, sh, 1, gate).permute(0, 2, 3, 1).reshape(rows, cin)
    g = grad_operand(dy, z, n, h, w, p_, q_, t_, rsc).permute(0, 2, 3, 1).reshape(rows, cout)
    ref = (g.t() @ a).view(cout, cin, 1, 1)
    assert rel_err(dw, ref) < DW_TOL


@pytest.mark.parametrize("cin,cout,n,h,w,bnz", [(64, 32, 2, 256, 256, True), (128, 64, 2, 128, 128, True),
                                                (256, 128, 2, 64, 64, True), (512, 256, 3, 32, 32, False),
                                                (2048, 512, 4, 16, 16, True)])
def test_wgradT_vs_torch(lib, cin, cout, n, h, w, bnz):
    rows = n * h * w
    x = mk((rows, cin), 251).to(BF)
    sc, sh = 1 + 0.1 * mk((cin,), 252), 0.1 * mk((cin,), 253)
    dy, z = mk((rows * 4, cout), 254).to(BF), mk((rows * 4, cout), 255).to(BF)
    p_, q_, t_ = 1 + 0.1 * mk((cout,), 256), 0.05 * mk((cout,), 257), 0.01 * mk((cout,), 258)
    lg = lazygrad(dy, cout, z, p_, q_, t_) if bnz else lazygrad(dy, cout)
    dw = torch.zeros(cin, cout, 2, 2, device="cuda")
    db = torch.zeros(cout, device="cuda")
    run_wgrad(lib, 1, 2, lazy(x, cin, sc, sh, act=2), lg, n, h, w, cin, cout, dw, db)
    a = operand(x, n, h, w, sc, sh, 2)                                              # [n, cin, h, w]
    g = grad_operand(dy, z if bnz else None, n, 2 * h, 2 * w, p_, q_, t_)           # [n, cout, 2h, 2w]
    g6 = g.view(n, cout, h, 2, w, 2)
    ref = torch.einsum("nihw,nohywx->ioyx", a, g6)
    assert rel_err(dw, ref) < DW_TOL
    assert db_err(db, dy, z if bnz else None, n, 2 * h, 2 * w, p_, q_, t_) < DW_TOL
