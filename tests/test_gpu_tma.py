"""The TMA-fed tcgen05 engine (csrc/tconv_tc.cu) on MATERIALISED operands against plain torch fp32 ops on the same
bf16-rounded operands: decoder 3x3 convs and their data gradients at <= 128 px wide rows (im2col by shifted TMA boxes,
zero padding = out-of-bounds fill), deep 1x1 convs, 2x2 transposed convs (scatter epilogue).  Every case asserts
through s2lc_tma_launches() that the call really went through TMA."""
import math

import pytest
import torch
import torch.nn.functional as F

from conftest import rel_err
from test_gpu_kernels import L, lazy, lazygrad, lib, mk, nchw, pack, run_conv  # noqa: F401
from test_gpu_stream_torch import act_bwd, bfr

pytestmark = pytest.mark.gpu
BF = torch.bfloat16
OUT_TOL = 4e-3


@pytest.fixture(autouse=True)
def take_all_shapes(lib):
    """The dispatcher hands the TMA engine only the shapes it wins on (csrc/tconv_tc.cu); the tests also cover the
    thin ones it leaves to the streaming engine in production (debug flag bit 4 = take every plain shape)."""
    old = lib.s2lc_set_debug_flags(0)
    lib.s2lc_set_debug_flags((old | 16) & ~(1 | 2 | 8))
    yield
    lib.s2lc_set_debug_flags(old)


def took_tma(lib, fn):
    lib.s2lc_tma_launches(1)
    fn()
    n = lib.s2lc_tma_launches(0)
    assert n >= 1, "the TMA engine declined this shape"


C3 = [(168, 168, 128, 2, 24, 128), (128, 128, 128, 1, 128, 128), (320, 320, 256, 2, 64, 64), (256, 256, 256, 1, 40, 64),
      (640, 640, 512, 2, 32, 32), (512, 512, 512, 1, 32, 32), (40, 48, 32, 3, 16, 16), (72, 72, 200, 2, 8, 16)]


@pytest.mark.parametrize("cin,cbuf,cout,n,h,w", C3)
def test_tma_conv3x3(lib, cin, cbuf, cout, n, h, w):
    rows = n * h * w
    x = mk((rows, cbuf), 301).to(BF)
    wt = mk((cout, cin, 3, 3), 304, 1 / math.sqrt(9 * cin))
    bias = 0.1 * mk((cout,), 305)
    pw = pack(lib, wt, BF, 1, 1, cout, 9, cin, 0, cin * 9, 1, 9)
    out = torch.zeros(rows, cout, dtype=BF, device="cuda")
    stats = torch.zeros(2 * cout, dtype=torch.float64, device="cuda")
    took_tma(lib, lambda: run_conv(lib, 1, 1, lazy(x, cbuf), n, h, w, cin, pw, out, cout, bias=bias, stats=stats, kind=1))
    ref = F.conv2d(nchw(x, n, h, w)[:, :cin], bfr(wt), bias, padding=1)
    got = nchw(out, n, h, w)
    assert rel_err(got, ref) < OUT_TOL
    assert rel_err(stats[:cout], got.double().sum((0, 2, 3))) < 1e-5
    assert rel_err(stats[cout:], (got.double() ** 2).sum((0, 2, 3))) < 1e-5
    # data gradient from a materialised gradient: act'(prev) epilogue + BNBWD statistics
    g = mk((rows, cout), 306).to(BF)
    sc, sh = 1 + 0.1 * mk((cbuf,), 302), 0.1 * mk((cbuf,), 303)
    ncol = (cin + 7) // 8 * 8
    pwd = pack(lib, wt, BF, 1, 1, cin, 9, cout, 0, 9, 1, cin * 9, flip=True)
    dx = torch.zeros(rows, ncol, dtype=BF, device="cuda")
    st2 = torch.zeros(2 * ncol, dtype=torch.float64, device="cuda")
    took_tma(lib, lambda: run_conv(lib, 1, 1, lazygrad(g, cout), n, h, w, cout, pwd, dx, ncol, nout=ncol,
                                   prev=lazy(x, cbuf, sc, sh, act=2), stats=st2, kind=2))
    u = nchw(x, n, h, w) * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1)
    refx = F.conv_transpose2d(nchw(g, n, h, w), bfr(wt), padding=1) * act_bwd(u[:, :cin], 2)
    gotx = nchw(dx, n, h, w)[:, :cin]
    assert rel_err(gotx, refx) < OUT_TOL
    assert rel_err(st2[:cin], gotx.double().sum((0, 2, 3))) < 1e-5
    assert rel_err(st2[ncol:ncol + cin], (gotx.double() * nchw(x, n, h, w)[:, :cin].double()).sum((0, 2, 3))) < 1e-5


C1 = [(304, 1824, 16, 16, 16), (1824, 304, 16, 16, 16), (3072, 512, 4, 16, 16), (512, 2048, 16, 16, 16),
      (176, 1056, 8, 32, 32), (24, 144, 2, 144, 144), (40, 240, 3, 50, 30), (1056, 176, 2, 9, 9)]


@pytest.mark.parametrize("cin,cout,n,h,w", C1)
def test_tma_conv1x1(lib, cin, cout, n, h, w):
    rows = n * h * w
    x = mk((rows, cin), 311).to(BF)
    wt = mk((cout, cin, 1, 1), 315, 1 / math.sqrt(cin))
    pw = pack(lib, wt, BF, 1, 1, cout, 1, cin, 0, cin, 0, 1)
    out = torch.zeros(rows, cout, dtype=BF, device="cuda")
    stats = torch.zeros(2 * cout, dtype=torch.float64, device="cuda")
    took_tma(lib, lambda: run_conv(lib, 1, 0, lazy(x, cin), n, h, w, cin, pw, out, cout, stats=stats, kind=1))
    ref = F.conv2d(nchw(x, n, h, w), bfr(wt))
    got = nchw(out, n, h, w)
    assert rel_err(got, ref) < OUT_TOL
    assert rel_err(stats[:cout], got.double().sum((0, 2, 3))) < 1e-5
    assert rel_err(stats[cout:], (got.double() ** 2).sum((0, 2, 3))) < 1e-5
    # data gradient cout -> cin from a materialised gradient, residual, SiLU'(prev), BNBWD statistics with row scale
    g = mk((rows, cout), 316).to(BF)
    sc, sh = 1 + 0.1 * mk((cin,), 312), 0.1 * mk((cin,), 313)
    pwd = pack(lib, wt, BF, 1, 1, cin, 1, cout, 0, 1, 0, cin)
    res = mk((rows, cin), 321).to(BF)
    dx = torch.zeros(rows, cin, dtype=BF, device="cuda")
    st2 = torch.zeros(2 * cin, dtype=torch.float64, device="cuda")
    took_tma(lib, lambda: run_conv(lib, 1, 0, lazygrad(g, cout), n, h, w, cout, pwd, dx, cin,
                                   prev=lazy(x, cin, sc, sh, act=1), stats=st2, kind=2, res1=res, ld_res1=cin))
    u = nchw(x, n, h, w) * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1)
    refx = (F.conv_transpose2d(nchw(g, n, h, w), bfr(wt)) + nchw(res, n, h, w)) * act_bwd(u, 1)
    gotx = nchw(dx, n, h, w)
    assert rel_err(gotx, refx) < OUT_TOL
    assert rel_err(st2[:cin], gotx.double().sum((0, 2, 3))) < 1e-5
    assert rel_err(st2[cin:], (gotx.double() * nchw(x, n, h, w).double()).sum((0, 2, 3))) < 1e-5


@pytest.mark.parametrize("cin,cout,n,h,w", [(2048, 512, 4, 16, 16), (512, 256, 2, 32, 32), (128, 64, 2, 128, 128),
                                            (64, 32, 1, 40, 24)])
def test_tma_convT(lib, cin, cout, n, h, w):
    rows = n * h * w
    x = mk((rows, cin), 331).to(BF)
    wt = mk((cin, cout, 2, 2), 332, 1 / math.sqrt(cin))
    bias = 0.1 * mk((cout,), 333)
    pw = pack(lib, wt, BF, 1, 4, cout, 1, cin, 1, 4, 0, cout * 4)
    ccat = cout + 8
    cat = torch.zeros(rows * 4, ccat, dtype=BF, device="cuda")
    took_tma(lib, lambda: run_conv(lib, 1, 2, lazy(x, cin), n, h, w, cin, pw, cat, ccat, bias=bias))
    ref = F.conv_transpose2d(nchw(x, n, h, w), bfr(wt), bias, stride=2)
    got = nchw(cat, n, 2 * h, 2 * w)
    assert rel_err(got[:, :cout], ref) < OUT_TOL
    assert float(got[:, cout:].abs().max()) == 0.0


# ------------------------------------------------------------------------------------------ weight gradients (twgrad_tc.cu)
from test_gpu_kernels import run_wgrad  # noqa: E402

DW_TOL = 1e-3

W3 = [(640, 512, 2, 32, 32), (512, 512, 1, 32, 32), (320, 256, 2, 64, 64), (256, 256, 1, 40, 64), (40, 32, 3, 16, 16),
      (72, 200, 2, 8, 16), (168, 128, 1, 24, 128),
      # thin layers: the three x-shifts of a tap row packed along N
      (32, 32, 1, 70, 256), (40, 32, 2, 12, 128), (64, 64, 1, 72, 256), (24, 128, 2, 16, 64)]


@pytest.mark.parametrize("cin,cout,n,h,w", W3)
def test_tma_wgrad3x3(lib, cin, cout, n, h, w):
    rows = n * h * w
    x = mk((rows, cin), 341).to(BF)
    g = mk((rows, cout), 342).to(BF)
    dw = torch.zeros(cout, cin, 3, 3, device="cuda")
    db = torch.zeros(cout, device="cuda")
    took_tma(lib, lambda: run_wgrad(lib, 1, 1, lazy(x, cin), lazygrad(g, cout), n, h, w, cin, cout, dw, db))
    ref = torch.nn.grad.conv2d_weight(nchw(x, n, h, w), (cout, cin, 3, 3), nchw(g, n, h, w), padding=1)
    assert rel_err(dw, ref) < DW_TOL
    assert rel_err(db, nchw(g, n, h, w).sum((0, 2, 3))) < DW_TOL


W1 = [(304, 1824, 16, 16, 16), (1824, 304, 16, 16, 16), (176, 1056, 8, 32, 32), (512, 2048, 4, 16, 16),
      (128, 768, 4, 32, 32), (40, 240, 3, 50, 30), (1056, 176, 2, 9, 9)]


@pytest.mark.parametrize("cin,cout,n,h,w", W1)
def test_tma_wgrad1x1(lib, cin, cout, n, h, w):
    rows = n * h * w
    x = mk((rows, cin), 351).to(BF)
    g = mk((rows, cout), 352).to(BF)
    dw = torch.zeros(cout, cin, 1, 1, device="cuda")
    took_tma(lib, lambda: run_wgrad(lib, 1, 0, lazy(x, cin), lazygrad(g, cout), n, h, w, cin, cout, dw))
    ref = (g.float().t() @ x.float()).view(cout, cin, 1, 1)
    assert rel_err(dw, ref) < DW_TOL


@pytest.mark.parametrize("cin,cout,n,h,w", [(2048, 512, 4, 16, 16), (512, 256, 2, 32, 32), (256, 128, 2, 64, 64),
                                            (64, 32, 1, 8, 16), (64, 32, 2, 40, 64), (128, 64, 1, 24, 128), (40, 48, 2, 8, 16)])
def test_tma_wgradT(lib, cin, cout, n, h, w):
    rows = n * h * w
    x = mk((rows, cin), 361).to(BF)
    g = mk((rows * 4, cout), 362).to(BF)
    dw = torch.zeros(cin, cout, 2, 2, device="cuda")
    db = torch.zeros(cout, device="cuda")
    took_tma(lib, lambda: run_wgrad(lib, 1, 2, lazy(x, cin), lazygrad(g, cout), n, h, w, cin, cout, dw, db))
    g6 = nchw(g, n, 2 * h, 2 * w).view(n, cout, h, 2, w, 2)
    ref = torch.einsum("nihw,nohywx->ioyx", nchw(x, n, h, w), g6)
    assert rel_err(dw, ref) < DW_TOL
    assert rel_err(db, nchw(g, n, 2 * h, 2 * w).sum((0, 2, 3))) < DW_TOL


@pytest.mark.parametrize("cin,cout,n,h,w", [(2048, 512, 2, 16, 16), (512, 256, 2, 32, 32), (128, 64, 2, 128, 128),
                                            (64, 32, 1, 24, 256), (40, 48, 2, 8, 16)])
def test_tma_convT_dgrad(lib, cin, cout, n, h, w):
    """Data gradient of ConvTranspose2d(k=2, s=2): each input pixel gathers its four children (phases of a 5-D TMA view
    of the 2H x 2W gradient), act'(prev) epilogue, BNBWD statistics."""
    rows = n * h * w
    x = mk((rows, cin), 381).to(BF)
    sc, sh = 1 + 0.1 * mk((cin,), 382), 0.1 * mk((cin,), 383)
    wt = mk((cin, cout, 2, 2), 384, 1 / math.sqrt(4 * cout))
    ccat = cout + 8
    dcat = torch.zeros(rows * 4, ccat, dtype=BF, device="cuda")
    dcat[:, :cout] = mk((rows * 4, cout), 385).to(BF)
    pwd = pack(lib, wt, BF, 1, 1, cin, 4, cout, 0, cout * 4, 1, 4)
    dx = torch.empty(rows, cin, dtype=BF, device="cuda")
    st = torch.zeros(2 * cin, dtype=torch.float64, device="cuda")
    took_tma(lib, lambda: run_conv(lib, 1, 3, lazygrad(dcat, ccat), n, h, w, cout, pwd, dx, cin,
                                   prev=lazy(x, cin, sc, sh, act=2), stats=st, kind=2))
    g = nchw(dcat, n, 2 * h, 2 * w)[:, :cout]
    u = nchw(x, n, h, w) * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1)
    want = F.conv2d(g, bfr(wt), stride=2) * act_bwd(u, 2)     # dgrad of convT(k=2, s=2) = strided conv with the same weight
    gotx = nchw(dx, n, h, w)
    assert rel_err(gotx, want) < OUT_TOL
    assert rel_err(st[:cin], gotx.double().sum((0, 2, 3))) < 1e-5
    assert rel_err(st[cin:], (gotx.double() * nchw(x, n, h, w).double()).sum((0, 2, 3))) < 1e-5
