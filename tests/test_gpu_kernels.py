"""Per-kernel numerics on a real B200: every libs2lc kernel family (called through the C ABI) against the
plain-PyTorch fp32 op it replaces, on seeded inputs.  Tolerances: fp32 storage 2e-5 relative (L2), bf16
storage 1.5e-2 (the output rounding alone is 2^-9 per element)."""
import ctypes as C
import math

import pytest
import torch
import torch.nn.functional as F

from conftest import rel_err

pytestmark = pytest.mark.gpu

DT = {"fp32": (torch.float32, 0, 2e-5), "bf16": (torch.bfloat16, 1, 1.5e-2)}


@pytest.fixture(scope="module")
def lib():
    from s2lc_b200 import _lib
    lib = _lib.load()
    _lib.check(lib.s2lc_arch_check(0))
    return lib


def L():
    from s2lc_b200 import _lib
    return _lib


def nhwc(x, dtype):  # [N,C,H,W] -> [N*H*W, C] contiguous
    return x.permute(0, 2, 3, 1).reshape(-1, x.shape[1]).to(dtype).contiguous()


def nchw(rows, n, h, w):
    return rows.float().view(n, h, w, -1).permute(0, 3, 1, 2)


def stream():
    return torch.cuda.current_stream().cuda_stream


def same_pad(x, k, s):
    h, w = x.shape[-2:]
    oh, ow = math.ceil(h / s), math.ceil(w / s)
    ph, pw = max((oh - 1) * s + k - h, 0), max((ow - 1) * s + k - w, 0)
    return F.pad(x, [pw // 2, pw - pw // 2, ph // 2, ph - ph // 2])


def mk(shape, seed, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(shape, generator=g, device="cuda") * scale


def lazy(t, ld, scale=None, shift=None, gate=None, act=0):
    return L().Lazy(t.data_ptr(), ld, L().ptr(scale), L().ptr(shift), L().ptr(gate), act, 0)


def lazygrad(dy, ld, z=None, p=None, q=None, t=None, rs=None):
    return L().LazyGrad(dy.data_ptr(), ld, L().ptr(z), ld, L().ptr(p), L().ptr(q), L().ptr(t), L().ptr(rs))


# ----------------------------------------------------------------------------------------------- dwconv
@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("k,s,h,w,c", [(3, 1, 24, 40, 48), (5, 1, 16, 16, 72), (3, 2, 32, 32, 144), (5, 2, 28, 28, 64),
                                       (5, 2, 7, 7, 24), (3, 2, 15, 17, 8),
                                       # rows >= 64 px with a last channel block <= half full: 32-channel blocks
                                       (3, 1, 20, 72, 24), (3, 2, 24, 136, 144), (3, 1, 10, 64, 32), (3, 2, 20, 64, 88)])
def test_dwconv_fwd_bwd(lib, prec, k, s, h, w, c):
    dtype, dt, tol = DT[prec]
    n = 2
    x = mk((n, c, h, w), 1)
    wt = mk((c, 1, k, k), 2, 0.3)
    sc, sh = 1 + 0.1 * mk((c,), 3), 0.1 * mk((c,), 4)
    xr = nhwc(x, dtype)
    xq = nchw(xr, n, h, w).requires_grad_(True)          # the stored (rounded) input
    wq = wt.clone().requires_grad_(True)
    a = F.silu(xq * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1))
    if dtype == torch.bfloat16:
        a = a + (a.to(dtype).float() - a).detach()        # the kernel stages activations in bf16
    ref = F.conv2d(same_pad(a, k, s), wq, stride=s, groups=c)
    ho, wo = ref.shape[-2:]
    out = torch.empty(n * ho * wo, c, dtype=dtype, device="cuda")
    stats = torch.zeros(2 * c, dtype=torch.float64, device="cuda")
    lz = lazy(xr, c, sc, sh, act=1)
    L().check(lib.s2lc_dwconv_fwd(C.byref(lz), wt.data_ptr(), out.data_ptr(), stats.data_ptr(), n, h, w, c, k, s, dt,
                                  stream()))
    got = nchw(out, n, ho, wo)
    assert rel_err(got, ref) < tol
    assert rel_err(stats[:c], got.double().sum((0, 2, 3))) < 1e-5
    assert rel_err(stats[c:], (got.double() ** 2).sum((0, 2, 3))) < 1e-5
    # backward: lazy gradient g = p*dy + q*z + t on the output side
    dy = mk((n, c, ho, wo), 5)
    p_, q_, t_ = 1 + 0.1 * mk((c,), 6), 0.05 * mk((c,), 7), 0.01 * mk((c,), 8)
    dyr = nhwc(dy, dtype)
    g_ref = (p_.view(1, -1, 1, 1) * nchw(dyr, n, ho, wo) + q_.view(1, -1, 1, 1) * got + t_.view(1, -1, 1, 1))
    if dtype == torch.bfloat16:
        g_stage = g_ref.to(dtype).float()                 # bwd-data kernel stages the gradient tile in bf16
    else:
        g_stage = g_ref
    gx, = torch.autograd.grad(ref, xq, g_stage, retain_graph=True)   # includes act'(u)*scale
    gw, = torch.autograd.grad(ref, wq, g_ref)
    # kernel returns d/d(BN output) i.e. without the *scale factor
    gx_bnout = gx / sc.view(1, -1, 1, 1)
    lg = lazygrad(dyr, c, out, p_, q_, t_)
    dx = torch.empty(n * h * w, c, dtype=dtype, device="cuda")
    st2 = torch.zeros(2 * c, dtype=torch.float64, device="cuda")
    L().check(lib.s2lc_dwconv_bwd_data(C.byref(lg), wt.data_ptr(), C.byref(lz), None, 0, dx.data_ptr(), c,
                                       st2.data_ptr(), n, h, w, c, k, s, dt, stream()))
    gotx = nchw(dx, n, h, w)
    assert rel_err(gotx, gx_bnout) < tol * 2
    assert rel_err(st2[:c], gotx.double().sum((0, 2, 3))) < 1e-5
    assert rel_err(st2[c:], (gotx.double() * nchw(xr, n, h, w).double()).sum((0, 2, 3))) < 1e-5
    dw = torch.zeros(c, 1, k, k, device="cuda")
    L().check(lib.s2lc_dwconv_bwd_weight(C.byref(lg), C.byref(lz), dw.data_ptr(), n, h, w, c, k, s, dt, stream()))
    assert rel_err(dw, gw) < tol * 2
    # fused streaming backward (one pass: dx + BNBWD statistics + dW) must agree with the same references
    dx2 = torch.empty(n * h * w, c, dtype=dtype, device="cuda")
    st3 = torch.zeros(2 * c, dtype=torch.float64, device="cuda")
    dw2 = torch.zeros(c, 1, k, k, device="cuda")
    L().check(lib.s2lc_dwconv_bwd(C.byref(lg), wt.data_ptr(), C.byref(lz), 1, None, 0, dx2.data_ptr(), c,
                                  st3.data_ptr(), dw2.data_ptr(), n, h, w, c, k, s, dt, stream()))
    gotx2 = nchw(dx2, n, h, w)
    assert rel_err(gotx2, gx_bnout) < tol * 2
    assert rel_err(st3[:c], gotx2.double().sum((0, 2, 3))) < 1e-5
    assert rel_err(st3[c:], (gotx2.double() * nchw(xr, n, h, w).double()).sum((0, 2, 3))) < 1e-5
    assert rel_err(dw2, gw) < tol * 2
    # tile-engine forward (kept for cross-checking) against the streaming forward
    out_t = torch.empty_like(out)
    L().check(lib.s2lc_dwconv_fwd_tile(C.byref(lz), wt.data_ptr(), out_t.data_ptr(), None, n, h, w, c, k, s, dt,
                                       stream()))
    assert rel_err(out_t.float(), out.float()) < tol


@pytest.mark.parametrize("k,s,h,w,c", [(3, 1, 64, 96, 24), (5, 1, 32, 32, 176), (5, 2, 64, 64, 240), (3, 2, 128, 64, 144),
                                       (5, 1, 16, 16, 304), (3, 1, 40, 24, 200)])
def test_dwconv_stream_plain_residual(lib, k, s, h, w, c):
    """Streaming dwconv on strip-spanning shapes: identity input (no BN/act, as in the expand-less blocks), residual
    added to dx, gradient without a BatchNorm behind it but with a per-sample rowscale; bf16 storage."""
    dtype, dt, tol = DT["bf16"]
    n = 3
    xr = nhwc(mk((n, c, h, w), 11), dtype)
    xq = nchw(xr, n, h, w).requires_grad_(True)
    wq = mk((c, 1, k, k), 12, 0.3).requires_grad_(True)
    ref = F.conv2d(same_pad(xq, k, s), wq, stride=s, groups=c)
    ho, wo = ref.shape[-2:]
    out = torch.empty(n * ho * wo, c, dtype=dtype, device="cuda")
    stats = torch.zeros(2 * c, dtype=torch.float64, device="cuda")
    lz = lazy(xr, c)
    L().check(lib.s2lc_dwconv_fwd(C.byref(lz), wq.data_ptr(), out.data_ptr(), stats.data_ptr(), n, h, w, c, k, s, dt,
                                  stream()))
    got = nchw(out, n, ho, wo)
    assert rel_err(got, ref) < tol
    assert rel_err(stats[:c], got.double().sum((0, 2, 3))) < 1e-5
    dyr = nhwc(mk((n, c, ho, wo), 13), dtype)
    rs = torch.tensor([0.5, 1.25, 2.0], device="cuda")
    g_ref = nchw(dyr, n, ho, wo) * rs.view(-1, 1, 1, 1)
    gx, gw = torch.autograd.grad(ref, (xq, wq), g_ref)
    resr = nhwc(mk((n, c, h, w), 14), dtype)
    lg = lazygrad(dyr, c, rs=rs)
    dx = torch.empty(n * h * w, c, dtype=dtype, device="cuda")
    dw = torch.zeros(c, 1, k, k, device="cuda")
    L().check(lib.s2lc_dwconv_bwd(C.byref(lg), wq.data_ptr(), C.byref(lz), 0, resr.data_ptr(), c, dx.data_ptr(), c,
                                  None, dw.data_ptr(), n, h, w, c, k, s, dt, stream()))
    assert rel_err(nchw(dx, n, h, w), gx + nchw(resr, n, h, w)) < tol * 2
    assert rel_err(dw, gw) < tol * 2


@pytest.mark.parametrize("silu", [True, False, None])     # None: plain input WITHOUT a residual
@pytest.mark.parametrize("k,s,n,h,w,c", [(3, 1, 3, 32, 64, 24),     # 32-channel blocks, last block 75 % full
                                         (3, 1, 2, 32, 32, 192),    # 64-channel blocks
                                         (5, 1, 2, 16, 64, 80),     # 5x5, 32-channel blocks
                                         (5, 1, 2, 32, 32, 72),     # 5x5, 64-channel blocks, ragged last block
                                         (3, 2, 2, 64, 128, 48),    # stride 2, 32-channel blocks
                                         (3, 2, 2, 32, 32, 96),     # stride 2, 64-channel blocks
                                         (5, 2, 2, 32, 64, 40),     # 5x5 stride 2
                                         (3, 1, 100, 64, 64, 24),   # more strips than resident CTAs: several per CTA
                                         (5, 1, 40, 64, 64, 136)])
def test_dwconv_bwd_fast(lib, silu, k, s, n, h, w, c):
    """dw2_bwdx_kernel (compile-time geometry / transforms, input tile staged by cp.async) against torch fp32 AND the
    generic dw2_bwd_kernel on the same operands.  silu: BatchNorm + SiLU input with act' applied, lazy BatchNorm-
    backward gradient with a per-sample rowscale; otherwise the plain input + residual of the expand-less blocks."""
    dtype, dt, tol = DT["bf16"]
    xr = nhwc(mk((n, c, h, w), 21), dtype)
    xq = nchw(xr, n, h, w).requires_grad_(True)
    wq = mk((c, 1, k, k), 22, 0.3).requires_grad_(True)
    sc, sh = 1 + 0.1 * mk((c,), 23), 0.1 * mk((c,), 24)
    a = F.silu(xq * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1)) if silu else xq
    ref = F.conv2d(same_pad(a, k, s), wq, stride=s, groups=c)
    ho, wo = ref.shape[-2:]
    dyr = nhwc(mk((n, c, ho, wo), 25), dtype)
    zr = nhwc(mk((n, c, ho, wo), 26), dtype)
    rs = 0.5 + torch.sigmoid(mk((n,), 97))          # seeded: (0.5, 1.5)
    if silu:
        p_, q_, t_ = 1 + 0.1 * mk((c,), 27), 0.05 * mk((c,), 28), 0.01 * mk((c,), 29)
        g_ref = (p_.view(1, -1, 1, 1) * rs.view(-1, 1, 1, 1) * nchw(dyr, n, ho, wo)
                 + q_.view(1, -1, 1, 1) * nchw(zr, n, ho, wo) + t_.view(1, -1, 1, 1))
        lg = lazygrad(dyr, c, zr, p_, q_, t_, rs=rs)
        lz = lazy(xr, c, sc, sh, act=1)
        resr = None
    else:
        g_ref = nchw(dyr, n, ho, wo) * rs.view(-1, 1, 1, 1)
        lg = lazygrad(dyr, c, rs=rs)
        lz = lazy(xr, c)
        resr = nhwc(mk((n, c, h, w), 30), dtype) if silu is False else None
    gx, gw = torch.autograd.grad(ref, (xq, wq), g_ref)
    want = gx / sc.view(1, -1, 1, 1) if silu else (gx + nchw(resr, n, h, w) if resr is not None else gx)

    def run():
        dx = torch.empty(n * h * w, c, dtype=dtype, device="cuda")
        st = torch.zeros(2 * c, dtype=torch.float64, device="cuda")
        dw = torch.zeros(c, 1, k, k, device="cuda")
        L().check(lib.s2lc_dwconv_bwd(C.byref(lg), wq.data_ptr(), C.byref(lz), 1 if silu else 0,
                                      resr.data_ptr() if resr is not None else None, c, dx.data_ptr(), c,
                                      st.data_ptr(), dw.data_ptr(), n, h, w, c, k, s, dt, stream()))
        torch.cuda.synchronize()
        return dx, st, dw

    dx, st, dw = run()
    old = lib.s2lc_set_debug_flags(0)
    lib.s2lc_set_debug_flags(old | 128)                   # DBG_DW_GENERIC: the generic kernel
    try:
        dx_g, st_g, dw_g = run()
    finally:
        lib.s2lc_set_debug_flags(old)
    got = nchw(dx, n, h, w)
    assert rel_err(got, want) < tol * 2
    assert rel_err(dw, gw) < tol * 2
    assert rel_err(st[:c], got.double().sum((0, 2, 3))) < 1e-5
    assert rel_err(st[c:], (got.double() * nchw(xr, n, h, w).double()).sum((0, 2, 3))) < 1e-5
    # same arithmetic up to the order of fp32 sums and the folded 1/2 of the tanh form
    assert rel_err(dx.float(), dx_g.float()) < 4e-3
    assert rel_err(dw, dw_g) < 1e-3
    assert rel_err(st, st_g) < 1e-3


@pytest.mark.parametrize("n,hw,c", [(3, 14 * 10, 48), (2, 32 * 32, 1056), (2, 16 * 16, 3072), (4, 64 * 64, 24)])
def test_se_bwd_apply_packed_is_bit_identical(lib, n, hw, c):
    """se_bwd_apply_p_kernel (bf16 + SiLU: packed fp32 pairs, register double buffer) against the generic kernel on
    the same operands: the rounded gradient must be bit-identical, the BNBWD statistics equal up to the order of
    the atomics."""
    dtype, dt, _ = DT["bf16"]
    z = (mk((n * hw, c), 71) * 1.5).to(dtype)
    da0 = mk((n * hw, c), 72).to(dtype)
    sc, sh = 1 + 0.1 * mk((c,), 73), 0.1 * mk((c,), 74)
    gate, dpool = torch.sigmoid(mk((n, c), 76)), mk((n, c), 75)
    lzg = lazy(z, c, sc, sh, gate, act=1)

    def run(flags):
        old = lib.s2lc_set_debug_flags(0)
        lib.s2lc_set_debug_flags(old | flags)
        try:
            da, st = da0.clone(), torch.zeros(2 * c, dtype=torch.float64, device="cuda")
            L().check(lib.s2lc_se_bwd_apply(da.data_ptr(), c, C.byref(lzg), dpool.data_ptr(), st.data_ptr(), n, hw, c, dt,
                                            stream()))
            torch.cuda.synchronize()
        finally:
            lib.s2lc_set_debug_flags(old)
        return da, st

    da_p, st_p = run(0)
    da_g, st_g = run(1 << 16)                              # DBG_STREAM_GENERIC
    assert torch.equal(da_p, da_g)
    assert rel_err(st_p, st_g) < 1e-6


@pytest.mark.parametrize("silu", [True, False])
@pytest.mark.parametrize("k,s,n,h,w,c", [(3, 1, 3, 32, 64, 24), (3, 1, 2, 32, 32, 192), (5, 1, 2, 16, 64, 80),
                                         (5, 1, 2, 32, 32, 72), (3, 2, 2, 64, 128, 48), (3, 2, 2, 64, 64, 96),
                                         (5, 2, 2, 32, 64, 40), (3, 1, 100, 64, 64, 24), (5, 1, 40, 64, 64, 136)])
def test_dwconv_fwd_fast(lib, silu, k, s, n, h, w, c):
    """dw2_fwdx_kernel (compile-time geometry / transform) against torch fp32 and the generic dw2_fwd_kernel."""
    dtype, dt, tol = DT["bf16"]
    xr = nhwc(mk((n, c, h, w), 41), dtype)
    xq = nchw(xr, n, h, w)
    wq = mk((c, 1, k, k), 42, 0.3)
    sc, sh = 1 + 0.1 * mk((c,), 43), 0.1 * mk((c,), 44)
    a = F.silu(xq * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1)) if silu else xq
    ref = F.conv2d(same_pad(a, k, s), wq, stride=s, groups=c)
    ho, wo = ref.shape[-2:]
    lz = lazy(xr, c, sc, sh, act=1) if silu else lazy(xr, c)

    def run(flags):
        old = lib.s2lc_set_debug_flags(0)
        lib.s2lc_set_debug_flags(old | flags)
        try:
            out = torch.empty(n * ho * wo, c, dtype=dtype, device="cuda")
            st = torch.zeros(2 * c, dtype=torch.float64, device="cuda")
            L().check(lib.s2lc_dwconv_fwd(C.byref(lz), wq.data_ptr(), out.data_ptr(), st.data_ptr(), n, h, w, c, k, s, dt,
                                          stream()))
            torch.cuda.synchronize()
        finally:
            lib.s2lc_set_debug_flags(old)
        return out, st

    out, st = run(0)
    out_g, st_g = run(128)                                # DBG_DW_GENERIC
    got = nchw(out, n, ho, wo)
    assert rel_err(got, ref) < tol
    assert rel_err(st[:c], got.double().sum((0, 2, 3))) < 1e-5
    assert rel_err(st[c:], (got.double() ** 2).sum((0, 2, 3))) < 1e-5
    assert rel_err(out.float(), out_g.float()) < 4e-3
    assert rel_err(st, st_g) < 1e-3


@pytest.mark.parametrize("ratio,bound", [(1.0, 1e-5), (10.0, 2e-4), (100.0, 2e-2)])
def test_batchnorm_moments_with_large_mean(lib, ratio, bound):
    """the producers accumulate one-pass moments (sum, sum of squares; fp32 per thread, fp64 across
    threads), so var = E[x^2] - mean^2 loses (mean/std)^2 of the fp32 partial-sum accuracy.  Measured here on the
    depthwise forward (fp32 storage, so the stored values are exact) for |mean|/std = 1, 10, 100 and bounded; BatchNorm
    inputs in this network are conv outputs with |mean|/std of order 1 (tests/test_gpu_trained.py runs the flat bars
    on trained weights)."""
    n, c, h, w, k = 8, 64, 64, 64, 3
    x = ratio + mk((n, c, h, w), 91)
    wt = torch.zeros(c, 1, k, k, device="cuda")
    wt[:, 0, 1, 1] = 1.0                                    # identity stencil: out == in away from nothing (centre tap)
    xr = nhwc(x, torch.float32)
    out = torch.empty(n * h * w, c, device="cuda")
    stats = torch.zeros(2 * c, dtype=torch.float64, device="cuda")
    lz = lazy(xr, c)
    L().check(lib.s2lc_dwconv_fwd(C.byref(lz), wt.data_ptr(), out.data_ptr(), stats.data_ptr(), n, h, w, c, k, 1, 0,
                                  stream()))
    torch.cuda.synchronize()
    cnt = n * h * w
    mean = stats[:c] / cnt
    var = stats[c:] / cnt - mean * mean
    o = out.double()
    var_ref = o.var(0, unbiased=False)
    err = float(((var - var_ref).abs() / var_ref).max())
    print(f"\n[moments |mean|/std = {ratio:g}] worst relative error of the batch variance: {err:.2e}")
    assert float(((mean - o.mean(0)).abs() / o.mean(0).abs().clamp_min(1e-3)).max()) < 1e-6
    assert err < bound


# ----------------------------------------------------------------------------------------------- implicit GEMM
def pack(lib, wt, dtype, dt, n1, n2, taps, c, s1, s2, st, sc, flip=False, cpad=None):
    from s2lc_b200._engine import PackedWeight
    pw = PackedWeight(wt, dtype, n1, n2, taps, c, s1, s2, st, sc, flip, cpad)
    L().check(lib.s2lc_weight_pack(*pw.args, stream()))
    return pw


def run_conv(lib, dt, mode, a, n, h, w, c, pw, out, ld_out, nout=None, bias=None, prev=None, stats=None, kind=0,
             res1=None, ld_res1=0):
    p = L().ConvParams()
    p.mode, p.dtype, p.N, p.H, p.W, p.C = mode, dt, n, h, w, c
    p.Cpad, p.Kpad, p.Nout, p.Npad = pw.cpad, pw.kpad, (pw.n_out if nout is None else nout), pw.npad
    if isinstance(a, L().LazyGrad):
        p.a_is_grad, p.a_bwd = 1, C.pointer(a)
    else:
        p.a_fwd = C.pointer(a)
    p.w_packed, p.bias, p.out, p.ld_out = pw.buf.data_ptr(), L().ptr(bias), out.data_ptr(), ld_out
    p.res1, p.ld_res1 = L().ptr(res1), ld_res1
    if prev is not None:
        p.prev = C.pointer(prev)
    p.stats, p.stats_kind = L().ptr(stats), kind
    L().check(lib.s2lc_conv_fwd(C.byref(p), stream()))


def run_wgrad(lib, dt, mode, a, g, n, h, w, cin, cout, dw, db=None):
    p = L().WgradParams(mode, dt, n, h, w, cin, cout, 0, C.pointer(a), C.pointer(g), dw.data_ptr(), L().ptr(db))
    L().check(lib.s2lc_conv_wgrad(C.byref(p), stream()))


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("cin,cout,h,w,gated", [(24, 144, 16, 24, False), (144, 40, 12, 12, True), (512, 2048, 7, 7, False),
                                                (32, 4, 32, 32, False), (1056, 176, 8, 8, True)])
def test_conv1x1(lib, prec, cin, cout, h, w, gated):
    dtype, dt, tol = DT[prec]
    n = 2
    x = mk((n, cin, h, w), 11)
    wt = mk((cout, cin, 1, 1), 12, 1 / math.sqrt(cin))
    sc, sh = 1 + 0.1 * mk((cin,), 13), 0.1 * mk((cin,), 14)
    gate = torch.sigmoid(mk((n, cin), 15)) if gated else None
    xr = nhwc(x, dtype)
    xq = nchw(xr, n, h, w).requires_grad_(True)
    wq = (wt.to(dtype).float() if dtype == torch.bfloat16 else wt.clone()).requires_grad_(True)
    a = F.silu(xq * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1))
    if gated:
        a = a * gate.view(n, cin, 1, 1)
    if dtype == torch.bfloat16:
        a = a + (a.to(dtype).float() - a).detach()     # tensor-core operand is rounded to bf16
    ref = F.conv2d(a, wq)
    pw = pack(lib, wt, dtype, dt, 1, cout, 1, cin, 0, cin, 0, 1)
    out = torch.empty(n * h * w, cout, dtype=dtype, device="cuda")
    stats = torch.zeros(2 * cout, dtype=torch.float64, device="cuda")
    lz = lazy(xr, cin, sc, sh, gate, act=1)
    run_conv(lib, dt, 0, lz, n, h, w, cin, pw, out, cout, stats=stats, kind=1)
    got = nchw(out, n, h, w)
    assert rel_err(got, ref) < tol
    assert rel_err(stats[:cout], got.double().sum((0, 2, 3))) < 1e-5
    assert rel_err(stats[cout:], (got.double() ** 2).sum((0, 2, 3))) < 1e-5
    # weight gradient and data gradient (plain dy)
    dy = mk((n, cout, h, w), 16)
    dyr = nhwc(dy, dtype)
    dyq = nchw(dyr, n, h, w)
    gw, ga = torch.autograd.grad(ref, [wq, a], dyq)
    lg = lazygrad(dyr, cout)
    dw = torch.zeros_like(wt)
    run_wgrad(lib, dt, 0, lz, lg, n, h, w, cin, cout, dw)
    assert rel_err(dw, gw) < tol * 2
    pwd = pack(lib, wt, dtype, dt, 1, cin, 1, cout, 0, 1, 0, cin)
    da = torch.empty(n * h * w, cin, dtype=dtype, device="cuda")
    run_conv(lib, dt, 0, lg, n, h, w, cout, pwd, da, cin)
    assert rel_err(nchw(da, n, h, w), ga) < tol * 2


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("cin,cbuf,cout,h,w", [(38, 40, 32, 32, 32), (88, 88, 64, 16, 16), (640, 640, 512, 7, 7),
                                               (64, 64, 64, 14, 14)])
def test_conv3x3(lib, prec, cin, cbuf, cout, h, w):
    dtype, dt, tol = DT[prec]
    n = 2
    x = mk((n, cbuf, h, w), 21)
    x[:, cin:] = 0
    wt = mk((cout, cin, 3, 3), 22, 1 / math.sqrt(9 * cin))
    bias = 0.1 * mk((cout,), 23)
    xr = nhwc(x, dtype)
    xq = nchw(xr, n, h, w)[:, :cin].clone().requires_grad_(True)
    wq = (wt.to(dtype).float() if dtype == torch.bfloat16 else wt.clone()).requires_grad_(True)
    ref = F.conv2d(xq, wq, bias, padding=1)
    pw = pack(lib, wt, dtype, dt, 1, cout, 9, cin, 0, cin * 9, 1, 9, cpad=cbuf)
    out = torch.empty(n * h * w, cout, dtype=dtype, device="cuda")
    stats = torch.zeros(2 * cout, dtype=torch.float64, device="cuda")
    lz = lazy(xr, cbuf)
    run_conv(lib, dt, 1, lz, n, h, w, cbuf, pw, out, cout, bias=bias, stats=stats, kind=1)
    got = nchw(out, n, h, w)
    assert rel_err(got, ref) < tol
    assert rel_err(stats[:cout], got.double().sum((0, 2, 3))) < 1e-5
    dy = mk((n, cout, h, w), 24)
    dyr = nhwc(dy, dtype)
    dyq = nchw(dyr, n, h, w)
    gw, gx = torch.autograd.grad(ref, [wq, xq], dyq)
    lg = lazygrad(dyr, cout)
    dw, db = torch.zeros_like(wt), torch.zeros(cout, device="cuda")
    run_wgrad(lib, dt, 1, lz, lg, n, h, w, cin, cout, dw, db)
    assert rel_err(dw, gw) < tol * 2
    assert rel_err(db, dyq.sum((0, 2, 3))) < tol
    pwd = pack(lib, wt, dtype, dt, 1, cin, 9, cout, 0, 9, 1, cin * 9, flip=True)
    ncol = (cin + 7) // 8 * 8
    dx = torch.zeros(n * h * w, ncol, dtype=dtype, device="cuda")
    run_conv(lib, dt, 1, lg, n, h, w, cout, pwd, dx, ncol)
    assert rel_err(nchw(dx, n, h, w)[:, :cin], gx) < tol * 2


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("cin,cout,h,w", [(64, 32, 16, 16), (2048, 512, 7, 7), (128, 64, 9, 12)])
def test_convT(lib, prec, cin, cout, h, w):
    dtype, dt, tol = DT[prec]
    n = 2
    x = mk((n, cin, h, w), 31)
    wt = mk((cin, cout, 2, 2), 32, 1 / math.sqrt(cin))
    bias = 0.1 * mk((cout,), 33)
    sc, sh = 1 + 0.1 * mk((cin,), 34), 0.1 * mk((cin,), 35)
    xr = nhwc(x, dtype)
    xq = nchw(xr, n, h, w).requires_grad_(True)
    wq = (wt.to(dtype).float() if dtype == torch.bfloat16 else wt.clone()).requires_grad_(True)
    a = F.relu(xq * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1))
    if dtype == torch.bfloat16:
        a = a + (a.to(dtype).float() - a).detach()
    ref = F.conv_transpose2d(a, wq, bias, stride=2)
    pw = pack(lib, wt, dtype, dt, 4, cout, 1, cin, 1, 4, 0, cout * 4)
    ccat = cout + 8
    cat = torch.zeros(n * 4 * h * w, ccat, dtype=dtype, device="cuda")
    lz = lazy(xr, cin, sc, sh, act=2)
    run_conv(lib, dt, 2, lz, n, h, w, cin, pw, cat, ccat, bias=bias)
    got = nchw(cat, n, 2 * h, 2 * w)[:, :cout]
    assert rel_err(got, ref) < tol
    assert float(cat[:, cout:].abs().max()) == 0.0
    dy = mk((n, cout, 2 * h, 2 * w), 36)
    dcat = torch.zeros(n * 4 * h * w, ccat, dtype=dtype, device="cuda")
    dcat[:, :cout] = nhwc(dy, dtype)
    dyq = nchw(dcat, n, 2 * h, 2 * w)[:, :cout]
    gw, ga = torch.autograd.grad(ref, [wq, a], dyq)
    lg = lazygrad(dcat, ccat)
    dw, db = torch.zeros_like(wt), torch.zeros(cout, device="cuda")
    run_wgrad(lib, dt, 2, lz, lg, n, h, w, cin, cout, dw, db)
    assert rel_err(dw, gw) < tol * 2
    assert rel_err(db, dyq.sum((0, 2, 3))) < tol
    pwd = pack(lib, wt, dtype, dt, 1, cin, 4, cout, 0, cout * 4, 1, 4)
    dx = torch.empty(n * h * w, cin, dtype=dtype, device="cuda")
    st = torch.zeros(2 * cin, dtype=torch.float64, device="cuda")
    run_conv(lib, dt, 3, lg, n, h, w, cout, pwd, dx, cin, prev=lz, stats=st, kind=2)
    u = nchw(xr, n, h, w) * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1)
    want = ga * (u > 0)
    gotx = nchw(dx, n, h, w)
    assert rel_err(gotx, want) < tol * 2
    assert rel_err(st[cin:], (gotx.double() * nchw(xr, n, h, w).double()).sum((0, 2, 3))) < 1e-5


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("mode,cin,cout,n,h,w", [(0, 40, 240, 3, 20, 20), (0, 304, 1824, 2, 9, 9), (0, 1824, 304, 2, 9, 9),
                                                 (1, 88, 64, 3, 20, 12), (1, 320, 256, 1, 16, 16), (1, 40, 32, 2, 50, 30),
                                                 (2, 64, 32, 3, 10, 14), (2, 512, 256, 1, 8, 8)])
def test_wgrad_lazy_operands(lib, prec, mode, cin, cout, n, h, w):
    """Weight gradients with BOTH operands lazy (BatchNorm-backward affine + drop-connect row scale on G;
    BatchNorm + SiLU + squeeze-excite gate on A), ragged pixel counts (partial 64-pixel blocks, several slabs)
    and channel counts that need several tensor-core tiles — the tcgen05 wgrad engine's whole feature set."""
    dtype, dt, tol = DT[prec]
    up = 2 if mode == 2 else 1
    x = mk((n, cin, h, w), 51)
    sc, sh = 1 + 0.1 * mk((cin,), 52), 0.1 * mk((cin,), 53)
    gate = torch.sigmoid(mk((n, cin), 54)) if mode == 0 else None
    xr = nhwc(x, dtype)
    a = F.silu(nchw(xr, n, h, w) * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1))
    if gate is not None:
        a = a * gate.view(n, cin, 1, 1)
    dy, z = mk((n, cout, up * h, up * w), 55), mk((n, cout, up * h, up * w), 56)
    dyr, zr = nhwc(dy, dtype), nhwc(z, dtype)
    use_bn = mode != 2                      # the transposed conv feeds the concat buffer: no BatchNorm behind it
    p_, q_, t_ = 1 + 0.1 * mk((cout,), 57), 0.05 * mk((cout,), 58), 0.01 * mk((cout,), 59)
    rs = 0.5 + torch.sigmoid(mk((n,), 97))          # seeded: (0.5, 1.5)
    g = nchw(dyr, n, up * h, up * w) * rs.view(n, 1, 1, 1)
    if use_bn:
        g = p_.view(1, -1, 1, 1) * g + q_.view(1, -1, 1, 1) * nchw(zr, n, up * h, up * w) + t_.view(1, -1, 1, 1)
    if dtype == torch.bfloat16:             # tensor-core operands are rounded to bf16
        a, g = a.to(dtype).float(), g.to(dtype).float()
    if mode == 0:
        want = torch.einsum("nohw,nihw->oi", g, a).view(cout, cin, 1, 1)
    elif mode == 1:
        ap = F.pad(a, [1, 1, 1, 1])
        want = torch.stack([torch.einsum("nohw,nihw->oi", g, ap[:, :, ky:ky + h, kx:kx + w])
                            for ky in range(3) for kx in range(3)], -1).view(cout, cin, 3, 3)
    else:
        want = torch.stack([torch.einsum("nihw,nohw->io", a, g[:, :, qy::2, qx::2])
                            for qy in range(2) for qx in range(2)], -1).view(cin, cout, 2, 2)
    lz = lazy(xr, cin, sc, sh, gate, act=1)
    lg = lazygrad(dyr, cout, zr if use_bn else None, p_ if use_bn else None, q_ if use_bn else None,
                  t_ if use_bn else None, rs)
    dw = torch.zeros_like(want)
    db = torch.zeros(cout, device="cuda") if mode != 0 else None
    run_wgrad(lib, dt, mode, lz, lg, n, h, w, cin, cout, dw, db)
    assert rel_err(dw, want) < (5e-5 if prec == "fp32" else 2e-3)   # bf16: fp32 accumulation of identical operands
    if db is not None:
        # bf16: the kernel sums the fp32 gradient before its bf16 rounding, the reference after it; over the 256
        # pixels of the smallest case that difference alone reached 2.3e-3 for some row scales
        assert rel_err(db, g.sum((0, 2, 3))) < (5e-5 if prec == "fp32" else 5e-3)


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("c,ncls,n,h,w", [(32, 4, 2, 40, 24), (32, 2, 1, 17, 9), (64, 8, 1, 16, 16)])
def test_outconv_fused(lib, prec, c, ncls, n, h, w):
    """Classifier 1x1 conv (+bias) on a lazy BatchNorm+ReLU input: forward, and the one-pass backward
    (dW, db, dz with ReLU', BatchNorm-backward moments of dz) against autograd."""
    dtype, dt, tol = DT[prec]
    z = mk((n, c, h, w), 61)
    wt = mk((ncls, c, 1, 1), 62, 1 / math.sqrt(c))
    bias = 0.1 * mk((ncls,), 63)
    sc, sh = 1 + 0.1 * mk((c,), 64), 0.1 * mk((c,), 65)
    zr = nhwc(z, dtype)
    zq = nchw(zr, n, h, w)
    u = zq * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1)
    a = F.relu(u)
    if dtype == torch.bfloat16:
        a = a.to(dtype).float()
    a.requires_grad_(True)
    wq = wt.clone().requires_grad_(True)
    bq = bias.clone().requires_grad_(True)
    ref = F.conv2d(a, wq, bq)
    rows = n * h * w
    out = torch.empty(rows, ncls, dtype=dtype, device="cuda")
    lz = lazy(zr, c, sc, sh, act=2)
    L().check(lib.s2lc_outconv_fwd(C.byref(lz), wt.data_ptr(), bias.data_ptr(), out.data_ptr(), ncls, rows, c, ncls, dt,
                                   stream()))
    assert rel_err(nchw(out, n, h, w), ref) < tol
    dl = mk((n, ncls, h, w), 66)
    dlr = nhwc(dl, dtype)
    gw, gb, ga = torch.autograd.grad(ref, [wq, bq, a], nchw(dlr, n, h, w))
    dw, db = torch.zeros_like(wt), torch.zeros(ncls, device="cuda")
    dz = torch.empty(rows, c, dtype=dtype, device="cuda")
    st = torch.zeros(2 * c, dtype=torch.float64, device="cuda")
    L().check(lib.s2lc_outconv_bwd(C.byref(lz), dlr.data_ptr(), ncls, wt.data_ptr(), dw.data_ptr(), db.data_ptr(),
                                   dz.data_ptr(), c, st.data_ptr(), rows, c, ncls, dt, stream()))
    want = ga * (u > 0)
    got = nchw(dz, n, h, w)
    assert rel_err(dw, gw) < tol and rel_err(db, gb) < tol
    assert rel_err(got, want) < tol
    assert rel_err(st[:c], got.double().sum((0, 2, 3))) < 1e-5
    assert rel_err(st[c:], (got.double() * zq.double()).sum((0, 2, 3))) < 1e-5


# ----------------------------------------------------------------------------------------------- stem
@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("h,w,co", [(64, 64, 48), (30, 34, 32)])
def test_stem(lib, prec, h, w, co):
    dtype, dt, tol = DT[prec]
    n, cin = 2, 6
    x = mk((n, cin, h, w), 41).requires_grad_(False)
    wt = mk((co, cin, 3, 3), 42, 0.2).requires_grad_(True)
    ref = F.conv2d(same_pad(x, 3, 2), wt, stride=2)
    ho, wo = ref.shape[-2:]
    out = torch.empty(n * ho * wo, co, dtype=dtype, device="cuda")
    stats = torch.zeros(2 * co, dtype=torch.float64, device="cuda")
    L().check(lib.s2lc_stem_conv_fwd(x.data_ptr(), wt.data_ptr(), out.data_ptr(), stats.data_ptr(), n, cin, h, w, co,
                                     dt, stream()))
    got = nchw(out, n, ho, wo)
    assert rel_err(got, ref) < (tol if prec == "fp32" else 4e-3)
    assert rel_err(stats[co:], (got.double() ** 2).sum((0, 2, 3))) < 1e-5
    dy = mk((n, co, ho, wo), 43)
    dyr = nhwc(dy, dtype)
    gw, = torch.autograd.grad(ref, wt, nchw(dyr, n, ho, wo))
    dw = torch.zeros_like(wt)
    lg = lazygrad(dyr, co)
    L().check(lib.s2lc_stem_conv_wgrad(x.data_ptr(), C.byref(lg), dw.data_ptr(), n, cin, h, w, co, dt, stream()))
    assert rel_err(dw, gw) < tol


# ----------------------------------------------------------------------------------------------- BN / SE / misc
def test_bn_finalize_and_backward_coefficients(lib):
    c, n, h, w = 40, 3, 9, 11
    z = mk((n, c, h, w), 51) * 2 + 0.5
    gamma, beta = 1 + 0.1 * mk((c,), 52), 0.1 * mk((c,), 53)
    rm, rv = 0.1 * mk((c,), 54), 1 + 0.1 * mk((c,), 55).abs()
    rm0, rv0 = rm.clone(), rv.clone()
    zz = z.clone().requires_grad_(True)
    y = F.batch_norm(zz, rm0, rv0, gamma, beta, training=True, momentum=0.01, eps=1e-3)
    stats = torch.cat([z.double().sum((0, 2, 3)), (z.double() ** 2).sum((0, 2, 3))]).contiguous()
    outs = [torch.empty(c, device="cuda") for _ in range(4)]
    cnt = float(n * h * w)
    L().check(lib.s2lc_bn_finalize(stats.data_ptr(), gamma.data_ptr(), beta.data_ptr(), cnt, 1e-3, 0.01, rm.data_ptr(),
                                   rv.data_ptr(), *[o.data_ptr() for o in outs], c, stream()))
    scale, shift, mean, invstd = outs
    assert rel_err(z * scale.view(1, -1, 1, 1) + shift.view(1, -1, 1, 1), y) < 1e-5
    assert rel_err(rm, rm0) < 1e-6 and rel_err(rv, rv0) < 1e-6
    dy = mk((n, c, h, w), 56)
    gz, gg, gb = torch.autograd.grad(y, [zz, gamma.requires_grad_(True), beta.requires_grad_(True)], dy, allow_unused=True) \
        if False else (None, None, None)
    zz2 = z.clone().requires_grad_(True)
    g2, b2 = gamma.detach().clone().requires_grad_(True), beta.detach().clone().requires_grad_(True)
    y2 = F.batch_norm(zz2, None, None, g2, b2, training=True, eps=1e-3)
    gz, gg, gb = torch.autograd.grad(y2, [zz2, g2, b2], dy)
    st = torch.cat([dy.double().sum((0, 2, 3)), (dy.double() * z.double()).sum((0, 2, 3))]).contiguous()
    p, q, t, dg, db = [torch.empty(c, device="cuda") for _ in range(5)]
    L().check(lib.s2lc_bn_bwd_finalize(st.data_ptr(), mean.data_ptr(), invstd.data_ptr(), gamma.data_ptr(), cnt,
                                       p.data_ptr(), q.data_ptr(), t.data_ptr(), dg.data_ptr(), db.data_ptr(), c,
                                       stream()))
    got = p.view(1, -1, 1, 1) * dy + q.view(1, -1, 1, 1) * z + t.view(1, -1, 1, 1)
    assert rel_err(got, gz) < 2e-5
    assert rel_err(dg, gg) < 2e-5 and rel_err(db, gb) < 2e-5


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
def test_squeeze_excite(lib, prec):
    dtype, dt, tol = DT[prec]
    n, c, s, h, w = 3, 48, 12, 10, 14
    z = mk((n, c, h, w), 61)
    sc, sh = 1 + 0.1 * mk((c,), 62), 0.1 * mk((c,), 63)
    w1, b1, w2, b2 = mk((s, c), 64, 0.2), 0.1 * mk((s,), 65), mk((c, s), 66, 0.2), 0.1 * mk((c,), 67)
    zr = nhwc(z, dtype)
    zq = nchw(zr, n, h, w).requires_grad_(True)
    ps = [t.clone().requires_grad_(True) for t in (w1, b1, w2, b2)]
    a = F.silu(zq * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1))
    pooled_ref = a.mean((2, 3))
    gate_ref = torch.sigmoid(F.linear(F.silu(F.linear(pooled_ref, ps[0], ps[1])), ps[2], ps[3]))
    out_ref = a * gate_ref.view(n, c, 1, 1)
    pooled, gate = torch.empty(n, c, device="cuda"), torch.empty(n, c, device="cuda")
    lz = lazy(zr, c, sc, sh, act=1)
    L().check(lib.s2lc_se_pool(C.byref(lz), pooled.data_ptr(), n, h * w, c, dt, stream()))
    L().check(lib.s2lc_se_fc_fwd(pooled.data_ptr(), w1.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr(),
                                 gate.data_ptr(), n, c, s, stream()))
    assert rel_err(pooled, pooled_ref) < 1e-5 and rel_err(gate, gate_ref) < 1e-5
    da_t = mk((n, c, h, w), 68)
    dar = nhwc(da_t, dtype)
    daq = nchw(dar, n, h, w)
    grads = torch.autograd.grad(out_ref, [zq, *ps], daq)
    dgate, dpool = torch.empty(n, c, device="cuda"), torch.empty(n, c, device="cuda")
    dws = [torch.empty_like(t) for t in (w1, b1, w2, b2)]
    scratch = torch.empty(n * (c + 2 * s), device="cuda")
    L().check(lib.s2lc_se_bwd_reduce(dar.data_ptr(), c, C.byref(lz), dgate.data_ptr(), n, h * w, c, dt, stream()))
    L().check(lib.s2lc_se_fc_bwd(pooled.data_ptr(), w1.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr(),
                                 dgate.data_ptr(), dpool.data_ptr(), *[d.data_ptr() for d in dws], scratch.data_ptr(),
                                 n, c, s, stream()))
    lzg = lazy(zr, c, sc, sh, gate, act=1)
    st = torch.zeros(2 * c, dtype=torch.float64, device="cuda")
    L().check(lib.s2lc_se_bwd_apply(dar.data_ptr(), c, C.byref(lzg), dpool.data_ptr(), st.data_ptr(), n, h * w, c, dt,
                                    stream()))
    got = nchw(dar, n, h, w)                       # d/d(BN2 output)
    want = grads[0] / sc.view(1, -1, 1, 1)
    assert rel_err(got, want) < tol
    for d, g in zip(dws, grads[1:]):
        assert rel_err(d, g) < 1e-4
    assert rel_err(st[:c], got.double().sum((0, 2, 3))) < 1e-5


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
def test_bn_apply_and_reduce(lib, prec):
    dtype, dt, tol = DT[prec]
    n, c, h, w = 2, 24, 12, 12
    z, res = mk((n, c, h, w), 71), mk((n, c, h, w), 72)
    sc, sh = 1 + 0.1 * mk((c,), 73), 0.1 * mk((c,), 74)
    rs = torch.tensor([0.0, 1.25], device="cuda")
    zr, rr = nhwc(z, dtype), nhwc(res, dtype)
    out = torch.empty_like(zr)
    lz = lazy(zr, c, sc, sh)
    L().check(lib.s2lc_bn_apply(C.byref(lz), rr.data_ptr(), c, rs.data_ptr(), out.data_ptr(), c, n * h * w, c, h * w, dt,
                                stream()))
    want = (nchw(zr, n, h, w) * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1)) * rs.view(n, 1, 1, 1) + nchw(rr, n, h, w)
    assert rel_err(nchw(out, n, h, w), want) < tol
    st = torch.zeros(2 * c, dtype=torch.float64, device="cuda")
    L().check(lib.s2lc_bn_bwd_reduce(rr.data_ptr(), c, zr.data_ptr(), c, rs.data_ptr(), st.data_ptr(), n * h * w, c,
                                     h * w, dt, stream()))
    d = nchw(rr, n, h, w).double() * rs.view(n, 1, 1, 1)
    assert rel_err(st[:c], d.sum((0, 2, 3))) < 1e-6
    assert rel_err(st[c:], (d * nchw(zr, n, h, w).double()).sum((0, 2, 3))) < 1e-6


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("with_z,with_rs", [(True, False), (True, True), (False, True), (False, False)])
def test_lazygrad_apply(lib, prec, with_z, with_rs):
    """g = p*dy*rowscale + q*z + t written once (s2lc_lazygrad_apply), out of place and in place over dy: one fp32
    fma chain + ONE rounding to the storage type -> bit-exact against the same chain in torch for fp32, 1 bf16 ulp
    (2^-8 relative) for bf16."""
    dtype, dt, tol = DT[prec]
    n, c, h, w = 3, 40, 9, 11
    dy, z = nhwc(mk((n, c, h, w), 91), dtype), nhwc(mk((n, c, h, w), 92), dtype)
    p_, q_, t_ = 1 + 0.1 * mk((c,), 93), 0.05 * mk((c,), 94), 0.01 * mk((c,), 95)
    rs = torch.tensor([0.0, 1.25, 0.8], device="cuda") if with_rs else None
    g = lazygrad(dy, c, z, p_, q_, t_, rs) if with_z else lazygrad(dy, c, rs=rs)
    out = torch.empty_like(dy)
    L().check(lib.s2lc_lazygrad_apply(C.byref(g), out.data_ptr(), c, n * h * w, c, h * w, dt, stream()))
    d = dy.float().view(n, h * w, c)
    if with_rs:
        d = d * rs.view(n, 1, 1)
    want = d
    if with_z:
        want = torch.addcmul(torch.addcmul(t_.view(1, 1, c).expand_as(d), q_.view(1, 1, c), z.float().view(n, h * w, c)),
                             p_.view(1, 1, c), d)
    want = want.reshape(n * h * w, c)
    assert rel_err(out.float(), want) <= (1e-6 if prec == "fp32" else 2 ** -8)
    assert float((out.float() - want).abs().max()) <= (1e-5 if prec == "fp32" else 2 ** -7 * float(want.abs().max()))
    keep = dy.clone()
    L().check(lib.s2lc_lazygrad_apply(C.byref(g), dy.data_ptr(), c, n * h * w, c, h * w, dt, stream()))   # in place
    assert torch.equal(dy, out)
    del keep


def test_loss_against_golden(golden_arrays):
    from s2lc_b200.losses import CrossEntropyLoss, FocalLoss
    from s2lc_b200 import synthetic
    alpha = torch.tensor(synthetic.CLASS_DISTRIBUTION)
    alpha[1:] = 1 - alpha[1:]
    y = torch.from_numpy(golden_arrays["loss/y"]).cuda()
    cases = {"focal_g2": FocalLoss(alpha, 2.0, 0.0, ignore_index=0),
             "focal_g0.5_noignore": FocalLoss(torch.ones(4), 0.5, 0.0),
             "focal_g2_sum": FocalLoss(alpha, 2.0, 0.0, ignore_index=0, reduce_type="sum"),
             "ce_weighted": CrossEntropyLoss(alpha, ignore_index=0), "ce_plain": CrossEntropyLoss()}
    for nm, fn in cases.items():
        lg = torch.from_numpy(golden_arrays["loss/logits"]).cuda().requires_grad_(True)
        v = fn(lg, y)
        v.backward()
        want_v = float(golden_arrays[f"loss/{nm}/value"])
        assert abs(float(v) - want_v) <= 2e-5 * max(1.0, abs(want_v)), nm
        assert rel_err(lg.grad.cpu(), torch.from_numpy(golden_arrays[f"loss/{nm}/dlogits"])) < 1e-4, nm


def test_loss_label_smoothing_matches_torch():
    from s2lc_b200.losses import CrossEntropyLoss, FocalLoss
    lg = (2 * mk((2, 4, 16, 16), 81))
    y = torch.randint(0, 4, (2, 16, 16), device="cuda", generator=torch.Generator(device="cuda").manual_seed(82))
    wts = torch.tensor([0.2, 0.5, 0.9, 0.7], device="cuda")
    a = lg.clone().requires_grad_(True)
    b = lg.clone().requires_grad_(True)
    v = CrossEntropyLoss(wts, label_smoothing=0.1, ignore_index=1)(a, y)
    r = F.cross_entropy(b, y, weight=wts, label_smoothing=0.1, ignore_index=1)
    v.backward()
    r.backward()
    assert abs(float(v) - float(r)) < 1e-5 and rel_err(a.grad, b.grad) < 1e-4
    a.grad = b.grad = None
    v = FocalLoss(wts, 1.5, 0.05, ignore_index=1)(a, y)
    ce = F.cross_entropy(b, y, label_smoothing=0.05, ignore_index=1, reduction="none")
    r = (wts[y] * (1 - torch.exp(-ce)) ** 1.5 * ce).mean()
    v.backward()
    r.backward()
    assert abs(float(v) - float(r)) < 1e-5 and rel_err(a.grad, b.grad) < 1e-4
