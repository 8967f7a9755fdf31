"""Launch-list engine: the EfficientNet-UNet forward pass and its hand-written reverse pass expressed as
two flat lists of libs2lc kernel launches over pre-allocated HBM buffers.

Design (B200-first):
  * activations live in HBM as NHWC bf16 (production) or fp32 (verification); only RAW conv outputs are
    stored — BatchNorm(train) + activation (+ squeeze-excite gate) are applied by the CONSUMER kernel on load
    ("lazy" tensors, include/s2lc.h) and the per-channel moments are produced by the PRODUCER's epilogue;
  * the backward mirrors it: a gradient w.r.t. a raw conv output is the lazy affine p*dy + q*z + t;
  * every buffer, ctypes struct and argument tuple is built once per (batch, H, W, dtype); a step is then a
    plain loop of ~10^3 C calls with no allocation, no host sync and no torch ops (CUDA-graph capturable);
  * parameter gradients are written straight into one flat fp32 buffer laid out in reverse execution order,
    so the data-parallel trainer can all-reduce bucket k while bucket k+1 is still being computed.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from . import _lib as L
from .efficientnet_unet import BlockSpec, EfficientnetUnet

_DT = {torch.float32: L.F32, torch.bfloat16: L.BF16}
# A/B switch for the decoder's materialised BatchNorm-backward gradients (see Workspace.materialize_grad)
MATERIALIZE_DECODER_GRADS = os.environ.get("S2LC_MATERIALIZE", "1") != "0"
# Operands of the compute-heavy contractions are written ONCE as plain bf16 tensors so that the TMA-fed engine
# (csrc/tconv_tc.cu) takes them; S2LC_NO_TMA=1 keeps every operand lazy (the loader-warp engines, for A/B runs).
TMA_OPERANDS = os.environ.get("S2LC_NO_TMA", "0") != "1"


# MBConv blocks up to this many pixels per sample get their project-conv operand / BatchNorm-backward gradients
# materialised (A/B knob: S2LC_TMA_BLOCK_HW; 128 x 128 measured 0.8 ms per step better than 64 x 64; 256 x 256 was a loss
# until the materialisation passes reached 3.7-5.8 TB/s and is 0.3 ms better since: profiles/r02s_bench_hw256.json)
TMA_BLOCK_HW = int(os.environ.get("S2LC_TMA_BLOCK_HW", str(256 * 256)))


def tma_takes_3x3(width: int, c: int) -> bool:
    """Mirror of the shape rule in csrc/tconv_tc.cu launch_tconv_tc."""
    return width <= 128 or (width == 256 and c > 64)


def _up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class PackedWeight:
    """A conv weight re-laid-out as the [Npad, Kpad] K-major B operand of the implicit GEMM."""

    def __init__(self, src: torch.Tensor, dtype: torch.dtype, n1: int, n2: int, taps: int, c: int, s1: int, s2: int,
                 st: int, sc: int, flip: bool = False, cpad: int | None = None) -> None:
        self.src = src
        self.dtype = dtype
        self.n_out = n1 * n2
        self.npad = _up(self.n_out, 16)
        self.cpad = _up(c, 8) if cpad is None else cpad
        self.kpad = _up(taps * self.cpad, 64)
        self.buf = torch.zeros(self.npad, self.kpad, dtype=dtype, device=src.device)
        self.geom = (n1, n2, taps, c, s1, s2, st, sc, bool(flip))
        self.args = (src.data_ptr(), self.buf.data_ptr(), _DT[dtype], n1, n2, taps, c, self.npad, self.cpad,
                     self.kpad, s1, s2, st, sc, int(flip))

    def entry(self, flat_base_ptr: int) -> "L.PackEntry":
        """This pack job as a row of the fused Adam + re-pack table (include/s2lc.h s2lc_pack_entry): source strides
        in descending order with their roles (0 = n1, 1 = n2, 2 = tap, 3 = c); extent-1 dimensions get stride 0."""
        n1, n2, taps, c, s1, s2, st, sc, flip = self.geom
        dims = [(s1 if n1 > 1 else 0, 0), (s2 if n2 > 1 else 0, 1), (st if taps > 1 else 0, 2), (sc if c > 1 else 0, 3)]
        dims.sort(key=lambda d: -d[0])
        e = L.PackEntry()
        off = self.src.data_ptr() - flat_base_ptr
        assert off >= 0 and off % 4 == 0
        e.start, e.numel, e.dst = off // 4, self.src.numel(), self.buf.data_ptr()
        for i, (stride, role) in enumerate(dims):
            e.stride[i], e.role[i] = stride, role
        e.N2, e.Tt, e.Cpad, e.Kpad, e.flip, e.dtype = n2, taps, self.cpad, self.kpad, int(flip), _DT[self.dtype]
        return e


class Engine:
    def __init__(self, model: EfficientnetUnet, dry_run: bool = False) -> None:
        """``dry_run`` builds buffers and launch lists on whatever device the parameters live on WITHOUT touching
        the GPU (host-logic tests on a CPU-only box); such an engine can never execute a launch list."""
        self.model = model
        self.lib = L.load()
        self.dry_run = dry_run
        dev = next(model.parameters()).device
        if not dry_run:
            if dev.type != "cuda":
                raise RuntimeError("move the model to a CUDA device before the first forward")
            L.check(self.lib.s2lc_arch_check(dev.index if dev.index is not None else torch.cuda.current_device()))
        self.device = dev
        self.used = model.used_parameters()
        # flat gradient buffer, reverse execution order (decoder first), each tensor 16 B aligned
        self.grad_offsets: dict[str, int] = {}
        off = 0
        for k, p in reversed(self.used):
            self.grad_offsets[k] = off
            off += _up(p.numel(), 4)
        self.flat_grad = torch.zeros(off, dtype=torch.float32, device=dev)
        self.grad_views = {k: self.flat_grad[o:o + p.numel()].view_as(p)
                           for (k, p), o in ((kp, self.grad_offsets[kp[0]]) for kp in self.used)}
        self.workspaces: dict[tuple, list[Workspace]] = {}
        self.packs: dict[torch.dtype, dict[str, PackedWeight]] = {}
        self._pack_sig: dict[torch.dtype, tuple] = {}
        self._param_ptrs = self._ptr_signature()
        self.weight_gen = 0        # bumped by every out-of-band parameter update (FusedAdam writes through raw pointers)
        self.stats_gen = 0         # bumped by every training-mode forward (BatchNorm running statistics change)
        self.packs_version = 0     # bumped whenever a packed weight is added (FusedAdam rebuilds its table)
        self.direct_grad = False   # DataParallelTrainer: leave grads in flat_grad instead of returning clones
        self.timing_hook = None    # {"fwd": [], "bwd": [], "only": {kernel-class labels}} — see Workspace.run_forward

    def _ptr_signature(self) -> int:
        return hash(tuple(p.data_ptr() for _, p in self.model.named_parameters()))

    # ---------------------------------------------------------------- packed weights
    def pack(self, dtype: torch.dtype, key: str, factory) -> PackedWeight:
        d = self.packs.setdefault(dtype, {})
        if key not in d:
            d[key] = factory()
            self._pack_sig.pop(dtype, None)
            self.packs_version += 1
        return d[key]

    def weight_signature(self) -> tuple:
        """(generation of out-of-band updates, sum of autograd versions): changes whenever any parameter changed."""
        return (self.weight_gen, sum(p._version for _, p in self.used))

    def repack(self, dtype: torch.dtype, stream: int) -> None:
        """Refresh the packed copies of ``dtype`` unless they were made from the current parameter values (tracked
        PER DTYPE: a bf16 forward must not mark stale fp32 copies clean)."""
        sig = self.weight_signature()
        if self._pack_sig.get(dtype) == sig:
            return
        fn = self.lib.s2lc_weight_pack
        for pw in self.packs.get(dtype, {}).values():
            L.check(fn(*pw.args, stream))
        self._pack_sig[dtype] = sig

    # ---------------------------------------------------------------- workspaces
    def workspace(self, x: torch.Tensor, dtype: torch.dtype, training: bool, need_grad: bool) -> "Workspace":
        if self._ptr_signature() != self._param_ptrs:   # parameters were re-allocated (.to(), load with assign)
            self.workspaces.clear()
            self.packs.clear()
            self._pack_sig.clear()
            self._param_ptrs = self._ptr_signature()
        key = (tuple(x.shape), dtype, training, need_grad)
        pool = self.workspaces.setdefault(key, [])
        for ws in pool:
            if not ws.in_flight:
                return ws
        if len(pool) >= 4:
            raise RuntimeError(
                "four forward passes of one shape are waiting for their backward: activations are saved per workspace "
                "and are not recycled while in flight — call backward() (or run the forward under torch.no_grad())")
        ws = Workspace(self, x.shape[0], x.shape[2], x.shape[3], dtype, training, need_grad)
        pool.append(ws)
        return ws


class _BN:
    def __init__(self, ws: "Workspace", mod: torch.nn.BatchNorm2d, count: int) -> None:
        c = mod.num_features
        self.mod, self.C, self.count = mod, c, float(count)
        self.scale, self.shift = ws.f32(c), ws.f32(c)
        if ws.training:
            self.mean, self.invstd = ws.f32(c), ws.f32(c)
            self.st_f = ws.stats_slot(c, fwd=True)
        if ws.need_grad:
            self.p, self.q, self.t = ws.f32(c), ws.f32(c), ws.f32(c)
            self.st_b = ws.stats_slot(c, fwd=False)


class Workspace:
    """All buffers + the forward / backward launch lists for one (batch, H, W, dtype, mode)."""

    def __init__(self, eng: Engine, n: int, h: int, w: int, dtype: torch.dtype, training: bool,
                 need_grad: bool) -> None:
        self.eng, self.lib, self.model = eng, eng.lib, eng.model
        self.N, self.H, self.W, self.dtype = n, h, w, dtype
        self.dt = _DT[dtype]
        self.es = 2 if dtype == torch.bfloat16 else 4
        self.training, self.need_grad = training, need_grad
        self.dev = eng.device
        self.in_flight = False
        self.lease_id = 0
        self._keep: list = []
        self._tensors: list[torch.Tensor] = []
        self._stats_f: list[int] = []
        self._stats_b: list[int] = []
        self._stats_f_total = 0
        self._stats_b_total = 0
        self.fwd_list: list[tuple] = []
        self.bwd_list: list[tuple] = []
        self.fwd_meta: list[list] = []     # per launch: [label, algorithmic bytes, flops]
        self.bwd_meta: list[list] = []
        self._grad_write_idx: dict[str, int] = {}   # parameter key -> index of the backward launch that writes it
        self._bucket_sched: dict[int, dict[int, list[int]]] = {}
        self.x_ptr = C.c_void_p(0)
        self.dlogits_ptr_holder: L.LazyGrad | None = None
        self.n_bn_trained: list[torch.Tensor] = []
        self.affine_list: list[tuple] = []          # eval mode: the s2lc_bn_eval_affine launches (see bn_forward)
        self._affine_buffers: list[torch.Tensor] = []
        self._affine_sig: tuple | None = None
        self._build()

    # ---------------------------------------------------------------- allocation helpers
    def act(self, rows: int, c: int) -> torch.Tensor:
        t = torch.empty(rows, c, dtype=self.dtype, device=self.dev)
        self._tensors.append(t)      # launch lists hold raw pointers: every buffer must outlive them
        return t

    def f32(self, n: int) -> torch.Tensor:
        t = torch.zeros(n, dtype=torch.float32, device=self.dev)
        self._tensors.append(t)
        return t

    def stats_slot(self, c: int, fwd: bool) -> int:
        """Offset (in doubles) of a [2*c] slot inside the forward / backward fp64 statistics arena."""
        if fwd:
            off = self._stats_f_total
            self._stats_f_total += 2 * c
        else:
            off = self._stats_b_total
            self._stats_b_total += 2 * c
        return off

    def lazy(self, t, ld: int, scale=None, shift=None, gate=None, act: int = L.ACT_NONE, col0: int = 0) -> L.Lazy:
        base = t if isinstance(t, int) else t.data_ptr()
        s = L.Lazy(base + col0 * self.es, ld, L.ptr(scale), L.ptr(shift), L.ptr(gate), act, 0)
        self._keep.append(s)
        return s

    def lazygrad(self, dy, ld_dy: int, z=None, ld_z: int = 0, bn: _BN | None = None, rowscale=None,
                 col0: int = 0) -> L.LazyGrad:
        base = dy if isinstance(dy, int) else dy.data_ptr()
        g = L.LazyGrad(base + col0 * self.es, ld_dy, L.ptr(z), ld_z, L.ptr(bn.p) if bn else None,
                       L.ptr(bn.q) if bn else None, L.ptr(bn.t) if bn else None,
                       rowscale if isinstance(rowscale, (int, type(None))) else rowscale.data_ptr())
        self._keep.append(g)
        return g

    def materialize_grad(self, g: L.LazyGrad, dy, ld: int, rows: int, c: int, hw: int) -> L.LazyGrad:
        """g = p*dy + q*z + t written once over ``dy`` (s2lc_lazygrad_apply) and handed on as a plain gradient: the
        decoder 3x3 layers cut their channels over many CTAs in both gradient engines, each of which would redo the
        transform on its copy of the operand."""
        self.B("s2lc_lazygrad_apply", C.byref(g), dy.data_ptr(), ld, rows, c, hw, self.dt, nbytes=3 * rows * c * self.es)
        return self.lazygrad(dy, ld)

    def materialize_act(self, a: L.Lazy, rows: int, c: int, hw: int, desc: str = "") -> L.Lazy:
        """act(BatchNorm(z)) (* squeeze-excite gate) written once as a plain bf16 tensor (s2lc_bn_apply): the operand
        of a TMA-fed contraction in the forward pass and of its weight gradient in the backward pass."""
        buf = self.act(rows, c)
        self.F("s2lc_bn_apply", C.byref(a), None, 0, None, buf.data_ptr(), c, rows, c, hw, self.dt,
               nbytes=2 * rows * c * self.es)
        self.fwd_meta[-1][0] = "operand_apply"
        self.fwd_meta[-1][3] = desc
        return self.lazy(buf, c)

    def F(self, name: str, *args, nbytes: float = 0, flops: float = 0) -> None:
        self.fwd_list.append((getattr(self.lib, name), args))
        self.fwd_meta.append([name[5:], nbytes, flops, ""])

    def B(self, name: str, *args, nbytes: float = 0, flops: float = 0) -> None:
        self.bwd_list.append((getattr(self.lib, name), args))
        self.bwd_meta.append([name[5:], nbytes, flops, ""])

    def conv(self, lst, mode: int, a, n: int, h: int, w: int, c: int, pw: PackedWeight, out, ld_out: int, *, nout=None,
             bias=None, res1=None, ld_res1: int = 0, res2=None, ld_res2: int = 0, prev=None, stats=None,
             stats_kind: int = L.STATS_NONE, stats_rowscale=None, out_col0: int = 0) -> None:
        p = L.ConvParams()
        p.mode, p.dtype, p.N, p.H, p.W, p.C = mode, self.dt, n, h, w, c
        p.Cpad, p.Kpad, p.Nout, p.Npad = pw.cpad, pw.kpad, (pw.n_out if nout is None else nout), pw.npad
        if isinstance(a, L.LazyGrad):
            p.a_is_grad, p.a_bwd = 1, C.pointer(a)
        else:
            p.a_is_grad, p.a_fwd = 0, C.pointer(a)
        p.w_packed = pw.buf.data_ptr()
        p.bias = L.ptr(bias)
        p.out = (out if isinstance(out, int) else out.data_ptr()) + out_col0 * self.es
        p.ld_out = ld_out
        p.res1 = res1 if isinstance(res1, (int, type(None))) else res1.data_ptr()
        p.ld_res1 = ld_res1
        p.res2 = res2 if isinstance(res2, (int, type(None))) else res2.data_ptr()
        p.ld_res2 = ld_res2
        if prev is not None:
            p.prev = C.pointer(prev)
        if isinstance(stats, tuple):
            p._stats_marker = stats      # resolved to an arena pointer in _resolve()
        else:
            p.stats = stats
        p.stats_kind = stats_kind
        p.stats_rowscale = stats_rowscale
        self._keep.append(p)
        lst.append((self.lib.s2lc_conv_fwd, (C.byref(p),)))
        # algorithmic traffic: every distinct operand once; flops = 2*M*N*K of the unpadded problem
        rows, es = n * h * w, self.es
        taps = {L.CONV_1X1: 1, L.CONV_3X3: 9, L.CONVT_FWD: 1, L.CONVT_DGRAD: 4}[mode]
        src_rows = rows * (4 if mode == L.CONVT_DGRAD else 1)
        a_bytes = src_rows * c * es * (2 if isinstance(a, L.LazyGrad) and a.z else 1)
        o_bytes = rows * p.Nout * es
        extra = (1 if prev is not None else 0) + (1 if res1 is not None else 0) + (1 if res2 is not None else 0)
        nbytes = a_bytes + o_bytes * (1 + extra) + pw.n_out * taps * c * es
        label = {L.CONV_1X1: "conv1x1", L.CONV_3X3: "conv3x3", L.CONVT_FWD: "convT", L.CONVT_DGRAD: "convT_dgrad"}[mode]
        if lst is self.bwd_list and mode != L.CONVT_DGRAD:
            label += "_dgrad"
        (self.fwd_meta if lst is self.fwd_list else self.bwd_meta).append(
            [label, nbytes, 2.0 * rows * taps * c * p.Nout, f"{n}x{h}x{w} {c}->{p.Nout}"])

    def wgrad(self, mode: int, a: L.Lazy, g: L.LazyGrad, n: int, h: int, w: int, cin: int, cout: int, dw: int,
              dbias: int | None = None) -> None:
        p = L.WgradParams(mode, self.dt, n, h, w, cin, cout, 0, C.pointer(a), C.pointer(g), dw, dbias)
        self._keep.append(p)
        self.bwd_list.append((self.lib.s2lc_conv_wgrad, (C.byref(p),)))
        rows, es = n * h * w, self.es
        taps = 9 if mode == L.CONV_3X3 else 1
        grp = 4 if mode == L.CONVT_FWD else 1
        nbytes = rows * cin * es + rows * grp * cout * es * (2 if g.z else 1) + 4 * taps * grp * cin * cout
        label = {L.CONV_1X1: "conv1x1_wgrad", L.CONV_3X3: "conv3x3_wgrad", L.CONVT_FWD: "convT_wgrad"}[mode]
        self.bwd_meta.append([label, nbytes, 2.0 * rows * taps * grp * cin * cout, f"{n}x{h}x{w} {cin}->{cout}"])

    def gptr(self, key: str) -> int:
        self._grad_write_idx[key] = len(self.bwd_list)   # the launch appended next writes this gradient
        return self.eng.grad_views[key].data_ptr()

    def meta(self, lst_meta: list, nbytes: float, flops: float = 0.0, desc: str = "") -> None:
        lst_meta[-1][1], lst_meta[-1][2], lst_meta[-1][3] = nbytes, flops, desc

    # ---------------------------------------------------------------- BatchNorm helpers
    def bn_forward(self, bn: _BN) -> None:
        m = bn.mod
        if self.training:
            self.F("s2lc_bn_finalize", ("SF", bn.st_f), m.weight.data_ptr(), m.bias.data_ptr(), bn.count, float(m.eps),
                   float(m.momentum), m.running_mean.data_ptr(), m.running_var.data_ptr(), bn.scale.data_ptr(),
                   bn.shift.data_ptr(), bn.mean.data_ptr(), bn.invstd.data_ptr(), bn.C)
            self.n_bn_trained.append(m.num_batches_tracked)
        else:
            # eval mode: scale / shift depend on parameters and running statistics only — computed once and again only
            # when those change (run_forward), not 126 times per forward
            self.affine_list.append((self.lib.s2lc_bn_eval_affine,
                                     (m.weight.data_ptr(), m.bias.data_ptr(), m.running_mean.data_ptr(),
                                      m.running_var.data_ptr(), float(m.eps), bn.scale.data_ptr(), bn.shift.data_ptr(),
                                      bn.C)))
            self._affine_buffers += [m.running_mean, m.running_var]

    def bn_backward(self, bn: _BN, key: str) -> None:
        m = bn.mod
        self.B("s2lc_bn_bwd_finalize", ("SB", bn.st_b), bn.mean.data_ptr(), bn.invstd.data_ptr(), m.weight.data_ptr(),
               bn.count, bn.p.data_ptr(), bn.q.data_ptr(), bn.t.data_ptr(), self.gptr(key + ".weight"),
               self.gptr(key + ".bias"), bn.C)

    def stats_f(self, bn: _BN):
        return ("SF", bn.st_f) if self.training else None

    # ---------------------------------------------------------------- build
    def _build(self) -> None:
        m, eng = self.model, self.eng
        cfg = m.config
        enc = m.encoder
        specs: list[BlockSpec] = enc.specs
        N, H, W, dt = self.N, self.H, self.W, self.dtype
        train, grad = self.training, self.need_grad
        nb = len(specs)
        tma = TMA_OPERANDS and dt == torch.bfloat16

        def pk(key, factory):
            return eng.pack(dt, key, factory)

        def pack_1x1(key, wt):  # [O, I, 1, 1]
            o, i = wt.shape[0], wt.shape[1]
            return pk(key + "/f", lambda: PackedWeight(wt, dt, 1, o, 1, i, 0, i, 0, 1))

        def pack_1x1_d(key, wt):
            o, i = wt.shape[0], wt.shape[1]
            return pk(key + "/d", lambda: PackedWeight(wt, dt, 1, i, 1, o, 0, 1, 0, i))

        def cpad3(c, width):
            """Per-tap K stride of a packed 3x3 weight: whole 64-channel chunks where the streaming engine
            (csrc/sconv_tc.cu: rows that are a multiple of 128 px wide, <= 128 channels) can take the layer."""
            return _up(c, 64) if (width % 128 == 0 and c <= 128) else _up(c, 8)

        def pack_3x3(key, wt, cbuf, width):  # [O, I, 3, 3]
            o, i = wt.shape[0], wt.shape[1]
            cp = cpad3(cbuf, width)
            return pk(f"{key}/f{cp}", lambda: PackedWeight(wt, dt, 1, o, 9, i, 0, i * 9, 1, 9, cpad=cp))

        def pack_3x3_d(key, wt, n_in, width):
            o, i = wt.shape[0], wt.shape[1]
            cp = cpad3(o, width)
            return pk(f"{key}/d{cp}", lambda: PackedWeight(wt, dt, 1, n_in, 9, o, 0, 9, 1, i * 9, flip=True, cpad=cp))

        def pack_ct(key, wt):  # [I, O, 2, 2]
            i, o = wt.shape[0], wt.shape[1]
            return pk(key + "/f", lambda: PackedWeight(wt, dt, 4, o, 1, i, 1, 4, 0, o * 4))

        def pack_ct_d(key, wt):
            i, o = wt.shape[0], wt.shape[1]
            return pk(key + "/d", lambda: PackedWeight(wt, dt, 1, i, 4, o, 0, o * 4, 1, 4))

        # ---- drop-connect factors [n_blocks, N] (1.0 when disabled / eval)
        use_dc = train and cfg.drop_connect_rate is not None and cfg.drop_connect_rate > 0
        self.dc = torch.ones(nb, N, dtype=torch.float32, device=self.dev) if use_dc else None
        self.dc_step = C.c_uint64(0)
        if use_dc:
            self.fwd_list.append(("DC", ()))   # resolved in run_forward (internal generator or injected masks)
            self.fwd_meta.append(["dropconnect_masks", 0, 0, ""])

        def dc_ptr(i: int):
            if self.dc is None or not specs[i].residual:
                return None
            return self.dc.data_ptr() + 4 * i * N

        # ================================================================== FORWARD
        h0, w0 = H // 2, W // 2
        stem_conv, stem_bn_mod = enc.stem[0], enc.stem[1]
        c0 = stem_conv.out_channels
        z0 = self.act(N * h0 * w0, c0)
        bn0 = _BN(self, stem_bn_mod, N * h0 * w0)
        k_stem = 9 * cfg.in_channels
        stem_tc = tma and k_stem <= 64
        if stem_tc:
            # bf16 production path: im2col once (kept for the weight gradient), then a K = 56 contraction on the
            # tensor-core engines; the direct CUDA-core kernels below ran at 0.43 TB/s (0.93 + 1.08 ms per step)
            self.stem_cols = self.act(N * h0 * w0, 64)
            self.F("s2lc_stem_im2col", self.x_ptr, self.stem_cols.data_ptr(), 64, N, cfg.in_channels, H, W, self.dt,
                   nbytes=N * cfg.in_channels * H * W * 4 + N * h0 * w0 * 64 * self.es)
            self.stem_lazy = self.lazy(self.stem_cols, 64)
            w_stem = pk("encoder.stem.0/f", lambda: PackedWeight(stem_conv.weight, dt, 1, c0, 1, k_stem, 0, k_stem, 0, 1))
            self.conv(self.fwd_list, L.CONV_1X1, self.stem_lazy, N, h0, w0, _up(k_stem, 8), w_stem, z0, c0,
                      stats=self.stats_f(bn0), stats_kind=L.STATS_MOMENTS if train else L.STATS_NONE)
            self.fwd_meta[-1][0] = "stem_conv_fwd"
        else:
            self.F("s2lc_stem_conv_fwd", self.x_ptr, stem_conv.weight.data_ptr(), z0.data_ptr(), self.stats_f(bn0), N,
                   cfg.in_channels, H, W, c0, self.dt)
            self.meta(self.fwd_meta, N * cfg.in_channels * H * W * 4 + N * h0 * w0 * c0 * self.es, 2.0 * N * h0 * w0 * c0 * 9 * cfg.in_channels)
        self.bn_forward(bn0)
        cur = self.lazy(z0, c0, bn0.scale, bn0.shift, act=L.ACT_SILU)   # lazy stem output
        cur_h, cur_w = h0, w0

        head_hw = (H // 32, W // 32)
        blocks = []          # per-block record for the backward pass
        skips = []           # (tensor Y, channels, h, w, block index), deepest first
        seen_hw = set()
        Y_prev = None
        for i, sp in enumerate(specs):
            bm = enc.blocks[i]
            rec = type("Rec", (), {})()
            rec.sp, rec.h_in, rec.w_in = sp, cur_h, cur_w
            bd = f"blk{i} {N}x{cur_h}x{cur_w} C{sp.c_exp} k{sp.kernel}s{sp.stride}"
            rows_in = N * cur_h * cur_w
            ho, wo = (cur_h + sp.stride - 1) // sp.stride, (cur_w + sp.stride - 1) // sp.stride
            rows_out = N * ho * wo
            rec.h_out, rec.w_out, rec.rows_in, rec.rows_out = ho, wo, rows_in, rows_out
            rec.x_lazy, rec.X = cur, Y_prev
            key = f"encoder.blocks.{i}"
            if sp.expand:
                conv1, bn1m, dwc, bn2m = bm.stem[0], bm.stem[1], bm.stem[3], bm.stem[4]
                rec.k1, rec.kdw, rec.kbn1, rec.kbn2 = key + ".stem.0", key + ".stem.3", key + ".stem.1", key + ".stem.4"
                rec.z1 = self.act(rows_in, sp.c_exp)
                rec.bn1 = _BN(self, bn1m, rows_in)
                rec.w1 = pack_1x1(rec.k1, conv1.weight)
                self.conv(self.fwd_list, L.CONV_1X1, cur, N, cur_h, cur_w, sp.c_in, rec.w1, rec.z1, sp.c_exp,
                          stats=self.stats_f(rec.bn1), stats_kind=L.STATS_MOMENTS if train else L.STATS_NONE)
                self.bn_forward(rec.bn1)
                rec.dw_in = self.lazy(rec.z1, sp.c_exp, rec.bn1.scale, rec.bn1.shift, act=L.ACT_SILU)
            else:
                dwc, bn2m = bm.stem[0], bm.stem[1]
                rec.kdw, rec.kbn2 = key + ".stem.0", key + ".stem.1"
                rec.dw_in = cur
            rec.dw_w = dwc.weight
            rec.z2 = self.act(rows_out, sp.c_exp)
            rec.bn2 = _BN(self, bn2m, rows_out)
            self.F("s2lc_dwconv_fwd", C.byref(rec.dw_in), dwc.weight.data_ptr(), rec.z2.data_ptr(),
                   self.stats_f(rec.bn2), N, cur_h, cur_w, sp.c_exp, sp.kernel, sp.stride, self.dt)
            self.meta(self.fwd_meta, (rows_in + rows_out) * sp.c_exp * self.es, 2.0 * rows_out * sp.c_exp * sp.kernel ** 2, desc=bd)
            self.bn_forward(rec.bn2)
            se1, se2 = bm.squeeze_excitation[1], bm.squeeze_excitation[3]
            rec.se1, rec.se2 = se1, se2
            rec.pooled, rec.gate = self.f32(N * sp.c_exp), self.f32(N * sp.c_exp)
            rec.s2_nogate = self.lazy(rec.z2, sp.c_exp, rec.bn2.scale, rec.bn2.shift, act=L.ACT_SILU)
            rec.s2_gated = self.lazy(rec.z2, sp.c_exp, rec.bn2.scale, rec.bn2.shift, gate=rec.gate, act=L.ACT_SILU)
            self.F("s2lc_se_pool", C.byref(rec.s2_nogate), rec.pooled.data_ptr(), N, ho * wo, sp.c_exp, self.dt)
            self.meta(self.fwd_meta, rows_out * sp.c_exp * self.es, desc=bd)
            self.F("s2lc_se_fc_fwd", rec.pooled.data_ptr(), se1.weight.data_ptr(), se1.bias.data_ptr(),
                   se2.weight.data_ptr(), se2.bias.data_ptr(), rec.gate.data_ptr(), N, sp.c_exp, sp.c_se)
            conv3, bn3m = bm.final_layer[0], bm.final_layer[1]
            rec.k3, rec.kbn3 = key + ".final_layer.0", key + ".final_layer.1"
            rec.z3 = self.act(rows_out, sp.c_out)
            rec.bn3 = _BN(self, bn3m, rows_out)
            rec.w3 = pack_1x1(rec.k3, conv3.weight)
            # project convs at <= 64 px: the gated operand is written once and fed by TMA (forward conv + weight gradient)
            rec.s2_op = rec.s2_gated
            if tma and ho * wo <= TMA_BLOCK_HW:
                rec.s2_op = self.materialize_act(rec.s2_gated, rows_out, sp.c_exp, ho * wo, bd)
            self.conv(self.fwd_list, L.CONV_1X1, rec.s2_op, N, ho, wo, sp.c_exp, rec.w3, rec.z3, sp.c_out,
                      stats=self.stats_f(rec.bn3), stats_kind=L.STATS_MOMENTS if train else L.STATS_NONE)
            self.bn_forward(rec.bn3)
            rec.Y = self.act(rows_out, sp.c_out)
            rec.z3_lazy = self.lazy(rec.z3, sp.c_out, rec.bn3.scale, rec.bn3.shift)
            self.F("s2lc_bn_apply", C.byref(rec.z3_lazy), Y_prev.data_ptr() if sp.residual else None, sp.c_out,
                   dc_ptr(i), rec.Y.data_ptr(), sp.c_out, rows_out, sp.c_out, ho * wo, self.dt)
            self.meta(self.fwd_meta, rows_out * sp.c_out * self.es * (3 if sp.residual else 2), desc=bd)
            rec.is_skip = (ho, wo) not in seen_hw and (ho, wo) != head_hw
            if rec.is_skip:
                seen_hw.add((ho, wo))
                skips.insert(0, rec)
            Y_prev = rec.Y
            cur = self.lazy(rec.Y, sp.c_out)
            cur_h, cur_w = ho, wo
            blocks.append(rec)

        # ---- head
        headc, head_bn_mod = enc.conv_head[0], enc.conv_head[1]
        ch = headc.out_channels
        rows_h = N * cur_h * cur_w
        zh = self.act(rows_h, ch)
        bnh = _BN(self, head_bn_mod, rows_h)
        wh = pack_1x1("encoder.conv_head.0", headc.weight)
        c_last = specs[-1].c_out
        self.conv(self.fwd_list, L.CONV_1X1, cur, N, cur_h, cur_w, c_last, wh, zh, ch, stats=self.stats_f(bnh),
                  stats_kind=L.STATS_MOMENTS if train else L.STATS_NONE)
        self.bn_forward(bnh)
        dec_in = self.lazy(zh, ch, bnh.scale, bnh.shift, act=L.ACT_SILU)
        dec_in_c, dec_h, dec_w = ch, cur_h, cur_w

        # ---- decoder
        levels = []
        n_levels = len(m.up_convs) + 1
        if len(skips) != len(m.up_convs):
            raise RuntimeError(f"expected {len(m.up_convs)} encoder skips, found {len(skips)} (input {H}x{W})")
        for j in range(n_levels):
            lv = type("Lvl", (), {})()
            last = j == n_levels - 1
            up = m.input_up_conv if last else m.up_convs[j]
            dcv = m.input_double_conv if last else m.double_convs[j]
            lv.kup = "input_up_conv" if last else f"up_convs.{j}"
            lv.kdc = "input_double_conv" if last else f"double_convs.{j}"
            lv.up, lv.dcv, lv.last = up, dcv, last
            lv.in_lazy, lv.in_c, lv.h_in, lv.w_in = dec_in, dec_in_c, dec_h, dec_w
            lv.h, lv.w = dec_h * 2, dec_w * 2
            lv.rows = N * lv.h * lv.w
            lv.c_up = up.out_channels
            lv.c_skip = cfg.in_channels if last else skips[j].sp.c_out
            lv.c_cat = lv.c_up + lv.c_skip               # logical concat width (== dcv[0].in_channels)
            if lv.c_cat != dcv[0].in_channels:
                raise RuntimeError(f"decoder level {j}: concat width {lv.c_cat} != conv in_channels {dcv[0].in_channels}")
            lv.c_buf = _up(lv.c_cat, 8)
            lv.cat = self.act(lv.rows, lv.c_buf)
            lv.wup = pack_ct(lv.kup, up.weight)
            lv.in_op = dec_in
            if tma and dec_in_c >= 64:
                lv.in_op = self.materialize_act(dec_in, N * dec_h * dec_w, dec_in_c, dec_h * dec_w,
                                                f"{N}x{dec_h}x{dec_w} C{dec_in_c} convT operand")
            self.conv(self.fwd_list, L.CONVT_FWD, lv.in_op, N, dec_h, dec_w, dec_in_c, lv.wup, lv.cat, lv.c_buf,
                      bias=up.bias)
            if last:
                self.F("s2lc_nchw_to_slice", self.x_ptr, lv.cat.data_ptr() + lv.c_up * self.es, lv.c_buf, N,
                       cfg.in_channels, H, W, lv.c_buf - lv.c_up, self.dt)
            else:
                sk = skips[j]
                lv.skip = sk
                self.F("s2lc_copy_slice", sk.Y.data_ptr(), sk.sp.c_out, lv.cat.data_ptr() + lv.c_up * self.es,
                       lv.c_buf, lv.rows, lv.c_skip, self.dt)
            co = dcv[0].out_channels
            lv.co = co
            lv.zc1, lv.zc2 = self.act(lv.rows, co), self.act(lv.rows, co)
            lv.bnA, lv.bnB = _BN(self, dcv[1], lv.rows), _BN(self, dcv[4], lv.rows)
            lv.wA = pack_3x3(lv.kdc + ".0", dcv[0].weight, lv.c_buf, lv.w)
            lv.wB = pack_3x3(lv.kdc + ".3", dcv[3].weight, co, lv.w)
            lv.cat_lazy = self.lazy(lv.cat, lv.c_buf)
            self.conv(self.fwd_list, L.CONV_3X3, lv.cat_lazy, N, lv.h, lv.w, lv.c_buf, lv.wA, lv.zc1, co,
                      bias=dcv[0].bias, stats=self.stats_f(lv.bnA),
                      stats_kind=L.STATS_MOMENTS if train else L.STATS_NONE)
            self.bn_forward(lv.bnA)
            lv.a1 = self.lazy(lv.zc1, co, lv.bnA.scale, lv.bnA.shift, act=L.ACT_RELU)
            lv.a1_op = lv.a1
            # (256 px rows with <= 64 channels: the forward conv stays on the rolling-row engine, but the packed-tap
            #  TMA weight gradient needs a plain operand and is 2x faster than the loader-warp one)
            if tma and (tma_takes_3x3(lv.w, co) or (lv.w == 256 and co <= 64 and co % 8 == 0)):
                lv.a1_op = self.materialize_act(lv.a1, lv.rows, co, lv.h * lv.w, f"{N}x{lv.h}x{lv.w} C{co} conv operand")
            self.conv(self.fwd_list, L.CONV_3X3, lv.a1_op, N, lv.h, lv.w, co, lv.wB, lv.zc2, co, bias=dcv[3].bias,
                      stats=self.stats_f(lv.bnB), stats_kind=L.STATS_MOMENTS if train else L.STATS_NONE)
            self.bn_forward(lv.bnB)
            lv.a2 = self.lazy(lv.zc2, co, lv.bnB.scale, lv.bnB.shift, act=L.ACT_RELU)
            dec_in, dec_in_c, dec_h, dec_w = lv.a2, co, lv.h, lv.w
            levels.append(lv)

        # ---- classifier
        oc = m.out_conv1x1
        ncls = oc.out_channels
        self.ncls = ncls
        self.logits = self.act(N * H * W, ncls)
        lpp = dec_in_c // 8
        self.fused_outconv = dec_in_c % 8 == 0 and 1 <= lpp <= 32 and lpp & (lpp - 1) == 0 and 2 <= ncls <= 8
        if self.fused_outconv:       # per-pixel HBM-bound kernel (N = num_classes is not a tensor-core shape)
            self.F("s2lc_outconv_fwd", C.byref(dec_in), oc.weight.data_ptr(), oc.bias.data_ptr(),
                   self.logits.data_ptr(), ncls, N * H * W, dec_in_c, ncls, self.dt)
            self.meta(self.fwd_meta, N * H * W * (dec_in_c + ncls) * self.es, 2.0 * N * H * W * dec_in_c * ncls,
                      f"{N}x{H}x{W} {dec_in_c}->{ncls}")
        else:
            wout = pack_1x1("out_conv1x1", oc.weight)
            self.conv(self.fwd_list, L.CONV_1X1, dec_in, N, H, W, dec_in_c, wout, self.logits, ncls, bias=oc.bias)

        self.stats_f_arena = torch.zeros(max(self._stats_f_total, 1), dtype=torch.float64, device=self.dev)
        self.blocks, self.levels = blocks, levels      # kept for introspection (tools/diag_inter.py)
        if not grad:
            self._resolve()
            return

        # ================================================================== BACKWARD
        rows_top = N * H * W
        self.dlogits_g = self.lazygrad(0, ncls)     # pointer patched per call
        top = levels[-1]
        self.dlogits_ptr = C.c_void_p(0)              # patched per call (fused classifier backward)
        dyB = self.act(rows_top, top.co)
        if self.fused_outconv:
            gw, gb = self.gptr("out_conv1x1.weight"), self.gptr("out_conv1x1.bias")
            self.B("s2lc_outconv_bwd", C.byref(top.a2), self.dlogits_ptr, ncls, oc.weight.data_ptr(), gw, gb,
                   dyB.data_ptr(), top.co, ("SB", top.bnB.st_b), rows_top, top.co, ncls, self.dt)
            self.meta(self.bwd_meta, rows_top * (2 * top.co + ncls) * self.es, 4.0 * rows_top * top.co * ncls,
                      f"{N}x{H}x{W} {ncls}->{top.co}")
        else:
            self.wgrad(L.CONV_1X1, top.a2, self.dlogits_g, N, H, W, top.co, ncls, self.gptr("out_conv1x1.weight"),
                       self.gptr("out_conv1x1.bias"))
            wout_d = pack_1x1_d("out_conv1x1", oc.weight)
            self.conv(self.bwd_list, L.CONV_1X1, self.dlogits_g, N, H, W, ncls, wout_d, dyB, top.co, prev=top.a2,
                      stats=("SB", top.bnB.st_b), stats_kind=L.STATS_BNBWD)
        for j in range(n_levels - 1, -1, -1):
            lv = levels[j]
            dcv = lv.dcv
            co = lv.co
            if j != n_levels - 1:
                dyB = lv.dyB_buf            # filled by the level above (convT dgrad)
            self.bn_backward(lv.bnB, lv.kdc + ".4")
            gB = self.lazygrad(dyB, co, lv.zc2, co, lv.bnB)
            if MATERIALIZE_DECODER_GRADS:
                gB = self.materialize_grad(gB, dyB, co, lv.rows, co, lv.h * lv.w)
            self.wgrad(L.CONV_3X3, lv.a1_op, gB, N, lv.h, lv.w, co, co, self.gptr(lv.kdc + ".3.weight"),
                       self.gptr(lv.kdc + ".3.bias"))
            dyA = self.act(lv.rows, co)
            lv.dyA, lv.dyB = dyA, dyB
            wB_d = pack_3x3_d(lv.kdc + ".3", dcv[3].weight, co, lv.w)
            self.conv(self.bwd_list, L.CONV_3X3, gB, N, lv.h, lv.w, co, wB_d, dyA, co, prev=lv.a1,
                      stats=("SB", lv.bnA.st_b), stats_kind=L.STATS_BNBWD)
            self.bn_backward(lv.bnA, lv.kdc + ".1")
            gA = self.lazygrad(dyA, co, lv.zc1, co, lv.bnA)
            if MATERIALIZE_DECODER_GRADS:
                gA = self.materialize_grad(gA, dyA, co, lv.rows, co, lv.h * lv.w)
            self.wgrad(L.CONV_3X3, lv.cat_lazy, gA, N, lv.h, lv.w, lv.c_cat, co, self.gptr(lv.kdc + ".0.weight"),
                       self.gptr(lv.kdc + ".0.bias"))
            n_in_used = lv.c_up if lv.last else lv.c_cat    # d(input tile) is not needed
            lv.c_dcat = _up(n_in_used, 8)
            lv.dcat = self.act(lv.rows, lv.c_dcat)
            wA_d = pack_3x3_d(lv.kdc + ".0", dcv[0].weight, n_in_used, lv.w)
            self.conv(self.bwd_list, L.CONV_3X3, gA, N, lv.h, lv.w, co, wA_d, lv.dcat, lv.c_dcat)
            # transposed conv: weight/bias gradient and data gradient into the level below
            gup = self.lazygrad(lv.dcat, lv.c_dcat)
            self.wgrad(L.CONVT_FWD, lv.in_op, gup, N, lv.h_in, lv.w_in, lv.in_c, lv.c_up,
                       self.gptr(lv.kup + ".weight"), self.gptr(lv.kup + ".bias"))
            wup_d = pack_ct_d(lv.kup, lv.up.weight)
            below = self.act(N * lv.h_in * lv.w_in, lv.in_c)
            below_bn = levels[j - 1].bnB if j > 0 else bnh
            self.conv(self.bwd_list, L.CONVT_DGRAD, gup, N, lv.h_in, lv.w_in, lv.c_up, wup_d, below, lv.in_c,
                      prev=lv.in_lazy, stats=("SB", below_bn.st_b), stats_kind=L.STATS_BNBWD)
            if j > 0:
                levels[j - 1].dyB_buf = below
            else:
                dyh = below

        # ---- head backward -> gradient w.r.t. the last block output
        self.bn_backward(bnh, "encoder.conv_head.1")
        gh = self.lazygrad(dyh, ch, zh, ch, bnh)
        last = blocks[-1]
        self.wgrad(L.CONV_1X1, self.lazy(last.Y, c_last), gh, N, cur_h, cur_w, c_last, ch,
                   self.gptr("encoder.conv_head.0.weight"))
        wh_d = pack_1x1_d("encoder.conv_head.0", headc.weight)
        max_y = max(r.rows_out * r.sp.c_out for r in blocks)
        dY = [self.act(max_y, 1) for _ in range(2)]
        da_buf = self.act(max(r.rows_out * r.sp.c_exp for r in blocks), 1)
        dy1_buf = self.act(max(r.rows_in * r.sp.c_exp for r in blocks), 1)
        g3m_buf = self.act(max(r.rows_out * r.sp.c_out for r in blocks), 1)
        se_scratch = self.f32(max(N * (r.sp.c_exp + 2 * r.sp.c_se) for r in blocks))
        dgate, dpool = self.f32(max(N * r.sp.c_exp for r in blocks)), self.f32(max(N * r.sp.c_exp for r in blocks))

        def skip_grad(rec):
            """(pointer, ld) of the decoder's gradient w.r.t. this block's output, if it is a skip."""
            if not rec.is_skip:
                return None, 0
            lv = next(l for l in levels if not l.last and l.skip is rec)
            return lv.dcat.data_ptr() + lv.c_up * self.es, lv.c_dcat

        sp_ptr, sp_ld = skip_grad(last)
        self.conv(self.bwd_list, L.CONV_1X1, gh, N, cur_h, cur_w, ch, wh_d, dY[(nb - 1) % 2], c_last,
                  res1=sp_ptr, ld_res1=sp_ld, prev=self._raw(last),
                  stats=("SB", last.bn3.st_b), stats_kind=L.STATS_BNBWD, stats_rowscale=dc_ptr(nb - 1))

        for i in range(nb - 1, -1, -1):
            rec = blocks[i]
            sp = rec.sp
            dYi = dY[i % 2]
            ho, wo, hi, wi = rec.h_out, rec.w_out, rec.h_in, rec.w_in
            bd = f"blk{i} {N}x{hi}x{wi} C{sp.c_exp} k{sp.kernel}s{sp.stride}"
            self.bn_backward(rec.bn3, rec.kbn3)
            g3 = self.lazygrad(dYi, sp.c_out, rec.z3, sp.c_out, rec.bn3, rowscale=dc_ptr(i))
            if tma and ho * wo <= TMA_BLOCK_HW:
                # deep blocks: the BatchNorm-backward gradient is written once (out of place: dYi still feeds the
                # residual path) and both gradient engines read it as a plain tensor, the data gradient through TMA
                self.B("s2lc_lazygrad_apply", C.byref(g3), g3m_buf.data_ptr(), sp.c_out, rec.rows_out, sp.c_out,
                       ho * wo, self.dt, nbytes=3 * rec.rows_out * sp.c_out * self.es)
                g3 = self.lazygrad(g3m_buf, sp.c_out)
            self.wgrad(L.CONV_1X1, rec.s2_op, g3, N, ho, wo, sp.c_exp, sp.c_out, self.gptr(rec.k3 + ".weight"))
            w3_d = pack_1x1_d(rec.k3, enc.blocks[i].final_layer[0].weight)
            self.conv(self.bwd_list, L.CONV_1X1, g3, N, ho, wo, sp.c_out, w3_d, da_buf, sp.c_exp)
            # squeeze-excite backward
            kse = f"encoder.blocks.{i}.squeeze_excitation"
            self.B("s2lc_se_bwd_reduce", da_buf.data_ptr(), sp.c_exp, C.byref(rec.s2_nogate), dgate.data_ptr(), N,
                   ho * wo, sp.c_exp, self.dt)
            self.meta(self.bwd_meta, 2 * rec.rows_out * sp.c_exp * self.es, desc=bd)
            self.B("s2lc_se_fc_bwd", rec.pooled.data_ptr(), rec.se1.weight.data_ptr(), rec.se1.bias.data_ptr(),
                   rec.se2.weight.data_ptr(), rec.se2.bias.data_ptr(), dgate.data_ptr(), dpool.data_ptr(),
                   self.gptr(kse + ".1.weight"), self.gptr(kse + ".1.bias"), self.gptr(kse + ".3.weight"),
                   self.gptr(kse + ".3.bias"), se_scratch.data_ptr(), N, sp.c_exp, sp.c_se)
            self.B("s2lc_se_bwd_apply", da_buf.data_ptr(), sp.c_exp, C.byref(rec.s2_gated), dpool.data_ptr(),
                   ("SB", rec.bn2.st_b), N, ho * wo, sp.c_exp, self.dt)
            self.meta(self.bwd_meta, 3 * rec.rows_out * sp.c_exp * self.es, desc=bd)
            self.bn_backward(rec.bn2, rec.kbn2)
            g2 = self.lazygrad(da_buf, sp.c_exp, rec.z2, sp.c_exp, rec.bn2)
            gdw = self.gptr(rec.kdw + ".weight")
            dw_bytes = (2 * rec.rows_out + 2 * rec.rows_in) * sp.c_exp * self.es
            dw_flops = 4.0 * rec.rows_out * sp.c_exp * sp.kernel ** 2      # data gradient + weight gradient
            prev_rec = blocks[i - 1] if i > 0 else None
            if prev_rec is not None:
                psk_ptr, psk_ld = skip_grad(prev_rec)
            if sp.expand:
                self.B("s2lc_dwconv_bwd", C.byref(g2), rec.dw_w.data_ptr(), C.byref(rec.dw_in), 1, None, 0,
                       dy1_buf.data_ptr(), sp.c_exp, ("SB", rec.bn1.st_b), gdw, N, hi, wi, sp.c_exp, sp.kernel,
                       sp.stride, self.dt)
                self.meta(self.bwd_meta, dw_bytes, dw_flops, desc=bd)
                self.bn_backward(rec.bn1, rec.kbn1)
                g1 = self.lazygrad(dy1_buf, sp.c_exp, rec.z1, sp.c_exp, rec.bn1)
                if tma and hi * wi <= TMA_BLOCK_HW:
                    g1 = self.materialize_grad(g1, dy1_buf, sp.c_exp, rec.rows_in, sp.c_exp, hi * wi)
                self.wgrad(L.CONV_1X1, rec.x_lazy, g1, N, hi, wi, sp.c_in, sp.c_exp, self.gptr(rec.k1 + ".weight"))
                if prev_rec is None:
                    raise RuntimeError("block 0 with an expand conv is not a reference architecture")
                w1_d = pack_1x1_d(rec.k1, enc.blocks[i].stem[0].weight)
                self.conv(self.bwd_list, L.CONV_1X1, g1, N, hi, wi, sp.c_exp, w1_d, dY[(i - 1) % 2], sp.c_in,
                          res1=dYi.data_ptr() if sp.residual else None, ld_res1=sp.c_out, res2=psk_ptr,
                          ld_res2=psk_ld, prev=self._raw(prev_rec), stats=("SB", prev_rec.bn3.st_b),
                          stats_kind=L.STATS_BNBWD, stats_rowscale=dc_ptr(i - 1))
            elif prev_rec is not None:
                # input is the materialised output of block i-1
                dYp = dY[(i - 1) % 2]
                self.B("s2lc_dwconv_bwd", C.byref(g2), rec.dw_w.data_ptr(), C.byref(rec.dw_in), 0,
                       dYi.data_ptr() if sp.residual else None, sp.c_out, dYp.data_ptr(), sp.c_in, None, gdw, N, hi, wi,
                       sp.c_exp, sp.kernel, sp.stride, self.dt)
                self.meta(self.bwd_meta, dw_bytes + (rec.rows_in * sp.c_exp * self.es if sp.residual else 0), dw_flops,
                          desc=bd)
                if psk_ptr is not None:
                    self.B("s2lc_add_slice", dYp.data_ptr(), sp.c_in, psk_ptr, psk_ld, dYp.data_ptr(), sp.c_in,
                           rec.rows_in, sp.c_in, self.dt)
                self.B("s2lc_bn_bwd_reduce", dYp.data_ptr(), sp.c_in, prev_rec.z3.data_ptr(), sp.c_in, dc_ptr(i - 1),
                       ("SB", prev_rec.bn3.st_b), rec.rows_in, sp.c_in, hi * wi, self.dt)
                self.meta(self.bwd_meta, 2 * rec.rows_in * sp.c_in * self.es, desc=bd)
            else:
                # block 0: input is the lazy stem output -> gradient w.r.t. the stem BatchNorm output
                dy0 = dy1_buf
                self.B("s2lc_dwconv_bwd", C.byref(g2), rec.dw_w.data_ptr(), C.byref(rec.dw_in), 1, None, 0,
                       dy0.data_ptr(), c0, ("SB", bn0.st_b), gdw, N, hi, wi, sp.c_exp, sp.kernel, sp.stride, self.dt)
                self.meta(self.bwd_meta, dw_bytes, dw_flops, desc=bd)
                self.bn_backward(bn0, "encoder.stem.1")
                g0 = self.lazygrad(dy0, c0, z0, c0, bn0)
                if stem_tc:
                    self.wgrad(L.CONV_1X1, self.stem_lazy, g0, N, h0, w0, k_stem, c0, self.gptr("encoder.stem.0.weight"))
                    self.bwd_meta[-1][0] = "stem_conv_wgrad"
                else:
                    self.B("s2lc_stem_conv_wgrad", self.x_ptr, C.byref(g0), self.gptr("encoder.stem.0.weight"), N,
                           cfg.in_channels, H, W, c0, self.dt)
        self.stats_b_arena = torch.zeros(max(self._stats_b_total, 1), dtype=torch.float64, device=self.dev)
        self._resolve()

    def _raw(self, rec) -> L.Lazy:
        """Lazy view of a block's raw project-conv output with NO transform (only supplies z for BNBWD statistics)."""
        if not hasattr(rec, "_raw_lazy"):
            rec._raw_lazy = self.lazy(rec.z3, rec.sp.c_out)
        return rec._raw_lazy

    def _resolve(self) -> None:
        """Replace ("SF"/"SB", offset) markers by absolute fp64 arena pointers."""
        basef = self.stats_f_arena.data_ptr()
        baseb = self.stats_b_arena.data_ptr() if hasattr(self, "stats_b_arena") else 0

        def fix(v):
            if isinstance(v, tuple) and len(v) == 2 and v[0] in ("SF", "SB"):
                return (basef if v[0] == "SF" else baseb) + 8 * v[1]
            return v

        for lst in (self.fwd_list, self.bwd_list):
            for idx, (fn, args) in enumerate(lst):
                lst[idx] = (fn, tuple(fix(a) for a in args))
        for obj in self._keep:
            if isinstance(obj, L.ConvParams) and isinstance(getattr(obj, "_stats_marker", None), tuple):
                obj.stats = fix(obj._stats_marker)

    # ---------------------------------------------------------------- execution
    def run_forward(self, x: torch.Tensor, events: list | None = None, only: set | None = None) -> torch.Tensor:
        """``events``: list that receives one (start, stop) CUDA-event pair per launch (recorded on the launching
        stream); ``only``: restrict that to the launches whose kernel-class label is in the set (cheap enough to
        leave on inside a timed region)."""
        if self.eng.dry_run:
            raise RuntimeError("dry-run engine cannot execute")
        st = torch.cuda.current_stream().cuda_stream
        lib = self.lib
        self.x_keep = x
        self.x_ptr.value = x.data_ptr()
        if self.training:
            L.check(lib.s2lc_memset_zero(self.stats_f_arena.data_ptr(), self.stats_f_arena.numel() * 8, st))
            self.eng.stats_gen += 1          # running statistics are about to change (written through raw pointers)
        self.eng.repack(self.dtype, st)
        if self.affine_list:
            sig = (self.eng.weight_signature(), self.eng.stats_gen, sum(b._version for b in self._affine_buffers),
                   hash(tuple(b.data_ptr() for b in self._affine_buffers)))
            if sig != self._affine_sig:
                for fn, args in self.affine_list:
                    rc = fn(*args, st)
                    if rc:
                        L.check(rc)
                self._affine_sig = sig
        if events is None and self.eng.timing_hook is not None:
            events, only = self.eng.timing_hook["fwd"], self.eng.timing_hook["only"]
        for i, (fn, args) in enumerate(self.fwd_list):
            timed = events is not None and (only is None or self.fwd_meta[i][0] in only)
            if timed:
                e0 = torch.cuda.Event(enable_timing=True)
                e0.record()
            if fn == "DC":
                self._drop_connect(st)
            else:
                rc = fn(*args, st)
                if rc:
                    L.check(rc)
            if timed:
                e1 = torch.cuda.Event(enable_timing=True)
                e1.record()
                events.append((e0, e1, i))
        if self.training and self.n_bn_trained:
            torch._foreach_add_(self.n_bn_trained, 1)
        n, h, w = self.N, self.H, self.W
        return self.logits.view(n, h, w, self.ncls).permute(0, 3, 1, 2)

    def _drop_connect(self, st: int) -> None:
        masks = self.model._dc_masks
        if masks is not None:
            self.dc.copy_(masks.to(self.dc.device, torch.float32))
            return
        cfg = self.model.config
        self.model._dc_step += 1
        L.check(self.lib.s2lc_dropconnect_masks(self.dc.data_ptr(), self.dc.shape[0], self.N,
                                                float(cfg.drop_connect_rate), 0x5eed, self.model._dc_step, st))

    def bucket_schedule(self, buckets) -> dict[int, list[int]]:
        """launch index -> gradient buckets that are complete once that launch has been enqueued."""
        key = id(buckets)
        if key not in self._bucket_sched:
            sched: dict[int, list[int]] = {}
            for b, (_, _, keys) in enumerate(buckets):
                last = max(self._grad_write_idx[k] for k in keys)
                sched.setdefault(last, []).append(b)
            self._bucket_sched[key] = sched
        return self._bucket_sched[key]

    def run_backward(self, dlogits_nhwc: torch.Tensor, events: list | None = None, only: set | None = None) -> None:
        st = torch.cuda.current_stream().cuda_stream
        lib = self.lib
        L.check(lib.s2lc_memset_zero(self.stats_b_arena.data_ptr(), self.stats_b_arena.numel() * 8, st))
        fg = self.eng.flat_grad
        L.check(lib.s2lc_memset_zero(fg.data_ptr(), fg.numel() * 4, st))
        self.dlogits_g.dy = dlogits_nhwc.data_ptr()
        self.dlogits_ptr.value = dlogits_nhwc.data_ptr()
        hook = getattr(self.eng, "bucket_hook", None)
        sched = self.bucket_schedule(self.eng.buckets) if hook is not None else None
        if events is None and self.eng.timing_hook is not None:      # bench.py: time one kernel class in the timed region
            events, only = self.eng.timing_hook["bwd"], self.eng.timing_hook["only"]
        for i, (fn, args) in enumerate(self.bwd_list):
            timed = events is not None and (only is None or self.bwd_meta[i][0] in only)
            if timed:
                e0 = torch.cuda.Event(enable_timing=True)
                e0.record()
            rc = fn(*args, st)
            if rc:
                L.check(rc)
            if timed:
                e1 = torch.cuda.Event(enable_timing=True)
                e1.record()
                events.append((e0, e1, i))
            if sched is not None and i in sched:
                for b in sched[i]:
                    hook(b)


class _Lease:
    """Marks a workspace as holding the saved activations of a live autograd graph.  Released by backward(), or when
    the graph is dropped without a backward (the lease dies with the autograd context), so a training-mode forward
    whose result is discarded does not pin its workspace."""

    def __init__(self, ws: "Workspace") -> None:
        self.ws = ws
        ws.lease_id += 1
        self.id = ws.lease_id
        ws.in_flight = True

    def release(self) -> None:
        if self.ws.lease_id == self.id:
            self.ws.in_flight = False

    __del__ = release


class _UnetFunction(torch.autograd.Function):
    """One autograd node for the whole network: forward = the launch list, backward = its hand-written reverse."""

    @staticmethod
    def forward(ctx, model: EfficientnetUnet, ws: Workspace, x: torch.Tensor, *params):
        out = ws.run_forward(x)
        ctx.lease = _Lease(ws)
        ctx.ws, ctx.model = ws, model
        # a trainer that owns the loop (direct_grad) consumes the logits before the next forward: hand out the
        # workspace's own buffer; everyone else gets a private copy, as a fresh tensor from the reference module is
        return out if ws.eng.direct_grad else out.clone()

    @staticmethod
    def backward(ctx, dlogits: torch.Tensor):
        ws: Workspace = ctx.ws
        n, h, w, c = ws.N, ws.H, ws.W, ws.ncls
        d = dlogits.permute(0, 2, 3, 1)
        if d.dtype != ws.dtype or not d.is_contiguous():
            d = d.to(ws.dtype).contiguous()
        if not ws.in_flight or ws.lease_id != ctx.lease.id:
            raise RuntimeError("the activations of this forward pass were overwritten by a later forward (workspace "
                               "lease lost): backward() must run before the workspace is reused")
        ws.run_backward(d)
        ctx.lease.release()
        eng = ws.eng
        if eng.direct_grad:
            for k, p in eng.used:                       # parameters see the flat buffer as their .grad
                if p.grad is None or p.grad.data_ptr() != eng.grad_views[k].data_ptr():
                    p.grad = eng.grad_views[k]
            grads = [None] * len(eng.model_param_order)
        else:
            grads = [eng.grad_views[k].clone() if k in eng.grad_views else None for k in eng.model_param_order]
        return (None, None, None, *grads)


def unet_apply(model: EfficientnetUnet, x: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    eng = model.engine()
    x = x.detach()
    if x.dtype != torch.float32 or not x.is_contiguous():
        x = x.float().contiguous()
    training = model.training
    need_grad = torch.is_grad_enabled() and any(p.requires_grad for _, p in eng.used)
    if need_grad and not training:
        raise RuntimeError("gradients in eval() mode are not built (the reference never trains in eval mode)")
    ws = eng.workspace(x, dtype, training, need_grad)
    if not need_grad:
        return ws.run_forward(x).clone()
    if not hasattr(eng, "model_param_order"):
        eng.model_param_order = [k for k, _ in model.named_parameters()]
    params = [p for _, p in model.named_parameters()]
    return _UnetFunction.apply(model, ws, x, *params)
