"""Host-side logic that needs no GPU: the C ABI surface, the launch-list builder (dry run), the gradient buckets
and the 2-rank data-parallel gradient averaging over gloo."""
import os
import re
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from s2lc_b200 import _lib, synthetic
from s2lc_b200.efficientnet_unet import EfficientNetConfig, EfficientnetUnet, architecture
from s2lc_b200.train_segmentation import BucketReducer, make_buckets

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "s2lc.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(s2lc_[a-z0-9_]+)\s*\(", src)))


def test_cabi_exports_every_declared_symbol():
    """libs2lc.so loads on a CPU-only box and exports exactly what include/s2lc.h declares (no compute calls)."""
    lib = _lib.load()
    declared = _declared_symbols()
    assert len(declared) >= 25
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/s2lc.h but not exported"
    assert set(_lib.EXPORTED_SYMBOLS) == set(declared), set(_lib.EXPORTED_SYMBOLS) ^ set(declared)
    assert lib.s2lc_version() >= 100
    assert lib.s2lc_last_error() is not None


def test_cabi_struct_layouts_match_header(tmp_path):
    """The ctypes mirrors in _lib.py have the size and field offsets a C compiler derives from include/s2lc.h."""
    import ctypes as C
    import subprocess
    structs = {"s2lc_lazy": _lib.Lazy, "s2lc_lazygrad": _lib.LazyGrad, "s2lc_conv_params": _lib.ConvParams,
               "s2lc_wgrad_params": _lib.WgradParams}
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{ROOT}/include/s2lc.h"', 'int main(void) {']
    for cname, cls in structs.items():
        lines.append(f'  printf("{cname} %zu\\n", sizeof({cname}));')
        for fname, _ in cls._fields_:
            lines.append(f'  printf("{cname}.{fname} %zu\\n", offsetof({cname}, {fname}));')
    lines += ['  return 0;', '}']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-std=c11", "-o", str(exe), str(src)])
    got = dict(l.split() for l in subprocess.check_output([str(exe)], text=True).splitlines())
    for cname, cls in structs.items():
        assert int(got[cname]) == C.sizeof(cls), cname
        for fname, _ in cls._fields_:
            assert int(got[f"{cname}.{fname}"]) == getattr(cls, fname).offset, f"{cname}.{fname}"


def _model(version="b0"):
    cfg = EfficientNetConfig(version=version, in_channels=6, num_classes=4,
                             class_distribution=list(synthetic.CLASS_DISTRIBUTION))
    return EfficientnetUnet(cfg)


@pytest.mark.parametrize("version,nblocks,stem,head", [("b0", 16, 32, 1280), ("b5", 39, 48, 2048), ("b7", 55, 64, 2560)])
def test_architecture_table(version, nblocks, stem, head):
    """Block list generation (reference efficientnet_unet.py:184-228): counts and the b5 stage table of SURVEY.md §8."""
    cfg = EfficientNetConfig(version=version, in_channels=6, num_classes=4)
    s, blocks, h = architecture(cfg)
    assert (s, len(blocks), h) == (stem, nblocks, head)
    if version == "b5":
        got = [(b.c_in, b.c_exp, b.kernel, b.stride, b.c_out, b.c_se) for b in blocks]
        assert got[0] == (48, 48, 3, 1, 24, 12) and got[1] == (24, 24, 3, 1, 24, 6)
        assert got[3] == (24, 144, 3, 2, 40, 6) and got[8] == (40, 240, 5, 2, 64, 10)
        assert got[-1] == (512, 3072, 3, 1, 512, 128)


def test_engine_dry_run_builds_launch_lists():
    """The whole forward/backward launch list is built without touching a GPU; every used parameter's gradient is
    written by exactly one backward launch and laid out in reverse execution order."""
    from s2lc_b200._engine import Engine
    model = _model("b0")
    eng = Engine(model, dry_run=True)
    x = torch.zeros(2, 6, 64, 64)
    ws = eng.workspace(x, torch.bfloat16, True, True)
    assert len(ws.fwd_list) == len(ws.fwd_meta) and len(ws.bwd_list) == len(ws.bwd_meta)
    assert len(ws.fwd_list) > 150 and len(ws.bwd_list) > 200
    used = dict(model.used_parameters())
    assert set(ws._grad_write_idx) == set(used)
    assert not any(k.startswith("encoder.fc.") for k in used)
    assert eng.flat_grad.numel() >= sum(p.numel() for p in used.values())
    # reverse execution order: the classifier's gradient sits at offset 0, the stem's at the end
    assert eng.grad_offsets["out_conv1x1.bias"] == 0
    assert eng.grad_offsets["encoder.stem.0.weight"] == max(eng.grad_offsets.values())
    with pytest.raises(RuntimeError):
        ws.run_forward(x)
    # algorithmic flops of the forward convs agree with the per-tile figure of SURVEY.md §8 (b0 @64x64 is tiny;
    # just require positive, finite metadata on every conv launch)
    for label, nbytes, flops, desc in ws.fwd_meta + ws.bwd_meta:
        assert nbytes >= 0 and flops >= 0
    # eval mode: the BatchNorm scale / shift launches sit in their own list (run once per parameter / running-statistic
    # state, Workspace.run_forward), one per BatchNorm layer the forward uses, and not in the per-call forward list
    ws_e = eng.workspace(x, torch.bfloat16, False, False)
    n_bn = sum(1 for m in model.modules() if isinstance(m, torch.nn.BatchNorm2d))
    assert 0 < len(ws_e.affine_list) <= n_bn and len(ws_e._affine_buffers) == 2 * len(ws_e.affine_list)
    assert not ws.affine_list                                        # training mode computes batch statistics instead
    assert all(meta[0] != "bn_eval_affine" for meta in ws_e.fwd_meta)
    assert len(ws_e.fwd_list) == len(ws_e.fwd_meta)


def test_bucket_schedule_covers_every_bucket_once():
    from s2lc_b200._engine import Engine
    model = _model("b0")
    eng = Engine(model, dry_run=True)
    ws = eng.workspace(torch.zeros(1, 6, 64, 64), torch.bfloat16, True, True)
    numels = {k: p.numel() for k, p in eng.used}
    buckets = make_buckets(eng.grad_offsets, numels, eng.flat_grad.numel(), 1 << 18)
    assert buckets[0][0] == 0 and buckets[-1][1] == eng.flat_grad.numel()
    assert all(a[1] == b[0] for a, b in zip(buckets, buckets[1:]))
    assert sorted(k for _, _, ks in buckets for k in ks) == sorted(numels)
    sched = ws.bucket_schedule(buckets)
    fired = [b for i in sorted(sched) for b in sched[i]]
    assert sorted(fired) == list(range(len(buckets)))
    # bucket b may fire only after the last launch that writes one of its gradients
    for i, bs in sched.items():
        for b in bs:
            assert i == max(ws._grad_write_idx[k] for k in buckets[b][2])
    # buckets complete in order (decoder first), so the all-reduce of bucket 0 overlaps the rest of backward
    # at the production granularity (a handful of buckets) they complete in order, decoder first, so the
    # all-reduce of bucket 0 overlaps the rest of backward
    coarse = make_buckets(eng.grad_offsets, numels, eng.flat_grad.numel(), eng.flat_grad.numel() // 4)
    sched = ws.bucket_schedule(coarse)
    first_ready = {b: i for i, bs in sched.items() for b in bs}
    assert all(first_ready[b] <= first_ready[b + 1] for b in range(len(coarse) - 1))
    assert first_ready[0] < len(ws.bwd_list) // 2


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _dp_worker(rank: int, world: int, port: int, ret) -> None:
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        offsets = {"a": 0, "b": 12, "c": 40, "d": 44}
        numels = {"a": 10, "b": 27, "c": 4, "d": 20}
        total = 64
        buckets = make_buckets(offsets, numels, total, 16)
        g = torch.Generator().manual_seed(100 + rank)
        flat = torch.randn(total, generator=g)
        mine = flat.clone()
        red = BucketReducer(flat, buckets)
        for b in range(len(buckets)):         # the order the backward launch list fires them in
            red.reduce(b)
        gathered = [torch.zeros(total) for _ in range(world)]
        dist.all_gather(gathered, mine)
        want = torch.stack(gathered).mean(0)
        ret[rank] = float((flat - want).abs().max())
    finally:
        dist.destroy_process_group()


def test_two_rank_gradient_average_gloo():
    """DDP semantics of the trainer's exchange step: after all buckets are reduced every rank holds
    the MEAN of the per-rank flat gradients.  world_size 2 over gloo on CPU."""
    world, port = 2, _free_port()
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_dp_worker, args=(world, port, ret), nprocs=world, join=True)
        assert len(ret) == world and all(v < 1e-6 for v in ret.values()), dict(ret)


def test_make_buckets_sizes():
    offsets = {f"p{i}": i * 100 for i in range(10)}
    numels = {f"p{i}": 100 for i in range(10)}
    b = make_buckets(offsets, numels, 1000, 250)
    assert [(x[0], x[1]) for x in b] == [(0, 300), (300, 600), (600, 900), (900, 1000)]


# ---------------------------------------------------------------- mosaic inference planning
@pytest.mark.parametrize("h,w,tile,margin", [(100, 130, 64, 16), (64, 64, 64, 0), (97, 33, 32, 8), (2048, 2048, 512, 64),
                                             (1, 1, 32, 0)])
def test_mosaic_plan_covers_scene_exactly_once(h, w, tile, margin):
    import numpy as np
    from oracle.mosaic_oracle import tile_origins, write_tile
    from s2lc_b200.inference import plan_mosaic
    origins = plan_mosaic(h, w, tile, margin)
    assert origins.dtype == torch.int32 and np.array_equal(origins.numpy(), tile_origins(h, w, tile, margin))
    hits = np.zeros((h, w), dtype=np.int32)
    keep = tile - 2 * margin
    for t, l in origins.tolist():
        y0, x0 = max(t + margin, 0), max(l + margin, 0)
        hits[y0:min(t + margin + keep, h), x0:min(l + margin + keep, w)] += 1
    assert (hits == 1).all()
    # the oracle writer agrees with that geometry: constant-class tiles fill the whole mosaic
    mosaic = np.full((h, w), 255, dtype=np.uint8)
    if h * w <= 20000:
        logits = np.zeros((3, tile, tile), dtype=np.float32)
        logits[2] = 1.0
        for t, l in origins.tolist():
            write_tile(mosaic, logits, t, l, margin)
        assert (mosaic == 2).all()


def test_mosaic_plan_rejects_bad_geometry():
    from s2lc_b200.inference import plan_mosaic, rank_batches
    for args in [(64, 64, 48, 0), (64, 64, 64, 32), (64, 64, 64, -1), (0, 64, 64, 0)]:
        with pytest.raises(ValueError):
            plan_mosaic(*args)
    with pytest.raises(ValueError):
        rank_batches(10, 4, 2, 2)


@pytest.mark.parametrize("n_tiles,batch,world", [(25, 8, 1), (25, 8, 2), (7, 8, 4), (64, 8, 8), (0, 4, 2)])
def test_rank_batches_partition_the_tiles(n_tiles, batch, world):
    from s2lc_b200.inference import rank_batches
    seen = []
    per_rank = []
    for r in range(world):
        mine = rank_batches(n_tiles, batch, r, world)
        per_rank.append(len(mine))
        for s, e in mine:
            assert 0 < e - s <= batch
            seen.extend(range(s, e))
    assert sorted(seen) == list(range(n_tiles))
    assert max(per_rank) - min(per_rank) <= 1


def _mosaic_worker(rank: int, world: int, port: int, ret) -> None:
    import numpy as np
    from oracle.mosaic_oracle import write_tile
    from s2lc_b200.inference import MosaicPredictor, plan_mosaic
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        h, w, tile, margin, batch = 150, 200, 96, 16, 4
        origins = plan_mosaic(h, w, tile, margin)
        index_of = {tuple(o): i for i, o in enumerate(origins.tolist())}

        def tile_logits(i):                     # class (i % 5) + 1 everywhere: which tile wrote a pixel stays visible
            lg = np.zeros((7, tile, tile), dtype=np.float32)
            lg[i % 5 + 1] = 1.0
            return lg

        class HostStages(MosaicPredictor):
            """MosaicPredictor with its two CUDA stages replaced by the numpy oracle, so that predict() itself — tile
            planning, dealing batches to ranks, parking the ragged last batch, engine chunks, the final all-reduce —
            runs on the CPU over gloo."""
            calls = 0

            def gather(self, scene, o):
                return o                         # the "network" below is the identity: hand the origins through

            def scatter_argmax(self, logits, o, mosaic):
                HostStages.calls += 1
                assert o.shape[0] <= self.engine_batch
                buf = mosaic.numpy()
                for t, l in o.tolist():
                    if (t, l) in index_of:       # parked tiles lie outside the scene and write nothing
                        write_tile(buf, tile_logits(index_of[(t, l)]), t, l, margin)

        full = np.zeros((h, w), dtype=np.uint8)
        for i, (t, l) in enumerate(origins.tolist()):
            write_tile(full, tile_logits(i), t, l, margin)
        pred = HostStages(torch.nn.Identity(), np.zeros(6, np.float32), np.ones(6, np.float32), tile=tile, margin=margin,
                          batch=batch, device="cpu", engine_batch=2)
        assert (pred.rank, pred.world_size) == (rank, world)
        mosaic = pred.predict(torch.zeros(6, h, w, dtype=torch.int16))
        ret[rank] = (bool(np.array_equal(mosaic.numpy(), full)) and bool((full > 0).all()) and HostStages.calls >= 2)
    finally:
        dist.destroy_process_group()


def test_two_rank_mosaic_assembly_gloo():
    """Mosaic inference at world_size 2 over gloo THROUGH MosaicPredictor.predict (its CUDA stages
    swapped for the numpy oracle): batches of tiles dealt round-robin to the ranks fill disjoint parts of
    zero-initialised uint8 buffers, and one SUM all-reduce assembles the scene on every rank."""
    world, port = 2, _free_port()
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_mosaic_worker, args=(world, port, ret), nprocs=world, join=True)
        assert len(ret) == world and all(ret.values()), dict(ret)


def test_fused_repack_table_stays_inside_every_pack():
    """The fused Adam + re-pack pass scatters every parameter element to (row, tap, channel) of its packed copies by
    peeling source strides (csrc/elementwise.cu pack_scatter).  Restated here on the entry table of a dry-run engine:
    elements inside the packed slice land on distinct addresses inside the buffer, and elements outside it (the
    data-gradient operand of the first decoder conv keeps 32 of the weight's 38 input channels) are skipped."""
    from s2lc_b200._engine import Engine
    model = _model("b0")
    eng = Engine(model, dry_run=True)
    eng.workspace(torch.zeros(2, 6, 128, 128), torch.bfloat16, True, True)
    sliced = 0
    for key, pw in eng.packs[torch.bfloat16].items():
        e = pw.entry(pw.src.data_ptr())
        n1, n2, taps, c = pw.geom[:4]
        r = torch.arange(pw.src.numel())
        idx = [torch.zeros_like(r) for _ in range(4)]
        for d in range(4):
            s = int(e.stride[d])
            if s > 0:
                q = r // s
                idx[int(e.role[d])] = q
                r = r - q * s
        keep = (idx[1] < e.N2) & (idx[3] < e.Cpad)
        sliced += int((~keep).any())
        n = idx[0] * e.N2 + idx[1]
        tap = (e.Tt - 1 - idx[2]) if e.flip else idx[2]
        di = (n * e.Kpad + tap * e.Cpad + idx[3])[keep]
        assert int(keep.sum()) == n1 * n2 * taps * c, key
        assert int(di.max()) < pw.buf.numel() and di.unique().numel() == di.numel(), key
    assert sliced >= 1
