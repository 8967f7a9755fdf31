"""Mosaic inference writer through the C ABI against oracle/mosaic_oracle.py.

Tile gathering and the argmax scatter are index / byte work -> BIT-EXACT on the same logits (normalisation: two fp32
operations, 1e-6).  The end-to-end mosaic goes through the network in eval mode: fp32 must agree with the torch oracle
on >= 99.9 % of the pixels (north_star's argmax-agreement bar; the remaining pixels are near-ties of random-init
logits), bf16 on >= 99 % here because random-init logits are far closer together than trained ones."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MEAN = np.array([5.1, 4.2, 3.9, 9.8, 7.7, 6.0], dtype=np.float32)
STD = np.array([2.0, 1.7, 1.9, 3.1, 2.6, 2.2], dtype=np.float32)


def _scene(c, h, w, seed=0):
    rng = np.random.default_rng(seed)
    # smooth + noise so neighbouring tiles differ and edge replication matters
    yy, xx = np.mgrid[0:h, 0:w]
    base = 1200 + 600 * np.sin(yy / 37.0)[None] + 500 * np.cos(xx / 23.0)[None] + 150 * np.arange(c)[:, None, None]
    return (base + rng.normal(0, 120, size=(c, h, w))).astype(np.int16)


@pytest.mark.parametrize("h,w,tile,margin", [(100, 130, 64, 16), (64, 64, 64, 0), (97, 33, 32, 8), (300, 200, 96, 32)])
def test_gather_tiles_matches_oracle(h, w, tile, margin):
    from oracle.mosaic_oracle import gather_tile, tile_origins
    from s2lc_b200.inference import MosaicPredictor, plan_mosaic
    scene = _scene(6, h, w)
    origins = plan_mosaic(h, w, tile, margin)
    assert np.array_equal(origins.numpy(), tile_origins(h, w, tile, margin))
    mp = MosaicPredictor(torch.nn.Identity(), MEAN, STD, tile=tile, margin=margin)
    x = mp.gather(torch.from_numpy(scene).cuda(), origins.cuda()).cpu().numpy()
    for i, (t, l) in enumerate(origins.tolist()):
        np.testing.assert_allclose(x[i], gather_tile(scene, t, l, tile, MEAN, STD), rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("h,w,tile,margin,classes", [(100, 130, 64, 16, 10), (64, 64, 64, 0, 2), (97, 33, 32, 8, 16),
                                                     (50, 70, 32, 0, 7)])
def test_argmax_mosaic_bit_exact(h, w, tile, margin, classes, dtype):
    from oracle.mosaic_oracle import write_tile
    from s2lc_b200.inference import MosaicPredictor, plan_mosaic
    origins = plan_mosaic(h, w, tile, margin)
    g = torch.Generator().manual_seed(3)
    logits = torch.randn(origins.shape[0], tile, tile, classes, generator=g).to(dtype)
    logits[0, :, :, 1] = logits[0, :, :, 0]                       # exact ties: the first maximum must win
    mp = MosaicPredictor(torch.nn.Identity(), MEAN, STD, tile=tile, margin=margin)
    mosaic = torch.full((h, w), 255, dtype=torch.uint8, device="cuda")
    mp.scatter_argmax(logits.cuda().permute(0, 3, 1, 2), origins.cuda(), mosaic)
    want = np.full((h, w), 255, dtype=np.uint8)
    for i, (t, l) in enumerate(origins.tolist()):
        write_tile(want, logits[i].float().permute(2, 0, 1).numpy(), t, l, margin)
    assert (want != 255).all()                                    # the kept centres cover the scene exactly
    assert np.array_equal(mosaic.cpu().numpy(), want)


def _oracle_mosaic(sd, scene, tile, margin, version):
    from oracle import unet_oracle as O
    from oracle.mosaic_oracle import predict_mosaic

    def logits_fn(x):
        with torch.no_grad():
            return O.unet_forward(sd, torch.from_numpy(x), version, training=False).numpy()
    return predict_mosaic(scene, tile, margin, MEAN, STD, logits_fn)


@pytest.mark.parametrize("precision,bar", [("fp32", 0.999), ("bf16", 0.99)])
def test_mosaic_end_to_end(precision, bar):
    from _parity import build_model, synthetic
    from s2lc_b200.inference import MosaicPredictor, plan_mosaic, rank_batches
    model, sd = build_model("b0", precision)
    scene = _scene(6, 150, 200, seed=5)
    # one training-mode pass so the running statistics are not the (0, 1) initial values
    with torch.no_grad():
        model(synthetic.tiles(2, 224).cuda())
    sd_now = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    mp = MosaicPredictor(model, MEAN, STD, tile=96, margin=16, batch=4)
    got = mp.predict(torch.from_numpy(scene))
    assert model.training                                          # the caller's mode is restored
    assert got.shape == (150, 200) and got.dtype == torch.uint8
    want = _oracle_mosaic(sd_now, scene, 96, 16, "b0")
    agree = float((got.cpu().numpy() == want).mean())
    assert agree >= bar, agree
    # rank striding: two "ranks" predicting without the final reduce fill disjoint halves whose sum is the mosaic
    parts = [MosaicPredictor(model, MEAN, STD, tile=96, margin=16, batch=4, rank=r, world_size=2)
             .predict(torch.from_numpy(scene), reduce=False) for r in range(2)]
    # eval mode is bit-reproducible (deterministic SE pooling): the two halves sum to exactly the one-rank mosaic
    assert torch.equal(parts[0] + parts[1], got)
    own = torch.zeros(150, 200, dtype=torch.int64)
    keep = 96 - 2 * 16
    for r in range(2):                                             # ownership is exact: batch i belongs to rank i % 2
        for s, e in rank_batches(plan_mosaic(150, 200, 96, 16).shape[0], 4, r, 2):
            for t, l in plan_mosaic(150, 200, 96, 16)[s:e].tolist():
                own[max(t + 16, 0):t + 16 + keep, max(l + 16, 0):l + 16 + keep] = r
    for r in range(2):
        assert int((parts[r].cpu()[own != r] != 0).sum()) == 0     # a rank never writes outside its own tiles


def test_full_size_scene_properties():
    """A 2048 x 2048 scene in 512-px tiles with 64 px of context (the reference's tile size) through b5 in bf16.
    A uint8 class map is byte work: eval mode has no RNG and no atomics (the SE pooling reduces in a fixed order),
    so two predictions must be BIT-IDENTICAL — also when the engine chunk differs (per-sample arithmetic only);
    plus every class index < n_classes and a non-constant map."""
    from _parity import build_model, synthetic
    from s2lc_b200.inference import MosaicPredictor
    model, _ = build_model("b5", "bf16")
    with torch.no_grad():
        model(synthetic.tiles(2, 512).cuda())                      # running statistics away from their (0, 1) init
    scene = torch.from_numpy(_scene(6, 2048, 2048, seed=9)).cuda()
    mp = MosaicPredictor(model, MEAN, STD, tile=512, margin=64, batch=8)
    a = mp.predict(scene)
    b = mp.predict(scene)
    assert torch.equal(a, b)
    c = MosaicPredictor(model, MEAN, STD, tile=512, margin=64, batch=8, engine_batch=4).predict(scene)
    assert torch.equal(a, c)
    assert int(a.max()) < model.config.num_classes
    assert len(torch.unique(a)) > 1
