"""Device input pipeline through the C ABI against the numpy oracle: crop / flips / label remap
are index and integer work -> BIT-EXACT; the normalisation is two fp32 operations -> tolerance 1 ulp-level (1e-6 rel)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

NAMES = {1: "urban", 2: "urban", 3: "crops", 4: "water", 7: "forest", 9: "glacier"}
NEW = ["other", "urban", "crops", "forest"]     # "water"/"glacier" are not kept -> 0


@pytest.mark.parametrize("n,c,h,w,crop,augment", [(4, 6, 64, 64, 32, True), (3, 6, 40, 56, 24, True), (2, 6, 32, 32, 32, False),
                                                   (1, 3, 17, 19, 8, True)])
def test_prepare_batch_matches_oracle(n, c, h, w, crop, augment):
    from oracle.input_oracle import prepare_sample, remap_labels
    from s2lc_b200.input_pipeline import DeviceBatchPreparer, build_label_lut, draw_augmentation
    rng = np.random.default_rng(0)
    raw = rng.integers(-200, 12000, size=(n, c, h, w), dtype=np.int16)
    lab = rng.integers(0, 12, size=(n, h, w), dtype=np.uint8)
    mean = rng.uniform(3.0, 9.0, size=c).astype(np.float32)
    std = rng.uniform(1.0, 4.0, size=c).astype(np.float32)
    g = torch.Generator().manual_seed(7)
    crop_yx, flips = draw_augmentation(n, (h, w), crop, augment, generator=g)
    prep = DeviceBatchPreparer(mean, std, crop, build_label_lut(NAMES, NEW))
    x, y = prep(torch.from_numpy(raw).cuda(), torch.from_numpy(lab).cuda(), crop_yx, flips)
    for i in range(n):
        xo, yo = prepare_sample(raw[i], remap_labels(lab[i], NAMES, NEW), int(crop_yx[i, 0]), int(crop_yx[i, 1]), crop,
                                bool(flips[i, 0]), bool(flips[i, 1]), mean, std)
        assert np.array_equal(y[i].cpu().numpy(), yo)
        np.testing.assert_allclose(x[i].cpu().numpy(), xo, rtol=1e-6, atol=1e-6)


def test_full_size_properties():
    """bs=32 of 512x512 6-band tiles cropped to 224 (the reference's random_crop_size): identity LUT and no flip
    reproduce the source window bit-exactly after de-normalisation is NOT assumed — check the label half exactly and
    the image half through linearity (x(raw) - x(0) is proportional to raw)."""
    from s2lc_b200.input_pipeline import DeviceBatchPreparer, draw_augmentation
    g = torch.Generator().manual_seed(1)
    raw = torch.randint(0, 10000, (32, 6, 512, 512), generator=g, dtype=torch.int16).cuda()
    lab = torch.randint(0, 23, (32, 512, 512), generator=g, dtype=torch.uint8).cuda()
    crop_yx, _ = draw_augmentation(32, (512, 512), 224, True, generator=g)
    prep = DeviceBatchPreparer([1.0] * 6, [2.0] * 6, 224)
    x, y = prep(raw, lab, crop_yx, None)
    x0, _ = prep(torch.zeros_like(raw), lab, crop_yx, None)
    for i in (0, 31):
        t, l = int(crop_yx[i, 0]), int(crop_yx[i, 1])
        assert torch.equal(y[i], lab[i, t:t + 224, l:l + 224].long())
        win = raw[i, :, t:t + 224, l:l + 224].float()
        assert torch.allclose(x[i] - x0[i], win / 510.0, rtol=1e-6, atol=1e-5)


def test_bad_arguments_raise():
    from s2lc_b200.input_pipeline import DeviceBatchPreparer
    prep = DeviceBatchPreparer([0.0] * 6, [1.0] * 6, 16)
    raw = torch.zeros(1, 6, 32, 32, dtype=torch.int16, device="cuda")
    lab = torch.zeros(1, 32, 32, dtype=torch.uint8, device="cuda")
    with pytest.raises(ValueError):
        prep(raw, lab, torch.tensor([[20, 0]]))          # window leaves the tile
    with pytest.raises(ValueError):
        prep(raw.float(), lab, torch.tensor([[0, 0]]))   # wrong dtype
