"""Device metrics through the C ABI against the oracle: the confusion counts and the uint8
class map are integer work and must be BIT-EXACT; the derived ratios agree to fp32 rounding."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(b, c, h, w, dtype, seed, channels_last):
    g = torch.Generator(device="cuda").manual_seed(seed)
    logits = torch.randn(b, c, h, w, generator=g, device="cuda").to(dtype)
    if channels_last:
        logits = logits.contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, c, (b, h, w), generator=g, device="cuda", dtype=torch.int64)
    return logits, y


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("b,c,h,w,cl", [(2, 4, 64, 64, True), (3, 4, 17, 23, False), (1, 2, 8, 8, True), (2, 7, 32, 48, True)])
def test_confusion_counts_and_metrics(dtype, b, c, h, w, cl):
    from oracle.metrics_oracle import confusion_counts, reference_metrics
    from s2lc_b200.metrics import SegmentationMetrics, argmax_uint8
    logits, y = _case(b, c, h, w, dtype, 3, cl)
    logits[0, :, 0, 0] = 1.0                      # a full tie: the first maximum must win (torch.argmax)
    y[0, 1, 1] = -100                             # out-of-range label (CrossEntropy's default ignore) is skipped
    m = SegmentationMetrics(c, ignore_index=0)
    m.update(logits, y)
    m.update(logits, y)                           # accumulates across steps
    ref = confusion_counts(logits, y, c)
    assert torch.equal(m.counts, 2 * ref)
    pred = argmax_uint8(logits)
    assert pred.dtype == torch.uint8 and torch.equal(pred.long(), logits.float().argmax(1))
    got, want = m.compute(), reference_metrics(logits, y, c, 0)
    for k in want:
        assert torch.allclose(got[k].cpu(), want[k].cpu(), atol=1e-6), k


def test_full_size_checksum():
    """BASELINE config 2 size (bs=32, 512x512, 4 classes): every pixel is counted exactly once."""
    from s2lc_b200.metrics import SegmentationMetrics
    logits, y = _case(32, 4, 512, 512, torch.bfloat16, 5, True)
    m = SegmentationMetrics(4)
    m.update(logits, y)
    assert int(m.counts.sum()) == 32 * 512 * 512
    assert torch.equal(m.counts.sum(1), torch.bincount(y.reshape(-1), minlength=4))


def test_step_updates_metrics():
    """SegmentationStep(metrics=...) mirrors the reference's update_metrics call inside _model_step."""
    import s2lc_b200  # noqa: F401
    from s2lc_b200 import synthetic
    from s2lc_b200.efficientnet_unet import EfficientNetConfig, EfficientnetUnet
    from s2lc_b200.losses import FocalLoss
    from s2lc_b200.metrics import SegmentationMetrics
    from s2lc_b200.train_segmentation import SegmentationStep
    cfg = EfficientNetConfig(version="b0", in_channels=6, num_classes=4, class_distribution=list(synthetic.CLASS_DISTRIBUTION))
    net = EfficientnetUnet(cfg).cuda().train().set_precision("bf16")
    alpha = torch.tensor(synthetic.CLASS_DISTRIBUTION)
    alpha[1:] = 1 - alpha[1:]
    step = SegmentationStep(net, FocalLoss(alpha, 2.0, 0.0, ignore_index=0), metrics=SegmentationMetrics(4, ignore_index=0))
    x = torch.randn(2, 6, 64, 64, device="cuda")
    y = torch.randint(0, 4, (2, 64, 64), device="cuda")
    step(x, y)
    assert int(step.metrics.counts.sum()) == 2 * 64 * 64
    out = step.metrics.compute()
    assert out["confusion_matrix"].shape == (4, 4) and float(out["confusion_matrix"][0].sum()) == 0.0
