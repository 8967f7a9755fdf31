"""Model-level parity on a real B200: the CUDA EfficientnetUnet + fused loss (through the public nn.Module /
loss call signatures, i.e. through the libs2lc C ABI) against the oracle on identical synthetic tiles + weights.

Tolerances (BASELINE.json north_star): fp32 mode rel-err (L2) <= 1e-4 on logits and on every parameter gradient;
bf16 mode rel-err <= 2e-2 on logits and >= 99.9 % argmax agreement."""
import pytest
import torch

from _parity import build_model, grad_report, rel, run_pair
from s2lc_b200 import synthetic

pytestmark = pytest.mark.gpu

FP32_TOL = 1e-4


@pytest.mark.parametrize("version,size,batch,loss,dc", [("b0", 224, 2, "focal", None), ("b0", 224, 2, "ce", None),
                                                        ("b0", 224, 2, "focal", 0.2), ("b5", 224, 1, "focal", None),
                                                        ("b0", 256, 1, "focal", None), ("b7", 256, 1, "focal", None)])
def test_fp32_matches_oracle(version, size, batch, loss, dc):
    res = run_pair(version, size, batch, "fp32", loss, dc)
    assert abs(res["loss"] - res["o_loss"]) <= 1e-5 * abs(res["o_loss"])
    assert rel(res["logits"], res["o_logits"]) <= FP32_TOL
    rows = grad_report(res)
    assert set(res["grads"]) == set(res["o_grads"])
    bad = [(r[0], r[1]) for r in rows if r[0] > FP32_TOL]
    assert not bad, bad[:10]


def test_fp32_matches_golden_fixture(golden_meta, golden_arrays):
    """The same run against the fixture produced by the LIVE reference (tier A, 224x224)."""
    case = golden_meta["cases"]["b0_224_A"]
    res = run_pair("b0", 224, 2, "fp32")
    assert abs(res["loss"] - case["loss"]) <= 1e-5 * abs(case["loss"])
    sub = torch.from_numpy(golden_arrays["b0_224_A/logits_sub"])
    assert rel(res["logits"][:, :, ::8, ::8], sub) <= FP32_TOL
    for k, want in case["grad_moments"].items():
        got = float(res["grads"][k].double().norm())
        assert abs(got - want[1]) <= 2e-4 * want[1] + 1e-6, (k, got, want[1])


def test_b5_512_fp32_tier_b():
    """BASELINE.json config[0]: b5, one 512x512 tile, fwd+bwd, against the generalised (tier B) oracle."""
    res = run_pair("b5", 512, 1, "fp32")
    assert rel(res["logits"], res["o_logits"]) <= FP32_TOL
    rows = grad_report(res)
    assert rows[0][0] <= FP32_TOL, rows[:5]


@pytest.mark.parametrize("version,size,batch", [("b0", 224, 2), ("b5", 224, 2), ("b5", 512, 1)])
def test_bf16_within_tolerance(version, size, batch):
    res = run_pair(version, size, batch, "bf16")
    assert rel(res["logits"], res["o_logits"]) <= 2e-2
    agree = float((res["logits"].argmax(1) == res["o_logits"].argmax(1)).float().mean())
    assert agree >= 0.999, agree
    assert abs(res["loss"] - res["o_loss"]) <= 2e-2 * abs(res["o_loss"])
    rows = grad_report(res, floor=1e-4)
    # gradients: report-level check — the bulk of the tensors within 2e-2, none wildly off
    frac_ok = sum(r[0] <= 2e-2 for r in rows) / len(rows)
    assert frac_ok >= 0.9 and rows[0][0] < 0.25, (frac_ok, rows[:5])


def test_running_stats_and_eval_mode(golden_meta):
    import zlib
    from oracle import unet_oracle as O
    model, sd = build_model("b0", "fp32")
    x = synthetic.tiles(2, 224)
    with torch.no_grad():
        model(x.cuda())
    case = golden_meta["cases"]["b0_224_A"]
    got = model.state_dict()
    for k, want in case["buffer_moments"].items():
        v = got[k].double().flatten().cpu()
        assert abs(float(v.norm()) - want[1]) <= 1e-5 * want[1] + 1e-7, k
    # eval mode: running statistics, no drop-connect; compare with the oracle in eval mode on the updated buffers
    model.eval()
    with torch.no_grad():
        ev = model(x.cuda()).float().cpu()
    sd_now = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    with torch.no_grad():
        ref = O.unet_forward(sd_now, x, "b0", training=False)
    assert rel(ev, ref) <= FP32_TOL


def test_state_dict_roundtrip(state_shapes):
    model, sd = build_model("b5", "bf16")
    got = model.state_dict()
    assert {k: list(v.shape) for k, v in got.items()} == state_shapes["b5"]
    for k in sd:
        assert torch.equal(got[k].cpu(), sd[k]), k


def test_training_curve_matches_reference(golden_meta):
    """First 20 steps of Adam(lr=1e-3, wd=0.05) on one fixed batch vs the curve of the live reference (fp32)."""
    from s2lc_b200.losses import FocalLoss
    from _parity import alpha_weights
    want = golden_meta["curve_b0_224_adam1e-3"]
    model, _ = build_model("b0", "fp32")
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, weight_decay=0.05)
    fn = FocalLoss(alpha_weights(), 2.0, 0.0, ignore_index=0)
    x, y = synthetic.tiles(2, 224).cuda(), synthetic.labels(2, 224).cuda()
    for step in range(20):
        opt.zero_grad()
        loss = fn(model(x), y)
        loss.backward()
        opt.step()
        assert abs(float(loss) - want[step]) <= 2e-3 * abs(want[step]) + 1e-4, (step, float(loss), want[step])


def test_cpu_tensor_is_rejected():
    model, _ = build_model("b0", "fp32")
    with pytest.raises(RuntimeError):
        model(torch.zeros(1, 6, 64, 64))
