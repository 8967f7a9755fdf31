"""The oracle (oracle/unet_oracle.py) against fixtures produced by the LIVE reference
(tests/golden/make_golden.py).  CPU only."""
import zlib

import numpy as np
import pytest
import torch

from oracle import unet_oracle as O
from s2lc_b200 import synthetic


def _moments(t, key):
    v = t.detach().double().flatten()
    h = zlib.crc32(key.encode())
    i = torch.arange(v.numel(), dtype=torch.float64)
    r = torch.sin(i * (0.001 * (1 + h % 97)) + (h % 13))
    return [float(v.sum()), float(v.norm()), float((v * r).sum())]


def _alpha():
    return O.loss_class_weights(list(synthetic.CLASS_DISTRIBUTION), masked_loss=True)


def _close_moments(got, want, tol, what):
    # l2 norms must agree relatively; sum/projection agree to tol * l2 (they can cancel to ~0).
    # FLOOR: some parameters have a theoretically-zero gradient (a per-channel bias feeding conv->BatchNorm,
    # e.g. the decoder conv biases) and hold only ~1e-8 rounding noise in the reference.
    floor = 1e-6
    assert abs(got[1] - want[1]) <= tol * want[1] + floor, (what, got, want)
    for g, w in ((got[0], want[0]), (got[2], want[2])):
        assert abs(g - w) <= 10 * tol * want[1] + floor, (what, got, want)


CASES = ["b0_224_A", "b0_224_A_ce", "b0_224_A_dc", "b5_224_A", "b0_256_B", "b5_512_B", "b7_256_B"]


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference(name, golden_meta, golden_arrays, state_shapes):
    case = golden_meta["cases"][name]
    v, size, batch = case["version"], case["size"], case["batch"]
    sd = synthetic.state_dict_like({k: tuple(s) for k, s in state_shapes[v].items()}, seed=0)
    x, y = synthetic.tiles(batch, size), synthetic.labels(batch, size)
    _, blocks, _ = O.blocks_for(v)
    masks = synthetic.drop_connect_masks(len(blocks), batch, case["dc_rate"])
    rule = "reference" if case["tier"] == "A" else "generalised"
    logits, loss, grads = O.loss_and_grads(sd, x, y, v, loss=case["loss_type"], alpha=_alpha(), gamma=2.0,
                                           ignore_index=0, skip_rule=rule, dc_masks=masks)
    assert abs(float(loss) - case["loss"]) <= 2e-6 * abs(case["loss"])
    sub = torch.from_numpy(golden_arrays[f"{name}/logits_sub"])
    assert torch.allclose(logits[:, :, ::8, ::8], sub, rtol=1e-5, atol=2e-5)
    _close_moments(_moments(logits, "logits"), case["logits_moments"], 1e-5, "logits")
    assert set(grads) == set(case["grad_moments"]), "set of parameters that receive a gradient differs"
    for k, g in grads.items():
        _close_moments(_moments(g, k), case["grad_moments"][k], 2e-4, k)
    if case["tier"] == "A" and case["dc_rate"] is None:
        # the declared generalisation is bit-identical to the literal rule at 224
        lg2, _, _ = O.loss_and_grads(sd, x, y, v, loss=case["loss_type"], alpha=_alpha(), skip_rule="generalised")
        assert torch.equal(lg2, logits)


def test_oracle_running_stats(golden_meta, state_shapes):
    case = golden_meta["cases"]["b0_224_A"]
    sd = synthetic.state_dict_like({k: tuple(s) for k, s in state_shapes["b0"].items()}, seed=0)
    with torch.no_grad():
        O.unet_forward(sd, synthetic.tiles(2, 224), "b0", training=True, update_buffers=True, skip_rule="reference")
    for k, want in case["buffer_moments"].items():
        _close_moments(_moments(sd[k], k), want, 1e-5, k)


def test_oracle_state_dict_contract(state_shapes):
    # the oracle consumes every non-fc key of the reference's state dict, for every version
    for v, shapes in state_shapes.items():
        stem, blocks, head = O.blocks_for(v)
        assert shapes["encoder.stem.0.weight"] == [stem, 6, 3, 3]
        assert shapes["encoder.conv_head.0.weight"][0] == head
        n_blocks = 1 + max(int(k.split(".")[2]) for k in shapes if k.startswith("encoder.blocks."))
        assert n_blocks == len(blocks)
        for i, b in enumerate(blocks):
            dw = f"encoder.blocks.{i}.stem.{3 if b.expand else 0}.weight"
            assert shapes[dw] == [b.cexp, 1, b.k, b.k], (v, i)
            assert shapes[f"encoder.blocks.{i}.squeeze_excitation.1.weight"] == [b.se, b.cexp, 1, 1], (v, i)
            assert shapes[f"encoder.blocks.{i}.final_layer.0.weight"] == [b.cout, b.cexp, 1, 1], (v, i)


@pytest.mark.parametrize("nm,kw", [
    ("focal_g2", dict(gamma=2.0, ignore_index=0, masked=True)),
    ("focal_g0.5_noignore", dict(gamma=0.5, ignore_index=-100, masked=None)),
    ("focal_g2_sum", dict(gamma=2.0, ignore_index=0, masked=True, reduce_type="sum")),
])
def test_oracle_focal(nm, kw, golden_arrays):
    lg = torch.from_numpy(golden_arrays["loss/logits"]).requires_grad_(True)
    y = torch.from_numpy(golden_arrays["loss/y"])
    alpha = _alpha() if kw["masked"] else torch.ones(4)
    v = O.focal_loss(lg, y, alpha, kw["gamma"], 0.0, kw["ignore_index"], kw.get("reduce_type", "mean"))
    v.backward()
    assert abs(float(v) - float(golden_arrays[f"loss/{nm}/value"])) < 1e-6 * max(1.0, abs(float(v)))
    assert torch.allclose(lg.grad, torch.from_numpy(golden_arrays[f"loss/{nm}/dlogits"]), rtol=1e-5, atol=1e-9)


@pytest.mark.parametrize("nm,weighted,ii", [("ce_weighted", True, 0), ("ce_plain", False, -100)])
def test_oracle_ce(nm, weighted, ii, golden_arrays):
    lg = torch.from_numpy(golden_arrays["loss/logits"]).requires_grad_(True)
    y = torch.from_numpy(golden_arrays["loss/y"])
    v = O.weighted_ce(lg, y, _alpha() if weighted else None, 0.0, ii)
    v.backward()
    assert abs(float(v) - float(golden_arrays[f"loss/{nm}/value"])) < 1e-6
    assert torch.allclose(lg.grad, torch.from_numpy(golden_arrays[f"loss/{nm}/dlogits"]), rtol=1e-5, atol=1e-9)


def test_oracle_training_curve(golden_meta, state_shapes):
    """First steps of the Adam(lr=1e-3, wd=0.05) loss curve of the reference loop (fp32)."""
    want = golden_meta["curve_b0_224_adam1e-3"]
    sd = synthetic.state_dict_like({k: tuple(s) for k, s in state_shapes["b0"].items()}, seed=0)
    keys = O.param_keys(sd)
    params = {k: sd[k].clone().requires_grad_(True) for k in keys}
    opt = torch.optim.Adam(list(params.values()), lr=1e-3, weight_decay=0.05)
    x, y = synthetic.tiles(2, 224), synthetic.labels(2, 224)
    state = dict(sd)
    state.update(params)
    for step in range(6):
        opt.zero_grad()
        loss = O.focal_loss(O.unet_forward(state, x, "b0", training=True, update_buffers=True, skip_rule="reference"),
                            y, _alpha(), 2.0, 0.0, 0)
        loss.backward()
        opt.step()
        assert abs(float(loss) - want[step]) <= 5e-4 * abs(want[step]), (step, float(loss), want[step])


def test_trained_recipe_first_steps_follow_the_reference_curve():
    """tests/_trained.py (the well-conditioned fixture): the oracle replays the first steps of the recipe on the CPU and
    lands on the loss curve the LIVE reference produced (tests/golden/make_trained.py -> trained.json)."""
    import json
    import os
    import _trained as T
    want = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "trained.json")))
    assert want["recipe"] == {"version": T.VERSION, "size": T.SIZE, "batch": T.BATCH, "steps": T.STEPS, "lr": T.LR,
                              "weight_decay": T.WD}
    torch.set_num_threads(max(torch.get_num_threads(), 4))
    _, curve = T.train_oracle("cpu", steps=3)
    for i in range(3):
        assert abs(curve[i] - want["curve"][i]) <= 1e-4 * abs(want["curve"][i]), (i, curve[i], want["curve"][i])
