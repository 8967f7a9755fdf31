"""The well-conditioned ("trained") parity fixture.

``synthetic.state_dict_like`` gives random, untrained weights with perturbed BatchNorm buffers: logits are near-tied
and every implementation's fp32 gradients sit ~5e-3 from the fp64 evaluation, so BASELINE.json's flat bars cannot be
shown on it.  This recipe TRAINS the network first — Adam(lr 1e-3, no decay: with the reference's L2 0.05 and this learning rate
Adam shrinks every weight by ~lr per step and most channels die within 300 steps), focal loss, b0, 224x224, bs 8, on
label-carrying tiles (``synthetic.signal_tiles``) — so that logits separate and the bars become meaningful.

The 12 M-parameter trained state is too large to commit, so the recipe is the fixture: ``tests/golden/make_trained.py``
runs it on the LIVE reference modules in the build container and commits the loss curve and end-state moments
(``tests/golden/trained.json``); the CPU test replays the first steps with the oracle against that curve, and the GPU
tests re-run the recipe with the oracle on the B200 (fp32, TF32 off) to obtain the trained weights."""
from __future__ import annotations

import torch

from oracle import unet_oracle as O
from s2lc_b200 import synthetic
from s2lc_b200.efficientnet_unet import EfficientNetConfig, EfficientnetUnet

VERSION, SIZE, BATCH, STEPS, LR, WD = "b0", 224, 8, 300, 1e-3, 0.0


def train_batch(batch: int = BATCH, size: int = SIZE):
    y = synthetic.labels(batch, size, seed=2235)
    return synthetic.signal_tiles(y, seed=4321), y


def eval_batch(batch: int = BATCH, size: int = SIZE):
    y = synthetic.labels(batch, size, seed=3235)
    return synthetic.signal_tiles(y, seed=5321), y


def initial_state() -> dict:
    cfg = EfficientNetConfig(version=VERSION, in_channels=6, num_classes=4, drop_connect_rate=None,
                             class_distribution=list(synthetic.CLASS_DISTRIBUTION))
    shapes = {k: tuple(v.shape) for k, v in EfficientnetUnet(cfg).state_dict().items()}
    return synthetic.state_dict_like(shapes, seed=3)


def alpha() -> torch.Tensor:
    return O.loss_class_weights(list(synthetic.CLASS_DISTRIBUTION), masked_loss=True)


def train_oracle(device: str = "cpu", steps: int = STEPS, dtype: torch.dtype = torch.float32):
    """Runs the recipe with the oracle; returns (trained state dict on ``device``, loss curve)."""
    dev = torch.device(device)
    sd = {k: (v.to(dev, dtype) if v.is_floating_point() else v.to(dev)) for k, v in initial_state().items()}
    keys = O.used_param_keys(sd)
    for k in keys:
        sd[k] = sd[k].clone().requires_grad_(True)
    opt = torch.optim.Adam([sd[k] for k in keys], lr=LR, weight_decay=WD)
    x, y = train_batch()
    x, y, a = x.to(dev, dtype), y.to(dev), alpha().to(dev, dtype)
    curve = []
    for _ in range(steps):
        opt.zero_grad(set_to_none=True)
        loss = O.focal_loss(O.unet_forward(sd, x, VERSION, training=True, update_buffers=True), y, a, 2.0, 0.0, 0)
        loss.backward()
        opt.step()
        curve.append(float(loss))
    return {k: v.detach() for k, v in sd.items()}, curve
