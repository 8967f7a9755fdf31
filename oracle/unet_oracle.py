"""ORACLE — TEST INFRASTRUCTURE ONLY.  Not shipped, not measured as the product.

A CPU/fp32 restatement, in plain functional PyTorch, of the reference's EfficientNet-UNet training
step (MaxWolf-01/sentinel2-landcover-classification).  Only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s cpu_baseline / ``--impl reference`` legs may import this package.

Parity status: **pinned against the live reference** — ``tests/golden/make_golden.py`` imports the
unmodified reference from a writable copy of ``/root/reference`` (with a 4-line ``timm`` stub, see
SURVEY.md F2/F3), runs it on the deterministic synthetic tiles/weights of ``synthetic.py`` and
commits the outputs under ``tests/golden/``; ``tests/test_oracle_golden.py`` checks this file
against those fixtures.  The reference itself ships no golden vectors or value-level tests
(SURVEY.md §4), so the fixtures are outputs of the reference run in the build container.

The arithmetic lives in PyTorch (``requirements.txt:1`` torch>=2.1.2, no lock file; installed
2.11.0+cu128): F.conv2d, F.batch_norm, F.silu, F.conv_transpose2d, F.cross_entropy.

Every function cites the reference file:line it follows (paths relative to /root/reference).
Weights come in as a ``state_dict`` with the reference's own keys, which also pins the key naming.
"""
from __future__ import annotations

import math
import re
from dataclasses import dataclass

import torch
import torch.nn.functional as F

# src/modules/efficientnet_unet.py:35-46 — (width, depth, resolution, dropout) per version
_SCALING = {
    "b0": (1.0, 1.0), "b1": (1.0, 1.1), "b2": (1.1, 1.2), "b3": (1.2, 1.4),
    "b4": (1.4, 1.8), "b5": (1.6, 2.2), "b6": (1.8, 2.6), "b7": (2.0, 3.1),
}
# src/modules/efficientnet_unet.py:201-209 — (repeat, kernel, stride, expand, in, out); se 0.25 everywhere
_STAGES = [(1, 3, 1, 1, 32, 16), (2, 3, 2, 6, 16, 24), (2, 5, 2, 6, 24, 40), (3, 3, 2, 6, 40, 80),
           (3, 5, 1, 6, 80, 112), (4, 5, 2, 6, 112, 192), (1, 3, 1, 6, 192, 320)]
# src/modules/efficientnet_unet.py:153-165 — concat widths per decoder level
_CAT = {"b0": [592, 296, 152, 80, 38, 32], "b1": [592, 296, 152, 80, 38, 32], "b2": [600, 304, 152, 80, 38, 32],
        "b3": [608, 304, 160, 88, 38, 32], "b4": [624, 312, 160, 88, 38, 32], "b5": [640, 320, 168, 88, 38, 32],
        "b6": [656, 328, 168, 96, 38, 32], "b7": [672, 336, 176, 96, 38, 32]}
BN_EPS_ENC = 1e-3        # efficientnet_unet.py:24
BN_MOM_ENC = 1 - 0.99    # efficientnet_unet.py:23,53 (flipped in __post_init__)
BN_EPS_DEC = 1e-5        # efficientnet_unet.py:171,174 (nn.BatchNorm2d defaults)
BN_MOM_DEC = 0.1


def round_filters(filters: int, width: float, divisor: int = 8) -> int:
    """efficientnet_unet.py:266-278."""
    f = filters * width
    new = max(divisor, int(f + divisor / 2) // divisor * divisor)
    if new < 0.9 * f:
        new += divisor
    return int(new)


@dataclass
class OBlock:
    k: int
    s: int
    cin: int
    cexp: int
    cout: int
    se: int
    expand: bool
    residual: bool


def blocks_for(version: str) -> tuple[int, list[OBlock], int]:
    """Architecture generation, efficientnet_unet.py:184-228 and MBConvBlock.__init__ :303-375.

    Note the quirk at :222-224 + :348: after the first block of a stage ``input_filters`` is
    overwritten with ``output_filters``, and SE width = max(1, int(input_filters*0.25)).
    """
    width, depth = _SCALING[version]
    stem = round_filters(32, width)
    out: list[OBlock] = []
    for r, k, s, e, i, o in _STAGES:
        ci, co = round_filters(i, width), round_filters(o, width)
        rep = int(math.ceil(depth * r))
        for j in range(rep):
            cin, stride = (ci, s) if j == 0 else (co, 1)
            out.append(OBlock(k=k, s=stride, cin=cin, cexp=cin * e, cout=co, se=max(1, int(cin * 0.25)),
                              expand=(e != 1), residual=(stride == 1 and cin == co)))
    return stem, out, round_filters(1280, width)


def same_pad(x: torch.Tensor, k: int, s: int) -> torch.Tensor:
    """TF-'same' asymmetric padding, Conv2dSamePadding.forward efficientnet_unet.py:288-296."""
    h, w = x.shape[-2:]
    oh, ow = math.ceil(h / s), math.ceil(w / s)
    ph = max((oh - 1) * s + k - h, 0)
    pw = max((ow - 1) * s + k - w, 0)
    if ph > 0 or pw > 0:
        x = F.pad(x, [pw // 2, pw - pw // 2, ph // 2, ph - ph // 2])
    return x


class _BN:
    """Training/eval BatchNorm over a state dict; optionally updates the running buffers in place."""

    def __init__(self, sd: dict, training: bool, update_buffers: bool):
        self.sd, self.training, self.update = sd, training, update_buffers

    def __call__(self, x: torch.Tensor, prefix: str, mom: float, eps: float) -> torch.Tensor:
        rm, rv = self.sd[prefix + ".running_mean"], self.sd[prefix + ".running_var"]
        if self.training and not self.update:
            rm, rv = rm.clone(), rv.clone()
        y = F.batch_norm(x, rm, rv, self.sd[prefix + ".weight"], self.sd[prefix + ".bias"],
                         training=self.training, momentum=mom, eps=eps)
        if self.training and self.update:
            self.sd[prefix + ".num_batches_tracked"] += 1
        return y


def mbconv(sd, bn: _BN, x, p: str, b: OBlock, dc):
    """MBConvBlock.forward, efficientnet_unet.py:377-387 (+ __init__ :312-375 for the layer list)."""
    identity = x
    if b.expand:
        x = F.conv2d(x, sd[f"{p}.stem.0.weight"])
        x = F.silu(bn(x, f"{p}.stem.1", BN_MOM_ENC, BN_EPS_ENC))
        dw, dwbn = f"{p}.stem.3", f"{p}.stem.4"
    else:
        dw, dwbn = f"{p}.stem.0", f"{p}.stem.1"
    x = F.conv2d(same_pad(x, b.k, b.s), sd[dw + ".weight"], stride=b.s, groups=b.cexp)
    x = F.silu(bn(x, dwbn, BN_MOM_ENC, BN_EPS_ENC))
    # squeeze-excite :346-362, applied at :381
    g = x.mean(dim=(2, 3), keepdim=True)
    g = F.silu(F.conv2d(g, sd[f"{p}.squeeze_excitation.1.weight"], sd[f"{p}.squeeze_excitation.1.bias"]))
    g = F.conv2d(g, sd[f"{p}.squeeze_excitation.3.weight"], sd[f"{p}.squeeze_excitation.3.bias"])
    x = x * torch.sigmoid(g)
    x = F.conv2d(x, sd[f"{p}.final_layer.0.weight"])
    x = bn(x, f"{p}.final_layer.1", BN_MOM_ENC, BN_EPS_ENC)
    if b.residual:
        if dc is not None and float(dc[1]) < 1.0:  # _drop_connect :390-398, mask injected: x / keep * floor(keep + U)
            x = x / float(dc[1]) * dc[0].view(-1, 1, 1, 1).to(x.dtype)
        x = x + identity
    return x


def unet_forward(sd: dict, x: torch.Tensor, version: str, *, training: bool = True,
                 update_buffers: bool = False, dc_masks=None,
                 skip_rule: str = "generalised", probe: dict | None = None) -> torch.Tensor:
    """EfficientnetUnet.forward efficientnet_unet.py:125-138 + EfficientNet.encode :251-263.

    ``skip_rule="reference"`` keeps the literal ``!= (7, 7)`` test of :259 (tier A: only valid at
    224x224); ``"generalised"`` drops the collected map whose HxW equals the head's (tier B,
    SURVEY.md F1) and is bit-identical to the reference at 224x224.
    ``dc_masks`` = (binary [n_blocks, B], keep [n_blocks]) as made by synthetic.drop_connect_masks;
    None means drop_connect_rate=None (a legal config value, :26).
    """
    stem_c, blocks, head_c = blocks_for(version)
    bn = _BN(sd, training, update_buffers)
    identity = x
    h_in, w_in = x.shape[-2:]
    x = F.conv2d(same_pad(x, 3, 2), sd["encoder.stem.0.weight"], stride=2)          # :191-197
    x = F.silu(bn(x, "encoder.stem.1", BN_MOM_ENC, BN_EPS_ENC))
    head_hw = (math.ceil(h_in / 32), math.ceil(w_in / 32))
    maps: list[torch.Tensor] = []
    for i, b in enumerate(blocks):                                                  # :255-260
        x = mbconv(sd, bn, x, f"encoder.blocks.{i}", b, None if dc_masks is None else (dc_masks[0][i], dc_masks[1][i]))
        hw = tuple(x.shape[-2:])
        drop = (hw == (7, 7)) if skip_rule == "reference" else (hw == head_hw)
        if hw not in [tuple(m.shape[-2:]) for m in maps] and not drop:
            maps.insert(0, x)
    x = F.conv2d(x, sd["encoder.conv_head.0.weight"])                               # :229-238, :261
    x = F.silu(bn(x, "encoder.conv_head.1", BN_MOM_ENC, BN_EPS_ENC))
    for j, skip in enumerate(maps):                                                 # :129-132
        x = F.conv_transpose2d(x, sd[f"up_convs.{j}.weight"], sd[f"up_convs.{j}.bias"], stride=2)
        x = torch.cat([x, skip], dim=1)
        x = _double_conv(sd, bn, x, f"double_convs.{j}", probe)
    x = F.conv_transpose2d(x, sd["input_up_conv.weight"], sd["input_up_conv.bias"], stride=2)   # :133-136
    x = torch.cat([x, identity], dim=1)
    x = _double_conv(sd, bn, x, "input_double_conv", probe)
    return F.conv2d(x, sd["out_conv1x1.weight"], sd["out_conv1x1.bias"])           # :137


def _double_conv(sd, bn: _BN, x, p: str, probe: dict | None = None):
    """_double_conv efficientnet_unet.py:168-176: 2x (Conv3x3 p=1 bias -> BN -> ReLU).
    ``probe`` (debug aid) records the concat input and the two raw conv outputs with retain_grad()."""
    def rec(name, t):
        if probe is not None:
            if t.requires_grad:
                t.retain_grad()
            probe[p + "/" + name] = t
        return t
    rec("cat", x)
    z1 = rec("z1", F.conv2d(x, sd[p + ".0.weight"], sd[p + ".0.bias"], padding=1))
    x = F.relu(bn(z1, p + ".1", BN_MOM_DEC, BN_EPS_DEC))
    z2 = rec("z2", F.conv2d(x, sd[p + ".3.weight"], sd[p + ".3.bias"], padding=1))
    return F.relu(bn(z2, p + ".4", BN_MOM_DEC, BN_EPS_DEC))


# --------------------------------------------------------------------------------------------- losses
def loss_class_weights(class_distribution, masked_loss: bool) -> torch.Tensor:
    """get_loss src/losses.py:26-29: w = p; w[int(masked):] = 1 - w[int(masked):]."""
    w = torch.tensor(class_distribution, dtype=torch.float32).clone()
    k = int(masked_loss)
    w[k:] = 1 - w[k:]
    return w


def focal_loss(logits, y, alpha, gamma: float, label_smoothing: float = 0.0, ignore_index: int = -100,
               reduce_type: str = "mean"):
    """FocalLoss.__call__ src/losses.py:77-89; mean over ALL pixels (_reduce :131-138)."""
    ce = F.cross_entropy(logits, y, label_smoothing=label_smoothing, ignore_index=ignore_index, reduction="none")
    pt = torch.exp(-ce)
    a = alpha.to(logits.device).gather(0, y.reshape(-1)).view(*y.shape)
    fl = a * (1 - pt) ** gamma * ce
    return fl.mean() if reduce_type == "mean" else fl.sum()


def weighted_ce(logits, y, weight, label_smoothing: float = 0.0, ignore_index: int = -100):
    """nn.CrossEntropyLoss(weight, label_smoothing, ignore_index) src/losses.py:36-40 (weight-normalised mean)."""
    return F.cross_entropy(logits, y, weight=None if weight is None else weight.to(logits.device),
                           label_smoothing=label_smoothing, ignore_index=ignore_index)


# --------------------------------------------------------------------------------------------- step
def used_param_keys(sd: dict) -> list[str]:
    """Trainable tensors that receive a gradient in the UNet forward (not encoder.fc.3.*)."""
    return [k for k in param_keys(sd) if not k.startswith("encoder.fc.")]


def param_keys(sd: dict) -> list[str]:
    return [k for k in sd if not re.search(r"running_mean|running_var|num_batches_tracked", k)]


def loss_and_grads(sd: dict, x, y, version: str, loss: str = "focal", *, alpha=None, gamma: float = 2.0,
                   ignore_index: int = 0, skip_rule: str = "generalised", dc_masks=None, autocast_bf16: bool = False):
    """One forward+loss+backward: SegmentationModule._model_step src/train_segmentation.py:129-147 minus logging.

    Returns (logits, loss, {key: grad}).  With ``autocast_bf16`` the net+loss run under
    torch.autocast(bfloat16) as Lightning's precision="bf16" does (src/configs/segmentation.py:153).
    """
    leaves = {k: v.detach().clone().requires_grad_(k in set(used_param_keys(sd))) if v.is_floating_point() else v.clone()
              for k, v in sd.items()}
    dev = x.device.type
    with torch.autocast(dev, dtype=torch.bfloat16, enabled=autocast_bf16):
        logits = unet_forward(leaves, x, version, training=True, skip_rule=skip_rule, dc_masks=dc_masks)
        if loss == "focal":
            val = focal_loss(logits, y, alpha, gamma, 0.0, ignore_index)
        else:
            val = weighted_ce(logits, y, alpha, 0.0, ignore_index)
    keys = used_param_keys(sd)
    grads = torch.autograd.grad(val, [leaves[k] for k in keys])
    return logits.detach(), val.detach(), dict(zip(keys, grads))
