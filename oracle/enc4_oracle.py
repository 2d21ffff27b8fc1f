"""CPU restatement (plain torch ops on a {name: tensor} dict) of the concept encoder's convolutional trunk,
``Encoder4.encoder`` (ldm/modules/diffusionmodules/openaimodel_enc.py:995-1014; EncResBlock :969-989; BatchNorm2d with the
torch defaults eps 1e-5 / momentum 0.1, batch statistics in training mode and running statistics in eval mode).  Autograd
through these ops gives the reference's parameter gradients.  TEST INFRASTRUCTURE ONLY: imported by tests/ and bench.py's
baseline legs, never by encdiff_b200/.  Pinned to tests/golden/enc4_trunk.npz, which oracle/make_enc4_golden.py produced by
running the UNMODIFIED reference ``Encoder4`` on the same seeded parameters (tests/test_enc4_oracle.py)."""
from __future__ import annotations

from typing import Dict, List, Tuple

import torch
import torch.nn.functional as F

BN_NAMES = ("1", "4", "7", "9", "11.convs.2", "12", "14.convs.2")       # BatchNorm2d modules in module order


def param_shapes(d: int = 128, latent_unit: int = 20, num_channels: int = 3) -> List[Tuple[str, Tuple[int, ...]]]:
    """(name, shape) of ``encoder.*`` in the reference's registration order (= named_parameters order)."""
    out = []

    def conv(n, co, ci, k):
        out.append((f"encoder.{n}.weight", (co, ci, k, k)))
        out.append((f"encoder.{n}.bias", (co,)))

    def bn(n):
        out.append((f"encoder.{n}.weight", (d,)))
        out.append((f"encoder.{n}.bias", (d,)))

    conv("0", d, num_channels, 4); bn("1")
    conv("3", d, d, 4); bn("4")
    conv("6", d, d, 4); bn("7")
    conv("8", d, d, 4); bn("9")
    for blk in ("11", "14"):
        conv(blk + ".convs.1", d, d, 3); bn(blk + ".convs.2"); conv(blk + ".convs.4", d, d, 1)
        if blk == "11":
            bn("12")
    out.append(("encoder.16.weight", (latent_unit, 16 * d)))
    out.append(("encoder.16.bias", (latent_unit,)))
    return out


def make_params(d=128, latent_unit=20, num_channels=3, seed=0) -> Dict[str, torch.Tensor]:
    """Seeded parameters at torch-default-like scales (fan-in uniform for weights, BN gamma around 1)."""
    g = torch.Generator().manual_seed(seed)
    p = {}
    for name, shape in param_shapes(d, latent_unit, num_channels):
        is_bn = name[len("encoder."):-len(".weight")] in BN_NAMES or name[len("encoder."):-len(".bias")] in BN_NAMES
        if is_bn:
            p[name] = (1.0 + 0.2 * torch.randn(shape, generator=g)) if name.endswith(".weight") else 0.1 * torch.randn(shape, generator=g)
        elif name.endswith(".weight"):
            fan_in = 1
            for s in shape[1:]:
                fan_in *= s
            p[name] = (torch.rand(shape, generator=g) * 2 - 1) / fan_in ** 0.5
        else:
            p[name] = (torch.rand(shape, generator=g) * 2 - 1) * 0.05
    return p


def with_relu_margin(p: Dict[str, torch.Tensor], seed=0) -> Dict[str, torch.Tensor]:
    """Copy of ``p`` whose BatchNorm affine parameters keep every ReLU input far from zero (gamma around 0.3, beta = +-2 by
    channel): each ReLU is the identity or zero for a whole channel, whatever the rounding, so parameter gradients can be
    compared entry by entry across precisions (see trunk_forward on mask flips)."""
    g = torch.Generator().manual_seed(seed + 991)
    q = {k: v.clone() for k, v in p.items()}
    for n in BN_NAMES:
        w = q[f"encoder.{n}.weight"]
        q[f"encoder.{n}.weight"] = 0.3 * (1.0 + 0.2 * torch.randn(w.shape, generator=g))
        q[f"encoder.{n}.bias"] = 2.0 * (torch.randint(0, 2, w.shape, generator=g).float() * 2 - 1)
    return q


def make_stats(d=128, seed=0) -> Dict[str, torch.Tensor]:
    """Seeded running statistics (so that eval mode is not the trivial mean 0 / var 1 case)."""
    g = torch.Generator().manual_seed(seed + 77)
    s = {}
    for n in BN_NAMES:
        s[f"encoder.{n}.running_mean"] = 0.1 * torch.randn(d, generator=g)
        s[f"encoder.{n}.running_var"] = 0.5 + torch.rand(d, generator=g)
    return s


def _ste_round(t, dtype):
    """Round to ``dtype`` and back with a straight-through gradient (the rounding has derivative 1)."""
    return t if dtype is None else t + (t.to(dtype).to(t.dtype) - t).detach()


def trunk_forward(p, stats, x, training: bool, momentum: float = 0.1, eps: float = 1e-5, round_to=None):
    """codes [B, latent_unit] = Encoder4.encoder(x).  ``stats`` is updated in place in training mode, as nn.BatchNorm2d does.

    ``round_to`` (None = the reference's fp32 arithmetic): emulate a 16-bit implementation by rounding the image, the
    convolution / Linear weights and every stored activation (convolution, BatchNorm+ReLU and block outputs) to that dtype,
    all arithmetic staying fp32.  ReLU makes the parameter gradients discontinuous in the activations: an element whose
    pre-activation is within rounding distance of zero flips its mask, and with it a whole upstream gradient term, so the
    gradients of ANY reduced-precision run (the reference's own TF32 default included) differ from the fp32 ones by
    percents in a few entries.  Checking an implementation's gradients entry by entry therefore needs the oracle evaluated
    at the same rounding points; the unrounded oracle still bounds the codes and the gradients in norm."""
    r = lambda t: _ste_round(t, round_to)        # noqa: E731

    def conv(n, h, stride, pad):
        return r(F.conv2d(h, r(p[f"encoder.{n}.weight"]), p[f"encoder.{n}.bias"], stride=stride, padding=pad))

    def bn(n, h):
        return F.batch_norm(h, stats[f"encoder.{n}.running_mean"], stats[f"encoder.{n}.running_var"], p[f"encoder.{n}.weight"],
                            p[f"encoder.{n}.bias"], training, momentum, eps)

    def resblock(n, h):                                  # EncResBlock.forward: x + convs(x), bn=True
        t = conv(n + ".convs.1", F.relu(h), 1, 1)
        t = r(F.relu(bn(n + ".convs.2", t)))
        w = r(p[f"encoder.{n}.convs.4.weight"])
        return r(h + F.conv2d(t, w, p[f"encoder.{n}.convs.4.bias"]))

    h = r(F.relu(bn("1", conv("0", r(x), 2, 1))))
    h = r(F.relu(bn("4", conv("3", h, 2, 1))))
    h = r(bn("7", conv("6", h, 2, 1)))
    h = r(F.relu(bn("9", conv("8", h, 2, 1))))
    h = resblock("11", h)
    h = r(F.relu(bn("12", h)))
    h = resblock("14", h)
    h = h.reshape(-1, 128 * 4 * 4)                       # View((-1, 128 * 4 * 4))
    return F.linear(h, r(p["encoder.16.weight"]), p["encoder.16.bias"])


def seeded_case(B: int, seed: int, latent_unit: int = 20):
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(B, 3, 64, 64, generator=g) * 2 - 1
    dcodes = torch.randn(B, latent_unit, generator=g) / B
    return x, dcodes


def forward_backward(p, stats, x, dcodes, training: bool, round_to=None):
    """codes and {name: grad} for the upstream gradient ``dcodes``; stats updated in place when training."""
    q = {k: v.clone().requires_grad_(True) for k, v in p.items()}
    codes = trunk_forward(q, stats, x, training, round_to=round_to)
    codes.backward(dcodes)
    return codes.detach(), {k: v.grad for k, v in q.items()}
