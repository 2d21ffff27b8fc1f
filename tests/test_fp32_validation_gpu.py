"""The north star's fp32 validation bar: eps-prediction max relative error <= 1e-4 against the reference's fp32 result.
``UNetModel.forward_fp32`` is the library's fp32 validation forward (csrc/fp32_ref.cu through the C ABI:
``encdiff_unet_forward_fp32``): fp32 tensors, CUDA-core FMA contractions, the fp32 master parameters in the reference's
layouts.  Checked here against (a) golden vectors produced by the UNMODIFIED reference UNetModel (tests/golden/, made by
oracle/make_golden.py), (b) the CPU oracle on seeded inputs for all three configurations, and then used (c) as the
GPU-side checker of the 16-bit product path at the benched shape B = 256, where the CPU oracle needs minutes."""
import os

import numpy as np
import pytest
import torch

from oracle import unet_oracle as UO
from oracle.make_golden import CONFIGS, seeded_inputs
from test_unet_gpu import GOLDEN, TOL, make_model, rel_err

pytestmark = pytest.mark.gpu
FP32_BAR = 1e-4      # north star: "<= 1e-4 in an fp32 validation build"


@pytest.mark.parametrize("fixture,cfgname", [("unet_tiny_b3.npz", "tiny"), ("unet_shapes3d_b2.npz", "shapes3d")])
def test_fp32_forward_vs_reference_golden(cuda_device, fixture, cfgname):
    g = np.load(os.path.join(GOLDEN, fixture))
    m, _ = make_model(CONFIGS[cfgname], cuda_device)
    x, t, ctx = (torch.tensor(g[k]).to(cuda_device) for k in ("x", "t", "ctx"))
    eps = m.forward_fp32(x, t, [ctx])
    err = rel_err(eps, g["eps"])
    print(f"{cfgname}: fp32 validation forward vs reference golden: eps rel err {err:.3e}")
    assert eps.shape == x.shape and eps.dtype == torch.float32
    assert err <= FP32_BAR


@pytest.mark.parametrize("cfgname,B", [("tiny", 5), ("shapes3d", 3), ("celeba64", 2)])
def test_fp32_forward_vs_oracle(cuda_device, cfgname, B):
    cfg = CONFIGS[cfgname]
    m, params = make_model(cfg, cuda_device)
    x, t, ctx, _ = seeded_inputs(cfg, B, seed=313)
    with torch.no_grad():
        want = UO.unet_forward(params, cfg, x, t, [ctx])
    eps = m.forward_fp32(x.to(cuda_device), t.to(cuda_device), [ctx.to(cuda_device)])
    err = rel_err(eps, want)
    print(f"{cfgname} B={B}: fp32 validation forward vs oracle: eps rel err {err:.3e}")
    assert err <= FP32_BAR


def test_product_path_vs_fp32_validation_at_benched_batch(cuda_device):
    """The 16-bit tensor-core forward against the fp32 validation forward on the GPU at the benched shape (Shapes3D,
    B = 256: persistent multi-tile GEMM variants, full-size attention grids) -- same 1e-2 bar as against the oracle."""
    cfg = CONFIGS["shapes3d"]
    m, _ = make_model(cfg, cuda_device)
    x, t, ctx, _ = seeded_inputs(cfg, 256, seed=77)
    x, t, ctx = x.to(cuda_device), t.to(cuda_device), ctx.to(cuda_device)
    with torch.no_grad():
        eps16 = m(x, t, context=[ctx])
    eps32 = m.forward_fp32(x, t, [ctx])
    err = rel_err(eps16, eps32)
    per_sample = ((eps16 - eps32).flatten(1).abs().amax(1) / eps32.flatten(1).abs().amax(1)).max().item()
    print(f"shapes3d B=256: 16-bit path vs fp32 validation path: eps rel err {err:.3e} (worst sample {per_sample:.3e})")
    assert torch.isfinite(eps32).all()
    assert err <= TOL.eps and per_sample <= 2 * TOL.eps
