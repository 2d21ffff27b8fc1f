"""Parity of the tcgen05 GEMM kernels against the reference call sites' math (nn.Conv2d / nn.Linear forward,
input gradient and weight gradient: openaimodel_enc.py:201-241, attention.py:37-64,152-193), computed by PyTorch in
fp32 on the same 16-bit-rounded inputs.  Tolerance: the only differences are fp32 accumulation order and the final
16-bit rounding of the output -> rel 2e-3 of the output scale in the fp16 build (11 significand bits), 1e-2 in the bf16
build (8 bits)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def ADT():
    """The library's 16-bit storage type (fp16 by default; bf16 in the -DEDB_ACT_BF16 build)."""
    from encdiff_b200 import _lib
    return _lib.act_dtype()


def _nhwc(B, H, W, C, dev, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return torch.randn(B, H, W, C, generator=g).to(dev).to(ADT())


def _close(got, want, tol=None):
    tol = tol if tol is not None else (1e-2 if ADT() == torch.bfloat16 else 2e-3)
    got, want = got.float(), want.float()
    scale = want.abs().max().item() + 1e-6
    err = (got - want).abs().max().item()
    assert err <= tol * scale, f"max err {err:.4g} vs scale {scale:.4g}"


@pytest.mark.parametrize("B,H,W,K,N", [(4, 16, 16, 64, 128), (3, 8, 8, 128, 64), (5, 4, 4, 256, 256),
                                       (7, 2, 2, 192, 384), (130, 1, 1, 256, 512), (2, 16, 16, 64, 3)])
def test_linear(cuda_device, B, H, W, K, N):
    from encdiff_b200 import ops
    dev = cuda_device
    a = _nhwc(B, H, W, K, dev, 1)
    g = torch.Generator().manual_seed(2)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(dev)
    bias = torch.randn(N, generator=g).to(dev)
    out = ops.sgemm([a], [1], ops.pack_linear(w), N, H=H, W=W, B=B, bias=bias)
    want = a.reshape(-1, K).float() @ w.to(ADT()).float().t() + bias
    _close(out, want)


@pytest.mark.parametrize("B,H,W,Ci,Co", [(2, 16, 16, 64, 64), (3, 8, 8, 128, 256), (9, 4, 4, 256, 128),
                                         (33, 2, 2, 512, 256), (2, 16, 16, 64, 3)])
def test_conv3x3_forward_and_dgrad(cuda_device, B, H, W, Ci, Co):
    from encdiff_b200 import ops
    dev = cuda_device
    x = _nhwc(B, H, W, Ci, dev, 3)
    g = torch.Generator().manual_seed(4)
    w = (torch.randn(Co, Ci, 3, 3, generator=g) / (9 * Ci) ** 0.5).to(dev)
    bias = torch.randn(Co, generator=g).to(dev)
    wq = w.to(ADT()).float()
    out = ops.sgemm([x], [9], ops.pack_conv3x3(w), Co, H=H, W=W, B=B, bias=bias, out_fp32=True)
    want = F.conv2d(x.float().permute(0, 3, 1, 2), wq, bias, padding=1).permute(0, 2, 3, 1).reshape(-1, Co)
    _close(out, want)
    if Co % 8 == 0:
        # input gradient: dX = conv_transpose(dY, W) through mirrored taps + transposed pack
        dy = _nhwc(B, H, W, Co, dev, 5)
        dx = ops.sgemm([dy], [9], ops.pack_conv3x3_t(w), Ci, H=H, W=W, B=B, tap_sign=-1)
        want_dx = F.conv_transpose2d(dy.float().permute(0, 3, 1, 2), wq, padding=1).permute(0, 2, 3, 1)
        _close(dx, want_dx.reshape(-1, Ci))


def test_fused_sources_residual_silu_split(cuda_device):
    """ResBlock tail (openaimodel_enc.py:267-275): conv3x3(a2) + 1x1 skip conv over cat([h, skip]) in one launch,
    plus residual / SiLU epilogues and the two-destination column split used by the concat input gradient."""
    from encdiff_b200 import ops
    dev = cuda_device
    B, H, W = 3, 8, 8
    a2, h, sk = _nhwc(B, H, W, 128, dev, 6), _nhwc(B, H, W, 128, dev, 7), _nhwc(B, H, W, 64, dev, 8)
    g = torch.Generator().manual_seed(9)
    w2 = (torch.randn(128, 128, 3, 3, generator=g) / 34.0).to(dev)
    ws = (torch.randn(128, 192, generator=g) / 14.0).to(dev)
    bias = torch.randn(128, generator=g).to(dev)
    res = _nhwc(B, H, W, 128, dev, 10).reshape(-1, 128)
    wp = torch.cat([ops.pack_conv3x3(w2), ops.pack_linear(ws)], dim=1).contiguous()
    out = ops.sgemm([a2, h, sk], [9, 1, 1], wp, 128, H=H, W=W, B=B, bias=bias, res=res, act=ops.ACT_SILU)
    xcat = torch.cat([h, sk], dim=-1).float().reshape(-1, 192)
    conv = F.conv2d(a2.float().permute(0, 3, 1, 2), w2.to(ADT()).float(), None, padding=1)
    pre = conv.permute(0, 2, 3, 1).reshape(-1, 128) + xcat @ ws.to(ADT()).float().t() + bias
    _close(out, F.silu(pre) + res.float())
    # column split: dX of the 1x1 skip conv lands in two tensors (grad of h, grad of skip)
    dy = _nhwc(B, H, W, 128, dev, 11)
    d0 = torch.empty(B * H * W, 128, device=dev, dtype=ADT())
    d1 = torch.empty(B * H * W, 64, device=dev, dtype=ADT())
    ops.sgemm([dy], [1], ops.pack_linear(ws.t().contiguous()), 192, H=H, W=W, B=B, out=d0, out1=d1, split_col=128,
              bn=64)
    want = dy.float().reshape(-1, 128) @ ws.to(ADT()).float()
    _close(d0, want[:, :128])
    _close(d1, want[:, 128:])


@pytest.mark.parametrize("B,H,W,Ci,Co,splits", [(4, 16, 16, 64, 64, 3), (6, 8, 8, 128, 256, 2), (40, 2, 2, 512, 256, 1),
                                                 (16, 4, 4, 192, 128, 2)])
def test_wgrad_conv3x3(cuda_device, B, H, W, Ci, Co, splits):
    from encdiff_b200 import ops
    dev = cuda_device
    x = _nhwc(B, H, W, Ci, dev, 12)
    dy = _nhwc(B, H, W, Co, dev, 13)
    dw = torch.zeros(Co, Ci, 3, 3, device=dev)
    ops.wgrad(dy, x, dw, H=H, W=W, B=B, taps=9, stride_p=Ci * 9, stride_q=9, stride_tap=1, splits=splits)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(False)
    want = torch.nn.grad.conv2d_weight(xr, (Co, Ci, 3, 3), dy.float().permute(0, 3, 1, 2), padding=1)
    _close(dw, want, tol=2e-3)


@pytest.mark.parametrize("B,H,W,K,N,splits", [(4, 16, 16, 64, 192, 4), (64, 1, 1, 256, 512, 1), (9, 4, 4, 256, 64, 2)])
def test_wgrad_linear(cuda_device, B, H, W, K, N, splits):
    from encdiff_b200 import ops
    dev = cuda_device
    x = _nhwc(B, H, W, K, dev, 14)
    dy = _nhwc(B, H, W, N, dev, 15)
    dw = torch.ones(N, K, device=dev)  # accumulates on top of existing content
    ops.wgrad(dy, x, dw, H=H, W=W, B=B, taps=1, stride_p=K, stride_q=1, stride_tap=0, splits=splits)
    want = dy.float().reshape(-1, N).t() @ x.float().reshape(-1, K) + 1.0
    _close(dw, want, tol=2e-3)

