"""The allocation-free entry point encdiff_op_cross_attention_bwd_noalloc (include/encdiff_b200.h; no counterpart in the
reference): its scratch is a static device array inside the library.
Checked through the C ABI against the existing caller-scratch entry point on the same inputs, on two calls in a row (the
array keeps the previous call's contents), and for the size limit.  Collected last on purpose."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _tol():
    from encdiff_b200 import _lib
    return 2 * (1e-2 if _lib.act_dtype() == torch.bfloat16 else 2e-3)


def _rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()


def _inputs(dev, dt, B, N, heads, D, T, Dc, seed):
    C = heads * D
    g = torch.Generator().manual_seed(seed)
    q = torch.randn(B, N, C, generator=g).to(dt).to(dev).contiguous()
    dout = (torch.randn(B, N, C, generator=g) * 2.0 ** -6).to(dt).to(dev).contiguous()
    ctx = torch.randn(B, T, Dc, generator=g).to(dev)
    wk = (torch.randn(C, Dc, generator=g) * 0.3).to(dev)
    wv = (torch.randn(C, Dc, generator=g) * 0.3).to(dev)
    return q, dout, ctx, wk, wv


def _run(lib, _lib, which, q, dout, ctx, wk, wv, kv, B, N, heads, D, T, Dc):
    dq = torch.zeros_like(q)
    dwk, dwv, dctx = torch.zeros_like(wk), torch.zeros_like(wv), torch.zeros_like(ctx)
    kv_ptr = 0 if kv is None else kv.data_ptr()
    if which == "noalloc":
        st = lib.encdiff_op_cross_attention_bwd_noalloc(q.data_ptr(), ctx.data_ptr(), wk.data_ptr(), wv.data_ptr(), kv_ptr,
                                                        dout.data_ptr(), dq.data_ptr(), dwk.data_ptr(), dwv.data_ptr(),
                                                        dctx.data_ptr(), B, N, heads, D, T, Dc, _lib.current_stream())
    else:
        scratch = torch.empty(lib.encdiff_op_cross_attention_bwd_scratch_bytes(B, heads, D, T) // 4, dtype=torch.float32,
                              device=q.device)
        st = lib.encdiff_op_cross_attention_bwd(q.data_ptr(), ctx.data_ptr(), wk.data_ptr(), wv.data_ptr(), kv_ptr,
                                                dout.data_ptr(), dq.data_ptr(), dwk.data_ptr(), dwv.data_ptr(),
                                                dctx.data_ptr(), scratch.data_ptr(), B, N, heads, D, T, Dc,
                                                _lib.current_stream())
    _lib.check(st, which)
    torch.cuda.synchronize()
    return dq, dwk, dwv, dctx


@pytest.mark.parametrize("use_kv", [False, True])      # False: CUDA-core path (accumulates into the scratch); True: mma path
def test_noalloc_matches_caller_scratch_entry_point_twice(cuda_device, use_kv):
    from encdiff_b200 import _lib
    lib = _lib.load()
    dt = _lib.act_dtype()
    heads, T, Dc = 8, 20, 16
    # two calls in a row with different shapes and data: the second must not see the first one's leftovers
    for call, (B, N, D, seed) in enumerate([(6, 64, 16, 1), (3, 256, 8, 2)]):
        C = heads * D
        q, dout, ctx, wk, wv = _inputs(cuda_device, dt, B, N, heads, D, T, Dc, seed)
        kv = None
        if use_kv:
            kv = torch.empty(B, T, 2 * C, dtype=torch.float32, device=cuda_device)
            tab = torch.empty(64, dtype=torch.uint8, device=cuda_device)
            _lib.check(lib.encdiff_op_cross_kv_project(ctx.data_ptr(), wk.data_ptr(), wv.data_ptr(), kv.data_ptr(),
                                                       tab.data_ptr(), B, C, T, Dc, _lib.current_stream()))
        want = _run(lib, _lib, "caller", q, dout, ctx, wk, wv, kv, B, N, heads, D, T, Dc)
        got = _run(lib, _lib, "noalloc", q, dout, ctx, wk, wv, kv, B, N, heads, D, T, Dc)
        for name, a, b in zip(("dq", "dwk", "dwv", "dctx"), got, want):
            err = _rel(a.float(), b.float())
            print(f"call {call} use_kv={use_kv} {name}: rel diff {err:.2e}")
            assert torch.isfinite(a.float()).all()
            assert err <= _tol(), (call, name, err)


def test_noalloc_rejects_sizes_beyond_the_static_scratch(cuda_device):
    from encdiff_b200 import _lib
    lib = _lib.load()
    dt = _lib.act_dtype()
    heads, D, T, Dc, N = 8, 32, 20, 16, 16
    B = (16 << 20) // (T * 2 * heads * D * 4) + 1            # one sample past ENCDIFF_NOALLOC_SCRATCH_BYTES
    q, dout, ctx, wk, wv = _inputs(cuda_device, dt, B, N, heads, D, T, Dc, 3)
    dq = torch.zeros_like(q)
    dwk, dwv, dctx = torch.zeros_like(wk), torch.zeros_like(wv), torch.zeros_like(ctx)
    st = lib.encdiff_op_cross_attention_bwd_noalloc(q.data_ptr(), ctx.data_ptr(), wk.data_ptr(), wv.data_ptr(), 0,
                                                    dout.data_ptr(), dq.data_ptr(), dwk.data_ptr(), dwv.data_ptr(),
                                                    dctx.data_ptr(), B, N, heads, D, T, Dc, _lib.current_stream())
    assert st != 0
    msg = lib.encdiff_last_error().decode()
    assert "ENCDIFF_NOALLOC_SCRATCH_BYTES" in msg and str(16 << 20) in msg
    torch.cuda.synchronize()
    assert dq.abs().max().item() == 0.0                      # nothing was launched
