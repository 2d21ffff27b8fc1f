"""First-stage VQ-f4 encode AND decode throughput at the training / sampling batch: encdiff_b200.VQEncoder (eager launches
and a replayed CUDA graph) next to the reference's own Encoder + quant_conv (oracle/_ref through oracle/make_vq_golden's
builder) in PyTorch eager mode on the same GPU (TF32, as the reference ships) and on the host cores (bounded sample).
Prints one JSON line.  usage: python tools/vq_bench.py [B]"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import bench  # noqa: E402
from oracle import vq_oracle as VO  # noqa: E402

GFLOP_IMG = 1.221          # VQ-f4 encode, hook-counted on the reference


def timed(fn, reps):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


GFLOP_IMG_DEC = 2.492      # VQ decode


def decode_leg(B, dev, pk):
    """VQDecoder.decode (quantise + concat + post_quant_conv + Decoder) at batch B, replayed graph, next to the reference
    Decoder + post_quant_conv in eager mode on the same GPU (TF32) on an already-quantised latent."""
    from test_vq_gpu import make_vqdec
    cfg = VO.SHAPES3D_VQ
    m, params = make_vqdec(cfg, dev, 2048, 20)
    g = torch.Generator().manual_seed(29)
    h = torch.randn(B, 3, 16, 16, generator=g).to(dev)
    s = torch.randn(B, 20, generator=g).to(dev)
    for _ in range(2):
        m.decode(h, disentangled_repr=s)
    ms_eager = timed(lambda: m.decode(h, disentangled_repr=s), 10)
    graph = torch.cuda.CUDAGraph()
    torch.cuda.synchronize()
    with torch.cuda.graph(graph):
        img = m.decode(h, disentangled_repr=s)
    for _ in range(3):
        graph.replay()
    ms_graph = timed(graph.replay, 20)
    out = {"metric": "vq_f4_decode_img_per_s", "batch": B, "value": B / (ms_graph / 1e3), "ms_per_batch_graph": ms_graph,
           "ms_per_batch_eager": ms_eager, "gflop_per_img": GFLOP_IMG_DEC,
           "tensor_frac": B / (ms_graph / 1e3) * GFLOP_IMG_DEC / 1e3 / pk["tf_sustained"],
           "workspace_gb": m._lib.encdiff_vqdec_workspace_bytes(m._h, B) / 1e9}
    try:
        from oracle.make_vq_golden import reference_decoder
        dec, pqc = reference_decoder(cfg, params, 20)
        dec, pqc = dec.to(dev), pqc.to(dev)
        torch.backends.cuda.matmul.allow_tf32 = True
        torch.backends.cudnn.allow_tf32 = True
        with torch.no_grad():
            q = torch.cat([h, s[:, :, None, None].expand(-1, -1, 16, 16)], dim=1)
            for _ in range(2):
                dec(pqc(q))
            ms_ref = timed(lambda: dec(pqc(q)), 5)
            out["ref_eager_gpu_tf32"] = {"img_per_s": B / (ms_ref / 1e3), "ms_per_batch": ms_ref,
                                         "note": "Decoder + post_quant_conv only (no quantiser)"}
            ref = dec(pqc(q))
            out["rel_err_vs_ref_gpu_tf32_no_quant"] = ((m.decode(h, force_not_quantize=True, disentangled_repr=s) - ref).abs().max()
                                                       / ref.abs().max()).item()
    except Exception as e:
        out["ref_error"] = repr(e)[:200]
    return out


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    dev = torch.device("cuda:0")
    from test_vq_gpu import make_vq
    cfg = VO.SHAPES3D_VQ
    m, params = make_vq(cfg, dev)
    g = torch.Generator().manual_seed(23)
    xs = [(torch.rand(B, 3, 64, 64, generator=g) * 2 - 1).to(dev) for _ in range(3)]
    for x in xs:
        m.encode(x)
    ms_eager = timed(lambda: m.encode(xs[0]), 10)
    static = xs[0].clone()
    graph = torch.cuda.CUDAGraph()
    torch.cuda.synchronize()
    with torch.cuda.graph(graph):
        z = m.encode(static)
    it = [0]

    def step():
        static.copy_(xs[it[0] % 3])
        it[0] += 1
        graph.replay()
    for _ in range(3):
        step()
    ms_graph = timed(step, 20)
    # per-family device time (eager, one event per launch)
    lib = m._lib
    lib.encdiff_unet_profile(m._h, 1)
    import ctypes as C
    buf = C.create_string_buffer(1 << 16)
    lib.encdiff_unet_profile_read(m._h, buf, len(buf))
    for _ in range(3):
        m.encode(xs[0])
    torch.cuda.synchronize()
    lib.encdiff_unet_profile_read(m._h, buf, len(buf))
    lib.encdiff_unet_profile(m._h, 0)
    fam = {}
    for line in buf.value.decode().splitlines():
        name, k, ms, fl, by = line.split(",")
        fam[name] = {"n": int(k) // 3, "ms": round(float(ms) / 3, 4), "gflop": round(float(fl) / 3 / 1e9, 2)}
    pk = bench.peaks()
    out = {"metric": "vq_f4_encode_img_per_s", "batch": B, "value": B / (ms_graph / 1e3), "ms_per_batch_graph": ms_graph,
           "ms_per_batch_eager": ms_eager, "gflop_per_img": GFLOP_IMG,
           "tensor_frac": B / (ms_graph / 1e3) * GFLOP_IMG / 1e3 / pk["tf_sustained"],
           "workspace_gb": lib.encdiff_vqenc_workspace_bytes(m._h, B) / 1e9, "families_eager": fam}
    # the reference's own modules: same GPU (eager, TF32 as shipped) and the host cores
    try:
        from oracle.make_vq_golden import reference_encoder
        enc, qc = reference_encoder(cfg, params)
        torch.backends.cuda.matmul.allow_tf32 = True
        torch.backends.cudnn.allow_tf32 = True
        enc_d, qc_d = enc.to(dev), qc.to(dev)
        with torch.no_grad():
            for _ in range(2):
                qc_d(enc_d(xs[0]))
            ms_ref = timed(lambda: qc_d(enc_d(xs[0])), 5)
            zr = qc_d(enc_d(xs[0]))
            out["ref_eager_gpu_tf32"] = {"img_per_s": B / (ms_ref / 1e3), "ms_per_batch": ms_ref}
            out["rel_err_vs_ref_gpu_tf32"] = ((m.encode(xs[0]) - zr).abs().max() / zr.abs().max()).item()
            enc_c, qc_c = enc.to("cpu"), qc.to("cpu")
            torch.set_num_threads(os.cpu_count() or 1)
            xc = xs[0][:32].cpu()
            qc_c(enc_c(xc))
            t0 = time.perf_counter()
            for _ in range(3):
                qc_c(enc_c(xc))
            sec = (time.perf_counter() - t0) / 3
            out["ref_cpu"] = {"img_per_s": 32 / sec, "cores": torch.get_num_threads(), "sample": "3 passes of a batch-32 sample, fp32"}
    except Exception as e:      # oracle/_ref missing on this box
        out["ref_error"] = repr(e)[:200]
    out["decode"] = decode_leg(B, dev, pk)
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
