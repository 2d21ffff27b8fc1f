"""CPU model of where bf16 rounding enters the CUDA plan: restates the plan's stores on top of the fp32 oracle and
rounds selected groups of tensors to bf16, to rank the error sources (diagnostic only).
groups: W weights, S residual stream (block outputs), A GEMM A-operands, I other intermediates, X input/temb"""
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import unet_oracle as UO  # noqa: E402
from oracle.make_golden import CONFIGS, seeded_inputs  # noqa: E402

ON = set()
FMT = {"W": torch.bfloat16, "S": torch.bfloat16, "A": torch.bfloat16, "I": torch.bfloat16, "X": torch.bfloat16}


def r(x, g):
    return x.to(FMT[g]).float() if g in ON else x


def resblock(p, l, x, emb):
    pre = l.prefix
    h = F.silu(UO.group_norm32(x, p[f'{pre}.in_layers.0.weight'], p[f'{pre}.in_layers.0.bias'], 1e-5))
    if l.down:
        h, x = F.avg_pool2d(h, 2), F.avg_pool2d(x, 2)
    elif l.up:
        h, x = F.interpolate(h, scale_factor=2, mode='nearest'), F.interpolate(x, scale_factor=2, mode='nearest')
    if l.up or l.down:
        x = r(x, "I")
    h = r(h, "A")
    h = r(F.conv2d(h, r(p[f'{pre}.in_layers.2.weight'], "W"), p[f'{pre}.in_layers.2.bias'], padding=1), "I")
    e = F.linear(r(F.silu(emb), "A"), r(p[f'{pre}.emb_layers.1.weight'], "W"), p[f'{pre}.emb_layers.1.bias'])
    scale, shift = e[:, :l.cout, None, None], e[:, l.cout:, None, None]
    h = UO.group_norm32(h, p[f'{pre}.out_layers.0.weight'], p[f'{pre}.out_layers.0.bias'], 1e-5) * (1 + scale) + shift
    h = F.conv2d(r(F.silu(h), "A"), r(p[f'{pre}.out_layers.3.weight'], "W"), p[f'{pre}.out_layers.3.bias'], padding=1)
    if l.cin != l.cout:
        x = F.conv2d(x, r(p[f'{pre}.skip_connection.weight'], "W"), p[f'{pre}.skip_connection.bias'])
    return r(x + h, "S")


def attention(p, pre, x, context, heads):
    q = r(F.linear(x, r(p[f'{pre}.to_q.weight'], "W")), "I")
    if context is None:
        k = r(F.linear(x, r(p[f'{pre}.to_k.weight'], "W")), "I")
        v = r(F.linear(x, r(p[f'{pre}.to_v.weight'], "W")), "I")
    else:
        k = F.linear(context, p[f'{pre}.to_k.weight'])
        v = F.linear(context, p[f'{pre}.to_v.weight'])
    b, n, c = q.shape
    d = c // heads
    sp = lambda t: t.reshape(b, t.shape[1], heads, d).permute(0, 2, 1, 3)
    q, k, v = sp(q), sp(k), sp(v)
    sim = torch.matmul(q, k.transpose(-1, -2)) * (d ** -0.5)
    out = torch.matmul(sim.softmax(dim=-1), v).permute(0, 2, 1, 3).reshape(b, n, c)
    return F.linear(r(out, "A"), r(p[f'{pre}.to_out.0.weight'], "W"), p[f'{pre}.to_out.0.bias'])


def st(p, l, x, ctx):
    pre = l.prefix
    b, c, hh, ww = x.shape
    h = r(F.group_norm(x, 32, p[f'{pre}.norm.weight'], p[f'{pre}.norm.bias'], 1e-6), "A")
    h = F.conv2d(h, r(p[f'{pre}.proj_in.weight'], "W"), p[f'{pre}.proj_in.bias'])
    h = r(h.permute(0, 2, 3, 1).reshape(b, hh * ww, c), "T")
    tp = f'{pre}.transformer_blocks.0'
    ln = lambda t, i: r(F.layer_norm(t, t.shape[-1:], p[f'{tp}.norm{i}.weight'], p[f'{tp}.norm{i}.bias'], 1e-5), "A")
    h = r(attention(p, f'{tp}.attn1', ln(h, 1), None, l.heads) + h, "T")
    h = r(attention(p, f'{tp}.attn2', ln(h, 2), ctx, l.heads) + h, "T")
    y = r(F.linear(ln(h, 3), r(p[f'{tp}.ff.net.0.proj.weight'], "W"), p[f'{tp}.ff.net.0.proj.bias']), "I")
    val, gate = y.chunk(2, dim=-1)
    h = r(F.linear(r(val * F.gelu(gate), "A"), r(p[f'{tp}.ff.net.2.weight'], "W"), p[f'{tp}.ff.net.2.bias']) + h, "T")
    h = r(h, "A")
    h = h.reshape(b, hh, ww, c).permute(0, 3, 1, 2)
    h = F.conv2d(h, r(p[f'{pre}.proj_out.weight'], "W"), p[f'{pre}.proj_out.bias'])
    return r(h + x, "S")


def forward(p, cfg, x, t, context):
    a = UO.build_arch(cfg)
    b = x.shape[0]
    emb = r(UO.timestep_embedding(t, cfg.model_channels), "X")
    emb = F.linear(emb, r(p['time_embed.0.weight'], "W"), p['time_embed.0.bias'])
    emb = F.linear(r(F.silu(emb), "A"), r(p['time_embed.2.weight'], "W"), p['time_embed.2.bias'])
    ctx = context[0].reshape(b, -1, cfg.context_dim)

    def run(layers, h):
        for l in layers:
            if l.kind == 'conv_in':
                h = r(F.conv2d(r(h, "X"), r(p[f'{l.prefix}.weight'], "W"), p[f'{l.prefix}.bias'], padding=1), "S")
            elif l.kind == 'res':
                h = resblock(p, l, h, emb)
            else:
                h = st(p, l, h, ctx)
        return h
    hs, h = [], x
    for blk in a.input_blocks:
        h = run(blk, h)
        hs.append(h)
    h = run(a.middle, h)
    for blk in a.output_blocks:
        h = run(blk, torch.cat([h, hs.pop()], dim=1))
    h = r(F.silu(UO.group_norm32(h, p['out.0.weight'], p['out.0.bias'], 1e-5)), "A")
    return F.conv2d(h, r(p['out.2.weight'], "W"), p['out.2.bias'], padding=1)


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "shapes3d"
    cfg = CONFIGS[name]
    params = UO.make_params(cfg, seed=1)
    x, t, ctx, _ = seeded_inputs(cfg, 4, seed=101)
    FMT["T"] = torch.bfloat16
    with torch.no_grad():
        ref = UO.unet_forward(params, cfg, x, t, [ctx])
        for groups in ["", "W", "S", "T", "A", "I", "X", "WA", "WAX", "WAXI", "WAXIT", "WAXITS"]:
            ON.clear()
            ON.update(groups)
            out = forward(params, cfg, x, t, [ctx])
            d = (out - ref).abs()
            print(f"{groups or '-':8s} maxerr/max {d.max() / ref.abs().max():.3e}  rmserr/rms "
                  f"{d.pow(2).mean().sqrt() / ref.pow(2).mean().sqrt():.3e}")


if __name__ == "__main__":
    main()
