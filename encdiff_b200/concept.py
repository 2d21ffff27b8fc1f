"""EncDiff's concept encoder: the per-unit MLPs as one CUDA launch (``ConceptMLP``) and the
drop-in ``Encoder4`` whose convolutional trunk runs forward and backward through ``encdiff_enc4_*``.

Reference: ``Encoder4`` (ldm/modules/diffusionmodules/openaimodel_enc.py:991-1041) ends in ``latent_unit`` independent
3-layer MLPs ``Linear(1, 64) - ELU - Linear(64, 128) - ELU - Linear(128, context_dim)`` held in ``self.net`` and run in a
Python loop, both in ``forward`` (:1024-1031, after the convolutional trunk) and in ``warp`` (:1037-1041, used by the
swap generation).  ``ConceptMLP`` owns the same parameters stacked over the units (so one kernel can walk them), reads
and writes the reference's ``state_dict`` keys (``{i}.0.weight`` ... ``{i}.4.bias`` under the module's prefix, i.e.
``net.{i}...`` when the module is registered as ``net``), and is differentiable w.r.t. codes and parameters.

    enc.net_fast = ConceptMLP.from_reference(enc.net)       # enc: the reference's Encoder4
    tokens = enc.net_fast(enc.encoding(images))             # == enc(images);   enc.net_fast(codes) == enc.warp(codes)

``Encoder4`` (below) is the whole module: trunk (4x4 stride-2 convolutions, BatchNorm2d, EncResBlocks, Linear) + ConceptMLP.
No CPU fallback: CPU tensors raise.
"""
from __future__ import annotations

import ctypes as C

import torch
from torch import nn

from . import _lib

H1, H2 = 64, 128          # hidden widths, fixed by the reference (openaimodel_enc.py:1018-1023)


class _ConceptFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, w3, b3):
        lib = _lib.load()
        B, L = x.shape
        Dc = w3.shape[1]
        y = torch.empty(B, L * Dc, dtype=torch.float32, device=x.device)
        _lib.check(lib.encdiff_concept_mlp_forward(x.data_ptr(), w1.data_ptr(), b1.data_ptr(), w2.data_ptr(),
                                                   b2.data_ptr(), w3.data_ptr(), b3.data_ptr(), y.data_ptr(), B, L, Dc,
                                                   _lib.current_stream()), "concept_mlp_forward")
        ctx.save_for_backward(x, w1, b1, w2, b2, w3, b3)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w1, b1, w2, b2, w3, b3 = ctx.saved_tensors
        lib = _lib.load()
        B, L = x.shape
        Dc = w3.shape[1]
        dy = dy.contiguous().float()
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        grads = [torch.zeros_like(t) for t in (w1, b1, w2, b2, w3, b3)]
        _lib.check(lib.encdiff_concept_mlp_backward(x.data_ptr(), w1.data_ptr(), b1.data_ptr(), w2.data_ptr(),
                                                    b2.data_ptr(), w3.data_ptr(), b3.data_ptr(), dy.data_ptr(),
                                                    _lib.ptr(dx), *[g.data_ptr() for g in grads], B, L, Dc,
                                                    _lib.current_stream()), "concept_mlp_backward")
        return (dx, *grads)


class ConceptMLP(nn.Module):
    """``latent_unit`` MLPs 1 -> 64 -> 128 -> context_dim with stacked parameters; forward(codes [B, L] or [B, L, 1])
    returns the tokens [B, L * context_dim] in the reference's ``th.cat(out, dim=1)`` order."""

    def __init__(self, latent_unit: int, context_dim: int):
        super().__init__()
        if not 1 <= context_dim <= 32:
            raise NotImplementedError("encdiff_b200.ConceptMLP: context_dim must be in 1..32")
        self.latent_unit, self.context_dim = int(latent_unit), int(context_dim)
        L, Dc = self.latent_unit, self.context_dim
        # same construction order as the reference's ModuleList => same draws from the global generator
        w1, b1, w2, b2, w3, b3 = [], [], [], [], [], []
        for _ in range(L):
            l0, l2, l4 = nn.Linear(1, H1), nn.Linear(H1, H2), nn.Linear(H2, Dc)
            w1.append(l0.weight.detach().reshape(H1)); b1.append(l0.bias.detach())
            w2.append(l2.weight.detach()); b2.append(l2.bias.detach())
            w3.append(l4.weight.detach()); b3.append(l4.bias.detach())
        self.w1 = nn.Parameter(torch.stack(w1))            # [L, 64]
        self.b1 = nn.Parameter(torch.stack(b1))
        self.w2 = nn.Parameter(torch.stack(w2))            # [L, 128, 64]
        self.b2 = nn.Parameter(torch.stack(b2))
        self.w3 = nn.Parameter(torch.stack(w3))            # [L, Dc, 128]
        self.b3 = nn.Parameter(torch.stack(b3))
        self._register_state_dict_hook(ConceptMLP._to_reference_keys)
        self._register_load_state_dict_pre_hook(self._from_reference_keys)

    # ---- the reference's state_dict layout: '{i}.0.weight' [64, 1], '{i}.0.bias', '{i}.2.*', '{i}.4.*' -------------------
    _MAP = (("w1", "0.weight"), ("b1", "0.bias"), ("w2", "2.weight"), ("b2", "2.bias"), ("w3", "4.weight"),
            ("b3", "4.bias"))

    @staticmethod
    def _to_reference_keys(module, state, prefix, local_metadata):
        stacked = {mine: state.pop(prefix + mine) for mine, _ in ConceptMLP._MAP}
        for i in range(module.latent_unit):                 # unit-major: the key ORDER of the reference's ModuleList too
            for mine, ref in ConceptMLP._MAP:
                t = stacked[mine][i]
                state[f"{prefix}{i}.{ref}"] = t.reshape(H1, 1) if mine == "w1" else t
        return state

    def _from_reference_keys(self, state, prefix, local_metadata, strict, missing, unexpected, errors):
        for mine, ref in ConceptMLP._MAP:
            keys = [f"{prefix}{i}.{ref}" for i in range(self.latent_unit)]
            if all(k in state for k in keys):
                parts = [state.pop(k) for k in keys]
                if mine == "w1":
                    parts = [p.reshape(H1) for p in parts]
                state[prefix + mine] = torch.stack(parts)

    @classmethod
    def from_reference(cls, net: nn.ModuleList) -> "ConceptMLP":
        """Build from (and copy the weights of) the reference's ``Encoder4.net`` ModuleList."""
        m = cls(len(net), net[0][4].out_features)
        m.load_state_dict(net.state_dict(), strict=True)
        return m.to(net[0][0].weight.device)

    def forward(self, codes: torch.Tensor) -> torch.Tensor:
        if codes.dim() == 3:
            codes = codes.reshape(codes.shape[0], -1)
        if codes.dim() != 2 or codes.shape[1] != self.latent_unit:
            raise ValueError(f"ConceptMLP: expected codes [B, {self.latent_unit}], got {tuple(codes.shape)}")
        if not codes.is_cuda or not self.w1.is_cuda:
            raise RuntimeError("encdiff_b200.ConceptMLP runs on CUDA tensors only (no CPU fallback)")
        return _ConceptFn.apply(codes.contiguous().float(), self.w1, self.b1, self.w2, self.b2, self.w3, self.b3)

    def warp(self, x: torch.Tensor) -> torch.Tensor:
        """Encoder4.warp (openaimodel_enc.py:1037-1041): [B, L, 1] codes give [B, L, context_dim], [B, L] codes give
        [B, L * context_dim] (what ``x[:, i][:, None]`` and ``th.cat(out, dim=1)`` produce in the reference)."""
        y = self.forward(x)
        return y.reshape(x.shape[0], self.latent_unit, self.context_dim) if x.dim() == 3 else y


class _EncResBlock(nn.Module):
    """Parameter / structure mirror of the reference's EncResBlock (openaimodel_enc.py:969-989): x + convs(x)."""

    def __init__(self, cin, cout, bn=False):
        super().__init__()
        layers = [nn.ReLU(), nn.Conv2d(cin, cout, 3, 1, 1), nn.ReLU(), nn.Conv2d(cout, cout, 1, 1, 0)]
        if bn:
            layers.insert(2, nn.BatchNorm2d(cout))
        self.convs = nn.Sequential(*layers)

    def forward(self, x):
        return x + self.convs(x)


class _TrunkFn(torch.autograd.Function):
    """Autograd node over ``encdiff_enc4_forward`` / ``encdiff_enc4_backward``: every trunk parameter is an input and gets
    its gradient as an output, so ``param.grad`` accumulates with torch semantics (hooks, DDP reducer)."""

    @staticmethod
    def forward(ctx, enc, x, *params):
        codes = enc._trunk_forward(x)
        ctx.enc, ctx.fwd_id, ctx.B = enc, enc._fwd_id, x.shape[0]
        ctx.need = tuple(p.requires_grad for p in params)
        return codes

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, dcodes):
        enc = ctx.enc
        if ctx.fwd_id != enc._fwd_id:
            raise RuntimeError("encdiff_b200.Encoder4: the activation workspace of this forward was overwritten by a later "
                               "forward; run backward before the next forward")
        grads = enc._trunk_backward(dcodes, ctx.B)
        return (None, None) + tuple(g if need else None for g, need in zip(grads, ctx.need))


class Encoder4(_lib.NativeStateMixin, nn.Module):
    """Drop-in for the reference's concept encoder (``cond_stage_config.target: encdiff_b200.Encoder4``; constructor,
    ``forward`` / ``encoding`` / ``warp`` and ``state_dict`` keys of openaimodel_enc.py:991-1041).  The 20 per-unit MLPs run
    as ONE CUDA launch (``ConceptMLP``; the reference loops over ``self.net[i]`` in Python, :1029-1031 and :1039-1040).  The
    convolutional trunk (four 4x4 stride-2 convolutions, BatchNorm2d in train or eval mode, two EncResBlocks, one Linear:
    0.1 GFLOP per image) runs forward and backward through ``encdiff_enc4_*`` (im2col + tcgen05 GEMMs, csrc/enc4.inc) for
    the shipped configuration (d = 128, bn=True, 64 x 64 images); ``self.encoder`` is the parameter / buffer holder with
    the reference's layer indices.  Other configurations raise at construction unless ``torch_trunk=True`` asks for the
    plain PyTorch ``nn.Sequential`` (a front end outside this package's kernels)."""

    _NATIVE_KEYS = ("_h", "_lib", "_bound_key", "_packed", "_ws", "_gflat", "_grad_views", "_fwd_id")

    def __init__(self, d, context_dim, latent_unit, bn=True, num_channels=3, torch_trunk=False):
        super().__init__()
        self._record_ctor(dict(d=d, context_dim=context_dim, latent_unit=latent_unit, bn=bn, num_channels=num_channels,
                               torch_trunk=torch_trunk))
        self.context_dim, self.latent_unit = context_dim, latent_unit
        self.encoder = nn.Sequential(
            nn.Conv2d(num_channels, d, 4, 2, 1), nn.BatchNorm2d(d), nn.ReLU(inplace=True),
            nn.Conv2d(d, d, 4, 2, 1), nn.BatchNorm2d(d), nn.ReLU(inplace=True),
            nn.Conv2d(d, d, 4, 2, 1), nn.BatchNorm2d(d),
            nn.Conv2d(d, d, 4, 2, 1), nn.BatchNorm2d(d), nn.ReLU(inplace=True),
            _EncResBlock(d, d, bn=bn), nn.BatchNorm2d(d), nn.ReLU(inplace=True),
            _EncResBlock(d, d, bn=bn),
            nn.Flatten(1),                              # the reference's View((-1, 128 * 4 * 4)): no parameters either
            nn.Linear(2048, latent_unit))
        self.net = ConceptMLP(latent_unit, context_dim)
        self.torch_trunk = bool(torch_trunk)
        self.num_channels, self.resolution = num_channels, 64
        self._h = None
        if not self.torch_trunk:
            if d != 128 or not bn or num_channels * 16 > 64 or latent_unit > 64:
                raise NotImplementedError("encdiff_b200.Encoder4: the CUDA trunk implements d=128, bn=True, num_channels<=4, "
                                          "latent_unit<=64 (pass torch_trunk=True for the PyTorch trunk)")
            self._lib = _lib.load()
            cfg = _lib.Enc4ConfigStruct(d, latent_unit, num_channels, 64, 1)
            h = C.c_void_p()
            _lib.check(self._lib.encdiff_enc4_create(C.byref(cfg), C.byref(h)), "enc4_create")
            self._h = h
            mine = list(self.encoder.named_parameters())
            n = self._lib.encdiff_unet_num_params(self._h)
            buf, shape, nd = C.create_string_buffer(256), (C.c_longlong * 4)(), C.c_int()
            if n != len(mine):
                raise RuntimeError(f"encdiff_b200.Encoder4: library expects {n} trunk tensors, module has {len(mine)}")
            for i, (name, p) in enumerate(mine):
                _lib.check(self._lib.encdiff_unet_param_info(self._h, i, buf, 256, shape, C.byref(nd)), "param_info")
                if buf.value.decode() != "encoder." + name or tuple(shape[k] for k in range(nd.value)) != tuple(p.shape):
                    raise RuntimeError(f"encdiff_b200.Encoder4: parameter {i} mismatch: module encoder.{name}{tuple(p.shape)} "
                                       f"vs library {buf.value.decode()}{tuple(shape[k] for k in range(nd.value))}")
        self._bound_key = None
        self._packed = self._ws = self._gflat = None
        self._fwd_id = 0

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self._lib.encdiff_unet_destroy(self._h)
                self._h = None
        except Exception:
            pass

    # ---- CUDA trunk ---------------------------------------------------------------------------------------------
    def _bns(self):
        return [m for m in self.encoder.modules() if isinstance(m, nn.BatchNorm2d)]

    def _bind(self):
        params = list(self.encoder.parameters())
        stats = [t for m in self._bns() for t in (m.running_mean, m.running_var)]
        dev = params[0].device
        if dev.type != "cuda":
            raise RuntimeError("encdiff_b200.Encoder4 runs on CUDA only (sm_100a kernels, no CPU fallback)")
        for p in params + stats:
            if p.dtype != torch.float32 or not p.is_contiguous():
                raise RuntimeError("encdiff_b200.Encoder4 keeps fp32 contiguous parameters and BatchNorm statistics")
        key = tuple(p.data_ptr() for p in params + stats)
        if key == self._bound_key:
            return
        if self._packed is None or self._packed.device != dev:
            self._packed = torch.empty(self._lib.encdiff_unet_packed_bytes(self._h) + 1024, dtype=torch.uint8, device=dev)
            self._gflat = torch.zeros(sum(p.numel() for p in params), dtype=torch.float32, device=dev)
        views, off = [], 0
        for p in params:
            views.append(self._gflat[off:off + p.numel()].view_as(p))
            off += p.numel()
        self._grad_views = views
        ptr = (self._packed.data_ptr() + 1023) // 1024 * 1024
        pp = (C.c_void_p * len(params))(*[p.data_ptr() for p in params])
        gp = (C.c_void_p * len(params))(*[v.data_ptr() for v in views])
        _lib.check(self._lib.encdiff_unet_bind(self._h, pp, gp, ptr, _lib.current_stream()), "enc4 bind")
        sp = (C.c_void_p * len(stats))(*[t.data_ptr() for t in stats])
        _lib.check(self._lib.encdiff_enc4_bind_stats(self._h, sp), "enc4 bind_stats")
        self._bound_key = key

    def _workspace(self, B, dev):
        nbytes = self._lib.encdiff_enc4_workspace_bytes(self._h, B)
        if self._ws is None or self._ws.numel() < nbytes + 1024 or self._ws.device != dev:
            self._ws = torch.empty(nbytes + 1024, dtype=torch.uint8, device=dev)
        return (self._ws.data_ptr() + 1023) // 1024 * 1024

    def _trunk_forward(self, x):
        B = x.shape[0]
        if tuple(x.shape[1:]) != (self.num_channels, self.resolution, self.resolution):
            raise RuntimeError(f"encdiff_b200.Encoder4: expected (B,{self.num_channels},64,64) images, got {tuple(x.shape)}")
        self._bind()
        _lib.check(self._lib.encdiff_unet_pack_weights(self._h, _lib.current_stream()), "enc4 pack")
        xf = x.contiguous().float()
        codes = torch.empty(B, self.latent_unit, dtype=torch.float32, device=x.device)
        bns = self._bns()
        training = bool(self.encoder.training)
        _lib.check(self._lib.encdiff_enc4_forward(self._h, B, xf.data_ptr(), codes.data_ptr(), self._workspace(B, x.device),
                                                  int(training), float(bns[0].momentum), _lib.current_stream()), "enc4_forward")
        if training:
            torch._foreach_add_([m.num_batches_tracked for m in bns], 1)
        self._fwd_id += 1
        return codes

    def _trunk_backward(self, dcodes, B):
        self._gflat.zero_()
        d = dcodes.contiguous().float()
        _lib.check(self._lib.encdiff_enc4_backward(self._h, B, d.data_ptr(), self._workspace(B, d.device),
                                                   _lib.current_stream()), "enc4_backward")
        flat = self._gflat.clone()
        out, off = [], 0
        for v in self._grad_views:
            out.append(flat[off:off + v.numel()].view_as(v))
            off += v.numel()
        return out

    def encoding(self, x):
        if self.torch_trunk:
            return self.encoder(x)
        if not x.is_cuda:
            raise RuntimeError("encdiff_b200.Encoder4 needs CUDA tensors (no CPU fallback)")
        if x.requires_grad:
            raise NotImplementedError("encdiff_b200.Encoder4: gradients w.r.t. the images are not implemented")
        params = list(self.encoder.parameters())
        if torch.is_grad_enabled() and any(p.requires_grad for p in params):
            return _TrunkFn.apply(self, x, *params)
        return self._trunk_forward(x)

    def warp(self, x):
        return self.net.warp(x)

    def forward(self, x):
        return self.net(self.encoding(x))
