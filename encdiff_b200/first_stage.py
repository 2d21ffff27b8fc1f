"""Host side of the first-stage VQ-f4 ENCODE path: ``z = quant_conv(encoder(x))`` as sm_100a kernels behind the C ABI
(``encdiff_vqenc_*``, include/encdiff_b200.h).  Mirrors, for the encode direction, the reference's
``VQModelInterface`` (ldm/models/autoencoder.py:308-316) built from ``Encoder`` (ldm/modules/diffusionmodules/model.py:
368-459): same ``ddconfig`` keywords, same parameter names (``encoder.*``, ``quant_conv.*``) in the same registration
order, so ``load_state_dict`` takes the first-stage entries of a reference checkpoint as they are.  The first stage is
frozen in EncDiff (ddpm_enc.py instantiate_first_stage: eval, requires_grad False), hence inference only; the decoder and
the quantiser stay the reference's PyTorch modules.  No CPU fallback."""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn

from . import _lib


def _norm(c):
    return nn.GroupNorm(num_groups=32, num_channels=c, eps=1e-6, affine=True)      # Normalize(), model.py:37-38


class _ResnetBlock(nn.Module):          # parameter holder; registration order of model.py:82-119 (temb_channels = 0)
    def __init__(self, cin, cout):
        super().__init__()
        self.norm1 = _norm(cin)
        self.conv1 = nn.Conv2d(cin, cout, 3, 1, 1)
        self.norm2 = _norm(cout)
        self.conv2 = nn.Conv2d(cout, cout, 3, 1, 1)
        if cin != cout:
            self.nin_shortcut = nn.Conv2d(cin, cout, 1, 1, 0)


class _AttnBlock(nn.Module):            # model.py:153-176
    def __init__(self, c):
        super().__init__()
        self.norm = _norm(c)
        self.q = nn.Conv2d(c, c, 1)
        self.k = nn.Conv2d(c, c, 1)
        self.v = nn.Conv2d(c, c, 1)
        self.proj_out = nn.Conv2d(c, c, 1)


class _Downsample(nn.Module):           # model.py:63-71 (with_conv)
    def __init__(self, c):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, 2, 0)


class _Encoder(nn.Module):              # model.py:368-435
    def __init__(self, ch, ch_mult, num_res_blocks, in_channels, z_channels, double_z):
        super().__init__()
        self.conv_in = nn.Conv2d(in_channels, ch, 3, 1, 1)
        in_mult = (1,) + tuple(ch_mult)
        self.down = nn.ModuleList()
        block_in = ch
        for lvl in range(len(ch_mult)):
            block_in, block_out = ch * in_mult[lvl], ch * ch_mult[lvl]
            down = nn.Module()
            down.block = nn.ModuleList()
            down.attn = nn.ModuleList()
            for _ in range(num_res_blocks):
                down.block.append(_ResnetBlock(block_in, block_out))
                block_in = block_out
            if lvl != len(ch_mult) - 1:
                down.downsample = _Downsample(block_in)
            self.down.append(down)
        self.mid = nn.Module()
        self.mid.block_1 = _ResnetBlock(block_in, block_in)
        self.mid.attn_1 = _AttnBlock(block_in)
        self.mid.block_2 = _ResnetBlock(block_in, block_in)
        self.norm_out = _norm(block_in)
        self.conv_out = nn.Conv2d(block_in, 2 * z_channels if double_z else z_channels, 3, 1, 1)


class VQEncoder(_lib.NativeStateMixin, nn.Module):
    """``VQEncoder(embed_dim=3, ddconfig={...})`` — the encode half of the reference's ``VQModelInterface``; ``encode(x)`` /
    ``forward(x)``: images [B, 3, R, R] in [-1, 1] -> latents [B, embed_dim, R/4, R/4] (f4)."""

    _NATIVE_KEYS = ("_h", "_lib", "_bound_key", "_packed", "_ws")

    def __init__(self, embed_dim, ddconfig=None, **ddkw):
        super().__init__()
        self._record_ctor(dict(embed_dim=embed_dim, ddconfig=ddconfig, **ddkw))
        dd = dict(ddconfig or {})
        dd.update(ddkw)
        ch, ch_mult = int(dd["ch"]), tuple(int(v) for v in dd.get("ch_mult", (1, 2, 4, 8)))
        nrb, in_ch = int(dd["num_res_blocks"]), int(dd.get("in_channels", 3))
        res, zc = int(dd["resolution"]), int(dd["z_channels"])
        if list(dd.get("attn_resolutions", [])):
            raise NotImplementedError("encdiff_b200.VQEncoder: attn_resolutions must be empty (the shipped first stages)")
        if dd.get("double_z", True):
            raise NotImplementedError("encdiff_b200.VQEncoder: double_z=True (KL first stage) is not implemented")
        if float(dd.get("dropout", 0.0)) != 0.0 or dd.get("use_linear_attn", False) or dd.get("attn_type", "vanilla") != "vanilla":
            raise NotImplementedError("encdiff_b200.VQEncoder: dropout / linear attention are not implemented")
        if not dd.get("resamp_with_conv", True):
            raise NotImplementedError("encdiff_b200.VQEncoder: resamp_with_conv=False is not implemented")
        self.embed_dim, self.resolution, self.in_channels = int(embed_dim), res, in_ch
        self.encoder = _Encoder(ch, ch_mult, nrb, in_ch, zc, False)
        self.quant_conv = nn.Conv2d(zc, int(embed_dim), 1)
        self.out_res = res // (2 ** (len(ch_mult) - 1))
        self._lib = _lib.load()
        cfg = _lib.VQEncConfigStruct()
        cfg.ch, cfg.n_ch_mult = ch, len(ch_mult)
        for i, v in enumerate(ch_mult):
            cfg.ch_mult[i] = v
        cfg.num_res_blocks, cfg.in_channels, cfg.resolution, cfg.z_channels, cfg.embed_dim = nrb, in_ch, res, zc, int(embed_dim)
        h = C.c_void_p()
        _lib.check(self._lib.encdiff_vqenc_create(C.byref(cfg), C.byref(h)), "vqenc_create")
        self._h = h
        n = self._lib.encdiff_unet_num_params(self._h)
        mine = list(self.named_parameters())
        if n != len(mine):
            raise RuntimeError(f"encdiff_b200.VQEncoder: library expects {n} parameter tensors, module has {len(mine)}")
        buf, shape, nd = C.create_string_buffer(256), (C.c_longlong * 4)(), C.c_int()
        for i, (name, p) in enumerate(mine):
            _lib.check(self._lib.encdiff_unet_param_info(self._h, i, buf, 256, shape, C.byref(nd)), "param_info")
            if buf.value.decode() != name or tuple(shape[k] for k in range(nd.value)) != tuple(p.shape):
                raise RuntimeError(f"encdiff_b200.VQEncoder: parameter {i} mismatch: module {name}{tuple(p.shape)} vs library "
                                   f"{buf.value.decode()}{tuple(shape[k] for k in range(nd.value))}")
        self._bound_key = None
        self._packed = None
        self._ws = None
        for p in self.parameters():
            p.requires_grad_(False)          # frozen first stage (ddpm_enc.py instantiate_first_stage)

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self._lib.encdiff_unet_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def refresh_weights(self):
        """(Re-)bind the parameter tensors and pack them into 16-bit GEMM operands; automatic on the first call and
        whenever a parameter tensor moved (``.to()``, ``load_state_dict`` keeps storage and needs an explicit call)."""
        params = [p for _, p in self.named_parameters()]
        dev = params[0].device
        if dev.type != "cuda":
            raise RuntimeError("encdiff_b200.VQEncoder runs on CUDA only (sm_100a kernels, no CPU fallback)")
        for p in params:
            if p.dtype != torch.float32 or not p.is_contiguous():
                raise RuntimeError("encdiff_b200.VQEncoder keeps fp32 contiguous parameters")
        if self._packed is None or self._packed.device != dev:
            self._packed = torch.empty(self._lib.encdiff_unet_packed_bytes(self._h) + 1024, dtype=torch.uint8, device=dev)
        ptr = (self._packed.data_ptr() + 1023) // 1024 * 1024
        pp = (C.c_void_p * len(params))(*[p.data_ptr() for p in params])
        _lib.check(self._lib.encdiff_unet_bind(self._h, pp, None, ptr, _lib.current_stream()), "vqenc bind")
        _lib.check(self._lib.encdiff_unet_pack_weights(self._h, _lib.current_stream()), "vqenc pack")
        self._bound_key = tuple(p.data_ptr() for p in params)

    def load_state_dict(self, *a, **k):
        out = super().load_state_dict(*a, **k)
        self._bound_key = None
        return out

    @torch.no_grad()
    def encode(self, x):
        """VQModelInterface.encode (autoencoder.py:313-316)."""
        if not x.is_cuda:
            raise RuntimeError("encdiff_b200.VQEncoder.encode needs CUDA tensors (no CPU fallback)")
        B = x.shape[0]
        if tuple(x.shape[1:]) != (self.in_channels, self.resolution, self.resolution):
            raise RuntimeError(f"encdiff_b200.VQEncoder: expected (B,{self.in_channels},{self.resolution},{self.resolution}), "
                               f"got {tuple(x.shape)}")
        if self._bound_key is None or self._bound_key[0] != next(self.parameters()).data_ptr():
            self.refresh_weights()
        xf = x.contiguous().float()
        nbytes = self._lib.encdiff_vqenc_workspace_bytes(self._h, B)
        if self._ws is None or self._ws.numel() < nbytes + 1024 or self._ws.device != x.device:
            self._ws = torch.empty(nbytes + 1024, dtype=torch.uint8, device=x.device)
        ws_ptr = (self._ws.data_ptr() + 1023) // 1024 * 1024
        z = torch.empty(B, self.embed_dim, self.out_res, self.out_res, dtype=torch.float32, device=x.device)
        _lib.check(self._lib.encdiff_vqenc_forward(self._h, B, xf.data_ptr(), z.data_ptr(), ws_ptr, _lib.current_stream()),
                   "vqenc_forward")
        return z.type(x.dtype)

    forward = encode


class _Upsample(nn.Module):             # model.py:41-51 (with_conv)
    def __init__(self, c):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, 1, 1)


class _Decoder(nn.Module):              # model.py:462-530: levels are created deepest first and PREPENDED to self.up
    def __init__(self, ch, ch_mult, num_res_blocks, z_channels, out_ch):
        super().__init__()
        L = len(ch_mult)
        block_in = ch * ch_mult[L - 1]
        self.conv_in = nn.Conv2d(z_channels, block_in, 3, 1, 1)
        self.mid = nn.Module()
        self.mid.block_1 = _ResnetBlock(block_in, block_in)
        self.mid.attn_1 = _AttnBlock(block_in)
        self.mid.block_2 = _ResnetBlock(block_in, block_in)
        self.up = nn.ModuleList()
        for lvl in reversed(range(L)):
            block_out = ch * ch_mult[lvl]
            up = nn.Module()
            up.block = nn.ModuleList()
            up.attn = nn.ModuleList()
            for _ in range(num_res_blocks + 1):
                up.block.append(_ResnetBlock(block_in, block_out))
                block_in = block_out
            if lvl != 0:
                up.upsample = _Upsample(block_in)
            self.up.insert(0, up)
        self.norm_out = _norm(block_in)
        self.conv_out = nn.Conv2d(block_in, out_ch, 3, 1, 1)


class _Quantize(nn.Module):             # holder of taming's VectorQuantizer2.embedding (n_e x e_dim)
    def __init__(self, n_embed, embed_dim):
        super().__init__()
        self.embedding = nn.Embedding(n_embed, embed_dim)


class VQDecoder(_lib.NativeStateMixin, nn.Module):
    """The decode half of the reference's ``VQModelInterface`` (ldm/models/autoencoder.py:328-369):
    ``decode(h, force_not_quantize=False, disentangled_repr=None)`` = nearest-code quantisation (unless told not to), concat of
    the disentangled representation over the latent grid (when built with ``use_disentangled_concat``; zeros if none is
    given), ``post_quant_conv`` and the ``Decoder``.  Parameter names ``decoder.*``, ``quantize.embedding.weight``,
    ``post_quant_conv.*`` as in a reference checkpoint.  Inference only, CUDA only."""

    _NATIVE_KEYS = ("_h", "_lib", "_bound_key", "_packed", "_ws")

    def __init__(self, embed_dim, n_embed=0, ddconfig=None, use_disentangled_concat=False, disentangled_dim=0, **ddkw):
        super().__init__()
        self._record_ctor(dict(embed_dim=embed_dim, n_embed=n_embed, ddconfig=ddconfig,
                               use_disentangled_concat=use_disentangled_concat, disentangled_dim=disentangled_dim, **ddkw))
        dd = dict(ddconfig or {})
        dd.update(ddkw)
        ch, ch_mult = int(dd["ch"]), tuple(int(v) for v in dd.get("ch_mult", (1, 2, 4, 8)))
        nrb, res, zc, out_ch = int(dd["num_res_blocks"]), int(dd["resolution"]), int(dd["z_channels"]), int(dd.get("out_ch", 3))
        if list(dd.get("attn_resolutions", [])):
            raise NotImplementedError("encdiff_b200.VQDecoder: attn_resolutions must be empty (the shipped first stages)")
        if float(dd.get("dropout", 0.0)) != 0.0 or dd.get("use_linear_attn", False) or dd.get("attn_type", "vanilla") != "vanilla":
            raise NotImplementedError("encdiff_b200.VQDecoder: dropout / linear attention are not implemented")
        if dd.get("give_pre_end", False) or dd.get("tanh_out", False) or not dd.get("resamp_with_conv", True):
            raise NotImplementedError("encdiff_b200.VQDecoder: give_pre_end / tanh_out / resamp_with_conv=False are not implemented")
        self.embed_dim, self.n_embed, self.resolution, self.out_ch = int(embed_dim), int(n_embed), res, out_ch
        self.use_disentangled_concat = bool(use_disentangled_concat)
        self.disentangled_dim = int(disentangled_dim) if use_disentangled_concat else 0
        self.latent_res = res // (2 ** (len(ch_mult) - 1))
        self.decoder = _Decoder(ch, ch_mult, nrb, zc, out_ch)
        if self.n_embed > 0:
            self.quantize = _Quantize(self.n_embed, self.embed_dim)
        self.post_quant_conv = nn.Conv2d(self.embed_dim + self.disentangled_dim, zc, 1)
        self._lib = _lib.load()
        cfg = _lib.VQDecConfigStruct()
        cfg.dd.ch, cfg.dd.n_ch_mult = ch, len(ch_mult)
        for i, v in enumerate(ch_mult):
            cfg.dd.ch_mult[i] = v
        cfg.dd.num_res_blocks, cfg.dd.in_channels, cfg.dd.resolution = nrb, int(dd.get("in_channels", 3)), res
        cfg.dd.z_channels, cfg.dd.embed_dim = zc, self.embed_dim
        cfg.out_ch, cfg.n_embed, cfg.disentangled_dim = out_ch, self.n_embed, self.disentangled_dim
        h = C.c_void_p()
        _lib.check(self._lib.encdiff_vqdec_create(C.byref(cfg), C.byref(h)), "vqdec_create")
        self._h = h
        n = self._lib.encdiff_unet_num_params(self._h)
        mine = list(self.named_parameters())
        if n != len(mine):
            raise RuntimeError(f"encdiff_b200.VQDecoder: library expects {n} parameter tensors, module has {len(mine)}")
        buf, shape, nd = C.create_string_buffer(256), (C.c_longlong * 4)(), C.c_int()
        for i, (name, p) in enumerate(mine):
            _lib.check(self._lib.encdiff_unet_param_info(self._h, i, buf, 256, shape, C.byref(nd)), "param_info")
            if buf.value.decode() != name or tuple(shape[k] for k in range(nd.value)) != tuple(p.shape):
                raise RuntimeError(f"encdiff_b200.VQDecoder: parameter {i} mismatch: module {name}{tuple(p.shape)} vs library "
                                   f"{buf.value.decode()}{tuple(shape[k] for k in range(nd.value))}")
        self._bound_key = None
        self._packed = None
        self._ws = None
        for p in self.parameters():
            p.requires_grad_(False)

    __del__ = VQEncoder.__del__
    refresh_weights = VQEncoder.refresh_weights

    def load_state_dict(self, *a, **k):
        out = super().load_state_dict(*a, **k)
        self._bound_key = None
        return out

    @torch.no_grad()
    def decode(self, h, force_not_quantize=False, disentangled_repr=None):
        if not h.is_cuda:
            raise RuntimeError("encdiff_b200.VQDecoder.decode needs CUDA tensors (no CPU fallback)")
        B = h.shape[0]
        if tuple(h.shape[1:]) != (self.embed_dim, self.latent_res, self.latent_res):
            raise RuntimeError(f"encdiff_b200.VQDecoder: expected (B,{self.embed_dim},{self.latent_res},{self.latent_res}), "
                               f"got {tuple(h.shape)}")
        if not force_not_quantize and self.n_embed <= 0:
            raise RuntimeError("encdiff_b200.VQDecoder: built without a codebook (n_embed=0): pass force_not_quantize=True")
        if self._bound_key is None or self._bound_key[0] != next(self.parameters()).data_ptr():
            self.refresh_weights()
        hf = h.contiguous().float()
        s = None
        if self.disentangled_dim > 0 and disentangled_repr is not None:
            s = disentangled_repr.reshape(B, -1).contiguous().float()
            if s.shape[1] != self.disentangled_dim:
                raise RuntimeError(f"encdiff_b200.VQDecoder: disentangled_repr has {s.shape[1]} dims, model has {self.disentangled_dim}")
        nbytes = self._lib.encdiff_vqdec_workspace_bytes(self._h, B)
        if self._ws is None or self._ws.numel() < nbytes + 1024 or self._ws.device != h.device:
            self._ws = torch.empty(nbytes + 1024, dtype=torch.uint8, device=h.device)
        ws_ptr = (self._ws.data_ptr() + 1023) // 1024 * 1024
        img = torch.empty(B, self.out_ch, self.resolution, self.resolution, dtype=torch.float32, device=h.device)
        _lib.check(self._lib.encdiff_vqdec_forward(self._h, B, hf.data_ptr(), _lib.ptr(s), 0 if force_not_quantize else 1,
                                                   img.data_ptr(), ws_ptr, _lib.current_stream()), "vqdec_forward")
        return img.type(h.dtype)

    forward = decode


class VQModelInterface(nn.Module):
    """Drop-in for the reference's frozen first stage as LatentDiffusion builds it (``first_stage_config.target:
    encdiff_b200.VQModelInterface``; ldm/models/autoencoder.py:20-75 constructor keywords, :308-369 ``encode`` / ``decode``):
    one module with the reference's sub-module names in the reference's registration order (``encoder``, ``decoder``,
    ``loss``, ``quantize``, ``quant_conv``, ``post_quant_conv``), so ``state_dict`` keys and their order are those of a
    reference checkpoint, over the two CUDA paths (``VQEncoder`` / ``VQDecoder`` share these parameter holders).  Inference
    only, as the first stage is used by ``LatentDiffusion`` (instantiate_first_stage: eval, requires_grad False); training
    the first stage itself, EMA, colorize, remap and per-batch resizing are the reference's business and raise here."""

    def __init__(self, embed_dim, ddconfig=None, lossconfig=None, n_embed=0, ckpt_path=None, ignore_keys=(), image_key="image",
                 colorize_nlabels=None, monitor=None, batch_resize_range=None, scheduler_config=None, lr_g_factor=1.0,
                 remap=None, sane_index_shape=False, use_ema=False, use_disentangled_concat=False, disentangled_dim=0):
        super().__init__()
        bad = {"colorize_nlabels": colorize_nlabels is not None, "batch_resize_range": batch_resize_range is not None,
               "remap": remap is not None, "use_ema": bool(use_ema)}
        if any(bad.values()):
            raise NotImplementedError("encdiff_b200.VQModelInterface does not implement: " +
                                      ", ".join(k for k, v in bad.items() if v))
        if lossconfig is not None and dict(lossconfig).get("target", "torch.nn.Identity") != "torch.nn.Identity":
            raise NotImplementedError("encdiff_b200.VQModelInterface is the frozen first stage: lossconfig must be "
                                      "torch.nn.Identity (as in the EncDiff yamls)")
        self.embed_dim, self.n_embed, self.image_key = int(embed_dim), int(n_embed), image_key
        self.use_disentangled_concat, self.disentangled_dim = bool(use_disentangled_concat), int(disentangled_dim)
        if monitor is not None:
            self.monitor = monitor
        enc = VQEncoder(embed_dim=embed_dim, ddconfig=ddconfig)
        dec = VQDecoder(embed_dim=embed_dim, n_embed=n_embed, ddconfig=ddconfig, use_disentangled_concat=use_disentangled_concat,
                        disentangled_dim=disentangled_dim)
        if self.n_embed <= 0:
            raise ValueError("encdiff_b200.VQModelInterface needs n_embed > 0 (the VQ codebook size)")
        # the reference's registration order (autoencoder.py:46-56); the two CUDA paths keep working on the same modules
        self.encoder = enc.encoder
        self.decoder = dec.decoder
        self.loss = nn.Identity()
        self.quantize = dec.quantize
        self.quant_conv = enc.quant_conv
        self.post_quant_conv = dec.post_quant_conv
        object.__setattr__(self, "_enc", enc)            # not registered: their parameters are the ones above
        object.__setattr__(self, "_dec", dec)
        if ckpt_path is not None:
            self.init_from_ckpt(ckpt_path, ignore_keys=ignore_keys)

    def init_from_ckpt(self, path, ignore_keys=()):
        """VQModel.init_from_ckpt (autoencoder.py:92-104): ``state_dict`` entry of a Lightning checkpoint, keys starting
        with one of ``ignore_keys`` dropped, non-strict."""
        sd = torch.load(path, map_location="cpu")["state_dict"]
        sd = {k: v for k, v in sd.items() if not any(k.startswith(ik) for ik in ignore_keys)}
        missing, unexpected = self.load_state_dict(sd, strict=False)
        print(f"Restored from {path} with {len(missing)} missing and {len(unexpected)} unexpected keys")

    def load_state_dict(self, *a, **k):
        out = super().load_state_dict(*a, **k)
        self._enc._bound_key = self._dec._bound_key = None          # same storages, new values: re-pack on the next call
        return out

    def encode(self, x):
        return self._enc.encode(x)

    def decode(self, h, force_not_quantize=False, disentangled_repr=None):
        return self._dec.decode(h, force_not_quantize=force_not_quantize, disentangled_repr=disentangled_repr)

    def forward(self, x):
        raise NotImplementedError("encdiff_b200.VQModelInterface is used through encode() / decode() (frozen first stage)")


# ---- reference seam ------------------------------------------------------------------------------------------------
def _ddconfig_of(fs):
    """ddconfig of a reference VQModel(Interface) instance, read back from its modules (the constructor does not keep it)."""
    enc = fs.encoder
    ch = enc.ch
    mult = [blk.block[-1].out_channels // ch for blk in enc.down]
    return dict(double_z=False, z_channels=fs.quant_conv.in_channels, resolution=enc.resolution, in_channels=enc.in_channels,
                out_ch=fs.decoder.conv_out.out_channels, ch=ch, ch_mult=mult, num_res_blocks=enc.num_res_blocks,
                attn_resolutions=[], dropout=0.0)


def from_reference(fs, device=None):
    """(VQEncoder, VQDecoder) holding the weights of a reference ``VQModelInterface`` instance (``model.first_stage_model``)."""
    dd = _ddconfig_of(fs)
    sd = fs.state_dict()
    dev = device if device is not None else next(fs.parameters()).device
    enc = VQEncoder(embed_dim=fs.embed_dim, ddconfig=dd)
    enc.load_state_dict({k: v for k, v in sd.items() if k.startswith(("encoder.", "quant_conv."))}, strict=True)
    cat = bool(getattr(fs, "use_disentangled_concat", False))
    dec = VQDecoder(embed_dim=fs.embed_dim, n_embed=fs.n_embed, ddconfig=dd, use_disentangled_concat=cat,
                    disentangled_dim=getattr(fs, "disentangled_dim", 0) if cat else 0)
    dec.load_state_dict({k: v for k, v in sd.items()
                         if k.startswith(("decoder.", "post_quant_conv.")) or k == "quantize.embedding.weight"}, strict=True)
    return enc.to(dev), dec.to(dev)


def install(fs):
    """Rebind ``fs.encode`` / ``fs.decode`` of a reference ``VQModelInterface`` instance to the CUDA paths (the frozen first
    stage of LatentDiffusion: ``encode_first_stage`` / ``decode_first_stage`` call exactly these two methods,
    ddpm_enc.py).  Returns (encoder, decoder); ``fs`` keeps its own modules and parameters (checkpoints keep loading; call
    ``install`` again after loading new first-stage weights)."""
    enc, dec = from_reference(fs)
    fs.encode = enc.encode
    fs.decode = dec.decode
    fs._encdiff_b200 = (enc, dec)
    return enc, dec
