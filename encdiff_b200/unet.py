"""Drop-in ``UNetModel`` (reference: ldm/modules/diffusionmodules/openaimodel_enc.py:413-748).

Select it through the reference's own factory (ldm/util.py:78-93) by setting
``unet_config.target: encdiff_b200.UNetModel`` — constructor parameters, parameter names/shapes/order, and the
``forward(x, timesteps, context)`` signature are the reference's.  All compute runs in libencdiff_b200.so (hand-written
sm_100a kernels) through the C ABI; there is no PyTorch/CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch
from torch import nn

from . import _lib
from .modules import (ResBlock, SpatialTransformer, TimestepEmbedSequential, normalization, zero_module)


def _as_list(v):
    # OmegaConf ListConfig / tuple / list all iterate; avoid importing omegaconf (the reference does, :477)
    return [int(i) for i in v]


class _UNetFn(torch.autograd.Function):
    """Autograd node over the C-ABI forward/backward.  EVERY parameter is an input of the node and receives its gradient
    as an output of ``backward`` (views of the flat fp32 bucket the library accumulated into), so each parameter's
    AccumulateGrad node runs: ``param.grad`` accumulates with torch semantics, gradient hooks fire, and
    ``torch.nn.parallel.DistributedDataParallel`` (the reference's Lightning 'ddp' strategy, main_val.py:656-660) sees
    all 734 tensors in its bucket reducer.  The fused trainer (train.DenoiserTrainer) does not come through here."""

    @staticmethod
    def forward(ctx, model, x, t, cond, *params):
        out = model._auto_forward(x, t, cond)
        ctx.model = model
        ctx.fwd_id = model._fwd_id
        ctx.save_for_backward(cond)
        ctx.need_dx = x.requires_grad
        ctx.need_param = tuple(p.requires_grad for p in params)
        return out

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, dout):
        model = ctx.model
        (cond,) = ctx.saved_tensors
        if ctx.fwd_id != model._fwd_id:
            raise RuntimeError("encdiff_b200.UNetModel: the activation workspace of this forward was overwritten by a "
                               "later forward; run backward before the next forward")
        dx, dctx, grads = model._auto_backward(dout, cond, ctx.need_dx)
        return (None, dx, None, dctx) + tuple(g if need else None for g, need in zip(grads, ctx.need_param))


class UNetModel(_lib.NativeStateMixin, nn.Module):
    """The full UNet with spatial transformers and timestep embedding; same constructor as the reference
    (openaimodel_enc.py:443-470).  Combinations the CUDA plan does not implement raise instead of silently differing."""

    _NATIVE_KEYS = ("_h", "_lib", "_bound_key", "_packed", "_gflat", "_grad_views", "_ws", "_fwd_id", "_planned", "_auto",
                    "_prof_on", "_packed_valid")

    def __init__(self, image_size, in_channels, model_channels, out_channels, num_res_blocks, attention_resolutions,
                 latent_unit, dropout=0, channel_mult=(1, 2, 4, 8), conv_resample=True, dims=2, num_classes=None,
                 use_checkpoint=False, use_fp16=False, num_heads=-1, num_head_channels=-1, num_heads_upsample=-1,
                 use_scale_shift_norm=False, resblock_updown=False, use_new_attention_order=False,
                 use_spatial_transformer=False, transformer_depth=1, context_dim=None, n_embed=None, legacy=True):
        kw = dict(locals())
        super().__init__()
        self._record_ctor(kw)
        if not use_spatial_transformer:
            raise NotImplementedError("encdiff_b200.UNetModel: use_spatial_transformer=False (AttentionBlock) is not supported")
        if context_dim is None:
            raise AssertionError("context_dim is required with use_spatial_transformer")
        if not isinstance(context_dim, int):
            context_dim = _as_list(context_dim)
            if len(context_dim) != 1:
                raise NotImplementedError("encdiff_b200.UNetModel: a single context_dim is supported")
            context_dim = context_dim[0]
        unsupported = {"dims != 2": dims != 2, "num_classes": num_classes is not None, "n_embed": n_embed is not None,
                       "use_fp16": bool(use_fp16), "dropout > 0": dropout != 0,
                       "use_scale_shift_norm=False": not use_scale_shift_norm,
                       "resblock_updown=False": not resblock_updown, "transformer_depth != 1": transformer_depth != 1,
                       "num_head_channels": num_head_channels != -1, "legacy=False": not legacy,
                       "num_heads_upsample": num_heads_upsample not in (-1, num_heads)}
        bad = [k for k, v in unsupported.items() if v]
        if bad:
            raise NotImplementedError("encdiff_b200.UNetModel does not implement: " + ", ".join(bad))
        if num_heads == -1:
            raise AssertionError("Either num_heads or num_head_channels has to be set")
        attention_resolutions = _as_list(attention_resolutions)
        channel_mult = _as_list(channel_mult)

        self.image_size = image_size
        self.in_channels = in_channels
        self.model_channels = model_channels
        self.out_channels = out_channels
        self.num_res_blocks = num_res_blocks
        self.attention_resolutions = attention_resolutions
        self.dropout = dropout
        self.channel_mult = channel_mult
        self.conv_resample = conv_resample
        self.num_classes = num_classes
        self.use_checkpoint = use_checkpoint
        self.dtype = torch.float32
        self.num_heads = num_heads
        self.num_head_channels = num_head_channels
        self.num_heads_upsample = num_heads
        self.predict_codebook_ids = False
        self.context_dim = context_dim
        self.latent_unit = latent_unit

        # ---- parameter tree, constructed in the reference's order (openaimodel_enc.py:506-688) ----
        ted = model_channels * 4
        self.time_embed = nn.Sequential(nn.Linear(model_channels, ted), nn.SiLU(), nn.Linear(ted, ted))
        self.input_blocks = nn.ModuleList(
            [TimestepEmbedSequential(nn.Conv2d(in_channels, model_channels, 3, padding=1))])
        chans = [model_channels]
        ch, ds = model_channels, 1

        def st(c):
            return SpatialTransformer(c, num_heads, c // num_heads, depth=transformer_depth, context_dim=context_dim)

        for level, mult in enumerate(channel_mult):
            for _ in range(num_res_blocks):
                layers = [ResBlock(ch, ted, dropout, out_channels=mult * model_channels)]
                ch = mult * model_channels
                if ds in attention_resolutions:
                    layers.append(st(ch))
                self.input_blocks.append(TimestepEmbedSequential(*layers))
                chans.append(ch)
            if level != len(channel_mult) - 1:
                self.input_blocks.append(TimestepEmbedSequential(ResBlock(ch, ted, dropout, out_channels=ch, down=True)))
                chans.append(ch)
                ds *= 2
        self.middle_block = TimestepEmbedSequential(ResBlock(ch, ted, dropout), st(ch), ResBlock(ch, ted, dropout))
        self.output_blocks = nn.ModuleList([])
        for level, mult in list(enumerate(channel_mult))[::-1]:
            for i in range(num_res_blocks + 1):
                ich = chans.pop()
                layers = [ResBlock(ch + ich, ted, dropout, out_channels=model_channels * mult)]
                ch = model_channels * mult
                if ds in attention_resolutions:
                    layers.append(st(ch))
                if level and i == num_res_blocks:
                    layers.append(ResBlock(ch, ted, dropout, out_channels=ch, up=True))
                    ds //= 2
                self.output_blocks.append(TimestepEmbedSequential(*layers))
        self.out = nn.Sequential(normalization(ch), nn.SiLU(),
                                 zero_module(nn.Conv2d(model_channels, out_channels, 3, padding=1)))

        # ---- native handle: architecture check against the library's own inventory ----
        self._lib = _lib.load()
        cfg = _lib.UNetConfigStruct()
        cfg.image_size, cfg.in_channels, cfg.model_channels = image_size, in_channels, model_channels
        cfg.out_channels, cfg.num_res_blocks = out_channels, num_res_blocks
        cfg.n_attention_resolutions = len(attention_resolutions)
        for i, v in enumerate(attention_resolutions):
            cfg.attention_resolutions[i] = v
        cfg.n_channel_mult = len(channel_mult)
        for i, v in enumerate(channel_mult):
            cfg.channel_mult[i] = v
        cfg.num_heads, cfg.context_dim, cfg.latent_unit = num_heads, context_dim, latent_unit
        cfg.transformer_depth = transformer_depth
        cfg.use_scale_shift_norm, cfg.resblock_updown = 1, 1
        h = C.c_void_p()
        _lib.check(self._lib.encdiff_unet_create(C.byref(cfg), C.byref(h)), "unet_create")
        self._h = h
        self._check_inventory()
        self._bound_key = None
        self._packed = None
        self._gflat = None
        self._grad_views = None
        self._ws = {}
        self._fwd_id = 0
        self._planned = None
        self.cache_packed_weights = False   # set by the sampler: parameters do not change inside its loop
        # The autograd path (forward() under grad mode + loss.backward(): what a reference training loop drives) replays
        # captured CUDA graphs of the library's forward / backward launch sequences from the third call of a shape on
        # (call 1 runs eagerly and lets the library build its device tables, call 2 captures).  ~330 / ~430 launches
        # become one graph launch each.  Set False to run every call eagerly.
        self.graph_autograd = True
        self._auto = {}
        self._prof_on = False
        self._packed_valid = False

    # ------------------------------------------------------------------------------------------------------------
    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and getattr(self, "_lib", None) is not None:
            try:
                self._lib.encdiff_unet_destroy(h)
            except Exception:
                pass

    def _check_inventory(self):
        n = self._lib.encdiff_unet_num_params(self._h)
        mine = list(self.named_parameters())
        if n != len(mine):
            raise RuntimeError(f"encdiff_b200: library expects {n} parameter tensors, module has {len(mine)}")
        buf = C.create_string_buffer(256)
        shape = (C.c_longlong * 4)()
        nd = C.c_int()
        for i, (name, p) in enumerate(mine):
            _lib.check(self._lib.encdiff_unet_param_info(self._h, i, buf, 256, shape, C.byref(nd)), "param_info")
            lname = buf.value.decode()
            lshape = tuple(shape[k] for k in range(nd.value))
            if lname != name or lshape != tuple(p.shape):
                raise RuntimeError(f"encdiff_b200: parameter {i} mismatch: module {name}{tuple(p.shape)} vs library "
                                   f"{lname}{lshape}")

    # ------------------------------------------------------------------------------------------------------------
    def _params(self):
        return [p for _, p in self.named_parameters()]

    def _bind(self, need_grads: bool):
        params = self._params()
        dev = params[0].device
        if dev.type != "cuda":
            raise RuntimeError("encdiff_b200.UNetModel runs on CUDA only (sm_100a kernels, no CPU fallback): "
                               "move the module with .cuda() / .to('cuda')")
        for p in params:
            if p.dtype != torch.float32 or not p.is_contiguous():
                raise RuntimeError("encdiff_b200.UNetModel keeps fp32 contiguous master parameters")
        key = (tuple(p.data_ptr() for p in params), need_grads or self._gflat is not None)
        if key == self._bound_key:
            return
        n = len(params)
        if self._packed is None or self._packed.device != dev:
            nbytes = self._lib.encdiff_unet_packed_bytes(self._h)
            self._packed = torch.empty(nbytes + 1024, dtype=torch.uint8, device=dev)
        packed_ptr = (self._packed.data_ptr() + 1023) // 1024 * 1024
        pp = (C.c_void_p * n)(*[p.data_ptr() for p in params])
        gp = None
        if key[1]:
            if self._gflat is None or self._gflat.device != dev:
                total = sum(p.numel() for p in params)
                self._gflat = torch.zeros(total, dtype=torch.float32, device=dev)
                views, off = [], 0
                for p in params:
                    views.append(self._gflat[off:off + p.numel()].view_as(p))
                    off += p.numel()
                self._grad_views = views
            gp = (C.c_void_p * n)(*[v.data_ptr() for v in self._grad_views])
        _lib.check(self._lib.encdiff_unet_bind(self._h, pp, gp, packed_ptr, _lib.current_stream()), "unet_bind")
        self._bound_key = key
        self._packed_valid = False

    def refresh_weights(self):
        """Re-pack the fp32 parameters into the bf16 GEMM operands.  Needed after in-place parameter updates that do
        not go through this module (optimizer.step, LitEma.copy_to, load_state_dict); UNetModel.forward does it on
        every call unless ``cache_packed_weights`` is set."""
        self._bind(need_grads=False)
        _lib.check(self._lib.encdiff_unet_pack_weights(self._h, _lib.current_stream()), "pack_weights")
        self._packed_valid = True

    def _workspace(self, B: int, training: bool, dev):
        key = (B, training)
        ws = self._ws.get(key)
        if ws is None or ws.device != dev:
            nbytes = self._lib.encdiff_unet_workspace_bytes(self._h, B, int(training))
            ws = torch.empty(nbytes + 1024, dtype=torch.uint8, device=dev)
            # two live workspaces (most recently used): a training loop that samples now and then (train at B = 256,
            # log_images at B = 128) keeps both instead of freeing and re-allocating 6 GB at every switch; a third
            # shape evicts the older one
            self._ws.pop(key, None)
            while len(self._ws) >= 2:
                self._ws.pop(next(iter(self._ws)))
            self._ws[key] = ws
        else:
            self._ws[key] = self._ws.pop(key)        # most recently used last
        return ws

    def plan_key(self):
        """Identity of everything a captured graph of forward() bakes in: workspace plan and parameter binding."""
        return (self._planned, self._bound_key, None if self._packed is None else self._packed.data_ptr())

    def _prepare_grads(self):
        """param.grad must be the views of the flat bucket the library accumulates into (zeroed when re-attached)."""
        params = self._params()
        for p, v in zip(params, self._grad_views):
            if p.grad is None:
                v.zero_()
                p.grad = v
            elif p.grad.data_ptr() != v.data_ptr():
                v.copy_(p.grad)
                p.grad = v

    def flat_grad(self) -> Optional[torch.Tensor]:
        """The flat fp32 gradient bucket (one NCCL all-reduce covers all 734 tensors)."""
        return self._gflat

    def _run_forward(self, x, t, cond, training: bool, noise=None, sqrt_acp=None, sqrt_1macp=None, x_t_out=None):
        B = x.shape[0]
        self._bind(need_grads=training)
        if not (self.cache_packed_weights and self._packed_valid):
            _lib.check(self._lib.encdiff_unet_pack_weights(self._h, _lib.current_stream()), "pack_weights")
            self._packed_valid = True
        ws = self._workspace(B, training, x.device)
        ws_ptr = (ws.data_ptr() + 1023) // 1024 * 1024
        plan_key = (B, training, ws_ptr)
        if self._planned != plan_key:
            _lib.check(self._lib.encdiff_unet_plan(self._h, B, ws_ptr, int(training)), "unet_plan")
            self._planned = plan_key
        out = torch.empty(B, self.out_channels, x.shape[2], x.shape[3], dtype=torch.float32, device=x.device)
        _lib.check(self._lib.encdiff_unet_forward(self._h, B, x.data_ptr(), _lib.ptr(noise), _lib.ptr(sqrt_acp),
                                                  _lib.ptr(sqrt_1macp), t.data_ptr(), cond.data_ptr(), out.data_ptr(),
                                                  _lib.ptr(x_t_out), _lib.current_stream()), "unet_forward")
        self._fwd_id += 1
        return out

    def _run_backward(self, dout, cond, need_dx: bool):
        """Backward of the last forward for the autograd node: the library accumulates the parameter gradients into the
        flat bucket, which is zeroed first; the returned views are handed to autograd, whose AccumulateGrad nodes copy
        / add them into ``param.grad``."""
        B = dout.shape[0]
        for p, v in zip(self._params(), self._grad_views):
            if p.grad is not None and p.grad.data_ptr() == v.data_ptr():
                p.grad = p.grad.clone()      # a fused-trainer step left param.grad aliasing the bucket: un-alias it
        self._gflat.zero_()
        dout = dout.contiguous().float()
        dx = torch.empty(B, self.in_channels, self.image_size, self.image_size, dtype=torch.float32,
                         device=dout.device) if need_dx else None
        dctx = torch.empty_like(cond)
        _lib.check(self._lib.encdiff_unet_backward(self._h, B, dout.data_ptr(), cond.data_ptr(), _lib.ptr(dx),
                                                   dctx.data_ptr(), _lib.current_stream()), "unet_backward")
        # (a no-op unless a data-parallel DenoiserTrainer set a gradient split on this handle)
        _lib.check(self._lib.encdiff_unet_backward_tail(self._h, _lib.current_stream()), "unet_backward_tail")
        # clones: AccumulateGrad may keep the tensor it is handed as param.grad, and the bucket is reused next backward
        parts = self._gflat.clone().split([v.numel() for v in self._grad_views])
        return dx, dctx, tuple(c.view_as(v) for c, v in zip(parts, self._grad_views))

    @torch.no_grad()
    def forward_fp32(self, x, timesteps, context):
        """The library's fp32 VALIDATION forward (csrc/fp32_ref.cu, ``encdiff_unet_forward_fp32``): fp32 tensors and
        CUDA-core FMA contractions over the fp32 master parameters -- the north star's "fp32 validation build"
        (eps max relative error <= 1e-4).  A checker for tests at shapes the CPU oracle cannot reach; not a product path
        (no autograd, ~100x slower than ``forward``)."""
        cond = context[0] if isinstance(context, (list, tuple)) else context
        if x.device.type != "cuda":
            raise RuntimeError("encdiff_b200.UNetModel.forward_fp32 runs on CUDA only")
        B = x.shape[0]
        x = x.contiguous().float()
        t = timesteps.to(device=x.device, dtype=torch.float32).contiguous()
        cond = cond.reshape(B, -1).contiguous().float()
        self._bind(need_grads=False)
        nbytes = self._lib.encdiff_unet_fp32_workspace_bytes(self._h, B)
        if nbytes < 0:
            raise RuntimeError("encdiff_b200: fp32_workspace_bytes failed")
        ws = torch.empty(nbytes + 256, dtype=torch.uint8, device=x.device)
        ws_ptr = (ws.data_ptr() + 255) // 256 * 256
        out = torch.empty(B, self.out_channels, x.shape[2], x.shape[3], dtype=torch.float32, device=x.device)
        _lib.check(self._lib.encdiff_unet_forward_fp32(self._h, B, x.data_ptr(), t.data_ptr(), cond.data_ptr(),
                                                       out.data_ptr(), ws_ptr, nbytes, _lib.current_stream()),
                   "unet_forward_fp32")
        return out

    # ---- graph replay of the autograd path -------------------------------------------------------------------------
    def _graph_ok(self):
        # cache_packed_weights (sampler loops) skips the re-pack: such calls stay eager so that every captured forward
        # contains the pack kernel and always sees the current master weights
        return (self.graph_autograd and not self._prof_on and not self.cache_packed_weights
                and not torch.cuda.is_current_stream_capturing())

    def _auto_forward(self, x, t, cond):
        if not self._graph_ok():
            return self._run_forward(x, t, cond, training=True)
        st = self._auto.setdefault(("fwd", x.shape[0]), {})
        if st.get("plan") is None or st["plan"] != self.plan_key():
            out = self._run_forward(x, t, cond, training=True)       # eager: the library builds its tables for this plan
            st.clear()
            st["plan"] = self.plan_key()
            return out
        if st.get("g") is None:
            sx, stt, sc = x.clone(), t.clone(), cond.clone()
            torch.cuda.current_stream().synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                so = self._run_forward(sx, stt, sc, training=True)
            if st["plan"] != self.plan_key():                        # the capture itself must not have re-planned
                raise RuntimeError("encdiff_b200.UNetModel: plan changed during graph capture")
            st.update(g=g, x=sx, t=stt, c=sc, out=so)
        else:
            st["x"].copy_(x)
            st["t"].copy_(t)
            st["c"].copy_(cond)
        st["g"].replay()
        self._fwd_id += 1
        return st["out"].clone()

    def _auto_backward(self, dout, cond, need_dx):
        if not self._graph_ok():
            return self._run_backward(dout, cond, need_dx)
        st = self._auto.setdefault(("bwd", dout.shape[0], bool(need_dx)), {})
        if st.get("plan") is None or st["plan"] != self.plan_key():
            res = self._run_backward(dout, cond, need_dx)            # eager: builds the weight-gradient tables
            st.clear()
            st["plan"] = self.plan_key()
            return res
        dout = dout.contiguous().float()
        if st.get("g") is None:
            for p, v in zip(self._params(), self._grad_views):
                if p.grad is not None and p.grad.data_ptr() == v.data_ptr():
                    p.grad = p.grad.clone()
            sd, sc = dout.clone(), cond.clone()
            B = dout.shape[0]
            dx = torch.empty(B, self.in_channels, self.image_size, self.image_size, dtype=torch.float32,
                             device=dout.device) if need_dx else None
            dctx = torch.empty_like(sc)
            torch.cuda.current_stream().synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._gflat.zero_()
                _lib.check(self._lib.encdiff_unet_backward(self._h, B, sd.data_ptr(), sc.data_ptr(), _lib.ptr(dx),
                                                           dctx.data_ptr(), _lib.current_stream()), "unet_backward")
                _lib.check(self._lib.encdiff_unet_backward_tail(self._h, _lib.current_stream()), "unet_backward_tail")
            st.update(g=g, d=sd, c=sc, dx=dx, dctx=dctx)
        else:
            for p, v in zip(self._params(), self._grad_views):
                if p.grad is not None and p.grad.data_ptr() == v.data_ptr():
                    p.grad = p.grad.clone()
            st["d"].copy_(dout)
            st["c"].copy_(cond)
        st["g"].replay()
        parts = self._gflat.clone().split([v.numel() for v in self._grad_views])
        dx = st["dx"].clone() if need_dx else None
        return dx, st["dctx"].clone(), tuple(c.view_as(v) for c, v in zip(parts, self._grad_views))

    def profile(self, enable: bool):
        """Switch the library's per-launch event timing on/off (bench.py roofline figures)."""
        self._prof_on = bool(enable)
        _lib.check(self._lib.encdiff_unet_profile(self._h, int(enable)), "unet_profile")

    def profile_read(self):
        """{family: (launches, total_ms, total_flops, total_algorithmic_bytes)} since the last read."""
        buf = C.create_string_buffer(1 << 16)
        n = self._lib.encdiff_unet_profile_read(self._h, buf, len(buf))
        if n < 0:
            raise RuntimeError("encdiff_b200: profile_read failed")
        out = {}
        for line in buf.value.decode().splitlines():
            name, k, ms, fl, by = line.split(",")
            out[name] = (int(k), float(ms), float(fl), float(by))
        return out

    def debug_block(self, section: int, index: int = 0) -> torch.Tensor:
        """Output of input_blocks[index] (section 0) / middle_block (1) / output_blocks[index] (2) of the last forward,
        as a (B, C, H, W) fp32 copy — what a forward hook on the reference module would see (parity tests only)."""
        ptr, c, h, w = C.c_void_p(), C.c_int(), C.c_int(), C.c_int()
        _lib.check(self._lib.encdiff_unet_debug_block(self._h, section, index, C.byref(ptr), C.byref(c), C.byref(h),
                                                      C.byref(w)), "debug_block")
        (B, _training), ws = list(self._ws.items())[-1]       # the workspace of the last forward
        off = ptr.value - ws.data_ptr()
        n = B * h.value * w.value * c.value
        t = ws[off:off + 2 * n].view(_lib.act_dtype()).view(B, h.value, w.value, c.value)
        return t.float().permute(0, 3, 1, 2).contiguous()

    # ------------------------------------------------------------------------------------------------------------
    def forward(self, x, timesteps=None, context=None, y=None, **kwargs):
        """eps prediction, openaimodel_enc.py:712-748.  ``context`` is a sequence; only ``context[0]`` is used,
        reshaped to (B, -1, context_dim) (:729).  Extra kwargs (``return_context``) are ignored as in the reference."""
        assert (y is not None) == (self.num_classes is not None), \
            "must specify y if and only if the model is class-conditional"
        if not x.is_cuda:
            raise RuntimeError("encdiff_b200.UNetModel.forward needs CUDA tensors (no CPU fallback)")
        b = x.shape[0]
        if x.shape[1] != self.in_channels or x.shape[2] != self.image_size or x.shape[3] != self.image_size:
            raise RuntimeError(f"encdiff_b200.UNetModel: expected input (B,{self.in_channels},{self.image_size},"
                               f"{self.image_size}), got {tuple(x.shape)}")
        cond = context[0].reshape(b, -1, self.context_dim)
        if cond.shape[1] != self.latent_unit:
            raise RuntimeError(f"encdiff_b200.UNetModel: context has {cond.shape[1]} tokens, latent_unit is {self.latent_unit}")
        cond = cond.reshape(b, -1)
        cond_f = cond.contiguous().float()
        xf = x.contiguous().float()
        tf = timesteps.to(device=x.device, dtype=torch.float32).contiguous()
        needs_grad = torch.is_grad_enabled() and (
            xf.requires_grad or cond_f.requires_grad or any(p.requires_grad for p in self.parameters()))
        if needs_grad:
            out = _UNetFn.apply(self, xf, tf, cond_f, *self._params())
        else:
            out = self._run_forward(xf, tf, cond_f, training=False)
        return out.type(x.dtype)
