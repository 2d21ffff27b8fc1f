"""ctypes binding of libencdiff_b200.so (include/encdiff_b200.h).

There is no CPU fallback: if the library cannot be loaded, every entry point raises.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

from . import _build

_lib = None

c_void_p, c_int, c_ll, c_float_p = C.c_void_p, C.c_int, C.c_longlong, C.c_void_p

# name -> (restype, argtypes).  Kept in sync with include/encdiff_b200.h; tests/test_abi.py checks that every
# symbol the header declares is exported and listed here.
PROTOTYPES = {
    "encdiff_last_error": (C.c_char_p, []),
    "encdiff_act_format": (c_int, []),
    "encdiff_op_sgemm": (c_int, [c_int, C.POINTER(c_void_p), C.POINTER(c_int), C.POINTER(c_int), c_int, c_int, c_int,
                                 c_void_p, c_ll, c_int, c_int, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p,
                                 c_int, c_int, c_void_p, c_int, c_int, c_void_p]),
    "encdiff_op_wgrad": (c_int, [c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_ll, c_ll,
                                 c_ll, c_int, c_int, c_int, c_void_p]),
    "encdiff_op_group_norm": (c_int, [c_void_p, c_void_p, c_void_p, C.c_float, c_int, c_void_p, c_void_p, c_int, c_int,
                                      c_int, c_int, c_void_p]),
    "encdiff_op_layer_norm": (c_int, [c_void_p, c_void_p, c_void_p, C.c_float, c_void_p, c_void_p, c_ll, c_int,
                                      c_void_p]),
    "encdiff_op_self_attention": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "encdiff_op_self_attention_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                              c_int, c_void_p]),
    "encdiff_op_cross_kv_project": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                            c_void_p]),
    "encdiff_op_cross_attention": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                           c_int, c_int, c_int, c_int, c_void_p]),
    "encdiff_op_cross_attention_bwd_scratch_bytes": (c_ll, [c_int, c_int, c_int, c_int]),
    "encdiff_op_cross_attention_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                               c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                               c_int, c_int, c_void_p]),
    "encdiff_op_cross_attention_bwd_noalloc": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                                       c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                                       c_int, c_int, c_void_p]),
    "encdiff_op_geglu": (c_int, [c_void_p, c_void_p, c_ll, c_int, c_void_p]),
    "encdiff_concept_mlp_forward": (c_int, [c_void_p] * 8 + [c_int, c_int, c_int, c_void_p]),
    "encdiff_concept_mlp_backward": (c_int, [c_void_p] * 15 + [c_int, c_int, c_int, c_void_p]),
    "encdiff_unet_create": (c_int, [c_void_p, C.POINTER(c_void_p)]),
    "encdiff_unet_destroy": (c_int, [c_void_p]),
    "encdiff_unet_num_params": (c_int, [c_void_p]),
    "encdiff_unet_param_info": (c_int, [c_void_p, c_int, C.c_char_p, c_int, C.POINTER(c_ll), C.POINTER(c_int)]),
    "encdiff_unet_packed_bytes": (c_ll, [c_void_p]),
    "encdiff_unet_workspace_bytes": (c_ll, [c_void_p, c_int, c_int]),
    "encdiff_unet_bind": (c_int, [c_void_p, C.POINTER(c_void_p), C.POINTER(c_void_p), c_void_p, c_void_p]),
    "encdiff_unet_pack_weights": (c_int, [c_void_p, c_void_p]),
    "encdiff_unet_plan": (c_int, [c_void_p, c_int, c_void_p, c_int]),
    "encdiff_unet_forward": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                     c_void_p, c_void_p, c_void_p]),
    "encdiff_unet_backward": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "encdiff_unet_set_grad_split": (c_int, [c_void_p, c_void_p]),
    "encdiff_unet_backward_tail": (c_int, [c_void_p, c_void_p]),
    "encdiff_vqenc_create": (c_int, [c_void_p, C.POINTER(c_void_p)]),
    "encdiff_vqenc_workspace_bytes": (c_ll, [c_void_p, c_int]),
    "encdiff_vqenc_forward": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "encdiff_vqdec_create": (c_int, [c_void_p, C.POINTER(c_void_p)]),
    "encdiff_vqdec_workspace_bytes": (c_ll, [c_void_p, c_int]),
    "encdiff_vqdec_forward": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "encdiff_enc4_create": (c_int, [c_void_p, C.POINTER(c_void_p)]),
    "encdiff_enc4_num_batchnorm": (c_int, [c_void_p]),
    "encdiff_enc4_bind_stats": (c_int, [c_void_p, c_void_p]),
    "encdiff_enc4_workspace_bytes": (c_ll, [c_void_p, c_int]),
    "encdiff_enc4_forward": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, C.c_float, c_void_p]),
    "encdiff_enc4_backward": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "encdiff_unet_fp32_workspace_bytes": (c_ll, [c_void_p, c_int]),
    "encdiff_unet_forward_fp32": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_ll, c_void_p]),
    "encdiff_unet_profile": (c_int, [c_void_p, c_int]),
    "encdiff_unet_profile_read": (c_ll, [c_void_p, C.c_char_p, c_ll]),
    "encdiff_unet_debug_block": (c_int, [c_void_p, c_int, c_int, C.POINTER(c_void_p), C.POINTER(c_int),
                                         C.POINTER(c_int), C.POINTER(c_int)]),
    "encdiff_l1_loss": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_ll, c_void_p]),
    "encdiff_ddim_update": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, C.c_float, c_void_p, c_void_p, c_ll,
                                    c_void_p]),
    "encdiff_ddim_select": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "encdiff_ddim_update_indexed": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, C.c_float, c_void_p,
                                            c_void_p, c_ll, c_void_p]),
    "encdiff_timestep_embedding": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "encdiff_adamw_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_ll, C.c_float, C.c_float,
                                   C.c_float, C.c_float, C.c_float, c_int, C.c_float, c_void_p]),
    "encdiff_adamw_step_dev": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_ll, c_void_p, C.c_float,
                                       C.c_float, C.c_float, C.c_float, C.c_float, c_void_p]),
    "encdiff_adamw_hyper_advance": (c_int, [c_void_p, c_void_p, c_void_p, C.c_float, C.c_float, C.c_float, c_void_p]),
}


class UNetConfigStruct(C.Structure):
    """Mirror of ``encdiff_unet_config`` (include/encdiff_b200.h)."""
    _fields_ = [("image_size", c_int), ("in_channels", c_int), ("model_channels", c_int), ("out_channels", c_int),
                ("num_res_blocks", c_int), ("n_attention_resolutions", c_int), ("attention_resolutions", c_int * 8),
                ("n_channel_mult", c_int), ("channel_mult", c_int * 8), ("num_heads", c_int), ("context_dim", c_int),
                ("latent_unit", c_int), ("transformer_depth", c_int), ("use_scale_shift_norm", c_int),
                ("resblock_updown", c_int)]


class VQEncConfigStruct(C.Structure):
    """Mirror of ``encdiff_vqenc_config`` (include/encdiff_b200.h)."""
    _fields_ = [("ch", c_int), ("n_ch_mult", c_int), ("ch_mult", c_int * 4), ("num_res_blocks", c_int),
                ("in_channels", c_int), ("resolution", c_int), ("z_channels", c_int), ("embed_dim", c_int)]


class VQDecConfigStruct(C.Structure):
    """Mirror of ``encdiff_vqdec_config`` (include/encdiff_b200.h)."""
    _fields_ = [("dd", VQEncConfigStruct), ("out_ch", c_int), ("n_embed", c_int), ("disentangled_dim", c_int)]


class Enc4ConfigStruct(C.Structure):
    """Mirror of ``encdiff_enc4_config`` (include/encdiff_b200.h)."""
    _fields_ = [("d", c_int), ("latent_unit", c_int), ("num_channels", c_int), ("resolution", c_int), ("bn", c_int)]


def lib_path() -> Path:
    """The in-tree library; ENCDIFF_B200_LIB points at another build of the same ABI (A/B timing of two builds on one box)."""
    alt = os.environ.get("ENCDIFF_B200_LIB")
    return Path(alt) if alt else _build.LIB_PATH


def load(build_if_missing: bool = True):
    """Load (building first if the sources changed and nvcc is available) and return the ctypes library."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if build_if_missing and os.environ.get("ENCDIFF_B200_NO_BUILD") != "1":
        try:
            _build.build()
        except Exception as e:  # a stale-but-present library on a box without nvcc is still usable
            if not path.exists():
                raise RuntimeError(f"encdiff_b200: cannot build the CUDA library and none is present: {e}") from e
    if not path.exists():
        raise RuntimeError(
            f"encdiff_b200: {path} is missing. Run `python -c 'import __graft_entry__ as g; g.build()'`; "
            "this package has no CPU fallback.")
    lib = C.CDLL(str(path))
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported: loud by design
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def act_dtype():
    """torch dtype of the library's 16-bit storage type (fp16 by default, bf16 in the -DEDB_ACT_BF16 build)."""
    import torch
    return torch.bfloat16 if load().encdiff_act_format() == 1 else torch.float16


def check(status: int, what: str = "") -> None:
    if status != 0:
        msg = load().encdiff_last_error().decode(errors="replace")
        raise RuntimeError(f"encdiff_b200 {what} failed (status {status}): {msg}")


def ptr(t) -> int:
    """Device pointer of a torch tensor (None -> NULL)."""
    return 0 if t is None else t.data_ptr()


def current_stream() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream


class NativeStateMixin:
    """``copy.deepcopy`` / ``pickle`` for modules that own a library handle.  Parameters, buffers and plain attributes travel
    as usual; the handle, the ctypes library object, device scratch and captured graphs (``_NATIVE_KEYS`` of the class) are
    per-instance and are rebuilt on the other side by constructing a fresh instance from the recorded constructor keywords
    and adopting ITS native members."""
    _NATIVE_KEYS = ()

    def _record_ctor(self, kw):
        kw = {k: v for k, v in kw.items() if k not in ("self", "__class__")}
        object.__setattr__(self, "_ctor_kw", kw)

    def __getstate__(self):
        d = self.__dict__.copy()
        for k in self._NATIVE_KEYS:
            d.pop(k, None)
        return d

    def __setstate__(self, d):
        super().__setstate__(d)
        fresh = type(self)(**self._ctor_kw)
        for k in self._NATIVE_KEYS:
            if k in fresh.__dict__:
                self.__dict__[k] = fresh.__dict__[k]
        fresh.__dict__["_h"] = None          # the handle now belongs to this instance
