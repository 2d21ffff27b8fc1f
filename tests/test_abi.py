"""CPU checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/encdiff_b200.h declares,
the ctypes prototype table covers exactly those symbols, and the product path refuses to run without a GPU (no CPU
fallback).  No compute call is made here."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "encdiff_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(encdiff_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_plain_c_abi():
    text = open(os.path.join(ROOT, "include", "encdiff_b200.h")).read()
    assert 'extern "C"' in text
    assert "torch" not in re.sub(r"/\*.*?\*/", "", text, flags=re.S).lower()
    assert "at::" not in text and "std::" not in text
    assert len(header_symbols()) >= 30


def test_library_exports_every_header_symbol():
    from encdiff_b200 import _lib
    path = _lib.lib_path()
    assert path.exists(), f"{path} missing: run __graft_entry__.build()"
    lib = ctypes.CDLL(str(path))
    missing = [s for s in header_symbols() if not hasattr(lib, s)]
    assert not missing, missing


def test_prototype_table_matches_header():
    from encdiff_b200 import _lib
    assert sorted(_lib.PROTOTYPES) == header_symbols()


def test_act_format_and_error_string():
    from encdiff_b200 import _lib
    lib = _lib.load(build_if_missing=False)
    assert lib.encdiff_act_format() in (0, 1)
    assert isinstance(lib.encdiff_last_error(), bytes)
    assert _lib.act_dtype() in (torch.float16, torch.bfloat16)


def test_handle_inventory_without_gpu():
    """create / param_info / destroy are host-only: the parameter inventory of the library equals the oracle's."""
    from encdiff_b200 import UNetModel
    from oracle import unet_oracle as UO
    cfg = UO.SHAPES3D
    m = UNetModel(image_size=cfg.image_size, in_channels=cfg.in_channels, out_channels=cfg.out_channels,
                  model_channels=cfg.model_channels, attention_resolutions=list(cfg.attention_resolutions),
                  num_res_blocks=cfg.num_res_blocks, channel_mult=list(cfg.channel_mult), num_heads=cfg.num_heads,
                  use_scale_shift_norm=True, resblock_updown=True, use_spatial_transformer=True,
                  context_dim=cfg.context_dim, latent_unit=cfg.latent_unit)
    specs = UO.param_specs(cfg)
    got = [(n, tuple(p.shape)) for n, p in m.named_parameters()]
    assert got == [(n, tuple(s)) for n, s in specs]
    assert sum(p.numel() for p in m.parameters()) == 37469635
    assert list(m.state_dict().keys()) == [n for n, _ in specs]
    assert m.latent_unit == cfg.latent_unit and m.context_dim == cfg.context_dim


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback():
    from encdiff_b200 import CrossAttention, SpatialTransformer, UNetModel
    from oracle import unet_oracle as UO
    from oracle.make_golden import TINY
    cfg = TINY
    m = UNetModel(image_size=cfg.image_size, in_channels=cfg.in_channels, out_channels=cfg.out_channels,
                  model_channels=cfg.model_channels, attention_resolutions=list(cfg.attention_resolutions),
                  num_res_blocks=cfg.num_res_blocks, channel_mult=list(cfg.channel_mult), num_heads=cfg.num_heads,
                  use_scale_shift_norm=True, resblock_updown=True, use_spatial_transformer=True,
                  context_dim=cfg.context_dim, latent_unit=cfg.latent_unit)
    x = torch.zeros(1, cfg.in_channels, cfg.image_size, cfg.image_size)
    with pytest.raises(RuntimeError, match="CUDA"):
        m(x, torch.zeros(1, dtype=torch.long), context=[torch.zeros(1, cfg.latent_unit * cfg.context_dim)])
    with pytest.raises(RuntimeError, match="CUDA"):
        with torch.no_grad():
            CrossAttention(64, None, 8, 8)(torch.zeros(1, 16, 64))
    with pytest.raises(RuntimeError, match="CUDA"):
        with torch.no_grad():
            SpatialTransformer(64, 8, 8, context_dim=16)(torch.zeros(1, 64, 4, 4), torch.zeros(1, 20, 16))


def test_unsupported_configurations_raise():
    from encdiff_b200 import UNetModel
    base = dict(image_size=16, in_channels=3, out_channels=3, model_channels=64, attention_resolutions=[1, 2, 4],
                num_res_blocks=2, channel_mult=[1, 2, 4, 4], num_heads=8, use_scale_shift_norm=True,
                resblock_updown=True, use_spatial_transformer=True, context_dim=16, latent_unit=20)
    for bad in (dict(use_spatial_transformer=False), dict(dropout=0.1), dict(dims=3), dict(num_classes=10),
                dict(use_fp16=True), dict(use_scale_shift_norm=False), dict(n_embed=8)):
        with pytest.raises((NotImplementedError, AssertionError)):
            UNetModel(**{**base, **bad})


def test_product_never_imports_oracle():
    """oracle/ is test infrastructure: nothing under encdiff_b200/ may import it."""
    pkg = os.path.join(ROOT, "encdiff_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inc", ".h")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), os.path.join(dirpath, f)


def test_build_dependency_tracking_sees_inc_files():
    """The in-tree build recompiles a translation unit when anything it #includes changes (transitively): the launch
    sequences live in .inc files included by unet_run.cu, and every API file sees the public header."""
    from encdiff_b200 import _build
    names = {p.name for p in _build._includes(_build.CSRC / "unet_run.cu")}
    assert {"unet_bwd.inc", "vq_enc.inc", "enc4.inc", "kernels.cuh", "common.cuh", "unet.cuh"} <= names
    assert "encdiff_b200.h" in {p.name for p in _build._includes(_build.CSRC / "api_unet.cu")}
    assert "enc4.inc" not in {p.name for p in _build._includes(_build.CSRC / "gemm2.cu")}
