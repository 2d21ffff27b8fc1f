"""copy.deepcopy / pickle of the modules that own a library handle (what a training framework does when it clones a model
for an EMA copy or ships it to a spawned worker): parameters travel, the native handle is rebuilt per instance, and shared
parameter holders (VQModelInterface and its two CUDA paths) stay shared.  No compute: runs without a GPU."""
import copy
import pickle

import torch

from oracle.make_golden import CONFIGS

VQ_DD = dict(double_z=False, z_channels=3, resolution=64, in_channels=3, out_ch=3, ch=32, ch_mult=[1, 2, 4], num_res_blocks=2,
             attn_resolutions=[], dropout=0.0)


def same_state(a, b):
    sa, sb = a.state_dict(), b.state_dict()
    return list(sa) == list(sb) and all(torch.equal(sa[k], sb[k]) for k in sa)


def test_unet_deepcopy_and_pickle():
    from encdiff_b200 import UNetModel
    cfg = CONFIGS["tiny"]
    m = UNetModel(image_size=cfg.image_size, in_channels=cfg.in_channels, out_channels=cfg.out_channels,
                  model_channels=cfg.model_channels, attention_resolutions=list(cfg.attention_resolutions),
                  num_res_blocks=cfg.num_res_blocks, channel_mult=list(cfg.channel_mult), num_heads=cfg.num_heads,
                  use_scale_shift_norm=True, resblock_updown=True, use_spatial_transformer=True,
                  context_dim=cfg.context_dim, latent_unit=cfg.latent_unit)
    for clone in (copy.deepcopy(m), pickle.loads(pickle.dumps(m))):
        assert same_state(m, clone) and clone._h.value and clone._h.value != m._h.value
        assert clone.latent_unit == m.latent_unit and clone._bound_key is None


def test_encoder4_and_first_stage_deepcopy():
    from encdiff_b200 import Encoder4, VQModelInterface
    e = Encoder4(d=128, context_dim=16, latent_unit=20)
    e2 = copy.deepcopy(e)
    assert same_state(e, e2) and e2._h.value != e._h.value
    v = VQModelInterface(embed_dim=3, n_embed=64, ddconfig=VQ_DD, use_disentangled_concat=True, disentangled_dim=20)
    v2 = pickle.loads(pickle.dumps(v))
    assert same_state(v, v2)
    assert v2._enc.encoder is v2.encoder and v2._dec.decoder is v2.decoder and v2._dec.quantize is v2.quantize
    assert v2._enc._h.value != v._enc._h.value
