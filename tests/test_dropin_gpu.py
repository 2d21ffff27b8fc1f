"""The drop-in boundary exercised through the REFERENCE'S OWN seam: the real ``LatentDiffusion``
(ldm/models/diffusion/ddpm_enc.py:484) is built by the reference's ``instantiate_from_config`` (ldm/util.py:78-93) from
the Shapes3D yaml values with ``unet_config.target: encdiff_b200.UNetModel``, and driven through the reference's own
``forward`` / ``p_losses`` / ``sample_log`` / ``log_images`` code.  The comparator is the same ``LatentDiffusion`` with
the reference's ``UNetModel`` in fp32 eager mode on the same GPU.  The reference package is imported from /root/reference
(build container) or oracle/_ref (GPU box, oracle/make_ref.py); those tests are skipped when neither exists."""
import os

import numpy as np
import pytest
import torch

from oracle import ref_shim
from oracle import unet_oracle as UO
from oracle.make_golden import CONFIGS

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(not ref_shim.available(), reason="no reference tree (/root/reference or oracle/_ref)")
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def psnr(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    peak = (b.max() - b.min()).item()
    mse = ((a - b) ** 2).mean().item()
    return float("inf") if mse == 0 else 10.0 * np.log10(peak * peak / mse)


@pytest.fixture(scope="module")
def pair(cuda_device):
    """(reference LatentDiffusion, the same with encdiff_b200.UNetModel as unet_config.target), identical weights."""
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    cfg = CONFIGS["shapes3d"]
    params = UO.make_params(cfg, seed=1)
    ref = ref_shim.reference_latent_diffusion(cfg, device=cuda_device, unet_params=params)
    mine = ref_shim.reference_latent_diffusion(cfg, unet_target="encdiff_b200.UNetModel", device=cuda_device,
                                               unet_params=params)
    # everything outside the denoiser (first stage, Encoder4, EMA shadow) comes from the reference-side model through
    # the plain state_dict: this is also the check that the keys under `model.diffusion_model.` are the reference's
    missing, unexpected = mine.load_state_dict(ref.state_dict(), strict=True)
    assert not missing and not unexpected
    return cfg, ref, mine


@needs_ref
def test_instantiate_from_config_builds_the_dropin(pair):
    import encdiff_b200
    cfg, ref, mine = pair
    assert type(mine.model.diffusion_model) is encdiff_b200.UNetModel
    assert type(mine).__name__ == "LatentDiffusion" and type(mine.model).__name__ == "DiffusionWrapper"
    assert list(mine.state_dict().keys()) == list(ref.state_dict().keys())
    assert mine.model.diffusion_model.latent_unit == cfg.latent_unit          # read by log_images, ddpm_enc.py:1525
    assert len(list(mine.model_ema.buffers())) == len(list(ref.model_ema.buffers()))      # LitEma snapshot, ema.py:16-21


@needs_ref
def test_instantiate_from_config_p_losses_forward_backward(pair, cuda_device):
    """LatentDiffusion.forward -> Encoder4 -> p_losses -> apply_model -> DiffusionWrapper -> UNet (ddpm_enc.py:1040-1213)
    and loss.backward(): loss, the gradient reaching the trainable concept encoder through d context, and the UNet's own
    parameter gradients, against the reference UNet in fp32 on the same GPU."""
    cfg, ref, mine = pair
    g = torch.Generator().manual_seed(23)
    batch = {"image": (torch.rand(8, 64, 64, 3, generator=g) * 2 - 1).to(cuda_device)}
    out = {}
    for name, model in (("ref", ref), ("mine", mine)):
        model.train()
        model.zero_grad(set_to_none=True)
        x, c = model.get_input(batch, "image")
        torch.manual_seed(77)                       # t = randint(...), noise = randn_like(...) inside forward / p_losses
        loss, loss_dict = model(x, c)
        loss.backward()
        out[name] = (loss.item(), loss_dict,
                     {k: p.grad.detach().clone() for k, p in model.cond_stage_model.named_parameters() if p.grad is not None},
                     {k: p.grad.detach().clone() for k, p in model.model.diffusion_model.named_parameters()})
    l_ref, l_mine = out["ref"][0], out["mine"][0]
    print(f"LatentDiffusion.forward loss: reference {l_ref:.6f}, drop-in {l_mine:.6f}")
    assert abs(l_ref - l_mine) <= 2e-3 * abs(l_ref)
    assert set(out["ref"][1]) == set(out["mine"][1])

    # L1's gradient is sign(eps - noise): a few signs flip between two precisions, so the bars are on direction
    def cos(da, db):
        a = torch.cat([da[k].flatten().double() for k in db])
        b = torch.cat([db[k].flatten().double() for k in db])
        return (torch.dot(a, b) / (a.norm() * b.norm())).item()
    assert set(out["mine"][2]) == set(out["ref"][2]) and len(out["ref"][2]) > 0
    c_enc, c_unet = cos(out["mine"][2], out["ref"][2]), cos(out["mine"][3], out["ref"][3])
    print(f"gradient cosine: concept encoder {c_enc:.5f}, UNet {c_unet:.5f}")
    assert c_enc > 0.99 and c_unet > 0.99


@needs_ref
def test_sample_log_after_install(pair, cuda_device):
    """``sampler.install()`` rebinds the name LatentDiffusion.sample_log resolves (ddpm_enc.py:1442-1455); the reference's
    own sample_log then runs the graph-replayed CUDA sampler.  Same x_T, eta 0: PSNR >= 40 dB against the reference
    DDIMSampler around the reference UNet."""
    import ldm.models.diffusion.ddpm_enc as ddpm_enc
    from encdiff_b200 import sampler as edb_sampler
    cfg, ref, mine = pair
    g = torch.Generator().manual_seed(3)
    B, S = 4, 20
    cond = torch.randn(B, cfg.latent_unit * cfg.context_dim, generator=g).to(cuda_device)
    x_T = torch.randn(B, 3, cfg.image_size, cfg.image_size, generator=g).to(cuda_device)
    ref.eval(), mine.eval()
    want, _ = ref.sample_log(cond=cond, batch_size=B, ddim=True, ddim_steps=S, eta=0., x_T=x_T)
    prev = edb_sampler.install(ddpm_enc)
    try:
        got, inter = mine.sample_log(cond=cond, batch_size=B, ddim=True, ddim_steps=S, eta=0., x_T=x_T)
    finally:
        ddpm_enc.DDIMSampler = prev
    p = psnr(got, want)
    print(f"sample_log (DDIM-{S}, installed sampler) vs reference: PSNR {p:.1f} dB")
    assert got.shape == want.shape and set(inter) == {"x_inter", "pred_x0"}
    assert p >= 40.0


def test_swap_vs_reference_log_images_golden(cuda_device):
    """SURVEY.md section 8 row a17 pinned on the reference side: tests/golden/swap_shapes3d_n2_s10.npz holds what the
    UNMODIFIED reference's ``log_images(sample_swap=True, ddim_steps=10, ddim_eta=0)`` produced
    (oracle/make_swap_golden.py): codes, Encoder4.net weights, the start noise of its 20 sequential runs and the 40
    latents it decoded.  The folded CUDA run (one batch of 20 x N samples) must reproduce those latents."""
    from encdiff_b200 import ConceptMLP, DDIMSampler, UNetModel
    from encdiff_b200.diffusion import DenoiserHost
    from encdiff_b200.swap import sample_swap
    fx = np.load(os.path.join(GOLDEN, "swap_shapes3d_n2_s10.npz"))
    cfg = CONFIGS["shapes3d"]
    m = UNetModel(image_size=cfg.image_size, in_channels=cfg.in_channels, out_channels=cfg.out_channels,
                  model_channels=cfg.model_channels, attention_resolutions=list(cfg.attention_resolutions),
                  num_res_blocks=cfg.num_res_blocks, channel_mult=list(cfg.channel_mult), num_heads=cfg.num_heads,
                  use_scale_shift_norm=True, resblock_updown=True, use_spatial_transformer=True,
                  context_dim=cfg.context_dim, latent_unit=cfg.latent_unit)
    m.load_state_dict(UO.make_params(cfg, seed=1), strict=True)
    host = DenoiserHost(m.eval()).to(cuda_device)
    net = ConceptMLP(cfg.latent_unit, cfg.context_dim)
    net.load_state_dict({k[len("net::"):]: torch.tensor(fx[k]) for k in fx.files if k.startswith("net::")}, strict=True)
    net = net.to(cuda_device)
    S, N = int(fx["S"]), int(fx["N"])
    orc = torch.tensor(fx["orc"]).to(cuda_device)
    x_T = torch.tensor(fx["x_T"]).to(cuda_device)                    # [20, N, 3, 16, 16], run-major
    got = sample_swap(DDIMSampler(host), orc, net.warp, ddim_steps=S, eta=0., shape=(3, cfg.image_size, cfg.image_size),
                      x_T=x_T)
    want = torch.tensor(fx["samples"])
    assert got.shape == want.shape == (cfg.latent_unit * N, 3, cfg.image_size, cfg.image_size)
    p = psnr(got, want)
    per_run = min(psnr(got[i * N:(i + 1) * N], want[i * N:(i + 1) * N]) for i in range(cfg.latent_unit))
    print(f"swap generation vs the reference's log_images: PSNR {p:.1f} dB (worst run {per_run:.1f} dB)")
    assert p >= 40.0 and per_run >= 40.0


@needs_ref
def test_first_stage_install_encode_decode(pair, cuda_device):
    """encdiff_b200.first_stage.install() on the reference's OWN first_stage_model (the real VQModelInterface built by
    instantiate_from_config; its quantiser is ref_shim's restatement of the absent taming module): encode and decode through
    the rebound methods against the reference modules in fp32 on the same GPU."""
    from encdiff_b200 import first_stage
    cfg, ref, mine = pair
    fs_ref, fs = ref.first_stage_model, mine.first_stage_model
    g = torch.Generator().manual_seed(41)
    x = (torch.rand(6, 3, 64, 64, generator=g) * 2 - 1).to(cuda_device)
    s = torch.randn(6, cfg.latent_unit, generator=g).to(cuda_device)
    with torch.no_grad():
        z_ref = fs_ref.encode(x)
        img_ref = fs_ref.decode(z_ref, disentangled_repr=s)
        img_ref_nq = fs_ref.decode(z_ref, force_not_quantize=True)
    enc, dec = first_stage.install(fs)
    z = fs.encode(x)
    img = fs.decode(z_ref, disentangled_repr=s)
    img_nq = fs.decode(z_ref, force_not_quantize=True)
    rel = lambda a, b: ((a - b).abs().max() / b.abs().max()).item()
    print(f"first stage through the reference seam: encode {rel(z, z_ref):.3e}, decode {rel(img, img_ref):.3e}, "
          f"decode w/o quantisation {rel(img_nq, img_ref_nq):.3e}")
    assert type(enc).__name__ == "VQEncoder" and type(dec).__name__ == "VQDecoder"
    assert rel(z, z_ref) <= 1e-2 and rel(img, img_ref) <= 1e-2 and rel(img_nq, img_ref_nq) <= 1e-2


@needs_ref
def test_ddim_decoded_images_vs_reference(pair, cuda_device):
    """DDIM-50 (eta 0, shared x_T) through this package end to end -- CUDA sampler around the drop-in
    UNet, then the CUDA first-stage decoder -- against the reference sampler, UNet and decoder: PSNR of the DECODED images.
    The bar is the north star's 40 dB, with and without the quantiser inside ``decode`` (measured: latent 86 dB, decoded
    77 dB without and 71 dB with quantisation, 99.9 % of the nearest-code indices equal)."""
    import ldm.models.diffusion.ddpm_enc as ddpm_enc
    from encdiff_b200 import first_stage
    from encdiff_b200 import sampler as edb_sampler
    cfg, ref, mine = pair
    g = torch.Generator().manual_seed(13)
    B, S = 8, 50
    cond = torch.randn(B, cfg.latent_unit * cfg.context_dim, generator=g).to(cuda_device)
    x_T = torch.randn(B, 3, cfg.image_size, cfg.image_size, generator=g).to(cuda_device)
    s_dis = torch.randn(B, cfg.latent_unit, generator=g).to(cuda_device)
    ref.eval(), mine.eval()
    with torch.no_grad():
        z_ref, _ = ref.sample_log(cond=cond, batch_size=B, ddim=True, ddim_steps=S, eta=0., x_T=x_T)
        img_ref_nq = ref.decode_first_stage(z_ref, force_not_quantize=True, disentangled_repr=s_dis)
        img_ref_q = ref.decode_first_stage(z_ref, disentangled_repr=s_dis)
        prev = edb_sampler.install(ddpm_enc)
        saved = mine.first_stage_model.encode, mine.first_stage_model.decode
        try:
            first_stage.install(mine.first_stage_model)
            z, _ = mine.sample_log(cond=cond, batch_size=B, ddim=True, ddim_steps=S, eta=0., x_T=x_T)
            img_nq = mine.decode_first_stage(z, force_not_quantize=True, disentangled_repr=s_dis)
            img_q = mine.decode_first_stage(z, disentangled_repr=s_dis)
        finally:
            ddpm_enc.DDIMSampler = prev
            mine.first_stage_model.encode, mine.first_stage_model.decode = saved
        book = ref.first_stage_model.quantize.embedding.weight.double()

        def codes(lat):                      # nearest-code index per position (what the quantiser inside decode() picks)
            zf = (lat.double() / ref.scale_factor).permute(0, 2, 3, 1).reshape(-1, book.shape[1])
            return ((zf ** 2).sum(1, keepdim=True) + (book ** 2).sum(1)[None] - 2 * zf @ book.t()).argmin(1)
        same = (codes(z_ref) == codes(z)).float().mean().item()
    p_lat, p_nq, p_q = psnr(z, z_ref), psnr(img_nq, img_ref_nq), psnr(img_q, img_ref_q)
    print(f"DDIM-{S} end to end vs reference: latent PSNR {p_lat:.1f} dB; decoded (no quantiser) {p_nq:.1f} dB; decoded "
          f"(quantised) {p_q:.1f} dB with {100 * same:.1f} % of the code indices equal")
    assert p_lat >= 40.0 and p_nq >= 40.0 and p_q >= 40.0 and same >= 0.99


@needs_ref
def test_all_targets_swapped_training_step_through_the_reference_seam(pair, cuda_device):
    """``unet_config.target: encdiff_b200.UNetModel``, ``cond_stage_config.target: encdiff_b200.Encoder4`` AND
    ``first_stage_config.target: encdiff_b200.VQModelInterface`` in the reference's own ``instantiate_from_config``; the
    reference's ``LatentDiffusion.get_input`` / ``forward`` / ``loss.backward()`` then run the frozen VQ-f4 encode, the
    concept encoder's trunk forward / backward and the UNet through this package's kernels.  State dict keys and order are
    the reference's; loss and gradient directions against the all-reference model in fp32 on the same GPU."""
    import encdiff_b200
    cfg, ref, _ = pair
    both = ref_shim.reference_latent_diffusion(cfg, unet_target="encdiff_b200.UNetModel", device=cuda_device,
                                               unet_params=UO.make_params(cfg, seed=1), cond_target="encdiff_b200.Encoder4",
                                               first_target="encdiff_b200.VQModelInterface")
    assert type(both.cond_stage_model) is encdiff_b200.Encoder4 and type(both.model.diffusion_model) is encdiff_b200.UNetModel
    assert type(both.first_stage_model) is encdiff_b200.VQModelInterface
    assert list(both.state_dict().keys()) == list(ref.state_dict().keys())
    missing, unexpected = both.load_state_dict(ref.state_dict(), strict=True)
    assert not missing and not unexpected
    g = torch.Generator().manual_seed(29)
    batch = {"image": (torch.rand(16, 64, 64, 3, generator=g) * 2 - 1).to(cuda_device)}
    out = {}
    for name, model in (("ref", ref), ("both", both)):
        model.train()
        model.zero_grad(set_to_none=True)
        x, c = model.get_input(batch, "image")
        torch.manual_seed(78)
        loss, _ = model(x, c)
        loss.backward()
        enc_grads = {k: p.grad.detach().clone() for k, p in model.cond_stage_model.named_parameters() if p.grad is not None}
        if name == "both":          # ConceptMLP holds the 20 MLPs stacked: back to the reference's per-unit names
            from encdiff_b200.concept import ConceptMLP
            for mine_name, ref_name in ConceptMLP._MAP:
                stacked = enc_grads.pop("net." + mine_name)
                for i in range(cfg.latent_unit):
                    gi = stacked[i]
                    enc_grads[f"net.{i}.{ref_name}"] = gi.reshape(-1, 1) if mine_name == "w1" else gi
        out[name] = (loss.item(), enc_grads,
                     {k: p.grad.detach().clone() for k, p in model.model.diffusion_model.named_parameters()})

    def cos(da, db):
        a = torch.cat([da[k].flatten().double() for k in db])
        b = torch.cat([db[k].flatten().double() for k in db])
        return (torch.dot(a, b) / (a.norm() * b.norm())).item()
    l_ref, l_both = out["ref"][0], out["both"][0]
    assert set(out["both"][1]) == set(out["ref"][1])
    c_enc, c_unet = cos(out["both"][1], out["ref"][1]), cos(out["both"][2], out["ref"][2])
    print(f"all three targets swapped: loss {l_both:.6f} vs reference {l_ref:.6f}; gradient cosine concept encoder {c_enc:.5f}, "
          f"UNet {c_unet:.5f}")
    assert abs(l_ref - l_both) <= 2e-3 * abs(l_ref)
    assert c_enc > 0.99 and c_unet > 0.99
    ref.zero_grad(set_to_none=True)
