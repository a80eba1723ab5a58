"""-m gpu: END-TO-END parity of the image tower + projector at the sizes the numbers are quoted on (VERDICT r1 #1b).

The oracle (oracle/vit_oracle.py — plain torch ops, bit-identical to HF CLIP/SigLIP + the reference MLPProjector on
the same device/dtype, pinned in tests/test_oracle_golden.py) runs ON THE GPU here: once in true fp32 (math SDPA, TF32
off) as the ground truth and once in pure bf16 (regime (i), what the reference itself executes on a GPU) to measure
the reference's own rounding error.  Stated tolerance:
    max|B200 - fp32| <= 2 x max|ref_bf16 - fp32|     and     max|B200 - fp32| / max|fp32| <= 2e-2.
Cases: cfg2 tower (CLIP ViT-L/14-336 -> 4096) on 8 images, cfg3 tower (SigLIP so400m/14-384 -> 3584, 27 layers, hd 72,
I = 4304, post-LN, streamed K/V ring) on 2 images, each also with planted massive-activation channels.
"""
import pytest
import torch

from multimeditron_b200 import synth
from multimeditron_b200.configs import TOWERS
from multimeditron_b200.modalities import AutoModality
from oracle import vit_oracle

pytestmark = pytest.mark.gpu
DEV = "cuda"

CASES = [  # tower, llm hidden, images, massive activations
    ("clip_l14_336", 4096, 8, False),
    ("clip_l14_336", 4096, 8, True),
    ("siglip_so400m_384", 3584, 2, False),
    ("siglip_so400m_384", 3584, 2, True),
]


@torch.no_grad()
def _oracle_gpu(sd, cfg, px, dtype):
    from torch.nn.attention import SDPBackend, sdpa_kernel
    sdd = {k: v.to(DEV, dtype) for k, v in sd.items()}
    tf32 = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    try:
        if dtype == torch.float32:
            with sdpa_kernel(SDPBackend.MATH):
                return vit_oracle.image_modality_forward(sdd, cfg, px.to(DEV)).float()
        return vit_oracle.image_modality_forward(sdd, cfg, px.to(DEV)).float()     # bf16: the library's default SDPA
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = tf32


@pytest.mark.parametrize("name,hidden,batch,massive", CASES)
def test_tower_and_projector_at_quoted_size(name, hidden, batch, massive):
    cfg = TOWERS[name]
    sd = synth.make_modality_state_dict(cfg, hidden, seed=3)
    if massive:
        sd = synth.add_massive_activations(sd, cfg, seed=4)
    px = synth.make_pixels(cfg, batch, seed=5)
    m = AutoModality.from_config(AutoModality.config_from_dict(
        {"model_type": "meditron_clip_b200", "hidden_size": hidden, "clip_name": name}))
    m.load_reference_state_dict(sd, DEV)
    y = m([px[i] for i in range(batch)]).float()
    ref32 = _oracle_gpu(sd, cfg, px, torch.float32)
    refbf = _oracle_gpu(sd, cfg, px, torch.bfloat16)
    assert y.shape == ref32.shape == (batch, cfg.num_patches, hidden)
    scale = ref32.abs().max().item()
    err = (y - ref32).abs().max().item()
    ref_err = (refbf - ref32).abs().max().item()
    mean_rel = ((y - ref32).abs().mean() / ref32.abs().mean()).item()
    ref_mean_rel = ((refbf - ref32).abs().mean() / ref32.abs().mean()).item()
    print(f"[fullsize {name}{' +massive' if massive else ''} x{batch}] max-abs {err:.3e} (reference bf16 on GPU: {ref_err:.3e}) "
          f"max-abs/absmax {err / scale:.3e} mean-rel {mean_rel:.3e} (reference bf16: {ref_mean_rel:.3e})")
    assert torch.isfinite(y).all()
    assert err <= 2.0 * ref_err
    assert err / scale <= 2e-2
    assert mean_rel <= 2.0 * ref_mean_rel


@pytest.mark.parametrize("name,hidden,batch", [("clip_b32", 512, 8), ("clip_l14_336", 4096, 4)])
def test_inference_regime_autocast(name, hidden, batch):
    """SURVEY §8 C5 regime (ii): the regime the reference SHIPS for inference — bf16 weights under
    `torch.autocast("cuda", torch.bfloat16)` (inference.py:36,88; ui/app.py:66-70,126).  There LayerNorm and softmax run
    and RETURN fp32 (so the residual stream stays fp32) and Linear / conv re-cast to bf16.  The B200 path keeps the
    pure-bf16 rounding points of regime (i); stated tolerance against regime (ii):
        max|B200 - autocast| <= 2 x max|reference regime (i) - autocast|   and   <= 2e-2 of range,
    i.e. switching a user from the reference's training regime to ours moves the inference result no more than the
    reference's own two regimes differ."""
    cfg = TOWERS[name]
    sd = synth.make_modality_state_dict(cfg, hidden, seed=13)
    px = synth.make_pixels(cfg, batch, seed=14)
    m = AutoModality.from_config(AutoModality.config_from_dict(
        {"model_type": "meditron_clip_b200", "hidden_size": hidden, "clip_name": name}))
    m.load_reference_state_dict(sd, DEV)
    y = m([px[i] for i in range(batch)]).float()
    sdb = {k: v.to(DEV, torch.bfloat16) for k, v in sd.items()}
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        ref_ac = vit_oracle.image_modality_forward(sdb, cfg, px.to(DEV)).float()
    ref_i = _oracle_gpu(sd, cfg, px, torch.bfloat16)          # regime (i): pure bf16 modules, no autocast
    ref32 = _oracle_gpu(sd, cfg, px, torch.float32)
    scale = ref_ac.abs().max().item()
    err = (y - ref_ac).abs().max().item()
    ref_err = (ref_i - ref_ac).abs().max().item()
    print(f"[autocast {name} x{batch}] B200 vs regime (ii): max-abs {err:.3e} /absmax {err / scale:.3e}; reference regime (i) vs (ii): "
          f"{ref_err:.3e}; regime (ii) vs fp32: {(ref_ac - ref32).abs().max().item():.3e}; B200 vs fp32: {(y - ref32).abs().max().item():.3e}")
    assert err <= 2.0 * ref_err and err / scale <= 2e-2
