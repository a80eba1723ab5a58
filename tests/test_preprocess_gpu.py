"""GPU parity (bit-exact): mm_preprocess_images vs the oracle / the HF-processor goldens, through the C-ABI."""
import hashlib
import os

import numpy as np
import pytest
import torch

from multimeditron_b200 import preprocess as pre, synth
from multimeditron_b200.configs import TOWERS
from multimeditron_b200.modalities import AutoModality
from oracle import preprocess_oracle as po

case_image = synth.make_raw_image


def golden_cases(golden_dir):
    g = np.load(os.path.join(golden_dir, "preprocess.npz"))
    return g, [tuple(int(v) for v in row) for row in g["cases"]]

pytestmark = pytest.mark.gpu
DEV = "cuda"


def test_kernel_matches_hf_processor_golden(golden_dir):
    g, cases = golden_cases(golden_dir)
    for size, square in sorted({(c[2], c[3]) for c in cases}):
        idx = [i for i, c in enumerate(cases) if (c[2], c[3]) == (size, square)]
        imgs = [case_image(i, cases[i][0], cases[i][1]) for i in idx]       # one ragged batch per processor config
        p = pre.DevicePreprocessor(size, (0.5,) * 3, (0.5,) * 3, square=True) if square else pre.DevicePreprocessor(size)
        out = p(imgs, torch.device(DEV), torch.float32).cpu().numpy()
        for k, i in enumerate(idx):
            assert hashlib.sha256(out[k].tobytes()).digest() == g[f"sha_{i}"].tobytes(), f"case {i} {cases[i]}"


@pytest.mark.parametrize("size,shapes", [
    (224, [(300, 400), (1000, 333), (224, 224), (90, 120), (225, 224), (224, 4000), (3, 3)]),
    (336, [(1536, 2048), (336, 336), (337, 600), (4001, 2999)]),
    (56, [(1, 1), (56, 57), (2000, 56), (13, 900)]),
])
def test_ragged_batch_bit_exact_vs_oracle(size, shapes):
    rng = np.random.default_rng(size)
    imgs = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for h, w in shapes]
    p = pre.DevicePreprocessor(size)
    out32 = p(imgs, torch.device(DEV), torch.float32)
    outbf = p([torch.from_numpy(im).to(DEV) for im in imgs], torch.device(DEV), torch.bfloat16)   # device-resident raw images
    torch.cuda.synchronize()
    for k, im in enumerate(imgs):
        want = torch.from_numpy(po.clip_preprocess(im, size))
        assert torch.equal(out32[k].cpu(), want), f"image {k} {im.shape}"
        assert torch.equal(outbf[k].cpu(), want.to(torch.bfloat16)), f"image {k} bf16"


def test_siglip_square_resize_and_empty_batch():
    rng = np.random.default_rng(5)
    imgs = [rng.integers(0, 256, s + (3,), dtype=np.uint8) for s in [(500, 700), (100, 80), (384, 384)]]
    p = pre.DevicePreprocessor.for_tower(TOWERS["siglip_so400m_384"])
    out = p(imgs, torch.device(DEV), torch.float32).cpu()
    for k, im in enumerate(imgs):
        assert torch.equal(out[k], torch.from_numpy(po.clip_preprocess(im, 384, (0.5,) * 3, (0.5,) * 3, square=True)))
    assert p([], torch.device(DEV)).shape == (0, 3, 384, 384)
    with pytest.raises(ValueError):
        p([np.zeros((4, 4), np.uint8)], torch.device(DEV))


def test_modality_forward_on_raw_uint8_equals_forward_on_processed_pixels():
    """The plugin call with raw images (device pre-processing) must give the very same embeddings as the call with
    the processor's fp32 pixel_values (reference flow: image_modality.py:96 then :130-137)."""
    cfg = AutoModality.config_from_dict({"model_type": "meditron_clip_b200", "hidden_size": 64, "clip_name": "tiny",
                                         "device_preprocess": True})
    m = AutoModality.from_config(cfg).load_reference_state_dict(synth.make_modality_state_dict(TOWERS["tiny"], 64, seed=3), DEV)
    rng = np.random.default_rng(9)
    raws = [rng.integers(0, 256, s + (3,), dtype=np.uint8) for s in [(40, 60), (32, 32), (100, 33)]]
    y_raw = m([torch.from_numpy(r) for r in raws])
    y_px = m([torch.from_numpy(po.clip_preprocess(r, TOWERS["tiny"].image_size)) for r in raws])
    assert y_raw.shape == (3, TOWERS["tiny"].num_patches, 64)
    assert torch.equal(y_raw, y_px)


def test_batched_uint8_tensor_path_equals_list_path():
    """[B,H,W,3] uint8 tensor (pinned host batch, no packing copy) == list-of-images path, through the modality."""
    cfg = AutoModality.config_from_dict({"model_type": "meditron_clip_b200", "hidden_size": 64, "clip_name": "tiny"})
    m = AutoModality.from_config(cfg).load_reference_state_dict(synth.make_modality_state_dict(TOWERS["tiny"], 64, seed=3), DEV)
    raw = torch.randint(0, 256, (5, 48, 64, 3), dtype=torch.uint8, generator=torch.Generator().manual_seed(2))
    y_list = m([raw[i] for i in range(5)])
    y_batch = m(raw.pin_memory())
    assert torch.equal(y_list, y_batch)
    p = pre.DevicePreprocessor(32)
    a = p(raw, torch.device(DEV), torch.float32).cpu()
    for i in range(5):
        assert torch.equal(a[i], torch.from_numpy(po.clip_preprocess(raw[i].numpy(), 32)))


@pytest.mark.parametrize("tower", ["clip_b32", "clip_l14_336"])
def test_real_photographs_bit_exact_vs_hf_cpu_processor(tower):
    """SURVEY D2 / VERDICT r1 #8: the reference's mock_dataset photographs (committed downscaled, tests/golden/mock)
    through the REAL CPU image processor (what ImageProcessor.process runs, image_modality.py:96) vs the device path."""
    tr = pytest.importorskip("transformers")
    cfg = TOWERS[tower]
    raws = synth.load_mock_images(4)
    ip = tr.CLIPImageProcessor(size={"shortest_edge": cfg.image_size},
                               crop_size={"height": cfg.image_size, "width": cfg.image_size})
    want = torch.stack([ip(images=r.numpy(), return_tensors="pt")["pixel_values"][0] for r in raws], 0)
    got = pre.DevicePreprocessor.from_hf_processor(ip, cfg)(raws, torch.device(DEV), torch.float32).cpu()
    assert got.shape == want.shape == (4, 3, cfg.image_size, cfg.image_size)
    assert torch.equal(got, want)
    assert want.std().item() > 0.3          # real image content, not a constant
