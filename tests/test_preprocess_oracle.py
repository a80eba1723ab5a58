"""CPU: the pre-processing oracle against the committed HF-processor goldens and, where torch/transformers are
importable (they are on both machines), against the live processors; host geometry of the device path."""
import hashlib
import os

import numpy as np
import pytest
import torch

from multimeditron_b200 import preprocess as pre, synth
from oracle import preprocess_oracle as po


case_image = synth.make_raw_image   # the seeded images oracle/make_golden.py:gen_preprocess used


def golden_cases(golden_dir):
    g = np.load(os.path.join(golden_dir, "preprocess.npz"))
    return g, [tuple(int(v) for v in row) for row in g["cases"]]


def test_oracle_matches_hf_processor_golden(golden_dir):
    g, cases = golden_cases(golden_dir)
    for i, (h, w, size, square) in enumerate(cases):
        mean, std = ((0.5,) * 3, (0.5,) * 3) if square else (po.OPENAI_CLIP_MEAN, po.OPENAI_CLIP_STD)
        got = po.clip_preprocess(case_image(i, h, w), size, mean, std, square=bool(square))
        assert got.shape == (3, size, size) and got.dtype == np.float32
        assert hashlib.sha256(got.tobytes()).digest() == g[f"sha_{i}"].tobytes(), f"case {i} {(h, w, size)}"
        if f"u8_{i}" in g.files:
            m, s = po.fused_mean_std(mean, std)
            want = (g[f"u8_{i}"].astype(np.float32) - m[:, None, None]) / s[:, None, None]
            assert np.array_equal(got.view(np.int32), want.view(np.int32))


@pytest.mark.parametrize("shape", [(300, 400, 224, 298), (64, 48, 224, 168), (511, 337, 336, 509), (37, 53, 56, 80),
                                   (2000, 900, 224, 100), (100, 100, 384, 384), (7, 5, 56, 56)])
def test_resize_oracle_bit_exact_vs_torch_uint8_antialias(shape):
    tvF = pytest.importorskip("torchvision.transforms.v2.functional")
    from torchvision.transforms import InterpolationMode
    h, w, oh, ow = shape
    img = np.random.default_rng(h * 7 + w).integers(0, 256, (h, w, 3), dtype=np.uint8)
    ref = tvF.resize(torch.from_numpy(img).permute(2, 0, 1).contiguous(), [oh, ow],
                     interpolation=InterpolationMode.BICUBIC, antialias=True).permute(1, 2, 0).numpy()
    assert np.array_equal(po.resize_bicubic_aa_u8(img, oh, ow), ref)


def test_plan_matches_oracle_geometry():
    shapes = [(300, 400), (1000, 333), (224, 224), (90, 120), (4000, 3000)]
    plan = pre.plan_batch(shapes, 224, square=False)
    off = tmp = 0
    for (h, w), d in zip(shapes, plan.desc):
        oh, ow = po.resize_output_size(h, w, 224)
        top, left = po.center_crop_origin(oh, ow, 224, 224)
        assert tuple(d) == (off, h, w, oh, ow, top, left, tmp)
        assert off % 4 == 0
        off += (h * w * 3 + 3) & ~3
        tmp += h * 224 * 3
        for a, b in ((h, oh), (w, ow)):
            assert po.aa_weights_int16(a, b)[2].shape[1] <= plan.kmax
    assert plan.src_bytes == off and plan.tmp_bytes == tmp and plan.max_h == 4000 and plan.max_w == 3000
    sq = pre.plan_batch([(500, 700)], 384, square=True)
    assert tuple(sq.desc[0][3:7]) == (384, 384, 0, 0)
    with pytest.raises(ValueError):
        pre.plan_batch([(0, 10)], 224, False)


def test_fused_mean_std_matches_hf_fusion():
    p = pre.DevicePreprocessor(224)
    m, s = po.fused_mean_std(po.OPENAI_CLIP_MEAN, po.OPENAI_CLIP_STD)
    assert np.array_equal(np.asarray(p.mean255, np.float32), m) and np.array_equal(np.asarray(p.std255, np.float32), s)


def test_device_preprocessor_follows_the_checkpoints_image_processor():
    """the device path must take mean/std, sizes, crop and resample from the checkpoint's HF image
    processor instead of assuming the OpenAI-CLIP defaults, and refuse what the kernels do not implement."""
    tr = pytest.importorskip("transformers")
    from multimeditron_b200.configs import TOWERS
    ip = tr.CLIPImageProcessor(size={"shortest_edge": 336}, crop_size={"height": 336, "width": 336},
                               image_mean=[0.1, 0.2, 0.3], image_std=[0.4, 0.5, 0.6])
    p = pre.DevicePreprocessor.from_hf_processor(ip, TOWERS["clip_l14_336"])
    m, s = po.fused_mean_std((0.1, 0.2, 0.3), (0.4, 0.5, 0.6))
    assert p.size == 336 and p.square is False
    assert np.array_equal(np.asarray(p.mean255, np.float32), m) and np.array_equal(np.asarray(p.std255, np.float32), s)
    sg = tr.SiglipImageProcessor(size={"height": 384, "width": 384})
    q = pre.DevicePreprocessor.from_hf_processor(sg, TOWERS["siglip_so400m_384"])
    assert q.size == 384 and q.square is True and np.allclose(q.mean255, [127.5] * 3)
    # the stock processors give exactly what for_tower assumes for the named synthetic towers
    d = pre.DevicePreprocessor.from_hf_processor(tr.CLIPImageProcessor(), TOWERS["clip_b32"])
    f = pre.DevicePreprocessor.for_tower(TOWERS["clip_b32"])
    assert (d.size, d.square, d.mean255, d.std255) == (f.size, f.square, f.mean255, f.std255)
    with pytest.raises(ValueError):   # wrong output size for the tower
        pre.DevicePreprocessor.from_hf_processor(tr.CLIPImageProcessor(), TOWERS["clip_l14_336"])
    with pytest.raises(ValueError):   # bilinear resampling is not implemented
        pre.DevicePreprocessor.from_hf_processor(tr.CLIPImageProcessor(resample=2), TOWERS["clip_b32"])
    with pytest.raises(ValueError):   # crop size != resize target
        pre.DevicePreprocessor.from_hf_processor(
            tr.CLIPImageProcessor(size={"shortest_edge": 256}, crop_size={"height": 224, "width": 224}), TOWERS["clip_b32"])
    with pytest.raises(ValueError):   # no normalisation
        pre.DevicePreprocessor.from_hf_processor(tr.CLIPImageProcessor(do_normalize=False), TOWERS["clip_b32"])
    with pytest.raises(ValueError):   # no center crop after a shortest-edge resize
        pre.DevicePreprocessor.from_hf_processor(tr.CLIPImageProcessor(do_center_crop=False), TOWERS["clip_b32"])


def test_oracle_bit_exact_on_the_real_mock_photographs():
    """The numpy restatement vs the HF CPU processor on real pixels (the two mock_dataset photographs, downscaled)."""
    tr = pytest.importorskip("transformers")
    raws = synth.load_mock_images(2)
    ip = tr.CLIPImageProcessor()     # shortest edge 224, center crop 224, OpenAI-CLIP mean/std
    for r in raws:
        want = ip(images=r.numpy(), return_tensors="pt")["pixel_values"][0].numpy()
        got = po.clip_preprocess(r.numpy(), 224, po.OPENAI_CLIP_MEAN, po.OPENAI_CLIP_STD, square=False)
        assert np.array_equal(got.view(np.int32), want.view(np.int32))
