"""-m gpu: the whole path (ViT + projector + splice) through the plugin mirror vs the reference goldens / oracle.

Stated tolerance for embedder/projector outputs (bf16 storage, fp32 accumulate), SURVEY.md §8 C4(ii):
  err_vs_fp32_reference <= 2 x (the reference's own bf16 run's error vs its fp32 run)   and
  max-abs error / max|ref| <= 2e-2.
"""
import os

import numpy as np
import pytest
import torch

from multimeditron_b200 import ops, synth
from multimeditron_b200.configs import TOWERS
from multimeditron_b200.modalities import AutoModality, B200ImageConfig, B200ImageModality
from multimeditron_b200.model import B200MultiModalEmbedder
from oracle import splice_oracle, vit_oracle

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _modality(name, hidden, wseed):
    cfg = AutoModality.config_from_dict({"model_type": "meditron_clip_b200", "hidden_size": hidden, "clip_name": name})
    sd = synth.make_modality_state_dict(TOWERS[name], hidden, seed=wseed)
    m = AutoModality.from_config(cfg)
    m.load_reference_state_dict(sd, DEV)
    return m, sd


@pytest.mark.parametrize("name", ["tiny", "clip_b32", "tiny_siglip"])
def test_modality_forward_vs_reference_golden(golden_dir, name):
    g = np.load(os.path.join(golden_dir, f"modality_{name}.npz"))
    hidden, batch = int(g["hidden"]), int(g["batch"])
    m, _ = _modality(name, hidden, int(g["wseed"]))
    px = synth.make_pixels(TOWERS[name], batch, seed=int(g["pseed"]))
    y = m([px[i] for i in range(batch)])                       # CPU fp32 tensors in, like the collator hands them over
    assert y.shape == (batch, TOWERS[name].num_patches, hidden) and y.dtype == torch.bfloat16 and y.is_cuda
    ref32 = torch.from_numpy(g["out_fp32"])
    refbf = torch.from_numpy(g["out_bf16_bits"]).view(torch.bfloat16).float()
    scale = ref32.abs().max().item()
    err = (y.float().cpu() - ref32).abs().max().item()
    ref_err = (refbf - ref32).abs().max().item()
    rel_mean = ((y.float().cpu() - ref32).abs().mean() / ref32.abs().mean()).item()
    print(f"[{name}] max-abs {err:.3e} (reference bf16 run: {ref_err:.3e}), max-abs/absmax {err / scale:.3e}, mean-rel {rel_mean:.3e}")
    assert err <= 2.0 * ref_err + 1e-6
    assert err / scale <= 2e-2


def test_gpu_pixels_and_bf16_pixels_accepted():
    m, sd = _modality("tiny", 64, 1)
    px = synth.make_pixels(TOWERS["tiny"], 3, seed=2)
    a = m([p for p in px])
    b = m([p.to(DEV) for p in px])
    assert torch.equal(a, b)
    with pytest.raises(ValueError):  # HF:models/clip/modeling_clip.py:204-207
        m([torch.zeros(3, 16, 16)])


def _make_batch(B, S, n_img, cfg, V, seed, side):
    att, start, end, pad = V - 4, V - 3, V - 2, V - 1
    ids, _ = synth.make_sequences(B, S, n_img, cfg.num_patches, V - 4, att, start, end, pad, seed=seed,
                                  padding_side=side, min_len=S - 8)
    ranges = [splice_oracle.compute_token_range(ids[b].numpy(), [cfg.num_patches] * n_img, att) for b in range(B)]
    bidx, trng = splice_oracle.collator_indices(ranges)
    px = synth.make_pixels(cfg, B * n_img, seed=seed)
    pmi = {"stacked": {"image": [px[i] for i in range(B * n_img)]},
           "batch_idx": {"image": torch.from_numpy(bidx)}, "token_range": {"image": torch.from_numpy(trng)}}
    return ids, pmi, att, px


@pytest.mark.parametrize("table_dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("side", ["right", "left"])
def test_embed_modalities_with_text_vs_oracle(side, table_dtype):
    """model.py:410-446 end to end: positions from the collator, from the device prefix sum, fused and unfused —
    all four must be bit-identical to each other, and the spliced rows bit-identical to the modality output."""
    cfg = TOWERS["tiny"]
    H, V, B, S, n_img = 64, 300, 3, 64, 2
    m, sd = _modality("tiny", H, 3)
    ids, pmi, att, px = _make_batch(B, S, n_img, cfg, V, 5, side)
    table = (torch.randn(V, H, generator=torch.Generator().manual_seed(1)) * 0.02).to(table_dtype)
    outs = []
    for fuse in (False, True):
        for pos in ("collator", "device"):
            e = B200MultiModalEmbedder(table.to(DEV), [m], att, fuse_splice=fuse)
            outs.append(e(ids, pmi, positions_from=pos))
            e.check()
    for o in outs[1:]:
        assert torch.equal(o, outs[0])
    out = outs[0].cpu()
    assert out.shape == (B, S, H) and out.dtype == table_dtype
    # text rows: exact embedding rows; image rows: exactly the modality output cast to the table dtype
    proj = m(pmi["stacked"]["image"]).cpu()
    want = splice_oracle.embed_modalities_with_text(table, ids, pmi["batch_idx"]["image"], pmi["token_range"]["image"], proj)
    assert torch.equal(out, want)
    # and the modality output itself is within tolerance of the fp32 oracle
    with torch.no_grad():
        ref = vit_oracle.image_modality_forward(sd, cfg, pmi["stacked"]["image"])
    assert (proj.float() - ref).abs().max().item() <= 2e-2 * ref.abs().max().item()


def test_lazy_check_never_returns_uninitialised_rows():
    """Device-resident ids with MORE attachment tokens than modality rows: the check is lazy, and until check() raises
    the un-spliced attachment position must hold the attachment token's real embedding row (what the reference's
    nn.Embedding wrote there, model.py:433), not whatever torch.empty returned."""
    cfg = TOWERS["tiny"]
    m, _ = _modality("tiny", 64, 3)
    ids, pmi, att, _ = _make_batch(2, 64, 2, cfg, 300, 6, "right")
    extra = (ids[1] != att).nonzero()[-1]
    ids[1, extra] = att                         # one attachment token more than modality rows
    table = torch.randn(300, 64, device=DEV).bfloat16()
    e = B200MultiModalEmbedder(table, [m], att)
    poison = torch.full((2 * 64 * 64 * 8,), float("nan"), device=DEV).bfloat16()
    del poison                                  # the caching allocator hands these NaN bytes to the next torch.empty
    out = e(ids.to(DEV), pmi, positions_from="device")
    assert torch.equal(out[1, extra].squeeze(0), table[att])
    assert not torch.isnan(out.float()).any()
    with pytest.raises(RuntimeError):
        e.check()
    # the explicit opt-in keeps the old contract (rows untouched until the splice) — only the shape is promised
    out = e(ids.to(DEV), pmi, positions_from="device", skip_attachment_rows=True)
    assert out.shape == (2, 64, 64)
    with pytest.raises(RuntimeError):
        e.check()


def test_count_mismatch_is_detected_not_truncated():
    cfg = TOWERS["tiny"]
    m, _ = _modality("tiny", 64, 3)
    ids, pmi, att, _ = _make_batch(2, 64, 2, cfg, 300, 6, "right")
    ids[0, (ids[0] == att).nonzero()[0]] = 1   # one attachment token fewer than modality rows
    e = B200MultiModalEmbedder(torch.zeros(300, 64, device=DEV).bfloat16(), [m], att)
    with pytest.raises(RuntimeError):          # HOST ids (the collator's format): verified eagerly, before any launch
        e(ids, pmi, positions_from="device")
    e(ids.to(DEV), pmi, positions_from="device")   # DEVICE ids: lazy, one sync in check()
    with pytest.raises(RuntimeError):
        e.check()
    e.check()                                  # the status word was cleared: one bad batch does not poison later checks
    pmi["batch_idx"]["image"] = pmi["batch_idx"]["image"][:-1]
    pmi["token_range"]["image"] = pmi["token_range"]["image"][:-1]
    e2 = B200MultiModalEmbedder(torch.zeros(300, 64, device=DEV).bfloat16(), [m], att, fuse_splice=False)
    with pytest.raises(RuntimeError):
        e2(ids, pmi)


def test_clip_l14_336_layerwise_sanity():
    """BASELINE configs[1] shapes (ViT-L/14-336 -> 4096), 2 images: CUDA path vs the fp32 oracle on CPU."""
    cfg = TOWERS["clip_l14_336"]
    m, sd = _modality("clip_l14_336", 4096, 0)
    px = synth.make_pixels(cfg, 2, seed=0)
    y = m([px[0], px[1]]).float().cpu()
    with torch.no_grad():
        ref = vit_oracle.image_modality_forward(sd, cfg, px)
        sdb = {k: v.to(torch.bfloat16) for k, v in sd.items()}
        refb = vit_oracle.image_modality_forward(sdb, cfg, px).float()
    scale = ref.abs().max().item()
    err, ref_err = (y - ref).abs().max().item(), (refb - ref).abs().max().item()
    print(f"[clip_l14_336] max-abs {err:.3e} (reference bf16 run: {ref_err:.3e}) max-abs/absmax {err / scale:.3e}")
    assert err <= 2.0 * ref_err and err / scale <= 2e-2


def test_full_size_cfg2_batch_independence_and_splice_placement():
    """BASELINE configs[1] at FULL size (256 images of 336 px, ViT-L/14-336 -> 4096, S = 642): size-independent
    properties — (i) an image's embeddings do not depend on the batch it travels in (bit-exact vs a batch of 3),
    (ii) every projected row lands exactly at its attachment position and every text row is the embedding row."""
    cfg = TOWERS["clip_l14_336"]
    m, _ = _modality("clip_l14_336", 4096, 0)
    B, P, H, V = 256, cfg.num_patches, 4096, 5000
    px = synth.make_pixels(cfg, B, seed=11).to(DEV)
    y = m(px)                                                        # [256, 576, 4096]
    pick = [0, 100, 255]
    y3 = m(px[pick].contiguous())
    assert torch.equal(y[pick], y3)
    att, start, end, pad = V - 2, V - 4, V - 3, V - 1
    S = P + 2 + 64
    ids, _ = synth.make_sequences(B, S, 1, P, V - 4, att, start, end, pad, seed=5)
    table = (torch.randn(V, H, generator=torch.Generator().manual_seed(2)) * 0.02).bfloat16().to(DEV)
    e = B200MultiModalEmbedder(table, [m], att)
    out = e(ids, {"stacked": {"image": px}, "batch_idx": {}, "token_range": {}}, positions_from="device")
    e.check()
    flat, idsd = out.view(B * S, H), ids.to(DEV).view(-1)
    is_att = idsd == att
    assert int(is_att.sum()) == B * P
    assert torch.equal(flat[is_att], y.view(B * P, H))               # row-major order == collator order (model.py:443-444)
    assert torch.equal(flat[~is_att], table[idsd[~is_att]])          # text rows untouched, bit-exact


def test_full_size_siglip_so400m_batch_independence():
    """BASELINE configs[2]'s tower at FULL size (SigLIP so400m-patch14-384 -> 3584, head_dim 72 on the two-stream
    attention kernel, no CLS token): 64 images — an image's embeddings do not depend on the batch it travels in, bit-exact
    against a batch of 3 (different tile schedules, same per-row arithmetic), and nothing overflows."""
    cfg = TOWERS["siglip_so400m_384"]
    m, _ = _modality("siglip_so400m_384", 3584, 0)
    px = synth.make_pixels(cfg, 64, seed=13).to(DEV)
    y = m(px)                                                        # [64, 729, 3584]
    assert y.shape == (64, cfg.num_patches, 3584) and torch.isfinite(y.float()).all()
    pick = [0, 31, 63]
    assert torch.equal(y[pick], m(px[pick].contiguous()))


def test_two_streams_and_two_host_threads_share_one_modality():
    """B4 / include/mm_b200.h: re-entrant across streams and host threads.  One modality instance is driven from two
    CUDA streams at once (launches interleaved by two Python threads, each with its own stream): per-stream activation
    workspaces and patch scratch, no process-wide switches — every result must equal the sequential single-stream one."""
    import threading
    cfg = TOWERS["clip_b32"]
    m, _ = _modality("clip_b32", 512, 11)
    batches = [synth.make_pixels(cfg, 6, seed=20 + i).to(DEV) for i in range(4)]
    want = [m(b).clone() for b in batches]
    torch.cuda.synchronize()
    got, errs = [None] * 4, []
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    start = threading.Barrier(2)

    def worker(t):
        try:
            with torch.cuda.stream(streams[t]):
                start.wait()
                for rep in range(3):
                    for i in (t, t + 2):
                        got[i] = m(batches[i]).clone()
            streams[t].synchronize()
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    for s in streams:
        s.wait_stream(torch.cuda.current_stream())
    ths = [threading.Thread(target=worker, args=(t,)) for t in range(2)]
    [t.start() for t in ths]
    [t.join() for t in ths]
    torch.cuda.synchronize()
    assert not errs, errs
    for i in range(4):
        assert torch.equal(got[i], want[i]), f"batch {i} differs when two streams run concurrently"
