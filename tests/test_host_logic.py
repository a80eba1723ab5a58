"""-m "not gpu": host-side mirror of the reference plugin interface, configs, synthetic data, sharding."""
import pytest
import torch

from multimeditron_b200 import synth
from multimeditron_b200.configs import TOWERS
from multimeditron_b200.modalities import (AutoModality, B200ImageConfig, B200ImageModality, BaseModality,
                                           BaseModalityConfig, BaseModalityProcessor)
from multimeditron_b200.model import B200MultiModalEmbedder


def test_algorithmic_flops_match_survey():
    # cfg2 407.29 GFLOP/img, cfg3 693.13 GFLOP/img, cfg1 8.94
    assert abs(TOWERS["clip_l14_336"].flops_per_image(4096) / 1e9 - 407.29) < 0.01
    assert abs(TOWERS["siglip_so400m_384"].flops_per_image(3584) / 1e9 - 693.13) < 0.01
    assert abs(TOWERS["clip_b32"].flops_per_image(512) / 1e9 - 8.94) < 0.01
    c = TOWERS["clip_l14_336"]
    assert (c.num_patches, c.num_tokens, c.head_dim) == (576, 577, 64)
    s = TOWERS["siglip_so400m_384"]
    assert (s.num_patches, s.num_tokens, s.head_dim, s.has_cls, s.conv_bias) == (729, 729, 72, False, True)


def test_registry_contract_and_errors():
    assert AutoModality._registry["meditron_clip_b200"] is B200ImageModality
    assert B200ImageConfig.model_type == "meditron_clip_b200"
    with pytest.raises(RuntimeError):
        AutoModality()
    with pytest.raises(ValueError):  # base.py:207-208
        AutoModality.preprocessor_from_name("nope")
    with pytest.raises(ValueError):  # base.py:216-217
        AutoModality.config_from_dict({"model_type": "nope"})
    with pytest.raises(AssertionError):
        AutoModality.config_from_dict({})
    with pytest.raises(ValueError):  # duplicate name, base.py:184-185
        AutoModality.register("meditron_clip_b200")(B200ImageModality)
    with pytest.raises(ValueError):  # not a BaseModality, base.py:182-183
        AutoModality.register("x1")(dict)

    class NoProc(BaseModality):
        config_class = BaseModalityConfig
        def freeze_modality_embedder(self): ...
        def unfreeze_modality_embedder(self): ...
        def unfreeze_projection(self): ...
    with pytest.raises(ValueError):  # missing preprocessor_class, base.py:186-187
        AutoModality.register("x2")(NoProc)


def test_config_roundtrip_and_processor_contract():
    cfg = AutoModality.config_from_dict({"model_type": "meditron_clip_b200", "hidden_size": 512,
                                         "clip_name": "clip_b32", "use_2d_position_ids": True})
    assert cfg.modality_type == "image" and cfg.hidden_size == 512
    d = cfg.to_dict()
    assert d["model_type"] == "meditron_clip_b200"
    cfg2 = AutoModality.config_from_dict(d)
    assert cfg2.to_dict() == d
    proc = AutoModality.preprocessor_from_name("meditron_clip_b200", cfg)
    assert isinstance(proc, BaseModalityProcessor)
    px = torch.zeros(3, 224, 224)
    rec = {"type": "image", "value": px}
    out = proc(rec)
    assert out is not rec and out["num_embeddings"] == 49 and out["value"] is px
    assert out["position_ids"].shape == (49, 2) and out["position_ids"][8].tolist() == [1, 1]
    m = AutoModality.from_config(cfg)
    assert isinstance(m, B200ImageModality) and m.embedding_size == 768 and m._num_patches_per_entry == 49
    with pytest.raises(RuntimeError):
        m([px])  # no weights loaded -> loud failure, never a CPU fallback
    m.freeze_modality_embedder()
    with pytest.raises(NotImplementedError):
        m.unfreeze_modality_embedder()   # ViT backward is not built; the projector's is (tests/test_train_gpu.py)
    with pytest.raises(RuntimeError):
        m.unfreeze_projection()          # no weights loaded yet


def test_embedder_modality_lookup_errors():
    cfg = B200ImageConfig(hidden_size=64, clip_name="tiny")
    m = B200ImageModality(cfg)
    emb = B200MultiModalEmbedder(torch.zeros(10, 64), [m], attachment_token_idx=3)
    assert emb._get_modality_by_name("image") is m
    with pytest.raises(KeyError):   # model.py:386-389
        emb._get_modality_by_name("audio")
    emb.modalities_by_type["bad"] = object()
    with pytest.raises(TypeError):  # model.py:392-394
        emb._get_modality_by_name("bad")
    with pytest.raises(ValueError):  # model.py:275-278 duplicate modality type
        B200MultiModalEmbedder(torch.zeros(10, 64), [m, B200ImageModality(cfg)], attachment_token_idx=3)
    assert emb.processors() == {}


def test_synth_state_dict_keys_and_determinism():
    cfg = TOWERS["tiny"]
    a = synth.make_modality_state_dict(cfg, 64, seed=1)
    b = synth.make_modality_state_dict(cfg, 64, seed=1)
    c = synth.make_modality_state_dict(cfg, 64, seed=2)
    assert a.keys() == b.keys() and all(torch.equal(a[k], b[k]) for k in a)
    assert any(not torch.equal(a[k], c[k]) for k in a)
    assert "feature_extractor.vision_model.pre_layrnorm.weight" in a  # HF's spelling
    assert a["projector.projection.4.weight"].shape == (64, 64)
    assert a["feature_extractor.vision_model.embeddings.patch_embedding.weight"].shape == (128, 3, 8, 8)
    s = synth.make_modality_state_dict(TOWERS["siglip_so400m_384"].__class__("siglip", 28, 14, 128, 256, 1, 2), 64)
    assert "feature_extractor.vision_model.embeddings.patch_embedding.bias" in s
    assert "feature_extractor.vision_model.embeddings.class_embedding" not in s


def test_synth_sequences_structure():
    ids, mask = synth.make_sequences(6, 128, 3, 9, 500, 3, 4, 5, 0, seed=4, padding_side="left", min_len=64)
    assert ids.shape == (6, 128) and ids.dtype == torch.int64
    assert int((ids == 3).sum()) == 6 * 3 * 9
    for b in range(6):
        row = ids[b][mask[b]]
        assert not mask[b][: 128 - int(mask[b].sum())].any()  # left padded
        starts = (row == 4).nonzero().flatten()
        assert starts.numel() == 3
        for s in starts.tolist():
            assert (row[s + 1: s + 10] == 3).all() and row[s + 10] == 5


def test_moe_plugin_contract_on_cpu():
    """MoE modalities: registry names, processor num_embeddings rule (image_modality_moe.py:55-87), gating class-name
    permutation (image_modality_moe.py:125-138) and error classes — no GPU needed."""
    import pytest
    import torch
    from multimeditron_b200.modalities import AutoModality, B200MOEImageModality, B200MOEImageModalityPEP
    names = ["a", "b", "c"]
    base = {"hidden_size": 64, "expert_clip_names": names, "image_processor": "tiny",
            "expert_towers": {n: "tiny" for n in names}}
    cfg = AutoModality.config_from_dict({"model_type": "moe_meditron_clip_b200", **base,
                                         "gating_class_names": ["c", "a", "b"]})
    m = AutoModality.from_config(cfg)
    assert isinstance(m, B200MOEImageModality) and m._perm_list == [2, 0, 1] and m.embedding_size == 128
    cfgp = AutoModality.config_from_dict({"model_type": "moe_meditron_clip_pep_b200", **base})
    mp = AutoModality.from_config(cfgp)
    assert isinstance(mp, B200MOEImageModalityPEP) and mp._perm_list == [0, 1, 2] and mp.embedding_size == 64
    proc = AutoModality.preprocessor_from_name("moe_meditron_clip_b200", cfg)
    assert proc.process({"type": "image", "value": torch.zeros(3, 32, 32)})["num_embeddings"] == 16
    cfg_app = AutoModality.config_from_dict({"model_type": "moe_meditron_clip_b200", **base,
                                             "fusion_method": "sequence_append", "top_k_experts": 3})
    assert AutoModality.preprocessor_from_name("moe_meditron_clip_b200", cfg_app).process(
        {"type": "image", "value": torch.zeros(3, 32, 32)})["num_embeddings"] == 48
    with pytest.raises(ValueError):   # gating class not among the experts (image_modality_moe.py:131-132)
        AutoModality.from_config(AutoModality.config_from_dict(
            {"model_type": "moe_meditron_clip_b200", **base, "gating_class_names": ["a", "b", "zzz"]}))
    with pytest.raises(ValueError):   # unknown fusion (processor :86-87)
        AutoModality.preprocessor_from_name("moe_meditron_clip_b200", AutoModality.config_from_dict(
            {"model_type": "moe_meditron_clip_b200", **base, "fusion_method": "nope"})).process(
            {"type": "image", "value": torch.zeros(3, 32, 32)})
    with pytest.raises(AssertionError):  # no experts (:105)
        AutoModality.from_config(AutoModality.config_from_dict(
            {"model_type": "moe_meditron_clip_b200", "hidden_size": 64, "expert_clip_names": []}))
    with pytest.raises(RuntimeError):    # no weights loaded -> loud failure, no fallback
        m([torch.zeros(3, 32, 32)])


def test_gate_batchnorm_fold_is_exact_in_fp32():
    """gating.ResNet50Gate folds eval-mode BatchNorm into the convolutions: same logits as the unfused torchvision net."""
    pytest.importorskip("torchvision")
    from multimeditron_b200.gating import ResNet50Gate
    torch.manual_seed(0)
    a = ResNet50Gate(3, device="cpu", dtype=torch.float32, fold_bn=False)
    for m in a.resnet.modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.running_mean.normal_(0, 0.1); m.running_var.uniform_(0.5, 1.5)
            m.weight.data.uniform_(0.5, 1.5); m.bias.data.normal_(0, 0.1)
    sd = {"resnet." + k: v for k, v in a.resnet.state_dict().items()}
    b = ResNet50Gate(3, device="cpu", dtype=torch.float32, state_dict=sd, fold_bn=True)
    x = torch.randn(1, 3, 224, 224)
    la, ta, wa = a(x)
    lb, tb, wb = b(x)
    assert (la - lb).abs().max().item() <= 1e-5 * la.abs().max().item() and torch.equal(ta, tb)
