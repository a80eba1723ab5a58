"""Build-container only (needs /root/reference): the reference-side binding of INTEGRATION.md §2.

Constructs `meditron_clip_b200` through the REFERENCE's own AutoModality / HF AutoModel route, and checks that it is a
drop-in for `meditron_clip`: same config fields, processor class, state-dict keys — and that without a CUDA device the
call fails loudly instead of falling back."""
import pytest
import torch

from oracle import ref_shim

pytestmark = pytest.mark.needs_reference


@pytest.fixture(scope="module")
def ref():
    return ref_shim.import_reference()


@pytest.fixture(scope="module")
def clip_dir(tmp_path_factory):
    from transformers import CLIPConfig, CLIPImageProcessor, CLIPModel
    d = str(tmp_path_factory.mktemp("clip_tiny"))
    vc = dict(hidden_size=128, intermediate_size=256, num_hidden_layers=2, num_attention_heads=2, image_size=32,
              patch_size=8)
    tc = dict(hidden_size=32, intermediate_size=64, num_hidden_layers=1, num_attention_heads=2, vocab_size=64,
              max_position_embeddings=16)
    CLIPModel(CLIPConfig(vision_config=vc, text_config=tc, projection_dim=32)).save_pretrained(d)
    CLIPImageProcessor(size={"shortest_edge": 32}, crop_size={"height": 32, "width": 32}).save_pretrained(d)
    return d


def test_registers_with_reference_automodality_and_matches_image_modality(ref, clip_dir):
    from multimeditron_b200 import integration
    cls = integration.register_with_reference()
    assert integration.register_with_reference() is cls            # idempotent
    AutoModality = ref.modalities.AutoModality
    assert AutoModality._registry["meditron_clip_b200"] is cls
    from multimeditron.model.modalities.image_modality import ImageConfig, ImageModality, ImageProcessor
    cfg = AutoModality.config_from_dict({"model_type": "meditron_clip_b200", "hidden_size": 64, "clip_name": clip_dir})
    assert isinstance(cfg, ImageConfig) and cfg.modality_type == "image" and cfg.hidden_size == 64
    from transformers import AutoModel
    m = AutoModel.from_config(cfg)                                   # model.py:271 route
    assert isinstance(m, cls) and isinstance(m, ImageModality)
    ref_m = ImageModality(ImageConfig(hidden_size=64, clip_name=clip_dir))
    assert list(m.state_dict().keys()) == list(ref_m.state_dict().keys())
    assert {k: tuple(v.shape) for k, v in m.state_dict().items()} == \
           {k: tuple(v.shape) for k, v in ref_m.state_dict().items()}
    proc = AutoModality.preprocessor_from_name("meditron_clip_b200", cfg)
    assert isinstance(proc, ImageProcessor) and proc._num_patches_per_entry == 16
    assert m.embedding_size == 128 and m._num_patches_per_entry == 16


def test_no_silent_fallback(ref, clip_dir):
    from multimeditron_b200 import integration, ops
    cls = integration.register_with_reference()
    from multimeditron.model.modalities.image_modality import ImageConfig
    m = cls(cls.config_class(hidden_size=64, clip_name=clip_dir))
    px = [torch.zeros(3, 32, 32)]
    with pytest.raises(NotImplementedError):     # grad-requiring call: backward is not part of this path
        m(px)
    if not torch.cuda.is_available():
        with torch.no_grad(), pytest.raises((ops.MMError, RuntimeError, AssertionError)):
            m(px)                                # CPU tensors / no device: loud failure, never a torch fallback
