"""Build the UNMODIFIED reference model stack offline (TEST INFRASTRUCTURE; SURVEY.md §8 C2 recipe).

Everything is random-init and local: a CLIP checkpoint directory of a named tower shape, a tiny Llama config
directory, a WordLevel tokenizer with the Llama-3 special tokens the reference's `ChatTemplate.llama()` expects
(model/model.py:48-62), the reference `MultiModalModelForCausalLM` (model/model.py:205-285) and the reference
`DataCollatorForMultimodal` (model/data_loader.py:14-199).  Used by the `-m gpu` tests that drive the reference model
through the B200 plugin and by bench.py's torch_gpu / cfg1 arms.
"""
from __future__ import annotations

import os
import tempfile
from typing import List, Optional

import torch

from . import ref_shim

ATTACHMENT_TOKEN = "<|reserved_special_token_0|>"
SPECIALS = ["<pad>", "<|begin_of_text|>", "<|eot_id|>", "<|start_header_id|>", "<|end_header_id|>", ATTACHMENT_TOKEN,
            "<|image_start|>", "<|image_end|>"]
# Llama-3 style; tokens are separated by blanks because the stand-in tokenizer is word-level
CHAT_TEMPLATE = (
    "{% for m in messages %}<|start_header_id|> {{ m['role'] }} <|end_header_id|> {{ m['content'] }} <|eot_id|> "
    "{% endfor %}{% if add_generation_prompt %}<|start_header_id|> assistant <|end_header_id|> {% endif %}")


def make_tokenizer(n_words: int = 190, padding_side: str = "left"):
    """WordLevel tokenizer wrapped in PreTrainedTokenizerFast: V = len(SPECIALS) + 4 + n_words."""
    from tokenizers import Tokenizer, models, pre_tokenizers
    from transformers import PreTrainedTokenizerFast
    words = ["<unk>", "system", "user", "assistant"] + [f"w{i}" for i in range(n_words)]
    vocab = {t: i for i, t in enumerate(SPECIALS + words)}
    tk = Tokenizer(models.WordLevel(vocab, unk_token="<unk>"))
    tk.pre_tokenizer = pre_tokenizers.WhitespaceSplit()
    tok = PreTrainedTokenizerFast(tokenizer_object=tk, pad_token="<pad>", eos_token="<|eot_id|>",
                                  bos_token="<|begin_of_text|>", unk_token="<unk>", padding_side=padding_side)
    tok.add_special_tokens({"additional_special_tokens": SPECIALS[3:]})
    tok.chat_template = CHAT_TEMPLATE
    return tok


def save_clip_checkpoint(cfg, directory: Optional[str] = None, seed: int = 0) -> str:
    """Random-init HF CLIPModel with `cfg`'s vision tower (multimeditron_b200.configs.TowerConfig) + its image processor."""
    from transformers import CLIPConfig, CLIPImageProcessor, CLIPModel
    d = directory or tempfile.mkdtemp(prefix="clip_ckpt_")
    vc = dict(hidden_size=cfg.hidden_size, intermediate_size=cfg.intermediate_size,
              num_hidden_layers=cfg.num_hidden_layers, num_attention_heads=cfg.num_attention_heads,
              image_size=cfg.image_size, patch_size=cfg.patch_size, layer_norm_eps=cfg.layer_norm_eps,
              hidden_act=cfg.hidden_act)
    tc = dict(hidden_size=32, intermediate_size=64, num_hidden_layers=1, num_attention_heads=2, vocab_size=64,
              max_position_embeddings=16)
    torch.manual_seed(seed)
    CLIPModel(CLIPConfig(vision_config=vc, text_config=tc, projection_dim=32)).save_pretrained(d)
    CLIPImageProcessor(size={"shortest_edge": cfg.image_size},
                       crop_size={"height": cfg.image_size, "width": cfg.image_size}).save_pretrained(d)
    return d


def save_tiny_llama(vocab: int, hidden: int = 512, layers: int = 2, heads: int = 8, directory: Optional[str] = None) -> str:
    from transformers import LlamaConfig
    d = directory or tempfile.mkdtemp(prefix="llama_cfg_")
    LlamaConfig(hidden_size=hidden, intermediate_size=2 * hidden, num_hidden_layers=layers, num_attention_heads=heads,
                num_key_value_heads=heads, vocab_size=vocab, max_position_embeddings=16384,
                tie_word_embeddings=False).save_pretrained(d)
    return d


def perturb_(module: torch.nn.Module, seed: int = 0):
    """Random-init reference models have all biases 0 and LN gamma 1 / beta 0: perturb them so the fused
    bias / LayerNorm epilogues are exercised (biases N(0,0.02), gamma 1+N(0,0.1), beta N(0,0.1))."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in module.named_parameters():
            leaf = name.rsplit(".", 1)[-1]
            is_ln = any(s in name for s in ("layer_norm", "layrnorm", "layernorm"))
            if leaf == "bias":
                p.copy_(torch.randn(p.shape, generator=g) * (0.1 if is_ln else 0.02))
            elif leaf == "weight" and is_ln and p.dim() == 1:
                p.copy_(1.0 + torch.randn(p.shape, generator=g) * 0.1)


def build_reference_model(ref, clip_dir: str, llama_dir: str, tokenizer, model_type: str = "meditron_clip",
                          hidden: int = 512, dtype: str = "float32", seed: int = 0, **modality_kw):
    """-> reference MultiModalModelForCausalLM with ONE image modality of `model_type` (model.py:226-285 route:
    AutoModality.config_from_dict -> AutoModel.from_config -> registered class)."""
    AutoModality = ref.modalities.AutoModality
    mcfg = AutoModality.config_from_dict(dict(model_type=model_type, hidden_size=hidden, clip_name=clip_dir,
                                              **modality_kw))
    cfg = ref.model.MultimodalConfig(
        vocab_size=len(tokenizer), modalities=[mcfg], pad_token_idx=tokenizer.pad_token_id,
        eos_token_idx=tokenizer.convert_tokens_to_ids(tokenizer.eos_token), padding_side=tokenizer.padding_side,
        llm_path=llama_dir, dtype=dtype)
    torch.manual_seed(seed)
    model = ref.model.MultiModalModelForCausalLM(cfg)
    perturb_(model.modalities_with_projection, seed=seed + 1)
    return model.to(getattr(torch, dtype)).eval()


def make_collator(ref, tokenizer, model, use_2d_position_ids: bool = False, add_generation_prompt: bool = False):
    """The reference collator exactly as cli/train.py:145-152 builds it (loaders: raw PIL images pass through)."""
    from multimeditron.dataset.loader import BaseModalityLoader

    class PassThroughLoader(BaseModalityLoader):       # images are handed over already decoded (PIL / tensor)
        def load(self, sample):
            return sample["value"]

    return ref.data_loader.DataCollatorForMultimodal(
        tokenizer=tokenizer, modality_processors=model.processors(), modality_loaders={"image": PassThroughLoader()},
        attachment_token=ATTACHMENT_TOKEN, chat_template=ref.model.ChatTemplate.from_name("llama"),
        add_generation_prompt=add_generation_prompt, use_2d_position_ids=use_2d_position_ids)


def mock_images() -> List:
    """The reference's mock_dataset JPEGs as PIL images (cfg1 real pixels)."""
    from PIL import Image
    d = ref_shim.mock_dataset_dir()
    if d is None:
        return []
    return [Image.open(os.path.join(d, f)).convert("RGB") for f in sorted(os.listdir(d))
            if f.lower().endswith((".jpg", ".jpeg", ".png"))]


def synthetic_images(n: int, size: int = 96, seed: int = 0) -> List:
    from PIL import Image
    g = torch.Generator().manual_seed(seed)
    return [Image.fromarray(torch.randint(0, 256, (size + 7 * i, size + 3 * i, 3), dtype=torch.uint8, generator=g).numpy())
            for i in range(n)]


def make_samples(images: List, images_per_sample: List[int], n_words: int = 190, seed: int = 0) -> List[dict]:
    """Conversation samples in the collator's input schema (data_loader.py:40-56): every sample has
    `images_per_sample[i]` attachment tokens interleaved with seeded words; the texts have different lengths so the
    batch is padded (the `conversations` path pads; `_tokenize_text` does not — SURVEY §0.6)."""
    g = torch.Generator().manual_seed(seed)
    out, k = [], 0
    for i, n_img in enumerate(images_per_sample):
        n_txt = 3 + int(torch.randint(0, 6, (1,), generator=g))
        words = [f"w{int(j)}" for j in torch.randint(0, n_words, (n_txt + n_img,), generator=g)]
        parts = []
        for j in range(n_img):
            parts += [words.pop(), ATTACHMENT_TOKEN]
        parts += words
        mods = []
        for _ in range(n_img):
            mods.append({"type": "image", "value": images[k % len(images)]})
            k += 1
        out.append({"conversations": [{"role": "user", "content": " ".join(parts)},
                                      {"role": "assistant", "content": f"w{i % n_words} w{(i + 1) % n_words}"}],
                    "modalities": mods})
    return out
