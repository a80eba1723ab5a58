"""Shapes of the image towers the hot path supports (SURVEY.md §8 config shapes).

Field names follow HF `CLIPVisionConfig` / `SiglipVisionConfig` (HF:models/clip/configuration_clip.py).
"""
from __future__ import annotations

from dataclasses import dataclass, asdict


@dataclass(frozen=True)
class TowerConfig:
    kind: str = "clip"              # "clip" | "siglip"
    image_size: int = 224
    patch_size: int = 32
    hidden_size: int = 768          # D
    intermediate_size: int = 3072   # I
    num_hidden_layers: int = 12     # L
    num_attention_heads: int = 12
    layer_norm_eps: float = 1e-5
    hidden_act: str = "quick_gelu"

    # derived ---------------------------------------------------------------------------------
    @property
    def grid(self) -> int:
        return self.image_size // self.patch_size

    @property
    def num_patches(self) -> int:   # P
        return self.grid * self.grid

    @property
    def has_cls(self) -> bool:      # CLIP prepends a class token (modeling_clip.py:212-213), SigLIP does not
        return self.kind == "clip"

    @property
    def num_tokens(self) -> int:    # T
        return self.num_patches + (1 if self.has_cls else 0)

    @property
    def head_dim(self) -> int:
        return self.hidden_size // self.num_attention_heads

    @property
    def conv_bias(self) -> bool:    # modeling_clip.py:148-154 bias=False; modeling_siglip.py:124-130 default bias
        return self.kind == "siglip"

    @property
    def patch_k(self) -> int:
        return 3 * self.patch_size * self.patch_size

    def to_dict(self):
        return asdict(self)

    def flops_per_image(self, llm_hidden: int) -> float:
        """Algorithmic FLOPs (2*M*N*K over every GEMM incl. QK^T and PV) — SURVEY.md §8 D4."""
        P, T, D, I, L = self.num_patches, self.num_tokens, self.hidden_size, self.intermediate_size, self.num_hidden_layers
        patch = 2 * P * D * self.patch_k
        layer = 2 * T * D * 3 * D + 2 * T * D * D + 2 * 2 * T * T * D + 2 * 2 * T * D * I
        proj = 2 * P * (D * D + D * llm_hidden + llm_hidden * llm_hidden)
        return float(patch + L * layer + proj)


TOWERS = {
    # test-sized tower (head_dim 64 like the real ones)
    "tiny": TowerConfig("clip", 32, 8, 128, 256, 2, 2),
    # test-sized SigLIP tower: head_dim 72, MLP width not a multiple of 64, conv bias, no CLS, post-LN, tanh-GELU
    "tiny_siglip": TowerConfig("siglip", 56, 14, 144, 304, 2, 2, 1e-6, "gelu_pytorch_tanh"),
    "clip_b32": TowerConfig("clip", 224, 32, 768, 3072, 12, 12),
    "clip_l14_336": TowerConfig("clip", 336, 14, 1024, 4096, 24, 16),
    "siglip_so400m_384": TowerConfig("siglip", 384, 14, 1152, 4304, 27, 16, 1e-6, "gelu_pytorch_tanh"),
}
