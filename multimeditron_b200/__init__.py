"""multimeditron_b200 — B200-native (sm_100a) implementation of MultiMeditron's modality-embedding hot path:
image-embedder ViT forward -> MLP projector -> splice into inputs_embeds, behind the reference's plugin API.

Only what the path needs lives here: `csrc/` (CUDA kernels + the C-ABI of include/mm_b200.h), `ops.py`
(tensor-level wrappers) and the host-side mirror of the reference interface (`modalities/`, `model.py`).
"""
__version__ = "0.1.0"
