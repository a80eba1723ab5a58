"""Import the UNMODIFIED reference from /root/reference in the build container (TEST INFRASTRUCTURE).

The package root `multimeditron/__init__.py:1` eagerly imports the CLI (ray, deepspeed, hydra, accelerate —
not installed), so we pre-seed an empty package module whose __path__ points at the reference source and
import only multimeditron.model.* (SURVEY.md §8 C1).  /root/reference does not exist on the GPU box: nothing
outside oracle/make_golden.py and the `needs_reference` tests may call this.
"""
from __future__ import annotations

import os
import sys
import types

REF_SRC = "/root/reference/src/multimeditron"


def available() -> bool:
    return os.path.isdir(REF_SRC)


def import_reference():
    if not available():
        raise RuntimeError("/root/reference is not present on this machine")
    os.environ.setdefault("HF_HUB_OFFLINE", "1")
    if "multimeditron" not in sys.modules:
        pkg = types.ModuleType("multimeditron")
        pkg.__path__ = [REF_SRC]
        sys.modules["multimeditron"] = pkg
    import multimeditron.model.modalities as modalities  # noqa: F401  registers meditron_clip etc.
    import multimeditron.model.model as model
    import multimeditron.model.projectors.mlp as mlp
    import multimeditron.model.prompt_tokenizers as prompt_tokenizers
    import multimeditron.model.data_loader as data_loader
    return types.SimpleNamespace(modalities=modalities, model=model, mlp=mlp,
                                 prompt_tokenizers=prompt_tokenizers, data_loader=data_loader)
