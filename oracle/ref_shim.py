"""Import the UNMODIFIED reference (TEST INFRASTRUCTURE).

Source of the package, in this order:
  1. /root/reference/src/multimeditron      (build container)
  2. oracle/_ref/multimeditron              (staged byte-for-byte by `python -m oracle.build_ref`; git-ignored, travels
                                             to the GPU box with the gpurun snapshot)
The package root `multimeditron/__init__.py:1` eagerly imports the CLI (ray, deepspeed, hydra, accelerate — not
installed), so we pre-seed an empty package module whose __path__ points at the source and import only
multimeditron.model.* / multimeditron.dataset.* (SURVEY.md §8 C1).  Only oracle/make_golden.py, tests/ and bench.py's
reference / torch_gpu arms may call this; nothing under multimeditron_b200/ does.
"""
from __future__ import annotations

import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference/src/multimeditron"
STAGED_SRC = os.path.join(HERE, "_ref", "multimeditron")
STAGED_MOCK = os.path.join(HERE, "_ref", "mock_dataset")


def source_dir():
    if os.path.isdir(os.path.join(REF_SRC, "model")):
        return REF_SRC
    if os.path.isdir(os.path.join(STAGED_SRC, "model")):
        return STAGED_SRC
    return None


def mock_dataset_dir():
    for d in ("/root/reference/mock_dataset", STAGED_MOCK):
        if os.path.isdir(d):
            return d
    return None


def available() -> bool:
    return source_dir() is not None


def import_reference():
    src = source_dir()
    if src is None:
        raise RuntimeError("the reference is neither at /root/reference nor staged under oracle/_ref "
                           "(run `python -m oracle.build_ref` in the build container)")
    os.environ.setdefault("HF_HUB_OFFLINE", "1")
    if "multimeditron" not in sys.modules:
        pkg = types.ModuleType("multimeditron")
        pkg.__path__ = [src]
        sys.modules["multimeditron"] = pkg
    import multimeditron.model.modalities as modalities  # noqa: F401  registers meditron_clip etc.
    import multimeditron.model.model as model
    import multimeditron.model.projectors.mlp as mlp
    import multimeditron.model.prompt_tokenizers as prompt_tokenizers
    import multimeditron.model.data_loader as data_loader
    return types.SimpleNamespace(modalities=modalities, model=model, mlp=mlp,
                                 prompt_tokenizers=prompt_tokenizers, data_loader=data_loader, source=src)
