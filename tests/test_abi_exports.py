"""-m "not gpu": the C-ABI library loads and exports every symbol include/mm_b200.h declares (no compute calls)."""
import ctypes
import re

import pytest

from multimeditron_b200 import _lib


def test_header_declares_expected_entry_points():
    syms = _lib.header_symbols()
    for name in ("mm_attach_positions", "mm_splice_scatter", "mm_splice_scatter_idx", "mm_embed_splice_fused",
                 "mm_layernorm", "mm_gemm_bias_act", "mm_patch_embed", "mm_attention_fwd"):
        assert name in syms


def test_library_exports_every_header_symbol():
    if not _lib.LIB_PATH.exists():
        import __graft_entry__
        __graft_entry__.build()
    lib = ctypes.CDLL(str(_lib.LIB_PATH))
    for name in _lib.header_symbols():
        assert hasattr(lib, name), f"{name} declared in include/mm_b200.h but not exported"


def test_binding_table_matches_header():
    assert set(_lib.SIGNATURES) == set(_lib.header_symbols())
    # argument counts of the ctypes table == parameter counts in the header
    text = re.sub(r"/\*.*?\*/", "", _lib.HEADER_PATH.read_text(), flags=re.S)
    for name, (_, args) in _lib.SIGNATURES.items():
        m = re.search(r"\b" + name + r"\s*\(([^;]*?)\)\s*;", text, flags=re.S)
        assert m, name
        params = m.group(1).strip()
        n = 0 if params in ("void", "") else params.count(",") + 1
        assert n == len(args), (name, n, len(args))


def test_no_compute_without_gpu_fails_loudly():
    import torch
    from multimeditron_b200 import ops
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    x = torch.zeros(4, 8, dtype=torch.bfloat16)
    with pytest.raises(ops.MMError):
        ops.layernorm(x, x[0], x[0], 1e-5)
    with pytest.raises(ops.MMError):
        ops.gemm_bias_act(x, x)


def test_version_string():
    lib = _lib.load()
    assert b"sm_100a" in lib.mm_version()
