"""ctypes binding of the C-ABI in include/mm_b200.h (csrc/libmm_b200.so).

The library is the product path: there is no CPU or PyTorch fallback.  If the shared object is
missing (not built) `load()` raises, and every op in `ops.py` goes through `load()`.
"""
from __future__ import annotations

import ctypes as C
import os
import re
from pathlib import Path

_HERE = Path(__file__).resolve().parent
# MM_B200_LIB: developer override used by tools/ to A/B two builds of the library in one GPU session
LIB_PATH = Path(os.environ["MM_B200_LIB"]) if os.environ.get("MM_B200_LIB") else _HERE / "csrc" / "libmm_b200.so"
HEADER_PATH = _HERE.parent / "include" / "mm_b200.h"

_lib = None

i64, i32, vp, f32 = C.c_int64, C.c_int, C.c_void_p, C.c_float

# name -> (restype, argtypes); mirrors include/mm_b200.h one to one
SIGNATURES = {
    "mm_version": (C.c_char_p, []),
    "mm_last_cuda_error": (i32, []),
    "mm_device_sm_count": (i32, [i32]),
    "mm_attach_positions_workspace_bytes": (i64, [i64]),
    "mm_attach_positions": (i32, [vp, i64, i64, vp, i64, vp, vp, i64, vp, vp, vp]),
    "mm_attach_positions_static": (i32, [vp, i64, i64, vp, i64, vp, vp, i64, vp, vp]),
    "mm_splice_scatter": (i32, [vp, i64, i64, i32, vp, i64, i32, vp, vp, vp]),
    "mm_splice_scatter_idx": (i32, [vp, i64, i64, i64, i32, vp, i64, i32, vp, vp, vp, vp]),
    "mm_embed_splice_fused": (i32, [vp, i64, i64, i32, vp, i64, vp, vp, vp, i64, i32, vp, vp]),
    "mm_embed_gather": (i32, [vp, i64, i64, i32, vp, i64, vp, vp, vp]),
    "mm_embed_gather_skip": (i32, [vp, i64, i64, i32, vp, i64, vp, i64, vp, vp]),
    "mm_gather_rows": (i32, [vp, i64, i64, i32, vp, i64, i32, vp, vp, vp]),
    "mm_gelu_fwd": (i32, [vp, vp, i64, vp]),
    "mm_colsum_workspace_bytes": (i64, [i64, i64]),
    "mm_gelu_bwd_colsum": (i32, [vp, vp, vp, vp, vp, i64, i64, vp]),
    "mm_transpose": (i32, [vp, vp, i64, i64, vp]),
    "mm_layernorm": (i32, [vp, vp, vp, vp, i64, i64, f32, i64, i64, i64, vp]),
    "mm_gemm_bias_act": (i32, [vp, i64, vp, i64, vp, vp, i64, i64, i64, i64, i32, vp, i64, i64, i64, vp, i64, vp]),
    "mm_gemm_row_stats_slots": (i64, [i64]),
    "mm_fold_layernorm": (i32, [vp, i64, vp, vp, vp, vp, i64, vp, i64, i64, vp]),
    "mm_gemm_ln_bias_act": (i32, [vp, i64, vp, i64, vp, vp, i64, i64, i64, i64, i32, vp, i64, vp, i64, vp, vp, f32, vp]),
    "mm_patch_embed": (i32, [vp, i32, i64, i64, i64, vp, i64, vp, vp, vp, vp, i64, vp]),
    "mm_patch_embed_workspace_bytes": (i64, [i64, i64, i64]),
    "mm_patch_embed_ws": (i32, [vp, i32, i64, i64, i64, vp, i64, vp, vp, vp, vp, i64, vp, i64, vp]),
    "mm_attention_head_stride": (i64, [i64]),
    "mm_attention_fwd": (i32, [vp, vp, i64, i64, i64, i64, f32, vp]),
    "mm_weighted_expert_sum": (i32, [vp, vp, vp, vp, i64, i64, i64, i64, vp]),
    "mm_gemm_grouped": (i32, [vp, i64, vp, i64, vp, vp, i64, i64, i64, i64, i64, i32, vp, i64, vp, i64, vp, vp, f32, vp]),
    "mm_conv_rows_nhwc": (i32, [vp, vp, i64, i64, i64, i64, i64, i64, i64, i64, i64, vp]),
    "mm_conv_rows_stem": (i32, [vp, i32, vp, i64, i64, i64, i64, i64, i64, i64, i64, vp]),
    "mm_maxpool_nhwc": (i32, [vp, vp, i64, i64, i64, i64, vp]),
    "mm_avgpool_nhwc": (i32, [vp, vp, i64, i64, i64, vp]),
    "mm_cross_attention": (i32, [vp, vp, vp, i64, i64, i64, i64, i64, C.c_float, vp]),
    "mm_position_ids": (i32, [vp, vp, i32, i64, i64, i64, i64, i64, i32, vp, vp, vp]),
    "mm_preprocess_workspace_bytes": (i64, [i64, i64, i64]),
    "mm_preprocess_images": (i32, [vp, vp, i64, i64, i64, i64, i64, vp, vp, C.POINTER(f32), C.POINTER(f32), vp, i32, vp]),
}


def header_symbols() -> list[str]:
    """Every function name declared in include/mm_b200.h (used by the ABI export test)."""
    text = HEADER_PATH.read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mm_[a-z0-9_]+)\s*\(", text)))


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  multimeditron_b200 has no CPU/PyTorch fallback path."
        )
    lib = C.CDLL(os.fspath(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name, None)
        if fn is None:
            if os.environ.get("MM_B200_LIB"):   # A/B against an older build: bind what it has
                continue
            raise RuntimeError(f"{LIB_PATH} does not export {name}: rebuild it (python __graft_entry__.py)")
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
