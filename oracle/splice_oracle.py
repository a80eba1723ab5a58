"""numpy/torch restatement of the splice and its index construction (TEST INFRASTRUCTURE; bit-exact work).

  * splice statement            src/multimeditron/model/model.py:441-444
  * embedding lookup            src/multimeditron/model/model.py:433
  * token ranges per sample     src/multimeditron/model/prompt_tokenizers.py:255-286  (compute_token_range)
  * collator index expansion    src/multimeditron/model/data_loader.py:124-141
"""
from __future__ import annotations

import numpy as np
import torch


def compute_token_range(sequence_input_ids, num_embeddings_per_modality, attachment_token_idx):
    """prompt_tokenizers.py:255-286: k-th modality starts at the (sum of previous lengths)-th attachment token."""
    if len(num_embeddings_per_modality) == 0:
        return []
    ids = np.asarray(sequence_input_ids)
    idx = np.argwhere(ids == attachment_token_idx).flatten()
    starts = idx[np.cumsum([0] + list(num_embeddings_per_modality[:-1]))].tolist()
    return [(s, s + n) for s, n in zip(starts, num_embeddings_per_modality)]


def collator_indices(token_ranges_per_sample):
    """data_loader.py:124-141: expand [(batch_idx, (start,end))...] to per-token index vectors (int64)."""
    bidx, tr = [], []
    for b, ranges in enumerate(token_ranges_per_sample):
        for (s, e) in ranges:
            bidx.append(np.full(e - s, b, dtype=np.int64))
            tr.append(np.arange(s, e, dtype=np.int64))
    if not bidx:
        return np.zeros(0, np.int64), np.zeros(0, np.int64)
    return np.concatenate(bidx), np.concatenate(tr)


def attach_positions(input_ids: np.ndarray, attachment_token_idx: int):
    """Row-major positions of attachment tokens == (batch_idx*S + token_range) of the collator."""
    flat = np.asarray(input_ids).reshape(-1)
    dest = np.flatnonzero(flat == attachment_token_idx).astype(np.int32)
    sop = np.full(flat.shape[0], -1, dtype=np.int32)
    sop[dest] = np.arange(dest.shape[0], dtype=np.int32)
    return dest, sop


def embed(table: torch.Tensor, input_ids: torch.Tensor) -> torch.Tensor:
    """model.py:433."""
    return torch.nn.functional.embedding(input_ids, table)


def splice(embedded_tokens: torch.Tensor, batch_idx: torch.Tensor, token_range: torch.Tensor,
           embedded_modality_stack: torch.Tensor) -> torch.Tensor:
    """model.py:441-444 verbatim semantics (in place, returns the same tensor)."""
    embedded_tokens[batch_idx, token_range] = embedded_modality_stack.view(
        -1, embedded_modality_stack.shape[-1]).to(embedded_tokens.dtype)
    return embedded_tokens


def embed_modalities_with_text(table, input_ids, batch_idx, token_range, projected):
    """model.py:410-446 for a single modality type."""
    return splice(embed(table, input_ids), batch_idx, token_range, projected)


def splice_rows_numpy(dest_bytes: np.ndarray, src_bytes: np.ndarray, dest_rows: np.ndarray) -> np.ndarray:
    """Byte-level row scatter (used as the timed CPU baseline of the splice): dest[dest_rows[k]] = src[k]."""
    dest_bytes[dest_rows] = src_bytes
    return dest_bytes


def position_ids(attention_mask: np.ndarray, token_ranges_per_sample, grid: int, use_2d: bool) -> np.ndarray:
    """data_loader.py:155-188: `cumsum(mask) - 1`, zero where masked; with use_2d_position_ids every image's tokens get
    (row, col) grid ids offset by the position at its first token, and everything after the image is shifted by
    (largest id of the image's last token) - (1-D id of its last token; 0 when the image starts the row, :178)."""
    m = np.asarray(attention_mask).astype(np.int64)
    pos = np.cumsum(m, -1) - 1
    pos[m == 0] = 0
    if not use_2d:
        return pos
    pos = np.repeat(pos[..., None], 2, -1)
    ii, jj = np.meshgrid(np.arange(grid), np.arange(grid), indexing="ij")
    grid_ids = np.stack([ii, jj], -1).reshape(grid * grid, 2).astype(np.int64)   # image_modality.py:101-108
    for b, ranges in enumerate(token_ranges_per_sample):
        for (s, e) in ranges:
            old_last = pos[b, e - 1].copy() if s > 0 else np.zeros(2, np.int64)
            mod = grid_ids + pos[b, s][None, :]
            next_last = np.full(2, mod[-1].max())
            pos[b, s:e] = mod
            pos[b, e:] += (next_last - old_last)[None, :]
    return pos
