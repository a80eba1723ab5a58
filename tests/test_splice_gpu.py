"""-m gpu: splice / attachment indexing through the C-ABI — BIT-EXACT against the oracle and the reference goldens."""
import os

import numpy as np
import pytest
import torch

from multimeditron_b200 import ops, synth
from oracle import splice_oracle
from oracle.make_golden import ATT, SPLICE_CASES, splice_case_inputs

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _bits(t):
    t = t.cpu().contiguous()
    return t.view(torch.int16).numpy() if t.dtype == torch.bfloat16 else t.numpy()


@pytest.mark.parametrize("case", SPLICE_CASES, ids=[c[0] for c in SPLICE_CASES])
def test_golden_cases_all_three_variants(golden_dir, case):
    g = np.load(os.path.join(golden_dir, "splice.npz"))
    name = case[0]
    ids, table, proj, tpi = splice_case_inputs(case)
    want = g[name + "__out"]
    B, S = ids.shape
    H = table.shape[1]
    ids_d, table_d, proj_d = ids.to(DEV), table.to(DEV), proj.to(DEV)
    st = ops.new_status(DEV)
    # (1) collator indices, like the reference statement
    bidx = torch.from_numpy(g[name + "__bidx"]).to(DEV)
    trng = torch.from_numpy(g[name + "__trng"]).to(DEV)
    emb = ops.embed_gather(table_d, ids_d, status=st)
    ops.splice_scatter_idx(emb, proj_d.reshape(-1, H), bidx, trng, status=st)
    assert np.array_equal(_bits(emb), want)
    # (2) device prefix-sum positions
    n_rows = proj.shape[0] * proj.shape[1]
    dest, sop, cnt = ops.attach_positions(ids_d, ATT, expected_rows=n_rows, status=st)
    assert int(cnt) == n_rows
    assert np.array_equal(dest.cpu().numpy().astype(np.int64), g[name + "__bidx"] * S + g[name + "__trng"])
    emb2 = ops.embed_gather(table_d, ids_d, status=st)
    ops.splice_scatter(emb2.view(B * S, H), proj_d.reshape(-1, H), dest, status=st)
    assert np.array_equal(_bits(emb2), want)
    # (3) fused embed+splice single pass
    emb3 = ops.embed_splice_fused(table_d, ids_d, sop, proj_d.reshape(-1, H), status=st)
    assert np.array_equal(_bits(emb3), want)
    assert int(st.item()) == 0


@pytest.mark.parametrize("n,p", [(0, 0.0), (1, 1.0), (7, 0.5), (2048, 0.3), (2049, 0.9), (100_003, 0.28), (8192 * 64, 0.28)])
def test_attach_positions_vs_oracle(n, p):
    gen = torch.Generator().manual_seed(n)
    ids = torch.randint(10, 1000, (n,), generator=gen)
    ids[torch.rand(n, generator=gen) < p] = 5
    dest_o, sop_o = splice_oracle.attach_positions(ids.numpy(), 5)
    dest, sop, cnt = ops.attach_positions(ids.to(DEV), 5)
    k = int(cnt)
    assert k == dest_o.shape[0]
    assert np.array_equal(dest[:k].cpu().numpy(), dest_o)
    if n:
        assert np.array_equal(sop.cpu().numpy(), sop_o)


def test_negative_indices_wrap_like_index_put():
    B, S, H = 3, 10, 16
    emb = torch.zeros(B, S, H, device=DEV)
    ref = emb.cpu().clone()
    src = torch.randn(4, H)
    bi = torch.tensor([0, -1, 2, -3])
    tr = torch.tensor([-1, 0, -9, 5])  # no duplicate targets (index_put_ with duplicates is undefined)
    ref[bi, tr] = src
    ops.splice_scatter_idx(emb, src.to(DEV), bi.to(DEV), tr.to(DEV))
    assert torch.equal(emb.cpu(), ref)


def test_error_reporting_matches_reference_classes():
    B, S, H = 2, 8, 16
    emb = torch.zeros(B, S, H, device=DEV, dtype=torch.bfloat16)
    src = torch.ones(3, H, device=DEV, dtype=torch.bfloat16)
    st = ops.new_status(DEV)
    # size mismatch is host-known -> RuntimeError at call time, like index_put_ (model.py:441)
    with pytest.raises(RuntimeError):
        ops.splice_scatter_idx(emb, src, torch.zeros(2, dtype=torch.int64, device=DEV),
                               torch.zeros(2, dtype=torch.int64, device=DEV), status=st)
    # out-of-range index -> IndexError (lazy)
    ops.splice_scatter_idx(emb, src, torch.tensor([0, 1, 2], device=DEV), torch.tensor([0, 1, 2], device=DEV), status=st)
    with pytest.raises(IndexError):
        ops.raise_for_status(st)
    # attachment-token count != number of modality rows -> RuntimeError (lazy)
    st.zero_()
    ids = torch.tensor([[1, 5, 5, 2], [5, 1, 1, 1]], device=DEV)
    ops.attach_positions(ids, 5, expected_rows=4, status=st)
    with pytest.raises(RuntimeError):
        ops.raise_for_status(st)
    st.zero_()
    ops.attach_positions(ids, 5, expected_rows=3, status=st)
    ops.raise_for_status(st)
    # embedding id out of vocab -> IndexError (lazy)
    ops.embed_gather(torch.zeros(4, 8, device=DEV), torch.tensor([[1, 9]], device=DEV), status=st)
    with pytest.raises(IndexError):
        ops.raise_for_status(st)


@pytest.mark.parametrize("H", [512, 3584, 4096, 20])
@pytest.mark.parametrize("dts", [("bfloat16", "bfloat16"), ("float32", "bfloat16"), ("bfloat16", "float32"),
                                 ("float32", "float32")])
def test_scatter_dtypes_and_hidden_sizes(H, dts):
    sdt, ddt = getattr(torch, dts[0]), getattr(torch, dts[1])
    gen = torch.Generator().manual_seed(H)
    n_dest, n = 300, 77
    dest = torch.randn(n_dest, H, generator=gen).to(ddt)
    src = torch.randn(n, H, generator=gen).to(sdt)
    rows = torch.randperm(n_dest, generator=gen)[:n]
    want = dest.clone()
    want[rows] = src.to(ddt)
    d = dest.to(DEV)
    ops.splice_scatter(d, src.to(DEV), rows.int().to(DEV))
    assert np.array_equal(_bits(d), _bits(want))


def test_full_size_cfg4_properties():
    """BASELINE configs[3]: 64 rows x 8192 tokens, 4 images/row (4*576 attachment tokens) into H=4096, bf16.
    Size-independent properties: positions sorted & exact; scatter->gather round trip; untouched rows bit-identical;
    checksum of fused == checksum of two-step."""
    B, S, H, P, V = 64, 8192, 4096, 576, 128_258
    ids, _ = synth.make_sequences(B, S, 4, P, V, 128_002, 128_256, 128_257, 128_001, seed=4, padding_side="left",
                                  min_len=6144)
    ids_d = ids.to(DEV)
    n_rows = B * 4 * P
    st = ops.new_status(DEV)
    dest, sop, cnt = ops.attach_positions(ids_d, 128_002, expected_rows=n_rows, status=st)
    assert int(cnt) == n_rows
    d = dest.long()
    assert bool((d[1:] > d[:-1]).all())
    assert bool((ids_d.view(-1)[d] == 128_002).all())
    table = (torch.randn(V, H, device=DEV) * 0.02).bfloat16()
    proj = torch.randn(n_rows, H, device=DEV).bfloat16()
    emb = ops.embed_gather(table, ids_d, status=st)
    before = emb.clone()
    ops.splice_scatter(emb.view(-1, H), proj, dest, status=st)
    flat = emb.view(-1, H)
    assert torch.equal(flat[d], proj)                       # round trip
    keep = torch.ones(B * S, dtype=torch.bool, device=DEV)
    keep[d] = False
    assert torch.equal(flat[keep], before.view(-1, H)[keep])  # everything else untouched
    assert torch.equal(before.view(-1, H)[keep][:1000], table[ids_d.view(-1)[keep][:1000]])
    fused = ops.embed_splice_fused(table, ids_d, sop, proj, status=st)
    assert torch.equal(fused, emb)
    assert int(st.item()) == 0


@pytest.mark.parametrize("H", [20, 520, 4096])
@pytest.mark.parametrize("dt", ["bfloat16", "float32"])
def test_scatter_guard_bands_stay_intact(H, dt):
    """compute-sanitizer is not available on the B200 pool, so out-of-bounds stores are caught here: the destination
    is the middle of a larger sentinel-filled allocation, first and last rows are both written, and the bands on
    either side must come back bit-identical."""
    ddt = getattr(torch, dt)
    n_dest, band = 257, 3
    gen = torch.Generator().manual_seed(H)
    big = torch.full((n_dest + 2 * band, H), 7.0, dtype=ddt, device=DEV)
    dest = big[band:band + n_dest]
    rows = torch.cat([torch.tensor([0, n_dest - 1]), torch.randperm(n_dest - 2, generator=gen)[:61] + 1])
    src = torch.randn(rows.numel(), H, generator=gen).to(ddt)
    st = ops.new_status(DEV)
    ops.splice_scatter(dest, src.to(DEV), rows.int().to(DEV), status=st)
    assert int(st.item()) == 0
    assert bool((big[:band] == 7.0).all()) and bool((big[band + n_dest:] == 7.0).all())
    want = torch.full((n_dest, H), 7.0, dtype=ddt)
    want[rows] = src
    assert np.array_equal(_bits(dest), _bits(want))
