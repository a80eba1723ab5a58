"""Position ids: oracle vs the goldens made by the reference collator (CPU), kernel vs both (GPU)."""
import os

import numpy as np
import pytest
import torch

from multimeditron_b200 import ops, synth
from oracle import splice_oracle

CASES = ["right_1img", "left_3img", "noimg", "edges", "grid4_left"]
ATT = 3   # attachment id used by oracle/make_golden.py


def _case(golden_dir, name):
    g = np.load(os.path.join(golden_dir, "position_ids.npz"))
    ids, mask, grid = g[name + "__ids"], g[name + "__mask"], int(g[name + "__grid"])
    ranges = [splice_oracle.compute_token_range(ids[b], [grid * grid] * (int((ids[b] == ATT).sum()) // (grid * grid)), ATT)
              for b in range(ids.shape[0])]
    return ids, mask, grid, ranges, g[name + "__pos1d"], g[name + "__pos2d"]


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_collator_golden(golden_dir, name):
    ids, mask, grid, ranges, p1, p2 = _case(golden_dir, name)
    assert np.array_equal(splice_oracle.position_ids(mask, ranges, grid, False), p1)
    assert np.array_equal(splice_oracle.position_ids(mask, ranges, grid, True), p2)


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_kernel_matches_reference_collator_golden(golden_dir, name):
    ids, mask, grid, _, p1, p2 = _case(golden_dir, name)
    ids_d, m64 = torch.from_numpy(ids).cuda(), torch.from_numpy(mask).long().cuda()
    st = ops.new_status("cuda")
    assert np.array_equal(ops.position_ids(m64).cpu().numpy(), p1)
    assert np.array_equal(ops.position_ids(torch.from_numpy(mask).cuda()).cpu().numpy(), p1)          # bool mask
    got = ops.position_ids(m64, ids_d, ATT, grid * grid, grid, use_2d=True, status=st)
    assert np.array_equal(got.cpu().numpy(), p2)
    assert int(st.item()) == 0


@pytest.mark.gpu
@pytest.mark.parametrize("B,S,n_img,grid,side", [(64, 8192, 4, 24, "right"), (64, 8192, 4, 24, "left"), (7, 1000, 11, 3, "left"),
                                                 (3, 5, 0, 2, "right"), (2, 33000, 1, 27, "right")])
def test_kernel_vs_oracle_large_rows(B, S, n_img, grid, side):
    """cfg4-sized rows (4 images of 576 tokens in 8192-token sequences) and odd shapes, bit-exact vs the oracle."""
    P, V = grid * grid, 1000
    ids, mask = synth.make_sequences(B, S, n_img, P, V - 4, V - 2, V - 4, V - 3, V - 1, seed=S + n_img, padding_side=side,
                                     min_len=max(n_img * (P + 2), (3 * S) // 4))
    ranges = [splice_oracle.compute_token_range(ids[b].numpy(), [P] * n_img, V - 2) for b in range(B)]
    m = mask.long().cuda()
    assert np.array_equal(ops.position_ids(m).cpu().numpy(), splice_oracle.position_ids(mask.numpy(), ranges, grid, False))
    got = ops.position_ids(m, ids.cuda(), V - 2, P, grid, use_2d=True)
    assert np.array_equal(got.cpu().numpy(), splice_oracle.position_ids(mask.numpy(), ranges, grid, True))


@pytest.mark.gpu
def test_kernel_flags_an_image_block_that_runs_past_the_row():
    ids = torch.full((1, 10), 5, dtype=torch.int64)
    ids[0, 7:] = 3                                   # only 3 of the 4 attachment tokens fit
    st = ops.new_status("cuda")
    ops.position_ids(torch.ones(1, 10, dtype=torch.int64).cuda(), ids.cuda(), 3, 4, 2, use_2d=True, status=st)
    assert int(st.item()) & ops.STATUS_COUNT_MISMATCH


@pytest.mark.gpu
def test_adjacent_images_without_delimiters_are_two_images():
    """The k-th image starts at the (k*P)-th attachment token of the row (compute_token_range, prompt_tokenizers.py:255-286),
    not at the start of a run of attachment ids: two images that touch are still two images."""
    grid, V = 3, 100
    P, att = grid * grid, V - 2
    ids = torch.randint(0, V - 4, (2, 40))
    ids[0, 5:5 + 2 * P] = att                 # two images back to back
    ids[1, 0:P] = att                         # one at the row start ...
    ids[1, P:2 * P] = att                     # ... immediately followed by another
    ids[1, 30:30 + P] = att
    mask = torch.ones(2, 40, dtype=torch.bool)
    ranges = [splice_oracle.compute_token_range(ids[0].numpy(), [P, P], att),
              splice_oracle.compute_token_range(ids[1].numpy(), [P, P, P], att)]
    assert ranges[0] == [(5, 5 + P), (5 + P, 5 + 2 * P)]
    want = splice_oracle.position_ids(mask.numpy(), ranges, grid, True)
    st = ops.new_status("cuda")
    got = ops.position_ids(mask.cuda(), ids.cuda(), att, P, grid, use_2d=True, status=st)
    assert np.array_equal(got.cpu().numpy(), want) and int(st.item()) == 0
