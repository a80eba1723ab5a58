"""mm_attach_positions_static — the attachment prefix sum for hosts that cannot allocate device memory.

Some programs that load libmm_b200.so are not able to hand it scratch memory: a server that gave the whole GPU
to its own memory pool when it started and forbids cudaMalloc from then on, or a plugin / foreign-language
binding that is given device pointers for inputs and outputs but has no allocator.  For them the library has one
entry point whose scratch is a fixed 32 KB array inside the library itself (include/mm_b200.h,
MM_ATTACH_STATIC_MAX_IDS).  It has no counterpart in the reference, so the checker here is the ordinary entry
point, mm_attach_positions, which the other tests pin to the oracle: both must produce the same bytes.

The file is named test_zz_* so that pytest collects it last — with `-x` a failure here cannot hide another test.
It must pass alone in a fresh process (the first call does the per-device set-up) and after the rest of the suite.
Every call goes straight through the C-ABI (ctypes); torch only owns the input/output buffers.
"""
import numpy as np
import pytest
import torch

from multimeditron_b200 import _lib

pytestmark = pytest.mark.gpu

ATT = 7                      # attachment token id used in these cases
MAX_IDS = 1 << 24            # MM_ATTACH_STATIC_MAX_IDS in include/mm_b200.h
MM_OK, MM_ERR_BAD_ARG = 0, 1


def _header_limit() -> int:
    import re
    m = re.search(r"#define\s+MM_ATTACH_STATIC_MAX_IDS\s+(\d+)", _lib.HEADER_PATH.read_text())
    return int(m.group(1))


def _ptr(t):
    return None if t is None else t.data_ptr()


def _ids(n, p, seed, dev):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(8, 32000, (n,), generator=g, dtype=torch.int64)
    ids[torch.rand(n, generator=g) < p] = ATT
    return ids.to(dev)


def _outputs(n, dev):
    # 0x5A fill: a position the kernels fail to write shows up as a difference between the two entry points
    dest = torch.full((max(n, 1),), 0x5A5A5A5A, dtype=torch.int32, device=dev)
    sop = torch.full((max(n, 1),), 0x5A5A5A5A, dtype=torch.int32, device=dev)
    cnt = torch.full((1,), -123, dtype=torch.int32, device=dev)
    st = torch.zeros(1, dtype=torch.int32, device=dev)
    return dest, sop, cnt, st


def _run_both(lib, ids, expected_rows, stream):
    """One call of each entry point on the same ids; returns their outputs (still on the device)."""
    n, dev = ids.numel(), ids.device
    ref = _outputs(n, dev)
    ws = torch.empty(max(int(lib.mm_attach_positions_workspace_bytes(n)), 4), dtype=torch.uint8, device=dev)
    rc = lib.mm_attach_positions(_ptr(ids), n, ATT, _ptr(ref[0]), n, _ptr(ref[1]), _ptr(ref[2]), expected_rows,
                                 _ptr(ref[3]), _ptr(ws), stream)
    assert rc == MM_OK
    got = _outputs(n, dev)
    rc = lib.mm_attach_positions_static(_ptr(ids), n, ATT, _ptr(got[0]), n, _ptr(got[1]), _ptr(got[2]),
                                        expected_rows, _ptr(got[3]), stream)
    assert rc == MM_OK, f"mm_attach_positions_static rc={rc} cuda={lib.mm_last_cuda_error()}"
    return ref, got


def _same(ref, got):
    for name, a, b in zip(("dest_rows", "src_of_pos", "count", "status"), ref, got):
        assert torch.equal(a, b), f"{name} differs between mm_attach_positions and mm_attach_positions_static"


def test_static_scratch_matches_workspace_entry_point_two_calls_in_a_row():
    """Bit-exact (integer work, tolerance 0) against mm_attach_positions, on two back-to-back calls.

    The second call is shorter and sparser than the first, so most of the shared array still holds the first
    call's block counters when it runs: a result that depended on stale contents would differ here."""
    lib = _lib.load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)          # the entry point resolves its array for the CURRENT device
    stream = torch.cuda.current_stream().cuda_stream
    assert _header_limit() == MAX_IDS

    first = _ids(64 * 8192 + 5, 0.30, 1, dev)       # ragged tail, 257 blocks, dense
    second = _ids(3 * 2048 + 1, 0.02, 2, dev)       # 4 blocks, sparse
    n_second = int((second == ATT).sum())
    torch.cuda.synchronize()

    ref1, got1 = _run_both(lib, first, -1, stream)
    ref2, got2 = _run_both(lib, second, n_second + 1, stream)   # wrong expectation -> COUNT_MISMATCH in both
    torch.cuda.synchronize()
    _same(ref1, got1)
    _same(ref2, got2)
    assert int(got1[2]) == int((first == ATT).sum()) and int(got1[3]) == 0
    assert int(got2[2]) == n_second and int(got2[3]) == 2        # MM_STATUS_COUNT_MISMATCH
    # and the positions themselves, against numpy
    assert np.array_equal(got2[0][:n_second].cpu().numpy(), np.flatnonzero(second.cpu().numpy() == ATT))


def test_static_scratch_calls_do_not_take_device_memory():
    """After the set-up call, a call leaves the GPU's free memory where it was (nothing allocated, nothing kept)."""
    lib = _lib.load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream().cuda_stream
    ids = _ids(1 << 20, 0.1, 3, dev)
    dest, sop, cnt, st = _outputs(ids.numel(), dev)
    args = (_ptr(ids), ids.numel(), ATT, _ptr(dest), ids.numel(), _ptr(sop), _ptr(cnt), -1, _ptr(st), stream)
    assert lib.mm_attach_positions_static(*args) == MM_OK     # set-up (first call in a fresh process)
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info(dev)
    for _ in range(3):
        assert lib.mm_attach_positions_static(*args) == MM_OK
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info(dev)
    assert free_after >= free_before, (free_before, free_after)
    assert int(cnt) == int((ids == ATT).sum())


def test_static_scratch_limit_and_edges():
    lib = _lib.load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream().cuda_stream
    cnt = torch.full((1,), -1, dtype=torch.int32, device=dev)
    # one id past the limit: refused on the host, nothing launched, outputs untouched
    rc = lib.mm_attach_positions_static(_ptr(cnt), MAX_IDS + 1, ATT, None, 0, None, _ptr(cnt), -1, None, stream)
    assert rc == MM_ERR_BAD_ARG
    assert lib.mm_attach_positions_static(None, -1, ATT, None, 0, None, _ptr(cnt), -1, None, stream) == MM_ERR_BAD_ARG
    torch.cuda.synchronize()
    assert int(cnt) == -1
    # empty input: count 0, like mm_attach_positions
    assert lib.mm_attach_positions_static(None, 0, ATT, None, 0, None, _ptr(cnt), -1, None, stream) == MM_OK
    torch.cuda.synchronize()
    assert int(cnt) == 0
    # exactly at the limit (128 MiB of ids): every one of the array's 8192 counters is used
    ids = torch.full((MAX_IDS,), 11, dtype=torch.int64, device=dev)
    ids[::4099] = ATT
    ref, got = _run_both(lib, ids, -1, stream)
    torch.cuda.synchronize()
    _same(ref, got)
    assert int(got[2]) == (MAX_IDS + 4098) // 4099
