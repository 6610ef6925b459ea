"""GPU test (pytest -m gpu) of gsr_sort_pairs_static: the onesweep sort with its scratch in a `static __device__` array
inside the library, for callers that cannot allocate (include/gsr_b200.h).  No counterpart in the reference; the check is
against the existing entry point gsr_sort_pairs (caller-owned scratch), which tests/test_gpu_parity.py pins against a
stable CPU sort.  Integer work: the two must agree bit for bit, on two calls in a row (the array keeps the previous
call's contents).  Collected last on purpose (file name)."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from dynamic3dgaussians_b200 import _lib  # noqa: E402

pytestmark = pytest.mark.gpu


def _inputs(key_bytes, n, bits, seed):
    g = torch.Generator().manual_seed(seed)
    hi = 2 ** 31 - 1 if key_bytes == 4 else 2 ** 62
    keys = torch.randint(0, hi, (n,), generator=g, dtype=torch.int64)
    if bits < key_bytes * 8:
        keys &= (1 << bits) - 1
    if n > 1000:
        keys[: n // 2] = keys[n // 2: 2 * (n // 2)]            # duplicates: stability shows in the values
    kd = (keys.to(torch.int32) if key_bytes == 4 else keys).cuda()
    vd = torch.arange(n, dtype=torch.int32).cuda()
    return keys, kd, vd


def _existing(lib, key_bytes, kd, vd, n, bits):
    ko, vo = torch.empty_like(kd), torch.empty_like(vd)
    tb = lib.gsr_sort_temp_bytes(n, key_bytes)
    tmp = torch.empty(max(tb, 1), dtype=torch.uint8, device="cuda")
    _lib.check(lib.gsr_sort_pairs(key_bytes, kd.data_ptr(), ko.data_ptr(), vd.data_ptr(), vo.data_ptr(), n, 0, bits,
                                  tmp.data_ptr(), tb, torch.cuda.current_stream().cuda_stream), "gsr_sort_pairs")
    return ko, vo


def _static(lib, key_bytes, kd, vd, n, bits):
    # outputs are the caller's; they are allocated BEFORE the call, which itself allocates nothing
    ko, vo = torch.full_like(kd, -1), torch.full_like(vd, -1)
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats()["allocation.all.allocated"]
    rc = lib.gsr_sort_pairs_static(key_bytes, kd.data_ptr(), ko.data_ptr(), vd.data_ptr(), vo.data_ptr(), n, 0, bits,
                                   torch.cuda.current_stream().cuda_stream)
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == before
    _lib.check(rc, "gsr_sort_pairs_static")
    return ko, vo


def test_static_scratch_sort_equals_the_existing_entry_point_twice_in_a_row():
    lib = _lib.load()
    nmax = int(lib.gsr_sort_pairs_static_max())
    assert nmax == 262144
    # two calls in a row with different sizes, key widths and bit ranges: the second sees the first one's leftovers in
    # the shared array (a larger sort first, then smaller ones whose look-back words lie inside the dirtied region)
    cases = [(8, nmax, 42, 1), (4, 100_000, 32, 2), (8, 3073, 64, 3), (4, nmax, 17, 4), (4, 1, 32, 5), (8, 12_289, 40, 6)]
    for key_bytes, n, bits, seed in cases:
        keys, kd, vd = _inputs(key_bytes, n, bits, seed)
        ko_ref, vo_ref = _existing(lib, key_bytes, kd, vd, n, bits)
        for call in range(2):
            ko, vo = _static(lib, key_bytes, kd, vd, n, bits)
            torch.cuda.synchronize()
            assert torch.equal(ko, ko_ref), (key_bytes, n, bits, call)
            assert torch.equal(vo, vo_ref), (key_bytes, n, bits, call)
        # and the pair is what a stable CPU sort gives (so both are right, not merely equal)
        rk, ri = torch.sort(keys, stable=True)
        assert torch.equal(ko.cpu().to(torch.int64), rk) and torch.equal(vo.cpu().to(torch.int64), ri)


def test_static_scratch_sort_rejects_sizes_beyond_the_maximum():
    lib = _lib.load()
    nmax = int(lib.gsr_sort_pairs_static_max())
    n = nmax + 1
    kd = torch.zeros(n, dtype=torch.int32, device="cuda")
    vd = torch.zeros(n, dtype=torch.int32, device="cuda")
    ko, vo = torch.full_like(kd, 7), torch.full_like(vd, 7)
    rc = lib.gsr_sort_pairs_static(4, kd.data_ptr(), ko.data_ptr(), vd.data_ptr(), vo.data_ptr(), n, 0, 32,
                                   torch.cuda.current_stream().cuda_stream)
    assert rc == -2 and lib.gsr_strerror(rc) == b"workspace too small"
    msg = lib.gsr_last_cuda_error().decode()
    assert "GSR_SORT_STATIC_MAX_PAIRS" in msg and str(n) in msg
    torch.cuda.synchronize()
    assert int(ko.min()) == 7 and int(vo.min()) == 7            # nothing ran
    with pytest.raises(RuntimeError):
        _lib.check(rc, "gsr_sort_pairs_static")
    # bad arguments are refused like gsr_sort_pairs refuses them; n = 0 is a no-op
    s = torch.cuda.current_stream().cuda_stream
    assert lib.gsr_sort_pairs_static(2, kd.data_ptr(), ko.data_ptr(), vd.data_ptr(), vo.data_ptr(), 8, 0, 16, s) == -1
    assert lib.gsr_sort_pairs_static(4, kd.data_ptr(), ko.data_ptr(), vd.data_ptr(), vo.data_ptr(), 8, 0, 33, s) == -1
    assert lib.gsr_sort_pairs_static(4, None, None, None, None, 0, 0, 32, s) == 0
    assert lib.gsr_sort_pairs_static(4, None, ko.data_ptr(), vd.data_ptr(), vo.data_ptr(), 8, 0, 32, s) == -1
