"""rt_sort_pairs_static: the reorder pass's sort with every byte of scratch in a static device array (no cudaMalloc in the
call). Collected LAST on purpose (file name).

Checked through the C ABI against the existing entry point rt_sort_pairs on the same pairs: bit-exact keys and values
(integer work), on two calls in a row that share the static array (the second smaller and narrower than the first, so stale
histograms / status flags / ping-pong contents of the first would show), plus the size limit's error."""
import ctypes as C

import numpy as np
import pytest

import cuda_raytracer_b200 as rt
from cuda_raytracer_b200.core import SORT_STATIC_MAX_PAIRS, lib

pytestmark = pytest.mark.gpu


def _sort_static(r, keys, vals, bits):
    """Device buffers from torch (the TEST allocates; the entry point under test does not)."""
    import torch
    n = len(keys)
    dk = torch.from_numpy(keys.view(np.int32)).cuda()
    dv = torch.from_numpy(vals.view(np.int32)).cuda()
    ok = torch.full((max(n, 1),), -1, dtype=torch.int32, device="cuda")
    ov = torch.full((max(n, 1),), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()  # the core runs on its own non-blocking stream
    r.sort_pairs_static(dk.data_ptr(), dv.data_ptr(), ok.data_ptr(), ov.data_ptr(), n, bits)
    return ok.cpu().numpy().view(np.uint32)[:n], ov.cpu().numpy().view(np.uint32)[:n]


def test_static_scratch_sort_matches_the_allocating_entry_point():
    r = rt.Renderer(0)
    try:
        rs = np.random.RandomState(11)
        for n, bits in ((SORT_STATIC_MAX_PAIRS, 32), (300_007, 15)):  # two calls in a row on the shared array
            keys = rs.randint(0, 2 ** 32, size=n, dtype=np.uint64).astype(np.uint32)
            if bits < 32:
                keys &= np.uint32((1 << bits) - 1)
            keys[: n // 5] = keys[0]  # a long run of equal keys: stability
            vals = rs.randint(0, 2 ** 32, size=n, dtype=np.uint64).astype(np.uint32)
            want_k, want_v = r.sort_pairs(keys, vals, bits)
            got_k, got_v = _sort_static(r, keys, vals, bits)
            assert got_k.tobytes() == want_k.tobytes(), f"keys differ (n {n}, {bits} bits)"
            assert got_v.tobytes() == want_v.tobytes(), f"values differ (n {n}, {bits} bits)"
            order = np.argsort(keys, kind="stable")
            assert (got_k == keys[order]).all() and (got_v == vals[order]).all()
        # a third, tiny call (one tile, one digit pass: in -> out directly) and the empty call
        keys = np.array([5, 1, 5, 0, 255, 1], np.uint32)
        vals = np.arange(6, dtype=np.uint32)
        k, v = _sort_static(r, keys, vals, 8)
        assert k.tolist() == [0, 1, 1, 5, 5, 255] and v.tolist() == [3, 1, 5, 0, 2, 4]
        k, v = _sort_static(r, keys[:0], vals[:0], 8)
        assert len(k) == 0
    finally:
        r.close()


def test_static_scratch_sort_rejects_sizes_beyond_its_array():
    r = rt.Renderer(0)
    try:
        with pytest.raises(rt.RtError, match="RT_SORT_STATIC_MAX_PAIRS"):
            r.sort_pairs_static(256, 512, 768, 1024, SORT_STATIC_MAX_PAIRS + 1, 32)  # rejected before any pointer is touched
        with pytest.raises(rt.RtError):
            r.sort_pairs_static(0, 512, 768, 1024, 16, 32)
    finally:
        r.close()
