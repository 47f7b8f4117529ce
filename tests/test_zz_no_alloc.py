"""humid_gpu_partition_by_slot_noalloc — the slot partition with its scratch in a static device array (no
counterpart in the reference, include/humid_gpu.h).

Through the C ABI, twice in a row on one handle with different sizes and tables, against the existing entry point
humid_gpu_debug_partition (kind 0) on the same inputs.  Integer / index work: the bar is bit-exact.  The second call
runs over the scratch the first one left behind, so stale status words or tickets would show.  Collected last on
purpose; passes alone in a fresh process and after the other tests.
"""
import numpy as np
import pytest
import torch

from humid_b200 import api

pytestmark = pytest.mark.gpu


def run_both(dd, keys: np.ndarray, mask: int, base: int):
    n = keys.size
    k_in = torch.from_numpy(keys.view(np.int64)).cuda()
    ref_k, new_k = torch.empty_like(k_in), torch.zeros_like(k_in)
    ref_v = torch.empty(n, dtype=torch.int32, device="cuda")
    new_v = torch.zeros(n, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    dd.debug_partition(0, k_in.data_ptr(), n, mask, 0, None, base, ref_k.data_ptr(), ref_v.data_ptr())  # synchronises
    free_before = torch.cuda.mem_get_info()[0]
    dd.partition_by_slot_noalloc(k_in.data_ptr(), n, mask, base, new_k.data_ptr(), new_v.data_ptr())  # enqueues only
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    return (ref_k.cpu().numpy(), ref_v.cpu().numpy()), (new_k.cpu().numpy(), new_v.cpu().numpy()), free_before - free_after


def test_two_calls_in_a_row_match_the_allocating_entry_point():
    rng = np.random.default_rng(20261105)
    with api.Dedup(word_length=8, distance=1, seg_take=[4, 4]) as dd:
        # first call: many tiles; second: fewer tiles, another table size and index base, over the first call's leftovers;
        # third: the largest size the static scratch allows
        for n, mask, base in ((1_000_003, (1 << 22) - 1, 7), (70_001, (1 << 17) - 1, 123_456),
                              (api.NOALLOC_MAX_ELEMENTS, (1 << 26) - 1, 0)):
            keys = rng.integers(0, 1 << 48, size=n, dtype=np.uint64)
            keys[::5] = keys[0]  # runs of equal keys: stability shows
            (rk, rv), (nk, nv), grew = run_both(dd, keys, mask, base)
            assert np.array_equal(nk, rk), (n, "keys")
            assert np.array_equal(nv, rv), (n, "values")
            assert np.array_equal(np.sort(nk.view(np.uint64)), np.sort(keys)), (n, "not a permutation of the input")
            assert grew <= 0, f"device memory in use grew by {grew} bytes across a call that must not allocate"


def test_sizes_beyond_the_static_scratch_are_refused():
    with api.Dedup(word_length=8, distance=1, seg_take=[4, 4]) as dd:
        n = api.NOALLOC_MAX_ELEMENTS + 1
        k = torch.zeros(n, dtype=torch.int64, device="cuda")
        v = torch.zeros(n, dtype=torch.int32, device="cuda")
        with pytest.raises(api.HumidGpuError) as e:
            dd.partition_by_slot_noalloc(k.data_ptr(), n, 1023, 0, k.data_ptr(), v.data_ptr())
        assert "(-4)" in str(e.value) and "at most" in str(e.value)
        dd.partition_by_slot_noalloc(0, 0, 1023, 0, 0, 0)  # nothing to do is not an error
