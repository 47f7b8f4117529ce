"""Device primitives (humid_b200/csrc/primitives.cu) against numpy, at sizes from one element to
many more tiles than are resident at once."""
import numpy as np
import pytest
import torch

from humid_b200 import api

pytestmark = pytest.mark.gpu

M64 = (1 << 64) - 1


def mix64(x: np.ndarray) -> np.ndarray:  # common.cuh
    x = x.astype(np.uint64).copy()
    with np.errstate(over="ignore"):
        x ^= x >> np.uint64(30)
        x *= np.uint64(0xBF58476D1CE4E5B9)
        x ^= x >> np.uint64(27)
        x *= np.uint64(0x94D049BB133111EB)
        x ^= x >> np.uint64(31)
    return x


@pytest.fixture(scope="module")
def dd():
    d = api.Dedup(word_length=8, distance=1, seg_take=[4, 4])
    yield d
    d.close()


def dev(a, dtype):
    return torch.from_numpy(a.view(dtype)).cuda()


SIZES = [1, 2, 31, 4096, 4097, 6144, 6145, 50_000, 3_000_001]  # 6144 = one radix tile, 4096 = one scan tile


@pytest.mark.parametrize("n", SIZES)
@pytest.mark.parametrize("bits", [(0, 48), (0, 12), (24, 48), (5, 19), (60, 64)])
@pytest.mark.parametrize("with_vals", [False, True])
def test_sort_pairs(dd, n, bits, with_vals):
    rng = np.random.default_rng(n + bits[0])
    keys = rng.integers(0, 1 << 62, size=n, dtype=np.uint64)
    if n > 10:
        keys[::3] = keys[0]  # many equal keys: stability shows
    vals = rng.integers(0, 1 << 32, size=n, dtype=np.uint32) if with_vals else np.arange(n, dtype=np.uint32)
    k_in, v_in = dev(keys, np.int64), dev(vals, np.int32)
    k_out, v_out = torch.empty_like(k_in), torch.empty_like(v_in)
    torch.cuda.synchronize()
    dd.debug_sort_pairs(k_in.data_ptr(), v_in.data_ptr() if with_vals else 0, n, bits[0], bits[1], k_out.data_ptr(),
                        v_out.data_ptr())
    field = (keys >> np.uint64(bits[0])) & np.uint64((1 << (bits[1] - bits[0])) - 1)
    order = np.argsort(field, kind="stable")
    assert np.array_equal(k_out.cpu().numpy().view(np.uint64), keys[order])
    assert np.array_equal(v_out.cpu().numpy().view(np.uint32), vals[order])


@pytest.mark.parametrize("n", SIZES)
def test_partition_by_block_hash(dd, n):
    rng = np.random.default_rng(n)
    keys = rng.integers(0, 1 << 48, size=n, dtype=np.uint64)
    mask, parts, base = 0xFFFFFF, 5, 1000
    k_in = dev(keys, np.int64)
    k_out, v_out = torch.empty_like(k_in), torch.empty(n, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    counts = dd.debug_partition(1, k_in.data_ptr(), n, mask, parts, None, base, k_out.data_ptr(), v_out.data_ptr())
    dest = (((mix64(keys & np.uint64(mask)) >> np.uint64(32)) * np.uint64(parts)) >> np.uint64(32)).astype(np.int64)
    order = np.argsort(dest, kind="stable")
    assert np.array_equal(counts.astype(np.int64), np.bincount(dest, minlength=parts))
    assert np.array_equal(k_out.cpu().numpy().view(np.uint64), keys[order])
    assert np.array_equal(v_out.cpu().numpy().view(np.uint32), (np.arange(n) + base).astype(np.uint32)[order])


@pytest.mark.parametrize("n", SIZES)
def test_partition_by_row_owner(dd, n):
    rng = np.random.default_rng(n)
    bases = np.array([0, 10, 10, 5000, 100000, 2_000_000], dtype=np.uint64)  # one empty part
    rows = rng.integers(0, int(bases[-1]), size=n, dtype=np.uint64)
    ent = (rows << np.uint64(32)) | rng.integers(0, 1 << 32, size=n, dtype=np.uint64)
    e_in = dev(ent, np.int64)
    e_out = torch.empty_like(e_in)
    torch.cuda.synchronize()
    counts = dd.debug_partition(2, e_in.data_ptr(), n, 0, 5, bases, 0, e_out.data_ptr(), 0)
    owner = np.searchsorted(bases[1:], rows, "right")
    order = np.argsort(owner, kind="stable")
    assert np.array_equal(counts.astype(np.int64), np.bincount(owner, minlength=5))
    assert np.array_equal(e_out.cpu().numpy().view(np.uint64), ent[order])


@pytest.mark.parametrize("n", [1, 4097, 3_000_001])
def test_partition_by_slot(dd, n):
    rng = np.random.default_rng(n)
    keys = rng.integers(0, 1 << 48, size=n, dtype=np.uint64)
    mask = (1 << 22) - 1
    k_in = dev(keys, np.int64)
    k_out, v_out = torch.empty_like(k_in), torch.empty(n, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    dd.debug_partition(0, k_in.data_ptr(), n, mask, 0, None, 7, k_out.data_ptr(), v_out.data_ptr())
    region = ((mix64(keys) & np.uint64(mask)) >> np.uint64(14)).astype(np.int64)
    order = np.argsort(region, kind="stable")
    assert np.array_equal(k_out.cpu().numpy().view(np.uint64), keys[order])
    assert np.array_equal(v_out.cpu().numpy().view(np.uint32), (np.arange(n) + 7).astype(np.uint32)[order])
