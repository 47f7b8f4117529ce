"""The data-parallel form of directional clustering (humid_b200/csrc/cluster_big.cu) as restated in
tests/directional_fixpoint.py against the reference's recursion (oracle.cluster_all restates src/cluster.cc:39-87
+ the seed loop of src/humid.cc:178-189): random graphs with sorted, shuffled and duplicated neighbour lists."""
import numpy as np
import pytest

import oracle
from directional_fixpoint import directional_by_fixpoint


def random_graph(rng, U, p, cmax, mode):
    count = rng.integers(1, cmax + 1, size=U)
    adj = [[] for _ in range(U)]
    for a in range(U):
        for b in range(a + 1, U):
            if rng.random() < p:
                adj[a].append(b)
                adj[b].append(a)
    for a in range(U):
        if mode == 0:
            adj[a].sort()  # Hamming lists: ascending
        elif mode == 1:
            rng.shuffle(adj[a])
        else:  # edit-distance lists: repeats, any order
            adj[a] = adj[a] + [x for x in adj[a] if rng.random() < 0.3]
            rng.shuffle(adj[a])
    row_ptr = np.zeros(U + 1, dtype=np.uint64)
    row_ptr[1:] = np.cumsum([len(a) for a in adj])
    col = np.array([x for a in adj for x in a], dtype=np.uint32)
    return count, row_ptr, col


@pytest.mark.parametrize("seed", range(6))
def test_fixpoint_equals_recursion(seed):
    rng = np.random.default_rng(seed)
    for trial in range(40):
        U = int(rng.integers(2, 50))
        count, row_ptr, col = random_graph(rng, U, rng.choice([0.05, 0.1, 0.2, 0.4]), int(rng.choice([2, 3, 5, 20, 1000])),
                                           trial % 3)
        cl, size, mx = oracle.cluster_all(count, row_ptr, col, False)
        label, seed_max, sizes = directional_by_fixpoint(count, row_ptr, col)
        seeds = sorted(set(label.tolist()))
        ids = {s: i + 1 for i, s in enumerate(seeds)}
        assert np.array_equal(np.array([ids[int(x)] for x in label]), cl)
        assert [sizes[s] for s in seeds] == size.tolist()
        assert [seed_max[s] for s in seeds] == mx.tolist()


def test_chain_of_doubling_counts():
    """1 - 2 - 4 - 8 ...: one cluster, top = the far end, whatever the rank order of the words."""
    for order in (list(range(8)), list(range(7, -1, -1)), [3, 0, 6, 1, 7, 2, 5, 4]):
        count = np.array([1 << order[i] for i in range(8)])
        pos = {order[i]: i for i in range(8)}
        adj = [[] for _ in range(8)]
        for k in range(7):
            a, b = pos[k], pos[k + 1]
            adj[a].append(b)
            adj[b].append(a)
        for a in adj:
            a.sort()
        row_ptr = np.zeros(9, dtype=np.uint64)
        row_ptr[1:] = np.cumsum([len(a) for a in adj])
        col = np.array([x for a in adj for x in a], dtype=np.uint32)
        cl, size, mx = oracle.cluster_all(count, row_ptr, col, False)
        label, seed_max, sizes = directional_by_fixpoint(count, row_ptr, col)
        assert len(set(label.tolist())) == 1 == len(size)
        assert seed_max[int(label[0])] == pos[7] == mx[0]
