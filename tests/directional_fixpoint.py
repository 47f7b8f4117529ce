"""TEST INFRASTRUCTURE — Python restatement of the data-parallel form of directional clustering that
humid_b200/csrc/cluster_big.cu runs on the words of large components (the product never imports this).

The reference (src/cluster.cc:39-87, driven by src/humid.cc:178-189) seeds clusters in rank order: an unassigned
word climbs to the first unassigned neighbour (list order) holding >= 2x its count until none is left (the "top"),
then the cluster takes the closure over unassigned neighbours holding <= half the count.  Call D the directed graph
u -> v  iff  u ~ v and count[u] >= 2 count[v] (counts strictly fall along D, so D is acyclic and shallow).  Then

  (A) between two clusters, an assigned word has only assigned D-descendants (the closure took them);
  (B) hence an unassigned word has only unassigned D-ancestors, so a climb never meets an assigned candidate:
      it follows the STATIC pointer up(x) = first neighbour in list order with count >= 2 count[x];
  (C) the tops are exactly the sources of D (words without such a neighbour), each the top of one cluster;
  (D) a word is taken by the first-processed source among its ancestor sources; the time of source t is the rank
      T(t) of its seed, and label(v) = M(v) = min over ancestor sources a of T(a);
  (E) T(t) = min { s : top(s) = t and M(s) >= s }   (s is still unassigned at its turn).

(E) is solved in rounds over all sources at once.  Every unconfirmed source holds a candidate (the smallest not yet
rejected word of its up-tree, a lower bound of T); M_lo / M_hi = min over ancestor sources of (confirmed T or
candidate) / (confirmed T or infinity) are propagated down D; a candidate s with M_hi(s) < s is rejected (a source
confirmed for an earlier time takes s first), one with M_lo(s) >= s is confirmed.  The source with the smallest
candidate always moves, so the rounds terminate.
"""
import numpy as np

INF = np.iinfo(np.int64).max


def directional_by_fixpoint(count, row_ptr, col, return_rounds=False):
    """-> (label[U] = rank of the cluster's seed, seed_max: dict seed -> top, sizes: dict seed -> Cluster::size)."""
    count = np.asarray(count, dtype=np.int64)
    U = count.size
    row_ptr = np.asarray(row_ptr, dtype=np.int64)
    col = np.asarray(col, dtype=np.int64)
    up = np.full(U, -1, dtype=np.int64)
    for x in range(U):
        for e in range(row_ptr[x], row_ptr[x + 1]):
            if count[col[e]] >= 2 * count[x]:
                up[x] = col[e]
                break
    top = np.arange(U)
    for x in range(U):
        t = x
        while up[t] >= 0:
            t = up[t]
        top[x] = t
    sources = np.nonzero(up < 0)[0]
    # members of every up-tree in rank order, a cursor per source
    order = np.lexsort((np.arange(U), top))
    tree_start = {int(t): int(np.searchsorted(top[order], t, side="left")) for t in sources}
    tree_end = {int(t): int(np.searchsorted(top[order], t, side="right")) for t in sources}
    cursor = dict(tree_start)
    T = {int(t): None for t in sources}
    # words by D-level (longest path from a source) so that one sweep per level propagates exactly
    level = np.zeros(U, dtype=np.int64)
    by_count = np.argsort(-count, kind="stable")
    for v in by_count:  # ancestors hold larger counts: they come first
        for e in range(row_ptr[v], row_ptr[v + 1]):
            u = col[e]
            if count[u] >= 2 * count[v]:
                level[v] = max(level[v], level[u] + 1)
    sweep_order = np.argsort(level, kind="stable")

    def propagate(val):
        m = np.full(U, INF, dtype=np.int64)
        for v in sweep_order:
            if up[v] < 0:
                m[v] = val[v]
                continue
            best = INF
            for e in range(row_ptr[v], row_ptr[v + 1]):
                u = col[e]
                if count[u] >= 2 * count[v]:
                    best = min(best, m[u])
            m[v] = best
        return m

    rounds = 0
    while any(v is None for v in T.values()):
        rounds += 1
        lo = np.full(U, INF, dtype=np.int64)
        hi = np.full(U, INF, dtype=np.int64)
        for t in sources:
            t = int(t)
            if T[t] is not None:
                lo[t] = hi[t] = T[t]
            else:
                lo[t] = order[cursor[t]]
        m_lo, m_hi = propagate(lo), propagate(hi)
        moved = False
        for t in sources:
            t = int(t)
            if T[t] is not None:
                continue
            while True:
                assert cursor[t] < tree_end[t], "a source is always a valid seed of its own tree at the latest"
                s = int(order[cursor[t]])
                if m_hi[s] < s:
                    cursor[t] += 1
                    moved = True
                    continue
                if m_lo[s] >= s:
                    T[t] = s
                    moved = True
                break
        assert moved, "no progress"
    val = np.full(U, INF, dtype=np.int64)
    for t, s in T.items():
        val[t] = s
    label = propagate(val)
    seed_max = {s: t for t, s in T.items()}
    sizes = {}
    for v in range(U):
        sizes[int(label[v])] = sizes.get(int(label[v]), 0) + int(count[v])
    if return_rounds:
        return label, seed_max, sizes, rounds
    return label, seed_max, sizes
