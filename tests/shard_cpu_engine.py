"""TEST INFRASTRUCTURE — a CPU stand-in for `humid_b200.sharded.CudaShardEngine`.

It implements the engine methods of include/humid_gpu_shard.h with numpy (word construction
and clustering come from the oracle), so that the orchestration in humid_b200/sharded.py —
tuple offsets, splitters, all-to-all-v routing, answer routing, rank bases, graph gather — can
be exercised with world_size 2 over gloo on a machine without a GPU.  It is a checker, never a
product path: nothing under humid_b200/ imports it.
"""
from __future__ import annotations

import numpy as np
import torch

import oracle

NONE32 = np.uint32(0xFFFFFFFF)


def _t64(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint64).view(np.int64).copy())


def _t32(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint32).view(np.int32).copy())


def _u64(t):
    return t.numpy().view(np.uint64)


def _u32(t):
    return t.numpy().view(np.uint32)


class CpuShardEngine:
    device = torch.device("cpu")
    stream = 0

    def __init__(self, word_length, distance, seg_take, edit=False, maximum=False, **_):
        assert not (edit and distance >= 2)
        self.n, self.d, self.seg_take, self.maximum = word_length, distance, list(seg_take), maximum
        self.word_length, self.distance = word_length, distance
        self.reset()

    def reset(self):
        self._keys = []
        self.n_tuples = 0

    def add_batch(self, chars, seg_len=None):
        self._keys.append(oracle.pack_words(chars, self.seg_take, seg_len))
        self.n_tuples += chars.shape[0]

    # ---- tuples to their owners ---------------------------------------------------------------
    def begin(self, tuple_offset):
        self.tuple_keys = np.concatenate(self._keys) if self._keys else np.zeros(0, np.uint64)
        self.tuple_offset = tuple_offset
        return int((self.tuple_keys != oracle.INVALID).sum())

    def sample_keys(self, n):
        R = self.tuple_keys.size
        if R == 0:
            return np.zeros(0, np.uint64)
        k = self.tuple_keys[[((i * R + R // 2) // n) % R for i in range(n)]]
        return k[k != oracle.INVALID]

    def partition(self, splitters, n_parts):
        ok = self.tuple_keys != oracle.INVALID
        self._own = np.searchsorted(np.asarray(splitters, np.uint64), self.tuple_keys, "right")
        self._own[~ok] = n_parts  # dropped
        return np.bincount(self._own, minlength=n_parts + 1)[:n_parts].astype(np.int64)

    def export_tuples(self, counts):
        order = np.argsort(self._own, kind="stable")[: int(np.sum(counts))]
        order = order[::-1].copy() if False else order  # any order inside a part is allowed
        return _t64(self.tuple_keys[order]), _t32((order + self.tuple_offset).astype(np.uint32))

    def import_tuples(self, keys, index):
        k, ix = _u64(keys), _u32(index)
        self.rk = k.copy()
        uk, inv, cnt = np.unique(k, return_inverse=True, return_counts=True)
        fst = np.full(uk.size, 0xFFFFFFFF, np.uint32)
        np.minimum.at(fst, inv, ix)
        self.sk, self.sc, self.sf = uk, cnt.astype(np.uint32), fst
        self.n_shard_words = uk.size
        return uk.size

    def set_layout(self, part, n_parts, splitters, bases):
        self.part, self.n_parts = part, n_parts
        self.splitters = np.asarray(splitters, np.uint64)
        self.base = int(bases[part])
        self.all_bases = np.asarray(bases, np.int64)

    # ---- neighbours: Hamming block join -------------------------------------------
    def _blocks(self):
        n, d = self.n, self.d
        out = []
        for b in range(d + 1):
            p0, p1 = b * n // (d + 1), (b + 1) * n // (d + 1)
            lo = 2 * (n - p1)
            out.append(np.uint64(((1 << (2 * (p1 - p0))) - 1) << lo) if p1 > p0 else np.uint64(0))
        return out

    def _join(self, keys, ranks, block):
        """All linked pairs (rank a < rank b) whose FIRST common block is `block`; keys in rank order."""
        masks = self._blocks()
        order = np.argsort(keys & masks[block], kind="stable")
        k, r = keys[order], ranks[order]
        g = k & masks[block]
        starts = np.nonzero(np.concatenate([[True], g[1:] != g[:-1]]))[0]
        ends = np.concatenate([starts[1:], [k.size]])
        low = np.uint64(0x5555555555555555)
        pa, pb = [], []
        for s0, e0 in zip(starts, ends):
            if e0 - s0 < 2:
                continue
            kk, rr = k[s0:e0], r[s0:e0]
            i, j = np.triu_indices(kk.size, 1)
            a, c = kk[i], kk[j]
            x = a ^ c
            diff = (x | (x >> np.uint64(1))) & low
            cnt = np.array([bin(int(v)).count("1") for v in diff])
            ok = (cnt >= 1) & (cnt <= self.d)
            ah, ch, al, cl = (a >> np.uint64(1)) & low, (c >> np.uint64(1)) & low, a & low, c & low
            a_gt = (ah & ~ch) | (~(ah ^ ch) & al & ~cl)
            ok &= (a_gt & diff) == 0
            for bb in range(block):
                ok &= (x & masks[bb]) != 0
            pa.append(rr[i][ok]); pb.append(rr[j][ok])
        if not pa:
            return np.zeros(0, np.uint64), np.zeros(0, np.uint64)
        return np.concatenate(pa).astype(np.uint64), np.concatenate(pb).astype(np.uint64)

    def join_local(self):
        self.entries = []
        if self.d == 0 or self.sk.size == 0:
            return 0
        base = np.uint64(self.base)
        a, b = self._join(self.sk, np.arange(self.sk.size, dtype=np.uint64), 0)
        self.entries = [(a << np.uint64(32)) | (base + b), (b << np.uint64(32)) | (base + a)]
        return 2 * a.size

    def join_export(self, block, n_parts):
        import hashlib  # any deterministic function of the block value will do as the meeting point
        m = self._blocks()[block]
        v = self.sk & m
        dest = np.array([int.from_bytes(hashlib.blake2b(int(x).to_bytes(8, "little"), digest_size=4).digest(), "little") % n_parts
                         for x in v], dtype=np.int64) if v.size else np.zeros(0, np.int64)
        order = np.argsort(dest, kind="stable")
        ranks = (np.arange(self.sk.size, dtype=np.uint32) + np.uint32(self.base))[order]
        return _t64(self.sk[order]), _t32(ranks), np.bincount(dest, minlength=n_parts).astype(np.int64)

    def join_match(self, block, keys, ranks, n_parts):
        a, b = self._join(_u64(keys), _u32(ranks).astype(np.uint64), block)
        ent = np.concatenate([(a << np.uint64(32)) | b, (b << np.uint64(32)) | a])
        rows = (ent >> np.uint64(32)).astype(np.int64)
        owner = np.searchsorted(self.all_bases[1:], rows, "right")
        order = np.argsort(owner, kind="stable")
        return _t64(ent[order]), np.bincount(owner, minlength=n_parts).astype(np.int64)

    def join_import(self, entries):
        e = _u64(entries)
        self.entries.append(((e >> np.uint64(32)) - np.uint64(self.base)) << np.uint64(32) | (e & np.uint64(0xFFFFFFFF)))

    def build_rows(self):
        e = np.sort(np.concatenate(self.entries)) if self.entries else np.zeros(0, np.uint64)
        self.deg = np.bincount((e >> np.uint64(32)).astype(np.int64), minlength=self.n_shard_words).astype(np.uint32)
        self.col = (e & np.uint64(0xFFFFFFFF)).astype(np.uint32)
        return e.size

    # ---- clusters ------------------------------------------------------------------------------
    @staticmethod
    def _cluster(count, deg, col, maximum):
        """oracle clustering of a compact graph -> (cluster id per word, seed word per cluster, size, maxLeaf)."""
        row = np.concatenate([[0], np.cumsum(deg.astype(np.uint64))]).astype(np.uint64)
        cl, size, mx = oracle.cluster_all(count.astype(np.uint64), row, col.astype(np.uint32), maximum)
        _, seed = np.unique(cl, return_index=True)  # the seed is the lowest-ranked member, ids follow seed order
        return cl, seed.astype(np.int64), size, mx

    def cluster_local(self):
        from scipy.sparse import coo_matrix
        from scipy.sparse.csgraph import connected_components

        U, base = self.n_shard_words, self.base
        src = np.repeat(np.arange(U), self.deg)
        loc = (self.col >= base) & (self.col < base + U)
        if U:
            gph = coo_matrix((np.ones(int(loc.sum()), np.int8), (src[loc], (self.col[loc] - base).astype(np.int64))), shape=(U, U))
            _, comp = connected_components(gph, directed=False)
        else:
            comp = np.zeros(0, np.int64)
        comp_open = np.zeros(comp.max() + 1 if U else 0, bool)
        comp_open[comp[src[~loc]]] = True
        self.is_open = comp_open[comp] if U else np.zeros(0, bool)
        closed = np.nonzero(~self.is_open)[0]
        cmap = np.full(U, -1, np.int64)
        cmap[closed] = np.arange(closed.size)
        keep_e = ~self.is_open[src]
        ccol = cmap[(self.col[keep_e] - base).astype(np.int64)]
        cl, seed, size, mx = self._cluster(self.sc[closed], self.deg[closed], ccol, self.maximum)
        # per local word: seed word (local rank) of its cluster; per seed: size, maxLeaf (global), first of maxLeaf
        self.seed_of = np.full(U, -1, np.int64)
        self.seed_of[closed] = closed[seed[cl.astype(np.int64) - 1]] if closed.size else []
        self.seed_info = {int(closed[sd]): (int(size[k]), base + int(closed[mx[k]]), int(self.sf[closed[mx[k]]]))
                          for k, sd in enumerate(seed)}
        self.open_words = np.nonzero(self.is_open)[0]
        self.n_open = self.open_words.size
        return self.n_open, int(self.deg[self.open_words].sum())

    def export_open(self):
        ow = self.open_words
        src = np.repeat(np.arange(self.n_shard_words), self.deg)
        return (_t32((ow + self.base).astype(np.uint32)), _t32(self.sc[ow]), _t32(self.sf[ow]), _t32(self.deg[ow]),
                _t32(self.col[self.is_open[src]]))

    def import_open(self, open_offset, rank, count, first, deg, col):
        rk, cnt, fst, dg, cl_ = _u32(rank), _u32(count), _u32(first), _u32(deg), _u32(col)
        ccol = np.searchsorted(rk, cl_)
        cl, seed, size, mx = self._cluster(cnt, dg, ccol, self.maximum)
        self.open_off, self.n_open_total = open_offset, rk.size
        self.o_seed_of = seed[cl.astype(np.int64) - 1] if rk.size else np.zeros(0, np.int64)  # compact index of the seed
        mine = range(open_offset, open_offset + self.n_open)
        self.o_is_seed = np.zeros(rk.size, bool)
        self.o_is_seed[seed] = True
        for k, sd in enumerate(seed):
            if open_offset <= sd < open_offset + self.n_open:
                self.seed_info[int(rk[sd]) - self.base] = (int(size[k]), int(rk[mx[k]]), int(fst[mx[k]]))
        self.local_seeds = np.array(sorted(self.seed_info), dtype=np.int64)
        self.seed_index = {int(w): i for i, w in enumerate(self.local_seeds)}
        self.o_rank = rk
        return self.local_seeds.size

    def assign_ids(self, id_offset):
        self.id_offset = id_offset
        ids = np.zeros(max(self.n_open_total, 1), np.uint32)
        for i in range(self.open_off, self.open_off + self.n_open):
            if self.o_is_seed[i]:
                ids[i] = id_offset + self.seed_index[int(self.o_rank[i]) - self.base] + 1
        return _t32(ids)

    def finish(self, open_ids, n_tuples, n_usable, n_words, n_clusters):
        ids = _u32(open_ids)
        U = self.n_shard_words
        self.word_cid = np.zeros(U, np.uint32)
        pos = 0
        for r in range(U):
            if self.is_open[r]:
                self.word_cid[r] = ids[self.o_seed_of[self.open_off + pos]]
                pos += 1
            else:
                self.word_cid[r] = self.id_offset + self.seed_index[int(self.seed_of[r])] + 1
        self.totals = dict(total=int(n_tuples), usable=int(n_usable), unique=int(n_words), clusters=int(n_clusters))
        info = [self.seed_info[int(w)] for w in self.local_seeds]
        self.c_size = np.array([x[0] for x in info], np.uint64)
        self.c_max = np.array([x[1] for x in info], np.uint32)
        self.c_first = np.array([x[2] for x in info], np.int64)

    def keep_bits32(self, n_tuples):
        keep = np.zeros(((n_tuples + 31) // 32 + 1) * 32, np.uint8)
        keep[self.c_first] = 1
        return _t32(np.packbits(keep, bitorder="little").view(np.uint32))

    def counts(self):
        return dict(self.totals)

    def histogram(self, which):
        v = {"counts": self.sc.astype(np.uint64), "neigh": self.deg.astype(np.uint64), "clusters": self.c_size}[which]
        k, c = np.unique(v, return_counts=True)
        return k.astype(np.uint64), c.astype(np.uint64)

    def tuple_ids(self):
        return _t32(self.word_cid[np.searchsorted(self.sk, self.rk)])

    def local_words(self):
        return self.sk, self.sc, self.sf, self.word_cid

    def local_rows(self):
        return self.deg, self.col

    def local_clusters(self, n):
        return self.c_size, self.c_max

    def stage_stats(self):
        return []

    def close(self):
        pass
