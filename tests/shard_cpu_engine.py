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
        assert not (edit and distance >= 2) and distance in (0, 1, 2)
        self.n, self.d, self.seg_take, self.maximum = word_length, distance, list(seg_take), maximum
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

    # ---- neighbours ----------------------------------------------------------------------------
    def _variants(self):
        """(source shard-local rank, upward variant) for every variant within distance d."""
        w = self.sk
        n = self.n
        src, var = [], []
        r = np.arange(w.size, dtype=np.uint32)

        def letter(x, p):
            return (x >> np.uint64(2 * (n - 1 - p))) & np.uint64(3)

        for p1 in range(n):
            for a in range(1, 4):
                m1 = letter(w, p1) + np.uint64(a) <= 3
                v1 = w[m1] + (np.uint64(a) << np.uint64(2 * (n - 1 - p1)))
                src.append(r[m1]); var.append(v1)
                if self.d >= 2:
                    for p2 in range(p1 + 1, n):
                        for b in range(1, 4):
                            m2 = letter(v1, p2) + np.uint64(b) <= 3
                            src.append(r[m1][m2]); var.append(v1[m2] + (np.uint64(b) << np.uint64(2 * (n - 1 - p2))))
        if not src or self.d == 0:
            return np.zeros(0, np.uint32), np.zeros(0, np.uint64)
        return np.concatenate(src), np.concatenate(var)

    def _lookup(self, keys):
        pos = np.searchsorted(self.sk, keys)
        pos_c = np.minimum(pos, max(self.sk.size - 1, 0))
        hit = (self.sk[pos_c] == keys) if self.sk.size else np.zeros(keys.size, bool)
        return np.where(hit, pos_c, 0xFFFFFFFF).astype(np.uint32)

    def probe_count(self, n_parts):
        self._src, self._var = self._variants()
        self._owner = np.searchsorted(self.splitters, self._var, "right") if self.n_parts > 1 else np.zeros(self._var.size, int)
        return np.bincount(self._owner[self._owner != self.part], minlength=n_parts).astype(np.int64)

    def probe(self, counts):
        local = self._owner == self.part
        hit = self._lookup(self._var[local])
        ok = hit != NONE32
        a, b = self._src[local][ok].astype(np.uint64), hit[ok].astype(np.uint64)
        base = np.uint64(self.base)
        self.entries = [(a << np.uint64(32)) | (base + b), (b << np.uint64(32)) | (base + a)]
        order = np.argsort(self._owner[~local], kind="stable")
        qk = self._var[~local][order]
        qs = (self._src[~local][order] + np.uint32(self.base)).astype(np.uint32)
        return _t64(qk), _t32(qs)

    def answer(self, qkeys, qsrc):
        hit = self._lookup(_u64(qkeys))
        ok = hit != NONE32
        self.entries.append((hit[ok].astype(np.uint64) << np.uint64(32)) | _u32(qsrc)[ok].astype(np.uint64))
        return _t32(np.where(ok, hit + np.uint32(self.base), NONE32))

    def apply_answers(self, qsrc, answers):
        a, s = _u32(answers), _u32(qsrc)
        ok = a != NONE32
        self.entries.append(((s[ok] - np.uint32(self.base)).astype(np.uint64) << np.uint64(32)) | a[ok].astype(np.uint64))

    def build_rows(self):
        e = np.sort(np.concatenate(self.entries)) if self.entries else np.zeros(0, np.uint64)
        self.deg = np.bincount((e >> np.uint64(32)).astype(np.int64), minlength=self.n_shard_words).astype(np.uint32)
        self.col = (e & np.uint64(0xFFFFFFFF)).astype(np.uint32)
        return e.size

    def export_graph(self):
        return _t64(self.sk), _t32(self.sc), _t32(self.sf), _t32(self.deg), _t32(self.col)

    # ---- whole job ----------------------------------------------------------------------------
    def import_graph(self, n_tuples, n_usable, keys, counts, first, degree, col):
        self.g_keys, self.g_count, self.g_first = _u64(keys).copy(), _u32(counts).copy(), _u32(first).copy()
        deg = _u32(degree).astype(np.uint64)
        self.g_row = np.concatenate([[0], np.cumsum(deg)]).astype(np.uint64)
        self.g_col = _u32(col).copy()
        self.total, self.usable = n_tuples, n_usable
        self.g_cluster, self.c_size, self.c_max = oracle.cluster_all(self.g_count, self.g_row, self.g_col, self.maximum)
        self.keep = np.zeros(n_tuples, bool)
        self.keep[self.g_first[self.c_max]] = True

    def counts(self):
        return dict(total=int(self.total), usable=int(self.usable), unique=int(self.g_keys.size),
                    clusters=int(self.c_size.size))

    def histogram(self, which):
        v = {"counts": self.g_count.astype(np.uint64), "neigh": np.diff(self.g_row), "clusters": self.c_size}[which]
        k, c = np.unique(v, return_counts=True)
        return k.astype(np.uint64), c.astype(np.uint64)

    def stats_text(self):
        out = {}
        for name, f in (("counts", "counts.dat"), ("neigh", "neigh.dat"), ("clusters", "clusters.dat")):
            k, v = self.histogram(name)
            out[f] = "".join(f"{int(a)} {int(b)}\n" for a, b in zip(k, v))
        c = self.counts()
        out["stats.dat"] = (f"total: {c['total']}\nusable: {c['usable']}\nunique: {c['unique']}\n"
                            f"clusters: {c['clusters']}\n")
        return out

    def global_keep_bits(self, first, n):
        return self.keep[first:first + n]

    def tuple_ids(self):
        lr = np.searchsorted(self.sk, self.rk)
        return _t32(self.g_cluster[self.base + lr])

    def words(self):
        return self.g_keys, self.g_count, self.g_first, self.g_cluster

    def neighbours(self):
        return self.g_row, self.g_col

    def clusters(self):
        return self.c_size, self.c_max

    def stage_stats(self):
        return []

    def close(self):
        pass
