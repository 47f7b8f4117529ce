"""TEST INFRASTRUCTURE ONLY — ctypes wrapper around oracle/humid_oracle.c (the CPU
restatement of the reference's hot path) and helpers to run the unmodified reference
binaries built into oracle/_ref/.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` leg
may import this package.  The product (humid_b200/, include/, host/) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_build", "liboracle.so")
REF_HUMID = os.path.join(HERE, "_ref", "humid")
REF_PROBE = os.path.join(HERE, "_ref", "trie_probe")
REF_HOTPATH = os.path.join(HERE, "_ref", "hotpath_probe")
REFERENCE_ROOT = "/root/reference"
INVALID = np.uint64(0xFFFFFFFFFFFFFFFF)
MAX_NARROW = 31  # longest word held in one uint64 (the spare top bit keeps INVALID unambiguous)

_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C")
_u16p = np.ctypeslib.ndpointer(np.uint16, flags="C")
_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C")
_u64p = np.ctypeslib.ndpointer(np.uint64, flags="C")


def build(ref: bool = True) -> None:
    """Compile the C restatement and, when /root/reference is present, oracle/_ref."""
    subprocess.run(["make", "-s", "-C", HERE, "oracle"], check=True)
    if ref and os.path.isdir(REFERENCE_ROOT):
        subprocess.run(["make", "-s", "-C", HERE, "-j8", "ref"], check=True)


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        src = os.path.join(HERE, "humid_oracle.c")
        if not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
            build(ref=False)
        L = C.CDLL(LIB_PATH)
        L.oracle_run.restype = C.c_void_p
        L.oracle_run.argtypes = [_u64p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int]
        L.oracle_free.argtypes = [C.c_void_p]
        for name in ("total", "usable", "unique", "clusters", "total_degree"):
            f = getattr(L, "oracle_" + name)
            f.restype = C.c_uint64
            f.argtypes = [C.c_void_p]
        for name in ("keys", "count", "first", "cluster_of", "cluster_size", "cluster_max", "keep", "cid"):
            f = getattr(L, "oracle_" + name)
            f.restype = C.c_void_p
            f.argtypes = [C.c_void_p]
        L.oracle_csr.argtypes = [C.c_void_p, _u64p, _u32p]
        L.oracle_histogram.restype = C.c_uint64
        L.oracle_histogram.argtypes = [C.c_void_p, C.c_int, _u64p, _u64p, C.c_uint64]
        L.oracle_pack_words.argtypes = [_u8p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, _u32p, _u64p]
        L.oracle_pack_words_wide.argtypes = [_u8p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, _u32p, _u64p, _u64p]
        L.oracle_run_wide.restype = C.c_void_p
        L.oracle_run_wide.argtypes = [_u64p, _u64p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int]
        L.oracle_keys_hi.restype = C.c_void_p
        L.oracle_keys_hi.argtypes = [C.c_void_p]
        L.oracle_nt_from_file.argtypes = [C.c_uint64, C.c_uint64, _u64p]
        L.oracle_pre_compute.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64), _u64p]
        L.oracle_extract_umi.argtypes = [C.c_char_p, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.oracle_trie_query.restype = C.c_uint64
        L.oracle_trie_query.argtypes = [_u8p, _u32p, C.c_uint64, _u8p, C.c_uint32, C.c_int, C.c_int, _u64p, C.c_uint64]
        L.oracle_cluster_graph.argtypes = [C.c_uint64, _u64p, _u64p, _u32p, C.c_int, _u32p, C.c_uint64,
                                           C.c_void_p, _u32p, _u64p, _u32p]
        L.oracle_cluster_all.restype = C.c_uint64
        L.oracle_cluster_all.argtypes = [C.c_uint64, _u64p, _u64p, _u32p, C.c_int, _u32p, _u64p, _u32p]
        L.oracle_max_neighbour.restype = C.c_uint32
        L.oracle_max_neighbour.argtypes = [C.c_uint64, _u64p, _u64p, _u32p, C.c_void_p, C.c_uint32]
        L.oracle_at_least_double.restype = C.c_int
        L.oracle_at_least_double.argtypes = [C.c_uint64, C.c_uint64]
        _lib = L
    return _lib


def _view(ptr: int, n: int, dtype) -> np.ndarray:
    if n == 0:
        return np.zeros(0, dtype=dtype)
    buf = (C.c_char * (n * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=n).copy()


# ---- word construction -----------------------------------------------------------------
def nt_from_file(files: int, length: int) -> List[int]:
    out = np.zeros(files, dtype=np.uint64)
    lib().oracle_nt_from_file(files, length, out)
    return [int(x) for x in out]


def pre_compute(umi_size_first: int, word_length: int, files: int) -> Tuple[int, List[int]]:
    out = np.zeros(files, dtype=np.uint64)
    h = C.c_uint64(0)
    lib().oracle_pre_compute(umi_size_first, word_length, files, C.byref(h), out)
    return int(h.value), [int(x) for x in out]


def extract_umi(header: str) -> str:
    b = header.encode()
    off, ln = C.c_uint64(0), C.c_uint64(0)
    lib().oracle_extract_umi(b, len(b), C.byref(off), C.byref(ln))
    return b[off.value: off.value + ln.value].decode()


def pack_words(chars: np.ndarray, seg_take: Sequence[int], seg_len: Optional[np.ndarray] = None):
    """chars uint8[R, n] -> packed keys uint64[R] (INVALID where filtered); for n > 31 a pair
    (lo, hi) of uint64[R] (word = hi << 64 | lo, both INVALID where filtered)."""
    chars = np.ascontiguousarray(chars, dtype=np.uint8)
    R, n = chars.shape
    take = np.asarray(seg_take, dtype=np.uint32)
    assert int(take.sum()) == n
    out = np.zeros(R, dtype=np.uint64)
    sl = None
    if seg_len is not None:
        seg_len = np.ascontiguousarray(seg_len, dtype=np.uint16)
        assert seg_len.shape == (R, take.size)
        sl = seg_len.ctypes.data_as(C.c_void_p)
    if n > MAX_NARROW:
        hi = np.zeros(R, dtype=np.uint64)
        lib().oracle_pack_words_wide(chars, sl, R, n, take.size, take, out, hi)
        return out, hi
    lib().oracle_pack_words(chars, sl, R, n, take.size, take, out)
    return out


# ---- whole path ------------------------------------------------------------------------
class Result:
    """Everything the reference computes between readData and writeStatistics."""

    def __init__(self, keys, n: int, d: int, edit: bool, maximum: bool, stop_after: int = 0):
        """keys: uint64[R], or for words longer than 31 letters a (lo, hi) pair of uint64[R]."""
        L = lib()
        if isinstance(keys, tuple):
            keys, keys_hi = (np.ascontiguousarray(k, dtype=np.uint64) for k in keys)
            h = L.oracle_run_wide(keys, keys_hi, keys.size, n, d, int(edit), int(maximum), stop_after)
        else:
            keys = np.ascontiguousarray(keys, dtype=np.uint64)
            h = L.oracle_run(keys, keys.size, n, d, int(edit), int(maximum), stop_after)
        try:
            self.total = L.oracle_total(h)
            self.usable = L.oracle_usable(h)
            self.unique = U = L.oracle_unique(h)
            self.clusters = Cn = L.oracle_clusters(h)
            self.keys = _view(L.oracle_keys(h), U, np.uint64)
            self.keys_hi = _view(L.oracle_keys_hi(h), U, np.uint64)
            self.count = _view(L.oracle_count(h), U, np.uint64)
            self.first = _view(L.oracle_first(h), U, np.uint64)
            self.cluster_of = _view(L.oracle_cluster_of(h), U, np.uint32)
            self.keep = _view(L.oracle_keep(h), keys.size, np.uint8)
            self.cid = _view(L.oracle_cid(h), keys.size, np.uint32)
            self.hist: Dict[str, Tuple[np.ndarray, np.ndarray]] = {}
            if stop_after != 1:
                td = L.oracle_total_degree(h)
                self.row_ptr = np.zeros(U + 1, dtype=np.uint64)
                self.col = np.zeros(max(td, 1), dtype=np.uint32)
                L.oracle_csr(h, self.row_ptr, self.col)
                self.col = self.col[:td]
            if stop_after == 0:
                self.cluster_size = _view(L.oracle_cluster_size(h), Cn, np.uint64)
                self.cluster_max = _view(L.oracle_cluster_max(h), Cn, np.uint32)
            which = {"counts": 0, "neigh": 1, "clusters": 2}
            for name, w in which.items():
                if (w == 1 and stop_after == 1) or (w == 2 and stop_after != 0):
                    continue
                cap = max(U, Cn, 1)
                hk = np.zeros(cap, dtype=np.uint64)
                hv = np.zeros(cap, dtype=np.uint64)
                m = L.oracle_histogram(h, w, hk, hv, cap)
                self.hist[name] = (hk[:m].copy(), hv[:m].copy())
        finally:
            L.oracle_free(h)

    def neighbours(self, r: int) -> np.ndarray:
        return self.col[int(self.row_ptr[r]): int(self.row_ptr[r + 1])]

    def stats_text(self) -> Dict[str, str]:
        """The four .dat files exactly as writeStatistics (src/humid.cc:328-357) prints them."""
        out = {}
        for name, f in (("counts", "counts.dat"), ("neigh", "neigh.dat"), ("clusters", "clusters.dat")):
            k, v = self.hist[name]
            out[f] = "".join(f"{int(a)} {int(b)}\n" for a, b in zip(k, v))
        out["stats.dat"] = (f"total: {self.total}\nusable: {self.usable}\n"
                            f"unique: {self.unique}\nclusters: {self.clusters}\n")
        return out


def trie_query(words: Sequence[Sequence[int]], query: Sequence[int], distance: int, kind: str) -> List[int]:
    """kind in {hamming, asymmetricHamming, levenshtein, asymmetricLevenshtein}; returns the
    indices (into `words`) of the yielded leaves in yield order."""
    kinds = {"hamming": 0, "asymmetricHamming": 1, "levenshtein": 2, "asymmetricLevenshtein": 3}
    letters = np.array([c for w in words for c in w] or [0], dtype=np.uint8)
    lens = np.array([len(w) for w in words], dtype=np.uint32)
    q = np.array(list(query) or [0], dtype=np.uint8)
    cap = 1 << 16
    out = np.zeros(cap, dtype=np.uint64)
    m = lib().oracle_trie_query(letters, lens, len(words), q, len(query), distance, kinds[kind], out, cap)
    assert m <= cap
    return [int(x) for x in out[:m]]


def cluster_graph(count, adj: Sequence[Sequence[int]], maximum: bool, seeds: Sequence[int],
                  pre_assigned: Optional[Sequence[int]] = None):
    U = len(count)
    count = np.asarray(count, dtype=np.uint64)
    row_ptr = np.zeros(U + 1, dtype=np.uint64)
    row_ptr[1:] = np.cumsum([len(a) for a in adj])
    col = np.array([x for a in adj for x in a] or [0], dtype=np.uint32)
    seeds_a = np.asarray(seeds, dtype=np.uint32)
    cl = np.zeros(U, dtype=np.uint32)
    size = np.zeros(len(seeds), dtype=np.uint64)
    mx = np.zeros(len(seeds), dtype=np.uint32)
    pa = None
    if pre_assigned is not None:
        pa_arr = np.asarray(pre_assigned, dtype=np.uint8)
        pa = pa_arr.ctypes.data_as(C.c_void_p)
    lib().oracle_cluster_graph(U, count, row_ptr, col, int(maximum), seeds_a, len(seeds), pa, cl, size, mx)
    return cl, size, mx


def cluster_all(count, row_ptr, col, maximum: bool):
    """findClusters on an explicit graph: (cluster id per word, size per cluster, maxLeaf per cluster)."""
    count = np.ascontiguousarray(count, dtype=np.uint64)
    row_ptr = np.ascontiguousarray(row_ptr, dtype=np.uint64)
    U = count.size
    col = np.ascontiguousarray(col if len(col) else [0], dtype=np.uint32)
    cl = np.zeros(max(U, 1), dtype=np.uint32)
    size = np.zeros(max(U, 1), dtype=np.uint64)
    mx = np.zeros(max(U, 1), dtype=np.uint32)
    n = int(lib().oracle_cluster_all(U, count if U else np.zeros(1, np.uint64), row_ptr, col, int(maximum), cl, size, mx))
    return cl[:U], size[:n], mx[:n]


def max_neighbour(count, adj, assigned, leaf: int) -> int:
    U = len(count)
    count = np.asarray(count, dtype=np.uint64)
    row_ptr = np.zeros(U + 1, dtype=np.uint64)
    row_ptr[1:] = np.cumsum([len(a) for a in adj])
    col = np.array([x for a in adj for x in a] or [0], dtype=np.uint32)
    a = np.asarray(assigned, dtype=np.uint8)
    return int(lib().oracle_max_neighbour(U, count, row_ptr, col, a.ctypes.data_as(C.c_void_p), leaf))


def at_least_double(a: int, b: int) -> bool:
    return bool(lib().oracle_at_least_double(a, b))


# ---- the unmodified reference, built into oracle/_ref ----------------------------------
def have_ref() -> bool:
    return os.access(REF_HUMID, os.X_OK) and os.access(REF_PROBE, os.X_OK)


def key_to_word(key: int, n: int) -> str:
    return "".join("ACGT"[(int(key) >> (2 * (n - 1 - p))) & 3] for p in range(n))


def word_to_key(word: str) -> int:
    k = 0
    for ch in word:
        k = (k << 2) | "ACGT".index(ch)
    return k


def run_ref_probe(words: Sequence[str], counts: Sequence[int], edit: bool, distance: int, maximum: bool,
                  queries: Sequence[Tuple[str, int, str]] = ()):
    """Run the reference trie + cluster code (oracle/_ref/trie_probe) on a word multiset.

    Returns (leaves, clusters, query_results): leaves = list of dicts in walk order with
    word/count/cluster/neighbours (ranks, in the reference's order, with multiplicity)."""
    lines = [f"{'E' if edit else 'H'} {distance} {'X' if maximum else 'D'}"]
    lines += [f"{w} {c}" for w, c in zip(words, counts)]
    lines += [f"?{k} {d} {w}" for (k, d, w) in queries]
    p = subprocess.run([REF_PROBE], input="\n".join(lines) + "\n", capture_output=True, text=True, check=True)
    leaves, clusters, qres = [], [], []
    for ln in p.stdout.splitlines():
        t = ln.split()
        if t[0] == "W":
            deg = int(t[5])
            leaves.append(dict(rank=int(t[1]), word=t[2], count=int(t[3]), cluster=int(t[4]),
                               neighbours=[int(x) for x in t[6:6 + deg]]))
        elif t[0] == "C":
            clusters.append(dict(id=int(t[1]), size=int(t[2]), max_leaf=int(t[3])))
        elif t[0] == "Q":
            qres.append(["" if x == "-" else x for x in t[3:3 + int(t[2])]])
    return leaves, clusters, qres


def run_ref_humid(files: Sequence[str], out_dir: str, flags: Sequence[str], stats: bool = True,
                  annotate: bool = False, timeout: Optional[float] = None) -> Dict[str, float]:
    """Run the unmodified reference CLI; returns {'wall_s': …}.  `ulimit -s unlimited` per
    docs/troubleshooting.rst:17-18 (recursive clustering)."""
    import resource
    import time

    os.makedirs(out_dir, exist_ok=True)
    cmd = [REF_HUMID, *flags, "-d", out_dir, "-l", os.path.join(out_dir, "humid.log")]
    if stats:
        cmd.append("-s")
    if annotate:
        cmd.append("-a")
    cmd += list(files)

    def pre():
        try:
            resource.setrlimit(resource.RLIMIT_STACK, (resource.RLIM_INFINITY, resource.RLIM_INFINITY))
        except (ValueError, OSError):
            pass

    t0 = time.monotonic()
    subprocess.run(cmd, check=True, preexec_fn=pre, timeout=timeout)
    return {"wall_s": time.monotonic() - t0}


def run_ref_hotpath(chars: np.ndarray, header_umi: int, nt_to_take: Sequence[int], distance: int, edit: bool,
                    maximum: bool, work_dir: str, timeout: Optional[float] = None) -> Dict:
    """The reference's hot path from host memory (oracle/_ref/hotpath_probe = oracle/ref_hotpath.cc around the
    unmodified src/humid.cc): word characters in the C ABI's layout in, counts / keep mask / statistics files and
    seconds per phase out.  Returns the probe's JSON plus 'keep' (bool[n]), 'stats' ({file: text}) and 'wall_s'."""
    import json
    import resource
    import time

    chars = np.ascontiguousarray(chars, dtype=np.uint8)
    n, wl = chars.shape
    os.makedirs(work_dir, exist_ok=True)
    path = os.path.join(work_dir, "chars.bin")
    chars.tofile(path)
    prefix = os.path.join(work_dir, "ref")
    cmd = [REF_HOTPATH, path, str(n), str(wl), str(distance), str(int(edit)), str(int(maximum)), str(header_umi), prefix,
           *[str(int(t)) for t in nt_to_take]]

    def pre():
        try:
            resource.setrlimit(resource.RLIMIT_STACK, (resource.RLIM_INFINITY, resource.RLIM_INFINITY))
        except (ValueError, OSError):
            pass

    t0 = time.monotonic()
    p = subprocess.run(cmd, check=True, preexec_fn=pre, timeout=timeout, capture_output=True, text=True)
    out = json.loads(p.stdout.strip().splitlines()[-1])
    out["wall_s"] = time.monotonic() - t0
    out["keep"] = np.unpackbits(np.fromfile(prefix + ".keep", dtype=np.uint8), bitorder="little")[:n].astype(bool)
    out["stats"] = {f: open(prefix + "." + f).read() for f in ("counts.dat", "neigh.dat", "clusters.dat")}
    out["stats"]["stats.dat"] = (f"total: {out['total']}\nusable: {out['usable']}\nunique: {out['unique']}\n"
                                 f"clusters: {out['clusters']}\n")
    return out


def get_nucleotides(header: str, seqs: Sequence[str], nt_to_take: Sequence[int], header_umi_size: int) -> str:
    """`getNucleotides` (src/fastq.cc:116-144) through the boundary layout: the host part
    (UMI location, copy of min(len, take) chars) is done here the way the host does it."""
    segs, takes = [], []
    if header_umi_size > 0:
        segs.append(extract_umi(header))
        takes.append(header_umi_size)
    for s, t in zip(seqs, nt_to_take):
        segs.append(s)
        takes.append(t)
    n = sum(takes)
    chars = np.zeros(max(n, 1), dtype=np.uint8)
    seg_len = np.zeros(max(len(takes), 1), dtype=np.uint16)
    p = 0
    for i, (s, t) in enumerate(zip(segs, takes)):
        m = min(len(s), t)
        chars[p:p + m] = np.frombuffer(s[:m].encode(), dtype=np.uint8)
        seg_len[i] = m
        p += t
    out = np.zeros(max(n, 1), dtype=np.uint8)
    L = lib()
    L.oracle_expand_word.argtypes = [_u8p, _u16p, C.c_uint32, _u32p, _u8p]
    L.oracle_expand_word(chars, seg_len, len(takes), np.asarray(takes or [0], dtype=np.uint32), out)
    return bytes(out[:n]).decode()
