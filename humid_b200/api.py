"""ctypes binding of libhumid_gpu (include/humid_gpu.h) — the Python host side of the hot path.

The reference has no plugin interface for this path: `humid()` (reference src/humid.cc:369-409)
drives a `Trie<4,NLeaf>` directly.  `Dedup` below mirrors that sequence of calls —
readData -> find*Neighbours -> findClusters -> writeFiltered/writeAnnotated/runStatistics —
with every step executed by the CUDA library.  There is no CPU implementation here: if the
shared library is missing or no CUDA device is usable, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HUMID_GPU_LIB") or os.path.join(HERE, "libhumid_gpu.so")  # (override: A/B runs of two builds)
MAX_SEGMENTS = 16
MAX_WORD_LENGTH = 63
MAX_NARROW_WORD_LENGTH = 31  # longer words: two 64-bit keys, Hamming / edit distance 1, one GPU
HIST = {"counts": 0, "neigh": 1, "clusters": 2}

EXPORTS = [
    "humid_gpu_abi_version", "humid_gpu_create", "humid_gpu_destroy", "humid_gpu_reset", "humid_gpu_last_error",
    "humid_gpu_add_batch", "humid_gpu_add_batch_device", "humid_gpu_finalize", "humid_gpu_counts",
    "humid_gpu_histogram", "humid_gpu_keep_mask", "humid_gpu_cluster_ids", "humid_gpu_debug_words",
    "humid_gpu_debug_neighbours", "humid_gpu_debug_clusters", "humid_gpu_stage_stats", "humid_gpu_stream",
    "humid_gpu_debug_sort_pairs", "humid_gpu_debug_partition", "humid_gpu_host_alloc", "humid_gpu_host_free",
    "humid_gpu_debug_words_hi", "humid_gpu_host_register", "humid_gpu_host_unregister",
    "humid_gpu_partition_by_slot_noalloc", "humid_gpu_add_batch_packed4",
]
NOALLOC_MAX_ELEMENTS = 1 << 22  # HUMID_NOALLOC_MAX_ELEMENTS


class HumidGpuError(RuntimeError):
    pass


class Config(C.Structure):
    _fields_ = [
        ("word_length", C.c_uint32), ("distance", C.c_uint32), ("n_segments", C.c_uint32),
        ("seg_take", C.c_uint32 * MAX_SEGMENTS), ("edit", C.c_uint8), ("maximum", C.c_uint8),
        ("defer_table", C.c_uint8), ("reserved", C.c_uint8), ("device", C.c_int32), ("expected_tuples", C.c_uint64),
    ]


class StageStat(C.Structure):
    _fields_ = [("name", C.c_char * 32), ("ms", C.c_float), ("launches", C.c_uint32), ("units", C.c_uint64),
                ("algo_bytes", C.c_uint64)]


_lib = None


def load_library() -> C.CDLL:
    """dlopen the in-tree library (never builds, never falls back)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise HumidGpuError(f"{LIB_PATH} is missing: run `python -m humid_b200.build` (needs nvcc); "
                            "there is no CPU fallback for the deduplication hot path")
    L = C.CDLL(LIB_PATH)
    vp, u64, u32 = C.c_void_p, C.c_uint64, C.c_uint32
    P = C.POINTER
    L.humid_gpu_abi_version.restype = C.c_int
    L.humid_gpu_create.argtypes = [P(Config), P(vp)]
    L.humid_gpu_destroy.argtypes = [vp]
    L.humid_gpu_destroy.restype = None
    L.humid_gpu_reset.argtypes = [vp]
    L.humid_gpu_last_error.argtypes = [vp]
    L.humid_gpu_last_error.restype = C.c_char_p
    L.humid_gpu_add_batch.argtypes = [vp, vp, vp, u64, u64]
    L.humid_gpu_add_batch_device.argtypes = [vp, vp, vp, u64, u64]
    L.humid_gpu_add_batch_packed4.argtypes = [vp, vp, vp, u64, u64]
    L.humid_gpu_finalize.argtypes = [vp]
    L.humid_gpu_counts.argtypes = [vp, P(u64), P(u64), P(u64), P(u64)]
    L.humid_gpu_histogram.argtypes = [vp, C.c_int, vp, vp, P(u64)]
    L.humid_gpu_keep_mask.argtypes = [vp, u64, u64, vp]
    L.humid_gpu_cluster_ids.argtypes = [vp, u64, u64, vp]
    L.humid_gpu_debug_words.argtypes = [vp, u64, vp, vp, vp, vp]
    L.humid_gpu_debug_words_hi.argtypes = [vp, u64, vp]
    L.humid_gpu_debug_neighbours.argtypes = [vp, vp, vp, P(u64)]
    L.humid_gpu_debug_clusters.argtypes = [vp, u64, vp, vp]
    L.humid_gpu_stage_stats.argtypes = [vp, P(StageStat), P(u32)]
    L.humid_gpu_stream.argtypes = [vp]
    L.humid_gpu_stream.restype = vp
    L.humid_gpu_host_alloc.argtypes = [vp, C.c_size_t, P(vp)]
    L.humid_gpu_host_free.argtypes = [vp, vp]
    L.humid_gpu_host_register.argtypes = [vp, vp, C.c_size_t]
    L.humid_gpu_host_unregister.argtypes = [vp, vp]
    L.humid_gpu_debug_sort_pairs.argtypes = [vp, vp, vp, u64, C.c_int, C.c_int, vp, vp]
    L.humid_gpu_debug_partition.argtypes = [vp, C.c_int, vp, u64, u64, u32, vp, u32, vp, vp, vp]
    L.humid_gpu_partition_by_slot_noalloc.argtypes = [vp, vp, u64, u64, u32, vp, vp]
    _lib = L
    return L


NIBBLE_OTHER = 4  # HUMID_NIBBLE_OTHER
_NIBBLE_OF = np.full(256, NIBBLE_OTHER, dtype=np.uint8)
_NIBBLE_OF[[ord(c) for c in "ACGT"]] = [0, 1, 2, 3]


def pack_nibbles(chars: np.ndarray) -> np.ndarray:
    """The host side of the 4-bit wire format (include/humid_gpu.h: humid_gpu_add_batch_packed4): uint8[n, word_length]
    ASCII -> uint8[n, (word_length + 1) // 2], character q in the high nibble of byte q // 2 when q is even.  A table
    look-up per byte; no validation, no cut / pad rule — those stay on the device."""
    chars = np.ascontiguousarray(chars, dtype=np.uint8)
    n, wl = chars.shape
    code = _NIBBLE_OF[chars]
    if wl & 1:
        code = np.concatenate([code, np.zeros((n, 1), dtype=np.uint8)], axis=1)
    return np.ascontiguousarray((code[:, 0::2] << 4) | code[:, 1::2])


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Dedup:
    """One deduplication job on one GPU.

    seg_take: header-UMI size first (only if > 0), then nucleotides per input file — what
    the reference's preCompute (src/humid.cc:38-59) returns.
    """

    def __init__(self, word_length: int = 24, distance: int = 1, seg_take: Sequence[int] = (12, 12),
                 edit: bool = False, maximum: bool = False, device: int = -1, expected_tuples: int = 0,
                 defer_table: bool = False):
        self._L = load_library()
        self._h = C.c_void_p()
        cfg = Config()
        cfg.word_length, cfg.distance = int(word_length), int(distance)
        seg_take = [int(x) for x in seg_take]
        if len(seg_take) > MAX_SEGMENTS:
            raise HumidGpuError("too many segments")
        cfg.n_segments = len(seg_take)
        for i, t in enumerate(seg_take):
            cfg.seg_take[i] = t
        cfg.edit, cfg.maximum, cfg.device = int(bool(edit)), int(bool(maximum)), int(device)
        cfg.expected_tuples = int(expected_tuples)
        cfg.defer_table = 1 if defer_table else 0
        self.word_length, self.n_segments = cfg.word_length, cfg.n_segments
        self.distance = int(distance)
        rc = self._L.humid_gpu_create(C.byref(cfg), C.byref(self._h))
        if rc != 0:
            msg = self._L.humid_gpu_last_error(None).decode()
            self._h = C.c_void_p()
            raise HumidGpuError(f"humid_gpu_create failed ({rc}): {msg}")
        self.n_tuples = 0

    # -- plumbing ---------------------------------------------------------------------
    def _check(self, rc: int, what: str):
        if rc != 0:
            raise HumidGpuError(f"{what} failed ({rc}): {self._L.humid_gpu_last_error(self._h).decode()}")

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._L.humid_gpu_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    @property
    def stream(self) -> int:
        """Raw cudaStream_t of the handle (e.g. for torch.cuda.ExternalStream)."""
        return int(self._L.humid_gpu_stream(self._h) or 0)

    # -- test hooks for the device primitives (device pointers as ints) --------------------------
    def debug_sort_pairs(self, d_keys_in: int, d_vals_in: int, n: int, begin_bit: int, end_bit: int, d_keys_out: int,
                         d_vals_out: int):
        self._check(self._L.humid_gpu_debug_sort_pairs(self._h, C.c_void_p(d_keys_in), C.c_void_p(d_vals_in or None), n,
                                                       begin_bit, end_bit, C.c_void_p(d_keys_out), C.c_void_p(d_vals_out)),
                    "debug_sort_pairs")

    def debug_partition(self, kind: int, d_keys_in: int, n: int, arg: int, n_parts: int, bases, index_base: int,
                        d_keys_out: int, d_vals_out: int) -> np.ndarray:
        counts = np.zeros(16, dtype=np.uint64)
        b = None if bases is None else np.ascontiguousarray(bases, dtype=np.uint64)
        self._check(self._L.humid_gpu_debug_partition(self._h, kind, C.c_void_p(d_keys_in), n, arg, n_parts, _ptr(b),
                                                      index_base, C.c_void_p(d_keys_out), C.c_void_p(d_vals_out or None),
                                                      _ptr(counts)), "debug_partition")
        return counts[:n_parts]

    def partition_by_slot_noalloc(self, d_keys_in: int, n: int, slot_mask: int, index_base: int, d_keys_out: int,
                                  d_vals_out: int):
        """debug_partition(kind 0) without any allocation (static scratch, at most NOALLOC_MAX_ELEMENTS elements);
        enqueued on the handle's stream, NOT synchronised."""
        self._check(self._L.humid_gpu_partition_by_slot_noalloc(self._h, C.c_void_p(d_keys_in), n, slot_mask, index_base,
                                                                C.c_void_p(d_keys_out), C.c_void_p(d_vals_out)),
                    "partition_by_slot_noalloc")

    def reset(self):
        self._check(self._L.humid_gpu_reset(self._h), "reset")
        self.n_tuples = 0

    # -- pass 1: readData -----------------------------------------------------------
    def add_batch(self, chars: np.ndarray, seg_len: Optional[np.ndarray] = None):
        """chars: uint8[n, word_length] host array; seg_len: uint16[n, n_segments] or None."""
        chars = np.ascontiguousarray(chars, dtype=np.uint8)
        n = chars.shape[0]
        if chars.ndim != 2 or chars.shape[1] != self.word_length:
            raise HumidGpuError("chars must be [n_tuples, word_length]")
        if seg_len is not None:
            seg_len = np.ascontiguousarray(seg_len, dtype=np.uint16)
            if seg_len.shape != (n, self.n_segments):
                raise HumidGpuError("seg_len must be [n_tuples, n_segments]")
        self._check(self._L.humid_gpu_add_batch(self._h, _ptr(chars), _ptr(seg_len), self.n_tuples, n), "add_batch")
        self.n_tuples += n

    def add_batch_packed4(self, nibbles: np.ndarray, seg_len: Optional[np.ndarray] = None):
        """nibbles: uint8[n, (word_length + 1) // 2] from pack_nibbles(); half the bytes of add_batch on the wire."""
        nibbles = np.ascontiguousarray(nibbles, dtype=np.uint8)
        n = nibbles.shape[0]
        if nibbles.ndim != 2 or nibbles.shape[1] != (self.word_length + 1) // 2:
            raise HumidGpuError("nibbles must be [n_tuples, (word_length + 1) // 2]")
        if seg_len is not None:
            seg_len = np.ascontiguousarray(seg_len, dtype=np.uint16)
            if seg_len.shape != (n, self.n_segments):
                raise HumidGpuError("seg_len must be [n_tuples, n_segments]")
        self._check(self._L.humid_gpu_add_batch_packed4(self._h, _ptr(nibbles), _ptr(seg_len), self.n_tuples, n),
                    "add_batch_packed4")
        self.n_tuples += n

    def add_batch_device(self, d_chars: int, n: int, d_seg_len: int = 0):
        """d_chars / d_seg_len: raw device addresses (e.g. torch.Tensor.data_ptr())."""
        self._check(self._L.humid_gpu_add_batch_device(self._h, C.c_void_p(d_chars),
                                                       C.c_void_p(d_seg_len) if d_seg_len else None,
                                                       self.n_tuples, n), "add_batch_device")
        self.n_tuples += n

    # -- neighbours + clusters ---------------------------------------------------------
    def finalize(self):
        self._check(self._L.humid_gpu_finalize(self._h), "finalize")

    # -- results ----------------------------------------------------------------------
    def counts(self) -> Dict[str, int]:
        v = [C.c_uint64() for _ in range(4)]
        self._check(self._L.humid_gpu_counts(self._h, *[C.byref(x) for x in v]), "counts")
        return dict(zip(("total", "usable", "unique", "clusters"), (int(x.value) for x in v)))

    def histogram(self, which: str) -> Tuple[np.ndarray, np.ndarray]:
        n = C.c_uint64(0)
        self._check(self._L.humid_gpu_histogram(self._h, HIST[which], None, None, C.byref(n)), "histogram")
        k = np.zeros(max(n.value, 1), dtype=np.uint64)
        v = np.zeros(max(n.value, 1), dtype=np.uint64)
        self._check(self._L.humid_gpu_histogram(self._h, HIST[which], _ptr(k), _ptr(v), C.byref(n)), "histogram")
        return k[: n.value], v[: n.value]

    def keep_mask(self, first: int = 0, n: Optional[int] = None) -> np.ndarray:
        """bool[n]: tuple is written to the *_dedup files."""
        n = self.n_tuples - first if n is None else n
        bits = np.zeros((n + 63) // 64 + 1, dtype=np.uint64)
        self._check(self._L.humid_gpu_keep_mask(self._h, first, n, _ptr(bits)), "keep_mask")
        return np.unpackbits(bits.view(np.uint8), bitorder="little")[:n].astype(bool)

    def keep_bits(self, first: int = 0, n: Optional[int] = None) -> np.ndarray:
        n = self.n_tuples - first if n is None else n
        bits = np.zeros((n + 63) // 64, dtype=np.uint64)
        self._check(self._L.humid_gpu_keep_mask(self._h, first, n, _ptr(bits)), "keep_mask")
        return bits

    def cluster_ids(self, first: int = 0, n: Optional[int] = None) -> np.ndarray:
        n = self.n_tuples - first if n is None else n
        ids = np.zeros(max(n, 1), dtype=np.uint32)
        self._check(self._L.humid_gpu_cluster_ids(self._h, first, n, _ptr(ids)), "cluster_ids")
        return ids[:n]

    def stats_text(self) -> Dict[str, str]:
        """The four .dat files as writeStatistics (reference src/humid.cc:328-357) prints them."""
        out = {}
        for name, f in (("counts", "counts.dat"), ("neigh", "neigh.dat"), ("clusters", "clusters.dat")):
            k, v = self.histogram(name)
            out[f] = "".join(f"{int(a)} {int(b)}\n" for a, b in zip(k, v))
        c = self.counts()
        out["stats.dat"] = (f"total: {c['total']}\nusable: {c['usable']}\n"
                            f"unique: {c['unique']}\nclusters: {c['clusters']}\n")
        return out

    # -- introspection (parity tests) -------------------------------------------------
    def words(self):
        U = self.counts()["unique"]
        keys = np.zeros(max(U, 1), dtype=np.uint64)
        cnt = np.zeros(max(U, 1), dtype=np.uint32)
        first = np.zeros(max(U, 1), dtype=np.uint32)
        cid = np.zeros(max(U, 1), dtype=np.uint32)
        self._check(self._L.humid_gpu_debug_words(self._h, max(U, 1), _ptr(keys), _ptr(cnt), _ptr(first), _ptr(cid)),
                    "debug_words")
        return keys[:U], cnt[:U], first[:U], cid[:U]

    def words_hi(self) -> np.ndarray:
        """Bits 64.. of every word in rank order (all zero up to 31 letters; `words()` has the low 64)."""
        U = self.counts()["unique"]
        hi = np.zeros(max(U, 1), dtype=np.uint64)
        self._check(self._L.humid_gpu_debug_words_hi(self._h, max(U, 1), _ptr(hi)), "debug_words_hi")
        return hi[:U]

    def neighbours(self) -> Tuple[np.ndarray, np.ndarray]:
        U = self.counts()["unique"]
        n = C.c_uint64(0)
        self._check(self._L.humid_gpu_debug_neighbours(self._h, None, None, C.byref(n)), "debug_neighbours")
        row_ptr = np.zeros(U + 1, dtype=np.uint64)
        col = np.zeros(max(n.value, 1), dtype=np.uint32)
        self._check(self._L.humid_gpu_debug_neighbours(self._h, _ptr(row_ptr), _ptr(col), C.byref(n)),
                    "debug_neighbours")
        return row_ptr, col[: n.value]

    def n_entries(self) -> int:
        """Directed neighbour-list entries (sum of NLeaf::neighbours sizes)."""
        n = C.c_uint64(0)
        self._check(self._L.humid_gpu_debug_neighbours(self._h, None, None, C.byref(n)), "debug_neighbours")
        return int(n.value)

    def clusters(self) -> Tuple[np.ndarray, np.ndarray]:
        Cn = self.counts()["clusters"]
        size = np.zeros(max(Cn, 1), dtype=np.uint64)
        mx = np.zeros(max(Cn, 1), dtype=np.uint32)
        self._check(self._L.humid_gpu_debug_clusters(self._h, max(Cn, 1), _ptr(size), _ptr(mx)), "debug_clusters")
        return size[:Cn], mx[:Cn]

    def stage_stats(self) -> List[dict]:
        n = C.c_uint32(0)
        self._check(self._L.humid_gpu_stage_stats(self._h, None, C.byref(n)), "stage_stats")
        arr = (StageStat * max(n.value, 1))()
        self._check(self._L.humid_gpu_stage_stats(self._h, arr, C.byref(n)), "stage_stats")
        return [dict(name=a.name.decode(), ms=float(a.ms), launches=int(a.launches), units=int(a.units),
                     algo_bytes=int(a.algo_bytes)) for a in arr[: n.value]]
