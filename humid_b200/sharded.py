"""One-process-per-GPU deduplication: the phases of include/humid_gpu_shard.h with the
exchanges between them done by torch.distributed (NCCL over NVLink on the GPU box, gloo in the
CPU tests).  DESIGN.md §5 describes the partitioning.

The reference is a single process with one global trie (src/humid.cc:369-409); the bar for
this module is that every rank ends up with exactly the results the single-GPU path (and the
reference) gives for the concatenation of all ranks' tuples in rank order.

`ShardedDedup` is the orchestration (splitters, all-to-all-v, all-gather); the per-rank compute
is an *engine*: `CudaShardEngine` binds the C ABI.  The engine methods take and return torch
tensors that live on the engine's device, so the same orchestration runs over gloo with a
CPU engine in tests/ (a checker built on the oracle — not part of the product).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.distributed as dist

from . import api

MAX_PARTS = 16
SAMPLES_PER_RANK = 4096
NONE32 = 0xFFFFFFFF


def _bind(L):
    if getattr(L, "_shard_bound", False):
        return L
    vp, u64, u32 = C.c_void_p, C.c_uint64, C.c_uint32
    P = C.POINTER
    L.humid_gpu_shard_begin.argtypes = [vp, u64, P(u64)]
    L.humid_gpu_shard_sample_keys.argtypes = [vp, u32, vp, P(u32)]
    L.humid_gpu_shard_partition.argtypes = [vp, vp, u32, vp]
    L.humid_gpu_shard_export_tuples.argtypes = [vp, vp, vp, vp]
    L.humid_gpu_shard_import_tuples.argtypes = [vp, vp, vp, u64, P(u64)]
    L.humid_gpu_shard_set_layout.argtypes = [vp, u32, u32, vp, vp]
    L.humid_gpu_shard_join_local.argtypes = [vp, P(u64)]
    L.humid_gpu_shard_join_export.argtypes = [vp, u32, vp, vp, vp]
    L.humid_gpu_shard_join_match.argtypes = [vp, u32, vp, vp, u64, vp]
    L.humid_gpu_shard_join_entries.argtypes = [vp, vp]
    L.humid_gpu_shard_join_import.argtypes = [vp, vp, u64]
    L.humid_gpu_shard_build_rows.argtypes = [vp, P(u64)]
    L.humid_gpu_shard_cluster_local.argtypes = [vp, P(u64), P(u64)]
    L.humid_gpu_shard_export_open.argtypes = [vp, vp, vp, vp, vp, vp]
    L.humid_gpu_shard_import_open.argtypes = [vp, u64, u64, u64, vp, vp, vp, vp, vp, P(u64)]
    L.humid_gpu_shard_assign_ids.argtypes = [vp, u64, vp]
    L.humid_gpu_shard_finish.argtypes = [vp, vp, u64, u64, u64, u64]
    L.humid_gpu_shard_keep_bits.argtypes = [vp, vp, u64]
    L.humid_gpu_shard_tuple_ids.argtypes = [vp, vp]
    L._shard_bound = True
    return L


SHARD_EXPORTS = [
    "humid_gpu_shard_begin", "humid_gpu_shard_sample_keys", "humid_gpu_shard_partition",
    "humid_gpu_shard_export_tuples", "humid_gpu_shard_import_tuples", "humid_gpu_shard_set_layout",
    "humid_gpu_shard_join_local", "humid_gpu_shard_join_export", "humid_gpu_shard_join_match",
    "humid_gpu_shard_join_entries", "humid_gpu_shard_join_import", "humid_gpu_shard_build_rows",
    "humid_gpu_shard_cluster_local",
    "humid_gpu_shard_export_open", "humid_gpu_shard_import_open", "humid_gpu_shard_assign_ids",
    "humid_gpu_shard_finish", "humid_gpu_shard_keep_bits", "humid_gpu_shard_tuple_ids",
]


def _dp(t: Optional[torch.Tensor]):
    return None if t is None or t.numel() == 0 else C.c_void_p(t.data_ptr())


class CudaShardEngine(api.Dedup):
    """The C ABI (include/humid_gpu.h + include/humid_gpu_shard.h) on one GPU."""

    def __init__(self, device: int = 0, **kw):
        super().__init__(device=device, defer_table=True, **kw)  # words are counted by their owners
        _bind(self._L)
        self.device = torch.device("cuda", device)

    def _empty(self, n, dtype):
        return torch.empty(max(int(n), 0), dtype=dtype, device=self.device)

    def _sync(self):
        """The library works on its own stream: tensors produced by torch / NCCL on torch's
        current stream must be complete before a call that reads them."""
        torch.cuda.current_stream(self.device).synchronize()

    # torch has no unsigned 64/32-bit arithmetic: keys travel as int64, ranks/counts as int32
    def begin(self, tuple_offset: int) -> int:
        u = C.c_uint64(0)
        self._check(self._L.humid_gpu_shard_begin(self._h, tuple_offset, C.byref(u)), "shard_begin")
        return int(u.value)

    def sample_keys(self, n: int) -> np.ndarray:
        out = np.zeros(n, dtype=np.uint64)
        got = C.c_uint32(0)
        self._check(self._L.humid_gpu_shard_sample_keys(self._h, n, out.ctypes.data_as(C.c_void_p), C.byref(got)),
                    "shard_sample_keys")
        return out[: got.value]

    def partition(self, splitters: np.ndarray, n_parts: int) -> np.ndarray:
        sp = np.ascontiguousarray(splitters, dtype=np.uint64)
        out = np.zeros(n_parts, dtype=np.uint64)
        self._check(self._L.humid_gpu_shard_partition(self._h, sp.ctypes.data_as(C.c_void_p), n_parts,
                                                      out.ctypes.data_as(C.c_void_p)), "shard_partition")
        return out.astype(np.int64)

    def export_tuples(self, counts: np.ndarray):
        off = np.zeros(len(counts) + 1, dtype=np.uint64)
        off[1:] = np.cumsum(counts)
        n = int(off[-1])
        k, ix = self._empty(n, torch.int64), self._empty(n, torch.int32)
        self._check(self._L.humid_gpu_shard_export_tuples(self._h, off.ctypes.data_as(C.c_void_p), _dp(k), _dp(ix)),
                    "shard_export_tuples")
        return k, ix

    def import_tuples(self, keys, index) -> int:
        self._sync()
        u = C.c_uint64(0)
        self._check(self._L.humid_gpu_shard_import_tuples(self._h, _dp(keys), _dp(index), keys.numel(), C.byref(u)),
                    "shard_import_tuples")
        self.n_shard_words, self.n_shard_tuples = int(u.value), keys.numel()
        return int(u.value)

    def set_layout(self, part: int, n_parts: int, splitters: np.ndarray, bases: np.ndarray):
        sp = np.ascontiguousarray(splitters, dtype=np.uint64)
        b = np.ascontiguousarray(bases, dtype=np.uint64)
        self._check(self._L.humid_gpu_shard_set_layout(self._h, part, n_parts, sp.ctypes.data_as(C.c_void_p),
                                                       b.ctypes.data_as(C.c_void_p)), "shard_set_layout")

    def join_local(self) -> int:
        e = C.c_uint64(0)
        self._check(self._L.humid_gpu_shard_join_local(self._h, C.byref(e)), "shard_join_local")
        return int(e.value)

    def join_export(self, block: int, n_parts: int):
        counts = np.zeros(MAX_PARTS, dtype=np.uint64)
        k, r = self._empty(self.n_shard_words, torch.int64), self._empty(self.n_shard_words, torch.int32)
        self._check(self._L.humid_gpu_shard_join_export(self._h, block, counts.ctypes.data_as(C.c_void_p), _dp(k), _dp(r)),
                    "shard_join_export")
        return k, r, counts[:n_parts].astype(np.int64)

    def join_match(self, block: int, keys, ranks, n_parts: int):
        """-> (entries grouped by the part owning their row, entries per part)"""
        self._sync()
        counts = np.zeros(MAX_PARTS, dtype=np.uint64)
        self._check(self._L.humid_gpu_shard_join_match(self._h, block, _dp(keys), _dp(ranks), keys.numel(),
                                                       counts.ctypes.data_as(C.c_void_p)), "shard_join_match")
        counts = counts[:n_parts].astype(np.int64)
        out = self._empty(int(counts.sum()), torch.int64)
        self._check(self._L.humid_gpu_shard_join_entries(self._h, _dp(out)), "shard_join_entries")
        return out, counts

    def join_import(self, entries):
        self._sync()
        self._check(self._L.humid_gpu_shard_join_import(self._h, _dp(entries), entries.numel()), "shard_join_import")

    def build_rows(self) -> int:
        e = C.c_uint64(0)
        self._check(self._L.humid_gpu_shard_build_rows(self._h, C.byref(e)), "shard_build_rows")
        self.n_shard_entries = int(e.value)
        return int(e.value)

    def cluster_local(self) -> Tuple[int, int]:
        a, b = C.c_uint64(0), C.c_uint64(0)
        self._check(self._L.humid_gpu_shard_cluster_local(self._h, C.byref(a), C.byref(b)), "shard_cluster_local")
        self.n_open, self.n_open_entries = int(a.value), int(b.value)
        return self.n_open, self.n_open_entries

    def export_open(self):
        n, e = self.n_open, self.n_open_entries
        r, c, f, d = (self._empty(n, torch.int32) for _ in range(4))
        col = self._empty(e, torch.int32)
        self._check(self._L.humid_gpu_shard_export_open(self._h, _dp(r), _dp(c), _dp(f), _dp(d), _dp(col)),
                    "shard_export_open")
        return r, c, f, d, col

    def import_open(self, open_offset: int, rank, count, first, deg, col) -> int:
        self._sync()
        n = C.c_uint64(0)
        self._check(self._L.humid_gpu_shard_import_open(self._h, open_offset, rank.numel(), col.numel(), _dp(rank),
                                                        _dp(count), _dp(first), _dp(deg), _dp(col), C.byref(n)),
                    "shard_import_open")
        self.n_open_total = rank.numel()
        return int(n.value)

    def assign_ids(self, id_offset: int):
        ids = torch.zeros(max(self.n_open_total, 1), dtype=torch.int32, device=self.device)
        self._sync()
        self._check(self._L.humid_gpu_shard_assign_ids(self._h, id_offset, C.c_void_p(ids.data_ptr())), "shard_assign_ids")
        return ids

    def finish(self, open_ids, n_tuples: int, n_usable: int, n_words: int, n_clusters: int):
        self._sync()
        self._check(self._L.humid_gpu_shard_finish(self._h, C.c_void_p(open_ids.data_ptr()), n_tuples, n_usable, n_words,
                                                   n_clusters), "shard_finish")

    def keep_bits32(self, n_tuples: int):
        """int32 tensor of ceil(n_tuples/32) words: this shard's contribution to the keep mask."""
        n32 = (n_tuples + 31) // 32 + 1
        bits = self._empty(n32, torch.int32)
        self._check(self._L.humid_gpu_shard_keep_bits(self._h, C.c_void_p(bits.data_ptr()), n32), "shard_keep_bits")
        return bits

    # this shard's part of the results (numpy)
    def local_words(self):
        U = self.n_shard_words
        keys, cnt, first, cid = (np.zeros(max(U, 1), dtype=t) for t in (np.uint64, np.uint32, np.uint32, np.uint32))
        self._check(self._L.humid_gpu_debug_words(self._h, max(U, 1), api._ptr(keys), api._ptr(cnt), api._ptr(first),
                                                  api._ptr(cid)), "debug_words")
        return keys[:U], cnt[:U], first[:U], cid[:U]

    def local_rows(self):
        U = self.n_shard_words
        n = C.c_uint64(0)
        self._check(self._L.humid_gpu_debug_neighbours(self._h, None, None, C.byref(n)), "debug_neighbours")
        row_ptr = np.zeros(U + 1, dtype=np.uint64)
        col = np.zeros(max(n.value, 1), dtype=np.uint32)
        self._check(self._L.humid_gpu_debug_neighbours(self._h, api._ptr(row_ptr), api._ptr(col), C.byref(n)),
                    "debug_neighbours")
        return np.diff(row_ptr).astype(np.uint32), col[: n.value]

    def local_clusters(self, n: int):
        size = np.zeros(max(n, 1), dtype=np.uint64)
        mx = np.zeros(max(n, 1), dtype=np.uint32)
        self._check(self._L.humid_gpu_debug_clusters(self._h, max(n, 1), api._ptr(size), api._ptr(mx)), "debug_clusters")
        return size[:n], mx[:n]

    def tuple_ids(self):
        """int32 tensor: cluster id of every tuple handed to import_tuples, in that order."""
        ids = self._empty(self.n_shard_tuples, torch.int32)
        self._check(self._L.humid_gpu_shard_tuple_ids(self._h, _dp(ids)), "shard_tuple_ids")
        return ids


# ----------------------------------------------------------------------------------------------
class TorchComm:
    """The two collectives the path needs, over a torch.distributed process group."""

    def __init__(self, group=None):
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)

    def all_to_all_single(self, out, inp, out_splits=None, in_splits=None):
        dist.all_to_all_single(out, inp, output_split_sizes=out_splits, input_split_sizes=in_splits, group=self.group)

    def all_gather_into_tensor(self, out, inp):
        dist.all_gather_into_tensor(out, inp, group=self.group)

    def all_reduce_sum(self, t):
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)


class ThreadComm:
    """Loop-back collectives between `world` threads of ONE process (tests: several logical
    shards on one GPU run the same orchestration without a process group).  Every engine call
    synchronises its stream before returning, so no kernel ever waits on another rank."""

    class Shared:
        def __init__(self, world: int):
            import threading

            self.world = world
            self.barrier = threading.Barrier(world)
            self.slots = [None] * world

    def __init__(self, shared: "ThreadComm.Shared", rank: int):
        self.sh, self.rank, self.world = shared, rank, shared.world

    def _exchange(self, item):
        self.sh.slots[self.rank] = item
        self.sh.barrier.wait()
        items = list(self.sh.slots)
        self.sh.barrier.wait()
        return items

    def all_to_all_single(self, out, inp, out_splits=None, in_splits=None):
        W = self.world
        if in_splits is None:
            in_splits = [inp.numel() // W] * W
        offs = np.concatenate([[0], np.cumsum(in_splits)])
        items = self._exchange((inp, offs))
        parts = [t[int(o[self.rank]): int(o[self.rank + 1])] for (t, o) in items]
        got = torch.cat(parts) if parts else inp[:0]
        out.copy_(got.to(out.device))
        self.sh.barrier.wait()  # senders may reuse their buffers only after everyone copied

    def all_gather_into_tensor(self, out, inp):
        items = self._exchange(inp)
        out.copy_(torch.cat([t.to(out.device) for t in items]))
        self.sh.barrier.wait()

    def all_reduce_sum(self, t):
        items = self._exchange(t.clone())
        t.copy_(torch.stack(items).sum(dim=0).to(t.dtype))
        self.sh.barrier.wait()


def all_to_all_v(comm, t: torch.Tensor, send_counts: Sequence[int]) -> Tuple[torch.Tensor, List[int]]:
    """t is laid out destination-major; returns (received, recv_counts) source-major."""
    sc = torch.tensor([int(x) for x in send_counts], dtype=torch.int64, device=t.device)
    rc = torch.empty(comm.world, dtype=torch.int64, device=t.device)
    comm.all_to_all_single(rc, sc)
    recv_counts = [int(x) for x in rc.tolist()]
    return all_to_all_fixed(comm, t, send_counts, recv_counts), recv_counts


def all_to_all_fixed(comm, t: torch.Tensor, send_counts: Sequence[int], recv_counts: Sequence[int]) -> torch.Tensor:
    out = torch.empty(int(sum(recv_counts)), dtype=t.dtype, device=t.device)
    comm.all_to_all_single(out, t, [int(x) for x in recv_counts], [int(x) for x in send_counts])
    return out


def all_gather_v(comm, t: torch.Tensor, sizes: Sequence[int]) -> torch.Tensor:
    """Concatenation of every rank's `t` (sizes[r] elements) in rank order."""
    m = max(max(sizes), 1)
    pad = torch.zeros(m, dtype=t.dtype, device=t.device)
    pad[: t.numel()] = t
    buf = torch.empty(comm.world * m, dtype=t.dtype, device=t.device)
    comm.all_gather_into_tensor(buf, pad)
    return torch.cat([buf[r * m: r * m + sizes[r]] for r in range(comm.world)])


GROUP_EXPORTS = [
    "humid_gpu_group_create", "humid_gpu_group_destroy", "humid_gpu_group_join", "humid_gpu_group_finalize",
    "humid_gpu_group_collect_ids", "humid_gpu_group_phases",
]
GROUP_SLICES, GROUP_BATCHES = 0, 1


def _bind_group(L):
    if getattr(L, "_group_bound", False):
        return L
    vp, u32 = C.c_void_p, C.c_uint32
    L.humid_gpu_group_create.argtypes = [u32]
    L.humid_gpu_group_create.restype = vp
    L.humid_gpu_group_destroy.argtypes = [vp]
    L.humid_gpu_group_destroy.restype = None
    L.humid_gpu_group_join.argtypes = [vp, vp, C.c_char_p, u32, u32, C.c_int]
    L.humid_gpu_group_finalize.argtypes = [vp]
    L.humid_gpu_group_collect_ids.argtypes = [vp]
    L.humid_gpu_group_phases.argtypes = [vp, C.POINTER(C.c_char_p), C.POINTER(C.c_float), C.POINTER(u32)]
    L._group_bound = True
    return L


class GroupDedup(api.Dedup):
    """One rank of a multi-GPU job run by the library itself (include/humid_gpu_shard.h, "Groups"): the exchanges are
    stores / copies into the peers' buffers over NVLink, issued from C++ — no NCCL, no Python between the phases.
    Same call sequence as `api.Dedup`; rank r's tuples are the r-th contiguous slice of the job's input order and the
    per-tuple results (keep mask, cluster ids) are this rank's slice; counts and statistics are the whole job's.

    group: handle from `new_group(world)` when the ranks are threads of one process, else None and `shm_name` names
    the POSIX shared memory segment the ranks (processes) meet in."""

    transport = "peer memory (CUDA IPC mappings / peer access over NVLink), stores from the producing kernels; no NCCL"

    def __init__(self, rank: int, world: int, group=None, shm_name: Optional[str] = None, device: int = 0, **kw):
        kw.pop("defer_table", None)
        super().__init__(device=device, defer_table=True, **kw)  # words are counted by their owners
        _bind_group(self._L)
        self.rank, self.world = rank, world
        self.phases: List[Tuple[str, float]] = []
        rc = self._L.humid_gpu_group_join(self._h, group, (shm_name or "").encode() or None, rank, world, GROUP_SLICES)
        if rc != 0:
            msg = self._L.humid_gpu_last_error(self._h).decode()
            self.close()
            raise api.HumidGpuError(f"humid_gpu_group_join failed ({rc}): {msg}")

    def finalize(self):
        self._check(self._L.humid_gpu_group_finalize(self._h), "group_finalize")
        n = C.c_uint32(64)
        names, ms = (C.c_char_p * 64)(), (C.c_float * 64)()
        if self._L.humid_gpu_group_phases(self._h, names, ms, C.byref(n)) == 0:
            self.phases = [(names[i].decode(), float(ms[i])) for i in range(n.value)]

    def cluster_ids(self, first: int = 0, n: Optional[int] = None) -> np.ndarray:
        self._check(self._L.humid_gpu_group_collect_ids(self._h), "group_collect_ids")  # collective
        return super().cluster_ids(first, n)


def new_group(world: int):
    """In-process group object (ranks = threads of this process); free with free_group()."""
    L = _bind_group(api.load_library())
    g = L.humid_gpu_group_create(world)
    if not g:
        raise api.HumidGpuError("humid_gpu_group_create failed")
    return C.c_void_p(g)


def free_group(g):
    _bind_group(api.load_library()).humid_gpu_group_destroy(g)


_group_serial = [0]


def make_sharded(**kw):
    """The multi-GPU path bench.py and the tests use: one instance per rank (torch.distributed already initialised).
    On CUDA the ranks form a group of the library (peer-memory exchanges); HUMID_SHARD_TRANSPORT=nccl selects the
    torch.distributed / NCCL orchestration (`ShardedDedup`), which is also the fallback when the group cannot be
    formed (no CUDA IPC between the ranks)."""
    import os
    import sys

    world, rank = dist.get_world_size(), dist.get_rank()
    if os.environ.get("HUMID_SHARD_TRANSPORT", "peer") != "nccl" and torch.cuda.is_available() and world <= MAX_PARTS:
        name = [None]
        if rank == 0:
            _group_serial[0] += 1
            name[0] = "/humid_%d_%d_%s" % (os.getpid(), _group_serial[0], os.urandom(4).hex())
        dist.broadcast_object_list(name, src=0)
        ok, dd = 1, None
        try:
            dd = GroupDedup(rank, world, None, name[0], **kw)
        except api.HumidGpuError as e:
            ok = 0
            sys.stderr.write(f"rank {rank}: no peer-memory group ({e}); falling back to NCCL\n")
        oks = [None] * world
        dist.all_gather_object(oks, ok)
        if all(oks):
            return dd
        if dd is not None:
            dd.close()
    return ShardedDedup(**kw)


class ShardedDedup:
    transport = "NCCL all-to-all-v / all-gather (torch.distributed)"
    """Same call sequence as `api.Dedup` (add_batch* … finalize … results), one instance per
    rank; rank r's tuples are the r-th contiguous slice of the job's input order."""

    def __init__(self, engine=None, group=None, comm=None, **kw):
        self.comm = comm if comm is not None else TorchComm(group)
        self.rank, self.world = self.comm.rank, self.comm.world
        if self.world > MAX_PARTS:
            raise api.HumidGpuError(f"at most {MAX_PARTS} shards")
        self.eng = engine if engine is not None else CudaShardEngine(**kw)
        self.n_local = 0
        self.stage_ms: Dict[str, float] = {}

    # -- pass 1 -------------------------------------------------------------------------------
    @property
    def stream(self) -> int:
        return self.eng.stream

    def reset(self):
        self.eng.reset()
        self.n_local = 0

    def add_batch(self, chars, seg_len=None):
        self.eng.add_batch(chars, seg_len)
        self.n_local = self.eng.n_tuples

    def add_batch_device(self, d_chars: int, n: int, d_seg_len: int = 0):
        self.eng.add_batch_device(d_chars, n, d_seg_len)
        self.n_local = self.eng.n_tuples

    def add_batch_packed4(self, nibbles, seg_len=None):
        self.eng.add_batch_packed4(nibbles, seg_len)
        self.n_local = self.eng.n_tuples

    def _gather_ints(self, *vals: int) -> np.ndarray:
        dev = self.eng.device
        t = torch.tensor(list(vals), dtype=torch.int64, device=dev)
        out = torch.empty(self.world * len(vals), dtype=torch.int64, device=dev)
        self.comm.all_gather_into_tensor(out, t)
        return out.cpu().numpy().reshape(self.world, len(vals))

    # -- everything between readData and the outputs ------------------------------------------------
    def _mark(self, name):
        """HUMID_SHARD_PHASES=1: wall time per phase with a device synchronize at each mark (diagnostic)."""
        if not self._phases_on:
            return
        import time
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        now = time.perf_counter()
        self.phases.append((name, (now - self._t_mark) * 1e3))
        self._t_mark = time.perf_counter()

    def finalize(self):
        eng, W, me, g = self.eng, self.world, self.rank, self.comm
        dev = eng.device
        import os, time
        self._phases_on = os.environ.get("HUMID_SHARD_PHASES") == "1"
        self.phases = []
        if self._phases_on:
            torch.cuda.synchronize()
            self._t_mark = time.perf_counter()
        # tuple offsets: rank order == input order
        n_tuples = self._gather_ints(self.n_local)[:, 0]
        offs = np.concatenate([[0], np.cumsum(n_tuples)])
        self.tuple_offset, self.n_total = int(offs[me]), int(offs[-1])
        usable = eng.begin(self.tuple_offset)
        self._mark("offsets+begin")

        # splitters: an equal number of evenly spaced tuple keys from every rank
        samp = eng.sample_keys(SAMPLES_PER_RANK)
        st = torch.zeros(SAMPLES_PER_RANK + 1, dtype=torch.int64, device=dev)
        st[0] = len(samp)
        if len(samp):
            st[1: 1 + len(samp)] = torch.from_numpy(samp.view(np.int64).copy()).to(dev)
        allst = torch.empty(W * (SAMPLES_PER_RANK + 1), dtype=torch.int64, device=dev)
        g.all_gather_into_tensor(allst, st)
        allst = allst.cpu().numpy().reshape(W, SAMPLES_PER_RANK + 1)
        pool = np.sort(np.concatenate([allst[r, 1: 1 + allst[r, 0]].view(np.uint64) for r in range(W)]))
        if len(pool) and W > 1:
            splitters = pool[[(len(pool) * p) // W for p in range(1, W)]]
        else:
            splitters = np.zeros(max(W - 1, 0), dtype=np.uint64)
        # block 0 of the join (the first word_length // (distance + 1) letters) must not straddle two
        # shards: clear every letter after it (include/humid_gpu_shard.h: join_local)
        n, d = int(eng.word_length), int(eng.distance)
        self.n_blocks = d + 1 if d >= 1 else 1
        tail_bits = 2 * (n - n // (d + 1))
        splitters = (splitters >> np.uint64(tail_bits)) << np.uint64(tail_bits) if tail_bits < 64 else splitters * np.uint64(0)
        self.splitters = splitters
        self._mark("splitters")

        # tuples to the owners of their key range; each owner then runs the single-GPU word stages
        send = eng.partition(splitters, W)
        k, ix = eng.export_tuples(send)
        self._mark("partition+export")
        rk, recv = all_to_all_v(g, k, send)
        rix = all_to_all_fixed(g, ix, send, recv)
        self._sent_index, self._send_counts, self._recv_counts = ix, send, recv
        self._mark("tuples all-to-all")
        n_shard = eng.import_tuples(rk, rix)
        self._mark("import_tuples")
        del k, rk, rix
        info = self._gather_ints(n_shard, usable)
        bases = np.concatenate([[0], np.cumsum(info[:, 0])]).astype(np.uint64)
        self.n_words_total, self.usable_total = int(bases[-1]), int(info[:, 1].sum())
        eng.set_layout(me, W, splitters, bases)
        self._mark("layout")

        # neighbour search: Hamming block join.  Block 0 groups lie inside a shard (aligned splitters);
        # for every other block the words meet at the part that owns their block value, and the
        # entries found there travel back to the owners of their rows.
        eng.join_local()
        self._mark("join_local")
        self.words_exchanged = 0
        for b in range(1, self.n_blocks):
            bk, br, bcounts = eng.join_export(b, W)
            rk_, brecv = all_to_all_v(g, bk, bcounts)
            rr_ = all_to_all_fixed(g, br, bcounts, brecv)
            self._mark("block words all-to-all")
            ent, ecounts = eng.join_match(b, rk_, rr_, W)
            self._mark("join_match")
            rent, _ = all_to_all_v(g, ent, ecounts)
            eng.join_import(rent)
            self._mark("entries all-to-all + import")
            self.words_exchanged += int(np.sum(bcounts))
            del bk, br, rk_, rr_, ent, rent
        n_entries = eng.build_rows()
        self._mark("apply+rows")

        # clusters: components inside a shard are finished there; the ones that cross shards
        # ("open") are gathered whole and replayed identically on every shard
        n_open, n_open_e = eng.cluster_local()
        self._mark("cluster_local")
        orank, ocount, ofirst, odeg, ocol = eng.export_open()
        sizes = self._gather_ints(n_open, n_open_e, n_shard)
        osz, esz = [int(x) for x in sizes[:, 0]], [int(x) for x in sizes[:, 1]]
        open_off = int(np.sum(sizes[:me, 0]))
        self._mark("export_open+sizes")
        n_cl = eng.import_open(open_off, all_gather_v(g, orank, osz), all_gather_v(g, ocount, osz),
                               all_gather_v(g, ofirst, osz), all_gather_v(g, odeg, osz), all_gather_v(g, ocol, esz))
        ncl = self._gather_ints(n_cl)[:, 0]
        self.cluster_offsets = np.concatenate([[0], np.cumsum(ncl)])
        self.n_clusters_local = int(n_cl)
        self._mark("open gather+import")
        ids = eng.assign_ids(int(self.cluster_offsets[me]))
        if hasattr(eng, "_sync"):
            eng._sync()
        g.all_reduce_sum(ids)
        eng.finish(ids, self.n_total, self.usable_total, self.n_words_total, int(self.cluster_offsets[-1]))
        self._mark("ids+finish")
        self.word_sizes = [int(x) for x in sizes[:, 2]]
        self.open_words_total = int(np.sum(sizes[:, 0]))

    # -- results: whole job, identical on every rank, unless stated; all of them are collective -----
    def counts(self):
        return self.eng.counts()

    def _gather_np(self, a: np.ndarray) -> np.ndarray:
        """Concatenation over the ranks (in rank order) of a 1-d numpy array."""
        dev = self.eng.device
        n = self._gather_ints(a.size)[:, 0]
        if a.dtype.itemsize == 8:
            t = torch.from_numpy(np.ascontiguousarray(a).view(np.int64).copy()).to(dev)
        else:
            t = torch.from_numpy(np.ascontiguousarray(a).view(np.int32).copy()).to(dev)
        if hasattr(self.eng, "_sync"):
            self.eng._sync()
        out = all_gather_v(self.comm, t, [int(x) for x in n]).cpu().numpy()
        return out.view(a.dtype)

    def histogram(self, which: str):
        """counts.dat / neigh.dat / clusters.dat: every shard histograms its own words / lists /
        clusters on its GPU, the (key, value) lists are merged here."""
        k, v = self.eng.histogram(which)
        ks, vs = self._gather_np(k.astype(np.uint64)), self._gather_np(v.astype(np.uint64))
        uk, inv = np.unique(ks, return_inverse=True)
        uv = np.zeros(uk.size, dtype=np.uint64)
        np.add.at(uv, inv, vs)
        return uk, uv

    def stats_text(self):
        out = {}
        for name, f in (("counts", "counts.dat"), ("neigh", "neigh.dat"), ("clusters", "clusters.dat")):
            k, v = self.histogram(name)
            out[f] = "".join(f"{int(a)} {int(b)}\n" for a, b in zip(k, v))
        c = self.counts()
        out["stats.dat"] = (f"total: {c['total']}\nusable: {c['usable']}\nunique: {c['unique']}\n"
                            f"clusters: {c['clusters']}\n")
        return out

    def keep_words32(self):
        """int32 tensor (device): the job's keep mask, bit t of word t//32 = tuple t is written."""
        bits = self.eng.keep_bits32(self.n_total)
        if hasattr(self.eng, "_sync"):
            self.eng._sync()
        self.comm.all_reduce_sum(bits)  # every tuple is kept by at most one cluster: sum == or
        return bits

    def keep_mask(self) -> np.ndarray:
        """bool[n_local]: this rank's tuples that are written to the *_dedup files."""
        w = self.keep_words32().cpu().numpy().view(np.uint32)
        lo, hi = self.tuple_offset, self.tuple_offset + self.n_local
        return np.unpackbits(w[lo // 32: (hi + 31) // 32 + 1].view(np.uint8), bitorder="little")[lo % 32: lo % 32 + self.n_local].astype(bool)

    def keep_bits(self) -> np.ndarray:
        return np.packbits(self.keep_mask(), bitorder="little")

    def cluster_ids(self) -> np.ndarray:
        """uint32[n_local]: cluster ids of this rank's tuples (the owners of the key ranges look
        the ids up and send them back along the route the tuples took)."""
        ids = self.eng.tuple_ids()
        if hasattr(self.eng, "_sync"):
            self.eng._sync()
        back = all_to_all_fixed(self.comm, ids, self._recv_counts, self._send_counts)
        out = torch.zeros(self.n_local, dtype=torch.int32, device=back.device)
        out[(self._sent_index.long() & 0xFFFFFFFF) - self.tuple_offset] = back
        return out.cpu().numpy().view(np.uint32)

    # introspection for the parity tests: the shards' pieces put back together
    def words(self):
        return tuple(self._gather_np(a) for a in self.eng.local_words())

    def neighbours(self):
        deg, col = self.eng.local_rows()
        deg, col = self._gather_np(deg), self._gather_np(col)
        return np.concatenate([[0], np.cumsum(deg.astype(np.uint64))]).astype(np.uint64), col

    def clusters(self):
        size, mx = self.eng.local_clusters(self.n_clusters_local)
        return self._gather_np(size), self._gather_np(mx)

    def stage_stats(self):
        return self.eng.stage_stats()

    def close(self):
        self.eng.close()
