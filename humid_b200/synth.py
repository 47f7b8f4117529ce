"""Seeded synthetic paired-end reads with UMIs, PCR duplicates and sequencing errors.

This is the workload generator of SURVEY.md §8(d) / BASELINE.json `configs`: it produces
either FastQ files (for the `humid` CLI and for the reference binary) or, for the large
configurations, the word-character batches the C ABI consumes directly
(`humid_gpu_add_batch`: `word_length` ASCII bytes per read tuple), so that hot-path timing
does not include disk I/O or FastQ parsing.

Model (all draws from one `numpy.random.Generator(PCG64(seed))`):
  * a fixed random "genome"; molecules M = R/3, each = (UMI, fragment start, insert size);
    mate 1 = genome[start:start+L], mate 2 = reverse complement of the fragment's far end;
  * copies per molecule 1 + Geometric(0.5), renormalised to R reads in total, shuffled;
  * substitution errors: 0.5 %/base on mates, 1 %/base on the UMI;
  * optional indels (config 4): at most one insertion or deletion per mate prefix,
    0.1 %/base each inside the first 30 nt, the mate re-extended from the genome so that
    sequence and quality lengths always match (the reference crashes otherwise);
  * 0.1 % of the tuples get an 'N' inside the word region (=> filtered, cluster id 0).
"""
from __future__ import annotations

import dataclasses
import os
from typing import List, Optional, Tuple

import numpy as np

_ASCII = np.frombuffer(b"ACGT", dtype=np.uint8)
READ_LEN = 150


@dataclasses.dataclass
class SynthConfig:
    """One of BASELINE.json's configurations (names follow `configs[i]`)."""

    name: str
    n_tuples: int
    umi_len: int = 8
    umi_where: str = "header"  # "header" | "file" (UMI as an extra FastQ) | "none"
    header_style: str = "bcl"  # "bcl" (…:UMI) | "underscore" (…_UMI)
    word_length: int = 24
    distance: int = 1
    edit: bool = False
    maximum: bool = False
    indels: bool = False
    seed: int = 1001
    genome_len: int = 50_000_000
    n_frac: float = 0.001

    @property
    def n_files(self) -> int:
        return 3 if self.umi_where == "file" else 2

    def cli_flags(self) -> List[str]:
        f = ["-n", str(self.word_length), "-m", str(self.distance)]
        if self.edit:
            f.append("-e")
        if self.maximum:
            f.append("-x")
        return f


def baseline_config(i: int, n_tuples: Optional[int] = None, **kw) -> SynthConfig:
    """BASELINE.json `configs[i]` (0-based), optionally at a reduced number of tuples."""
    table = [
        dict(name="c1_1M_umi8hdr_ham1_max", n_tuples=1_000_000, umi_len=8, umi_where="header",
             header_style="bcl", distance=1, maximum=True, seed=1001),
        dict(name="c2_50M_umi12file_ham1_dir", n_tuples=50_000_000, umi_len=12, umi_where="file",
             distance=1, maximum=False, seed=1002),
        dict(name="c3_50M_noumi_ham2", n_tuples=50_000_000, umi_len=0, umi_where="none",
             distance=2, maximum=False, seed=1003),
        dict(name="c4_20M_indel_edit", n_tuples=20_000_000, umi_len=0, umi_where="none",
             distance=2, edit=True, indels=True, seed=1004),
        dict(name="c5_500M_umi10hdr_ham1_dir", n_tuples=500_000_000, umi_len=10,
             umi_where="header", header_style="underscore", distance=1, maximum=False, seed=1005),
    ]
    d = dict(table[i])
    if n_tuples is not None:
        d["n_tuples"] = int(n_tuples)
    d.update(kw)
    return SynthConfig(**d)


def nt_from_file(files: int, length: int) -> List[int]:
    """`ntFromFile` (reference src/fastq.cc:220-230): remainder goes to the LAST file."""
    div = length // files
    return [div] * (files - 1) + [div + length % files]


def pre_compute(umi_size_first: int, word_length: int, files: int) -> Tuple[int, List[int]]:
    """`preCompute` (reference src/humid.cc:38-59)."""
    from_file = max(0, word_length - umi_size_first)
    return min(umi_size_first, word_length), nt_from_file(files, from_file)


class Synth:
    """Lazily generated read tuples; chunked so 500 M tuples never sit in host RAM at once."""

    def __init__(self, cfg: SynthConfig):
        self.cfg = cfg
        rng = np.random.Generator(np.random.PCG64(cfg.seed))
        R = cfg.n_tuples
        self.M = M = max(1, R // 3)
        # small explicit genomes (tests) are honoured; the default is capped by the molecule count
        glen = int(cfg.genome_len if cfg.genome_len < 100_000 else min(cfg.genome_len, max(100_000, 40 * M)))
        self.genome = rng.integers(0, 4, size=glen + 1024, dtype=np.uint8)
        self.glen = glen
        self.mol_umi = rng.integers(0, 4, size=(M, max(cfg.umi_len, 1)), dtype=np.uint8)
        self.mol_start = rng.integers(0, glen - 600, size=M, dtype=np.int64)
        ins = np.rint(rng.normal(300.0, 30.0, size=M)).astype(np.int64)
        self.mol_ins = np.clip(ins, READ_LEN + 10, 500)
        w = rng.geometric(0.5, size=M).astype(np.float64)
        counts = rng.multinomial(R, w / w.sum())
        mol = np.repeat(np.arange(M, dtype=np.int64), counts)
        rng.shuffle(mol)
        self.mol_of_tuple = mol
        self._chunk_seed = int(rng.integers(0, 2**62))

    # -- per-chunk generation ---------------------------------------------------------
    BLOCK = 1 << 18  # tuples per independently seeded block => any [lo,hi) gives the same reads

    def _mates(self, lo: int, hi: int, length: int):
        """ASCII arrays (umi[R,u], m1[R,length], m2[R,length]) for tuples lo..hi."""
        out = []
        for b in range(lo // self.BLOCK, (max(hi, lo + 1) - 1) // self.BLOCK + 1):
            blo, bhi = b * self.BLOCK, min(self.cfg.n_tuples, (b + 1) * self.BLOCK)
            au, a1, a2 = self._block(blo, bhi, length)
            s = slice(max(lo, blo) - blo, min(hi, bhi) - blo)
            out.append((au[s], a1[s], a2[s]))
        if len(out) == 1:
            return out[0]
        return tuple(np.concatenate([o[k] for o in out], axis=0) for k in range(3))

    def _block(self, lo: int, hi: int, length: int):
        cfg = self.cfg
        rng = np.random.Generator(np.random.PCG64([self._chunk_seed, lo]))
        want, length = length, max(length, 32)  # draws for the first 32 nt never depend on `length`
        mol = self.mol_of_tuple[lo:hi]
        r = hi - lo
        start = self.mol_start[mol]
        end = start + self.mol_ins[mol]  # one past the fragment
        j = np.arange(length, dtype=np.int64)[None, :]
        idx1 = start[:, None] + j
        idx2 = end[:, None] - 1 - j
        ins_base1 = ins_base2 = None
        if cfg.indels:
            idx1, ins_base1 = self._indel(rng, idx1, +1)
            idx2, ins_base2 = self._indel(rng, idx2, -1)
        m1 = self.genome[np.clip(idx1, 0, self.glen + 1000)]
        m2 = 3 - self.genome[np.clip(idx2, 0, self.glen + 1000)]
        if ins_base1 is not None:
            rows, cols, vals = ins_base1
            m1[rows, cols] = vals
            rows, cols, vals = ins_base2
            m2[rows, cols] = vals
        def subst(view, p):
            err = rng.random(view.shape) < p
            view[err] = (view[err] + rng.integers(1, 4, size=int(err.sum()), dtype=np.uint8)) & 3

        subst(m1[:, :32], 0.005)
        subst(m2[:, :32], 0.005)
        umi = self.mol_umi[mol].copy()
        if cfg.umi_len:
            subst(umi, 0.01)
        bad = np.nonzero(rng.random(r) < cfg.n_frac)[0]
        bad_pos = rng.integers(0, 6, size=bad.size)
        if length > 32:  # tails last, so the word region is identical for every `length`
            subst(m1[:, 32:], 0.005)
            subst(m2[:, 32:], 0.005)
        a1, a2, au = _ASCII[m1], _ASCII[m2], _ASCII[umi]
        # 'N' inside the word region of mate 1 for ~0.1 % of the tuples
        if bad.size:
            a1[bad, bad_pos] = ord("N")
        return au, a1[:, :want], a2[:, :want]

    @staticmethod
    def _indel(rng, idx, step):
        """At most one indel inside the first 30 nt of each row of genome indices."""
        r, length = idx.shape
        span = min(30, length)
        p_any = 1.0 - (1.0 - 0.002) ** span
        hit = np.nonzero(rng.random(r) < p_any)[0]
        pos = rng.integers(0, span, size=hit.size)
        is_ins = rng.random(hit.size) < 0.5
        j = np.arange(length, dtype=np.int64)[None, :]
        # deletion at pos: positions >= pos read one base further along the fragment
        d_rows = hit[~is_ins]
        idx[d_rows] += step * (j >= pos[~is_ins][:, None])
        # insertion at pos: positions > pos read one base earlier; pos itself is random
        i_rows = hit[is_ins]
        idx[i_rows] -= step * (j > pos[is_ins][:, None])
        vals = rng.integers(0, 4, size=i_rows.size, dtype=np.uint8)
        return idx, (i_rows, pos[is_ins], vals)

    # -- word batches (the C-ABI input) -----------------------------------------------
    def layout(self) -> Tuple[int, List[int]]:
        """(header_umi_size, nt_to_take per file) exactly as the reference's preCompute."""
        cfg = self.cfg
        hdr = cfg.umi_len if cfg.umi_where == "header" else 0
        return pre_compute(hdr, cfg.word_length, cfg.n_files)

    def word_chars(self, lo: int = 0, hi: Optional[int] = None) -> np.ndarray:
        """uint8[hi-lo, word_length]: UMI-from-header first, then each file's prefix."""
        cfg = self.cfg
        hi = cfg.n_tuples if hi is None else hi
        hdr, take = self.layout()
        need = max(take[0], take[1], 8)
        au, a1, a2 = self._mates(lo, hi, need)
        parts = []
        if hdr:
            parts.append(au[:, :hdr])
        parts.append(a1[:, : take[0]])
        parts.append(a2[:, : take[1]])
        if cfg.umi_where == "file":
            u = au[:, : take[2]]
            if u.shape[1] < take[2]:  # UMI read shorter than requested => 'N' padding
                pad = np.full((u.shape[0], take[2] - u.shape[1]), ord("N"), dtype=np.uint8)
                u = np.concatenate([u, pad], axis=1)
            parts.append(u)
        out = np.ascontiguousarray(np.concatenate(parts, axis=1))
        assert out.shape[1] == cfg.word_length, (out.shape, cfg.word_length)
        return out

    def word_chars_chunks(self, chunk: int = 4_000_000):
        for lo in range(0, self.cfg.n_tuples, chunk):
            hi = min(self.cfg.n_tuples, lo + chunk)
            yield lo, self.word_chars(lo, hi)

    # -- FastQ files ------------------------------------------------------------------
    def write_fastq(self, directory: str, prefix: Optional[str] = None, chunk: int = 500_000,
                    gzip_output: bool = False) -> List[str]:
        """Write R1/R2 (and the UMI file for umi_where == 'file'); returns the paths in the
        order they are passed to `humid`."""
        import gzip

        cfg = self.cfg
        prefix = prefix or cfg.name
        os.makedirs(directory, exist_ok=True)
        ext = ".fastq.gz" if gzip_output else ".fastq"
        names = [os.path.join(directory, f"{prefix}_R{i + 1}{ext}") for i in range(2)]
        if cfg.umi_where == "file":
            names.append(os.path.join(directory, f"{prefix}_UMI{ext}"))
        opener = (lambda p: gzip.open(p, "wb", compresslevel=1)) if gzip_output else (lambda p: open(p, "wb"))
        files = [opener(p) for p in names]
        try:
            for lo in range(0, cfg.n_tuples, chunk):
                hi = min(cfg.n_tuples, lo + chunk)
                au, a1, a2 = self._mates(lo, hi, READ_LEN)
                ids = np.arange(lo, hi, dtype=np.int64)
                for fi, seq in enumerate((a1, a2)):
                    files[fi].write(_records(_headers(cfg, ids, au, fi + 1), seq).tobytes())
                if cfg.umi_where == "file":
                    files[2].write(_records(_headers(cfg, ids, au, 3), au[:, : cfg.umi_len]).tobytes())
        finally:
            for f in files:
                f.close()
        return names


def _digits(x: np.ndarray, width: int) -> np.ndarray:
    out = np.empty((x.size, width), dtype=np.uint8)
    for k in range(width):
        out[:, width - 1 - k] = ord("0") + (x // 10**k) % 10
    return out


def _const(s: bytes, r: int) -> np.ndarray:
    return np.broadcast_to(np.frombuffer(s, dtype=np.uint8), (r, len(s)))


def _headers(cfg: SynthConfig, ids: np.ndarray, au: np.ndarray, mate: int) -> np.ndarray:
    r = ids.size
    in_header = cfg.umi_where == "header" and cfg.umi_len > 0
    if cfg.header_style == "underscore":
        parts = [_const(b"@SIM.", r), _digits(ids, 10)]
        if in_header:
            parts += [_const(b"_", r), au[:, : cfg.umi_len]]
        parts.append(_const(b" %d" % mate, r))
    else:
        parts = [_const(b"@SIM:1:FC:1:", r), _digits(ids // 10_000_000 + 1101, 4), _const(b":", r),
                 _digits((ids // 3163) % 100_000, 5), _const(b":", r), _digits(ids % 100_000, 5)]
        if in_header:
            parts += [_const(b":", r), au[:, : cfg.umi_len]]
        parts.append(_const(b" %d:N:0:1" % mate, r))
    return np.concatenate(parts, axis=1)


def _records(hdr: np.ndarray, seq: np.ndarray) -> np.ndarray:
    r, L = seq.shape
    qual = np.full((r, L), ord("I"), dtype=np.uint8)
    return np.ascontiguousarray(np.concatenate(
        [hdr, _const(b"\n", r), seq, _const(b"\n+\n", r), qual, _const(b"\n", r)], axis=1))


# ======================================================================================
# Stateless generator for the large configurations (50 M – 500 M tuples)
# ======================================================================================
def _i64(x: int) -> int:
    """Wrap a Python int into the int64 range (two's complement)."""
    return ((x + (1 << 63)) % (1 << 64)) - (1 << 63)


class HashSynth:
    """Same read model as `Synth`, but every quantity is a pure function of
    (seed, tuple index) built from 64-bit integer hashing, evaluated with torch integer ops.

    Consequences: any tuple range can be produced on any device (each rank of a multi-GPU
    run generates just its own slice, directly in HBM), CPU and CUDA give identical bytes,
    and no per-tuple state (the 8 B/tuple molecule table of `Synth`) is ever materialised.
    torch is used here as plumbing for the *workload generator* only; nothing on the
    deduplication path depends on it.

      molecule of tuple t  : weighted pick; weights 1+Geometric(0.5)… precisely
                             Geometric(0.5) on {1,2,…} as in `Synth`, arranged in a repeating
                             pattern of PATTERN molecules so the inverse CDF is a 4096-entry
                             search; M = R/3 molecules (rounded up to whole patterns)
      molecule m           : UMI, fragment start, insert size ~ N(300, 30) (Irwin–Hall of 4 bytes)
      genome               : base(p) = hash(p) & 3 over `glen` positions (never stored)
      errors               : 0.5 %/base on the first 32 nt of each mate, 1 %/base on the UMI,
                             at most one indel per mate prefix when cfg.indels, 'N' in 0.1 %
    """

    PATTERN = 4096

    def __init__(self, cfg: SynthConfig):
        import torch  # local: the numpy generator above must stay importable without torch

        self.torch = torch
        self.cfg = cfg
        R = cfg.n_tuples
        M = max(1, R // 3)
        self.n_patterns = (M + self.PATTERN - 1) // self.PATTERN
        self.M = self.n_patterns * self.PATTERN
        rng = np.random.Generator(np.random.PCG64(cfg.seed))
        w = rng.geometric(0.5, size=self.PATTERN).astype(np.int64)
        self._cdf = np.cumsum(w)  # inclusive
        self.glen = int(cfg.genome_len if cfg.genome_len < 100_000 else min(cfg.genome_len, max(100_000, 40 * self.M)))
        s = [int(x) for x in rng.integers(1, 2**62, size=8)]
        self.k_pick, self.k_umi, self.k_start, self.k_ins, self.k_gen, self.k_err, self.k_n, self.k_indel = s
        self._cdf_dev = {}

    # splitmix64 finaliser on int64 tensors (wrapping multiply, logical shifts emulated)
    def _mix(self, x):
        x = x ^ ((x >> 30) & 0x3FFFFFFFF)
        x = x * -4658895280553007687  # 0xbf58476d1ce4e5b9
        x = x ^ ((x >> 27) & 0x1FFFFFFFFF)
        x = x * -7723592293110705685  # 0x94d049bb133111eb
        x = x ^ ((x >> 31) & 0x1FFFFFFFF)
        return x

    def layout(self) -> Tuple[int, List[int]]:
        cfg = self.cfg
        hdr = cfg.umi_len if cfg.umi_where == "header" else 0
        return pre_compute(hdr, cfg.word_length, cfg.n_files)

    def _molecule(self, t):
        torch = self.torch
        dev = t.device
        if dev not in self._cdf_dev:
            self._cdf_dev[dev] = torch.from_numpy(self._cdf).to(dev)
        cdf = self._cdf_dev[dev]
        h = self._mix(t ^ self.k_pick)
        block = ((h >> 32) & 0xFFFFFFFF) % self.n_patterns
        slot = (h & 0xFFFFFFFF) % int(self._cdf[-1])
        idx = torch.searchsorted(cdf, slot, right=True)
        return block * self.PATTERN + idx

    def _genome(self, p):
        return self._mix(p ^ self.k_gen) & 3

    def _mates(self, lo: int, hi: int, length: int, device):
        """Letter codes (0..3, 4 = 'N') umi[r,u], m1[r,length], m2[r,length] as uint8 tensors."""
        torch = self.torch
        cfg = self.cfg
        t = torch.arange(lo, hi, dtype=torch.int64, device=device)
        m = self._molecule(t)
        start = (self._mix(m ^ self.k_start) & 0x7FFFFFFFFFFFFFFF) % (self.glen - 600)
        hb = self._mix(m ^ self.k_ins)
        ih = (hb & 255) + ((hb >> 8) & 255) + ((hb >> 16) & 255) + ((hb >> 24) & 255)  # mean 510, sd 147.8
        ins = torch.clamp(300 + ((ih - 510) * 13) // 64, READ_LEN + 10, 500)
        end = start + ins
        j = torch.arange(length, dtype=torch.int64, device=device)[None, :]
        idx1 = start[:, None] + j
        idx2 = end[:, None] - 1 - j
        ins1 = ins2 = None
        if cfg.indels:
            idx1, ins1 = self._indel(t, idx1, +1, 0)
            idx2, ins2 = self._indel(t, idx2, -1, 1)
        m1 = self._genome(idx1)
        m2 = 3 - self._genome(idx2)
        if ins1 is not None:
            for mm, (rows, cols, vals) in ((m1, ins1), (m2, ins2)):
                ok = cols < length
                mm[rows[ok], cols[ok]] = vals[ok]
        span = min(32, length)
        for which, mm in ((0, m1), (1, m2)):
            self._subst(t, mm, span, 328, which)  # 328/65536 = 0.5 %
        u = max(cfg.umi_len, 1)
        hu = self._mix(m ^ self.k_umi)
        ju = torch.arange(u, dtype=torch.int64, device=device)[None, :]
        umi = (hu[:, None] >> (2 * ju)) & 3
        if cfg.umi_len:
            self._subst(t, umi, u, 655, 2)  # 655/65536 = 1 %
        hn = self._mix(t ^ self.k_n)
        bad = (hn & 0xFFFFF) < int(cfg.n_frac * (1 << 20))
        if length >= 6:
            pos = ((hn >> 20) & 0xFFFF) % 6
            rows = torch.nonzero(bad)[:, 0]
            m1[rows, pos[rows]] = 4
        return umi.to(torch.uint8), m1.to(torch.uint8), m2.to(torch.uint8)

    def _subst(self, t, mat, span, thresh, which):
        """Substitution errors on mat[:, :span]: one 16-bit draw per base, 4 draws per hash."""
        torch = self.torch
        for q in range((span + 3) // 4):
            h = self._mix(t ^ _i64(self.k_err + 0x9E3779B97F4A7C15 * (which * 16 + q + 1)))
            for k in range(4):
                col = 4 * q + k
                if col >= span:
                    break
                draw = (h >> (16 * k)) & 0xFFFF
                err = draw < thresh
                # a different letter: + (1 + draw % 3)
                mat[:, col] = torch.where(err, (mat[:, col] + 1 + draw % 3) & 3, mat[:, col])

    def _indel(self, t, idx, step, which):
        torch = self.torch
        length = idx.shape[1]
        span = min(30, length)
        h = self._mix(t ^ _i64(self.k_indel + which))
        p_any = 1.0 - (1.0 - 0.002) ** span
        hit = (h & 0xFFFFF) < int(p_any * (1 << 20))
        pos = ((h >> 20) & 0xFFFF) % span
        is_ins = ((h >> 36) & 1) == 1
        val = (h >> 37) & 3
        j = torch.arange(length, dtype=torch.int64, device=idx.device)[None, :]
        dele = (hit & ~is_ins)[:, None] & (j >= pos[:, None])
        inse = (hit & is_ins)[:, None] & (j > pos[:, None])
        idx = idx + step * dele.to(torch.int64) - step * inse.to(torch.int64)
        rows = torch.nonzero(hit & is_ins)[:, 0]
        return idx, (rows, pos[rows], val[rows])

    _LETTERS = b"ACGTN"

    def word_chars(self, lo: int, hi: int, device="cpu", chunk: int = 2_000_000):
        """uint8 tensor [hi-lo, word_length] of ASCII word characters on `device`."""
        torch = self.torch
        cfg = self.cfg
        hdr, take = self.layout()
        need = max(take[0], take[1], 8)
        table = torch.tensor(list(self._LETTERS), dtype=torch.uint8, device=device)
        out = torch.empty((hi - lo, cfg.word_length), dtype=torch.uint8, device=device)
        for a in range(lo, hi, chunk):
            b = min(hi, a + chunk)
            umi, m1, m2 = self._mates(a, b, need, device)
            parts = []
            if hdr:
                parts.append(umi[:, :hdr])
            parts += [m1[:, : take[0]], m2[:, : take[1]]]
            if cfg.umi_where == "file":
                uu = umi[:, : take[2]]
                if uu.shape[1] < take[2]:
                    pad = torch.full((uu.shape[0], take[2] - uu.shape[1]), 4, dtype=torch.uint8, device=device)
                    uu = torch.cat([uu, pad], dim=1)
                parts.append(uu)
            out[a - lo: b - lo] = table[torch.cat(parts, dim=1).long()]
        return out

    def write_fastq(self, directory: str, prefix: Optional[str] = None, chunk: int = 250_000, device="cpu") -> List[str]:
        """FastQ files of the whole configuration, for the executables (the same bytes whatever `device` the
        letters are drawn on)."""
        cfg = self.cfg
        prefix = prefix or cfg.name
        os.makedirs(directory, exist_ok=True)
        names = [os.path.join(directory, f"{prefix}_R{i + 1}.fastq") for i in range(2)]
        if cfg.umi_where == "file":
            names.append(os.path.join(directory, f"{prefix}_UMI.fastq"))
        files = [open(p, "wb") for p in names]
        letters = np.frombuffer(self._LETTERS, dtype=np.uint8)
        try:
            for lo in range(0, cfg.n_tuples, chunk):
                hi = min(cfg.n_tuples, lo + chunk)
                umi, m1, m2 = (letters[x.cpu().numpy()] for x in self._mates(lo, hi, READ_LEN, device))
                ids = np.arange(lo, hi, dtype=np.int64)
                for fi, seq in enumerate((m1, m2)):
                    files[fi].write(_records(_headers(cfg, ids, umi, fi + 1), seq).tobytes())
                if cfg.umi_where == "file":
                    files[2].write(_records(_headers(cfg, ids, umi, 3), umi[:, : cfg.umi_len]).tobytes())
        finally:
            for f in files:
                f.close()
        return names
