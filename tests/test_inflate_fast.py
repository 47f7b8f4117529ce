"""host/inflate_fast.h (the `humid` executable's own deflate decoder + gzip reader) and
host/inflate_parallel.h (several threads on one gzip member: speculative block starts, place holders for the unknown
window, verified chain) against zlib; every case runs through both readers, the parallel one with 3 threads and spans
of 4 KiB so that small files already are many spans: the inflated bytes of every valid file are zlib's, whatever the compressor settings, header fields, member
layout or read size; damaged, cut and trailing-garbage files end with an error message and a non-zero exit status —
never a crash (the helper is built with the address and undefined-behaviour sanitizers).  The reference reads .gz
through isa-l (lib/fastp/src/fastqreader.cpp:79-133); what it observes is the inflated byte stream and "anything that
is not a gzip header behind a member is an error", which is what is pinned here."""
import gzip
import os
import struct
import subprocess
import zlib

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def tool(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("bin") / "inflate_check")
    subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-pthread", "-fsanitize=address,undefined", "-fno-sanitize-recover=all", "-o", exe,
                    os.path.join(HERE, "inflate_check.cc"), "-lz"], check=True)
    return exe


@pytest.fixture(params=["fast", "par3"])
def mode(request):
    return request.param


def fastq_text(n, rng):
    out = []
    for i in range(n):
        L = int(rng.integers(30, 151))
        out.append(b"@M0:%d:%d 1:N:0:ACGT\n" % (i, rng.integers(0, 99999)))
        out.append(bytes(rng.choice(list(b"ACGT"), L).tolist()) + b"\n+\n")
        out.append(bytes(rng.choice(list(b"FFFFFF:,#"), L).tolist()) + b"\n")
    return b"".join(out)


def skewed_bytes(n, rng):
    """All 256 byte values with geometrically falling frequencies: code lengths up to 15 bits, sub-tables in use."""
    p = 0.93 ** np.arange(256)
    return rng.choice(256, n, p=p / p.sum()).astype(np.uint8).tobytes()


DATA = {
    "empty": lambda rng: b"",
    "one_byte": lambda rng: b"A",
    "fastq": lambda rng: fastq_text(6000, rng),
    "zeros": lambda rng: bytes(700_000),
    "short_periods": lambda rng: b"".join(bytes(rng.integers(65, 70, p).tolist()) * int(rng.integers(2, 400)) for p in range(1, 12)) * 20,
    "random": lambda rng: rng.integers(0, 256, 300_000, dtype=np.uint8).tobytes(),
    "skewed": lambda rng: skewed_bytes(900_000, rng),
    "large_fastq": lambda rng: fastq_text(25000, rng),  # inflated size beyond the reader's output buffer several times
}


def deflate(data, level=6, strategy=zlib.Z_DEFAULT_STRATEGY, mem=8, chunks=1, flush=zlib.Z_NO_FLUSH):
    c = zlib.compressobj(level, zlib.DEFLATED, -15, mem, strategy)
    out = []
    step = max(1, (len(data) + chunks - 1) // chunks)
    for at in range(0, len(data), step):
        out.append(c.compress(data[at:at + step]))
        if flush != zlib.Z_NO_FLUSH:
            out.append(c.flush(flush))
    out.append(c.flush())
    return b"".join(out)


def member(data, raw=None, name=None, comment=None, extra=None, hcrc=False):
    flg = (4 if extra is not None else 0) | (8 if name is not None else 0) | (16 if comment is not None else 0) | (2 if hcrc else 0)
    head = b"\x1f\x8b\x08" + bytes([flg]) + b"\0\0\0\0\x00\x03"
    if extra is not None:
        head += struct.pack("<H", len(extra)) + extra
    if name is not None:
        head += name + b"\0"
    if comment is not None:
        head += comment + b"\0"
    if hcrc:
        head += struct.pack("<H", zlib.crc32(head) & 0xFFFF)
    return head + (deflate(data) if raw is None else raw) + struct.pack("<II", zlib.crc32(data), len(data) & 0xFFFFFFFF)


def run(tool, path, mode="fast", read_size=None, span=4096, stats=False):
    cmd = [tool, mode, path] + ([str(read_size)] if read_size else [])
    env = dict(os.environ, HUMID_INFLATE_SPAN=str(span))
    if stats:
        env["HUMID_INFLATE_STATS"] = "1"
    return subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)


def expect_ok(tool, mode, tmp_path, blob, data, read_size=None):
    path = str(tmp_path / "x.gz")
    with open(path, "wb") as f:
        f.write(blob)
    p = run(tool, path, mode, read_size=read_size)
    assert p.returncode == 0, p.stderr[-2000:]
    size, crc, _ = p.stdout.split()
    assert (int(size), int(crc, 16)) == (len(data), zlib.crc32(data))


def expect_error(tool, mode, tmp_path, blob, what=None):
    path = str(tmp_path / "bad.gz")
    with open(path, "wb") as f:
        f.write(blob)
    p = run(tool, path, mode)
    assert p.returncode > 0, (p.returncode, p.stdout, p.stderr[-2000:])  # > 0: exit(), not a signal (< 0)
    assert "Sanitizer" not in p.stderr and "runtime error" not in p.stderr, p.stderr[-3000:]
    assert p.stderr.startswith("ERROR: "), p.stderr[:300]
    if what:
        assert what in p.stderr, p.stderr[:300]


@pytest.mark.parametrize("kind", sorted(DATA))
def test_every_compressor_setting(tool, mode, tmp_path, kind):
    rng = np.random.default_rng(len(kind))
    data = DATA[kind](rng)
    settings = [dict(level=l) for l in (0, 1, 4, 6, 9)] + [
        dict(strategy=zlib.Z_FIXED), dict(strategy=zlib.Z_HUFFMAN_ONLY), dict(strategy=zlib.Z_RLE), dict(strategy=zlib.Z_FILTERED),
        dict(mem=1), dict(chunks=9, flush=zlib.Z_SYNC_FLUSH), dict(chunks=5, flush=zlib.Z_FULL_FLUSH), dict(level=0, chunks=3, flush=zlib.Z_SYNC_FLUSH)]
    for kw in settings:
        raw = deflate(data, **kw)
        assert zlib.decompress(raw, -15) == data
        expect_ok(tool, mode, tmp_path, member(data, raw=raw), data)


def test_read_sizes(tool, mode, tmp_path):
    rng = np.random.default_rng(5)
    data = fastq_text(3000, rng)
    blob = gzip.compress(data, 6)
    for read_size in (1, 7, 4096, 70_001, 1 << 22):
        expect_ok(tool, mode, tmp_path, blob, data, read_size=read_size)


def test_header_fields_and_members(tool, mode, tmp_path):
    rng = np.random.default_rng(6)
    a, b, c = fastq_text(500, rng), fastq_text(3, rng), skewed_bytes(50_000, rng)
    blob = (member(a, name=b"reads.fastq") + member(b"", comment=b"an empty member in the middle") +
            member(b, extra=b"AB\x03\x00xyz" + bytes(300), name=b"n", comment=b"c", hcrc=True) + member(c, hcrc=True) + member(b""))
    assert gzip.decompress(blob) == a + b + c
    expect_ok(tool, mode, tmp_path, blob, a + b + c)
    # a header with a name longer than one input buffer is still a header
    expect_ok(tool, mode, tmp_path, member(a, name=b"n" * 700_000) + member(b), a + b)
    # an empty file reads as nothing, like gzread does
    expect_ok(tool, mode, tmp_path, b"", b"")


def test_matches_do_not_reach_into_the_member_before(tool, mode, tmp_path):
    """A distance that points in front of the member's own output is an error even though an earlier member's output
    still sits in the reader's buffer."""
    a = b"ACGT" * 100
    bad = bytes([0b011, 0b00000010, 0b00100000]) + b"\0" * 2  # fixed block: a length-3 match at distance 1 as its first symbol
    expect_error(tool, mode, tmp_path, member(a) + member(b"AAA", raw=bad), "too far back")


def test_trailing_garbage_is_an_error_like_the_reference(tool, mode, tmp_path):
    data = b"@r\nACGT\n+\nIIII\n"
    expect_error(tool, mode, tmp_path, gzip.compress(data) + b"garbage", "invalid gzip header")
    expect_error(tool, mode, tmp_path, gzip.compress(data) + b"\0" * 512, "invalid gzip header")
    expect_error(tool, mode, tmp_path, b"not a gzip file at all", "invalid gzip header")


def test_cut_files(tool, mode, tmp_path):
    rng = np.random.default_rng(7)
    data = fastq_text(40, rng)
    one = gzip.compress(data, 6)
    blob = one + gzip.compress(b"tail\n")
    for cut in list(range(1, len(blob), 37)) + [len(blob) - 1, len(one) - 1, len(one) + 1]:
        if cut == len(one):
            continue
        expect_error(tool, mode, tmp_path, blob[:cut])
    expect_ok(tool, mode, tmp_path, blob[:len(one)], data)  # a cut between members leaves a valid file
    # stored blocks (level 0) and a file cut inside one
    stored = gzip.compress(rng.integers(0, 256, 200_000, dtype=np.uint8).tobytes(), 0)
    for cut in (20, 70_000, len(stored) - 9, len(stored) - 3):
        expect_error(tool, mode, tmp_path, stored[:cut])


def test_damage_never_crashes(tool, mode, tmp_path):
    rng = np.random.default_rng(8)
    texts = [fastq_text(300, rng), skewed_bytes(40_000, rng), bytes(50_000), rng.integers(0, 256, 30_000, dtype=np.uint8).tobytes()]
    blobs = [gzip.compress(t, lvl) for t in texts for lvl in (1, 6)] + [member(texts[0], raw=deflate(texts[0], strategy=zlib.Z_FIXED))]
    path = str(tmp_path / "d.gz")
    detected = 0
    trials = 0
    for blob in blobs:
        for _ in range(60):
            bad = bytearray(blob)
            for _ in range(int(rng.integers(1, 4))):
                at = int(rng.integers(0, len(bad)))
                bad[at] ^= 1 << int(rng.integers(0, 8))
            with open(path, "wb") as f:
                f.write(bad)
            p = run(tool, path, mode)
            trials += 1
            assert p.returncode >= 0, ("killed by a signal", p.stderr[-3000:])
            assert "Sanitizer" not in p.stderr and "runtime error" not in p.stderr, p.stderr[-3000:]
            if p.returncode:
                assert p.stderr.startswith("ERROR: "), p.stderr[:300]
                detected += 1
            else:  # harmless damage (MTIME, XFL, OS ... of the header): the data must still be right
                assert gzip.decompress(bytes(bad)) is not None
    assert detected > trials * 0.9


def test_same_bytes_as_zlib_on_a_dump(tool, tmp_path):
    rng = np.random.default_rng(9)
    data = fastq_text(2000, rng) + skewed_bytes(100_000, rng) + bytes(100_000)
    path, out = str(tmp_path / "x.gz"), str(tmp_path / "x.out")
    with open(path, "wb") as f:
        f.write(gzip.compress(data[:150_000], 9) + gzip.compress(data[150_000:], 1))
    subprocess.run([tool, "dump", path, out], check=True)
    assert open(out, "rb").read() == data


def test_speculations_are_used(tool, tmp_path):
    """On FastQ text from an ordinary compressor nearly every span is decoded ahead of the chain and used; stored and
    fixed-code blocks (nothing for the block finder) and a deflate stream hidden inside stored blocks (headers the finder
    does find, in the wrong place) fall back to decoding again from the verified position — same bytes either way."""
    rng = np.random.default_rng(11)
    data = fastq_text(30000, rng)
    path = str(tmp_path / "x.gz")

    def stats(blob, want):
        with open(path, "wb") as f:
            f.write(blob)
        p = run(tool, path, "par4", span=1 << 16, stats=True)
        assert p.returncode == 0, p.stderr[-2000:]
        size, crc, _ = p.stdout.split()
        assert (int(size), int(crc, 16)) == (len(want), zlib.crc32(want))
        words = p.stderr.split(":")[-1].replace(",", "").split()  # "<n> spans <u> speculations used <r> decoded again"
        return int(words[0]), int(words[2]), int(words[5])

    spans, used, redone = stats(gzip.compress(data, 6), data)
    assert spans > 10 and used >= spans - 2 and redone <= 1
    spans, used, redone = stats(member(data, raw=deflate(data, strategy=zlib.Z_FIXED)), data)
    assert spans > 10 and used == 0
    inner = gzip.compress(data, 6)
    spans, used, redone = stats(gzip.compress(inner, 0), inner)
    assert spans > 10 and used == 0 and redone >= spans - 2
    two = gzip.compress(data[:len(data) // 3], 6) + gzip.compress(data[len(data) // 3:], 1)
    spans, used, redone = stats(two, data)
    assert used >= spans - 4
