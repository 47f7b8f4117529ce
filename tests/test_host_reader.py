"""host/fastq_block.h (block parser of the `humid` executable) against host/fastq_io.h (the line-by-line
reader whose behaviour was pinned against the reference binary): same records, whatever the line
endings, junk lines, truncation or block size."""
import gzip
import os
import struct
import subprocess
import zlib

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def checker(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("bin") / "host_reader_check")
    subprocess.run(["g++", "-O1", "-std=c++17", "-pthread", "-DHUMID_REWRITE_CHUNK=7", "-o", exe,
                    os.path.join(HERE, "host_reader_check.cc"), "-lz"], check=True)
    return exe


def records(n, rng, eol=b"\n", umi=True):
    out = []
    for i in range(n):
        L = int(rng.integers(0, 40))
        seq = bytes(rng.choice(list(b"ACGTN"), L).tolist())
        name = b"@r%d:%d" % (i, rng.integers(0, 99)) + (b"_" + bytes(rng.choice(list(b"ACGT"), 8).tolist()) if umi and i % 3 else b"") + \
            (b" 1:N:0" if i % 2 else b"")
        out.append(name + eol + seq + eol + b"+" + eol + b"I" * L + eol)
    return out


CASES = {
    "plain": lambda rng: b"".join(records(300, rng)),
    "crlf": lambda rng: b"".join(records(300, rng, b"\r\n")),
    "lone_cr": lambda rng: b"".join(records(50, rng, b"\r")),
    "mixed_eol": lambda rng: b"".join(r for k in range(60) for r in records(1, rng, [b"\n", b"\r\n", b"\r"][k % 3])),
    "no_final_newline": lambda rng: b"".join(records(20, rng))[:-1],
    "junk_between": lambda rng: b"\n\nhello\n" + b"garbage\n\n".join(records(40, rng)) + b"\n\ntrailing junk",
    "truncated_record": lambda rng: b"".join(records(10, rng)) + b"@last\nACGT\n+\n",
    "length_mismatch": lambda rng: b"".join(records(5, rng)) + b"@bad\nACGT\n+\nII\n" + b"".join(records(5, rng)),
    "empty": lambda rng: b"",
    "only_newlines": lambda rng: b"\n\n\r\n\n",
    "long_lines": lambda rng: b"@x\n" + b"A" * 5000 + b"\n+\n" + b"I" * 5000 + b"\n" + b"".join(records(3, rng)),
}


@pytest.mark.parametrize("case", sorted(CASES))
@pytest.mark.parametrize("gz", [False, True])
def test_block_reader_matches_line_reader(checker, tmp_path, case, gz):
    data = CASES[case](np.random.default_rng(len(case)))
    path = str(tmp_path / ("in.fastq.gz" if gz else "in.fastq"))
    with (gzip.open(path, "wb") if gz else open(path, "wb")) as f:
        f.write(data)
    ref = subprocess.run([checker, "old", path], capture_output=True, check=True)
    for block in (16, 64, 1000, 1 << 20):
        got = subprocess.run([checker, "block", path, str(block)], capture_output=True, check=True)
        assert got.stdout == ref.stdout, (case, block)
        assert (b"different length" in got.stderr) == (b"different length" in ref.stderr)
    if case in ("plain", "crlf"):
        assert ref.stdout.endswith(b"records 300\n")


def test_lone_cr_files_are_streamed(checker, tmp_path):
    """A file whose lines end with '\\r' only has no '\\n' to wait for: the block reader must hand its records out block
    by block instead of growing its buffer to the size of the file."""
    data = b"".join(records(4000, np.random.default_rng(3), b"\r"))
    path = str(tmp_path / "in.fastq")
    with open(path, "wb") as f:
        f.write(data)
    ref = subprocess.run([checker, "old", path], capture_output=True, check=True)
    got = subprocess.run([checker, "block", path, "256"], capture_output=True, check=True)
    assert got.stdout == ref.stdout and ref.stdout.endswith(b"records 4000\n")
    grown_to = int(got.stderr.split(b"buffer ")[1].split()[0])
    assert grown_to <= 1024 < len(data) // 100, grown_to


@pytest.mark.parametrize("case", ["plain", "crlf", "mixed_eol", "junk_between", "no_final_newline", "long_lines", "empty",
                                  "truncated_record"])
@pytest.mark.parametrize("annotate", [0, 1])
def test_parallel_rewrite_matches_one_thread(checker, tmp_path, case, annotate):
    """Pass 2 cut into chunks of 7 records over 1..5 threads writes the bytes of the one-thread pass
    (host/fastq_rewrite.h), whatever the line endings and junk between records."""
    data = CASES[case](np.random.default_rng(len(case) + 3))
    path = str(tmp_path / "in.fastq")
    with open(path, "wb") as f:
        f.write(data)
    gathered = {}
    for threads in (1, 2, 5):
        a, b = str(tmp_path / "one.fastq"), str(tmp_path / f"par{threads}.fastq")
        env = dict(os.environ, HUMID_NO_MMAP="1") if threads == 2 else None  # (gathered with pread instead of from a mapping)
        p = subprocess.run([checker, "rewrite", path, a, b, str(threads + 10 * annotate), str(annotate), str(threads)], check=True,
                           capture_output=True, text=True, env=env)
        assert open(a, "rb").read() == open(b, "rb").read(), (case, threads)
        words = p.stdout.split()
        gathered = dict(chunks=int(words[3]), gathered=int(words[5]))
    if case == "plain" and annotate:
        assert open(a, "rb").read().count(b"\n") == 4 * 300  # every record, with ":id" behind its name
    # chunks stored exactly as they are written back are gathered by record lengths (no second parse); any other goes
    # through the parser: all of "plain", none of "crlf", some of the mixed ones
    if case in ("plain", "long_lines"):
        assert gathered["gathered"] == gathered["chunks"] > 0
    if case == "crlf":
        assert gathered["gathered"] == 0
    if case in ("mixed_eol", "junk_between", "no_final_newline"):
        assert gathered["gathered"] < gathered["chunks"]


@pytest.mark.parametrize("case", ["plain", "mixed_eol", "empty", "truncated_record"])
@pytest.mark.parametrize("gz_in,gz_out", [(False, True), (True, False), (True, True)])
def test_parallel_rewrite_of_compressed_files(checker, tmp_path, case, gz_in, gz_out):
    """A .gz on either side of pass 2: one thread inflates + parses a .gz input, the pool deflates a .gz output as
    one gzip member per chunk of 7 records — the inflated bytes are those of the one-thread pass."""
    data = CASES[case](np.random.default_rng(len(case) + 5))
    path = str(tmp_path / ("in.fastq.gz" if gz_in else "in.fastq"))
    with (gzip.open(path, "wb") if gz_in else open(path, "wb")) as f:
        f.write(data)
    ext = ".fastq.gz" if gz_out else ".fastq"
    read = (lambda p: gzip.open(p, "rb").read()) if gz_out else (lambda p: open(p, "rb").read())
    for annotate in (0, 1):
        for threads in (2, 4):
            a, b = str(tmp_path / ("one" + ext)), str(tmp_path / (f"par{threads}" + ext))
            subprocess.run([checker, "rewrite", path, a, b, str(threads + 10 * annotate), str(annotate), str(threads)],
                           check=True)
            assert read(a) == read(b), (case, annotate, threads)
            if gz_out and case == "plain":
                raw = open(b, "rb").read()
                assert raw.count(b"\x1f\x8b\x08") >= 300 * (1 if annotate else 0) // 7  # several members


def bgzf_bytes(data, block=700):
    """`data` as a BGZF file (the layout bgzip writes): gzip members with a 'BC' extra field holding the member size,
    at most `block` inflated bytes each, and the empty end-of-file member."""
    out = []
    for a in list(range(0, len(data), block)) + [None]:
        chunk = b"" if a is None else data[a:a + block]
        c = zlib.compressobj(6, zlib.DEFLATED, -15)
        body = c.compress(chunk) + c.flush()
        bsize = 12 + 6 + len(body) + 8 - 1
        out.append(b"\x1f\x8b\x08\x04" + b"\0" * 4 + b"\x00\xff" + struct.pack("<H", 6) + b"BC" + struct.pack("<HH", 2, bsize) +
                   body + struct.pack("<II", zlib.crc32(chunk), len(chunk)))
    return b"".join(out)


@pytest.mark.parametrize("case", ["plain", "mixed_eol", "junk_between", "long_lines", "empty", "truncated_record"])
def test_bgzf_inflated_by_several_threads(checker, tmp_path, case):
    """A BGZF input (host/bgzf.h: members inflated side by side) reads as the same records as the same bytes through
    the one-stream path, with 2..5 inflate threads and jobs of one or many members."""
    data = CASES[case](np.random.default_rng(len(case) + 11))
    if case == "plain":
        data = data * 40  # several jobs' worth of members
    path = str(tmp_path / "in.fastq.gz")
    with open(path, "wb") as f:
        f.write(bgzf_bytes(data, 700 if case != "plain" else 60000))
    assert gzip.open(path, "rb").read() == data  # a valid multi-member gzip
    env = dict(os.environ, HUMID_INFLATE_THREADS="1")
    ref = subprocess.run([checker, "block", path, "4096"], capture_output=True, check=True, env=env)
    old = subprocess.run([checker, "old", path], capture_output=True, check=True, env=env)
    assert ref.stdout == old.stdout
    for threads in (2, 3, 5):
        env = dict(os.environ, HUMID_INFLATE_THREADS=str(threads))
        got = subprocess.run([checker, "block", path, "4096"], capture_output=True, check=True, env=env)
        assert got.stdout == ref.stdout, (case, threads)


def test_bgzf_damage_is_an_error(checker, tmp_path):
    data = CASES["plain"](np.random.default_rng(5))
    raw = bytearray(bgzf_bytes(data, 700))
    raw[len(raw) // 2] ^= 0x55  # somewhere inside a member
    path = str(tmp_path / "in.fastq.gz")
    with open(path, "wb") as f:
        f.write(bytes(raw))
    env = dict(os.environ, HUMID_INFLATE_THREADS="3")
    got = subprocess.run([checker, "block", path, "4096"], capture_output=True, env=env)
    assert got.returncode != 0 and b"decompressing" in got.stderr


def test_bgzf_followed_by_plain_gzip_members(checker, tmp_path):
    """`cat a.bgz b.gz`: valid gzip that starts like BGZF.  It cannot be cut at member headers, so it takes the one-stream
    path whatever the number of inflate threads (the reference reads such files, lib/fastp/src/fastqreader.cpp:79-133)."""
    rng = np.random.default_rng(17)
    a, b = b"".join(records(200, rng)), b"".join(records(150, rng))
    raw = bgzf_bytes(a, 700)
    raw = raw[:-28] + gzip.compress(b)  # without the empty end-of-file member, then an ordinary member
    path = str(tmp_path / "in.fastq.gz")
    with open(path, "wb") as f:
        f.write(raw)
    assert gzip.open(path, "rb").read() == a + b
    outs = []
    for threads in (1, 4):
        env = dict(os.environ, HUMID_INFLATE_THREADS=str(threads))
        got = subprocess.run([checker, "block", path, "4096"], capture_output=True, check=True, env=env)
        outs.append(got.stdout)
    assert outs[0] == outs[1] and outs[0].endswith(b"records 350\n")


def test_bgzf_member_claiming_more_than_64k_is_not_bgzf(checker, tmp_path):
    """A trailer with ISIZE above the BGZF bound: the file is not treated as BGZF (no 4 GiB buffer is sized from it); the
    one-stream reader then reports the damage like any other."""
    data = CASES["plain"](np.random.default_rng(5))
    raw = bytearray(bgzf_bytes(data, 700))
    first = 12 + 6 + struct.unpack("<H", raw[16:18])[0] + 1 - 18  # end of the first member
    raw[first - 4:first] = struct.pack("<I", 0xFFFFFFF0)
    path = str(tmp_path / "in.fastq.gz")
    with open(path, "wb") as f:
        f.write(bytes(raw))
    env = dict(os.environ, HUMID_INFLATE_THREADS="3")
    got = subprocess.run([checker, "block", path, "4096"], capture_output=True, env=env)
    assert got.returncode != 0 and b"decompressing" in got.stderr


@pytest.mark.parametrize("case", sorted(CASES))
def test_parallel_parser_matches_the_sequential_one(checker, tmp_path, case):
    """host/ingest_parallel.h (several threads on one plain file: guessed record starts, commit in file order) gives the
    records of the sequential block reader — same order, offsets, stored lengths, fields — whatever the range size, for
    every odd input; the message of a malformed record appears exactly when the sequential reader prints it."""
    data = CASES[case](np.random.default_rng(len(case) + 7))
    if case in ("plain", "crlf", "junk_between"):
        data = data * 5
    path = str(tmp_path / "in.fastq")
    with open(path, "wb") as f:
        f.write(data)
    ref = subprocess.run([checker, "hash_seq", path], capture_output=True, check=True)
    for rng_bytes, threads, no_mmap in ((64, 3, "0"), (333, 2, "1"), (4096, 5, "0"), (1 << 25, 2, "0")):
        env = dict(os.environ, HUMID_INGEST_RANGE=str(rng_bytes))
        if no_mmap == "1":  # the file read in blocks instead of parsed out of one shared mapping
            env["HUMID_NO_MMAP"] = "1"
        got = subprocess.run([checker, "hash_par", path, str(threads)], capture_output=True, check=True, env=env)
        assert got.stdout == ref.stdout, (case, rng_bytes)
        assert (b"different length" in got.stderr) == (b"different length" in ref.stderr), (case, rng_bytes)
        if case == "plain" and rng_bytes == 4096:  # well-formed text: every guess is right
            assert b" reparsed 0" in got.stderr and b"ranges 1 " not in got.stderr, got.stderr


def test_parallel_parser_with_misleading_lines(checker, tmp_path):
    """Quality lines that start with '@' or '+', strand lines that repeat the name, sequences of one letter, records of
    very different sizes, junk lines between records: whatever a range's guessed first record is ('@' line whose second-next
    line starts with '+'), the commit only uses a range that was entered where the range before it says the next record starts
    and parses it again otherwise, so the records are the sequential reader's."""
    rng = np.random.default_rng(2024)
    out = []
    for i in range(6000):
        L = int(rng.choice([0, 1, 2, 5, 40, 150, 700]))
        seq = bytes(rng.choice(list(b"ACGTN"), L).tolist())
        first = rng.choice(list(b"@+IF#"))
        qual = (bytes([first]) + bytes(rng.choice(list(b"@+IF#:,"), L - 1).tolist())) if L else b""
        strand = b"+" + (b"r%d" % i if i % 4 == 0 else b"")
        out.append(b"@r%d extra@+fields\n" % i + seq + b"\n" + strand + b"\n" + qual + b"\n")
        if i % 97 == 0:
            out.append(b"+junk line\n\nmore junk @ here\n")  # skipped by the reader (no '@' at a line's start)
    data = b"".join(out)
    path = str(tmp_path / "in.fastq")
    with open(path, "wb") as f:
        f.write(data)
    ref = subprocess.run([checker, "hash_seq", path], capture_output=True, check=True)
    assert ref.stdout.count(b"\n") > 100
    for rng_bytes, threads in ((200, 4), (1000, 3), (4096, 6), (50_000, 2)):
        got = subprocess.run([checker, "hash_par", path, str(threads)], capture_output=True, check=True,
                             env=dict(os.environ, HUMID_INGEST_RANGE=str(rng_bytes)))
        assert got.stdout == ref.stdout, rng_bytes
        assert (b"different length" in got.stderr) == (b"different length" in ref.stderr)
    # '\r' line ends leave nothing to guess from (no '\n'): every range is parsed in its turn, same records
    cr = str(tmp_path / "cr.fastq")
    with open(cr, "wb") as f:
        f.write(data.replace(b"\n", b"\r"))
    ref = subprocess.run([checker, "hash_seq", cr], capture_output=True, check=True)
    got = subprocess.run([checker, "hash_par", cr, "4"], capture_output=True, check=True, env=dict(os.environ, HUMID_INGEST_RANGE="3000"))
    assert got.stdout == ref.stdout
    assert int(got.stderr.split(b" reparsed ")[1].split()[0]) > 0
