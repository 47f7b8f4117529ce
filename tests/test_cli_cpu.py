"""The HOST side of the `humid` executable (host/humid_main.cc: FastQ block parsing, .gz / BGZF input, batch assembly
across files, the parallel output pass, file names, statistics files) against the UNMODIFIED reference binary, on a
machine without a GPU: the executable is linked against tests/stub_humid_gpu.c, a test-only stand-in for libhumid_gpu
that answers the C ABI from the CPU oracle.  The device path is checked by tests/test_gpu_cli.py through the real
library; this file only makes the host logic part of the CPU suite."""
import os
import subprocess

import pytest

import oracle
import test_gpu_cli as G

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPU_CASES = [c for c in G.CASES if c[0] in ("c1_max", "c1_dir_gz", "c2_umifile", "c4_edit2", "n36_umifile", "no_dedup_output",
                                            "c2_bgzf")]


@pytest.fixture(scope="module")
def humid_cpu(tmp_path_factory):
    d = str(tmp_path_factory.mktemp("cpu_cli"))
    lib = os.path.join(d, "libhumid_gpu.so")
    subprocess.run(["gcc", "-O2", "-shared", "-fPIC", "-w", "-o", lib, os.path.join(ROOT, "tests", "stub_humid_gpu.c"),
                    os.path.join(ROOT, "oracle", "humid_oracle.c")], check=True)
    exe = os.path.join(d, "humid_cpu")
    # small batches and output chunks: every case runs over several of both
    subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", "-DHUMID_BATCH=2048", "-DHUMID_REWRITE_CHUNK=512", "-o", exe, os.path.join(ROOT, "host", "humid_main.cc"),
                    "-L" + d, "-lhumid_gpu", "-lz", "-Wl,-rpath," + d], check=True)
    return exe


@pytest.mark.skipif(not oracle.have_ref(), reason="reference binary not built on this machine")
@pytest.mark.parametrize("name,ci,R,kw,flags,gz", CPU_CASES, ids=[c[0] for c in CPU_CASES])
def test_host_side_matches_reference_binary(tmp_path, humid_cpu, name, ci, R, kw, flags, gz):
    G.test_cli_matches_reference_binary(tmp_path, humid_cpu, name, ci, R, kw, flags, gz)


def test_host_side_with_many_inflate_and_writer_threads(tmp_path, humid_cpu, monkeypatch):
    """150 000 tuples (73 batches of 2048, 293 output chunks of 512 in this build) in BGZF files and in ordinary .gz files with
    3 inflate threads per file, against the same input as plain files."""
    from humid_b200.synth import Synth, baseline_config
    from test_host_reader import bgzf_bytes
    cfg = baseline_config(1, 150_000)
    files = list(Synth(cfg).write_fastq(str(tmp_path / "in"), prefix="x", gzip_output=False))
    outs = {}
    for kind in ("plain", "bgzf", "gz", "plain_parallel"):
        use = files
        if kind == "plain_parallel":  # 4 parser threads per file, ranges of 64 KiB
            monkeypatch.setenv("HUMID_INGEST_THREADS", "4")
            monkeypatch.setenv("HUMID_INGEST_RANGE", "65536")
        if kind == "gz":  # ordinary one-member files, 3 threads on each (host/inflate_parallel.h), spans of 16 KiB
            import gzip
            use = []
            for f in files:
                with open(f + ".1.gz", "wb") as out:
                    out.write(gzip.compress(open(f, "rb").read(), 6))
                use.append(f + ".1.gz")
            monkeypatch.setenv("HUMID_INFLATE_THREADS", "3")
            monkeypatch.setenv("HUMID_INFLATE_SPAN", "16384")
            monkeypatch.setenv("HUMID_INFLATE_STATS", "1")
        if kind == "bgzf":
            use = []
            for f in files:
                with open(f + ".gz", "wb") as out:
                    out.write(bgzf_bytes(open(f, "rb").read(), 65000))
                use.append(f + ".gz")
            monkeypatch.setenv("HUMID_INFLATE_THREADS", "3")
        d = str(tmp_path / kind)
        p = subprocess.run([humid_cpu, *cfg.cli_flags(), "-s", "-a", "-d", d, "-l", str(tmp_path / (kind + ".log")), *use],
                           check=True, timeout=600, capture_output=True, text=True)
        if kind == "gz":  # both passes of every file went through the parallel reader and used its speculations
            lines = [l for l in p.stderr.splitlines() if l.startswith("inflate ")]
            assert len(lines) == 3 * len(files), p.stderr[-2000:]  # ingest, dedup output, annotated output: per file
            assert all(int(l.split(", ")[1].split()[0]) > 5 for l in lines), lines
        outs[kind] = {f.replace(".1.gz", "").replace(".gz", ""): G.read_any(os.path.join(d, f)) for f in sorted(os.listdir(d))}
    assert outs["plain"].keys() == outs["bgzf"].keys() == outs["gz"].keys() and len(outs["plain"]) == 4 + 2 * len(files)
    for f in outs["plain"]:
        assert outs["plain"][f] == outs["bgzf"][f], f
        assert outs["plain"][f] == outs["gz"][f], f
        assert outs["plain"][f] == outs["plain_parallel"][f], f
    dedup = [f for f in outs["plain"] if "_dedup" in f][0]
    assert 0 < outs["plain"][dedup].count(b"\n") // 4 < 150_000


ODD_INPUTS = ["crlf", "mixed_eol", "junk_between", "no_final_newline", "truncated_record", "length_mismatch", "long_lines",
              "only_newlines", "empty"]


PARALLEL_INGEST = dict(HUMID_INGEST_THREADS="3", HUMID_INGEST_RANGE="300")  # host/ingest_parallel.h on these small files


@pytest.mark.skipif(not oracle.have_ref(), reason="reference binary not built on this machine")
@pytest.mark.parametrize("case", ODD_INPUTS)
@pytest.mark.parametrize("gz", [False, True, "parallel"], ids=["plain", "gz", "plain_parallel_ingest"])
def test_odd_fastq_files_like_the_reference_binary(tmp_path, humid_cpu, case, gz):
    """Line endings, junk between records, truncated and malformed records (the inputs of tests/test_host_reader.py), two
    files each, through both executables: same exit status, same output files."""
    import gzip

    import numpy as np
    from test_host_reader import CASES

    env = dict(os.environ, **PARALLEL_INGEST) if gz == "parallel" else None
    gz = gz is True
    files = []
    for i in range(2):
        data = CASES[case](np.random.default_rng(len(case) + 7))  # both files hold the same records
        f = str(tmp_path / (f"r{i + 1}.fastq" + (".gz" if gz else "")))
        with (gzip.open(f, "wb") if gz else open(f, "wb")) as fh:
            fh.write(data)
        files.append(f)
    flags = ["-n", "12", "-m", "1"]
    ours, ref = str(tmp_path / "ours"), str(tmp_path / "ref")
    os.makedirs(ref)
    a = subprocess.run([humid_cpu, *flags, "-s", "-a", "-d", ours, "-l", str(tmp_path / "ours.log"), *files],
                       capture_output=True, timeout=300, env=env)
    b = subprocess.run([oracle.REF_HUMID, *flags, "-s", "-a", "-d", ref, "-l", os.path.join(ref, "humid.log"), *files],
                       capture_output=True, timeout=300)
    if b.returncode < 0:  # the reference dies of a signal on inputs without any record (SIGSEGV in preCompute's first
        assert a.returncode >= 0 and case in ("empty", "only_newlines")  # read): nothing to compare, ours must end normally
        return
    assert (a.returncode == 0) == (b.returncode == 0), (a.returncode, b.returncode, a.stderr[-300:], b.stderr[-300:])
    if b.returncode != 0:
        return
    names = sorted(os.listdir(ref))
    names.remove("humid.log")
    assert sorted(os.listdir(ours)) == names
    for f in names:
        assert G.read_any(os.path.join(ours, f)) == G.read_any(os.path.join(ref, f)), f


LAYOUTS = [  # (name, records per file, drop the first record (so that the first header carries a UMI), flags)
    ("one_file", [200], False, ["-n", "12"]),
    ("one_file_header_umi", [200], True, ["-n", "14"]),
    ("two_files_m2", [200, 200], True, ["-n", "16", "-m", "2"]),
    ("three_files_third_shorter_max", [150, 150, 90], True, ["-n", "21", "-x"]),
    ("two_files_edit", [150, 150], False, ["-n", "10", "-e"]),
    ("two_files_edit2_header_umi", [120, 120], True, ["-n", "13", "-e", "-m", "2"]),
    ("wide_header_umi", [150, 150], True, ["-n", "40"]),
    ("distance_zero", [200], False, ["-n", "8", "-m", "0"]),
    ("umi_longer_than_word", [100, 100], True, ["-n", "5"]),
    ("annotate_only", [100, 100], False, ["-n", "12", "-q"]),
]


@pytest.mark.skipif(not oracle.have_ref(), reason="reference binary not built on this machine")
@pytest.mark.parametrize("name,counts,drop_first,flags", LAYOUTS, ids=[c[0] for c in LAYOUTS])
@pytest.mark.parametrize("ingest", ["one_thread_per_file", "parallel_ingest"])
def test_word_layouts_like_the_reference_binary(tmp_path, humid_cpu, name, counts, drop_first, flags, ingest):
    """preCompute's split of the word over header UMI and files (remainder to the last file), header UMI found or
    not, files of different lengths (the tuple stream ends with the shortest), reads shorter than their share and
    reads with N (filtered): every output file as the reference binary writes it."""
    import numpy as np
    from test_host_reader import records

    rng = np.random.default_rng(len(name))
    files = []
    for i, n in enumerate(counts):
        f = str(tmp_path / f"r{i + 1}.fastq")
        with open(f, "wb") as fh:
            fh.write(b"".join(records(n + 1, rng)[1 if drop_first else 0:][:n]))
        files.append(f)
    ours, ref = str(tmp_path / "ours"), str(tmp_path / "ref")
    os.makedirs(ref)
    env = dict(os.environ, **PARALLEL_INGEST, HUMID_TIMING="1") if ingest == "parallel_ingest" else None
    p = subprocess.run([humid_cpu, *flags, "-s", "-a", "-d", ours, "-l", str(tmp_path / "ours.log"), *files], check=True, timeout=300,
                       env=env, capture_output=True, text=True)
    if env:  # every file went through the parser pool, in many ranges
        lines = [l for l in p.stderr.splitlines() if l.startswith("ingest ")]
        assert len(lines) == len(files) and all(int(l.split(": ")[1].split()[0]) > 10 for l in lines), p.stderr[-1500:]
    subprocess.run([oracle.REF_HUMID, *flags, "-s", "-a", "-d", ref, "-l", os.path.join(ref, "humid.log"), *files], check=True,
                   timeout=300)
    names = sorted(os.listdir(ref))
    names.remove("humid.log")
    assert sorted(os.listdir(ours)) == names
    for f in names:
        assert open(os.path.join(ours, f), "rb").read() == open(os.path.join(ref, f), "rb").read(), f


@pytest.mark.skipif(not oracle.have_ref(), reason="reference binary not built on this machine")
@pytest.mark.parametrize("fname,outdir", [("sample.R1.fastq", "out/"), ("a.b.c.fq", "deep/er/dir"), (".hidden.fastq", "out")])
def test_output_file_names_like_the_reference_binary(tmp_path, humid_cpu, fname, outdir):
    """makeFileName (src/fastq.cc:170-181): the suffix goes in front of the FIRST dot of the base name; the output
    directory is created with its parents.  (A name without any dot makes the reference throw std::out_of_range from
    substr; this executable writes <name>_dedup there — not compared.)"""
    import numpy as np
    from test_host_reader import records

    src = tmp_path / "in"
    src.mkdir()
    f = str(src / fname)
    with open(f, "wb") as fh:
        fh.write(b"".join(records(60, np.random.default_rng(4))))
    listing = {}
    for who, exe in (("ours", humid_cpu), ("ref", oracle.REF_HUMID)):
        top = tmp_path / who
        top.mkdir()
        subprocess.run([exe, "-n", "10", "-s", "-a", "-d", outdir, "-l", "run.log", f], check=True, timeout=120, cwd=str(top))
        listing[who] = sorted(os.path.relpath(os.path.join(d, x), str(top)) for d, _, fs in os.walk(str(top)) for x in fs)
    assert listing["ours"] == listing["ref"]
    for rel in listing["ref"]:
        if not rel.endswith("run.log"):
            assert open(tmp_path / "ours" / rel, "rb").read() == open(tmp_path / "ref" / rel, "rb").read(), rel


@pytest.mark.parametrize("flags,what", [(["-n", "64"], "word length"), (["-m", "9"], "Hamming distance"),
                                        (["-e", "-m", "4"], "edit distance"), (["-e", "-m", "2", "-n", "40"], "edit distance"),
                                        (["-n", "0"], "word length")])
def test_limits_are_refused_before_any_file_is_read(humid_cpu, tmp_path, flags, what):
    """What the device path does not cover (the reference takes any -n / -m) fails at once with a message that names the
    option — not after the first pass over the inputs, and not with an error about the (missing) file."""
    p = subprocess.run([humid_cpu, *flags, "-d", str(tmp_path / "out"), "-l", str(tmp_path / "log"), str(tmp_path / "missing.fastq")],
                       capture_output=True, text=True)
    assert p.returncode != 0 and "ERROR" in p.stderr and what in p.stderr and "missing.fastq" not in p.stderr
    assert not os.path.exists(tmp_path / "out")
    q = subprocess.run([humid_cpu, "-h"], capture_output=True, text=True)
    assert "limits of this implementation" in q.stdout
