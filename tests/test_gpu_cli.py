"""End-to-end: the `humid` executable of this repo (host/humid_main.cc over libhumid_gpu)
against the UNMODIFIED reference CLI (oracle/_ref/humid, built from the reference's sources;
the binary travels to the GPU box) on the same FastQ files — every output file byte for byte
(`.gz` outputs compared after decompression, as the deflate encoders differ)."""
import filecmp
import gzip
import os
import subprocess

import numpy as np
import pytest

import oracle
from humid_b200 import build
from humid_b200.synth import Synth, baseline_config

pytestmark = pytest.mark.gpu

CASES = [  # (name, config, tuples, overrides, extra flags, gz inputs)
    ("c1_max", 0, 20000, {}, ["-a"], False),
    ("c1_dir_gz", 0, 8000, dict(maximum=False), ["-a"], True),
    ("c2_umifile", 1, 15000, {}, ["-a"], False),
    ("c3_ham2", 2, 8000, {}, [], False),
    ("c4_edit2", 3, 4000, {}, ["-a"], False),
    ("c5_umi10", 4, 15000, {}, [], False),
    ("n30", 1, 5000, dict(word_length=30), ["-a"], False),
    ("n36_umifile", 1, 15000, dict(word_length=36), ["-a"], False),  # all 12 UMI letters used (docs/usage.rst:18-20)
    ("n50_ham2_max", 2, 6000, dict(word_length=50, distance=2, maximum=True), ["-a"], False),
    ("no_dedup_output", 0, 3000, {}, ["-q"], False),
    ("c2_bgzf", 1, 12000, {}, ["-a"], "bgzf"),  # blocked gzip inputs: inflated by several threads (host/bgzf.h)
]


def read_any(path):
    return gzip.open(path, "rb").read() if path.endswith(".gz") else open(path, "rb").read()


@pytest.fixture(scope="module")
def humid_bin():
    return build.build_host()


@pytest.mark.parametrize("name,ci,R,kw,flags,gz", CASES, ids=[c[0] for c in CASES])
def test_cli_matches_reference_binary(tmp_path, humid_bin, name, ci, R, kw, flags, gz):
    cfg = baseline_config(ci, R, **kw)
    s = Synth(cfg)
    files = list(s.write_fastq(str(tmp_path / "in"), prefix="x", gzip_output=gz is True))
    if gz == "bgzf":
        from test_host_reader import bgzf_bytes
        for i, f in enumerate(files):
            with open(f + ".gz", "wb") as out:
                out.write(bgzf_bytes(open(f, "rb").read(), 60000))
            os.remove(f)
            files[i] = f + ".gz"
    ours = str(tmp_path / "ours" / "nested")
    cmd = [humid_bin, *cfg.cli_flags(), "-s", *flags, "-d", ours, "-l", str(tmp_path / "ours.log"), *files]
    subprocess.run(cmd, check=True, timeout=600)
    log = open(tmp_path / "ours.log").read()
    assert "Reading data... done." in log and "clusters... done." in log

    if oracle.have_ref():
        ref = str(tmp_path / "ref")
        ref_flags, skip = [], False
        for f in flags:  # `-g N` is not an option of the reference
            if skip or f == "-g":
                skip = f == "-g"
                continue
            if f != "-a":
                ref_flags.append(f)
        oracle.run_ref_humid(files, ref, cfg.cli_flags() + ref_flags, stats=True, annotate="-a" in flags)
        names = sorted(os.listdir(ref))
        names.remove("humid.log")
        assert sorted(os.listdir(ours)) == names
        for f in names:
            if f.endswith(".gz"):
                assert read_any(os.path.join(ours, f)) == read_any(os.path.join(ref, f)), f
            else:
                assert filecmp.cmp(os.path.join(ours, f), os.path.join(ref, f), shallow=False), f
    else:  # no reference binary on this machine: the oracle's statistics files
        hdr, take = s.layout()
        seg_take = ([hdr] if hdr else []) + take
        r = oracle.Result(oracle.pack_words(s.word_chars(), seg_take), cfg.word_length, cfg.distance, cfg.edit, cfg.maximum)
        for f, text in r.stats_text().items():
            assert open(os.path.join(ours, f)).read() == text, f
        if "-q" not in flags:
            kept = read_any(os.path.join(ours, os.path.basename(files[0]).replace(".", "_dedup.", 1))).count(b"\n") // 4
            assert kept == r.clusters


def test_cli_help_and_errors(humid_bin, tmp_path):
    p = subprocess.run([humid_bin, "-h"], capture_output=True, text=True)
    assert p.returncode == 0 and "-x" in p.stdout and "word length" in p.stdout
    p = subprocess.run([humid_bin, str(tmp_path / "missing.fastq")], capture_output=True, text=True)
    assert p.returncode != 0 and "ERROR" in p.stderr


# ---- several GPUs behind the same executable (`-g N` / $HUMID_GPUS; not in the reference) --------------------------
@pytest.fixture(scope="module")
def humid_small_batches(tmp_path_factory):
    """The executable built with batches of 2048 tuples (and output chunks of 512 records), so that a job of a few
    thousand tuples is dealt out over all the GPUs of a group."""
    build.build_host()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib_dir = os.path.join(root, "humid_b200")
    exe = os.path.join(str(tmp_path_factory.mktemp("cli_small")), "humid_small")
    subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", "-DHUMID_BATCH=2048", "-DHUMID_REWRITE_CHUNK=512", "-o", exe,
                    os.path.join(root, "host", "humid_main.cc"), "-L" + lib_dir, "-lhumid_gpu", "-lz",
                    "-Wl,-rpath," + lib_dir], check=True)
    return exe


MULTI_CASES = [c for c in CASES if c[0] in ("c1_max", "c2_umifile", "c3_ham2", "c5_umi10", "c1_dir_gz", "n36_umifile", "c4_edit2")]


@pytest.mark.parametrize("n_gpus", [2, 3])
@pytest.mark.parametrize("name,ci,R,kw,flags,gz", MULTI_CASES, ids=[c[0] for c in MULTI_CASES])
def test_cli_on_several_gpus_matches_reference_binary(tmp_path, humid_small_batches, monkeypatch, name, ci, R, kw, flags, gz,
                                                      n_gpus):
    """Same files, same outputs, with the word table shared over `n_gpus` handles: distinct GPUs when the box has them,
    else all on device 0 (the group's exchanges work through one device's memory just the same).  Words of more than 31
    letters and edit distance 2 fall back to one GPU inside the executable."""
    import torch

    have = torch.cuda.device_count()
    monkeypatch.setenv("HUMID_GPUS", str(n_gpus))
    monkeypatch.setenv("HUMID_GPU_DEVICES", ",".join(str(i % max(have, 1)) for i in range(n_gpus)))
    test_cli_matches_reference_binary(tmp_path, humid_small_batches, name, ci, R, kw, flags, gz)


def test_cli_gpu_option(tmp_path, humid_small_batches, monkeypatch):
    """`-g 2` on the command line does what $HUMID_GPUS does."""
    monkeypatch.setenv("HUMID_GPU_DEVICES", "0,0")
    name, ci, R, kw, flags, gz = CASES[2]
    test_cli_matches_reference_binary(tmp_path, humid_small_batches, name, ci, R, kw, ["-g", "2"] + flags, gz)
