"""CPU-only checks of the drop-in boundary: the library builds, loads, exports every symbol
include/humid_gpu.h declares, and refuses to run without a CUDA device (no CPU fallback)."""
import ctypes
import os
import re

import pytest

from humid_b200 import api, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build.build()
    return api.load_library()


def declared_symbols():
    import glob

    syms = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        text = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        syms |= set(re.findall(r"\b(humid_gpu_[a-z_0-9]+)\s*\(", text))
    return sorted(syms)


def test_header_symbols_exported(lib):
    syms = declared_symbols()
    assert len(syms) >= 16
    for s in syms:
        assert hasattr(lib, s), s
    from humid_b200 import sharded

    assert sorted(api.EXPORTS + sharded.SHARD_EXPORTS + sharded.GROUP_EXPORTS) == syms


def test_abi_version(lib):
    assert lib.humid_gpu_abi_version() == 1


def test_config_struct_layout():
    # must match humid_gpu_config in include/humid_gpu.h
    assert ctypes.sizeof(api.Config) == 96  # 12 + 64 + 4 flags + 4 device + 4 pad + 8
    assert api.Config.seg_take.offset == 12
    assert api.Config.edit.offset == 76
    assert api.Config.device.offset == 80
    assert api.Config.expected_tuples.offset == 88
    assert ctypes.sizeof(api.StageStat) == 32 + 4 + 4 + 8 + 8


def test_no_cpu_fallback(lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(api.HumidGpuError) as e:
        api.Dedup(word_length=24, distance=1, seg_take=[12, 12])
    assert "failed (-2)" in str(e.value)


def test_product_never_touches_oracle():
    """The oracle is test infrastructure: nothing under humid_b200/, include/ or host/ may
    import, include or link it."""
    bad = []
    for base in ("humid_b200", "include", "host"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            if "build" in dp.split(os.sep):
                continue
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".cpp", "Makefile")):
                    t = open(os.path.join(dp, f), errors="ignore").read()
                    if re.search(r"^\s*(import|from)\s+oracle\b|oracle/|liboracle", t, flags=re.M):
                        bad.append(os.path.join(dp, f))
    assert not bad, bad
