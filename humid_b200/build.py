"""Builds humid_b200/libhumid_gpu.so from humid_b200/csrc/*.cu with nvcc for sm_100a.

In-tree output so that the library travels with the repo snapshot to the GPU box.
"""
from __future__ import annotations

import glob
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libhumid_gpu.so")
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-O2,-Wall,-Wno-unused-function", "--expt-relaxed-constexpr",
]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


STAMP = os.path.join(HERE, "build", "flags.stamp")


def extra_flags():
    """A/B builds: HUMID_NVCC_EXTRA="-DHUMID_RS_ITEMS=16 ..." — part of the staleness check below."""
    return os.environ.get("HUMID_NVCC_EXTRA", "").split()


def flags_signature() -> str:
    return " ".join(NVCC_FLAGS + extra_flags())


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    try:
        if open(STAMP).read() != flags_signature():
            return True
    except OSError:
        return True
    t = os.path.getmtime(LIB)
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(HERE, "..", "include", "*.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    objs, procs = [], []
    for src in sources():
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        cmd = [nvcc, *NVCC_FLAGS, *extra_flags(), "-c", src, "-o", obj] + (["-Xptxas", "-v"] if verbose else [])
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for cmd, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode:
            sys.stderr.write(out)
        if p.returncode:
            raise RuntimeError("nvcc failed: " + " ".join(cmd))
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB, *objs, "-cudart", "static"]
    subprocess.run(cmd, check=True)
    with open(STAMP, "w") as f:
        f.write(flags_signature())
    return LIB


HOST_BIN = os.path.join(HERE, "humid")


def build_host(force: bool = False) -> str:
    """The `humid` executable (host/humid_main.cc): same CLI as the reference, linked against
    libhumid_gpu.so (found next to the binary through $ORIGIN)."""
    root = os.path.dirname(HERE)
    import glob

    srcs = [os.path.join(root, "host", "humid_main.cc")] + sorted(glob.glob(os.path.join(root, "host", "*.h"))) + \
        sorted(glob.glob(os.path.join(root, "include", "*.h")))
    if not force and os.path.exists(HOST_BIN) and all(os.path.getmtime(s) <= os.path.getmtime(HOST_BIN) for s in srcs):
        return HOST_BIN
    build()
    cmd = ["g++", "-O2", "-std=c++17", "-Wall", "-pthread", "-o", HOST_BIN, srcs[0], "-L" + HERE, "-lhumid_gpu", "-lz",
           "-Wl,-rpath,$ORIGIN"]
    subprocess.run(cmd, check=True)
    return HOST_BIN


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_host(force="--force" in sys.argv))
