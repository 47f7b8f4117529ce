"""A/B builds of libhumid_gpu.so: one library per set of -D build parameters, all measured in ONE gpurun call.

    python tools/ab_build.py name1:file:-DMACRO=1,-DOTHER=2 name2:file:-D...      (here, no GPU needed)
    gpurun -- 'bash tools/ab_run.sh 1 50000000 "" name1 name2'                    (on the box)

`file` is the csrc/*.cu that reads the macros (primitives, words, neighbours, cluster, ...): only that object is
recompiled, the others come from humid_b200/build/ (run `python -m humid_b200.build` first).  Libraries land in
ab/lib_<name>.so (git-ignored, shipped by gpurun); `HUMID_GPU_LIB` makes humid_b200/api.py load one of them.  The build
parameters and their defaults are listed in INTEGRATION.md."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from humid_b200 import build  # noqa: E402

OBJ = os.path.join(ROOT, "humid_b200", "build")
AB = os.path.join(ROOT, "ab")


def make(spec: str) -> str:
    name, src, flags = spec.split(":", 2)
    flags = [f for f in flags.split(",") if f]
    obj = os.path.join(AB, name + ".o")
    lib = os.path.join(AB, "lib_" + name + ".so")
    nvcc = "/usr/local/cuda/bin/nvcc"
    cmd = [nvcc, *build.NVCC_FLAGS, *flags, "-Xptxas", "-v", "-c", os.path.join(build.CSRC, src + ".cu"), "-o", obj]
    out = subprocess.run(cmd, check=True, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True).stdout
    spills = [l.strip() for l in out.splitlines() if "spill stores" in l and not l.strip().startswith("0 bytes stack frame, 0 bytes spill")]
    others = [os.path.join(OBJ, f) for f in sorted(os.listdir(OBJ)) if f.endswith(".o") and f != src + ".o"]
    subprocess.run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", lib, *others, obj, "-cudart", "static"],
                   check=True)
    return f"{name}: {' '.join(flags)} -> {os.path.relpath(lib, ROOT)} ({len(spills)} kernels spill)"


if __name__ == "__main__":
    if len(sys.argv) < 2:
        sys.exit(__doc__)
    build.build()
    os.makedirs(AB, exist_ok=True)
    with ThreadPoolExecutor(max_workers=8) as pool:
        for line in pool.map(make, sys.argv[1:]):
            print(line)
