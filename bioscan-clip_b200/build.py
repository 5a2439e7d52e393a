"""Build libbsc_hot.so in-tree with nvcc for sm_100a (no torch extension machinery: the library is a
plain C-ABI shared object loaded with ctypes)."""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "bsc", "libbsc_hot.so")
SOURCES = ["api.cu", "prep.cu", "simt_f32.cu", "tc_sim.cu", "tc_bwd2.cu", "tc_bwd3.cu", "taxo.cu", "rescore.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
]


def _newer(a: str, b: str) -> bool:
    return not os.path.exists(b) or os.path.getmtime(a) > os.path.getmtime(b)


def build_lib(force: bool = False, verbose: bool = False, variant: str = "", defines=()) -> str:
    """variant / defines: an A/B build with extra -D flags, e.g. build_lib(variant="v2", defines=["-DBSC_SCAN_V2=1"]) ->
    bsc/libbsc_hot_v2.so (objects under build_v2/); BSC_LIB_PATH selects which library the package loads."""
    objdir = os.path.join(HERE, "build" + ("_" + variant if variant else ""))
    out = OUT if not variant else OUT.replace(".so", "_" + variant + ".so")
    os.makedirs(objdir, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(HERE, "..", "include", "bsc_hot.h"))
    objs = []
    jobs = []
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(objdir, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _newer(s, o) or any(_newer(h, o) for h in headers):
            cmd = [NVCC, *FLAGS, *defines, "-c", s, "-o", o]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            jobs.append(cmd)

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))
        return r.stderr

    with ThreadPoolExecutor(max_workers=min(4, max(1, len(jobs)))) as ex:
        for msg in ex.map(run, jobs):
            if verbose and msg:
                print(msg, file=sys.stderr)
    if jobs or not os.path.exists(out):
        link = [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", out, *objs,
                "-Xcompiler", "-fPIC", "-cudart", "static"]
        run(link)
    return out


if __name__ == "__main__":
    # python build.py [--force] [-v] [--variant NAME -DFLAG=1 ...]
    name = sys.argv[sys.argv.index("--variant") + 1] if "--variant" in sys.argv else ""
    print(build_lib(force="--force" in sys.argv, verbose="-v" in sys.argv, variant=name,
                    defines=[a for a in sys.argv[1:] if a.startswith("-D")]))
