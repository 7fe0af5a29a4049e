"""Build libaccel_ps.so (the C-ABI library of include/accel_ps.h) in-tree for sm_100a.

nvcc is called directly — one object per source, in parallel — and the objects
are linked into kubeflow-controller_b200/libaccel_ps.so. The .so is git-ignored
but travels to the GPU box with the repo snapshot. No GPU is needed to build.
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libaccel_ps.so")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = [
    "-O3", "-std=c++17", "-lineinfo", *ARCH,
    "-Xcompiler", "-fPIC,-ffp-contract=off,-Wall",
    "-I", os.path.join(ROOT, "include"), "-I", CSRC,
]


def nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found; libaccel_ps.so cannot be built")
    return exe


def sources() -> list[str]:
    return sorted(
        os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp"))
    )


def headers_mtime() -> float:
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".hpp", ".cuh"))]
    hs.append(os.path.join(ROOT, "include", "accel_ps.h"))
    hs.append(os.path.abspath(__file__))
    return max(os.path.getmtime(h) for h in hs)


def compile_one(src: str, hdr_m: float, verbose: bool) -> tuple[str, bool, str]:
    obj = os.path.join(OBJ, os.path.basename(src) + ".o")
    if os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(src), hdr_m):
        return obj, False, ""
    cmd = [nvcc(), *NVCC_FLAGS, "-x", "cu", "-c", src, "-o", obj]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    p = subprocess.run(cmd, capture_output=True, text=True)
    if p.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src}:\n{p.stdout}\n{p.stderr}")
    return obj, True, p.stderr


def build(verbose: bool = False, force: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    if force:
        for f in os.listdir(OBJ):
            os.unlink(os.path.join(OBJ, f))
    hdr_m = headers_mtime()
    srcs = sources()
    rebuilt = False
    logs = []
    with cf.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        objs = []
        for obj, did, log in ex.map(lambda s: compile_one(s, hdr_m, verbose), srcs):
            objs.append(obj)
            rebuilt |= did
            if log:
                logs.append(log)
    if rebuilt or not os.path.exists(LIB):
        cmd = [nvcc(), "-shared", *ARCH, "-o", LIB, *objs, "-cudart", "static",
               "-Xlinker", "--no-undefined"]
        p = subprocess.run(cmd, capture_output=True, text=True)
        if p.returncode != 0:
            raise RuntimeError(f"link failed:\n{p.stdout}\n{p.stderr}")
    if verbose:
        sys.stderr.write("\n".join(logs))
    build_probe()
    return LIB


def build_probe() -> str:
    """tools/probe/libaccel_probe.so: link-rate probes for development, kept OUT of the product
    library (nothing in the package loads it)."""
    src = os.path.join(ROOT, "tools", "probe", "ps_probe.cu")
    out = os.path.join(ROOT, "tools", "probe", "libaccel_probe.so")
    if not os.path.exists(src):
        return ""
    if os.path.exists(out) and os.path.getmtime(out) > os.path.getmtime(src):
        return out
    cmd = [nvcc(), "-O3", "-std=c++17", "-lineinfo", *ARCH, "-Xcompiler", "-fPIC", "-shared",
           "-cudart", "static", "-o", out, src]
    p = subprocess.run(cmd, capture_output=True, text=True)
    if p.returncode != 0:
        raise RuntimeError(f"nvcc failed on {src}:\n{p.stdout}\n{p.stderr}")
    return out


if __name__ == "__main__":
    path = build(verbose="-v" in sys.argv, force="-f" in sys.argv)
    print(path)
