"""Build libyc_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libyc_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC,-fvisibility=hidden", "--expt-relaxed-constexpr", "-ccbin", "/usr/bin/g++"]


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp")))


def _stale(obj, src):
    if not os.path.exists(obj):
        return True
    t = os.path.getmtime(obj)
    deps = [src] + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h", ".hpp"))]
    deps.append(os.path.join(HERE, "..", "include", "yc_b200.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(verbose: bool = False, force: bool = False) -> str:
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    jobs = []
    objs = []
    for f in sources():
        src = os.path.join(CSRC, f)
        obj = os.path.join(objdir, f.rsplit(".", 1)[0] + ".o")
        objs.append(obj)
        if force or _stale(obj, src):
            cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-x", "cu", "-c", src, "-o", obj]
            jobs.append(cmd)

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        return r.stderr

    if jobs:
        with ThreadPoolExecutor(max_workers=8) as ex:
            for out in ex.map(run, jobs):
                if verbose and out:
                    print(out)
    if jobs or not os.path.exists(OUT):
        cmd = [NVCC, "-shared", "-o", OUT] + objs + ["-ccbin", "/usr/bin/g++", "-lcudart_static", "-lpthread", "-ldl", "-lrt"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    return OUT


SHIM_SRC = os.path.join(HERE, "shim", "yolo_core_b200.cpp")
SHIM_OUT = os.path.join(HERE, "libyolo_core_b200.so")
REF = "/root/reference"


def build_shim(force: bool = False):
    """libyolo_core_b200.so: the reference's own entry points (LoadNetwork, NetworkPredict, GetNetworkBoxes, NmsSort,
    Forward*LayerGpu(layer*, NetworkState)) on top of libyc_b200.so, compiled against the reference's public header
    where it lies (<reference>/src/yolo_core.h; nothing of the reference is copied or linked).  Only possible where the
    reference tree exists (this container); the built .so travels to the GPU box with the snapshot.  Returns the path,
    or None when the header is not available and no prebuilt library exists."""
    hdr = os.path.join(REF, "src", "yolo_core.h")
    if not os.path.exists(hdr):
        return SHIM_OUT if os.path.exists(SHIM_OUT) else None
    deps = [SHIM_SRC, os.path.join(HERE, "..", "include", "yc_b200.h"), OUT]
    if not force and os.path.exists(SHIM_OUT) and all(os.path.getmtime(d) <= os.path.getmtime(SHIM_OUT) for d in deps):
        return SHIM_OUT
    cmd = ["/usr/bin/g++", "-std=c++14", "-O2", "-fPIC", "-shared", "-fvisibility=hidden", "-w", "-fpermissive",
           "-I", os.path.join(REF, "src"), "-I", os.path.join(REF, "3rdparty", "stb", "include"),
           SHIM_SRC, "-o", SHIM_OUT, "-L", HERE, "-l:libyc_b200.so", "-Wl,-rpath,$ORIGIN", "-ldl", "-lpthread"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("shim build failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    return SHIM_OUT


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
