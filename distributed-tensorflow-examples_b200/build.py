"""Build the native libraries in-tree.

    python -m psb200.build            # everything
    python -m psb200.build kernels    # libpsb200_kernels.so  (nvcc, sm_100a)
    python -m psb200.build ps         # libpsb200_ps.so       (g++, parameter-server runtime)

nvcc cross-compiles for sm_100a without a GPU, so this runs on the CPU box; the built
``lib/*.so`` are git-ignored and travel to the GPU box with the working tree.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from typing import List

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "lib")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

KERNEL_SOURCES = ["lib.cu", "gemm.cu", "lstm.cu", "losses.cu", "push.cu", "peer.cu", "lstm_cell.cu", "ps_service.cu"]
PS_SOURCES = ["ps_server.cpp", "tf_bundle.cpp"]

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "--use_fast_math" if False else "-DNDEBUG"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _digest(paths: List[str], extra: str) -> str:
    h = hashlib.sha256(extra.encode())
    for p in sorted(paths):
        with open(p, "rb") as f:
            h.update(p.encode())
            h.update(f.read())
    return h.hexdigest()


def _deps(sources: List[str]) -> List[str]:
    deps = [os.path.join(CSRC, s) for s in sources]
    deps += [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    deps += [os.path.join(INCLUDE, f) for f in os.listdir(INCLUDE) if f.endswith(".h")]
    return deps


def _uptodate(out: str, digest: str) -> bool:
    stamp = out + ".stamp"
    return os.path.exists(out) and os.path.exists(stamp) and open(stamp).read().strip() == digest


def build_kernels(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(LIB, exist_ok=True)
    out = os.path.join(LIB, "libpsb200_kernels.so")
    sources = [s for s in KERNEL_SOURCES if os.path.exists(os.path.join(CSRC, s))]
    digest = _digest(_deps(sources), " ".join(NVCC_FLAGS))
    if not force and _uptodate(out, digest):
        return out
    objs = []
    procs = []
    for s in sources:
        obj = os.path.join(LIB, s.replace(".cu", ".o"))
        cmd = [_nvcc()] + NVCC_FLAGS + ["-c", os.path.join(CSRC, s), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
        objs.append(obj)
    for s, p in procs:
        log = p.communicate()[0].decode()
        if verbose and log:
            print(log)
        if p.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s" % (s, log))
    cmd = [_nvcc(), "-shared", "-o", out] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
    subprocess.run(cmd, check=True)
    with open(out + ".stamp", "w") as f:
        f.write(digest)
    return out


def build_ps(force: bool = False) -> str:
    os.makedirs(LIB, exist_ok=True)
    out = os.path.join(LIB, "libpsb200_ps.so")
    sources = [s for s in PS_SOURCES if os.path.exists(os.path.join(CSRC, s))]
    if not sources:
        return out
    flags = ["-O2", "-std=c++17", "-fPIC", "-shared", "-pthread", "-Wall"]
    digest = _digest(_deps(sources), " ".join(flags))
    if not force and _uptodate(out, digest):
        return out
    cmd = ["g++"] + flags + [os.path.join(CSRC, s) for s in sources] + ["-o", out]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if r.returncode != 0:
        raise RuntimeError("g++ failed:\n%s" % r.stdout.decode())
    with open(out + ".stamp", "w") as f:
        f.write(digest)
    return out


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_ps(force)
    build_kernels(force, verbose)


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    force = "--force" in sys.argv
    verbose = "-v" in sys.argv
    if what == "kernels":
        print(build_kernels(force, verbose))
    elif what == "ps":
        print(build_ps(force))
    else:
        build_all(force, verbose)
        print("built", LIB)
