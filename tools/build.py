#!/usr/bin/env python
"""Build b200dp/lib/libb200dp.so from csrc/*.cu for sm_100a (nvcc cross-compiles without a GPU).

    python tools/build.py [--force] [--verbose]

Objects go to build/, ptxas resource usage to build/ptxas.log.  The .so is git-ignored but ships
to the GPU box with the gpurun snapshot.
"""
import argparse
import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "csrc")
BUILD = os.path.join(ROOT, "build")
OUT = os.path.join(ROOT, "b200dp", "lib", "libb200dp.so")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xptxas", "-v", "--expt-relaxed-constexpr",
         "-diag-suppress", "177,550"]


def nvcc() -> str:
    for c in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def _digest(paths):
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(p.encode())
        h.update(open(p, "rb").read())
    h.update(" ".join(ARCH + FLAGS).encode())
    return h.hexdigest()


def sources():
    return sorted(os.path.join(SRC, f) for f in os.listdir(SRC) if f.endswith((".cu", ".cpp")))


def headers():
    hs = [os.path.join(SRC, f) for f in os.listdir(SRC) if f.endswith((".cuh", ".h"))]
    hs += [os.path.join(ROOT, "include", f) for f in os.listdir(os.path.join(ROOT, "include"))]
    return sorted(hs)


def compile_one(src, verbose):
    obj = os.path.join(BUILD, os.path.basename(src) + ".o")
    stamp = obj + ".sha"
    dig = _digest([src] + headers())
    if os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == dig:
        return obj, "", True
    cmd = [nvcc(), *ARCH, *FLAGS, "-I", os.path.join(ROOT, "include"), "-c", src, "-o", obj]
    if src.endswith(".cpp"):
        cmd = [nvcc(), "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-I", os.path.join(ROOT, "include"), "-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    open(stamp, "w").write(dig)
    return obj, r.stderr, False


def build(force=False, verbose=False) -> str:
    os.makedirs(BUILD, exist_ok=True)
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    if force:
        for f in os.listdir(BUILD):
            if f.endswith((".o", ".sha")):
                os.remove(os.path.join(BUILD, f))
    srcs = sources()
    logs = []
    objs = []
    all_cached = True
    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        for obj, log, cached in ex.map(lambda s: compile_one(s, verbose), srcs):
            objs.append(obj)
            all_cached &= cached
            if log:
                logs.append(log)
    if logs:
        with open(os.path.join(BUILD, "ptxas.log"), "a" if not force else "w") as f:
            f.write("\n".join(logs))
    if all_cached and os.path.exists(OUT):
        return OUT
    tmp = OUT + ".tmp"
    cmd = [nvcc(), *ARCH, "-shared", "-o", tmp, *objs, "-lcudart_static", "-ldl", "-lpthread", "-lrt"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    os.replace(tmp, OUT)
    if verbose:
        print("\n".join(logs))
    return OUT


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(a.force, a.verbose))
