"""In-tree build of the native libraries (sm_100a only).

    python -m starpu_inference_server_b200.build [--force]

Produces, next to this file:
  libspu_b200.so        CUDA kernels + the C-ABI of include/spu_b200.h
  libspu_host.so        C++ host layer (StarPU-shaped codelet, slot pools, replica registry, mini runtime)
  libspu_loader.so      start-up-only TorchScript weight loader (the one library that links LibTorch)
nvcc cross-compiles without a GPU, so this runs in the CPU-only build container;
the .so files travel to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
CORE = os.path.join(HERE, "core")
LOADER = os.path.join(HERE, "loader")
OBJ = os.path.join(HERE, "_obj")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ARCH + ["-std=c++17", "-O3", "-lineinfo", "-Xcompiler", "-fPIC",
                     "--expt-relaxed-constexpr", "-split-compile=0",
                     "-I", os.path.join(ROOT, "include")]
if os.environ.get("SPU_DEBUG_HOOKS") == "1":  # measurement builds only (include/spu_b200_debug.h)
    NVCC_FLAGS.append("-DSPU_DEBUG_HOOKS=1")

CUDA_SOURCES = ["replica.cu", "gemm_sm100.cu", "attention.cu", "attention_sm100.cu", "rowops.cu", "vit.cu"]
HOST_SOURCES = ["inference_codelet.cpp", "slot_pools.cpp", "replica_registry.cpp", "task_runner.cpp",
                "mini_starpu.cpp", "host_capi.cpp", "host_test_hooks.cpp"]
CXX = os.environ.get("CXX", "g++")
CUDA_HOME = os.environ.get("CUDA_HOME", "/usr/local/cuda")
CXX_FLAGS = ["-std=c++20", "-O2", "-g", "-fPIC", "-Wall", "-Wextra", "-pthread",
             "-I", os.path.join(ROOT, "include"), "-I", os.path.join(CUDA_HOME, "include")]


def _digest(paths):
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(p.encode())
        with open(p, "rb") as f:
            h.update(f.read())
    h.update(" ".join(NVCC_FLAGS + CXX_FLAGS).encode())
    return h.hexdigest()


def _deps(directory):
    return [os.path.join(directory, f) for f in os.listdir(directory)
            if f.endswith((".cu", ".cuh", ".h", ".hpp", ".cpp"))] + \
           [os.path.join(ROOT, "include", f) for f in os.listdir(os.path.join(ROOT, "include"))]


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))
    return r.stdout + r.stderr


def _build_lib(name, directory, sources, extra_link, force):
    out = os.path.join(HERE, name)
    stamp = os.path.join(OBJ, name + ".sha256")
    dig = _digest(_deps(directory) + ([os.path.join(HERE, "libspu_b200.so")] if False else []))
    if not force and os.path.exists(out) and os.path.exists(stamp) and open(stamp).read() == dig:
        return out, False
    os.makedirs(OBJ, exist_ok=True)
    objs = []

    def compile_one(src):
        o = os.path.join(OBJ, name + "." + src + ".o")
        if src.endswith(".cu"):
            _run([NVCC] + NVCC_FLAGS + ["-c", os.path.join(directory, src), "-o", o])
        else:  # plain C++ host code: no device code, C++20 (std::span)
            _run([CXX] + CXX_FLAGS + ["-c", os.path.join(directory, src), "-o", o])
        return o

    with ThreadPoolExecutor(max_workers=min(8, len(sources))) as ex:
        objs = list(ex.map(compile_one, sources))
    _run([NVCC] + ARCH + ["-shared", "-o", out] + objs + ["--cudart", "static"] + extra_link)
    with open(stamp, "w") as f:
        f.write(dig)
    return out, True


def _torch_dir():
    """site-packages/torch without importing it (its headers and libs ship in the wheel)."""
    import importlib.util

    spec = importlib.util.find_spec("torch")
    if spec is None or not spec.submodule_search_locations:
        return None
    d = list(spec.submodule_search_locations)[0]
    return d if os.path.isdir(os.path.join(d, "include", "torch")) else None


def _build_loader(force):
    """libspu_loader.so: plain g++ against the wheel's LibTorch (C++17, cxx11 ABI as torch >= 2.x wheels use)."""
    out = os.path.join(HERE, "libspu_loader.so")
    src = os.path.join(LOADER, "torchscript_loader.cpp")
    tdir = _torch_dir()
    if tdir is None or not os.path.exists(src):
        return None, False
    stamp = os.path.join(OBJ, "libspu_loader.so.sha256")
    dig = _digest([src, os.path.join(ROOT, "include", "spu_loader.h"), os.path.join(ROOT, "include", "spu_b200.h")])
    if not force and os.path.exists(out) and os.path.exists(stamp) and open(stamp).read() == dig:
        return out, False
    os.makedirs(OBJ, exist_ok=True)
    tlib = os.path.join(tdir, "lib")
    _run([CXX, "-std=c++17", "-O2", "-fPIC", "-shared", "-D_GLIBCXX_USE_CXX11_ABI=1",
          "-I", os.path.join(ROOT, "include"), "-I", os.path.join(tdir, "include"),
          "-I", os.path.join(tdir, "include", "torch", "csrc", "api", "include"), src, "-o", out,
          "-L" + HERE, "-lspu_b200", "-L" + tlib, "-ltorch", "-ltorch_cpu", "-lc10",
          "-Wl,-rpath,$ORIGIN", "-Wl,-rpath," + tlib])
    with open(stamp, "w") as f:
        f.write(dig)
    return out, True


def build(force: bool = False, verbose: bool = False):
    """Compile every CUDA/C++ source for sm_100a; returns the library paths."""
    libs = []
    p, rebuilt = _build_lib("libspu_b200.so", CSRC, CUDA_SOURCES, [], force)
    libs.append(p)
    if verbose:
        print(("built " if rebuilt else "up to date ") + p)
    if os.path.isdir(CORE) and all(os.path.exists(os.path.join(CORE, s)) for s in HOST_SOURCES):
        p2, rebuilt2 = _build_lib("libspu_host.so", CORE, HOST_SOURCES,
                                  ["-L" + HERE, "-lspu_b200", "-Xlinker", "-rpath=$ORIGIN",
                                   "-lpthread"], force or rebuilt)
        libs.append(p2)
        if verbose:
            print(("built " if rebuilt2 else "up to date ") + p2)
    p3, rebuilt3 = _build_loader(force or rebuilt)
    if p3 is not None:
        libs.append(p3)
        if verbose:
            print(("built " if rebuilt3 else "up to date ") + p3)
    return libs


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
