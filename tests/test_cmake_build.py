"""CPU: the top-level CMakeLists.txt (north_star: "CMake: sm_100a build") configures and builds the three libraries
with nvcc cross-compiling for sm_100a, and what it builds exports the symbols the headers declare -- the same check
tests/test_capi_symbols.py runs on the in-tree build."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def cmake_build(tmp_path_factory):
    if shutil.which("cmake") is None or shutil.which("nvcc") is None and not os.path.exists("/usr/local/cuda/bin/nvcc"):
        pytest.skip("cmake / nvcc not available")
    build = tmp_path_factory.mktemp("cmake_build")
    env = dict(os.environ, CUDACXX=os.environ.get("CUDACXX", "/usr/local/cuda/bin/nvcc"))
    gen = ["-G", "Ninja"] if shutil.which("ninja") else []
    cfg = subprocess.run(["cmake", "-S", ROOT, "-B", str(build), *gen], capture_output=True, text=True, env=env)
    assert cfg.returncode == 0, cfg.stdout[-2000:] + cfg.stderr[-2000:]
    bld = subprocess.run(["cmake", "--build", str(build), "--parallel", str(os.cpu_count() or 4)],
                         capture_output=True, text=True, env=env)
    assert bld.returncode == 0, bld.stdout[-3000:] + bld.stderr[-3000:]
    return str(build)


def _declared(header):
    text = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", header)).read(), flags=re.S)
    return set(re.findall(r"\b(spu_[a-z0-9_]+)\s*\(", text)) - {"spu_status"}


def _exported(lib):
    out = subprocess.run(["nm", "-D", "--defined-only", lib], capture_output=True, text=True, check=True).stdout
    return {line.split()[-1] for line in out.splitlines() if line.strip()}


@pytest.mark.parametrize("lib,header", [("libspu_b200.so", "spu_b200.h"), ("libspu_host.so", "spu_host.h"),
                                        ("libspu_loader.so", "spu_loader.h")])
def test_cmake_libraries_export_the_declared_abi(cmake_build, lib, header):
    path = os.path.join(cmake_build, lib)
    assert os.path.exists(path), os.listdir(cmake_build)
    missing = _declared(header) - _exported(path)
    assert not missing, missing


def test_cmake_kernel_library_is_sm100a_only_and_has_no_debug_hooks(cmake_build):
    path = os.path.join(cmake_build, "libspu_b200.so")
    elf = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-lelf", path], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", elf))
    assert archs == {"sm_100a"}, archs
    # the measurement hooks of include/spu_b200_debug.h are not in a product build
    assert not {s for s in _exported(path) if s.startswith("spu_debug_")}
    needed = subprocess.run(["readelf", "-d", path], capture_output=True, text=True).stdout
    assert "libtorch" not in needed and "libcublas" not in needed and "libcudart" not in needed
