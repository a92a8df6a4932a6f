"""A torch-free C++ host (examples/cabi_mapping.cpp) built against include/isogs_rast.h + libisogs_rast.so:
the boundary really is a C ABI (plain pointers and sizes), usable from the reference's C++ tree
(cpp/src/rendering/renderer.cpp, cpp/src/slam/mapper.cpp) without torch."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "examples", "cabi_mapping.cpp")
EXE = os.path.join(ROOT, "examples", "cabi_mapping")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")


def _build(pkg):
    pkg.lib()   # makes sure libisogs_rast.so exists
    cmd = ["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(CUDA, "include"), SRC,
           "-L", os.path.join(ROOT, "isogs-slam-cpp_b200"), "-lisogs_rast", "-L", os.path.join(CUDA, "lib64"), "-lcudart",
           "-Wl,-rpath," + os.path.join(ROOT, "isogs-slam-cpp_b200"), "-Wl,-rpath," + os.path.join(CUDA, "lib64"),
           "-o", EXE]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return EXE


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++")
def test_cpp_host_compiles_and_links_without_torch(pkg):
    exe = _build(pkg)
    deps = subprocess.run(["ldd", exe], capture_output=True, text=True).stdout
    assert "libisogs_rast.so" in deps and "libcudart" in deps
    assert "torch" not in deps and "c10" not in deps


@pytest.mark.gpu
@pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++")
def test_cpp_host_runs_mapping_iterations(pkg):
    exe = _build(pkg)
    r = subprocess.run([exe, "25"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "no torch involved" in r.stdout
