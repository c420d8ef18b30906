"""The C++ host-side mirror (rust-hadoop-generated-by-llm_b200/cpp/gpu_ec.hpp) of the reference's Rust interface.
CPU: it compiles and links against libgpu_ec.so. GPU: tests/cpp/test_mirror.cpp replays the
reference's own unit tests through it (and therefore through the C ABI)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "rust-hadoop-generated-by-llm_b200")
SRC = os.path.join(ROOT, "tests", "cpp", "test_mirror.cpp")
EXE = os.path.join(ROOT, "tests", "cpp", "test_mirror")


def _build():
    deps = [SRC, os.path.join(PKG, "cpp", "gpu_ec.hpp"), os.path.join(ROOT, "include", "gpu_ec.h"),
            os.path.join(PKG, "libgpu_ec.so")]
    if not os.path.exists(EXE) or os.path.getmtime(EXE) < max(os.path.getmtime(d) for d in deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-o", EXE, SRC, "-L" + PKG, "-lgpu_ec",
                               "-Wl,-rpath," + PKG, "-lpthread"])
    return EXE


def test_cpp_mirror_compiles_and_links():
    assert os.path.exists(_build())


@pytest.mark.gpu
def test_cpp_mirror_replays_reference_unit_tests(tmp_path):
    exe = _build()
    r = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.startswith("ok ")
