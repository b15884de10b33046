"""The C++ veneer (include/mundy_b200/*.hpp): compiles and links against libmundy_b200.so with plain g++ (CPU check),
and its reference-style test program passes on a GPU."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "test_veneer.cpp")
LIBDIR = os.path.join(ROOT, "mundy_b200")
EXE = os.path.join(ROOT, "build", "test_veneer")


def _build():
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cmd = [cxx, "-std=c++17", "-O1", "-Wall", "-Wextra", "-I", os.path.join(ROOT, "include"), SRC, "-L", LIBDIR,
           "-lmundy_b200", f"-Wl,-rpath,{LIBDIR}", "-o", EXE]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "warning" not in r.stderr, r.stderr


def test_veneer_compiles_and_links_without_cuda_headers():
    _build()
    # without a device the program must refuse to run (exit 77), never compute on the host
    r = subprocess.run([EXE], capture_output=True, text=True)
    import torch
    if not torch.cuda.is_available():
        assert r.returncode == 77 and "no CPU path" in r.stdout


@pytest.mark.gpu
def test_veneer_reference_style_checks(cuda):
    _build()
    r = subprocess.run([EXE], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert " 0 failed" in r.stdout
