"""CPU tests of the drop-in boundary: the C-ABI library loads without a GPU, exports every symbol the header declares,
the ctypes table covers the header, and compute calls fail loudly (no CPU fallback) when there is no device."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "mundy_b200.h")


def _declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mb200_[A-Za-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from mundy_b200 import _capi
    syms = _declared_symbols()
    assert len(syms) >= 35
    out = subprocess.run(["nm", "-D", "--defined-only", _capi.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r" T (mb200_[A-Za-z0-9_]+)", out))
    assert set(syms) <= exported, sorted(set(syms) - exported)
    assert set(syms) == set(_capi.SIGNATURES), sorted(set(syms) ^ set(_capi.SIGNATURES))
    # nothing but the C ABI is exported
    assert all(s.startswith("mb200_") for s in re.findall(r" T (\S+)", out))


def test_abi_version_and_struct_sizes():
    from mundy_b200 import _capi
    assert _capi.lib.mb200_abi_version() == 1
    assert C.sizeof(_capi.Bodies) == 48 and C.sizeof(_capi.ContactOp) == 88
    assert C.sizeof(_capi.CollisionResult) == 24 and C.sizeof(_capi.StepStats) == 40


def test_product_does_not_touch_the_oracle():
    """The product path (package + C sources + header) must not reference anything under oracle/."""
    for base in ("mundy_b200", "include"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                    text = open(os.path.join(dp, f)).read()
                    assert "liboracle" not in text and "orc_" not in text and "import oracle" not in text, f


def test_compute_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from mundy_b200 import _capi
    assert _capi.lib.mb200_device_count() == 0
    a = np.zeros(6)
    st = _capi.lib.mb200_compute_aabb_sphere(None, 1, a.ctypes.data, a.ctypes.data, a.ctypes.data)
    assert st == -2 and b"no CPU fallback" in _capi.lib.mb200_last_error_string()
    # argument validation happens before the device check
    assert _capi.lib.mb200_compute_aabb_sphere(None, -1, None, None, None) == -1
    h = C.c_void_p()
    assert _capi.lib.mb200_neighbor_links_create(C.byref(h)) == 0
    assert _capi.lib.mb200_neighbor_links_set_search_buffer(h, -1.0) == -1
    assert _capi.lib.mb200_neighbor_links_set_search_buffer(h, 0.5) == 0
    assert _capi.lib.mb200_neighbor_links_count(h) == 0
    assert _capi.lib.mb200_neighbor_links_destroy(h) == 0


def test_header_constants_match_the_python_side():
    """Limits that the header states and the harness repeats."""
    from mundy_b200 import dist as mbd
    from mundy_b200 import ops
    text = open(HEADER).read()
    assert int(re.search(r"#define MB200_NO_ALLOC_MAX_BODIES (\d+)", text).group(1)) == ops.NO_ALLOC_MAX_BODIES
    src = open(os.path.join(ROOT, "mundy_b200", "csrc", "repartition.cu")).read()
    assert int(re.search(r"kHilbertBits = (\d+)", src).group(1)) == mbd.HILBERT_BITS
    assert 3 * mbd.HILBERT_BITS - int(re.search(r"kBucketBits = (\d+)", src).group(1)) == mbd.BUCKET_SHIFT


def test_documented_environment_switches_exist_in_the_sources():
    """INTEGRATION.md's table of switches against the getenv calls of the library and of bench.py."""
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    table = doc[doc.index("## Environment switches"):doc.index("## Stage timers")]
    documented = set(re.findall(r"`(MB200_[A-Z_0-9]+)", table))
    used = set()
    for dp, _, files in os.walk(os.path.join(ROOT, "mundy_b200", "csrc")):
        for f in files:
            used |= set(re.findall(r'getenv\("(MB200_[A-Z_0-9]+)"\)', open(os.path.join(dp, f)).read()))
    used |= {v for v in re.findall(r'environ\.get\("(MB200_[A-Z_0-9]+)"\)', open(os.path.join(ROOT, "bench.py")).read())
             if v != "MB200_BENCH_VERBOSE"}
    assert documented == used, sorted(documented ^ used)
