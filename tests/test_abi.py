"""CPU suite: the C-ABI library loads, exports every symbol include/orbx.h declares, keeps the
reference's error conventions and FAILS LOUDLY without a CUDA device (no CPU fallback)."""
import ctypes as C
import pathlib
import re

import numpy as np
import pytest

ROOT = pathlib.Path(__file__).resolve().parents[1]


def header_symbols():
    txt = (ROOT / "include" / "orbx.h").read_text()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(orbx_[a-z0-9_]+)\s*\(", txt)))


def test_header_declares_what_binding_binds(pkg):
    decl = header_symbols()
    assert len(decl) >= 30
    assert sorted(pkg.binding.SYMBOLS) == decl


def test_library_exports_every_declared_symbol(built_lib):
    lib = C.CDLL(str(built_lib))
    for name in header_symbols():
        assert hasattr(lib, name), name


def test_sass_is_sm100a_only(built_lib):
    import shutil, subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    out = subprocess.run([cuobjdump, "-lelf", str(built_lib)], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_keypoint_layout_matches_cv_keypoint(pkg):
    assert pkg.KP_DTYPE.itemsize == 28
    assert [pkg.KP_DTYPE.fields[n][1] for n in ("x", "y", "size", "angle", "response", "octave", "class_id")] == [0, 4, 8, 12, 16, 20, 24]


def test_no_cpu_fallback(pkg):
    lib = pkg.binding.load()
    if lib.orbx_device_count() > 0:
        pytest.skip("a CUDA device is visible; the loud-failure path is for CPU-only hosts")
    with pytest.raises(pkg.OrbxError) as ei:
        pkg.ORBextractor(1000)
    assert ei.value.code == pkg.binding.ORBX_ERR_CUDA and "no CPU path" in str(ei.value)
    q = np.zeros((4, 32), np.uint8)
    with pytest.raises(pkg.OrbxError) as ei:
        pkg.knn2_hamming(q, q)
    assert ei.value.code == pkg.binding.ORBX_ERR_CUDA


def test_descriptor_distance_host(pkg, oracle):
    """ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:2911-2931) is a host inline in the ABI."""
    rng = np.random.default_rng(3)
    for _ in range(300):
        a = rng.integers(0, 256, 32, dtype=np.uint8); b = rng.integers(0, 256, 32, dtype=np.uint8)
        assert pkg.DescriptorDistance(a, b) == oracle.hamming(a, b) == int(np.unpackbits(a ^ b).sum())
    z = np.zeros(32, np.uint8); f = np.full(32, 255, np.uint8)
    assert pkg.DescriptorDistance(z, z) == 0 and pkg.DescriptorDistance(z, f) == 256


def test_argument_errors_do_not_need_a_gpu(pkg):
    lib = pkg.binding.load()
    h = C.c_void_p()
    assert lib.orbx_create(1000, 1.2, 99, 20, 7, 0, C.byref(h)) == pkg.binding.ORBX_ERR_ARG
    assert lib.orbx_create(1000, 1.0, 8, 20, 7, 0, C.byref(h)) == pkg.binding.ORBX_ERR_ARG
    assert b"bad parameters" in lib.orbx_last_error()
    assert lib.orbx_knn2_hamming(0, None, 5, None, 5, None, None) == pkg.binding.ORBX_ERR_ARG
