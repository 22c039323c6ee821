"""ctypes binding of oracle/_ref/libref*.so — the UNMODIFIED reference ORBextractor.cc compiled against
oracle/cvshim (oracle/build_ref.sh).  TEST INFRASTRUCTURE ONLY: tests/, smoke() and bench.py's CPU legs.

Flavours (see oracle/ref_wrap.cc):
  "literal"  libref.so         system allocator: the reference's own pointer tie-break (src/ORBextractor.cc:927)
  "seq"      libref_seq.so     creation-ordered arena: the canonical tie-break, still the reference's code
  "native"   libref_native.so  -O3 + AVX2/FMA like the reference's CMakeLists.txt:10-13 (timed baseline)
"""
from __future__ import annotations

import ctypes as C
import pathlib
import subprocess

import numpy as np

_DIR = pathlib.Path(__file__).resolve().parent
_LIBS = {"literal": "libref.so", "seq": "libref_seq.so", "native": "libref_native.so"}
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])


def available(flavour: str = "literal") -> bool:
    return (_DIR / "_ref" / _LIBS[flavour]).exists()


def build() -> None:
    """Compile from /root/reference when it is present (build container); a no-op on the GPU box."""
    subprocess.run(["sh", str(_DIR / "build_ref.sh")], check=True, capture_output=True)


_loaded: dict = {}


def _lib(flavour: str):
    if flavour in _loaded:
        return _loaded[flavour]
    path = _DIR / "_ref" / _LIBS[flavour]
    if not path.exists():
        build()
    if not path.exists():
        raise FileNotFoundError(f"{path}: oracle/_ref not built (needs /root/reference; run oracle/build_ref.sh)")
    L = C.CDLL(str(path))  # RTLD_LOCAL + -Bsymbolic: the arena operator new of libref_seq.so stays private to it
    L.ref_create.restype = C.c_void_p
    L.ref_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
    L.ref_destroy.argtypes = [C.c_void_p]
    L.ref_extract.restype = C.c_int
    L.ref_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int,
                              C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
    L.ref_levels.argtypes = [C.c_void_p]
    L.ref_scale_factor.restype = C.c_float
    L.ref_scale_factor.argtypes = [C.c_void_p]
    L.ref_tables.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    L.ref_level_dims.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.ref_level_copy.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t]
    _loaded[flavour] = L
    return L


class RefExtractor:
    """ORB_SLAM3::ORBextractor of the reference (include/ORBextractor.h:49-83), called through its own operator()."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, flavour="seq"):
        self.L = _lib(flavour)
        self.flavour = flavour
        self.nlevels = nlevels
        self.cap = max(4 * nfeatures, 4096)
        self.h = self.L.ref_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_destroy(self.h)
            self.h = None

    def tables(self, which: int) -> np.ndarray:
        out = np.zeros(self.nlevels, np.float32)
        self.L.ref_tables(self.h, which, out.ctypes.data_as(C.c_void_p))
        return out

    def __call__(self, image: np.ndarray, lapping=(0, 0)):
        """returns (monoIndex, keypoints[KP_DTYPE], descriptors[n,32])"""
        if image.size == 0:
            rows = cols = 0
            step = 0
            ptr = None
        else:
            assert image.dtype == np.uint8 and image.ndim == 2 and image.strides[1] == 1
            rows, cols = image.shape
            step = image.strides[0]
            ptr = image.ctypes.data_as(C.c_void_p)
        while True:
            kps = np.zeros(self.cap, KP_DTYPE)
            desc = np.zeros((self.cap, 32), np.uint8)
            n = C.c_int(0)
            mono = self.L.ref_extract(self.h, ptr, rows, cols, step, int(lapping[0]), int(lapping[1]),
                                      kps.ctypes.data_as(C.c_void_p), desc.ctypes.data_as(C.c_void_p), self.cap, C.byref(n))
            if mono == -2:
                self.cap = max(2 * self.cap, n.value)
                continue
            return mono, kps[:n.value].copy(), desc[:n.value].copy()

    def extract_raw(self, image: np.ndarray, lapping, kps: np.ndarray, desc: np.ndarray) -> int:
        """timing path: caller-owned output arrays, no copies; returns the keypoint count"""
        n = C.c_int(0)
        self.L.ref_extract(self.h, image.ctypes.data_as(C.c_void_p), image.shape[0], image.shape[1], image.strides[0],
                           int(lapping[0]), int(lapping[1]), kps.ctypes.data_as(C.c_void_p),
                           desc.ctypes.data_as(C.c_void_p), len(kps), C.byref(n))
        return n.value

    def level_plane(self, level: int, padded: bool = False) -> np.ndarray:
        r, c = C.c_int(0), C.c_int(0)
        self.L.ref_level_dims(self.h, level, C.byref(r), C.byref(c))
        b = 19 if padded else 0
        out = np.zeros((r.value + 2 * b, c.value + 2 * b), np.uint8)
        self.L.ref_level_copy(self.h, level, int(padded), out.ctypes.data_as(C.c_void_p), out.strides[0])
        return out
