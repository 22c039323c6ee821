"""CPU + GPU: the header-only C++ shims (reference class names over the C ABI) compile and link
against liborbx.so; on a GPU box the example runs the stereo Frame front-end sequence."""
import pathlib
import shutil
import subprocess

import pytest

ROOT = pathlib.Path(__file__).resolve().parents[1]


def _build(tmp_path, built_lib):
    exe = tmp_path / "example_stereo"
    cmd = [shutil.which("g++") or "g++", "-O1", "-std=c++17", "-Wall", "-Werror",
           str(ROOT / "orb-slam3_b200" / "cpp" / "example_stereo.cc"),
           f"-L{built_lib.parent}", "-lorbx", f"-Wl,-rpath,{built_lib.parent}", "-pthread", "-o", str(exe)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_shim_compiles_and_fails_loudly_without_gpu(tmp_path, built_lib, pkg):
    exe = _build(tmp_path, built_lib)
    if pkg.binding.load().orbx_device_count() > 0:
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 1 and "no CPU path" in r.stderr


@pytest.mark.gpu
def test_shim_example_runs_on_gpu(tmp_path, built_lib):
    exe = _build(tmp_path, built_lib)
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "stereo matches" in r.stdout and "SearchForTriangulation" in r.stdout
    assert "one-call stereo identical: yes" in r.stdout
    print(r.stdout)


def _build_opencv_branch(tmp_path, built_lib):
    """cpp/ORBextractor.h with ORBX_WITH_OPENCV, against the OpenCV-shaped cv:: namespace of oracle/cvshim (the real OpenCV is
    not in this image); -Wall -Werror as the other shim build."""
    exe = tmp_path / "test_extract_orb"
    cmd = [shutil.which("g++") or "g++", "-O1", "-std=c++17", "-Wall", "-Werror", "-DORBX_WITH_OPENCV", f"-I{ROOT / 'oracle' / 'cvshim'}",
           str(ROOT / "orb-slam3_b200" / "cpp" / "test_extract_orb.cc"),
           f"-L{built_lib.parent}", "-lorbx", f"-Wl,-rpath,{built_lib.parent}", "-pthread", "-o", str(exe)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_opencv_branch_compiles(tmp_path, built_lib):
    _build_opencv_branch(tmp_path, built_lib)


@pytest.mark.gpu
@pytest.mark.parametrize("rows,cols,nf,lap,seed", [(480, 752, 1000, (0, 1000), 3), (512, 512, 1000, (128, 383), 4)])
def test_extract_orb_caller_matches_oracle(tmp_path, built_lib, pkg, orc, oracle, rows, cols, nf, lap, seed):
    """Frame::ExtractORB (src/Frame.cc:482-492) over the shim with the reference's operator() signatures: what lands in
    mvKeys / mDescriptors / monoLeft (and the opt-in host mvImagePyramid) must equal the oracle - and the reference's own
    ORBextractor.cc (oracle/_ref) where it is built."""
    import numpy as np
    exe = _build_opencv_branch(tmp_path, built_lib)
    img = pkg.synth.synth_frame(rows, cols, seed)
    step = cols + 24                                       # the caller's Mat is a view with a larger row step
    buf = np.zeros((rows, step), np.uint8); buf[:, :cols] = img
    (tmp_path / "frame.raw").write_bytes(buf.tobytes())
    out = tmp_path / "out.bin"
    r = subprocess.run([str(exe), str(tmp_path / "frame.raw"), str(rows), str(cols), str(step), str(nf), str(lap[0]), str(lap[1]), str(out)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    raw = out.read_bytes()
    mono, n, nlev = np.frombuffer(raw, np.int32, 3)
    kps = np.frombuffer(raw, pkg.KP_DTYPE, n, 12)
    desc = np.frombuffer(raw, np.uint8, n * 32, 12 + 28 * n).reshape(n, 32)
    E = orc.Extractor(nf, oracle=oracle)
    mo, ko, do = E(img, lap)
    assert mono == mo and n == len(ko) and np.array_equal(kps, ko) and np.array_equal(desc, do)
    off = 12 + 60 * n
    assert nlev == 8
    for l in range(nlev):
        h, w = np.frombuffer(raw, np.int32, 2, off); off += 8
        plane = np.frombuffer(raw, np.uint8, h * w, off).reshape(h, w); off += h * w
        assert np.array_equal(plane, E.level_plane(l)[19:-19, 19:-19]), f"mvImagePyramid[{l}]"
    import _pkg
    _pkg.load_oracle()
    import ref_py
    if ref_py.available("seq"):
        mr, kr, dr = ref_py.RefExtractor(nf, flavour="seq")(img, lap)
        assert mono == mr and np.array_equal(kps, kr) and np.array_equal(desc, dr)
