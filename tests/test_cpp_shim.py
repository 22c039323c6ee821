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
