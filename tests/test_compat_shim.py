"""The header-compatible C++ shim (include/voxelrt_compat) must keep compiling the reference's call
sequence (examples/headless_main.cu = VoxelApp/main.cu minus SDL) and link against libvxrt.so."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "examples", "headless")


def build_example():
    cmd = ["nvcc", "-std=c++17", "-arch=sm_100a", "-w", "-I" + os.path.join(ROOT, "include", "voxelrt_compat"),
           os.path.join(ROOT, "examples", "headless_main.cu"), "-L" + os.path.join(ROOT, "voxelengine_b200", "lib"), "-lvxrt",
           "-Xlinker", "-rpath=" + os.path.join(ROOT, "voxelengine_b200", "lib"), "-o", EXE]
    subprocess.check_call(cmd)


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="nvcc not available")
def test_reference_call_sequence_compiles_against_shim(vx):
    build_example()
    assert os.path.exists(EXE)


@pytest.mark.gpu
def test_headless_app_runs(vx, tmp_path):
    if not os.path.exists(EXE):
        build_example()
    out = subprocess.run([EXE, "32", "2"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-1000:]
    assert "frame 1" in out.stdout
    assert os.path.getsize(tmp_path / "frame.bgra") == 1280 * 720 * 4
