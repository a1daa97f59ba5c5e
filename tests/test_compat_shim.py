"""The header-compatible C++ shim (include/voxelrt_compat) must keep compiling the reference's call
sequence (examples/headless_main.cu = VoxelApp/main.cu minus SDL) and link against libvxrt.so."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "examples", "headless")
VIEW = os.path.join(ROOT, "examples", "vxview")


def build_example(src="headless_main.cu", exe=EXE):
    cmd = ["nvcc", "-std=c++17", "-arch=sm_100a", "-w", "-I" + os.path.join(ROOT, "include", "voxelrt_compat"),
           "-I" + os.path.join(ROOT, "examples"),
           os.path.join(ROOT, "examples", src), "-L" + os.path.join(ROOT, "voxelengine_b200", "lib"), "-lvxrt",
           "-Xlinker", "-rpath=" + os.path.join(ROOT, "voxelengine_b200", "lib"), "-o", exe]
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


# ---- the scripted frame loop (examples/vxview_main.cu + input_replay.h = VoxelApp/main.cu:70-190 without SDL) ----
SCRIPT = """# keys mouse_x mouse_y click wheel
- 100 100 0 0
W 100 100 1 0
W^ 130 90 1 1
AD 130 90 0 0
SQ 200 150 1 -2
E^ 180 170 1 0
d 180 170 0 0
"""

REPLAY_DRIVER = r"""
#include "input_replay.h"
#include <fstream>
#include <iostream>
int main(int argc, char** argv) {
    std::ifstream in(argv[1]);
    std::string line;
    replay::CameraRig rig;
    rig.forward = {0.25f, -0.5f, 0.8125f}; rig.right = {0.75f, 0.0f, -0.375f}; rig.up = {0.1f, 0.9f, 0.3f};
    replay::FrameClock clock;
    while (std::getline(in, line)) {
        if (line.empty() || line[0] == '#') continue;
        replay::InputFrame f;
        if (!replay::parse_line(line.c_str(), f)) return 3;
        replay::apply_input(rig, f);
        clock.add(4.0);
        printf("%a %a %a %a %a %a %a %a\n", rig.pos.x, rig.pos.y, rig.pos.z, rig.eular.x, rig.eular.y, rig.ortho.x, rig.ortho.y, clock.avg_ms);
    }
    return 0;
}
"""


def replay_restatement(lines):
    """main.cu:93-158 in numpy binary32: wheel, speed, W/S, A/D, Q/E with last frame's basis, mouse look."""
    import numpy as np
    f32 = np.float32
    pos = np.array([256, 256, 256], f32)
    fwd, right, up = np.array([0.25, -0.5, 0.8125], f32), np.array([0.75, 0.0, -0.375], f32), np.array([0.1, 0.9, 0.3], f32)
    eular, ortho, last, avg = np.zeros(2, f32), np.array([10, 10], f32), [0, 0], 0.0
    out = []
    for line in lines:
        keys, mx, my, click, wheel = line.split()
        mx, my, click, wheel = int(mx), int(my), int(click), int(wheel)
        keys = keys.upper()
        for _ in range(abs(wheel)):  # one SDL_MOUSEWHEEL event per notch; scrolling up shrinks the window
            ortho = (ortho + f32(-10.0 if wheel > 0 else 10.0)).astype(f32)
        speed = f32(0.2)
        if "^" in keys:
            speed = f32(speed * f32(10))
        for k, vec, sign in (("W", fwd, 1), ("S", fwd, -1), ("A", right, -1), ("D", right, 1), ("Q", up, -1), ("E", up, 1)):
            if k in keys:
                step = (vec * speed).astype(f32)
                pos = (pos + step).astype(f32) if sign > 0 else (pos - step).astype(f32)
        if click:
            eular = np.array([eular[0] + f32(my - last[1]) * f32(0.004), eular[1] + f32(mx - last[0]) * f32(0.004)], f32)
        last = [mx, my]
        avg = avg * (1 - (1 / 100.0)) + 4.0 * (1 / 100.0)
        out.append([float(v) for v in pos] + [float(eular[0]), float(eular[1]), float(ortho[0]), float(ortho[1]), avg])
    return out


def test_input_replay_follows_the_reference_loop(tmp_path):
    """The camera a script produces equals a numpy restatement of main.cu's update order, bit for bit."""
    cuda_inc = "/usr/local/cuda/include"
    if shutil.which("g++") is None or not os.path.exists(os.path.join(cuda_inc, "vector_types.h")):
        pytest.skip("g++ or CUDA headers not available")
    (tmp_path / "script.txt").write_text(SCRIPT)
    (tmp_path / "driver.cpp").write_text(REPLAY_DRIVER)
    exe = str(tmp_path / "driver")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-I" + cuda_inc, "-I" + os.path.join(ROOT, "examples"),
                           str(tmp_path / "driver.cpp"), "-o", exe])
    got = [[float.fromhex(v) for v in l.split()] for l in subprocess.check_output([exe, str(tmp_path / "script.txt")], text=True).splitlines()]
    want = replay_restatement([l for l in SCRIPT.splitlines() if l and not l.startswith("#")])
    assert len(got) == len(want) == 7
    assert got == want
    assert got[-1][5:7] == [20.0, 20.0]          # one notch up, two down
    assert got[0][:3] == [256.0, 256.0, 256.0]   # no key held: the camera stays


def test_write_ppm_swaps_bgra_to_rgb(tmp_path):
    cuda_inc = "/usr/local/cuda/include"
    if shutil.which("g++") is None or not os.path.exists(os.path.join(cuda_inc, "vector_types.h")):
        pytest.skip("g++ or CUDA headers not available")
    (tmp_path / "ppm.cpp").write_text(r"""
#include "input_replay.h"
int main(int, char** argv) { const unsigned char px[] = {1, 2, 3, 255, 10, 20, 30, 255}; return replay::write_ppm(argv[1], px, 2, 1) ? 0 : 1; }
""")
    exe = str(tmp_path / "ppm")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-I" + cuda_inc, "-I" + os.path.join(ROOT, "examples"), str(tmp_path / "ppm.cpp"), "-o", exe])
    subprocess.check_call([exe, str(tmp_path / "a.ppm")])
    assert (tmp_path / "a.ppm").read_bytes() == b"P6\n2 1\n255\n" + bytes([3, 2, 1, 30, 20, 10])


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="nvcc not available")
def test_scripted_viewer_compiles_against_shim(vx):
    build_example("vxview_main.cu", VIEW)
    assert os.path.exists(VIEW)


@pytest.mark.gpu
def test_scripted_viewer_runs_and_fills_the_checkerboard(vx, tmp_path):
    """Two frames with the camera at rest fill the whole image (checkerboard halves of Renderer.cu:186-194 alternate with
    the frame number into a persistent buffer that starts at 255); the dumps are PPMs of the frame size."""
    if not os.path.exists(VIEW):
        build_example("vxview_main.cu", VIEW)
    (tmp_path / "script.txt").write_text("- 0 0 0 0\n- 0 0 0 0\n- 0 0 0 0\nW^ 0 0 0 0\n")
    out = subprocess.run([VIEW, "--size", "256", "128", "256", "--factor", "8", "--script", str(tmp_path / "script.txt"), "--dump", "1",
                          "--width", "320", "--height", "192", "--out", str(tmp_path / "f")], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-1000:]
    assert "frame 3:" in out.stdout and "Avg FPS" in out.stdout
    frames = []
    for k in range(4):
        data = (tmp_path / ("f_%04d.ppm" % k)).read_bytes()
        head = b"P6\n320 192\n255\n"
        assert data.startswith(head) and len(data) == len(head) + 320 * 192 * 3
        frames.append(data[len(head):])
    assert frames[0] != frames[1]      # the second frame writes the other half of the pixels
    assert frames[1] == frames[2]      # camera at rest: the third frame rewrites the first half with the same values
    assert frames[3] != frames[2]      # the camera moved


# ---- the frame loop on N GPUs in one process, plain C++ over the C ABI (examples/multi_gpu_main.cu) ----
MULTI = os.path.join(ROOT, "examples", "multi_gpu")


def build_multi():
    subprocess.check_call(["nvcc", "-std=c++17", "-arch=sm_100a", "-w", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "examples", "multi_gpu_main.cu"), "-L" + os.path.join(ROOT, "voxelengine_b200", "lib"), "-lvxrt",
                           "-Xlinker", "-rpath=" + os.path.join(ROOT, "voxelengine_b200", "lib"), "-o", MULTI])


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="nvcc not available")
def test_multi_gpu_driver_compiles(vx):
    build_multi()
    assert os.path.exists(MULTI)


@pytest.mark.gpu
@pytest.mark.parametrize("devices", ["0", "0,0", "0,0,0", "0,1", "0,1,2,3"])
def test_multi_gpu_driver_assembles_the_single_gpu_frame(vx, tmp_path, devices):
    """One process, N devices (or N band sets on one device), every device's kernels storing its bands into the first
    device's frame: the assembled frame must be the whole-frame render of one device, byte for byte (the program compares
    them itself and returns 3 otherwise)."""
    if max(int(d) for d in devices.split(",")) >= vx.device_count():
        pytest.skip("needs %d GPUs" % (max(int(d) for d in devices.split(",")) + 1))
    if not os.path.exists(MULTI):
        build_multi()
    out = subprocess.run([MULTI, "--devices", devices, "--world", "256", "128", "256", "--frame", "640", "360", "--frames", "3",
                          "--out", str(tmp_path / "frame.bgra")], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-1000:] + out.stderr[-1000:]
    assert "frame_identical=1" in out.stdout
    assert os.path.getsize(tmp_path / "frame.bgra") == 640 * 360 * 4
