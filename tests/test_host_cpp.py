"""The C++ host layer above the C ABI (efficient-path-planner_b200/host): World, StateValidator,
MotionValidator, PathPlanner::checkTrajectoryValidity, BatchPlanner, poly_traj::generateTrajectory, OnlineTrajGenerator,
PathWriter — built by host/Makefile and
exercised by host_selftest the way the reference's own code uses these classes
(src/PathPlanner.cpp:27-58, src/StateValidator.cpp, src/MotionValidator.cpp)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "efficient-path-planner_b200", "host")
CONFIG = os.path.join(ROOT, "efficient-path-planner_b200", "config", "config.json")


def _exe():
    exe = os.path.join(HOST, "host_selftest")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "efficient-path-planner_b200", "csrc")])
        subprocess.check_call(["make", "-s", "-C", HOST])
    return exe


def _selftest(mode):
    import tempfile
    env = dict(os.environ, EPP_PATH_SEGMENTS_DIR=os.path.join(tempfile.mkdtemp(prefix="epp_host_"), "path_segments"))
    return subprocess.run([_exe(), CONFIG, mode], capture_output=True, text=True, timeout=300, env=env)


def test_cpp_path_writer_dumps_equal_the_python_writer(tmp_path):
    """host/src/PathWriter.cpp and online_traj_planner.PathWriter (checked against the reference's own PathWriter.cpp in
    tests/test_path_writer.py) write the same bytes for the call sequence of host_selftest --writer."""
    import online_traj_planner as otp
    r = subprocess.run([_exe(), "--writer", str(tmp_path / "cpp")], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Folder created: " + str(tmp_path / "cpp") in r.stdout
    w = otp.PathWriter(str(tmp_path / "py"))
    w.updateGatePos(0, [0.45, -1.0, 0, 0, 0, 2.35, 1])
    w.updateGatePos(1, [1.0, -1.55, 0, 0, 0, -0.78, 0])
    w.updateObstaclePos(0, [1.0, -0.5, 0, 0, 0, 0])
    pts = [[1.0, 1.0, 0.05], [1e-7, -123456.789, 1e6], [1.0 / 3.0, 2e10, -0.0], [-0.5, 0.29, 1.0000004]]
    w.writeCheckpoints(pts)
    w.writePath(pts)
    w.writePath(pts[:-1])
    w.updateGatePos(1, [0.123456789, -0.09, 0, 0, 0, 1e-3, 1])
    names = sorted(os.listdir(tmp_path / "py"))
    assert names == ["checkpoints.txt", "gates.txt", "obstacles.txt", "path_0.txt", "path_1.txt"]
    assert sorted(os.listdir(tmp_path / "cpp")) == names
    for name in names:
        assert (tmp_path / "cpp" / name).read_bytes() == (tmp_path / "py" / name).read_bytes(), name


def test_host_layer_api_and_errors_without_gpu():
    r = _selftest("--no-gpu")
    assert r.returncode == 0, r.stdout + r.stderr
    assert "host_selftest: ok" in r.stdout


def test_host_layer_fails_loudly_without_device():
    """No CPU fallback: with no usable device the --gpu run must fail, not answer."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    r = _selftest("--gpu")
    assert r.returncode != 0
    assert "CUDA" in r.stdout or "device" in r.stdout


@pytest.mark.gpu
def test_host_layer_kat_through_validators_on_gpu():
    r = _selftest("--gpu")
    assert r.returncode == 0, r.stdout + r.stderr
    assert "host_selftest: ok" in r.stdout
