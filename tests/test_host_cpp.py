"""The C++ host layer above the C ABI (efficient-path-planner_b200/host): World, StateValidator,
MotionValidator, PathPlanner::checkTrajectoryValidity, BatchPlanner — built by host/Makefile and
exercised by host_selftest the way the reference's own code uses these classes
(src/PathPlanner.cpp:27-58, src/StateValidator.cpp, src/MotionValidator.cpp)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "efficient-path-planner_b200", "host")
CONFIG = os.path.join(ROOT, "efficient-path-planner_b200", "config", "config.json")


def _selftest(mode):
    exe = os.path.join(HOST, "host_selftest")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "efficient-path-planner_b200", "csrc")])
        subprocess.check_call(["make", "-s", "-C", HOST])
    return subprocess.run([exe, CONFIG, mode], capture_output=True, text=True, timeout=300)


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
