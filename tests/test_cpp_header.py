"""The drop-in C++ header (include/delaunay_2_interp.h): a port of the reference's demo driver with assertions."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "test_delaunay_2.cpp")
SRC_D = os.path.join(ROOT, "tests", "cpp", "test_delaunay_d.cpp")


def _build(tmp_path, SRC=SRC):
    import delaunay_linterp_b200 as dl
    exe = str(tmp_path / os.path.basename(SRC)[:-4])
    libdir = os.path.dirname(dl.LIB_PATH)
    env = dict(os.environ)
    env.pop("CXX", None)
    subprocess.check_call(["g++", "-std=c++11", "-O1", "-I", os.path.join(ROOT, "include"), SRC, "-o", exe,
                           "-L", libdir, "-ldli", "-Wl,-rpath," + libdir], env=env)
    return exe


def test_header_compiles_and_host_side_runs(tmp_path):
    exe = _build(tmp_path)
    r = subprocess.run([exe, "--host"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_reference_demo_driver_with_assertions(gpu, tmp_path):
    exe = _build(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "regular grid: sum of squared errors: 5.800002" in r.stdout      # README.md:130
    assert "adaptive grid: sum of squared errors: 0.044019" in r.stdout     # README.md:131


def test_dd_header_compiles_and_host_side_runs(tmp_path):
    """include/delaunay_d_interp.h: port of src/delaunay_d_test2.cpp:33-115 for N = 2 and N = 3 (host part: native dD
    triangulation, adaptive queue, per-point interp answered on the host, timed against the reference's published rate)"""
    exe = _build(tmp_path, SRC_D)
    r = subprocess.run([exe, "--host"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_dd_reference_demo_driver_with_assertions(gpu, tmp_path):
    exe = _build(tmp_path, SRC_D)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout + r.stderr
