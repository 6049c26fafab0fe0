"""Builds and runs the C++ host API tests (include/mini_opt_b200/qp.hpp and nonlinear.hpp over the C ABI)."""
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def build(tmp_path, name="test_host_api"):
    from mini_opt_b200 import lib
    lib.load()
    exe = tmp_path / name
    cmd = ["g++", "-std=c++17", "-O1", f"-I{ROOT / 'include'}", str(ROOT / "tests" / "cpp" / f"{name}.cc"),
           f"-L{ROOT / 'mini_opt_b200'}", "-lminiopt_b200", f"-Wl,-rpath,{ROOT / 'mini_opt_b200'}",
           "-Wl,-rpath,/usr/local/cuda/lib64", "-o", str(exe)]
    subprocess.run(cmd, check=True, capture_output=True, text=True)
    return exe


def test_cpp_host_api_cpu(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu variant")
    r = subprocess.run([str(build(tmp_path)), "cpu"], capture_output=True, text=True)
    assert r.returncode == 0 and "cpu ok" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_cpp_host_api_gpu(tmp_path):
    r = subprocess.run([str(build(tmp_path)), "gpu"], capture_output=True, text=True)
    assert r.returncode == 0 and "gpu ok" in r.stdout, r.stdout + r.stderr


def test_cpp_nls_host_api_cpu(tmp_path):
    """Residual / LinearizeAndFillQP / EvaluateNonlinearErrors of nonlinear.hpp (no device calls)."""
    r = subprocess.run([str(build(tmp_path, "test_nls_host_api")), "cpu"], capture_output=True, text=True)
    assert r.returncode == 0 and "cpu ok" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_cpp_nls_host_api_gpu(tmp_path):
    """The reference's own nonlinear test problems (test/nonlinear_test.cc) through
    ConstrainedNonlinearLeastSquares::SolveBatch: host Residual callbacks, batched QP solves on the device."""
    r = subprocess.run([str(build(tmp_path, "test_nls_host_api")), "gpu"], capture_output=True, text=True)
    assert r.returncode == 0 and "gpu ok" in r.stdout, r.stdout + r.stderr
