"""The C++17 host mirror (include/sgz/TransformPair.hpp) compiles against the C ABI; on a CPU box it
must fail loudly with SGZ_ECUDA (no fallback), on a GPU it must produce the analytic answer."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build(tmp_path):
    exe = str(tmp_path / "cpp_mirror_smoke")
    libdir = os.path.join(ROOT, "signalizer_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "cpp_mirror_smoke.cpp"),
                           "-L" + libdir, "-lsgz_cuda", "-Wl,-rpath," + libdir, "-o", exe])
    return exe


def test_cpp_mirror_compiles_and_refuses_without_gpu(tmp_path):
    import torch
    exe = build(tmp_path)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; covered by the gpu-marked test")
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 77, (r.returncode, r.stdout)
    assert "no CUDA device" in r.stdout


@pytest.mark.gpu
def test_cpp_mirror_runs_on_gpu(tmp_path):
    exe = build(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    assert "ok" in r.stdout
