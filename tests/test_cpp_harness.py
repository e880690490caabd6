"""tools/sgz_harness.cpp — the C++ offline harness (one host thread, handle and stream per device over the C ABI):
it must build against include/sgz.h, refuse to run without a GPU (no CPU fallback), and on a GPU process a small batch
and report a rate and a finite checksum."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")


def build(tmp_path):
    exe = str(tmp_path / "sgz_harness")
    libdir = os.path.join(ROOT, "signalizer_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(CUDA, "include"),
                           os.path.join(ROOT, "tools", "sgz_harness.cpp"), "-L" + libdir, "-lsgz_cuda", "-L" + os.path.join(CUDA, "lib64"), "-lcudart",
                           "-lpthread", "-Wl,-rpath," + libdir, "-o", exe])
    return exe


def test_harness_builds_and_refuses_without_gpu(tmp_path):
    import torch
    exe = build(tmp_path)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; covered by the gpu-marked test")
    r = subprocess.run([exe, "8", "2", "1"], capture_output=True, text=True)
    assert r.returncode == 77 and "no CUDA device" in r.stdout


@pytest.mark.gpu
def test_harness_runs_on_gpu(tmp_path):
    exe = build(tmp_path)
    r = subprocess.run([exe, "40", "6", "2"], capture_output=True, text=True)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["streams"] == 40 and line["frames_per_s"] > 0 and line["devices"] >= 1
    assert line["checksum"] == line["checksum"] and abs(line["checksum"]) < 1e9   # finite
