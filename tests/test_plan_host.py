"""The coarse pass's work-item plan (TcPlan, tiler_b200/csrc/search_tc.h) checked on the host: the header is plain
__host__ __device__ code, so a small C++ program compiled with nvcc exercises decode / list indexing for the single,
split, tail and chained plans without a GPU."""
import os
import shutil
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def test_tcplan_items_and_lists(tmp_path):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    exe = str(tmp_path / "plan_check")
    build = subprocess.run([nvcc, "-std=c++17", "-O1", "-o", exe, os.path.join(HERE, "host", "plan_check.cu"), "-lcudart"],
                           capture_output=True, text=True, timeout=600)
    assert build.returncode == 0, build.stderr[-3000:]
    run = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert run.returncode == 0 and run.stdout.strip().endswith("ok"), run.stdout[-3000:] + run.stderr[-1000:]
