"""The GEMM engine's work order (decode_work, dplasma_b200/csrc/gemm_dmma.cuh) covers every CTA tile of a launch exactly
once, for every grouping, grid position, ragged edge and mask -- checked on the host: the decode function is
__host__ __device__ and tests/c/work_order_check.cu walks all block indices of ~6000 descriptors."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_work_order_covers_every_block_once(tmp_path):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not found")
    exe = str(tmp_path / "work_order_check")
    subprocess.run([nvcc, "-std=c++17", "-O1", "-gencode", "arch=compute_100a,code=sm_100a", "-o", exe,
                    os.path.join(ROOT, "tests", "c", "work_order_check.cu")], check=True, timeout=600)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert " 0 failures" in r.stdout
