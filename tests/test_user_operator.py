"""User-written element callbacks (include/b2f/user_operator.cuh): the reference's `assembleBilinear(Function)` template
(femutils/BSRFormat.h:786-787,1105-1111) for callbacks the library does not ship.  The driver tests/cpp/user_operator_driver.cu
is compiled with nvcc for sm_100a against the installed headers + libb200fem.so, like a module of the reference would be."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NVCC = "/usr/local/cuda/bin/nvcc"


def build_driver(tmp_path):
    exe = str(tmp_path / "user_operator_driver")
    libdir = os.path.join(ROOT, "arcanefem_b200", "lib")
    cmd = [NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O2", "-lineinfo", f"-I{ROOT}/include",
           os.path.join(ROOT, "tests", "cpp", "user_operator_driver.cu"), "-o", exe, f"-L{libdir}", "-lb200fem", "-Xlinker", f"-rpath,{libdir}"]
    subprocess.check_call(cmd)
    return exe


def test_user_operator_driver_compiles_and_refuses_to_run_without_a_device(tmp_path):
    exe = build_driver(tmp_path)
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    if has_gpu:
        assert r.returncode == 0, r.stdout + r.stderr
    else:
        assert r.returncode == 3, r.stdout + r.stderr
        assert "no CUDA device" in r.stdout
    assert "element-kernel vocabulary (host): bad=0" in r.stdout  # FixedMatrix / Real4 / computeGradient* checked on the host


@pytest.mark.gpu
def test_user_operator_matches_host_loop_on_device(tmp_path):
    exe = build_driver(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "user_operator_driver:" in r.stdout and r.stdout.rstrip().endswith("bad=0"), r.stdout
