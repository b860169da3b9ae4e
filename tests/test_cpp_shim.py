"""The header-only C++ shim (include/arcanefem_b200/*.h: the reference's class/method names over the C ABI) compiles with
plain g++ and drives the library like the reference's Poisson module does (tests/cpp/poisson_driver.cc)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CUDA = "/usr/local/cuda"


def build_driver(tmp_path, name="poisson_driver"):
    exe = str(tmp_path / name)
    libdir = os.path.join(ROOT, "arcanefem_b200", "lib")
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Werror", f"-I{ROOT}/include", f"-I{CUDA}/include", os.path.join(ROOT, "tests", "cpp", name + ".cc"), "-o", exe,
           f"-L{libdir}", "-lb200fem", f"-L{CUDA}/lib64", "-lcudart", f"-Wl,-rpath,{libdir}", f"-Wl,-rpath,{CUDA}/lib64"]
    subprocess.check_call(cmd)
    return exe


def test_shim_compiles_and_refuses_to_run_without_a_device(tmp_path):
    exe = build_driver(tmp_path)
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    if has_gpu:
        assert r.returncode == 0, r.stdout + r.stderr
    else:
        assert r.returncode == 3, r.stdout + r.stderr  # initialize() threw: no CPU fallback
        assert "no CUDA device" in r.stdout


@pytest.mark.gpu
def test_poisson_driver_on_device(tmp_path):
    exe = build_driver(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "bad=0" in r.stdout


NVCC = os.path.join(CUDA, "bin", "nvcc")
BODY_BLOCKS = None


def build_unchanged_module(tmp_path, name="poisson_module_unchanged"):
    """tests/cpp/poisson_module_unchanged.cu: the reference's Poisson module text (tests/cpp/poisson_module_body.inc) against the
    Arcane-typed drop-in headers include/arcanefem_b200/arcane/*.h and the model of the Arcane types (tests/cpp/arcane_stub),
    compiled as CUDA with warnings as errors - the reference compiles its modules as CUDA too (femutils/CMakeLists.txt:3-7)."""
    exe = str(tmp_path / name)
    libdir = os.path.join(ROOT, "arcanefem_b200", "lib")
    cmd = [NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O1", "--extended-lambda", "-Xcompiler", "-Wall,-Werror", f"-I{ROOT}/include",
           f"-I{ROOT}/tests/cpp/arcane_stub", f"-I{ROOT}/tests/cpp", os.path.join(ROOT, "tests", "cpp", name + ".cu"), "-o", exe, f"-L{libdir}", "-lb200fem",
           "-Xlinker", f"-rpath={libdir}"]
    subprocess.check_call(cmd)
    return exe


@pytest.mark.parametrize("name", ["poisson_module_unchanged", "elasticity_callback_unchanged"])
def test_unchanged_reference_text_compiles_against_the_drop_in_headers(tmp_path, name):
    exe = build_unchanged_module(tmp_path, name)
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    if has_gpu:
        assert r.returncode == 0 and "bad=0" in r.stdout, r.stdout + r.stderr
    else:
        assert r.returncode == 3 and "no CUDA device" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_unchanged_poisson_module_on_device(tmp_path):
    """startInit() + compute() of the reference's module text: BSRFormat<1>(ITraceMng*, RunQueue&, FemDoFsOnNodes), initialize(IMesh*, ...),
    computeSparsity(), assembleBilinear(device lambda CellLocalId -> FixedMatrix), toLinearSystem, the GPU boundary functions, solve():
    matrix, rhs and Dirichlet diagonals against a host evaluation of the reference formulas, and against the built-in operator."""
    exe = build_unchanged_module(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.rstrip().endswith("bad=0"), r.stdout
    print(r.stdout)


@pytest.mark.gpu
def test_unchanged_elasticity_callback_on_device(tmp_path):
    """modules/elasticity/FemModule.h:112-152 as the callback of BSRFormat<2>::assembleBilinear, both linear-system kinds."""
    exe = build_unchanged_module(tmp_path, "elasticity_callback_unchanged")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.rstrip().endswith("bad=0"), r.stdout
    print(r.stdout)


@pytest.mark.parametrize("inc,nblocks", [("poisson_module_body.inc", 10), ("elasticity_callback_body.inc", 2)])
def test_module_text_is_verbatim(inc, nblocks):
    """Every block of tests/cpp/*_body.inc equals the cited lines of the reference (skipped where /root/reference is
    absent: the GPU box)."""
    ref_root = "/root/reference"
    if not os.path.isdir(ref_root):
        pytest.skip("the reference tree is not on this machine")
    text = open(os.path.join(ROOT, "tests", "cpp", inc), encoding="utf-8").read().split("\n")
    blocks, cur = [], None
    for line in text:
        if line.startswith("// ---- "):
            name, rng = line[len("// ---- "):].rsplit(":", 1)
            a, b = (int(x) for x in rng.split("-"))
            cur = [name, a, b, []]
            blocks.append(cur)
        elif cur is not None:
            cur[3].append(line)
    assert len(blocks) == nblocks
    for name, a, b, lines in blocks:
        ref = open(os.path.join(ref_root, name), encoding="utf-8-sig").read().split("\n")[a - 1:b]
        assert lines[:len(ref)] == ref, f"{name}:{a}-{b} differs from the reference"
        assert all(x == "" for x in lines[len(ref):])


def test_heat_driver_compiles_and_refuses_to_run_without_a_device(tmp_path):
    """tests/cpp/heat_driver.cc: the heat module's time step through the shim (Operator::Heat, applyConvectionToMatrix / ToRhs,
    applyNodalSourceToRhs, the module's a_ii += P Dirichlet)."""
    exe = build_driver(tmp_path, "heat_driver")
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    if has_gpu:
        assert r.returncode == 0, r.stdout + r.stderr
    else:
        assert r.returncode == 3, r.stdout + r.stderr
        assert "no CUDA device" in r.stdout


@pytest.mark.gpu
def test_heat_driver_on_device(tmp_path):
    exe = build_driver(tmp_path, "heat_driver")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.rstrip().endswith("bad=0"), r.stdout
