"""The C++ drop-in classes (include/zkgpu_dropin: NTT_Goldilocks, PoseidonGoldilocks, MerkleTreeGL with the reference's
signatures) compile against the C ABI, link libzkgpu.so and behave like the reference's own gtest cases
(tests/cpp/dropin_test.cpp restates them).  The CPU part needs no device; the GPU part runs the KATs."""
import pathlib
import subprocess

import pytest

ROOT = pathlib.Path(__file__).resolve().parent.parent
EXE = ROOT / "tests" / "cpp" / "dropin_test"


def build_exe():
    lib_dir = ROOT / "zkevm_prover_b200"
    assert (lib_dir / "libzkgpu.so").exists(), "build libzkgpu.so first (__graft_entry__.build())"
    src = ROOT / "tests" / "cpp" / "dropin_test.cpp"
    if EXE.exists() and EXE.stat().st_mtime > max(f.stat().st_mtime for f in [src] + list((ROOT / "include").rglob("*.h*"))):
        return
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-I", str(ROOT / "include" / "zkgpu_dropin"), str(src), "-o", str(EXE),
                           "-L", str(lib_dir), "-lzkgpu", "-Wl,-rpath," + str(lib_dir)])


def run(mode):
    build_exe()
    r = subprocess.run([str(EXE), mode], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "dropin_test: ok" in r.stdout, r.stdout + r.stderr


def test_dropin_headers_compile_and_validate_arguments():
    run("cpu")


@pytest.mark.gpu
def test_dropin_classes_reproduce_reference_kats():
    run("gpu")
