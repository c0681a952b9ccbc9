"""The C++ drop-in classes (include/zkgpu_dropin: NTT_Goldilocks, PoseidonGoldilocks, MerkleTreeGL with the reference's
signatures) compile against the C ABI, link libzkgpu.so and behave like the reference's own gtest cases
(tests/cpp/dropin_test.cpp restates them).  The CPU part needs no device; the GPU part runs the KATs."""
import pathlib
import subprocess

import pytest

ROOT = pathlib.Path(__file__).resolve().parent.parent
EXE = ROOT / "tests" / "cpp" / "dropin_test"


EXE_REF = ROOT / "tests" / "cpp" / "dropin_test_refhdr"
REF_SRC = pathlib.Path("/root/reference/src/goldilocks/src")


def build_exe(with_reference_headers=False):
    """with_reference_headers: the configuration INTEGRATION.md section 2 prescribes — the reference's own
    goldilocks_base_field.hpp (guard GOLDILOCKS_BASE) comes first and the drop-in headers must compile after it.  Needs
    /root/reference (this container); the binary travels to the GPU box."""
    lib_dir = ROOT / "zkevm_prover_b200"
    assert (lib_dir / "libzkgpu.so").exists(), "build libzkgpu.so first (__graft_entry__.build())"
    src = ROOT / "tests" / "cpp" / "dropin_test.cpp"
    exe = EXE_REF if with_reference_headers else EXE
    if exe.exists() and exe.stat().st_mtime > max(f.stat().st_mtime for f in [src] + list((ROOT / "include").rglob("*.h*"))):
        return exe
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-I", str(ROOT / "include" / "zkgpu_dropin")]
    if with_reference_headers:
        cmd += ["-w", "-fopenmp", "-mavx2", "-I", str(ROOT / "oracle" / "shim"), "-I", str(REF_SRC), "-include", "cstring", "-include", "cstdlib",
                "-include", "stdexcept", "-include", str(REF_SRC / "goldilocks_base_field.hpp"), str(REF_SRC / "goldilocks_base_field.cpp")]
    subprocess.check_call(cmd + [str(src), "-o", str(exe), "-L", str(lib_dir), "-lzkgpu", "-Wl,-rpath," + str(lib_dir)])
    return exe


def run(mode, with_reference_headers=False):
    exe = build_exe(with_reference_headers) if (not with_reference_headers or REF_SRC.exists()) else EXE_REF
    r = subprocess.run([str(exe), mode], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "dropin_test: ok" in r.stdout, r.stdout + r.stderr


def test_dropin_headers_compile_and_validate_arguments():
    run("cpu")


@pytest.mark.gpu
def test_dropin_classes_reproduce_reference_kats():
    run("gpu")


@pytest.mark.skipif(not REF_SRC.exists() and not EXE_REF.exists(), reason="needs /root/reference or the prebuilt binary")
def test_dropin_headers_compile_after_the_reference_field_header():
    run("cpu", with_reference_headers=True)


@pytest.mark.gpu
@pytest.mark.skipif(not REF_SRC.exists() and not EXE_REF.exists(), reason="needs /root/reference or the prebuilt binary")
def test_dropin_with_reference_field_header_on_gpu():
    run("gpu", with_reference_headers=True)


@pytest.mark.gpu
def test_dropin_two_gpus_one_process():
    import zkevm_prover_b200 as zk
    if zk.lib().zkgpu_device_count() < 2:
        pytest.skip("needs 2 GPUs")
    run("gpu2")
