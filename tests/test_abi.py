"""CPU-side checks of the drop-in boundary: libzkgpu.so loads, exports every symbol include/zkgpu.h declares, the
Python mirror binds exactly that set, argument validation that needs no device behaves like the reference, and the
product fails loudly (no CPU fallback) when no CUDA device is present."""
import ctypes
import pathlib
import re
import subprocess

import numpy as np
import pytest

import zkevm_prover_b200 as zk

ROOT = pathlib.Path(__file__).resolve().parent.parent


def header_symbols():
    txt = (ROOT / "include" / "zkgpu.h").read_text()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return set(re.findall(r"\b(zkgpu_\w+)\s*\(", txt))


def test_header_python_and_library_agree():
    hs = header_symbols()
    assert hs == set(zk.ABI_SYMBOLS), (hs ^ set(zk.ABI_SYMBOLS))
    L = ctypes.CDLL(str(zk.lib_path))
    for name in hs:
        assert hasattr(L, name), name
    nm = subprocess.run(["nm", "-D", "--defined-only", str(zk.lib_path)], capture_output=True, text=True).stdout
    exported = set(re.findall(r"\bT (zkgpu_\w+)", nm))
    assert hs <= exported


def test_library_contains_sm100a_code_only():
    out = subprocess.run(["cuobjdump", "-lelf", str(zk.lib_path)], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_argument_validation_without_device():
    # size 0 / ncols 0 are silent no-ops (ntt_goldilocks.cpp:168-171, 310-313; poseidon_goldilocks.cpp:268-271)
    ntt = zk.NTT_Goldilocks(16)
    z = np.zeros(4, dtype=np.uint64)
    ntt.NTT(z, z, 0, 1)
    ntt.INTT(z, z, 4, 0)
    ntt.extendPol(z, z, 8, 4, 0)
    zk.PoseidonGoldilocks.merkletree(z, z, 3, 0)
    with pytest.raises(zk.ZkGpuError):
        ntt.NTT(z, z, 3, 1)  # the reference asserts on non-power-of-two sizes (ntt_goldilocks.cpp:28)
    with pytest.raises(OverflowError):
        zk.NTT_Goldilocks(1 << 33)  # std::range_error in the reference (ntt_goldilocks.hpp:94-97)


def test_no_cpu_fallback():
    if zk.lib().zkgpu_device_count() > 0:
        pytest.skip("a CUDA device is present")
    a = np.arange(16, dtype=np.uint64)
    with pytest.raises(zk.ZkGpuError, match="no usable CUDA device"):
        zk.NTT_Goldilocks(16).NTT(a, a, 16, 1)
    with pytest.raises(zk.ZkGpuError, match="no usable CUDA device"):
        zk.PoseidonGoldilocks.hash_full_result(np.zeros(12, dtype=np.uint64), np.zeros(12, dtype=np.uint64))


def test_product_does_not_reference_the_oracle():
    for f in list((ROOT / "zkevm_prover_b200").rglob("*.py")) + list((ROOT / "zkevm_prover_b200" / "csrc").glob("*.cu*")) + \
            [f for f in (ROOT / "include").rglob("*") if f.is_file()]:
        txt = f.read_text()
        assert "liboracle" not in txt and "gl_oracle" not in txt and "oracle_lib" not in txt, f
