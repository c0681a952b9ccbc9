"""zkevm_prover_b200 — host-side Python mirror of the reference's commit-stage classes over libzkgpu.so.

The product is the CUDA library (csrc/ -> libzkgpu.so, C ABI in include/zkgpu.h) and the C++ drop-in headers in
include/.  This module binds the same C ABI with ctypes and mirrors the reference's class surface
(NTT_Goldilocks, PoseidonGoldilocks, MerkleTreeGL: same method names, argument order and defaults as
src/goldilocks/src/ntt_goldilocks.hpp:49,166-173, src/goldilocks/src/poseidon_goldilocks.hpp:43-81 and
src/starkpil/merkleTree/merkleTreeGL.hpp:17-78) so that tests and the benchmark read like the reference's.

There is no CPU fallback: importing works anywhere (so that the CPU test-suite can check the ABI), but every
compute call fails with ZkGpuError when the library or a CUDA device is missing.

Buffers may be numpy uint64 arrays (host memory; staged through the device by the library), torch CUDA tensors
of dtype int64/uint64 (used in place) or raw integer addresses.
"""
import ctypes
import pathlib

import numpy as np

__all__ = ["ZkGpuError", "lib", "lib_path", "ABI_SYMBOLS", "NTT_Goldilocks", "PoseidonGoldilocks", "MerkleTreeGL", "FRIProve",
           "extend_and_merkelize", "stage4_commit", "invalidate", "cache_stats", "group_proofs", "build_const_tree", "fill", "stage_ms", "set_stream", "sync", "GOLDILOCKS_PRIME"]

GOLDILOCKS_PRIME = 0xFFFFFFFF00000001
import os

_HERE = pathlib.Path(__file__).resolve().parent
# ZKGPU_LIB: an alternative build of the same library (kernel-variant experiments); the product is libzkgpu.so beside this file
lib_path = pathlib.Path(os.environ["ZKGPU_LIB"]) if os.environ.get("ZKGPU_LIB") else _HERE / "libzkgpu.so"

_u64 = ctypes.c_uint64
_vp = ctypes.c_void_p
_int = ctypes.c_int

# every symbol include/zkgpu.h declares: (restype, argtypes)
ABI_SYMBOLS = {
    "zkgpu_init": (_int, [_int]),
    "zkgpu_set_stream": (_int, [_vp]),
    "zkgpu_sync": (_int, []),
    "zkgpu_last_error": (ctypes.c_char_p, []),
    "zkgpu_device_count": (_int, []),
    "zkgpu_launch_count": (_u64, []),
    "zkgpu_tune": (_int, [_int, _int]),
    "zkgpu_mem_info": (_int, [ctypes.POINTER(_u64), ctypes.POINTER(_u64)]),
    "zkgpu_malloc": (_int, [ctypes.POINTER(_vp), _u64]),
    "zkgpu_free": (_int, [_vp]),
    "zkgpu_malloc_host": (_int, [ctypes.POINTER(_vp), _u64]),
    "zkgpu_free_host": (_int, [_vp]),
    "zkgpu_memcpy": (_int, [_vp, _vp, _u64]),
    "zkgpu_stage_ms": (_int, [ctypes.POINTER(ctypes.c_float)]),
    "zkgpu_fill": (_int, [_vp, _u64, _u64, _u64]),
    "zkgpu_field_op": (_int, [_int, _vp, _vp, _vp, _u64]),
    "zkgpu_ntt_create": (_vp, [_u64, _int]),
    "zkgpu_ntt_destroy": (None, [_vp]),
    "zkgpu_ntt": (_int, [_vp, _vp, _vp, _u64, _u64, _vp, _int, _int]),
    "zkgpu_extend_pol": (_int, [_vp, _vp, _vp, _u64, _u64, _u64, _vp]),
    "zkgpu_reverse_permutation": (_int, [_vp, _vp, _vp, _u64, _u64, _u64, _u64]),
    "zkgpu_group_proof": (_int, [_vp, _vp, _vp, _u64, _u64, _u64]),
    "zkgpu_group_proofs": (_int, [_vp, _vp, _vp, _u64, _u64, _vp, _u64]),
    "zkgpu_hash_full_result": (_int, [_vp, _vp, _u64]),
    "zkgpu_linear_hash": (_int, [_vp, _vp, _u64]),
    "zkgpu_merkletree": (_int, [_vp, _vp, _u64, _u64, _u64]),
    "zkgpu_merkletree_batch": (_int, [_vp, _vp, _u64, _u64, _u64, _u64]),
    "zkgpu_merkletree_batch_noalloc": (_int, [_vp, _vp, _u64, _u64, _u64, _u64]),
    "zkgpu_extend_and_merkelize": (_int, [_vp, _vp, _vp, _vp, _u64, _u64, _u64]),
    "zkgpu_build_const_tree": (_int, [_vp, _vp, _u64, _u64, _u64]),
    "zkgpu_lde_columns": (_int, [_vp, _u64, _vp, _u64, _u64, _u64, _u64]),
    "zkgpu_merkle_leaves": (_int, [_vp, _vp, _u64, _u64, _u64]),
    "zkgpu_merkle_absorb": (_int, [_vp, _vp, _u64, _u64, _u64, _int]),
    "zkgpu_merkle_absorb_copy": (_int, [_vp, _vp, _u64, _u64, _u64, _int, _vp, _u64]),
    "zkgpu_fri_fold": (_int, [_vp, _vp, _u64, _u64, _u64, _vp]),
    "zkgpu_fri_transpose": (_int, [_vp, _vp, _u64, _u64]),
    "zkgpu_merkle_levels": (_int, [_vp, _u64]),
    "zkgpu_copy_block": (_int, [_vp, _u64, _vp, _u64, _u64, _u64]),
    "zkgpu_invalidate": (_int, [_vp]),
    "zkgpu_pipe_probe": (_int, [ctypes.POINTER(ctypes.c_double)]),
    "zkgpu_cache_stats": (_int, [ctypes.POINTER(_u64)]),
    "zkgpu_stage4_commit": (_int, [_vp, _vp, _vp, _vp, _u64, _u64, _u64, _u64]),
    "zkgpu_init_multi": (_int, [_int]),
    "zkgpu_nccl_unique_id": (_int, [_vp]),
    "zkgpu_dist_init": (_int, [_vp, _int, _int]),
    "zkgpu_dist_info": (_int, [ctypes.POINTER(_int), ctypes.POINTER(_int), ctypes.POINTER(_int)]),
    "zkgpu_dist_plan": (_int, [_u64, _int, _int, ctypes.POINTER(_int), ctypes.POINTER(_u64), ctypes.POINTER(_u64)]),
    "zkgpu_dist_commit": (_int, [_vp, _vp, ctypes.POINTER(_vp), ctypes.POINTER(_u64), _int, _int, _u64, _u64, _u64]),
    "zkgpu_dist_finalize": (_int, []),
    "zkgpu_dist_rows": (_int, [ctypes.POINTER(_vp), ctypes.POINTER(_u64), ctypes.POINTER(_u64), ctypes.POINTER(_vp)]),
}


class ZkGpuError(RuntimeError):
    pass


_lib = None


def lib():
    """The loaded libzkgpu.so (built by __graft_entry__.build() / make -C zkevm_prover_b200/csrc)."""
    global _lib
    if _lib is None:
        if not lib_path.exists():
            raise ZkGpuError("%s is missing: build it with `make -C zkevm_prover_b200/csrc` "
                             "(there is no CPU fallback)" % lib_path)
        L = ctypes.CDLL(str(lib_path))
        for name, (res, args) in ABI_SYMBOLS.items():
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        _lib = L
    return _lib


def _check(rc):
    if rc is not None and rc < 0:
        raise ZkGpuError(lib().zkgpu_last_error().decode("utf-8", "replace"))


def _ptr(x):
    """Address of a numpy array / torch tensor / int / None."""
    if x is None:
        return None
    if isinstance(x, int):
        return x
    if isinstance(x, np.ndarray):
        if x.dtype != np.uint64 or not x.flags["C_CONTIGUOUS"]:
            raise TypeError("host buffers must be C-contiguous numpy uint64 arrays")
        return x.ctypes.data
    if hasattr(x, "data_ptr"):  # torch tensor
        if x.element_size() != 8 or not x.is_contiguous():
            raise TypeError("device buffers must be contiguous 64-bit torch tensors")
        return x.data_ptr()
    raise TypeError("unsupported buffer type %r" % type(x))


def init(device=-1):
    _check(lib().zkgpu_init(device))


def set_stream(stream=None):
    """Enqueue on `stream` (a torch.cuda.Stream, a raw cudaStream_t integer, or None for the library's own)."""
    h = ctypes.c_void_p(-1) if stream is None else ctypes.c_void_p(getattr(stream, "cuda_stream", stream))
    _check(lib().zkgpu_set_stream(h))


def sync():
    _check(lib().zkgpu_sync())


def stage_ms():
    """{'lde', 'leaves', 'nodes', 'unused', 'total'} device times in ms of the last commit-type call (zkgpu.h: slot [3] is
    unused, slot [4] is the whole call including host<->device copies when any pointer was a host pointer)."""
    out = (ctypes.c_float * 5)()
    _check(lib().zkgpu_stage_ms(out))
    return dict(zip(("lde", "leaves", "nodes", "unused", "total"), [float(v) for v in out]))


def fill(dst, n, seed=0x9E3779B97F4A7C15, offset=0):
    """dst[i] = splitmix64(seed + offset + i) mod p — the synthetic trace generator shared with the oracle."""
    _check(lib().zkgpu_fill(_ptr(dst), n, seed & (2**64 - 1), offset))


def field_op(op, a, b):
    """Element-wise device field arithmetic (KAT hook): op in {'add','sub','mul','mul_nc','canon'}."""
    code = {"add": 0, "sub": 1, "mul": 2, "mul_nc": 3, "canon": 4}[op]
    a = np.ascontiguousarray(a, dtype=np.uint64)
    b = np.ascontiguousarray(b, dtype=np.uint64)
    out = np.zeros_like(a)
    _check(lib().zkgpu_field_op(code, _ptr(out), _ptr(a), _ptr(b), a.size))
    return out


class NTT_Goldilocks:
    """Mirror of the reference class (src/goldilocks/src/ntt_goldilocks.hpp:49-173)."""

    def __init__(self, maxDomainSize, nThreads=0, extension=1):
        self._h = lib().zkgpu_ntt_create(maxDomainSize, extension)
        if not self._h:
            # the reference throws std::range_error("Domain size too big for the curve")
            raise OverflowError(lib().zkgpu_last_error().decode())
        self.maxDomainSize, self.nThreads, self.extension = maxDomainSize, nThreads, extension

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                lib().zkgpu_ntt_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def NTT(self, dst, src, size, ncols=1, buffer=None, nphase=3, nblock=1, inverse=False, extend=False):
        _check(lib().zkgpu_ntt(self._h, _ptr(dst), _ptr(src), size, ncols, _ptr(buffer), int(inverse), int(extend)))

    def INTT(self, dst, src, size, ncols=1, buffer=None, nphase=3, nblock=1, extend=False):
        _check(lib().zkgpu_ntt(self._h, _ptr(dst), _ptr(src), size, ncols, _ptr(buffer), 1, int(extend)))

    def reversePermutation(self, dst, src, size, offset_cols, ncols, ncols_all):
        _check(lib().zkgpu_reverse_permutation(self._h, _ptr(dst), _ptr(src), size, offset_cols, ncols, ncols_all))

    def extendPol(self, output, input, N_Extended, N, ncols, buffer=None, nphase=3, nblock=1):
        _check(lib().zkgpu_extend_pol(self._h, _ptr(output), _ptr(input), N_Extended, N, ncols, _ptr(buffer)))


class PoseidonGoldilocks:
    """Mirror of the reference's static class (src/goldilocks/src/poseidon_goldilocks.hpp:43-81)."""
    RATE, CAPACITY, HASH_SIZE, SPONGE_WIDTH = 8, 4, 4, 12

    @staticmethod
    def hash_full_result(out, inp, n=1):
        _check(lib().zkgpu_hash_full_result(_ptr(out), _ptr(inp), n))

    @staticmethod
    def hash(out4, inp12):
        full = np.zeros(12, dtype=np.uint64)
        PoseidonGoldilocks.hash_full_result(full, inp12)
        out4[:4] = full[:4]

    @staticmethod
    def linear_hash(out4, inp, size):
        _check(lib().zkgpu_linear_hash(_ptr(out4), _ptr(inp), size))

    @staticmethod
    def merkletree(tree, inp, num_cols, num_rows, nThreads=0, dim=1):
        _check(lib().zkgpu_merkletree(_ptr(tree), _ptr(inp), num_cols, num_rows, dim))

    merkletree_avx = merkletree
    merkletree_seq = merkletree

    @staticmethod
    def merkletree_batch(tree, inp, num_cols, num_rows, batch_size, nThreads=0, dim=1):
        _check(lib().zkgpu_merkletree_batch(_ptr(tree), _ptr(inp), num_cols, num_rows, batch_size, dim))


class MerkleTreeGL:
    """Mirror of src/starkpil/merkleTree/merkleTreeGL.hpp:9-79 / .cpp:5-44 for host (numpy) sources."""

    def __init__(self, height, width, source=None, nodes=None):
        self.height, self.width = height, width
        self.source = source if source is not None else np.zeros(height * width, dtype=np.uint64)
        # nodes given: wrap caller-owned buffers (e.g. the device-resident tree the commit stage left behind)
        self.nodes = nodes if nodes is not None else np.zeros(self.getTreeNumElements(), dtype=np.uint64)

    def getTreeNumElements(self):
        # MerklehashGoldilocks::getTreeNumElements, merklehash_goldilocks.hpp:15-19
        return self.height * 4 + (self.height - 1) * 4

    def MerkleProofSize(self):
        h, n = self.height, 0
        while h > 1:
            h >>= 1
            n += 1
        return n

    def merkelize(self):
        PoseidonGoldilocks.merkletree(self.nodes, self.source, self.width, self.height)

    def getRoot(self):
        root = np.zeros(4, dtype=np.uint64)
        _check(lib().zkgpu_memcpy(_ptr(root), _ptr(self.nodes) + 8 * (self.getTreeNumElements() - 4), 32))
        return root

    def getGroupProof(self, idx):
        """merkleTreeGL.cpp:12-35: the source row followed by the sibling path (4 words per level); source and nodes may
        live on the host (numpy) or on the device (torch tensors)."""
        proof = np.zeros(self.width + 4 * self.MerkleProofSize(), dtype=np.uint64)
        _check(lib().zkgpu_group_proof(_ptr(proof), _ptr(self.source), _ptr(self.nodes), self.width, self.height, idx))
        return proof


def group_proofs(tree, idxs):
    """MerkleTreeGL::getGroupProof for a batch of row indices (one launch, one copy): array of len(idxs) proofs."""
    idx = np.ascontiguousarray(idxs, dtype=np.uint64)
    out = np.zeros((idx.size, tree.width + 4 * tree.MerkleProofSize()), dtype=np.uint64)
    _check(lib().zkgpu_group_proofs(_ptr(out), _ptr(tree.source), _ptr(tree.nodes), tree.width, tree.height, _ptr(idx), idx.size))
    return out


def build_const_tree(image, const_pols, nBits, nBitsExt, nPols):
    """buildConstTree, Goldilocks branch (tools/starkpil/bctree/build_const_tree.cpp:533-596): fills `image` with
    [nPols, nExt, extended constant polynomials, tree nodes]."""
    _check(lib().zkgpu_build_const_tree(_ptr(image), _ptr(const_pols), nBits, nBitsExt, nPols))


class FRIProve:
    """The data-parallel pieces of FRIProve::prove (src/starkpil/fri/friProve.cpp); the transcript and the proof
    object stay with the caller."""

    @staticmethod
    def fold(pol2_e, friPol, polBits, nextBits, shiftInv, special_x):
        """Folding step si > 0 (friProve.cpp:44-107): friPol (2^polBits x 3) -> pol2_e (2^nextBits x 3)."""
        x = np.ascontiguousarray(special_x, dtype=np.uint64)
        if x.size != 3:
            raise ValueError("special_x is one cubic-extension element (3 words)")
        _check(lib().zkgpu_fri_fold(_ptr(pol2_e), _ptr(friPol), polBits, nextBits, int(shiftInv), _ptr(x)))

    @staticmethod
    def getTransposed(aux, pol, degree, trasposeBits):
        """friProve.cpp:252-271 on 3-word elements."""
        _check(lib().zkgpu_fri_transpose(_ptr(aux), _ptr(pol), degree, trasposeBits))


def stage4_commit(ntt_extended, cm4_2ns, tree, q_2ns, N_Extended, N, qDim, qDeg):
    """Starks::genProof stage 4 (starks.cpp:261-293): INTT, shift/regroup, NTT, merkelize — one device-resident call."""
    _check(lib().zkgpu_stage4_commit(ntt_extended._h if ntt_extended is not None else None, _ptr(cm4_2ns), _ptr(tree), _ptr(q_2ns),
                                     N_Extended, N, qDim, qDeg))


def invalidate(host_buffer=None):
    """Drop commit-cache entries for this host buffer (None: all of them)."""
    _check(lib().zkgpu_invalidate(_ptr(host_buffer)))


def cache_stats():
    """Commit-cache counters: dict(hits, misses, stale, dropped, entries, device_bytes)."""
    out = (_u64 * 6)()
    _check(lib().zkgpu_cache_stats(out))
    return dict(zip(("hits", "misses", "stale", "dropped", "entries", "device_bytes"), (int(x) for x in out)))


def extend_and_merkelize(ntt, out_ext, tree, inp, N_Extended, N, ncols):
    """Starks::genProof commit stage k (starks.cpp:51-58): extendPol then MerkleTreeGL::merkelize, one call."""
    _check(lib().zkgpu_extend_and_merkelize(ntt._h if ntt is not None else None, _ptr(out_ext), _ptr(tree), _ptr(inp),
                                            N_Extended, N, ncols))
