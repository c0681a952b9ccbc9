"""ctypes loader for the CPU oracle (oracle/liboracle.so, built from oracle/gl_oracle.c) and, when it
exists, the compiled unmodified reference (oracle/_ref/libzkref.so).  TEST INFRASTRUCTURE: imported
only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
import ctypes, hashlib, json, pathlib, subprocess
import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent.parent
P = 0xFFFFFFFF00000001
SEED = 0x9E3779B97F4A7C15
u64 = ctypes.c_uint64
_vp = lambda a: a.ctypes.data_as(ctypes.c_void_p)


def _build():
    so = ROOT / "oracle" / "liboracle.so"
    src = ROOT / "oracle" / "gl_oracle.c"
    if not so.exists() or so.stat().st_mtime < src.stat().st_mtime:
        subprocess.check_call(["make", "-s", "-f", str(ROOT / "oracle" / "Makefile"), str(so)])
    return so


class Oracle:
    def __init__(self):
        self.L = ctypes.CDLL(str(_build()))
        for f in ("oracle_add", "oracle_sub", "oracle_mul"):
            getattr(self.L, f).restype = u64
            getattr(self.L, f).argtypes = [u64, u64]
        self.L.oracle_w.restype = u64
        self.L.oracle_w.argtypes = [ctypes.c_uint]

    def canon(self, a):
        self.L.oracle_canon(_vp(a), u64(a.size))
        return a

    def fill(self, n, seed=SEED, offset=0):
        a = np.empty(n, dtype=np.uint64)
        self.L.oracle_fill(_vp(a), u64(n), u64(seed & (2**64 - 1)), u64(offset))
        return a

    def ntt(self, a):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        n, c = a.shape
        o = np.zeros_like(a)
        self.L.oracle_ntt(_vp(o), _vp(a), u64(n), u64(c))
        return self.canon(o)

    def intt(self, a, extend=False):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        n, c = a.shape
        o = np.zeros_like(a)
        self.L.oracle_intt(_vp(o), _vp(a), u64(n), u64(c), int(extend))
        return self.canon(o)

    def extend_pol(self, a, n_ext):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        n, c = a.shape
        o = np.zeros((n_ext, c), dtype=np.uint64)
        self.L.oracle_extend_pol(_vp(o), _vp(a), u64(n_ext), u64(n), u64(c))
        return self.canon(o)

    def hash_full_result(self, st):
        st = np.ascontiguousarray(st, dtype=np.uint64)
        o = np.zeros(12, dtype=np.uint64)
        self.L.oracle_hash_full_result(_vp(o), _vp(st))
        return self.canon(o)

    def linear_hash(self, x):
        x = np.ascontiguousarray(x, dtype=np.uint64).reshape(-1)
        o = np.zeros(4, dtype=np.uint64)
        self.L.oracle_linear_hash(_vp(o), _vp(x), u64(x.size))
        return self.canon(o)

    def merkletree(self, a, ncols, nrows, dim=1, batch=0):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        tree = np.zeros((2 * nrows - 1) * 4, dtype=np.uint64)
        if batch:
            self.L.oracle_merkletree_batch(_vp(tree), _vp(a), u64(ncols), u64(nrows), u64(batch), u64(dim))
        else:
            self.L.oracle_merkletree(_vp(tree), _vp(a), u64(ncols), u64(nrows), u64(dim))
        return self.canon(tree)


    def stage4(self, q_2ns, n_ext, n, q_dim, q_deg):
        """Starks::genProof stage 4 transforms (starks.cpp:261-285): q_2ns (NExt x qDim) -> cm4_2ns (NExt x qDim*qDeg)."""
        q = np.ascontiguousarray(q_2ns, dtype=np.uint64).reshape(-1)
        out = np.zeros((n_ext, q_dim * q_deg), dtype=np.uint64)
        self.L.oracle_stage4(_vp(out), _vp(q), u64(n_ext), u64(n), u64(q_dim), u64(q_deg))
        return self.canon(out)

    def fri_fold(self, fri_pol, pol_bits, next_bits, shift_inv, special_x):
        """One FRI folding step (friProve.cpp:89-106): fri_pol (2^pol_bits x 3) -> pol2_e (2^next_bits x 3)."""
        a = np.ascontiguousarray(fri_pol, dtype=np.uint64).reshape(-1)
        x = np.ascontiguousarray(special_x, dtype=np.uint64)
        o = np.zeros((1 << next_bits) * 3, dtype=np.uint64)
        self.L.oracle_fri_fold(_vp(o), _vp(a), u64(pol_bits), u64(next_bits), u64(shift_inv), _vp(x))
        return self.canon(o).reshape(-1, 3)

    def fri_transpose(self, pol, transpose_bits):
        a = np.ascontiguousarray(pol, dtype=np.uint64).reshape(-1)
        o = np.zeros_like(a)
        self.L.oracle_fri_transpose(_vp(o), _vp(a), u64(a.size // 3), u64(transpose_bits))
        return o.reshape(-1, 3)


class Reference:
    """The compiled unmodified reference (None-like when oracle/_ref/libzkref.so is absent)."""

    def __init__(self):
        p = ROOT / "oracle" / "_ref" / "libzkref.so"
        self.L = ctypes.CDLL(str(p)) if p.exists() else None
        if self.L is not None:
            for f in ("zkref_add", "zkref_sub", "zkref_mul"):
                getattr(self.L, f).restype = u64
                getattr(self.L, f).argtypes = [u64, u64]

    def __bool__(self):
        return self.L is not None

    def canon(self, a):
        self.L.zkref_canon(_vp(a), u64(a.size))
        return a

    def ntt(self, a, inverse=False, extend=False):
        a = np.ascontiguousarray(a, dtype=np.uint64).copy()
        o = np.zeros_like(a); buf = np.zeros_like(a)
        self.L.zkref_ntt(_vp(o), _vp(a), u64(a.shape[0]), u64(a.shape[1]), _vp(buf), u64(3), u64(1), int(inverse), int(extend))
        return self.canon(o)

    def extend_pol(self, a, n_ext):
        a = np.ascontiguousarray(a, dtype=np.uint64).copy()
        o = np.zeros((n_ext, a.shape[1]), dtype=np.uint64); buf = np.zeros_like(o)
        self.L.zkref_extend_pol(_vp(o), _vp(a), u64(n_ext), u64(a.shape[0]), u64(a.shape[1]), _vp(buf))
        return self.canon(o)

    def hash_full_result(self, st, seq=False):
        st = np.ascontiguousarray(st, dtype=np.uint64)
        o = np.zeros(12, dtype=np.uint64)
        (self.L.zkref_hash_full_result_seq if seq else self.L.zkref_hash_full_result)(_vp(o), _vp(st))
        return self.canon(o)

    def linear_hash(self, x):
        x = np.ascontiguousarray(x, dtype=np.uint64).reshape(-1).copy()
        o = np.zeros(4, dtype=np.uint64)
        self.L.zkref_linear_hash(_vp(o), _vp(x), u64(x.size))
        return self.canon(o)

    def merkletree(self, a, ncols, nrows, dim=1, batch=0):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        tree = np.zeros((2 * nrows - 1) * 4, dtype=np.uint64)
        if batch:
            self.L.zkref_merkletree_batch(_vp(tree), _vp(a), u64(ncols), u64(nrows), u64(batch), u64(dim))
        else:
            self.L.zkref_merkletree(_vp(tree), _vp(a), u64(ncols), u64(nrows), u64(dim))
        return self.canon(tree)

    def reverse_permutation(self, src, size, offset_cols, ncols, ncols_all, extension=1):
        src = np.ascontiguousarray(src, dtype=np.uint64).copy()
        dst = np.zeros((size, ncols), dtype=np.uint64)
        self.L.zkref_reverse_permutation(_vp(dst), _vp(src), u64(size), u64(offset_cols), u64(ncols), u64(ncols_all), int(extension))
        return dst

    def stage4(self, q_2ns, n_ext, n, q_dim, q_deg):
        """Starks::genProof stage 4 (starks.cpp:261-285): q_2ns (NExt x qDim) -> cm4_2ns (NExt x qDim*qDeg)."""
        q = np.ascontiguousarray(q_2ns, dtype=np.uint64).reshape(-1).copy()
        out = np.zeros((n_ext, q_dim * q_deg), dtype=np.uint64)
        self.L.zkref_stage4(_vp(out), _vp(q), u64(n_ext), u64(n), u64(q_dim), u64(q_deg))
        return self.canon(out)

    def fri_fold(self, fri_pol, pol_bits, next_bits, shift_inv, special_x):
        a = np.ascontiguousarray(fri_pol, dtype=np.uint64).reshape(-1).copy()
        x = np.ascontiguousarray(special_x, dtype=np.uint64).copy()
        o = np.zeros((1 << next_bits) * 3, dtype=np.uint64)
        self.L.zkref_fri_fold(_vp(o), _vp(a), u64(pol_bits), u64(next_bits), u64(shift_inv), _vp(x))
        return self.canon(o).reshape(-1, 3)


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype="<u8").tobytes()).hexdigest()


def load_golden(name):
    return json.loads((ROOT / "tests" / "golden" / name).read_text())


def hexwords(lst):
    return np.array([int(x, 16) for x in lst], dtype=np.uint64)


def field_fib(n, a0, a1):
    """Field Fibonacci with canonical words (the reference builds these with Goldilocks::add)."""
    v = [a0 % P, a1 % P]
    for _ in range(2, n):
        v.append((v[-1] + v[-2]) % P)
    return np.array(v[:n], dtype=np.uint64)


def merkle_kat_matrix(nrows, ncols):
    """tests.cpp:1887-1900 / bench.cpp:409-424: rows 0,1 = 1..ncols; row j = row j-1 + row j-2 (mod p)."""
    m = np.zeros((nrows, ncols), dtype=np.uint64)
    base = np.arange(1, ncols + 1, dtype=np.uint64)
    m[0] = base
    if nrows > 1:
        m[1] = base
    Pn = np.uint64(P)
    for j in range(2, nrows):
        s = m[j - 1] + m[j - 2]  # may wrap 2^64
        wrapped = s < m[j - 1]
        s = np.where(wrapped, s + np.uint64(0xFFFFFFFF), s)  # 2^64 = 2^32-1 mod p
        s = np.where(s >= Pn, s - Pn, s)
        m[j] = s
    return m
