#!/usr/bin/env python3
"""Generate tests/golden/*.json.  Runs ONLY in the build container (needs /root/reference and the
compiled unmodified reference oracle/_ref/libzkref.so); the JSON files are committed and are what
the GPU box and the CPU test-suite read.

1. reference_kats.json  — every known-answer vector the reference's own tests hold for the hot path,
   extracted textually from src/goldilocks/tests/tests.cpp and src/goldilocks/benchs/bench.cpp
   (the expected words inside the named TEST bodies; file:line recorded per entry).
2. reference_diff.json  — outputs of the compiled reference on seeded synthetic inputs for shapes the
   reference's tests do not cover (odd ncols, blowup 16, dim 3, batch): sha256 over the canonical
   little-endian words + a few sample words + Merkle roots.
"""
import ctypes, hashlib, json, pathlib, re
import numpy as np

HERE = pathlib.Path(__file__).resolve().parent
ROOT = HERE.parent.parent
REFT = pathlib.Path("/root/reference/src/goldilocks/tests/tests.cpp")
REFB = pathlib.Path("/root/reference/src/goldilocks/benchs/bench.cpp")
P = 0xFFFFFFFF00000001


def test_body(txt, name):
    """(body, first_line) of `TEST(GOLDILOCKS_TEST, name)`."""
    m = re.search(r"TEST\(GOLDILOCKS_TEST, %s\)" % re.escape(name), txt)
    start = m.start()
    nxt = re.search(r"\n(TEST\(|#ifdef __AVX512__|#endif)", txt[m.end():])
    end = m.end() + (nxt.start() if nxt else len(txt))
    return txt[start:end], txt[:start].count("\n") + 1


def asserts(body):
    return [int(x, 16) for x in re.findall(r"ASSERT_EQ\(Goldilocks::toU64\([^)]*\)\)?,\s*(0[xX][0-9a-fA-F]+)\)", body)]


def kats():
    t = REFT.read_text()
    out = {}

    def put(key, name, sl, desc):
        body, line = test_body(t, name)
        vals = asserts(body)[sl]
        out[key] = {"source": "src/goldilocks/tests/tests.cpp:%d TEST(%s)" % (line, name), "input": desc,
                    "expected": ["%016x" % v for v in vals]}
        return vals

    put("poseidon_fibonacci", "poseidon_full_avx", slice(0, 12), "state = 0,1,1,2,...,89 (field Fibonacci)")
    put("poseidon_zero", "poseidon_full_avx", slice(12, 24), "state = 12 zeros")
    assert asserts(test_body(t, "poseidon_full_seq")[0])[:24] == asserts(test_body(t, "poseidon_full_avx")[0])[:24]
    put("linear_hash_fib128", "linear_hash_avx", slice(0, 4), "128 field-Fibonacci words starting 0,1")
    put("merkle_root_64x128", "merkletree_avx", slice(0, 4), "64 rows x 128 cols; rows 0,1 = 1..128; row j = row j-1 + row j-2")
    put("merkle_root_ncols0_64", "merkletree_avx", slice(4, 8), "64 rows, ncols = 0")
    put("merkle_root_ncols0_131072", "merkletree_avx", slice(8, 12), "2^17 rows, ncols = 0")
    assert asserts(test_body(t, "merkletree_seq")[0])[:12] == asserts(test_body(t, "merkletree_avx")[0])[:12]
    put("merkle_batch_root_64x128_b32", "merkletree_batch", slice(0, 4), "same 64x128 matrix, batch_size 32")
    put("lde_shift49_16to32", "LDE", slice(0, 32), "16-pt Fibonacci 1,1,..; INTT; *49^i; zero-pad; NTT 32")
    put("lde_block_16x8_col0", "LDE_block", slice(0, 32), "16x8, col j Fibonacci from (1+j,1+j); shift 7; column 0 of the 32 output rows")
    put("extendpol_16x8_col0", "extendePol", slice(0, 32), "same 16x8 input through extendPol(a,a,32,16,8,b); column 0")
    assert out["lde_block_16x8_col0"]["expected"] == out["extendpol_16x8_col0"]["expected"]

    b = REFB.read_text()
    for key, fn in (("merkle_root_8388608x128", "MERKLETREE_BENCH_AVX"), ("merkle_batch_root_8388608x128_b32", "MERKLETREE_BATCH_BENCH_AVX")):
        m = re.search(r"static void %s\(" % fn, b)
        seg = b[m.start():]
        seg = seg[:seg.index("\n}\n")]
        vals = re.findall(r"toU64\(root\[\d\]\)\s*==\s*(0[xX][0-9a-fA-F]+)", seg)
        assert len(vals) == 4, (fn, vals)
        out[key] = {"source": "src/goldilocks/benchs/bench.cpp:%d %s" % (b[:m.start()].count("\n") + 1, fn),
                    "input": "2^23 rows x 128 cols, same generator as merkle_root_64x128",
                    "expected": ["%016x" % int(v, 16) for v in vals]}
    return out


def splitmix_fill(n, seed, offset=0):
    x = (np.arange(n, dtype=np.uint64) + np.uint64((seed + offset) & (2**64 - 1)))
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        x = x ^ (x >> np.uint64(31))
    return x % np.uint64(P)


def digest(a):
    a = np.ascontiguousarray(a, dtype="<u8")
    return hashlib.sha256(a.tobytes()).hexdigest()


def diffs():
    L = ctypes.CDLL(str(ROOT / "oracle" / "_ref" / "libzkref.so"))
    vp = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    u = ctypes.c_uint64
    SEED = 0x9E3779B97F4A7C15
    out = {"generator": "in[i] = splitmix64(seed + i) mod p, seed = 0x9E3779B97F4A7C15 (oracle_fill / zkgpu_fill)", "cases": []}

    def samples(a):
        f = a.reshape(-1)
        idx = sorted(set([0, 1, len(f) // 3, len(f) // 2, len(f) - 2, len(f) - 1]))
        return {str(i): "%016x" % int(f[i]) for i in idx if 0 <= i < len(f)}

    for (n, c, blow) in [(1 << 10, 5, 2), (1 << 8, 665, 2), (1 << 7, 12, 16), (1 << 12, 1, 2), (1 << 9, 33, 4), (4, 3, 2), (1, 3, 2)]:
        a = splitmix_fill(n * c, SEED).reshape(n, c)
        o = np.zeros((n * blow, c), dtype=np.uint64)
        buf = np.zeros((n * blow, c), dtype=np.uint64)
        L.zkref_extend_pol(vp(o), vp(a.copy()), u(n * blow), u(n), u(c), vp(buf))
        L.zkref_canon(vp(o), u(o.size))
        tree = np.zeros((2 * n * blow - 1) * 4, dtype=np.uint64)
        L.zkref_merkletree(vp(tree), vp(o), u(c), u(n * blow), u(1))
        L.zkref_canon(vp(tree), u(tree.size))
        out["cases"].append({"op": "extend_and_merkelize", "n": n, "ncols": c, "blowup": blow,
                             "ext_sha256": digest(o), "ext_samples": samples(o),
                             "tree_sha256": digest(tree), "root": ["%016x" % int(x) for x in tree[-4:]]})
    for (n, c, inv) in [(1 << 12, 3, 0), (1 << 12, 3, 1), (1 << 9, 7, 0), (1 << 5, 2, 1), (1 << 16, 1, 0)]:
        a = splitmix_fill(n * c, SEED + 1).reshape(n, c)
        o = np.zeros_like(a)
        buf = np.zeros_like(a)
        L.zkref_ntt(vp(o), vp(a.copy()), u(n), u(c), vp(buf), u(3), u(1), inv, 0)
        L.zkref_canon(vp(o), u(o.size))
        out["cases"].append({"op": "intt" if inv else "ntt", "n": n, "ncols": c, "seed_offset": 1,
                             "sha256": digest(o), "samples": samples(o)})
    for (rows, c, dim, batch) in [(1 << 9, 665, 1, 0), (1 << 8, 7, 3, 0), (1 << 10, 4, 1, 0), (1 << 10, 9, 1, 0),
                                  (1 << 6, 100, 1, 32), (1 << 6, 37, 1, 8), (1, 20, 1, 0), (2, 20, 1, 0)]:
        a = splitmix_fill(rows * c * dim, SEED + 2).reshape(rows, c * dim)
        tree = np.zeros((2 * rows - 1) * 4, dtype=np.uint64)
        if batch:
            L.zkref_merkletree_batch(vp(tree), vp(a), u(c), u(rows), u(batch), u(dim))
        else:
            L.zkref_merkletree(vp(tree), vp(a), u(c), u(rows), u(dim))
        L.zkref_canon(vp(tree), u(tree.size))
        out["cases"].append({"op": "merkletree_batch" if batch else "merkletree", "rows": rows, "ncols": c, "dim": dim,
                             "batch": batch, "seed_offset": 2, "tree_sha256": digest(tree),
                             "root": ["%016x" % int(x) for x in tree[-4:]]})
    # non-canonical inputs are legal words for the API: results must equal the reference's canonicalised output
    n, c = 1 << 6, 5
    a = splitmix_fill(n * c, SEED + 3).reshape(n, c)
    a[0, :] = np.uint64(P); a[1, :] = np.uint64(P + 1); a[2, :] = np.uint64(2**64 - 1); a[3, :] = np.uint64(P - 1)
    o = np.zeros((2 * n, c), dtype=np.uint64); buf = np.zeros_like(o)
    L.zkref_extend_pol(vp(o), vp(a.copy()), u(2 * n), u(n), u(c), vp(buf))
    L.zkref_canon(vp(o), u(o.size))
    tree = np.zeros((2 * n - 1) * 4, dtype=np.uint64)
    L.zkref_merkletree(vp(tree), vp(a.copy()), u(c), u(n), u(1))
    L.zkref_canon(vp(tree), u(tree.size))
    out["cases"].append({"op": "noncanonical", "n": n, "ncols": c, "seed_offset": 3,
                         "rows_override": {"0": "p", "1": "p+1", "2": "2^64-1", "3": "p-1"},
                         "ext_sha256": digest(o), "tree_of_input_sha256": digest(tree)})
    # one FRI folding step (friProve.cpp:89-106) through the reference's NTT_Goldilocks + Polinomial element ops
    # (oracle/ref_wrap.cpp zkref_fri_fold); shift_inv and special_x come from the same generator
    for (pol_bits, next_bits) in [(10, 6), (12, 8), (8, 3), (9, 9), (6, 0), (16, 12)]:
        a = splitmix_fill((1 << pol_bits) * 3, SEED + 4)
        aux = splitmix_fill(4, SEED + 5)
        o = np.zeros((1 << next_bits) * 3, dtype=np.uint64)
        L.zkref_fri_fold(vp(o), vp(a.copy()), u(pol_bits), u(next_bits), u(int(aux[3])), vp(aux[:3].copy()))
        L.zkref_canon(vp(o), u(o.size))
        out["cases"].append({"op": "fri_fold", "pol_bits": pol_bits, "next_bits": next_bits, "seed_offset": 4,
                             "shift_inv_and_x": "splitmix(seed+5)[3], splitmix(seed+5)[0:3]", "sha256": digest(o), "samples": samples(o)})
    # stage 4 of Starks::genProof (starks.cpp:261-285) through the reference's NTT_Goldilocks + Goldilocks3::mul
    # (oracle/ref_wrap.cpp zkref_stage4)
    for (log_n, log_b, q_dim, q_deg) in [(10, 1, 3, 2), (8, 2, 3, 3), (12, 1, 3, 1), (6, 3, 3, 8), (16, 1, 3, 2)]:
        n, n_ext = 1 << log_n, 1 << (log_n + log_b)
        q = splitmix_fill(n_ext * q_dim, SEED + 6)
        o = np.zeros(n_ext * q_dim * q_deg, dtype=np.uint64)
        L.zkref_stage4(vp(o), vp(q.copy()), u(n_ext), u(n), u(q_dim), u(q_deg))
        L.zkref_canon(vp(o), u(o.size))
        tree = np.zeros((2 * n_ext - 1) * 4, dtype=np.uint64)
        L.zkref_merkletree(vp(tree), vp(o), u(q_dim * q_deg), u(n_ext), u(1))
        L.zkref_canon(vp(tree), u(tree.size))
        out["cases"].append({"op": "stage4", "log_n": log_n, "log_blowup": log_b, "q_dim": q_dim, "q_deg": q_deg, "seed_offset": 6,
                             "sha256": digest(o), "samples": samples(o), "root": ["%016x" % int(x) for x in tree[-4:]]})
    return out


if __name__ == "__main__":
    k = kats()
    (HERE / "reference_kats.json").write_text(json.dumps(k, indent=1) + "\n")
    d = diffs()
    (HERE / "reference_diff.json").write_text(json.dumps(d, indent=1) + "\n")
    print("kats:", list(k), "\ndiff cases:", len(d["cases"]))
