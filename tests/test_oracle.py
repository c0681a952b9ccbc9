"""Pins the CPU oracle (oracle/gl_oracle.c): against every known-answer vector the reference's own
tests hold for the hot path (tests/golden/reference_kats.json), against the committed outputs of the
compiled reference on seeded inputs (tests/golden/reference_diff.json) and, when the compiled
reference is present (oracle/_ref), word for word against it.  CPU only."""
import numpy as np
import pytest
from oracle_lib import (Oracle, Reference, P, SEED, digest, load_golden, hexwords, field_fib, merkle_kat_matrix)

O = Oracle()
R = Reference()
KATS = load_golden("reference_kats.json")
DIFF = load_golden("reference_diff.json")


def test_field_ops_edge_values():
    """goldilocks_base_field_scalar.hpp:16-160 on canonical and non-canonical words (class mod p)."""
    vals = [0, 1, 2, 0xFFFFFFFF, 0x100000000, P - 1, P, P + 1, 2**64 - 1, 2**63, 0xFFFFFFFF00000000, 7, 2**32 - 2]
    for a in vals:
        for b in vals:
            assert O.L.oracle_add(a, b) % P == (a + b) % P
            assert O.L.oracle_sub(a, b) % P == (a - b) % P
            assert O.L.oracle_mul(a, b) % P == (a * b) % P
            if R:
                assert O.L.oracle_add(a, b) == R.L.zkref_add(a, b)
                assert O.L.oracle_sub(a, b) == R.L.zkref_sub(a, b)
                assert O.L.oracle_mul(a, b) % P == R.L.zkref_mul(a, b) % P


def test_roots_of_unity_table():
    """goldilocks_base_field.cpp:11-44: W[1]=p-1, W[2]=2^48, W[3]=2^24, ... W[6]=8, W[k]^2=W[k-1]."""
    assert O.L.oracle_w(0) == 1 and O.L.oracle_w(1) == P - 1
    assert O.L.oracle_w(2) == 2**48 and O.L.oracle_w(3) == 2**24 and O.L.oracle_w(4) == 4096
    assert O.L.oracle_w(5) == 64 and O.L.oracle_w(6) == 8 and O.L.oracle_w(7) == 2198989700608
    assert O.L.oracle_w(32) == 7277203076849721926
    for k in range(1, 33):
        assert pow(O.L.oracle_w(k), 2, P) == O.L.oracle_w(k - 1)


def test_poseidon_kats():
    fib = field_fib(12, 0, 1)
    assert (O.hash_full_result(fib) == hexwords(KATS["poseidon_fibonacci"]["expected"])).all()
    assert (O.hash_full_result(np.zeros(12, dtype=np.uint64)) == hexwords(KATS["poseidon_zero"]["expected"])).all()


def test_linear_hash_kat_and_short_inputs():
    assert (O.linear_hash(field_fib(128, 0, 1)) == hexwords(KATS["linear_hash_fib128"]["expected"])).all()
    # size <= 4: output = input zero padded, not hashed (poseidon_goldilocks.cpp:234-239)
    for n in range(0, 5):
        x = np.arange(10, 10 + n, dtype=np.uint64)
        assert (O.linear_hash(x) == np.concatenate([x, np.zeros(4 - n, dtype=np.uint64)])).all()


def test_merkle_root_kats():
    m = merkle_kat_matrix(64, 128)
    assert (O.merkletree(m, 128, 64)[-4:] == hexwords(KATS["merkle_root_64x128"]["expected"])).all()
    assert (O.merkletree(m, 128, 64, batch=32)[-4:] == hexwords(KATS["merkle_batch_root_64x128_b32"]["expected"])).all()
    e = np.zeros(1, dtype=np.uint64)
    assert (O.merkletree(e, 0, 64)[-4:] == hexwords(KATS["merkle_root_ncols0_64"]["expected"])).all()
    assert (O.merkletree(e, 0, 1 << 17)[-4:] == hexwords(KATS["merkle_root_ncols0_131072"]["expected"])).all()


def test_ntt_roundtrip_and_sizes():
    """tests.cpp:2225-2384: NTT then INTT is the identity; size 1 and 2 work."""
    for n, c in [(16, 1), (16, 8), (1, 3), (2, 3), (64, 5)]:
        a = O.fill(n * c).reshape(n, c)
        assert (O.intt(O.ntt(a)) == a).all()


def test_lde_kats():
    # LDE with shift 49 (tests.cpp:2385-2461): INTT, scale by 49^i, zero pad, NTT 32
    a = field_fib(16, 1, 1).reshape(16, 1)
    c = O.intt(a)
    r = np.array([pow(49, i, P) for i in range(16)], dtype=object)
    c = np.array([(int(c[i, 0]) * int(r[i])) % P for i in range(16)] + [0] * 16, dtype=np.uint64).reshape(32, 1)
    assert (O.ntt(c)[:, 0] == hexwords(KATS["lde_shift49_16to32"]["expected"])).all()
    # extendPol 16x8 -> 32x8 (tests.cpp:2554-2611)
    m = np.stack([field_fib(16, 1 + j, 1 + j) for j in range(8)], axis=1)
    e = O.extend_pol(m, 32)
    assert (e[:, 0] == hexwords(KATS["extendpol_16x8_col0"]["expected"])).all()


def _gen(n, off):
    return O.fill(n, SEED, off)


def test_against_committed_reference_outputs():
    """reference_diff.json was produced by the compiled reference (tests/golden/make_golden.py)."""
    for case in DIFF["cases"]:
        op = case["op"]
        if op == "extend_and_merkelize":
            n, c, b = case["n"], case["ncols"], case["blowup"]
            ext = O.extend_pol(_gen(n * c, 0).reshape(n, c), n * b)
            assert digest(ext) == case["ext_sha256"], case
            for i, w in case["ext_samples"].items():
                assert int(ext.reshape(-1)[int(i)]) == int(w, 16)
            tree = O.merkletree(ext, c, n * b)
            assert digest(tree) == case["tree_sha256"], case
            assert (tree[-4:] == hexwords(case["root"])).all()
        elif op in ("ntt", "intt"):
            n, c = case["n"], case["ncols"]
            a = _gen(n * c, 1).reshape(n, c)
            o = O.intt(a) if op == "intt" else O.ntt(a)
            assert digest(o) == case["sha256"], case
        elif op in ("merkletree", "merkletree_batch"):
            rows, c, dim, batch = case["rows"], case["ncols"], case["dim"], case["batch"]
            a = _gen(rows * c * dim, 2).reshape(rows, c * dim)
            tree = O.merkletree(a, c, rows, dim=dim, batch=batch)
            assert digest(tree) == case["tree_sha256"], case
        elif op == "noncanonical":
            n, c = case["n"], case["ncols"]
            a = _gen(n * c, 3).reshape(n, c)
            a[0, :] = P; a[1, :] = P + 1; a[2, :] = 2**64 - 1; a[3, :] = P - 1
            assert digest(O.extend_pol(a, 2 * n)) == case["ext_sha256"]
            assert digest(O.merkletree(a, c, n)) == case["tree_of_input_sha256"]
        elif op == "fri_fold":
            pb, nb = case["pol_bits"], case["next_bits"]
            a, aux = _gen((1 << pb) * 3, 4), _gen(4, 5)
            o = O.fri_fold(a, pb, nb, int(aux[3]), aux[:3])
            assert digest(o) == case["sha256"], case
            for i, w in case["samples"].items():
                assert int(o.reshape(-1)[int(i)]) == int(w, 16)
        elif op == "stage4":
            n, n_ext = 1 << case["log_n"], 1 << (case["log_n"] + case["log_blowup"])
            q = _gen(n_ext * case["q_dim"], 6)
            o = O.stage4(q, n_ext, n, case["q_dim"], case["q_deg"])
            assert digest(o) == case["sha256"], case
            assert [("%016x" % int(x)) for x in O.merkletree(o, case["q_dim"] * case["q_deg"], n_ext)[-4:]] == case["root"]
        else:
            raise AssertionError(op)


@pytest.mark.skipif(not R, reason="compiled reference (oracle/_ref/libzkref.so) not present")
def test_against_live_reference():
    rng = np.random.default_rng(7)
    for n, c, b in [(1 << 6, 3, 2), (1 << 11, 9, 2), (1 << 5, 12, 16), (1 << 8, 17, 8)]:
        a = (rng.integers(0, 2**63, size=(n, c), dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=(n, c), dtype=np.uint64))
        assert (O.extend_pol(a, n * b) == R.extend_pol(a, n * b)).all()
        assert (O.ntt(a) == R.ntt(a)).all()
        assert (O.intt(a) == R.ntt(a, inverse=True)).all()
        assert (O.intt(a, extend=True) == R.ntt(a, inverse=True, extend=True)).all()
    for _ in range(20):
        st = rng.integers(0, 2**63, size=12, dtype=np.uint64) * np.uint64(2) + np.uint64(1)
        assert (O.hash_full_result(st) == R.hash_full_result(st)).all()
        assert (O.hash_full_result(st) == R.hash_full_result(st, seq=True)).all()
    for n in [0, 1, 4, 5, 8, 9, 16, 17, 665, 84 * 8]:
        x = rng.integers(0, 2**63, size=n, dtype=np.uint64)
        assert (O.linear_hash(x) == R.linear_hash(x)).all()
    for rows, c, dim, batch in [(64, 13, 1, 0), (32, 5, 3, 0), (128, 3, 1, 0), (16, 50, 1, 16), (8, 0, 1, 0)]:
        a = rng.integers(0, 2**63, size=(rows, max(c * dim, 1)), dtype=np.uint64)
        assert (O.merkletree(a, c, rows, dim, batch) == R.merkletree(a, c, rows, dim, batch)).all()
    for ln, lb, qd in [(5, 1, 2), (7, 2, 3), (4, 3, 5), (9, 1, 1)]:
        q = rng.integers(0, 2**63, size=(1 << (ln + lb)) * 3, dtype=np.uint64) * np.uint64(2) + np.uint64(1)
        assert (O.stage4(q, 1 << (ln + lb), 1 << ln, 3, qd) == R.stage4(q, 1 << (ln + lb), 1 << ln, 3, qd)).all()
    for size, ncols_all, off, nc, ext in ((64, 7, 2, 3, 1), (128, 5, 0, 5, 4)):
        src = rng.integers(0, P, size=(size, ncols_all), dtype=np.uint64)
        got = R.reverse_permutation(src, size, off, nc, ncols_all, ext)
        for i in range(size):
            r = int(format(i, "0%db" % (size.bit_length() - 1))[::-1], 2)
            want = src[r, off:off + nc] if (ext == 1 or r * ncols_all + off < (size // ext) * ncols_all) else np.zeros(nc, dtype=np.uint64)
            assert (got[i] == want).all()
    for pb, nb in [(4, 2), (8, 4), (10, 5), (5, 5), (7, 0), (12, 8)]:
        a = rng.integers(0, 2**63, size=((1 << pb), 3), dtype=np.uint64) * np.uint64(2) + np.uint64(1)  # includes words >= p
        x = rng.integers(0, P, size=3, dtype=np.uint64)
        sh = int(rng.integers(1, P, dtype=np.uint64))
        assert (O.fri_fold(a, pb, nb, sh, x) == R.fri_fold(a, pb, nb, sh, x)).all()
