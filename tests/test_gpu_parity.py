"""GPU parity tests: libzkgpu.so (CUDA, through the C ABI / the Python mirror of the reference classes) against the
CPU oracle (oracle/gl_oracle.c), the reference's own known-answer vectors (tests/golden/reference_kats.json) and the
committed outputs of the compiled reference (tests/golden/reference_diff.json).  Bit-exact on canonical words.
Mirrors src/goldilocks/tests/tests.cpp: poseidon_full_avx (:1627), linear_hash_avx (:1767), merkletree_avx (:1883),
ntt / ntt_block (:2225-2384), LDE (:2385), LDE_block (:2467), extendePol (:2554)."""
import numpy as np
import pytest

import zkevm_prover_b200 as zk
from oracle_lib import Oracle, P, SEED, digest, load_golden, hexwords, field_fib, merkle_kat_matrix

pytestmark = pytest.mark.gpu

O = Oracle()
KATS = load_golden("reference_kats.json")
DIFF = load_golden("reference_diff.json")


def u64(x):
    return np.ascontiguousarray(x, dtype=np.uint64)


def first_diff(a, b):
    a, b = np.asarray(a).reshape(-1), np.asarray(b).reshape(-1)
    idx = np.nonzero(a != b)[0]
    if idx.size == 0:
        return "equal"
    i = int(idx[0])
    return "%d of %d words differ; first at %d: got %016x want %016x" % (idx.size, a.size, i, int(a[i]), int(b[i]))


def addmod(a, b):
    """(a + b) mod p for canonical uint64 arrays."""
    s = a + b
    s = np.where(s < a, s + np.uint64(0xFFFFFFFF), s)
    return np.where(s >= np.uint64(P), s - np.uint64(P), s)


def assert_eq(got, want, what=""):
    if not np.array_equal(np.asarray(got).reshape(-1), np.asarray(want).reshape(-1)):
        pytest.fail(what + " " + first_diff(got, want), pytrace=False)


def gpu_ntt(a, inverse=False, extend=False):
    a = u64(a)
    n, c = a.shape
    ntt = zk.NTT_Goldilocks(n)
    out = np.zeros_like(a)
    ntt.NTT(out, a, n, c, None, 3, 1, inverse, extend)
    return out


def gpu_extend(a, n_ext):
    a = u64(a)
    n, c = a.shape
    ntt = zk.NTT_Goldilocks(n)
    out = np.zeros((n_ext, c), dtype=np.uint64)
    ntt.extendPol(out, a, n_ext, n, c)
    return out


def gpu_tree(a, ncols, nrows, dim=1, batch=0):
    tree = np.zeros((2 * nrows - 1) * 4, dtype=np.uint64)
    a = u64(a)
    if batch:
        zk.PoseidonGoldilocks.merkletree_batch(tree, a, ncols, nrows, batch, 0, dim)
    else:
        zk.PoseidonGoldilocks.merkletree(tree, a, ncols, nrows, 0, dim)
    return tree


# ------------------------------------------------------------------------------------------------ field
def test_field_ops_edges_and_random():
    """goldilocks_base_field_scalar.hpp:16-160 semantics (value class mod p) incl. non-canonical operands."""
    edge = [0, 1, 2, 7, 0xFFFFFFFF, 0x100000000, 0x100000001, 2**32 - 2, P - 1, P, P + 1, P - 2, 2**63, 2**63 - 1,
            0xFFFFFFFF00000000, 0xFFFFFFFEFFFFFFFF, 2**64 - 1, 2**64 - 2, 0x00000001FFFFFFFF, 0x7FFFFFFF80000001]
    A = [a for a in edge for _ in edge]
    B = [b for _ in edge for b in edge]
    rng = np.random.default_rng(3)
    ra = rng.integers(0, 2**63, size=200000, dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=200000, dtype=np.uint64)
    rb = rng.integers(0, 2**63, size=200000, dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=200000, dtype=np.uint64)
    # bias some operands to the top of the range where the carry fix-ups trigger
    ra[:50000] |= np.uint64(0xFFFFFFFF00000000)
    rb[25000:75000] |= np.uint64(0xFFFFFFFF00000000)
    a = np.concatenate([np.array(A, dtype=np.uint64), ra])
    b = np.concatenate([np.array(B, dtype=np.uint64), rb])
    ai = [int(x) for x in a]
    bi = [int(x) for x in b]
    for op, f in (("add", lambda x, y: (x + y) % P), ("sub", lambda x, y: (x - y) % P), ("mul", lambda x, y: (x * y) % P),
                  ("mul_nc", lambda x, y: (x * y) % P), ("canon", lambda x, y: x % P)):
        got = zk.field_op(op, a, b)
        want = np.array([f(x, y) for x, y in zip(ai, bi)], dtype=np.uint64)
        assert_eq(got, want, op)


# ------------------------------------------------------------------------------------------------ poseidon
def test_poseidon_kats():
    out = np.zeros(12, dtype=np.uint64)
    zk.PoseidonGoldilocks.hash_full_result(out, field_fib(12, 0, 1))
    assert_eq(out, hexwords(KATS["poseidon_fibonacci"]["expected"]), "fibonacci")
    zk.PoseidonGoldilocks.hash_full_result(out, np.zeros(12, dtype=np.uint64))
    assert_eq(out, hexwords(KATS["poseidon_zero"]["expected"]), "zero")


def test_poseidon_random_and_noncanonical_states():
    rng = np.random.default_rng(11)
    n = 4096
    st = rng.integers(0, 2**63, size=(n, 12), dtype=np.uint64) * np.uint64(2) + np.uint64(1)
    st[0, :] = np.uint64(2**64 - 1)
    st[1, :] = np.uint64(P)
    st[2, :] = np.uint64(P - 1)
    st[3, :] = 0
    out = np.zeros_like(st)
    zk.PoseidonGoldilocks.hash_full_result(out, st, n)
    want = np.stack([O.hash_full_result(st[i]) for i in range(n)])
    assert_eq(out, want)


def test_linear_hash_kat_and_lengths():
    out = np.zeros(4, dtype=np.uint64)
    zk.PoseidonGoldilocks.linear_hash(out, field_fib(128, 0, 1), 128)
    assert_eq(out, hexwords(KATS["linear_hash_fib128"]["expected"]))
    rng = np.random.default_rng(5)
    for n in [0, 1, 2, 3, 4, 5, 7, 8, 9, 15, 16, 17, 84 * 8, 665, 751, 1000]:
        x = rng.integers(0, 2**63, size=max(n, 1), dtype=np.uint64) * np.uint64(2) + np.uint64(1)
        zk.PoseidonGoldilocks.linear_hash(out, x, n)
        assert_eq(out, O.linear_hash(x[:n]), "n=%d" % n)


def test_merkle_root_kats():
    m = merkle_kat_matrix(64, 128)
    t = gpu_tree(m, 128, 64)
    assert_eq(t[-4:], hexwords(KATS["merkle_root_64x128"]["expected"]))
    assert_eq(t, O.merkletree(m, 128, 64), "all nodes")
    tb = gpu_tree(m, 128, 64, batch=32)
    assert_eq(tb[-4:], hexwords(KATS["merkle_batch_root_64x128_b32"]["expected"]))
    e = np.zeros(1, dtype=np.uint64)
    assert_eq(gpu_tree(e, 0, 64)[-4:], hexwords(KATS["merkle_root_ncols0_64"]["expected"]))
    assert_eq(gpu_tree(e, 0, 1 << 17)[-4:], hexwords(KATS["merkle_root_ncols0_131072"]["expected"]))


def test_merkletree_shapes_vs_oracle():
    rng = np.random.default_rng(9)
    for rows, c, dim, batch in [(1, 8, 1, 0), (2, 3, 1, 0), (64, 13, 1, 0), (32, 5, 3, 0), (128, 3, 1, 0), (128, 4, 1, 0),
                                (16, 50, 1, 16), (8, 0, 1, 0), (256, 665, 1, 0), (64, 12, 3, 5), (1024, 9, 1, 0)]:
        a = rng.integers(0, 2**63, size=(rows, max(c * dim, 1)), dtype=np.uint64) * np.uint64(2)
        assert_eq(gpu_tree(a, c, rows, dim, batch), O.merkletree(a, c, rows, dim, batch), str((rows, c, dim, batch)))


def test_merkle_tree_class_and_group_proof():
    """MerkleTreeGL (merkleTreeGL.cpp:12-44): merkelize, getRoot, getGroupProof verify against the root."""
    h, w = 256, 20
    src = O.fill(h * w, SEED, 5)
    mt = zk.MerkleTreeGL(h, w, src)
    mt.merkelize()
    assert_eq(mt.nodes, O.merkletree(src, w, h))
    for idx in (0, 1, 77, 255):
        proof = mt.getGroupProof(idx)
        assert proof.size == w + 4 * mt.MerkleProofSize()
        node = O.linear_hash(proof[:w])
        i = idx
        for lvl in range(mt.MerkleProofSize()):
            sib = proof[w + 4 * lvl: w + 4 * lvl + 4]
            pair = [node, sib] if i % 2 == 0 else [sib, node]
            st = np.concatenate(pair + [np.zeros(4, dtype=np.uint64)])
            node = O.hash_full_result(st)[:4]
            i >>= 1
        assert_eq(node, mt.getRoot(), "proof idx %d" % idx)


# ------------------------------------------------------------------------------------------------ ntt
def test_ntt_all_pass_sizes_vs_oracle():
    """forward / inverse / inverse+extend for every log size 0..13 (every in-tile radix split and 1-2 pass plans)."""
    for lg in range(0, 14):
        n = 1 << lg
        for c in (1, 3, 17):
            a = O.fill(n * c, SEED, lg * 100 + c).reshape(n, c)
            assert_eq(gpu_ntt(a), O.ntt(a), "ntt lg=%d c=%d" % (lg, c))
            assert_eq(gpu_ntt(a, inverse=True), O.intt(a), "intt lg=%d c=%d" % (lg, c))
            assert_eq(gpu_ntt(a, inverse=True, extend=True), O.intt(a, extend=True), "intt-extend lg=%d c=%d" % (lg, c))


def test_ntt_three_pass_plans_vs_oracle():
    for lg, c in [(17, 5), (18, 2), (20, 1)]:
        n = 1 << lg
        a = O.fill(n * c, SEED, lg).reshape(n, c)
        assert_eq(gpu_ntt(a), O.ntt(a), "ntt lg=%d" % lg)
        assert_eq(gpu_ntt(a, inverse=True), O.intt(a), "intt lg=%d" % lg)


def test_ntt_roundtrip_reference_cases():
    """tests.cpp:2225-2384: NTT then INTT is the identity; dst NULL / dst == src / dst != src; size 0, ncols 0 no-ops."""
    ntt = zk.NTT_Goldilocks(1 << 12)
    for n, c in [(16, 1), (16, 8), (1, 3), (2, 3), (1 << 12, 7)]:
        a = O.fill(n * c, SEED, 77).reshape(n, c)
        b = a.copy()
        ntt.NTT(None, b, n, c)          # in place (dst NULL)
        ntt.INTT(b, b, n, c)            # dst == src
        assert_eq(b, a, "in-place roundtrip %s" % ((n, c),))
        d = np.zeros_like(a)
        ntt.NTT(d, a, n, c, None, 4, 3)  # nphase / nblock accepted
        e = np.zeros_like(a)
        ntt.INTT(e, d, n, c, None, 3, 5)
        assert_eq(e, a, "out-of-place roundtrip")
    z = np.zeros(4, dtype=np.uint64)
    ntt.NTT(z, z, 0, 1)
    ntt.NTT(z, z, 4, 0)
    ntt.extendPol(z, z, 8, 4, 0)
    with pytest.raises(zk.ZkGpuError):
        ntt.NTT(z, z, 3, 1)
    with pytest.raises(OverflowError):
        zk.NTT_Goldilocks(1 << 33)


def test_lde_kats():
    # LDE with shift 49 (tests.cpp:2385-2461): INTT, scale by 49^i, zero pad, NTT 32
    a = field_fib(16, 1, 1).reshape(16, 1)
    c = gpu_ntt(a, inverse=True)
    c = np.array([(int(c[i, 0]) * pow(49, i, P)) % P for i in range(16)] + [0] * 16, dtype=np.uint64).reshape(32, 1)
    assert_eq(gpu_ntt(c)[:, 0], hexwords(KATS["lde_shift49_16to32"]["expected"]), "LDE shift 49")
    # extendPol 16x8 -> 32x8 (tests.cpp:2554-2611), out of place and in place (a, a)
    m = np.stack([field_fib(16, 1 + j, 1 + j) for j in range(8)], axis=1)
    e = gpu_extend(m, 32)
    assert_eq(e[:, 0], hexwords(KATS["extendpol_16x8_col0"]["expected"]), "extendPol KAT")
    buf = np.zeros((32, 8), dtype=np.uint64)
    buf[:16] = m
    zk.NTT_Goldilocks(16).extendPol(buf, buf, 32, 16, 8)
    assert_eq(buf, e, "extendPol in place")


def test_extend_pol_vs_oracle_shapes():
    rng = np.random.default_rng(21)
    for n, c, b in [(1, 3, 2), (2, 1, 2), (64, 3, 2), (1 << 11, 9, 2), (32, 12, 16), (256, 17, 8), (1 << 10, 665, 2), (1 << 9, 5, 1),
                    (1 << 13, 33, 4), (1 << 16, 8, 2)]:
        a = rng.integers(0, 2**63, size=(n, c), dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=(n, c), dtype=np.uint64)
        assert_eq(gpu_extend(a, n * b), O.extend_pol(a, n * b), "extendPol %s" % ((n, c, b),))


def test_committed_reference_outputs():
    """tests/golden/reference_diff.json: outputs of the compiled unmodified reference on seeded inputs."""
    for case in DIFF["cases"]:
        op = case["op"]
        if op == "extend_and_merkelize":
            n, c, b = case["n"], case["ncols"], case["blowup"]
            a = O.fill(n * c, SEED, 0).reshape(n, c)
            ext = np.zeros((n * b, c), dtype=np.uint64)
            tree = np.zeros((2 * n * b - 1) * 4, dtype=np.uint64)
            zk.extend_and_merkelize(zk.NTT_Goldilocks(n), ext, tree, a, n * b, n, c)
            assert digest(ext) == case["ext_sha256"], case
            assert digest(tree) == case["tree_sha256"], case
            assert_eq(tree[-4:], hexwords(case["root"]))
        elif op in ("ntt", "intt"):
            n, c = case["n"], case["ncols"]
            a = O.fill(n * c, SEED, 1).reshape(n, c)
            assert digest(gpu_ntt(a, inverse=(op == "intt"))) == case["sha256"], case
        elif op in ("merkletree", "merkletree_batch"):
            rows, c, dim, batch = case["rows"], case["ncols"], case["dim"], case["batch"]
            a = O.fill(rows * c * dim, SEED, 2).reshape(rows, c * dim)
            assert digest(gpu_tree(a, c, rows, dim, batch)) == case["tree_sha256"], case
        elif op == "noncanonical":
            n, c = case["n"], case["ncols"]
            a = O.fill(n * c, SEED, 3).reshape(n, c)
            a[0, :] = P
            a[1, :] = P + 1
            a[2, :] = 2**64 - 1
            a[3, :] = P - 1
            assert digest(gpu_extend(a, 2 * n)) == case["ext_sha256"]
            assert digest(gpu_tree(a, c, n)) == case["tree_of_input_sha256"]


# ------------------------------------------------------------------------------------------------ device-resident path
def _dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).cuda()


def _host(t):
    return t.cpu().numpy().view(np.uint64)


def test_device_pointer_commit_matches_oracle():
    """config #1 of BASELINE.json (2^16 x 8, blowup 2) through the fused call with DEVICE pointers, plus the on-device
    generator: extended trace, every tree node and the root equal the oracle."""
    import torch
    n, c, b = 1 << 16, 8, 2
    d_in = torch.empty(n * c, dtype=torch.int64, device="cuda")
    zk.fill(d_in, n * c, SEED, 0)
    zk.sync()
    a = O.fill(n * c, SEED, 0)
    assert_eq(_host(d_in), a, "generator")
    d_ext = torch.empty(n * b * c, dtype=torch.int64, device="cuda")
    d_tree = torch.empty((2 * n * b - 1) * 4, dtype=torch.int64, device="cuda")
    zk.extend_and_merkelize(zk.NTT_Goldilocks(n), d_ext, d_tree, d_in, n * b, n, c)
    ms = zk.stage_ms()
    assert ms["lde"] > 0 and ms["leaves"] > 0
    ext = O.extend_pol(a.reshape(n, c), n * b)
    assert_eq(_host(d_ext), ext, "extended trace")
    assert_eq(_host(d_tree), O.merkletree(ext, c, n * b), "tree")


def test_narrow_high_blowup_shape():
    """BASELINE.json configs[3] at its stated size: 2^17 -> 2^21 rows x 12 columns, blowup 16 (16 coset transforms)."""
    n, c, b = 1 << 17, 12, 16
    a = O.fill(n * c, SEED, 4).reshape(n, c)
    ext = np.zeros((n * b, c), dtype=np.uint64)
    tree = np.zeros((2 * n * b - 1) * 4, dtype=np.uint64)
    zk.extend_and_merkelize(zk.NTT_Goldilocks(n), ext, tree, a, n * b, n, c)
    want = O.extend_pol(a, n * b)
    assert_eq(ext, want, "ext")
    assert_eq(tree, O.merkletree(want, c, n * b), "tree")


def test_strip_path_and_properties_large():
    """A shape wide/tall enough to take several column strips when the scratch budget is small is covered by forcing
    strided shards: LDE of a column shard with ld > ncols equals the same columns of the full LDE (columns are
    independent), and linearity LDE(a + b) = LDE(a) + LDE(b) holds at 2^18 x 24."""
    import torch
    n, c, b = 1 << 18, 24, 2
    x = O.fill(n * c, SEED, 8)
    y = O.fill(n * c, SEED, 9)
    s = addmod(x, y)
    ex, ey, es = gpu_extend(x.reshape(n, c), n * b), gpu_extend(y.reshape(n, c), n * b), gpu_extend(s.reshape(n, c), n * b)
    want = addmod(ex, ey)
    assert_eq(es, want, "linearity")
    # column shard with strides
    d_in = _dev(x)
    d_out = torch.zeros(n * b * 7, dtype=torch.int64, device="cuda")
    rc = zk.lib().zkgpu_lde_columns(d_out.data_ptr(), 7, d_in.data_ptr() + 8 * 5, c, n * b, n, 7)
    assert rc == 0, zk.lib().zkgpu_last_error()
    zk.sync()
    assert_eq(_host(d_out).reshape(n * b, 7), ex[:, 5:12], "column shard")
    # spot-check against the oracle on 3 columns
    assert_eq(ex[:, :3], O.extend_pol(np.ascontiguousarray(x.reshape(n, c)[:, :3]), n * b), "oracle spot check")


def test_reverse_permutation_matches_reference_semantics():
    """NTT_Goldilocks::reversePermutation (ntt_goldilocks.cpp:243-305): bit-reversed row gather of a column block,
    zero fill beyond size/extension, and the in-place swap form."""
    def br(i, bits):
        return int(format(i, "0%db" % bits)[::-1], 2) if bits else 0
    size, ncols_all, off, nc = 64, 7, 2, 3
    src = O.fill(size * ncols_all, SEED, 11).reshape(size, ncols_all)
    for ext in (1, 4):
        dst = np.full((size, nc), 123, dtype=np.uint64)
        zk.NTT_Goldilocks(size, 0, ext).reversePermutation(dst, src, size, off, nc, ncols_all)
        want = np.zeros_like(dst)
        for i in range(size):
            r = br(i, 6)
            if ext == 1 or r * ncols_all + off < (size // ext) * ncols_all:
                want[i] = src[r, off:off + nc]
        assert_eq(dst, want, "reversePermutation ext=%d" % ext)
    a = src.copy()
    zk.NTT_Goldilocks(size).reversePermutation(a, a, size, 0, ncols_all, ncols_all)
    assert_eq(a, src[[br(i, 6) for i in range(size)]], "in place")
    with pytest.raises(zk.ZkGpuError):
        zk.NTT_Goldilocks(size, 0, 2).reversePermutation(a, a, size, 0, ncols_all, ncols_all)
    # and against the compiled unmodified reference (oracle/_ref, zkref_reverse_permutation) where it is present
    from oracle_lib import Reference
    R = Reference()
    if R:
        for size, ncols_all, off, nc, ext in ((1 << 10, 9, 0, 9, 1), (1 << 12, 13, 5, 4, 1), (1 << 11, 6, 1, 3, 2), (1 << 9, 8, 0, 8, 8)):
            src = O.fill(size * ncols_all, SEED, 12).reshape(size, ncols_all)
            dst = np.full((size, nc), 7, dtype=np.uint64)
            zk.NTT_Goldilocks(size, 0, ext).reversePermutation(dst, src, size, off, nc, ncols_all)
            assert_eq(dst, R.reverse_permutation(src, size, off, nc, ncols_all, ext), "vs reference %r" % ((size, ncols_all, off, nc, ext),))


def test_group_proof_from_device_resident_tree():
    """the extended trace and the tree stay on the device after the commit; MerkleTreeGL wraps them
    and serves getRoot / getGroupProof without copying the trace back."""
    import torch
    n, c, b = 1 << 10, 21, 2
    a = O.fill(n * c, SEED, 12).reshape(n, c)
    d_ext = torch.empty(n * b * c, dtype=torch.int64, device="cuda")
    d_tree = torch.empty((2 * n * b - 1) * 4, dtype=torch.int64, device="cuda")
    zk.extend_and_merkelize(zk.NTT_Goldilocks(n), d_ext, d_tree, a, n * b, n, c)
    ext = O.extend_pol(a, n * b)
    tree = O.merkletree(ext, c, n * b)
    host = zk.MerkleTreeGL(n * b, c, ext.reshape(-1), tree)
    dev = zk.MerkleTreeGL(n * b, c, d_ext, d_tree)
    mixed = zk.MerkleTreeGL(n * b, c, d_ext, tree)
    assert_eq(dev.getRoot(), tree[-4:], "root")
    for idx in (0, 1, 1000, n * b - 1):
        want = host.getGroupProof(idx)
        assert_eq(want[:c], ext[idx], "row")
        assert_eq(dev.getGroupProof(idx), want, "device proof %d" % idx)
        assert_eq(mixed.getGroupProof(idx), want, "mixed proof %d" % idx)
    # the batch form (FRIProve::prove asks every tree for nQueries proofs, friProve.cpp:144-160): one launch, one copy
    idxs = [0, 1, 1000, n * b - 1, 77, 77, 513]
    want = np.stack([host.getGroupProof(i) for i in idxs])
    assert_eq(zk.group_proofs(dev, idxs), want, "batch, device buffers")
    assert_eq(zk.group_proofs(host, idxs), want, "batch, host buffers")
    assert_eq(zk.group_proofs(mixed, idxs), want, "batch, mixed buffers")
    with pytest.raises(zk.ZkGpuError):
        zk.group_proofs(dev, [0, n * b])


# ------------------------------------------------------------------------------------------------ large shapes, host pipelines
def _cpu_commit(a, n_ext, c):
    """Extended trace and tree from the compiled reference when it travelled to this box, else from the C oracle."""
    from oracle_lib import Reference
    R = Reference()
    X = R if R else O
    ext = X.extend_pol(a, n_ext)
    return ext, X.merkletree(ext, c, n_ext)


def test_full_width_665_host_pipeline():
    """SURVEY §8(d) parity case (i): the headline width (665 columns, 83 sponge blocks + 1 ragged word) at 2^16 rows,
    all three pointers on the HOST: exercises the strip pipeline (6 strips of 112 columns, the last ragged) with
    host->device and device->host strip copies.  Every word of the extended trace and of the tree is compared."""
    n, c, b = 1 << 16, 665, 2
    a = O.fill(n * c, SEED, 21).reshape(n, c)
    ext = np.zeros((n * b, c), dtype=np.uint64)
    tree = np.zeros((2 * n * b - 1) * 4, dtype=np.uint64)
    zk.extend_and_merkelize(zk.NTT_Goldilocks(n), ext, tree, a, n * b, n, c)
    want_ext, want_tree = _cpu_commit(a, n * b, c)
    assert_eq(ext, want_ext, "extended trace")
    assert_eq(tree, want_tree, "tree")
    # the same through the two separate reference calls (extendPol, then merkletree) with host pointers
    ext2 = np.zeros_like(ext)
    zk.NTT_Goldilocks(n).extendPol(ext2, a, n * b, n, c)
    assert_eq(ext2, want_ext, "extendPol host strips")
    assert_eq(gpu_tree(ext2, c, n * b), want_tree, "merkletree host")


def test_full_height_2p23_rows():
    """SURVEY §8(d) parity case (ii): the headline height (2^23 -> 2^24 rows, three 8/8/7-bit NTT passes, 24-level tree)
    at 24 columns, host trace in (2 strips), extended trace left on the device."""
    import torch
    n, c, b = 1 << 23, 24, 2
    a = O.fill(n * c, SEED, 22).reshape(n, c)
    d_ext = torch.empty(n * b * c, dtype=torch.int64, device="cuda")
    tree = np.zeros((2 * n * b - 1) * 4, dtype=np.uint64)
    zk.extend_and_merkelize(zk.NTT_Goldilocks(n), d_ext, tree, a, n * b, n, c)
    want_ext, want_tree = _cpu_commit(a, n * b, c)
    assert_eq(_host(d_ext), want_ext, "extended trace")
    assert_eq(tree, want_tree, "tree")


def test_merkletree_host_row_chunks():
    """PoseidonGoldilocks::merkletree with a host source larger than one staging chunk (2^19 x 300 words = 1.26 GB ->
    two row chunks copied while the previous one is hashed)."""
    from oracle_lib import Reference
    rows, c = 1 << 19, 300
    a = O.fill(rows * c, SEED, 23)
    R = Reference()
    want = (R if R else O).merkletree(a, c, rows)
    assert_eq(gpu_tree(a, c, rows), want, "tree")


def test_build_const_tree_image():
    """buildConstTree's Goldilocks branch (build_const_tree.cpp:533-596): the file image
    [nPols, nExt, extended constant pols, nodes] equals header + oracle LDE + oracle tree; MerkleTreeGL views it."""
    nbits, nbits_ext, npols = 10, 12, 19
    n, n_ext = 1 << nbits, 1 << nbits_ext
    a = O.fill(n * npols, SEED, 31).reshape(n, npols)
    words = 2 + npols * n_ext + (2 * n_ext - 1) * 4
    img = np.zeros(words, dtype=np.uint64)
    zk.build_const_tree(img, a, nbits, nbits_ext, npols)
    ext = O.extend_pol(a, n_ext)
    assert int(img[0]) == npols and int(img[1]) == n_ext
    assert_eq(img[2:2 + npols * n_ext], ext, "extended constant pols")
    assert_eq(img[2 + npols * n_ext:], O.merkletree(ext, npols, n_ext), "tree")
    import torch
    d_img = torch.zeros(words, dtype=torch.int64, device="cuda")
    zk.build_const_tree(d_img, _dev(a), nbits, nbits_ext, npols)
    zk.sync()
    assert_eq(_host(d_img), img, "device image")


def test_adversarial_value_sets():
    """all-zero, all p-1, rows of non-canonical words (p, p+1, 2^64-1), single impulse, widths 1..9."""
    n, b = 1 << 10, 2
    for c in (1, 2, 3, 4, 5, 7, 8, 9):
        cases = {
            "zero": np.zeros((n, c), dtype=np.uint64),
            "p-1": np.full((n, c), P - 1, dtype=np.uint64),
            "noncanonical": np.tile(np.array([P, P + 1, 2**64 - 1, 2**64 - 2**32], dtype=np.uint64), n * c // 4 + 1)[:n * c].reshape(n, c),
            "impulse": np.zeros((n, c), dtype=np.uint64),
        }
        cases["impulse"][n // 3, c // 2] = 1
        for name, a in cases.items():
            ext = np.zeros((n * b, c), dtype=np.uint64)
            tree = np.zeros((2 * n * b - 1) * 4, dtype=np.uint64)
            zk.extend_and_merkelize(zk.NTT_Goldilocks(n), ext, tree, a, n * b, n, c)
            want = O.extend_pol(a, n * b)
            assert_eq(ext, want, "%s c=%d ext" % (name, c))
            assert_eq(tree, O.merkletree(want, c, n * b), "%s c=%d tree" % (name, c))


def test_sharded_building_blocks_absorb_copy():
    """The multi-GPU building blocks on one device (GpuOps, dist_commit.py): the row sponge fed piece by piece, with
    the fused scatter of each piece into the row-major matrix, equals linear_hash of the assembled rows; ragged last
    piece, 8-aligned pieces before it."""
    import torch
    from zkevm_prover_b200.dist_commit import GpuOps
    ops = GpuOps()  # hands torch's current stream to the library: no manual synchronisation below
    rows, widths = 4096, [16, 24, 8, 13]
    ncols = sum(widths)
    a = O.fill(rows * ncols, SEED, 5).reshape(rows, ncols)
    leaves = torch.zeros(rows * 4, dtype=torch.int64, device="cuda")
    full = torch.zeros(rows * ncols, dtype=torch.int64, device="cuda")
    c0 = 0
    for i, w in enumerate(widths):
        piece = torch.from_numpy(np.ascontiguousarray(a[:, c0:c0 + w]).view(np.int64).reshape(-1)).cuda()
        ops.absorb(leaves, piece, 0, w, w, rows, i == 0, full, c0, ncols)
        del piece  # the caching allocator may hand the block out again: only stream order protects it
        c0 += w
    assert_eq(_host(full), a, "scattered rows")
    assert_eq(_host(leaves), O.merkletree(a.reshape(-1), ncols, rows)[:rows * 4], "leaf digests")
    # without the copy the digests are the same
    leaves2 = torch.zeros(rows * 4, dtype=torch.int64, device="cuda")
    c0 = 0
    for i, w in enumerate(widths):
        ops.absorb(leaves2, full, c0, w, ncols, rows, i == 0)
        c0 += w
    assert_eq(_host(leaves2), _host(leaves), "absorb without copy")
    zk.set_stream(None)  # back to the library's own stream for the tests that follow


def test_fri_fold_layer_vs_oracle_and_reference_outputs():
    """one FRI folding step + getTransposed + merkelize of the next layer (friProve.cpp:44-127).
    Committed outputs of the reference's own NTT_Goldilocks + Polinomial arithmetic (reference_diff.json, op fri_fold),
    then larger shapes against the oracle, host and device pointers, non-canonical input words."""
    import torch
    for case in DIFF["cases"]:
        if case["op"] != "fri_fold":
            continue
        pb, nb = case["pol_bits"], case["next_bits"]
        a, aux = O.fill((1 << pb) * 3, SEED, 4), O.fill(4, SEED, 5)
        out = np.zeros((1 << nb) * 3, dtype=np.uint64)
        zk.FRIProve.fold(out, a, pb, nb, int(aux[3]), aux[:3])
        assert digest(out) == case["sha256"], case
    rng = np.random.default_rng(11)
    for pb, nb in [(14, 10), (18, 14), (20, 15), (12, 12), (9, 0), (13, 5)]:
        a = rng.integers(0, 2**63, size=(1 << pb) * 3, dtype=np.uint64) * np.uint64(2) + np.uint64(1)
        x = rng.integers(0, 2**63, size=3, dtype=np.uint64) * np.uint64(2) + np.uint64(1)
        sh = int(rng.integers(1, P, dtype=np.uint64))
        want = O.fri_fold(a, pb, nb, sh, x % np.uint64(P))
        out = np.zeros((1 << nb) * 3, dtype=np.uint64)
        zk.FRIProve.fold(out, a, pb, nb, sh, x)
        assert_eq(out, want, "fri_fold host %d->%d" % (pb, nb))
        d_a = torch.from_numpy(a.view(np.int64)).cuda()
        d_o = torch.zeros((1 << nb) * 3, dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        zk.FRIProve.fold(d_o, d_a, pb, nb, sh, x)
        zk.sync()  # device pointers: the call returns after enqueueing on the library's stream
        assert_eq(_host(d_o), want, "fri_fold device %d->%d" % (pb, nb))
        # next layer's tree: transpose into nGroups x (groupSize*3) and merkelize
        tb = max(nb - 4, 0)
        auxh = np.zeros_like(out)
        zk.FRIProve.getTransposed(auxh, out, 1 << nb, tb)
        assert_eq(auxh, O.fri_transpose(want, tb), "getTransposed")
        n_groups, group = 1 << tb, (1 << nb) >> tb
        tree = zk.MerkleTreeGL(n_groups, group * 3, auxh)
        tree.merkelize()
        assert_eq(tree.nodes, O.merkletree(auxh, group * 3, n_groups), "FRI layer tree")


# ------------------------------------------------------------------------------------------------ pinned at full size
def test_reference_bench_golden_root_2p23x128():
    """The reference's only large-shape golden vector: the root of the 2^23-row x 128-column Fibonacci matrix asserted in
    src/goldilocks/benchs/bench.cpp:406-443 (MERKLETREE_BENCH_AVX) and its batch twin :584-587 (batch_size (128+3)/4 = 32)."""
    import ctypes
    nrows, ncols = 1 << 23, 128
    m = np.empty(nrows * ncols, dtype=np.uint64)
    O.L.oracle_merkle_kat_matrix(m.ctypes.data_as(ctypes.c_void_p), ctypes.c_uint64(nrows), ctypes.c_uint64(ncols))
    small = merkle_kat_matrix(64, ncols)
    assert_eq(m[:64 * ncols], small, "generator")
    d = _dev(m)
    tree = np.zeros((2 * nrows - 1) * 4, dtype=np.uint64)
    import torch
    d_tree = torch.zeros(tree.size, dtype=torch.int64, device="cuda")
    zk.PoseidonGoldilocks.merkletree(d_tree, d, ncols, nrows)
    zk.sync()
    assert_eq(_host(d_tree)[-4:], hexwords(KATS["merkle_root_8388608x128"]["expected"]), "bench.cpp:440-443 root")
    zk.PoseidonGoldilocks.merkletree_batch(d_tree, d, ncols, nrows, (ncols + 3) // 4)
    zk.sync()
    assert_eq(_host(d_tree)[-4:], hexwords(KATS["merkle_batch_root_8388608x128_b32"]["expected"]), "bench.cpp:584-587 batch root")


def test_headline_against_reference_pin():
    """BASELINE.json configs[1] at full size (2^23 x 665, blowup 2) against tests/golden/headline_pin.json, which
    tools/pin_headline_root.py computed with the COMPILED REFERENCE strip by strip (SURVEY 8(d) option iii): the root,
    1026 leaf digests, nodes of every level and 28 full rows of the extended trace.  Device-resident, the bench's path."""
    import torch
    pin = load_golden("headline_pin.json")
    n, c, b = 1 << pin["log_rows"], pin["cols"], 1 << pin["log_blowup"]
    torch.cuda.empty_cache()
    free, _ = torch.cuda.mem_get_info()
    if free < 142e9:
        pytest.skip("needs 135 GB of device memory plus LDE scratch")
    n_ext = n * b
    d_in = torch.empty(n * c, dtype=torch.int64, device="cuda")
    d_ext = torch.empty(n_ext * c, dtype=torch.int64, device="cuda")
    d_tree = torch.empty((2 * n_ext - 1) * 4, dtype=torch.int64, device="cuda")
    zk.fill(d_in, n * c, SEED, 0)
    zk.extend_and_merkelize(zk.NTT_Goldilocks(n), d_ext, d_tree, d_in, n_ext, n, c)
    zk.sync()
    assert_eq(_host(d_tree[-4:]), hexwords(pin["root"]), "root")
    off = 0
    for lv in pin["levels"]:
        idx = torch.tensor(lv["index"], device="cuda")
        got = d_tree[off: off + lv["nodes"] * 4].view(-1, 4)[idx]
        assert_eq(_host(got), hexwords([w for dgt in lv["digests"] for w in dgt]), "level %d nodes" % lv["level"])
        off += lv["nodes"] * 4
    idx = torch.tensor(pin["rows"]["index"], device="cuda")
    got = d_ext.view(n_ext, c)[idx]
    assert_eq(_host(got), hexwords([w for r in pin["rows"]["words"] for w in r]), "sampled rows of the extended trace")
    del d_in, d_ext, d_tree
    torch.cuda.empty_cache()


def test_stage4_commit_on_device():
    """SURVEY 8(f) N1, second half: stage 4 of Starks::genProof (starks.cpp:261-293) as one device-resident call —
    INTT(NExt x qDim), shift/regroup qq2[k*qDeg+p] = qq1[p*N+k] * shiftIn^p, NTT(NExt x qDim*qDeg), merkelize — against the
    committed outputs of the compiled reference (zkref_stage4: the reference's NTT_Goldilocks + Goldilocks3::mul), the
    oracle, and the live reference where it travelled; host and device pointers."""
    import torch
    from oracle_lib import Reference
    R = Reference()
    for case in [c for c in DIFF["cases"] if c["op"] == "stage4"]:
        ln, lb, qd, qg = case["log_n"], case["log_blowup"], case["q_dim"], case["q_deg"]
        n, ne = 1 << ln, 1 << (ln + lb)
        q = O.fill(ne * qd, SEED + 6, 0)
        cm4 = np.zeros((ne, qd * qg), dtype=np.uint64)
        tree = np.zeros((2 * ne - 1) * 4, dtype=np.uint64)
        zk.stage4_commit(zk.NTT_Goldilocks(ne), cm4, tree, q, ne, n, qd, qg)
        assert digest(cm4) == case["sha256"], case
        assert ["%016x" % int(x) for x in tree[-4:]] == case["root"], case
        want = O.stage4(q, ne, n, qd, qg)
        assert_eq(cm4, want, "cm4_2ns vs oracle")
        assert_eq(tree, O.merkletree(want, qd * qg, ne), "tree vs oracle")
        d_q, d_cm4 = _dev(q), torch.zeros(ne * qd * qg, dtype=torch.int64, device="cuda")
        d_tree = torch.zeros(tree.size, dtype=torch.int64, device="cuda")
        zk.stage4_commit(None, d_cm4, d_tree, d_q, ne, n, qd, qg)
        zk.sync()
        assert_eq(_host(d_cm4), want, "device pointers")
        assert_eq(_host(d_tree), tree, "device tree")
    # a larger shape against the live reference / oracle: 2^18 -> 2^20, qDim 3, qDeg 2 (random non-canonical words)
    rng = np.random.default_rng(5)
    n, ne = 1 << 18, 1 << 20
    q = rng.integers(0, 2**63, size=ne * 3, dtype=np.uint64) * np.uint64(2) + np.uint64(1)
    cm4 = np.zeros((ne, 6), dtype=np.uint64)
    zk.stage4_commit(None, cm4, None, q, ne, n, 3, 2)
    assert_eq(cm4, (R if R else O).stage4(q, ne, n, 3, 2), "2^20 x 3, qDeg 2")
    with pytest.raises(zk.ZkGpuError):
        zk.stage4_commit(None, cm4, None, q, ne, n, 3, 8)  # qDeg > blowup


def test_commit_cache_serves_the_unchanged_call_pair():
    """starks.cpp:53-57 on one host buffer: extendPol then merkelize.  The second call must return the reference's tree
    without a second upload (ZKGPU_PIPE_TRACE shows it), also after the cache was switched off and on again."""
    n, c, b = 1 << 13, 45, 2
    a = O.fill(n * c, SEED, 77).reshape(n, c)
    want_ext = O.extend_pol(a, n * b)
    want_tree = O.merkletree(want_ext, c, n * b)
    ntt = zk.NTT_Goldilocks(n)
    for fuse in (1, 0, 1):
        zk._check(zk.lib().zkgpu_tune(2, fuse))
        ext = np.zeros((n * b, c), dtype=np.uint64)
        ntt.extendPol(ext, a, n * b, n, c)
        assert_eq(ext, want_ext, "extendPol")
        launches = zk.lib().zkgpu_launch_count()
        mt = zk.MerkleTreeGL(n * b, c, ext)
        mt.merkelize()
        assert_eq(mt.nodes, want_tree, "merkelize after extendPol (fuse=%d)" % fuse)
        assert (zk.lib().zkgpu_launch_count() == launches) == bool(fuse), "cache hit must launch nothing, a miss must hash"
        mt.nodes[:] = 0
        mt.merkelize()  # entry consumed: plain path
        assert_eq(mt.nodes, want_tree, "second merkelize")
    ext = np.zeros((n * b, c), dtype=np.uint64)
    ntt.extendPol(ext, a, n * b, n, c)
    ext[0, 0] ^= np.uint64(1)  # a fingerprint row changed: the stale tree must not be served
    mt = zk.MerkleTreeGL(n * b, c, ext)
    mt.merkelize()
    assert_eq(mt.nodes, O.merkletree(ext, c, n * b), "modified buffer")



def test_commit_cache_notices_any_word_and_a_new_matrix_at_the_same_address():
    """VERDICT r2 item 5 / ADVICE: a word changed anywhere in the buffer (not one of the two rows kept verbatim) between
    extendPol and merkelize() must give the recomputed tree; so must another matrix of the same shape at the same address;
    an entry nobody claims holds no device memory; the counters say what happened."""
    n, c, b = 1 << 13, 45, 2
    a = O.fill(n * c, SEED, 78).reshape(n, c)
    ntt = zk.NTT_Goldilocks(n)
    zk._check(zk.lib().zkgpu_tune(2, 1))
    zk.invalidate()
    s0 = zk.cache_stats()
    ext = np.zeros((n * b, c), dtype=np.uint64)
    ntt.extendPol(ext, a, n * b, n, c)
    s1 = zk.cache_stats()
    assert s1["entries"] == 1 and s1["device_bytes"] == 0, s1
    ext[n * b // 5 + 3, 17] ^= np.uint64(1 << 40)  # an arbitrary interior word
    mt = zk.MerkleTreeGL(n * b, c, ext)
    mt.merkelize()
    assert_eq(mt.nodes, O.merkletree(ext, c, n * b), "interior word changed")
    s2 = zk.cache_stats()
    assert s2["stale"] == s0["stale"] + 1 and s2["entries"] == 0, (s0, s2)
    # same address, same shape, another matrix that agrees with the first on rows 0 and n-1
    ntt.extendPol(ext, a, n * b, n, c)
    other = O.fill(n * b * c, SEED, 79).reshape(n * b, c)
    other[0, :] = ext[0, :]
    other[-1, :] = ext[-1, :]
    ext[:] = other
    mt = zk.MerkleTreeGL(n * b, c, ext)
    mt.merkelize()
    assert_eq(mt.nodes, O.merkletree(other, c, n * b), "another matrix at the same address")
    # an untouched buffer is still a hit
    ntt.extendPol(ext, a, n * b, n, c)
    launches = zk.lib().zkgpu_launch_count()
    mt = zk.MerkleTreeGL(n * b, c, ext)
    mt.merkelize()
    assert zk.lib().zkgpu_launch_count() == launches
    s3 = zk.cache_stats()
    assert s3["hits"] == s0["hits"] + 1 and s3["stale"] == s0["stale"] + 2, (s0, s3)


def test_commit_cache_gives_up_on_a_caller_that_never_merkelizes():
    """StarkRecursiveF extends and then builds a BN128 tree (starkRecursiveF.cpp:198-200): after two trees of one
    NTT_Goldilocks object went unclaimed, its extendPol stops hashing (no sponge launches, no entries)."""
    n, c, b = 1 << 10, 24, 2
    a = O.fill(n * c, SEED, 80).reshape(n, c)
    zk._check(zk.lib().zkgpu_tune(2, 1))
    zk.invalidate()
    ntt = zk.NTT_Goldilocks(n)
    bufs = [np.zeros((n * b, c), dtype=np.uint64) for _ in range(4)]
    per_call = []
    for ext in bufs:
        l0 = zk.lib().zkgpu_launch_count()
        ntt.extendPol(ext, a, n * b, n, c)
        per_call.append(zk.lib().zkgpu_launch_count() - l0)
    assert zk.cache_stats()["entries"] == 0, zk.cache_stats()
    assert per_call[2] < per_call[0] and per_call[3] == per_call[2], per_call
    want = O.extend_pol(a, n * b)
    for ext in bufs:
        assert_eq(ext, want, "extendPol")
    # another object is unaffected, and merkelize() on its buffer is a hit
    ntt2 = zk.NTT_Goldilocks(n)
    ntt2.extendPol(bufs[0], a, n * b, n, c)
    assert zk.cache_stats()["entries"] == 1
    mt = zk.MerkleTreeGL(n * b, c, bufs[0])
    mt.merkelize()
    assert_eq(mt.nodes, O.merkletree(want, c, n * b), "tree")
    zk.invalidate()
