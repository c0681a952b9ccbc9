// dropin_test.cpp — exercises the C++ drop-in classes (include/zkgpu_dropin) the way the reference's own gtest file
// does (src/goldilocks/tests/tests.cpp: poseidon_full_avx :1644, merkletree_avx :1883, extendePol :2554, ntt :2225),
// restated as a plain main().  `dropin_test cpu` runs what needs no device; `dropin_test gpu` runs the KATs.
#include "ntt_goldilocks.hpp"
#include "poseidon_goldilocks.hpp"
#include "merkleTreeGL.hpp"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

static int fails = 0;
#define EXPECT(cond)                                                      \
    do {                                                                  \
        if (!(cond)) { printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #cond); fails++; } \
    } while (0)

static uint64_t addmod(uint64_t a, uint64_t b) { return (uint64_t)(((unsigned __int128)a + b) % GOLDILOCKS_PRIME); }

static void cpu_part()
{
    // size 0 / ncols 0 are silent no-ops (ntt_goldilocks.cpp:168-171, 310-313; poseidon_goldilocks.cpp:268-271)
    NTT_Goldilocks ntt(16);
    Goldilocks::Element z[4] = {};
    ntt.NTT(z, z, 0);
    ntt.INTT(z, z, 4, 0);
    ntt.extendPol(z, z, 8, 4, 0);
    PoseidonGoldilocks::merkletree(z, z, 3, 0);
    bool thrown = false;
    try { NTT_Goldilocks big((uint64_t)1 << 33); } catch (const std::range_error &) { thrown = true; }
    EXPECT(thrown); // ntt_goldilocks.hpp:94-97
    thrown = false;
    try { ntt.NTT(z, z, 3); } catch (const std::runtime_error &) { thrown = true; }
    EXPECT(thrown); // reference asserts on non power-of-two sizes
    EXPECT(Goldilocks::toU64(ntt.root(1, 1)) == GOLDILOCKS_PRIME - 1);     // W[1] = -1
    EXPECT(Goldilocks::toU64(ntt.root(2, 1)) == 281474976710656ULL);       // W[2] = 2^48
    EXPECT(Goldilocks::toU64(ntt.root(3, 2)) == 281474976710656ULL);
    MerkleTreeGL mt(8, 3, NULL);
    EXPECT(mt.getTreeNumElements() == 8 * 4 + 7 * 4);
    EXPECT(mt.MerkleProofSize() == 3);
}

static void gpu_part()
{
    // tests.cpp:1644 poseidon_full_avx: Fibonacci state
    Goldilocks::Element st[12], out[12];
    st[0].fe = 0; st[1].fe = 1;
    for (int i = 2; i < 12; i++) st[i].fe = addmod(st[i - 1].fe, st[i - 2].fe);
    PoseidonGoldilocks::hash_full_result(out, st);
    EXPECT(out[0].fe == 0x3095570037F4605DULL);
    EXPECT(out[11].fe == 0xF58A41D4028C66A5ULL);

    // tests.cpp:1883 merkletree_avx: 64 rows x 128 cols, rows 0,1 = 1..128, row j = row j-1 + row j-2
    const uint64_t ncols = 128, nrows = 64;
    std::vector<Goldilocks::Element> cols(ncols * nrows);
    for (uint64_t i = 0; i < ncols; i++) { cols[i].fe = i + 1; cols[i + ncols].fe = i + 1; }
    for (uint64_t j = 2; j < nrows; j++)
        for (uint64_t i = 0; i < ncols; i++) cols[j * ncols + i].fe = addmod(cols[(j - 1) * ncols + i].fe, cols[(j - 2) * ncols + i].fe);
    MerkleTreeGL mt(nrows, ncols, cols.data());
    mt.merkelize();
    Goldilocks::Element root[4];
    mt.getRoot(root);
    EXPECT(root[0].fe == 0x918F7CD0C3E8701FULL);
    EXPECT(root[3].fe == 0xBD2B98A57B748BF4ULL);
    // proof of row 5 re-hashes to the root
    std::vector<Goldilocks::Element> proof(ncols + 4 * mt.MerkleProofSize());
    mt.getGroupProof(proof.data(), 5);
    EXPECT(memcmp(proof.data(), &cols[5 * ncols], ncols * 8) == 0);
    Goldilocks::Element node[4];
    PoseidonGoldilocks::linear_hash(node, proof.data(), ncols);
    uint64_t idx = 5;
    for (uint64_t l = 0; l < mt.MerkleProofSize(); l++, idx >>= 1)
    {
        Goldilocks::Element in[12] = {}, full[12];
        const Goldilocks::Element *sib = &proof[ncols + 4 * l];
        for (int k = 0; k < 4; k++) { in[(idx & 1) ? 4 + k : k] = node[k]; in[(idx & 1) ? k : 4 + k] = sib[k]; }
        PoseidonGoldilocks::hash_full_result(full, in);
        for (int k = 0; k < 4; k++) node[k] = full[k];
    }
    EXPECT(memcmp(node, root, 32) == 0);

    // tests.cpp:2554 extendePol: N = 16 -> 32, 8 columns of Fibonacci (1+j, 1+j)
    const uint64_t N = 16, NE = 32, C = 8;
    std::vector<Goldilocks::Element> a(NE * C), b(NE * C);
    for (uint64_t j = 0; j < C; j++) { a[j].fe = 1 + j; a[C + j].fe = 1 + j; }
    for (uint64_t i = 2; i < N; i++)
        for (uint64_t j = 0; j < C; j++) a[i * C + j].fe = addmod(a[(i - 1) * C + j].fe, a[(i - 2) * C + j].fe);
    NTT_Goldilocks ntt(N);
    ntt.extendPol(a.data(), a.data(), NE, N, C, b.data());
    EXPECT(a[0].fe == 0x3E7CA26D67147C31ULL);
    EXPECT(a[31 * C].fe == 0xA0705C72DD9F9A2FULL);

    // tests.cpp:2225 ntt: INTT(NTT(x)) == x, in place and with dst == NULL semantics of INTT
    const uint64_t M = 1 << 10;
    std::vector<Goldilocks::Element> x(M * 3), y(M * 3);
    for (uint64_t i = 0; i < M * 3; i++) x[i].fe = (i * 0x9E3779B97F4A7C15ULL) % GOLDILOCKS_PRIME;
    y = x;
    NTT_Goldilocks g(M);
    g.NTT(y.data(), y.data(), M, 3);
    EXPECT(memcmp(x.data(), y.data(), M * 24) != 0);
    g.INTT(NULL, y.data(), M, 3);
    EXPECT(memcmp(x.data(), y.data(), M * 24) == 0);
}

static uint64_t splitmix(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

// The call pair of UNCHANGED starks.cpp (:53-57, starks.hpp:186): ntt.extendPol(cm_2ns, cm_n, NExt, N, nCols, buffer) and
// then MerkleTreeGL(NExt, nCols, cm_2ns).merkelize() on the same host buffer.  The tree must equal the one computed from a
// COPY of the extended trace (a different pointer: no commit-cache hit, the plain upload-and-hash path), the second
// merkelize on the same buffer (entry consumed) must still be right, and a buffer modified in between must not get the
// stale tree.
static void commit_pair(uint64_t N, uint64_t NE, uint64_t C)
{
    std::vector<Goldilocks::Element> in(N * C), ext(NE * C), buf(NE * C);
    for (uint64_t i = 0; i < N * C; i++) in[i].fe = splitmix(0x9E3779B97F4A7C15ULL + i) % GOLDILOCKS_PRIME;
    NTT_Goldilocks ntt(N);
    ntt.extendPol(ext.data(), in.data(), NE, N, C, buf.data());
    MerkleTreeGL mt(NE, C, ext.data());
    mt.merkelize(); // served by the commit cache
    std::vector<Goldilocks::Element> copy(ext);
    MerkleTreeGL ref(NE, C, copy.data());
    ref.merkelize(); // uploaded and hashed
    EXPECT(memcmp(mt.nodes, ref.nodes, mt.getTreeNumElements() * 8) == 0);
    memset(mt.nodes, 0, mt.getTreeNumElements() * 8);
    mt.merkelize(); // entry already consumed: plain path
    EXPECT(memcmp(mt.nodes, ref.nodes, mt.getTreeNumElements() * 8) == 0);
    // modified between the two calls: the fingerprint (or an explicit zkgpu_invalidate) must keep the stale tree away
    ntt.extendPol(ext.data(), in.data(), NE, N, C, buf.data());
    ext[0].fe ^= 1;
    copy = ext;
    mt.merkelize();
    ref.merkelize();
    EXPECT(memcmp(mt.nodes, ref.nodes, mt.getTreeNumElements() * 8) == 0);
    ntt.extendPol(ext.data(), in.data(), NE, N, C, buf.data());
    ext[(NE / 2 + 3) * C + 1].fe ^= 1; // an interior word: covered by the checksum when the buffer has <= 2^20 words (or with
                                       // ZKGPU_CACHE_VERIFY=full); a few words of a larger buffer are the caller's to announce
    if (NE * C > ((uint64_t)1 << 20)) zkgpu_invalidate(ext.data());
    copy = ext;
    mt.merkelize();
    ref.merkelize();
    EXPECT(memcmp(mt.nodes, ref.nodes, mt.getTreeNumElements() * 8) == 0);
}

// `dropin_test gpu2`: one process, two GPUs (zkgpu_init_multi) behind the same classes; trees and extended traces must
// equal the 1-GPU ones word for word.
static void gpu2_part()
{
    const uint64_t N = 1 << 16, NE = 1 << 17, C = 665;
    std::vector<Goldilocks::Element> in(N * C), ext1(NE * C), ext2(NE * C), buf(NE * C);
    for (uint64_t i = 0; i < N * C; i++) in[i].fe = splitmix(0x9E3779B97F4A7C15ULL + i) % GOLDILOCKS_PRIME;
    NTT_Goldilocks ntt(N);
    MerkleTreeGL t1(NE, C, ext1.data()), t2(NE, C, ext2.data());
    zkgpu_dropin::check(zkgpu_extend_and_merkelize(ntt.handle(), zkgpu_dropin::words(ext1.data()), zkgpu_dropin::words(t1.nodes),
                                                   zkgpu_dropin::words(in.data()), NE, N, C), "1 gpu");
    zkgpu_dropin::check(zkgpu_init_multi(2), "zkgpu_init_multi");
    zkgpu_dropin::check(zkgpu_extend_and_merkelize(ntt.handle(), zkgpu_dropin::words(ext2.data()), zkgpu_dropin::words(t2.nodes),
                                                   zkgpu_dropin::words(in.data()), NE, N, C), "2 gpus");
    EXPECT(memcmp(ext1.data(), ext2.data(), NE * C * 8) == 0);
    EXPECT(memcmp(t1.nodes, t2.nodes, t1.getTreeNumElements() * 8) == 0);
    // and the unchanged call pair, now served by both GPUs
    memset(ext2.data(), 0, NE * C * 8);
    memset(t2.nodes, 0, t2.getTreeNumElements() * 8);
    ntt.extendPol(ext2.data(), in.data(), NE, N, C, buf.data());
    t2.merkelize();
    EXPECT(memcmp(ext1.data(), ext2.data(), NE * C * 8) == 0);
    EXPECT(memcmp(t1.nodes, t2.nodes, t1.getTreeNumElements() * 8) == 0);
    int rank = -1, world = 0, peer = -1;
    zkgpu_dist_info(&rank, &world, &peer);
    printf("gpu2: world %d, peer stores %d\n", world, peer);
    EXPECT(world == 2);
}

int main(int argc, char **argv)
{
    const bool gpu = argc > 1 && strcmp(argv[1], "gpu") == 0;
    const bool gpu2 = argc > 1 && strcmp(argv[1], "gpu2") == 0;
    try
    {
        cpu_part();
        if (gpu)
        {
            gpu_part();
            commit_pair(1 << 12, 1 << 13, 45);
            commit_pair(1 << 14, 1 << 15, 665);
        }
        if (gpu2) gpu2_part();
    }
    catch (const std::exception &e)
    {
        printf("FAIL exception: %s\n", e.what());
        return 2;
    }
    printf(fails ? "dropin_test: %d failure(s)\n" : "dropin_test: ok\n", fails);
    return fails ? 1 : 0;
}
