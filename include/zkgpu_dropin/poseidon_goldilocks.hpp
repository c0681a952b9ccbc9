// poseidon_goldilocks.hpp — drop-in for the reference's src/goldilocks/src/poseidon_goldilocks.hpp (+ .cpp).
//
// Same class name, static methods, argument order and defaults as reference poseidon_goldilocks.hpp:43-81.  The
// _seq / _avx / _avx512 spellings all name the same GPU implementation (the reference's variants compute the same
// function, tests.cpp:1505-1994).  Outputs are canonical words (Goldilocks::toU64 of the reference's output).
// `nThreads` is accepted and ignored.  Pointers may be host or device pointers.
#ifndef POSEIDON_GOLDILOCKS
#define POSEIDON_GOLDILOCKS
#include "goldilocks_element.hpp"

#define RATE 8
#define CAPACITY 4
#define HASH_SIZE 4
#define SPONGE_WIDTH (RATE + CAPACITY)
#define HALF_N_FULL_ROUNDS 4
#define N_FULL_ROUNDS_TOTAL (2 * HALF_N_FULL_ROUNDS)
#define N_PARTIAL_ROUNDS 22
#define N_ROUNDS (N_FULL_ROUNDS_TOTAL + N_PARTIAL_ROUNDS)

class PoseidonGoldilocks
{
    typedef Goldilocks::Element E;

public:
    // poseidon_goldilocks.cpp:166-228 (scalar statement :5-37): one permutation, all 12 words out
    static void hash_full_result(E *state, const E *input)
    {
        zkgpu_dropin::check(zkgpu_hash_full_result(zkgpu_dropin::words(state), zkgpu_dropin::words(input), 1), "PoseidonGoldilocks::hash_full_result");
    }
    static void hash(E (&state)[CAPACITY], const E (&input)[SPONGE_WIDTH])
    {
        E full[SPONGE_WIDTH];
        hash_full_result(full, input);
        for (int i = 0; i < CAPACITY; i++) state[i] = full[i];
    }
    // poseidon_goldilocks.cpp:229-265
    static void linear_hash(E *output, E *input, uint64_t size)
    {
        zkgpu_dropin::check(zkgpu_linear_hash(zkgpu_dropin::words(output), zkgpu_dropin::words(input), size), "PoseidonGoldilocks::linear_hash");
    }
    // poseidon_goldilocks.hpp:74-81 -> poseidon_goldilocks.cpp:266-301
    static void merkletree(E *tree, E *input, uint64_t num_cols, uint64_t num_rows, int nThreads = 0, uint64_t dim = 1)
    {
        (void)nThreads;
        zkgpu_dropin::check(zkgpu_merkletree(zkgpu_dropin::words(tree), zkgpu_dropin::words(input), num_cols, num_rows, dim), "PoseidonGoldilocks::merkletree");
    }
    // poseidon_goldilocks.cpp:302-352
    static void merkletree_batch(E *tree, E *input, uint64_t num_cols, uint64_t num_rows, uint64_t batch_size, int nThreads = 0, uint64_t dim = 1)
    {
        (void)nThreads;
        zkgpu_dropin::check(zkgpu_merkletree_batch(zkgpu_dropin::words(tree), zkgpu_dropin::words(input), num_cols, num_rows, batch_size, dim),
                            "PoseidonGoldilocks::merkletree_batch");
    }

    static void hash_full_result_seq(E *s, const E *i) { hash_full_result(s, i); }
    static void hash_seq(E (&s)[CAPACITY], const E (&i)[SPONGE_WIDTH]) { hash(s, i); }
    static void linear_hash_seq(E *o, E *i, uint64_t n) { linear_hash(o, i, n); }
    static void merkletree_seq(E *t, E *i, uint64_t c, uint64_t r, int nt = 0, uint64_t d = 1) { merkletree(t, i, c, r, nt, d); }
    static void merkletree_avx(E *t, E *i, uint64_t c, uint64_t r, int nt = 0, uint64_t d = 1) { merkletree(t, i, c, r, nt, d); }
    static void merkletree_batch_seq(E *t, E *i, uint64_t c, uint64_t r, uint64_t b, int nt = 0, uint64_t d = 1) { merkletree_batch(t, i, c, r, b, nt, d); }
    static void merkletree_batch_avx(E *t, E *i, uint64_t c, uint64_t r, uint64_t b, int nt = 0, uint64_t d = 1) { merkletree_batch(t, i, c, r, b, nt, d); }
#ifdef __AVX512__
    // poseidon_goldilocks.hpp:64-66 (dead in the shipped AVX2 build, Makefile:31-37; present so that an AVX-512 build links)
    static void hash_full_result_avx512(E *s, const E *i) { hash_full_result(s, i); }
    static void hash_avx512(E (&s)[CAPACITY], const E (&i)[SPONGE_WIDTH]) { hash(s, i); }
    static void linear_hash_avx512(E *o, E *i, uint64_t n) { linear_hash(o, i, n); }
    static void merkletree_avx512(E *t, E *i, uint64_t c, uint64_t r, int nt = 0, uint64_t d = 1) { merkletree(t, i, c, r, nt, d); }
    static void merkletree_batch_avx512(E *t, E *i, uint64_t c, uint64_t r, uint64_t b, int nt = 0, uint64_t d = 1) { merkletree_batch(t, i, c, r, b, nt, d); }
#endif
};
#endif
