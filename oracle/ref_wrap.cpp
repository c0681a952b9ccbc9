// oracle build glue (TEST INFRASTRUCTURE, not product code).
// extern "C" entry points around the UNMODIFIED reference classes, compiled from the sources
// where they lie under /root/reference/src/goldilocks/src (see oracle/Makefile, target _ref).
// Used (a) to pin oracle/gl_oracle.c, (b) to generate tests/golden fixtures, (c) as the timed
// CPU baseline ("kind": "reference") in bench.py.  Nothing in the product links this.
#include "goldilocks_base_field.hpp"
#include "ntt_goldilocks.hpp"
#include "poseidon_goldilocks.hpp"
#include <omp.h>
#include <cstdint>

typedef Goldilocks::Element E;

extern "C" {

int zkref_threads(void) { return omp_get_max_threads(); }

// Goldilocks::toU64 (goldilocks_base_field_tools.hpp:152-163) over an array, in place.
void zkref_canon(uint64_t *x, uint64_t n)
{
#pragma omp parallel for
    for (uint64_t i = 0; i < n; i++) x[i] = Goldilocks::toU64(((E *)x)[i]);
}
uint64_t zkref_add(uint64_t a, uint64_t b) { E x{a}, y{b}; return (x + y).fe; }
uint64_t zkref_sub(uint64_t a, uint64_t b) { E x{a}, y{b}; return (x - y).fe; }
uint64_t zkref_mul(uint64_t a, uint64_t b) { E x{a}, y{b}; return (x * y).fe; }

// NTT_Goldilocks::NTT / INTT (ntt_goldilocks.cpp:166-232, 307-324)
void zkref_ntt(uint64_t *dst, uint64_t *src, uint64_t size, uint64_t ncols, uint64_t *buffer,
               uint64_t nphase, uint64_t nblock, int inverse, int extend)
{
    NTT_Goldilocks ntt(size);
    if (extend) ntt.computeR(size);
    if (inverse) ntt.INTT((E *)dst, (E *)src, size, ncols, (E *)buffer, nphase, nblock, extend != 0);
    else ntt.NTT((E *)dst, (E *)src, size, ncols, (E *)buffer, nphase, nblock);
}
// NTT_Goldilocks::extendPol (ntt_goldilocks.cpp:326-352)
void zkref_extend_pol(uint64_t *out, uint64_t *in, uint64_t n_ext, uint64_t n, uint64_t ncols, uint64_t *buffer)
{
    NTT_Goldilocks ntt(n);
    ntt.extendPol((E *)out, (E *)in, n_ext, n, ncols, (E *)buffer);
}
// PoseidonGoldilocks::hash_full_result (AVX2, poseidon_goldilocks.cpp:166-228) and _seq (:5-37)
void zkref_hash_full_result(uint64_t *out12, const uint64_t *in12) { PoseidonGoldilocks::hash_full_result((E *)out12, (const E *)in12); }
void zkref_hash_full_result_seq(uint64_t *out12, const uint64_t *in12) { PoseidonGoldilocks::hash_full_result_seq((E *)out12, (const E *)in12); }
// PoseidonGoldilocks::linear_hash (poseidon_goldilocks.cpp:229-265)
void zkref_linear_hash(uint64_t *out4, uint64_t *in, uint64_t n) { PoseidonGoldilocks::linear_hash((E *)out4, (E *)in, n); }
void zkref_linear_hash_seq(uint64_t *out4, uint64_t *in, uint64_t n) { PoseidonGoldilocks::linear_hash_seq((E *)out4, (E *)in, n); }
// PoseidonGoldilocks::merkletree (wrapper poseidon_goldilocks.hpp:74-81 -> merkletree_avx :266-301)
void zkref_merkletree(uint64_t *tree, uint64_t *input, uint64_t ncols, uint64_t nrows, uint64_t dim)
{ PoseidonGoldilocks::merkletree((E *)tree, (E *)input, ncols, nrows, 0, dim); }
void zkref_merkletree_seq(uint64_t *tree, uint64_t *input, uint64_t ncols, uint64_t nrows, uint64_t dim)
{ PoseidonGoldilocks::merkletree_seq((E *)tree, (E *)input, ncols, nrows, 0, dim); }
// PoseidonGoldilocks::merkletree_batch (poseidon_goldilocks.cpp:302-352)
void zkref_merkletree_batch(uint64_t *tree, uint64_t *input, uint64_t ncols, uint64_t nrows, uint64_t batch, uint64_t dim)
{ PoseidonGoldilocks::merkletree_batch((E *)tree, (E *)input, ncols, nrows, batch, 0, dim); }

} // extern "C"
