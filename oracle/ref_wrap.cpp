// oracle build glue (TEST INFRASTRUCTURE, not product code).
// extern "C" entry points around the UNMODIFIED reference classes, compiled from the sources
// where they lie under /root/reference/src/goldilocks/src (see oracle/Makefile, target _ref).
// Used (a) to pin oracle/gl_oracle.c, (b) to generate tests/golden fixtures, (c) as the timed
// CPU baseline ("kind": "reference") in bench.py.  Nothing in the product links this.
#include "goldilocks_base_field.hpp"
#include "ntt_goldilocks.hpp"
#include "poseidon_goldilocks.hpp"
#include "polinomial.hpp" // src/starkpil/polinomial.hpp: the element arithmetic FRIProve uses (cubic extension)
#include <omp.h>
#include <cstdint>

typedef Goldilocks::Element E;

// link-time glue for what polinomial.hpp declares but this build never calls (src/utils/*.cpp are not compiled)
zkLog zklog;
zkLog::zkLog() {}
void zkLog::log(const zkLogType, const string &, const vector<LogTag> *) {}
void exitProcess(void) { abort(); }
bool CompareFeImpl(const Goldilocks::Element &a, const Goldilocks::Element &b) { return Goldilocks::toU64(a) < Goldilocks::toU64(b); }
bool CompareFeVectorImpl(const std::vector<Goldilocks::Element> &a, const std::vector<Goldilocks::Element> &b) { return a.size() < b.size(); }

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

// One FRI folding step, si > 0, with the reference's own pieces: the loop body of FRIProve::prove
// (friProve.cpp:89-106) around NTT_Goldilocks::INTT and Polinomial::{copyElement, mulElement, addElement};
// polMulAxi (:192-200) and evalPol (:201-217) are members of a class whose translation unit needs the whole prover
// (transcript, stark_info, json), so their few lines are spelled here over the same Polinomial calls.
// The per-thread starting value of sinv (:64-81) is shiftInv * wi^g; here it is walked sequentially.
void zkref_fri_fold(uint64_t *pol2_e_, uint64_t *fri_pol_, uint64_t pol_bits, uint64_t next_bits, uint64_t shift_inv, uint64_t *special_x_)
{
    const uint64_t pol2N = 1ull << next_bits, nX = (1ull << pol_bits) / pol2N;
    Polinomial friPol((E *)fri_pol_, 1ull << pol_bits, FIELD_EXTENSION, FIELD_EXTENSION);
    Polinomial pol2_e((E *)pol2_e_, pol2N, FIELD_EXTENSION, FIELD_EXTENSION);
    Polinomial special_x((E *)special_x_, 1, FIELD_EXTENSION, FIELD_EXTENSION);
    E wi = Goldilocks::inv(Goldilocks::w(pol_bits));
    E sinv_ = Goldilocks::fromU64(shift_inv);
    for (uint64_t g = 0; g < pol2N; g++)
    {
        Polinomial ppar(nX, FIELD_EXTENSION);
        Polinomial ppar_c(nX, FIELD_EXTENSION);
        for (uint64_t i = 0; i < nX; i++) Polinomial::copyElement(ppar, i, friPol, (i * pol2N) + g);
        NTT_Goldilocks ntt(nX, 1);
        ntt.INTT(ppar_c.address(), ppar.address(), nX, FIELD_EXTENSION);
        E r = Goldilocks::one(); // polMulAxi(ppar_c, one, sinv_)
        for (uint64_t i = 0; i < ppar_c.degree(); i++)
        {
            Polinomial::mulElement(ppar_c, i, ppar_c, i, r);
            r = r * sinv_;
        }
        Polinomial::copyElement(pol2_e, g, ppar_c, ppar_c.degree() - 1); // evalPol(pol2_e, g, ppar_c, special_x)
        for (int64_t i = (int64_t)ppar_c.degree() - 2; i >= 0; i--)
        {
            Polinomial aux(1, 3);
            Polinomial::mulElement(aux, 0, pol2_e, g, special_x, 0);
            Polinomial::addElement(pol2_e, g, aux, 0, ppar_c, i);
        }
        sinv_ = sinv_ * wi;
    }
}

} // extern "C"
