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
void zkref_set_threads(int n) { omp_set_dynamic(0); omp_set_num_threads(n > 0 ? n : 1); }

// NTT_Goldilocks::reversePermutation (ntt_goldilocks.cpp:243-305) on an object built with the given extension
void zkref_reverse_permutation(uint64_t *dst, uint64_t *src, uint64_t size, uint64_t offset_cols, uint64_t ncols, uint64_t ncols_all, int extension)
{
    NTT_Goldilocks ntt(size, 0, extension);
    ntt.reversePermutation((E *)dst, (E *)src, size, offset_cols, ncols, ncols_all);
}

// One column strip of the row sponge, for pinning shapes whose extended trace does not fit host RAM as a whole: the loop
// body of PoseidonGoldilocks::linear_hash (poseidon_goldilocks.cpp:240-256) over the reference's own hash_full_result,
// with the running state[0..3] of row i kept in cap[4*i..] between strips ("capacity <- previous out[0..3]", :246-249).
// Every strip but the last must be a multiple of RATE words so that the chunking equals the whole-row chunking.
void zkref_sponge_absorb(uint64_t *cap, uint64_t *rows, uint64_t strip_words, uint64_t row_ld, uint64_t nrows, int first)
{
#pragma omp parallel for
    for (uint64_t i = 0; i < nrows; i++)
    {
        E state[SPONGE_WIDTH];
        E *in = (E *)rows + i * row_ld;
        if (first) memset(state + RATE, 0, CAPACITY * sizeof(E));
        else std::memcpy(state + RATE, cap + 4 * i, CAPACITY * sizeof(E));
        uint64_t remaining = strip_words;
        while (remaining)
        {
            uint64_t n = (remaining < RATE) ? remaining : RATE;
            memset(&state[n], 0, (RATE - n) * sizeof(E));
            std::memcpy(state, in + (strip_words - remaining), n * sizeof(E));
            PoseidonGoldilocks::hash_full_result(state, state);
            remaining -= n;
            if (remaining) std::memcpy(state + RATE, state, CAPACITY * sizeof(E));
        }
        std::memcpy(cap + 4 * i, state, CAPACITY * sizeof(E));
    }
}

// Stage 4 of Starks::genProof (src/starkpil/starks.cpp:261-285) with the reference's own pieces:
// nttExtended.INTT(qq1, q_2ns, NExt, qDim, NULL, 2, 1); the shift/regroup loop over Goldilocks3::mul(Element&, Element&,
// Goldilocks::Element&) (goldilocks_cubic_extension.hpp:1887-1892); nttExtended.NTT(cm4_2ns, qq2, NExt, qDim*qDeg).
// qq2 is a zero-initialised Polinomial(NExt*qDeg, qDim) (:233).  qDim must be 3 (Goldilocks3).
void zkref_stage4(uint64_t *cm4_2ns, uint64_t *q_2ns, uint64_t n_ext, uint64_t n, uint64_t q_dim, uint64_t q_deg)
{
    NTT_Goldilocks nttExtended(n_ext);
    Polinomial qq1(n_ext, q_dim);
    Polinomial qq2(n_ext * q_deg, q_dim);
    memset(qq2.address(), 0, n_ext * q_deg * q_dim * sizeof(E));
    nttExtended.INTT(qq1.address(), (E *)q_2ns, n_ext, q_dim, NULL, 2, 1);
    E shiftIn = Goldilocks::exp(Goldilocks::inv(Goldilocks::shift()), n);
    uint64_t stride = 2048;
#pragma omp parallel for
    for (uint64_t ii = 0; ii < n; ii += stride)
    {
        E curS = Goldilocks::one();
        for (uint64_t p = 0; p < q_deg; p++)
        {
            for (uint64_t k = ii; k < std::min(n, ii + stride); ++k)
                Goldilocks3::mul((Goldilocks3::Element &)*qq2[k * q_deg + p], (Goldilocks3::Element &)*qq1[p * n + k], curS);
            curS = Goldilocks::mul(curS, shiftIn);
        }
    }
    nttExtended.NTT((E *)cm4_2ns, qq2.address(), n_ext, q_dim * q_deg);
}


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
