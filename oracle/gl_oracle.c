/* oracle/gl_oracle.c — TEST INFRASTRUCTURE ONLY (the checker, never the product).
 *
 * A plain-C CPU restatement of the reference's STARK-commit hot path: Goldilocks field ops,
 * NTT / INTT / extendPol, Poseidon permutation, linear_hash, merkletree, merkletree_batch.
 * Each function cites the reference file:line it follows (paths relative to
 * /root/reference/src/goldilocks/src/).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this; the product (libzkgpu.so) never does.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this file against every known-answer
 * vector the reference's own tests hold for this path (tests/golden/reference_kats.json, extracted
 * from src/goldilocks/tests/tests.cpp and benchs/bench.cpp) and, in the build container, word for
 * word against the compiled unmodified reference (oracle/_ref/libzkref.so).
 *
 * Words are raw reference-style (possibly non-canonical, in [0,2^64)); oracle_canon() applies
 * Goldilocks::toU64.  All comparisons in tests are made on canonical words.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "poseidon_constants.h"

typedef uint64_t u64;
typedef unsigned __int128 u128;

#define GL_P 0xFFFFFFFF00000001ULL
#define GL_EPS 0xFFFFFFFFULL /* 2^64 mod p  (Goldilocks::CQ, goldilocks_base_field.cpp:8) */

/* goldilocks_base_field_scalar.hpp:16-35 — add with up to two carry corrections, NOT canonical */
static inline u64 gl_add(u64 a, u64 b)
{
    u64 s = a + b;
    if (s < a) { u64 t = s + GL_EPS; if (t < s) t += GL_EPS; s = t; }
    return s;
}
/* goldilocks_base_field_scalar.hpp:62-80 */
static inline u64 gl_sub(u64 a, u64 b)
{
    u64 d = a - b;
    if (a < b) { u64 t = d - GL_EPS; if (t > d) t -= GL_EPS; d = t; }
    return d;
}
/* goldilocks_base_field_scalar.hpp:127-154 — 128-bit product folded with 2^64 = 2^32-1, 2^96 = -1.
 * The reference computes  lo + (hl*(2^32-1) + 2^32) - (hh + 2^32)  with carry fix-ups; the value
 * class mod p is what matters to every consumer (all reference asserts go through toU64), but we
 * mirror the same fold so raw words agree with the scalar reference path. */
static inline u64 gl_mul(u64 a, u64 b)
{
    u128 pr = (u128)a * b;
    u64 lo = (u64)pr, hi = (u64)(pr >> 64);
    u64 hh = hi >> 32, hl = hi & 0xFFFFFFFFULL;
    u64 m = hl * GL_EPS + 0x100000000ULL; /* < 2^64 */
    u64 t = lo + m;
    if (t < lo) t += GL_EPS;
    u64 s = hh + 0x100000000ULL;
    u64 r = t - s;
    if (t < s) r -= GL_EPS;
    return r;
}
/* goldilocks_base_field_tools.hpp:152-163 — Goldilocks::toU64 */
static inline u64 gl_canon(u64 x) { return x >= GL_P ? x - GL_P : x; }

static u64 gl_pow(u64 b, u64 e)
{
    u64 r = 1;
    while (e) { if (e & 1) r = gl_mul(r, b); b = gl_mul(b, b); e >>= 1; }
    return r;
}
static u64 gl_inv(u64 a) { return gl_pow(a, GL_P - 2); }

/* goldilocks_base_field.cpp:11-44 — W[k] is a primitive 2^k-th root of unity with W[k]^2 = W[k-1].
 * Only the seed W[32] is restated; the rest follow by squaring (checked against the table in
 * tests/test_oracle.py through oracle_w()). */
#define GL_W32 7277203076849721926ULL
static u64 gl_w(unsigned k)
{
    u64 w = GL_W32;
    for (unsigned i = 32; i > k; --i) w = gl_mul(w, w);
    return gl_canon(w);
}

u64 oracle_add(u64 a, u64 b) { return gl_add(a, b); }
u64 oracle_sub(u64 a, u64 b) { return gl_sub(a, b); }
u64 oracle_mul(u64 a, u64 b) { return gl_mul(a, b); }
u64 oracle_w(unsigned k) { return gl_w(k); }
void oracle_canon(u64 *x, u64 n)
{
#pragma omp parallel for
    for (u64 i = 0; i < n; i++) x[i] = gl_canon(x[i]);
}

static unsigned ilog2(u64 n) { unsigned r = 0; while (n > 1) { n >>= 1; r++; } return r; }
static u64 bitrev(u64 x, unsigned bits)
{
    u64 r = 0;
    for (unsigned i = 0; i < bits; i++) { r = (r << 1) | (x & 1); x >>= 1; }
    return r;
}

/* Forward transform of every column of a row-major size x ncols matrix:
 *   out[k] = sum_n in[n] * w^(n*k),  w = W[log2 size], natural order in and out.
 * Follows ntt_goldilocks.cpp:12-164 (NTT_iters): bit-reversal gather (reversePermutation,
 * :243-305, rows >= nrows_valid read as zero as in :259-277) then log2(size) radix-2 DIT stages
 * t = w*a[j+m/2]; a[j] = u+t; a[j+m/2] = u-t (:84-112) with twiddle root(s, j) = w_m^j.  The
 * reference's phase/batch blocking only reorders independent butterflies; the dataflow per
 * element is this network exactly. */
static void ntt_forward(u64 *dst, const u64 *src, u64 size, u64 ncols, u64 nrows_valid)
{
    unsigned lg = ilog2(size);
    u64 *roots = (u64 *)malloc((size / 2 + 1) * sizeof(u64));
    u64 w = gl_w(lg);
    roots[0] = 1;
    for (u64 i = 1; i < size / 2; i++) roots[i] = gl_mul(roots[i - 1], w); /* ntt_goldilocks.hpp:120-123 */
    u64 *a = (dst == src) ? (u64 *)malloc(size * ncols * sizeof(u64)) : dst;
#pragma omp parallel for
    for (u64 i = 0; i < size; i++)
    {
        u64 r = bitrev(i, lg);
        if (r < nrows_valid) memcpy(a + i * ncols, src + r * ncols, ncols * sizeof(u64));
        else memset(a + i * ncols, 0, ncols * sizeof(u64));
    }
    for (unsigned s = 1; s <= lg; s++)
    {
        u64 m = 1ULL << s, half = m >> 1, step = size >> s;
#pragma omp parallel for
        for (u64 b = 0; b < size / 2; b++)
        {
            u64 k = (b / half) * m, j = b % half;
            u64 wj = roots[j * step];
            u64 *lo = a + (k + j) * ncols, *hi = a + (k + j + half) * ncols;
            for (u64 c = 0; c < ncols; c++)
            {
                u64 t = gl_mul(wj, hi[c]);
                u64 u = lo[c];
                lo[c] = gl_add(t, u);
                hi[c] = gl_sub(u, t);
            }
        }
    }
    if (a != dst) { memcpy(dst, a, size * ncols * sizeof(u64)); free(a); }
    free(roots);
}

/* NTT_Goldilocks::NTT (ntt_goldilocks.cpp:166-232): size==0 or ncols==0 is a no-op. */
void oracle_ntt(u64 *dst, const u64 *src, u64 size, u64 ncols)
{
    if (size == 0 || ncols == 0) return;
    ntt_forward(dst, src, size, ncols, size);
}

/* NTT_Goldilocks::INTT (ntt_goldilocks.cpp:307-324) = forward network, then row x goes to row
 * (N-x) mod N (intt_idx, ntt_goldilocks.hpp:38-46) scaled by powTwoInv[log N] (:139-146) or, if
 * extend, by r_[dst] = 7^dst / N (:124-135; computeR ntt_goldilocks.hpp:153-165). */
void oracle_intt(u64 *dst, const u64 *src, u64 size, u64 ncols, int extend)
{
    if (size == 0 || ncols == 0) return;
    u64 *tmp = (u64 *)malloc(size * ncols * sizeof(u64));
    ntt_forward(tmp, src, size, ncols, size);
    u64 ninv = gl_inv(size % GL_P);
    u64 *sc = (u64 *)malloc(size * sizeof(u64));
    sc[0] = ninv;
    u64 r = 1;
    for (u64 i = 1; i < size; i++) { r = gl_mul(r, 7); sc[i] = extend ? gl_mul(r, ninv) : ninv; }
#pragma omp parallel for
    for (u64 x = 0; x < size; x++)
    {
        u64 d = (size - x) % size;
        for (u64 c = 0; c < ncols; c++) dst[d * ncols + c] = gl_mul(tmp[x * ncols + c], sc[d]);
    }
    free(sc);
    free(tmp);
}

/* NTT_Goldilocks::extendPol (ntt_goldilocks.cpp:326-352): INTT(extend=true) over N rows, then a
 * forward NTT over N_Extended rows whose input rows >= N are zero (extension>1 gather, :259-277).
 * Result: evaluations of each column's polynomial on the coset 7*<w_NExt>, natural order. */
void oracle_extend_pol(u64 *out, const u64 *in, u64 n_ext, u64 n, u64 ncols)
{
    if (n == 0 || ncols == 0 || n_ext == 0) return;
    u64 *coef = (u64 *)malloc(n * ncols * sizeof(u64));
    oracle_intt(coef, in, n, ncols, 1);
    ntt_forward(out, coef, n_ext, ncols, n);
    free(coef);
}


/* Stage 4 of Starks::genProof (src/starkpil/starks.cpp:261-285):
 *   nttExtended.INTT(qq1, q_2ns, NExt, qDim);
 *   qq2[k*qDeg + p] = qq1[p*N + k] * shiftIn^p, shiftIn = (1/7)^N, every word of the qDim-word element (:264-281,
 *   Goldilocks3::mul(Element&, Element&, Goldilocks::Element&), goldilocks_cubic_extension.hpp:1887-1892);
 *   qq2 is zero beyond element N*qDeg (:233, calloc'ed Polinomial);  nttExtended.NTT(cm4_2ns, qq2, NExt, qDim*qDeg). */
void oracle_stage4(u64 *cm4_2ns, const u64 *q_2ns, u64 n_ext, u64 n, u64 q_dim, u64 q_deg)
{
    u64 w = q_dim * q_deg;
    u64 *qq1 = (u64 *)malloc(n_ext * q_dim * sizeof(u64));
    u64 *qq2 = (u64 *)calloc(n_ext * w, sizeof(u64));
    oracle_intt(qq1, q_2ns, n_ext, q_dim, 0);
    u64 shift_in = gl_pow(gl_inv(7), n), cur = 1;
    for (u64 p = 0; p < q_deg; p++)
    {
        for (u64 k = 0; k < n; k++)
            for (u64 d = 0; d < q_dim; d++) qq2[(k * q_deg + p) * q_dim + d] = gl_mul(qq1[(p * n + k) * q_dim + d], cur);
        cur = gl_mul(cur, shift_in);
    }
    ntt_forward(cm4_2ns, qq2, n_ext, w, n_ext);
    free(qq1);
    free(qq2);
}

/* ---------------- Poseidon (poseidon_goldilocks.cpp:5-37, hash_full_result_seq) ---------------- */
static inline u64 pow7(u64 x)
{ /* poseidon_goldilocks.hpp:91-97 */
    u64 x2 = gl_mul(x, x), x3 = gl_mul(x, x2), x4 = gl_mul(x2, x2);
    return gl_mul(x3, x4);
}
static void mvp(u64 *st, const unsigned long long *mat)
{ /* poseidon_goldilocks.hpp:148-162: new[i] = sum_j mat[j][i] * old[j] */
    u64 old[12];
    memcpy(old, st, sizeof(old));
    for (int i = 0; i < 12; i++)
    {
        u64 acc = gl_mul(mat[i], old[0]);
        for (int j = 1; j < 12; j++) acc = gl_add(acc, gl_mul(mat[j * 12 + i], old[j]));
        st[i] = acc;
    }
}
void oracle_hash_full_result(u64 *out, const u64 *in)
{
    u64 st[12];
    memcpy(st, in, sizeof(st));
    for (int i = 0; i < 12; i++) st[i] = gl_add(st[i], PG_C[i]);
    for (int r = 0; r < 3; r++)
    {
        for (int i = 0; i < 12; i++) st[i] = gl_add(pow7(st[i]), PG_C[(r + 1) * 12 + i]);
        mvp(st, PG_M);
    }
    for (int i = 0; i < 12; i++) st[i] = gl_add(pow7(st[i]), PG_C[48 + i]);
    mvp(st, PG_P);
    for (int r = 0; r < 22; r++)
    {
        const unsigned long long *S = PG_S + 23 * r;
        st[0] = gl_add(pow7(st[0]), PG_C[60 + r]);
        u64 s0 = gl_mul(st[0], S[0]);
        for (int i = 1; i < 12; i++) s0 = gl_add(s0, gl_mul(st[i], S[i]));
        for (int i = 1; i < 12; i++) st[i] = gl_add(st[i], gl_mul(st[0], S[11 + i]));
        st[0] = s0;
    }
    for (int r = 0; r < 3; r++)
    {
        for (int i = 0; i < 12; i++) st[i] = gl_add(pow7(st[i]), PG_C[82 + r * 12 + i]);
        mvp(st, PG_M);
    }
    for (int i = 0; i < 12; i++) st[i] = pow7(st[i]);
    mvp(st, PG_M);
    memcpy(out, st, sizeof(st));
}

/* poseidon_goldilocks.cpp:38-74 / :229-265 — sponge, rate 8, capacity 4; size<=4 is NOT hashed. */
void oracle_linear_hash(u64 *out4, const u64 *in, u64 size)
{
    if (size <= 4)
    {
        memcpy(out4, in, size * sizeof(u64));
        memset(out4 + size, 0, (4 - size) * sizeof(u64));
        return;
    }
    u64 st[12];
    u64 remaining = size;
    while (remaining)
    {
        if (remaining == size) memset(st + 8, 0, 4 * sizeof(u64));
        else memcpy(st + 8, st, 4 * sizeof(u64));
        u64 n = remaining < 8 ? remaining : 8;
        memset(st + n, 0, (8 - n) * sizeof(u64));
        memcpy(st, in + (size - remaining), n * sizeof(u64));
        oracle_hash_full_result(st, st);
        remaining -= n;
    }
    memcpy(out4, st, 4 * sizeof(u64));
}

/* node levels, poseidon_goldilocks.cpp:283-300 (identical in _seq :92-111 and _batch) */
static void merkle_levels(u64 *tree, u64 num_rows)
{
    u64 pending = num_rows, next_n = (pending - 1) / 2 + 1, next_index = 0;
    while (pending > 1)
    {
#pragma omp parallel for
        for (u64 i = 0; i < next_n; i++)
        {
            u64 st[12];
            memcpy(st, tree + next_index + i * 8, 8 * sizeof(u64));
            memset(st + 8, 0, 4 * sizeof(u64));
            oracle_hash_full_result(st, st);
            memcpy(tree + next_index + (pending + i) * 4, st, 4 * sizeof(u64));
        }
        next_index += pending * 4;
        pending /= 2;
        next_n = (pending - 1) / 2 + 1;
    }
}

/* PoseidonGoldilocks::merkletree_avx (poseidon_goldilocks.cpp:266-301) */
void oracle_merkletree(u64 *tree, const u64 *input, u64 num_cols, u64 num_rows, u64 dim)
{
    if (num_rows == 0) return;
#pragma omp parallel for
    for (u64 i = 0; i < num_rows; i++) oracle_linear_hash(tree + i * 4, input + i * num_cols * dim, num_cols * dim);
    merkle_levels(tree, num_rows);
}

/* PoseidonGoldilocks::merkletree_batch_avx (poseidon_goldilocks.cpp:302-352) */
void oracle_merkletree_batch(u64 *tree, const u64 *input, u64 num_cols, u64 num_rows, u64 batch_size, u64 dim)
{
    if (num_rows == 0) return;
    u64 nbatches = 1;
    if (num_cols > 0) nbatches = (num_cols + batch_size - 1) / batch_size;
    u64 nlastb = num_cols - (nbatches - 1) * batch_size;
#pragma omp parallel for
    for (u64 i = 0; i < num_rows; i++)
    {
        u64 *buff = (u64 *)malloc(nbatches * 4 * sizeof(u64));
        for (u64 j = 0; j < nbatches; j++)
        {
            u64 nn = (j == nbatches - 1) ? nlastb : batch_size;
            oracle_linear_hash(buff + j * 4, input + i * num_cols * dim + j * batch_size * dim, nn * dim);
        }
        oracle_linear_hash(tree + i * 4, buff, nbatches * 4);
        free(buff);
    }
    merkle_levels(tree, num_rows);
}

/* the synthetic-input generator shared by tests and bench:
 *   in[i] = splitmix64(seed + i) mod p   — uniform, canonical. */
static inline u64 splitmix64(u64 x)
{
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}
void oracle_set_threads(int n)
{
#ifdef _OPENMP
    omp_set_dynamic(0);
    omp_set_num_threads(n > 0 ? n : 1);
#else
    (void)n;
#endif
}
void oracle_fill(u64 *x, u64 n, u64 seed, u64 offset)
{
#pragma omp parallel for
    for (u64 i = 0; i < n; i++) x[i] = splitmix64(seed + offset + i) % GL_P;
}


/* column block [c0, c0+w) of the row-major synthetic trace of ncols columns: dst[i*w + j] = gen(i*ncols + c0 + j) */
void oracle_fill_block(u64 *dst, u64 nrows, u64 c0, u64 w, u64 ncols, u64 seed)
{
#pragma omp parallel for
    for (u64 i = 0; i < nrows; i++)
        for (u64 j = 0; j < w; j++) dst[i * w + j] = splitmix64(seed + i * ncols + c0 + j) % GL_P;
}

/* The matrix of the reference's Merkle known-answer tests (tests.cpp:1887-1900, benchs/bench.cpp:409-424): rows 0 and 1
 * are 1..ncols, row j = row j-2 + row j-1.  Canonical words (the reference's are Goldilocks::add results, equal mod p). */
void oracle_merkle_kat_matrix(u64 *m, u64 nrows, u64 ncols)
{
    for (u64 i = 0; i < ncols; i++)
    {
        m[i] = i + 1;
        if (nrows > 1) m[ncols + i] = i + 1;
    }
    for (u64 j = 2; j < nrows; j++)
        for (u64 i = 0; i < ncols; i++) m[j * ncols + i] = gl_canon(gl_add(m[(j - 2) * ncols + i], m[(j - 1) * ncols + i]));
}

/* ---------------- FRI folding step (src/starkpil/fri/friProve.cpp:26-107) ----------------
 * Elements of the cubic extension F_p[x]/(x^3 - x - 1) are 3 consecutive words.
 * Polinomial::mulElement, dim 3 x dim 3 (src/starkpil/polinomial.hpp:193-206). */
static void gl3_mul(u64 *o, const u64 *a, const u64 *b)
{
    u64 A = gl_mul(gl_add(a[0], a[1]), gl_add(b[0], b[1]));
    u64 B = gl_mul(gl_add(a[0], a[2]), gl_add(b[0], b[2]));
    u64 C = gl_mul(gl_add(a[1], a[2]), gl_add(b[1], b[2]));
    u64 D = gl_mul(a[0], b[0]), E = gl_mul(a[1], b[1]), F = gl_mul(a[2], b[2]);
    u64 G = gl_sub(D, E);
    u64 o0 = gl_sub(gl_add(C, G), F);
    u64 o1 = gl_sub(gl_sub(gl_sub(gl_add(A, C), E), E), D);
    u64 o2 = gl_sub(B, G);
    o[0] = o0; o[1] = o1; o[2] = o2;
}

/* One folding step for si > 0: for every g < pol2N = 2^next_bits
 *   ppar[i] = friPol[i*pol2N + g], i < nX = 2^(pol_bits - next_bits)          (friProve.cpp:95-98)
 *   ppar_c = INTT_nX(ppar)                                                      (:99-101)
 *   ppar_c[i] *= sinv_g^i, sinv_g = shiftInv * w(polBits)^(-g)  (polMulAxi :192-200; sinv walk :36, :64-81, :105)
 *   pol2_e[g] = ppar_c evaluated at special_x by Horner                        (evalPol :201-217)
 * out: pol2N x 3 words. */
void oracle_fri_fold(u64 *pol2_e, const u64 *fri_pol, u64 pol_bits, u64 next_bits, u64 shift_inv, const u64 *special_x)
{
    const u64 pol2n = (u64)1 << next_bits, nx = (u64)1 << (pol_bits - next_bits);
    const u64 wi = gl_inv(gl_w((unsigned)pol_bits));
    u64 *ppar = (u64 *)malloc(nx * 3 * sizeof(u64)), *ppar_c = (u64 *)malloc(nx * 3 * sizeof(u64));
    u64 sinv = shift_inv;
    for (u64 g = 0; g < pol2n; g++)
    {
        for (u64 i = 0; i < nx; i++)
            for (int k = 0; k < 3; k++) ppar[i * 3 + k] = fri_pol[(i * pol2n + g) * 3 + k];
        oracle_intt(ppar_c, ppar, nx, 3, 0);
        u64 r = 1;
        for (u64 i = 0; i < nx; i++)
        {
            for (int k = 0; k < 3; k++) ppar_c[i * 3 + k] = gl_mul(ppar_c[i * 3 + k], r);
            r = gl_mul(r, sinv);
        }
        u64 res[3] = {ppar_c[(nx - 1) * 3], ppar_c[(nx - 1) * 3 + 1], ppar_c[(nx - 1) * 3 + 2]};
        for (u64 i = nx - 1; i-- > 0;)
        {
            u64 aux[3];
            gl3_mul(aux, res, special_x);
            for (int k = 0; k < 3; k++) res[k] = gl_add(aux[k], ppar_c[i * 3 + k]);
        }
        for (int k = 0; k < 3; k++) pol2_e[g * 3 + k] = res[k];
        sinv = gl_mul(sinv, wi);
    }
    free(ppar);
    free(ppar_c);
}

/* FRIProve::getTransposed (friProve.cpp:252-271): aux[i*h + j] = pol[j*w + i], w = 2^transpose_bits, h = degree / w */
void oracle_fri_transpose(u64 *aux, const u64 *pol, u64 degree, u64 transpose_bits)
{
    const u64 w = (u64)1 << transpose_bits, h = degree / w;
    for (u64 i = 0; i < w; i++)
        for (u64 j = 0; j < h; j++)
            for (int k = 0; k < 3; k++) aux[(i * h + j) * 3 + k] = pol[(j * w + i) * 3 + k];
}
