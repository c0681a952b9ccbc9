/* oracle build glue (TEST INFRASTRUCTURE, not product code).
 * Minimal stand-in for <gmp.h>: the GMP development headers are absent in this image, and the
 * reference's NTT_Goldilocks constructor (src/goldilocks/src/ntt_goldilocks.hpp:61-135) only
 * needs a handful of mpz_* calls on values below 2^65.  They are restated on unsigned __int128.
 * None of these values reach NTT/Merkle outputs except 2^-1 mod p (mpz_invert). */
#ifndef ZKGPU_ORACLE_GMP_SHIM_H
#define ZKGPU_ORACLE_GMP_SHIM_H
#include <stdint.h>
#include <stddef.h>
typedef unsigned __int128 zk_u128;
typedef struct { zk_u128 v; } __mpz_struct;
typedef __mpz_struct mpz_t[1];
static inline void mpz_init(mpz_t x) { x->v = 0; }
static inline void mpz_clear(mpz_t x) { (void)x; }
static inline void mpz_set_ui(mpz_t x, unsigned long u) { x->v = u; }
static inline void mpz_set(mpz_t x, const mpz_t y) { x->v = y->v; }
static inline unsigned long mpz_get_ui(const mpz_t x) { return (unsigned long)x->v; }
static inline void mpz_add_ui(mpz_t r, const mpz_t a, unsigned long u) { r->v = a->v + u; }
static inline void mpz_fdiv_q_2exp(mpz_t r, const mpz_t a, unsigned long b) { r->v = a->v >> b; }
static inline int mpz_tstbit(const mpz_t a, unsigned long b) { return (int)((a->v >> b) & 1); }
static inline int mpz_cmp_ui(const mpz_t a, unsigned long u) { return a->v < u ? -1 : (a->v > u ? 1 : 0); }
/* only the (count=1, order=1, size=8, endian=0, nails=0) form is used by the reference */
static inline void mpz_import(mpz_t r, size_t count, int order, size_t size, int endian, size_t nails, const void *op)
{ (void)count; (void)order; (void)size; (void)endian; (void)nails; r->v = *(const uint64_t *)op; }
static inline zk_u128 zk_mulmod(zk_u128 a, zk_u128 b, zk_u128 m)
{ /* a,b < m < 2^65: double-and-add keeps everything below 2^67 */
    zk_u128 r = 0; a %= m;
    while (b) { if (b & 1) { r += a; if (r >= m) r -= m; } a += a; if (a >= m) a -= m; b >>= 1; }
    return r;
}
static inline void mpz_powm(mpz_t r, const mpz_t b, const mpz_t e, const mpz_t m)
{
    zk_u128 acc = 1 % m->v, base = b->v % m->v, ex = e->v;
    while (ex) { if (ex & 1) acc = zk_mulmod(acc, base, m->v); base = zk_mulmod(base, base, m->v); ex >>= 1; }
    r->v = acc;
}
/* modulus is prime wherever the reference calls this: Fermat inverse */
static inline int mpz_invert(mpz_t r, const mpz_t a, const mpz_t m)
{
    __mpz_struct e[1]; e->v = m->v - 2; mpz_powm(r, a, e, m); return 1;
}
#endif
