// gl64.cuh — Goldilocks field (p = 2^64 - 2^32 + 1) device arithmetic for sm_100a.
//
// Replaces the reference's x86 inline-asm scalar ops (src/goldilocks/src/goldilocks_base_field_scalar.hpp:
// add :16-35, sub :62-80, mul :103-160) and Goldilocks::toU64 (goldilocks_base_field_tools.hpp:152-163).
// The reference leaves words un-normalised in [0, 2^64); every consumer compares through toU64.  Here:
//   * gl_mul      -> canonical word (< p)
//   * gl_mul_nc   -> word in [0, 2^64) congruent to a*b (two instructions cheaper; for mul-only chains)
//   * gl_add(a,b) -> requires b < p  (or a < p); result in [0, 2^64)
//   * gl_sub(a,b) -> requires b <= p;            result in [0, 2^64)
//   * gl_canon    -> toU64
// Everything the library stores to caller-visible memory is canonical.
//
// Built from 32-bit IMAD.WIDE / IADD3 carry chains (there is no 64-bit integer multiplier on the SM):
// product = 4 IMAD.WIDE.U32 + 3 carry ops, reduction uses 2^64 = 2^32 - 1 and 2^96 = -1 (mod p).
#pragma once
#include <cstdint>

typedef uint64_t u64;
typedef uint32_t u32;

#define GL_P 0xFFFFFFFF00000001ull
#define GL_EPS 0xFFFFFFFFu

namespace gl {

__host__ __device__ __forceinline__ u64 pack(u32 lo, u32 hi) { return ((u64)hi << 32) | lo; }

#ifdef __CUDACC__
#define GL_DEV __device__ __forceinline__

// x = (r1:r0) + r2*(2^32-1) - r3  folded into [0, 2^64).   [hi word = r3:r2]
// CANON: additionally x < p.
// 0/1 carry -> 0/0xFFFFFFFF mask.  bfe.s32 (SGXT, ALU pipe) keeps the negation off the multiplier pipe, where ptxas
// would otherwise place it as IMAD.MOV.
#ifndef ZK_MASK_BFE
#define ZK_MASK_BFE 0
#endif
#if ZK_MASK_BFE
#define ZK_NEG_MASK "bfe.s32 c, c, 0, 1;\n\t"
#else
#define ZK_NEG_MASK "sub.u32 c, 0, c;\n\t"
#endif
#ifndef ZK_RED_ALU
#define ZK_RED_ALU 0
#endif
template <bool CANON>
GL_DEV u64 reduce128(u32 r0, u32 r1, u32 r2, u32 r3)
{
    u32 x0, x1;
    if (CANON)
    {
        asm("{\n\t.reg .u32 m, c, u0, u1;\n\t.reg .pred q;\n\t"
            "sub.cc.u32 %0, %2, %5;\n\t"   // t = lo - r3
            "subc.cc.u32 %1, %3, 0;\n\t"
            "subc.u32 m, 0, 0;\n\t"        // m = borrow ? 0xFFFFFFFF : 0
            "sub.cc.u32 %0, %0, m;\n\t"    // t -= eps  (lo + 1, hi - 1 + carry)  when borrow
            "subc.u32 %1, %1, 0;\n\t"
            "mad.lo.cc.u32 %0, %4, 0xFFFFFFFF, %0;\n\t" // r = t + r2*eps
            "madc.hi.cc.u32 %1, %4, 0xFFFFFFFF, %1;\n\t"
            "addc.u32 c, 0, 0;\n\t"        // c = carry out of r
            "add.cc.u32 u0, %0, 0xFFFFFFFF;\n\t" // u = r + eps = r - p (mod 2^64)
            "addc.cc.u32 u1, %1, 0;\n\t"
            "addc.u32 c, c, 0;\n\t"        // c != 0  <=>  carry (r wrapped) or r >= p
            "setp.ne.u32 q, c, 0;\n\t"
            "selp.u32 %0, u0, %0, q;\n\t"
            "selp.u32 %1, u1, %1, q;\n\t}"
            : "=&r"(x0), "=&r"(x1)
            : "r"(r0), "r"(r1), "r"(r2), "r"(r3));
    }
#if ZK_RED_ALU
    else
    {
        // r2 * eps = (r2 << 32) - r2 spelled with subtract/add only: nothing for ptxas to put on the multiplier pipe
        asm("{\n\t.reg .u32 m, c, u0, u1;\n\t"
            "sub.cc.u32 %0, %2, %5;\n\t"
            "subc.cc.u32 %1, %3, 0;\n\t"
            "subc.u32 m, 0, 0;\n\t"
            "sub.cc.u32 %0, %0, m;\n\t"
            "subc.u32 %1, %1, 0;\n\t"
            "sub.cc.u32 u0, 0, %4;\n\t"
            "subc.u32 u1, %4, 0;\n\t"
            "add.cc.u32 %0, %0, u0;\n\t"
            "addc.cc.u32 %1, %1, u1;\n\t"
            "addc.u32 c, 0, 0;\n\t"
            ZK_NEG_MASK
            "add.cc.u32 %0, %0, c;\n\t"
            "addc.u32 %1, %1, 0;\n\t}"
            : "=&r"(x0), "=&r"(x1)
            : "r"(r0), "r"(r1), "r"(r2), "r"(r3));
    }
#else
    else
    {
        asm("{\n\t.reg .u32 m, c;\n\t"
            "sub.cc.u32 %0, %2, %5;\n\t"
            "subc.cc.u32 %1, %3, 0;\n\t"
            "subc.u32 m, 0, 0;\n\t"
            "sub.cc.u32 %0, %0, m;\n\t"
            "subc.u32 %1, %1, 0;\n\t"
            "mad.lo.cc.u32 %0, %4, 0xFFFFFFFF, %0;\n\t"
            "madc.hi.cc.u32 %1, %4, 0xFFFFFFFF, %1;\n\t"
            "addc.u32 c, 0, 0;\n\t"
            ZK_NEG_MASK
            "add.cc.u32 %0, %0, c;\n\t"    // r += eps when carry
            "addc.u32 %1, %1, 0;\n\t}"
            : "=&r"(x0), "=&r"(x1)
            : "r"(r0), "r"(r1), "r"(r2), "r"(r3));
    }
#endif
    return pack(x0, x1);
}

template <bool CANON>
GL_DEV u64 mul_t(u64 a, u64 b)
{
    unsigned __int128 pr = (unsigned __int128)a * b;
    u64 lo = (u64)pr, hi = (u64)(pr >> 64);
    return reduce128<CANON>((u32)lo, (u32)(lo >> 32), (u32)hi, (u32)(hi >> 32));
}
GL_DEV u64 mul(u64 a, u64 b) { return mul_t<true>(a, b); }
GL_DEV u64 mul_nc(u64 a, u64 b) { return mul_t<false>(a, b); }
#ifndef ZK_SQR3
#define ZK_SQR3 0
#endif
#if ZK_SQR3
// a^2 with three products: a0^2 + 2 a0 a1 2^32 + a1^2 2^64; the doubling is a funnel shift (ALU pipe)
GL_DEV u64 sqr_nc(u64 a)
{
    const u32 a0 = (u32)a, a1 = (u32)(a >> 32);
    const u64 p0 = (u64)a0 * a0, p1 = (u64)a0 * a1, p3 = (u64)a1 * a1;
    const u32 d0 = (u32)p1 << 1, d1 = __funnelshift_l((u32)p1, (u32)(p1 >> 32), 1), d2 = (u32)(p1 >> 63);
    u32 r1, r2, r3;
    asm("add.cc.u32 %0, %3, %4;\n\t"
        "addc.cc.u32 %1, %5, %6;\n\t"
        "addc.u32 %2, %7, %8;"
        : "=r"(r1), "=&r"(r2), "=&r"(r3)
        : "r"((u32)(p0 >> 32)), "r"(d0), "r"((u32)p3), "r"(d1), "r"((u32)(p3 >> 32)), "r"(d2));
    return reduce128<false>((u32)p0, r1, r2, r3);
}
#else
GL_DEV u64 sqr_nc(u64 a) { return mul_t<false>(a, a); }
#endif

// a + b, requires min(a,b) < p (then a single 2^64 -> eps fold cannot overflow again)
GL_DEV u64 add(u64 a, u64 b)
{
    u32 s0, s1;
    asm("{\n\t.reg .u32 c;\n\t"
        "add.cc.u32 %0, %2, %4;\n\t"
        "addc.cc.u32 %1, %3, %5;\n\t"
        "addc.u32 c, 0, 0;\n\t"
        ZK_NEG_MASK
        "add.cc.u32 %0, %0, c;\n\t"
        "addc.u32 %1, %1, 0;\n\t}"
        : "=&r"(s0), "=&r"(s1)
        : "r"((u32)a), "r"((u32)(a >> 32)), "r"((u32)b), "r"((u32)(b >> 32)));
    return pack(s0, s1);
}
// a - b, requires b <= p
GL_DEV u64 sub(u64 a, u64 b)
{
    u32 s0, s1;
    asm("{\n\t.reg .u32 m;\n\t"
        "sub.cc.u32 %0, %2, %4;\n\t"
        "subc.cc.u32 %1, %3, %5;\n\t"
        "subc.u32 m, 0, 0;\n\t"
        "sub.cc.u32 %0, %0, m;\n\t"
        "subc.u32 %1, %1, 0;\n\t}"
        : "=&r"(s0), "=&r"(s1)
        : "r"((u32)a), "r"((u32)(a >> 32)), "r"((u32)b), "r"((u32)(b >> 32)));
    return pack(s0, s1);
}
GL_DEV u64 canon(u64 x) { return x >= GL_P ? x - GL_P : x; }

// a * 2^K mod p for 0 < K < 96, canonical result; a may be any 64-bit word.  2 is a 192-th root of unity in this
// field (2^96 = -1), so the twiddles of every transform of size <= 64 — and in particular w_16 = 2^12, which the
// in-register radix-16 networks use — are shifts: no multiplier instruction at all.
// With x = a << (K mod 32) = (x2:x1:x0) the product is  sum x_i * 2^(32 (i + K/32))  and 2^64 = 2^32 - 1, 2^96 = -1,
// 2^128 = -2^32.
template <int K>
GL_DEV u64 mul_pow2(u64 a)
{
    static_assert(K > 0 && K < 96, "shift out of range");
    constexpr int q = K / 32, s = K % 32;
    const u32 a0 = (u32)a, a1 = (u32)(a >> 32);
    u32 x0, x1, x2;
    if (s == 0)
    {
        x0 = a0;
        x1 = a1;
        x2 = 0;
    }
    else
    {
        x0 = a0 << s;
        x1 = __funnelshift_l(a0, a1, s);
        x2 = a1 >> (32 - s);
    }
    if (q == 0) return reduce128<true>(x0, x1, x2, 0u);      // (x1:x0) + x2 * eps
    if (q == 1) return reduce128<true>(0u, x0, x1, x2);      // x0 * 2^32 + x1 * eps - x2
    // q == 2:  x0 * eps - x1 - x2 * 2^32  =  (x0:0) - ((x2:x1) + x0);  the minuend is <= p - 1 and the subtrahend < 2^53,
    // so one conditional "+ p" (= - eps after the 2^64 wrap) gives the canonical value
    u32 r0, r1;
    asm("{\n\t.reg .u32 d0, d1, m;\n\t"
        "add.cc.u32 d0, %3, %2;\n\t"   // d = (x2:x1) + x0
        "addc.u32 d1, %4, 0;\n\t"
        "sub.cc.u32 %0, 0, d0;\n\t"    // (x0:0) - d
        "subc.cc.u32 %1, %2, d1;\n\t"
        "subc.u32 m, 0, 0;\n\t"        // borrow ? 0xFFFFFFFF : 0
        "sub.cc.u32 %0, %0, m;\n\t"    // - eps when it wrapped
        "subc.u32 %1, %1, 0;\n\t}"
        : "=&r"(r0), "=&r"(r1)
        : "r"(x0), "r"(x1), "r"(x2));
    return pack(r0, r1);
}
// -x for canonical x
GL_DEV u64 neg(u64 x) { return x ? GL_P - x : 0; }

#undef GL_DEV
#endif // __CUDACC__

// ---- host/device constexpr-ish helpers (table building, host-side constants) ----
__host__ __device__ inline u64 mul_slow(u64 a, u64 b)
{
    unsigned __int128 pr = (unsigned __int128)a * b;
    return (u64)(pr % GL_P);
}
__host__ __device__ inline u64 pow_slow(u64 b, u64 e)
{
    u64 r = 1;
    b %= GL_P;
    while (e)
    {
        if (e & 1) r = mul_slow(r, b);
        b = mul_slow(b, b);
        e >>= 1;
    }
    return r;
}
// primitive 2^32-th root of unity of the reference's table: Goldilocks::W[32]
// (src/goldilocks/src/goldilocks_base_field.cpp:11-44); W[k] = W32^(2^(32-k)).
#define GL_W32 7277203076849721926ull
#define GL_SHIFT 7ull // Goldilocks::SHIFT, goldilocks_base_field.cpp:9 (coset generator)
__host__ __device__ inline u64 root_of_unity(unsigned log2n)
{
    u64 w = GL_W32;
    for (unsigned i = 32; i > log2n; --i) w = mul_slow(w, w);
    return w;
}

} // namespace gl
