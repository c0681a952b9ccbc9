// poseidon.cu — Poseidon-Goldilocks permutation, row sponge (linear_hash) and Merkle tree kernels (sm_100a).
//
// Replaces PoseidonGoldilocks::{hash_full_result, linear_hash, merkletree_avx, merkletree_batch_avx}
// (reference src/goldilocks/src/poseidon_goldilocks.cpp:166-228, 229-265, 266-301, 302-352; scalar statement of
// the same math :5-112) and the AVX2 field helpers they use (goldilocks_base_field_avx.hpp:176-573).
//
// One thread owns one 12-word state in registers.  The permutation (4 full + 22 partial + 4 full rounds, x^7) is
// evaluated in its DENSE form — every round is "S-box layer, add constants, multiply by the MDS matrix" — which is
// algebraically the reference's "optimised" sparse form (same outputs word for word) but maps to the SM differently:
//   * S-box: 4 un-normalised 64-bit modular multiplies (gl::mul_nc) on the integer-multiply pipe;
//   * MDS (entries <= 41): the state is split in 32-bit halves, converted to doubles, and the circulant product is done
//     EXACTLY on the FP64 pipe through a size-4 DFT over the Gaussian integers (78 DADD/DFMA per half-vector); one
//     fold per output brings the 74-bit result back to a 64-bit word;
//   * partial rounds go in pairs with the eleven passive words kept in the FP / DFT domain between the two products.
// The sparse form on the integer pipes only (ZK_POS_FP64=0: IMAD.WIDE MDS, 160-bit accumulators for the dense P and
// the sparse rows) is kept for comparison (tools/pos_bench.cu): 849 against 1438 Mperm/s.
// Constants live in __constant__ memory (every thread of a warp reads the same word: broadcast).
#include "gl64.cuh"
#include "zk_internal.h"
#include "poseidon_constants.inc"
#include "../../include/zkgpu.h"

namespace zk {

__constant__ u64 c_C[118];
__constant__ u64 c_PT[144]; // transposed: c_PT[i*12 + j] = P[j][i]
__constant__ u64 c_RC[96]; // constants added after the S-box layer of full round r = 0..7 (row 7 is zero)
__constant__ u64 c_S[507];

// MDS as the reference applies it: new[i] = sum_j M[j][i] * old[j]  (poseidon_goldilocks.hpp:148-162).
// M is circulant: M[j][i] = circ[(i - j) mod 12] plus 8 on M[0][0].
__host__ __device__ constexpr u32 mds_entry(int j, int i)
{
    constexpr u32 row0[12] = {17, 20, 34, 18, 39, 13, 13, 28, 2, 16, 41, 15}; // M[0][i] (without the +8 at [0][0])
    return row0[(i - j + 12) % 12] + ((i == 0 && j == 0) ? 8u : 0u);
}

// a*b + c folded to [0, 2^64)   (a*b + c < 2^128)
__device__ __forceinline__ u64 mul_add_nc(u64 a, u64 b, u64 c)
{
    unsigned __int128 pr = (unsigned __int128)a * b + c;
    u64 l = (u64)pr, h = (u64)(pr >> 64);
    return gl::reduce128<false>((u32)l, (u32)(l >> 32), (u32)h, (u32)(h >> 32));
}

// x^7 + c with the round constant riding in the addend of the last product (one fold instead of a separate add)
__device__ __forceinline__ u64 pow7_add(u64 x, u64 c)
{
    u64 x2 = gl::sqr_nc(x);
    u64 x3 = gl::mul_nc(x, x2);
    u64 x4 = gl::sqr_nc(x2);
    return mul_add_nc(x3, x4, c);
}

__device__ __forceinline__ u64 pow7(u64 x)
{
    u64 x2 = gl::sqr_nc(x);
    u64 x3 = gl::mul_nc(x, x2);
    u64 x4 = gl::sqr_nc(x2);
    return gl::mul_nc(x3, x4);
}

// x = lo64 + top * 2^64 with top < 2^32  ->  [0, 2^64)
__device__ __forceinline__ u64 fold96(u32 r0, u32 r1, u32 r2)
{
    u32 x0, x1;
    asm("{\n\t.reg .u32 c;\n\t"
        "mad.lo.cc.u32 %0, %4, 0xFFFFFFFF, %2;\n\t"
        "madc.hi.cc.u32 %1, %4, 0xFFFFFFFF, %3;\n\t"
        "addc.u32 c, 0, 0;\n\t"
        ZK_NEG_MASK
        "add.cc.u32 %0, %0, c;\n\t"
        "addc.u32 %1, %1, 0;\n\t}"
        : "=&r"(x0), "=&r"(x1)
        : "r"(r0), "r"(r1), "r"(r2));
    return gl::pack(x0, x1);
}

// st <- M^T-style product of the reference's mvp_ with the small MDS.
// Written with explicit mad.wide.u32 (one IMAD.WIDE per 32-bit half and matrix entry, immediates from the template
// recursion): left to itself nvcc strength-reduces the constant multiplies into shift/LEA/uniform-add sequences that
// came to 700 SASS instructions per call (profiles/r1_poseidon_notes.md); this form is 288 IMAD.WIDE + 8 per output.
template <int I, int J>
struct MdsDot
{
    static __device__ __forceinline__ void run(u64 &al, u64 &ah, const u32 (&lo)[12], const u32 (&hi)[12])
    {
        asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(al) : "r"(lo[J]), "n"(mds_entry(J, I)));
        asm("mad.wide.u32 %0, %1, %2, %0;" : "+l"(ah) : "r"(hi[J]), "n"(mds_entry(J, I)));
        MdsDot<I, J + 1>::run(al, ah, lo, hi);
    }
};
template <int I>
struct MdsDot<I, 12>
{
    static __device__ __forceinline__ void run(u64 &, u64 &, const u32 (&)[12], const u32 (&)[12]) {}
};
// value = al + ah * 2^32 with al, ah < 2^42  ->  [0, 2^64)
__device__ __forceinline__ u64 mds_combine(u64 al, u64 ah)
{
    u32 r1, r2;
    asm("add.cc.u32 %0, %2, %3;\n\taddc.u32 %1, %4, 0;" : "=r"(r1), "=r"(r2) : "r"((u32)(al >> 32)), "r"((u32)ah), "r"((u32)(ah >> 32)));
    return fold96((u32)al, r1, r2);
}
template <int I>
struct MdsRow
{
    static __device__ __forceinline__ void run(u64 (&st)[12], const u32 (&lo)[12], const u32 (&hi)[12])
    {
        u64 al = 0, ah = 0;
        MdsDot<I, 0>::run(al, ah, lo, hi);
        st[I] = mds_combine(al, ah);
        MdsRow<I + 1>::run(st, lo, hi);
    }
};
template <>
struct MdsRow<12>
{
    static __device__ __forceinline__ void run(u64 (&)[12], const u32 (&)[12], const u32 (&)[12]) {}
};
#ifndef ZK_MDS_ASM
#define ZK_MDS_ASM 1
#endif
#ifndef ZK_SBOX_ROLL
#define ZK_SBOX_ROLL 0
#endif
__device__ __forceinline__ void mds_small(u64 (&st)[12])
{
    u32 lo[12], hi[12];
#pragma unroll
    for (int j = 0; j < 12; j++)
    {
        lo[j] = (u32)st[j];
        hi[j] = (u32)(st[j] >> 32);
    }
#if ZK_MDS_ASM
    MdsRow<0>::run(st, lo, hi);
#else
#pragma unroll
    for (int i = 0; i < 12; i++)
    {
        u64 al = 0, ah = 0;
#pragma unroll
        for (int j = 0; j < 12; j++)
        {
            al += (u64)lo[j] * mds_entry(j, i);
            ah += (u64)hi[j] * mds_entry(j, i);
        }
        st[i] = mds_combine(al, ah);
    }
#endif
}

// 160-bit accumulator for sums of <= 2^32 products of 64-bit words
struct Acc160
{
    u32 w0, w1, w2, w3, w4;
    __device__ __forceinline__ void zero() { w0 = w1 = w2 = w3 = w4 = 0; }
    __device__ __forceinline__ void mac(u64 a, u64 b)
    {
        unsigned __int128 pr = (unsigned __int128)a * b;
        u64 l = (u64)pr, h = (u64)(pr >> 64);
        asm("add.cc.u32 %0, %0, %5;\n\t"
            "addc.cc.u32 %1, %1, %6;\n\t"
            "addc.cc.u32 %2, %2, %7;\n\t"
            "addc.cc.u32 %3, %3, %8;\n\t"
            "addc.u32 %4, %4, 0;"
            : "+r"(w0), "+r"(w1), "+r"(w2), "+r"(w3), "+r"(w4)
            : "r"((u32)l), "r"((u32)(l >> 32)), "r"((u32)h), "r"((u32)(h >> 32)));
    }
    __device__ __forceinline__ void add64(u64 a)
    {
        asm("add.cc.u32 %0, %0, %5;\n\t"
            "addc.cc.u32 %1, %1, %6;\n\t"
            "addc.cc.u32 %2, %2, 0;\n\t"
            "addc.cc.u32 %3, %3, 0;\n\t"
            "addc.u32 %4, %4, 0;"
            : "+r"(w0), "+r"(w1), "+r"(w2), "+r"(w3), "+r"(w4)
            : "r"((u32)a), "r"((u32)(a >> 32)));
    }
    // value = W + w4 * 2^128, 2^128 = -2^32 (mod p).  w4 is tiny (< number of terms).
    __device__ __forceinline__ u64 fold() const
    {
        u64 x = gl::reduce128<false>(w0, w1, w2, w3);         // [0, 2^64)
        u64 y = (u64)w4 << 32;                                  // < p (w4 < 2^31)
        return gl::sub(x, y);
    }
};

// ---------------------------------------------------------------------------------------------------
// MDS on the FP64 pipe.  The MDS entries are <= 41, so with the state split in 32-bit halves every partial sum
// sum_j M[j][i]*half_j is an integer below 2^42: exactly representable in a double.  One DFMA per matrix entry and
// half replaces one IMAD.WIDE on the integer-multiply pipe (the pipe the S-boxes saturate), and the FP64 pipe runs
// beside it.  B200 has a full-rate FP64 pipe (64 lanes/clk/SM); results are exact integers, there is no rounding
// anywhere on this path.
// ---------------------------------------------------------------------------------------------------
#ifndef ZK_POS_FP64
#define ZK_POS_FP64 1
#endif
#ifndef ZK_CONV_IN
#define ZK_CONV_IN 1 // 0: 2^52 magic (DADD), 1: I2F.F64.U32
#endif
#ifndef ZK_CONV_OUT
#define ZK_CONV_OUT 1 // 0: 2^52 magic (DADD + mask), 1: F2I.U64.F64
#endif
__device__ __forceinline__ double u32_to_f64(u32 x)
{
#if ZK_CONV_IN
    return __uint2double_rn(x);
#else
    return __hiloint2double(0x43300000, (int)x) - 4503599627370496.0;
#endif
}
// a is an exact integer in [0, 2^52)
__device__ __forceinline__ u64 f64_to_u64(double a)
{
#if ZK_CONV_OUT
    return __double2ull_rz(a);
#else
    const double t = a + 4503599627370496.0;
    return gl::pack((u32)__double2loint(t), (u32)__double2hiint(t) & 0xFFFFFu);
#endif
}
template <int I, int J>
struct MdsDotF
{
    static __device__ __forceinline__ void run(double &al, double &ah, const double (&lo)[12], const double (&hi)[12])
    {
        constexpr double m = (double)mds_entry(J, I);
        if (J == 0)
        {
            al = lo[0] * m;
            ah = hi[0] * m;
        }
        else
        {
            al = fma(lo[J], m, al);
            ah = fma(hi[J], m, ah);
        }
        MdsDotF<I, J + 1>::run(al, ah, lo, hi);
    }
};
template <int I>
struct MdsDotF<I, 12>
{
    static __device__ __forceinline__ void run(double &, double &, const double (&)[12], const double (&)[12]) {}
};
template <int I>
struct MdsRowF
{
    static __device__ __forceinline__ void run(u64 (&st)[12], const double (&lo)[12], const double (&hi)[12])
    {
        double al, ah;
        MdsDotF<I, 0>::run(al, ah, lo, hi);
        st[I] = mds_combine(f64_to_u64(al), f64_to_u64(ah));
        MdsRowF<I + 1>::run(st, lo, hi);
    }
};
template <>
struct MdsRowF<12>
{
    static __device__ __forceinline__ void run(u64 (&)[12], const double (&)[12], const double (&)[12]) {}
};
// The same product through the CRT split z^12 - 1 = prod_f (z^3 - i^f): a size-4 DFT along the stride-3 subsequences
// (i = sqrt(-1): Gaussian integers, exact), three 3-term twisted convolutions whose constants (MDS spectrum / 4) are
// all +-1 or powers of two — K0 = [16, 32, 16], K2 = [-1, -8, 2], K1 / 2 = [2 + i, -4 - i, 16 - i] — and the inverse
// DFT.  78 DADD/DFMA per half-vector instead of 144 DFMA; every intermediate is an integer below 2^45.
// y[i] = sum_j M[j][i] x[j], in place.
struct Spectrum
{
    double f0[3], f2[3], u[3], v[3]; // z^3-1 block, z^3+1 block, real and imaginary part of the z^3-i block
};
__device__ __forceinline__ void mds_dft(Spectrum &s, const double (&x)[12])
{
#pragma unroll
    for (int r = 0; r < 3; r++)
    {
        const double s02 = x[r] + x[6 + r], s13 = x[3 + r] + x[9 + r];
        s.u[r] = x[r] - x[6 + r];
        s.v[r] = x[3 + r] - x[9 + r];
        s.f0[r] = s02 + s13;
        s.f2[r] = s02 - s13;
    }
}
// the three twisted 3-term convolutions, in place; f0 comes out divided by 16 (mds_idft multiplies it back)
__device__ __forceinline__ void mds_blocks(Spectrum &s)
{
    const double T = s.f0[0] + s.f0[1] + s.f0[2];
    const double a0 = s.f0[0], a1 = s.f0[1], a2 = s.f0[2];
    s.f0[0] = T + a2;
    s.f0[1] = T + a0;
    s.f0[2] = T + a1;
    const double b0 = s.f2[0], b1 = s.f2[1], b2 = s.f2[2];
    s.f2[0] = fma(8.0, b2, fma(-2.0, b1, -b0));
    s.f2[1] = fma(-8.0, b0, fma(-2.0, b2, -b1));
    s.f2[2] = fma(2.0, b0, fma(-8.0, b1, -b2));
    const double U0 = s.u[0], U1 = s.u[1], U2 = s.u[2], V0 = s.v[0], V1 = s.v[1], V2 = s.v[2];
    s.u[0] = fma(4.0, V2, fma(-16.0, V1, fma(2.0, U0, -V0) + U1) + U2);
    s.v[0] = fma(-4.0, U2, fma(16.0, U1, fma(2.0, V0, U0) + V1) + V2);
    s.u[1] = fma(-16.0, V2, fma(2.0, U1, fma(-4.0, U0, V0) - V1) + U2);
    s.v[1] = fma(16.0, U2, fma(2.0, V1, fma(-4.0, V0, -U0) + U1) + V2);
    s.u[2] = fma(2.0, U2, fma(-4.0, U1, fma(16.0, U0, V0) + V1) - V2);
    s.v[2] = fma(2.0, V2, fma(-4.0, V1, fma(16.0, V0, -U0) - U1) + U2);
}
// x[3k + r] = C0 f0[r] + (-1)^k f2[r] + C1 Re((u[r] + i v[r]) (-i)^k);  then the diagonal term 8 x0 on word 0
template <int C0, int C1>
__device__ __forceinline__ void mds_idft(double (&x)[12], const Spectrum &s, double x0)
{
#pragma unroll
    for (int r = 0; r < 3; r++)
    {
        const double A = fma((double)C0, s.f0[r], s.f2[r]), B = fma((double)C0, s.f0[r], -s.f2[r]);
        x[r] = fma((double)C1, s.u[r], A);
        x[3 + r] = fma((double)C1, s.v[r], B);
        x[6 + r] = fma((double)-C1, s.u[r], A);
        x[9 + r] = fma((double)-C1, s.v[r], B);
    }
    x[0] = fma(8.0, x0, x[0]);
}
__device__ __forceinline__ void mds_half_fft(double (&x)[12])
{
    Spectrum s;
    const double x0 = x[0];
    mds_dft(s, x);
    mds_blocks(s);
    mds_idft<16, 1>(x, s, x0);
}
#ifndef ZK_MDS_FFT
#define ZK_MDS_FFT 1
#endif
__device__ __forceinline__ void mds_fp64(u64 (&st)[12])
{
#if ZK_MDS_FFT
    double h[12];
    u64 al[12];
#pragma unroll
    for (int j = 0; j < 12; j++) h[j] = u32_to_f64((u32)st[j]);
    mds_half_fft(h);
#pragma unroll
    for (int j = 0; j < 12; j++) al[j] = f64_to_u64(h[j]);
#pragma unroll
    for (int j = 0; j < 12; j++) h[j] = u32_to_f64((u32)(st[j] >> 32));
    mds_half_fft(h);
#pragma unroll
    for (int j = 0; j < 12; j++) st[j] = mds_combine(al[j], f64_to_u64(h[j]));
#else
    double lo[12], hi[12];
#pragma unroll
    for (int j = 0; j < 12; j++)
    {
        lo[j] = u32_to_f64((u32)st[j]);
        hi[j] = u32_to_f64((u32)(st[j] >> 32));
    }
    MdsRowF<0>::run(st, lo, hi);
#endif
}

#if ZK_POS_FP64
// Dense form of the same permutation.  The reference's sparse partial rounds (matrices P and S) are an algebraic
// refactoring of "x0 <- x0^7 + k_q ; x <- MDS x" made for CPUs, where a full 64x64 multiply costs one instruction and
// 23 of them beat 144 small-constant products.  Here the small-constant products ride the FP64 pipe, so the dense form
// (same scalar constants C[60+q], plain MDS in place of P and of every S_q: S_q = B_{q+1}^-1 M B_q with block-diagonal
// B that commute with the x0 S-box, P = B_0^-1 M, B_22 = I) costs 4 integer multiplies per partial round instead of 27.
// All 30 rounds share one loop body; outputs are word-for-word those of the sparse form (tests/test_gpu_parity.py).
#ifndef ZK_LAZY2
#define ZK_LAZY2 2
#endif
#if ZK_LAZY2
// Partial rounds go in pairs: the eleven words that skip the S-box stay in the FP domain between the two MDS products
// of a pair (halves grow 32 -> 41 -> 49 bits, still exact in a double), only word 0 makes the trip to the integer
// side for its S-box.  Saves 44 conversions and 11 recombinations per pair.
__device__ __forceinline__ void poseidon_permute_state(u64 (&st)[12])
{
#pragma unroll
    for (int i = 0; i < 12; i++) st[i] = gl::add(st[i], c_C[i]);
#pragma unroll 1
    for (int it = 0; it < 19; it++)
    {
        if (it < 4 || it >= 15)
        {
            const u64 *rc = c_RC + (it < 4 ? it : it - 11) * 12;
#pragma unroll
            for (int i = 0; i < 12; i++) st[i] = pow7_add(st[i], rc[i]);
            mds_fp64(st);
            continue;
        }
        const u64 *k = c_C + 60 + 2 * (it - 4);
        st[0] = pow7_add(st[0], k[0]);
#if ZK_LAZY2 == 2
        // ... and between the two products the state stays in the DFT domain as well: the inverse transform of the
        // first product and the forward transform of the second cancel up to powers of two (the DFT of the inverse
        // DFT of (f0, f2, u + iv) is (64 f0, 4 f2, 2 (u + iv)) in the scaling mds_blocks leaves), and replacing word 0
        // by its S-box image adds d = new - old to f0[0], f2[0] and u[0].
        Spectrum sl, sh;
        double h[12];
#pragma unroll
        for (int j = 0; j < 12; j++) h[j] = u32_to_f64((u32)st[j]);
        mds_dft(sl, h);
        mds_blocks(sl);
        const double yl = fma(16.0, sl.f0[0], sl.f2[0]) + sl.u[0]; // word 0 of the circulant part
        const double wl = fma(8.0, h[0], yl);
#pragma unroll
        for (int j = 0; j < 12; j++) h[j] = u32_to_f64((u32)(st[j] >> 32));
        mds_dft(sh, h);
        mds_blocks(sh);
        const double yh = fma(16.0, sh.f0[0], sh.f2[0]) + sh.u[0];
        const double wh = fma(8.0, h[0], yh);
        const u64 w0 = pow7_add(mds_combine(f64_to_u64(wl), f64_to_u64(wh)), k[1]);
        const double xl = u32_to_f64((u32)w0), xh = u32_to_f64((u32)(w0 >> 32));
        {
            const double d = xl - yl;
            sl.f0[0] = fma(d, 0.015625, sl.f0[0]);
            sl.f2[0] = fma(4.0, sl.f2[0], d);
            sl.f2[1] *= 4.0;
            sl.f2[2] *= 4.0;
            sl.u[0] = fma(d, 0.5, sl.u[0]);
        }
        mds_blocks(sl);
        mds_idft<1024, 2>(h, sl, xl);
        u64 al[12];
#pragma unroll
        for (int j = 0; j < 12; j++) al[j] = f64_to_u64(h[j]);
        {
            const double d = xh - yh;
            sh.f0[0] = fma(d, 0.015625, sh.f0[0]);
            sh.f2[0] = fma(4.0, sh.f2[0], d);
            sh.f2[1] *= 4.0;
            sh.f2[2] *= 4.0;
            sh.u[0] = fma(d, 0.5, sh.u[0]);
        }
        mds_blocks(sh);
        mds_idft<1024, 2>(h, sh, xh);
#pragma unroll
        for (int j = 0; j < 12; j++) st[j] = mds_combine(al[j], f64_to_u64(h[j]));
#else
        double hl[12], hh[12];
#pragma unroll
        for (int j = 0; j < 12; j++) hl[j] = u32_to_f64((u32)st[j]);
        mds_half_fft(hl);
#pragma unroll
        for (int j = 0; j < 12; j++) hh[j] = u32_to_f64((u32)(st[j] >> 32));
        mds_half_fft(hh);
        const u64 w0 = pow7_add(mds_combine(f64_to_u64(hl[0]), f64_to_u64(hh[0])), k[1]);
        hl[0] = u32_to_f64((u32)w0);
        hh[0] = u32_to_f64((u32)(w0 >> 32));
        mds_half_fft(hl);
        u64 al[12];
#pragma unroll
        for (int j = 0; j < 12; j++) al[j] = f64_to_u64(hl[j]);
        mds_half_fft(hh);
#pragma unroll
        for (int j = 0; j < 12; j++) st[j] = mds_combine(al[j], f64_to_u64(hh[j]));
#endif
    }
}
#else
__device__ __forceinline__ void poseidon_permute_state(u64 (&st)[12])
{
#pragma unroll
    for (int i = 0; i < 12; i++) st[i] = gl::add(st[i], c_C[i]);
#pragma unroll 1
    for (int r = 0; r < 30; r++)
    {
        if (r < 4 || r >= 26)
        {
            const u64 *rc = c_RC + (r < 4 ? r : r - 22) * 12;
#pragma unroll
            for (int i = 0; i < 12; i++) st[i] = pow7_add(st[i], rc[i]);
        }
        else
            st[0] = pow7_add(st[0], c_C[56 + r]); // C[60 + q], q = r - 4
        mds_fp64(st);
    }
}
#endif // ZK_LAZY2
#else
// The reference's hash_full_result on one state (poseidon_goldilocks.cpp:5-37 is the scalar statement).
// Input words may be any 64-bit value; output words are in [0, 2^64) (callers canonicalise what they store).
//
// Code-size matters as much as instruction count here: the SM's L1.5 instruction cache is 32 KB (2048 SASS
// instructions) and every warp of the SM sits at a different point of the permutation, so a fully unrolled body
// (7.4k instructions in the first version) spends ~9 cycles per issue waiting for instruction fetch
// (ncu: stall no_instruction, profiles/r1_*).  All 8 full rounds therefore share ONE loop body
//     st = pow7(st) + RC[r];  st = MDS * st
// (RC[7] = 0: the last round adds nothing), with the dense P product and the 22 sparse rounds hanging off
// iteration r == 3, and the dense P product is a rolled loop over output words.
__device__ __forceinline__ void poseidon_permute_state(u64 (&st)[12])
{
#pragma unroll
    for (int i = 0; i < 12; i++) st[i] = gl::add(st[i], c_C[i]);
#pragma unroll 1
    for (int r = 0; r < 8; r++)
    {
        const u64 *rc = c_RC + r * 12;
#if ZK_SBOX_ROLL
        // S-box layer as 3 iterations over 4 words with a register rotation: a third of the code of the unrolled
        // layer (the three hot bodies then fit the 32 KB instruction cache together) for 24 moves per iteration
#pragma unroll 1
        for (int it = 0; it < 3; it++)
        {
#pragma unroll
            for (int k = 0; k < 4; k++) st[k] = pow7_add(st[k], rc[4 * it + k]);
            u64 t0 = st[0], t1 = st[1], t2 = st[2], t3 = st[3];
#pragma unroll
            for (int k = 0; k < 8; k++) st[k] = st[k + 4];
            st[8] = t0;
            st[9] = t1;
            st[10] = t2;
            st[11] = t3;
        }
#else
#pragma unroll
        for (int i = 0; i < 12; i++) st[i] = pow7_add(st[i], rc[i]);
#endif
        if (r != 3)
        {
            mds_small(st);
            continue;
        }
        // dense P: new[i] = sum_j P[j][i] * old[j]   (c_PT[i][j] = P[j][i])
        u64 nw[12];
#pragma unroll 1
        for (int i = 0; i < 12; i++)
        {
            const u64 *pr = c_PT + i * 12;
            Acc160 acc;
            acc.zero();
#pragma unroll
            for (int j = 0; j < 12; j++) acc.mac(st[j], pr[j]);
            nw[i] = acc.fold();
        }
#pragma unroll
        for (int i = 0; i < 12; i++) st[i] = nw[i];
#pragma unroll 1
        for (int q = 0; q < 22; q++)
        {
            const u64 *S = c_S + 23 * q;
            const u64 s0 = pow7_add(st[0], c_C[60 + q]);
            Acc160 acc;
            acc.zero();
            acc.mac(s0, S[0]);
#pragma unroll
            for (int i = 1; i < 12; i++) acc.mac(st[i], S[i]);
#pragma unroll
            for (int i = 1; i < 12; i++) st[i] = mul_add_nc(s0, S[11 + i], st[i]);
            st[0] = acc.fold();
        }
    }
}

#endif // ZK_POS_FP64

// ---------------------------------------------------------------------------------------------------
// kernels
// ---------------------------------------------------------------------------------------------------
// ZK_POS_MIN_CTAS (experiment knob, tools/pos_bench.cu): minimum resident CTAs per SM asked of ptxas for the 128-thread hash
// kernels.  Unset, ptxas is free to use up to 255 registers and settles on 72 (7 CTAs per SM) — it is not squeezed.
#ifndef ZK_POS_MIN_CTAS
#define ZK_POS_MIN_CTAS 1
#endif
__global__ void __launch_bounds__(128, ZK_POS_MIN_CTAS) permute_kernel(u64 *out, const u64 *in, u64 n)
{
    const u64 t = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    u64 st[12];
#pragma unroll
    for (int i = 0; i < 12; i++) st[i] = in[t * 12 + i];
    poseidon_permute_state(st);
#pragma unroll
    for (int i = 0; i < 12; i++) out[t * 12 + i] = gl::canon(st[i]);
}

// sponge absorption of `count` (>= 1) words at p, rate 8, starting from capacity st[8..11]; on return st[0..3] is the
// running digest (un-normalised).  The last chunk is zero padded (poseidon_goldilocks.cpp:244-256).
// COPY: the words are also written to q[0..count) as they are read (the multi-GPU path assembles the row-major
// extended trace from the exchanged pieces this way instead of with a separate copy pass).
template <bool COPY = false>
__device__ __forceinline__ void sponge_absorb(u64 (&st)[12], const u64 *__restrict__ p, u64 count, u64 *__restrict__ q = nullptr)
{
    u64 done = 0;
    while (done < count)
    {
        const u64 n = count - done < 8 ? count - done : 8;
        if (n == 8)
        {
#pragma unroll
            for (int i = 0; i < 8; i++) st[i] = p[done + i];
            if (COPY)
            {
#pragma unroll
                for (int i = 0; i < 8; i++) q[done + i] = st[i];
            }
        }
        else
        {
#pragma unroll
            for (int i = 0; i < 8; i++) st[i] = (u64)i < n ? p[done + i] : 0;
            if (COPY)
            {
#pragma unroll
                for (int i = 0; i < 8; i++)
                    if ((u64)i < n) q[done + i] = st[i];
            }
        }
        poseidon_permute_state(st);
        done += n;
#pragma unroll
        for (int i = 0; i < 4; i++) st[8 + i] = st[i];
    }
}

// linear_hash of `count` words at p into out[0..3] (canonical).  poseidon_goldilocks.cpp:229-265:
// count <= 4 is copied (zero padded) and NOT hashed; otherwise sponge with rate 8, capacity = previous out[0..3].
__device__ __forceinline__ void linear_hash_words(u64 (&out)[4], const u64 *__restrict__ p, u64 count)
{
    if (count <= 4)
    {
#pragma unroll
        for (int i = 0; i < 4; i++) out[i] = (u64)i < count ? gl::canon(p[i]) : 0;
        return;
    }
    u64 st[12];
#pragma unroll
    for (int i = 8; i < 12; i++) st[i] = 0;
    sponge_absorb(st, p, count);
#pragma unroll
    for (int i = 0; i < 4; i++) out[i] = gl::canon(st[i]);
}

// One column strip of the row sponge: row r continues from the digest in leaves[4r..4r+3] (zero for the first strip)
// and absorbs strip_words words.  Every strip but the last must be a multiple of 8 words wide, so the chunking is the
// one linear_hash would use on the whole row.  Lets leaf hashing follow the LDE strip by strip (and the host->device
// copy of the next strip) instead of waiting for the whole extended trace.
template <int THREADS, bool COPY = false>
__global__ void __launch_bounds__(THREADS) absorb_kernel(u64 *leaves, const u64 *rows, u64 strip_words, u64 row_ld, u64 nrows, int first,
                                                         u64 *copy_to = nullptr, u64 copy_ld = 0)
{
    const u64 row = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
    u64 st[12];
    ulonglong2 *o = reinterpret_cast<ulonglong2 *>(leaves + row * 4);
    if (first)
    {
#pragma unroll
        for (int i = 8; i < 12; i++) st[i] = 0;
    }
    else
    {
        const ulonglong2 a = o[0], b = o[1];
        st[8] = a.x;
        st[9] = a.y;
        st[10] = b.x;
        st[11] = b.y;
    }
    sponge_absorb<COPY>(st, rows + row * row_ld, strip_words, COPY ? copy_to + row * copy_ld : nullptr);
    o[0] = make_ulonglong2(gl::canon(st[0]), gl::canon(st[1]));
    o[1] = make_ulonglong2(gl::canon(st[2]), gl::canon(st[3]));
}

__global__ void __launch_bounds__(128) leaves_kernel(u64 *leaves, const u64 *input, u64 row_words, u64 row_ld, u64 nrows)
{
    const u64 row = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
    u64 h[4];
    linear_hash_words(h, input + row * row_ld, row_words);
    ulonglong2 *o = reinterpret_cast<ulonglong2 *>(leaves + row * 4);
    o[0] = make_ulonglong2(h[0], h[1]);
    o[1] = make_ulonglong2(h[2], h[3]);
}

// merkletree_batch leaves (poseidon_goldilocks.cpp:302-352): linear_hash of every `batch`-column group, then
// linear_hash over the concatenated 4-word digests.
__global__ void __launch_bounds__(128) leaves_batch_kernel(u64 *leaves, u64 *digests, const u64 *input, u64 ncols, u64 dim,
                                                           u64 batch, u64 nbatches, u64 nrows)
{
    const u64 row = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
    u64 *dg = digests + row * nbatches * 4;
    const u64 *in = input + row * ncols * dim;
    for (u64 j = 0; j < nbatches; j++)
    {
        const u64 nn = (j == nbatches - 1) ? ncols - (nbatches - 1) * batch : batch;
        u64 h[4];
        linear_hash_words(h, in + j * batch * dim, nn * dim);
#pragma unroll
        for (int i = 0; i < 4; i++) dg[j * 4 + i] = h[i];
    }
    u64 h[4];
    linear_hash_words(h, dg, nbatches * 4);
#pragma unroll
    for (int i = 0; i < 4; i++) leaves[row * 4 + i] = h[i];
}

// one tree level: node i = hash(children 2i, 2i+1 | capacity 0)[0..3]   (poseidon_goldilocks.cpp:283-300)
__global__ void __launch_bounds__(128) level_kernel(u64 *parents, const u64 *children, u64 nparents)
{
    const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nparents) return;
    u64 st[12];
    const ulonglong2 *ch = reinterpret_cast<const ulonglong2 *>(children + i * 8);
#pragma unroll
    for (int k = 0; k < 4; k++)
    {
        ulonglong2 v = ch[k];
        st[2 * k] = v.x;
        st[2 * k + 1] = v.y;
    }
#pragma unroll
    for (int k = 8; k < 12; k++) st[k] = 0;
    poseidon_permute_state(st);
    ulonglong2 *o = reinterpret_cast<ulonglong2 *>(parents + i * 4);
    o[0] = make_ulonglong2(gl::canon(st[0]), gl::canon(st[1]));
    o[1] = make_ulonglong2(gl::canon(st[2]), gl::canon(st[3]));
}



// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------
int poseidon_init_constants()
{
    Ctx &C = *t_ctx; // once per device: __constant__ arrays exist per device
    if (C.consts_ready) return 0;
    ZK_CUDA(cudaMemcpyToSymbol(c_C, PG_C, sizeof(u64) * 118));
    {
        u64 pt[144], rc[96];
        for (int i = 0; i < 12; i++)
            for (int j = 0; j < 12; j++) pt[i * 12 + j] = PG_P[j * 12 + i];
        for (int i = 0; i < 12; i++)
        {
            for (int r = 0; r < 3; r++) rc[r * 12 + i] = PG_C[(r + 1) * 12 + i];
            rc[3 * 12 + i] = PG_C[48 + i];
            for (int r = 0; r < 3; r++) rc[(4 + r) * 12 + i] = PG_C[82 + r * 12 + i];
            rc[7 * 12 + i] = 0;
        }
        ZK_CUDA(cudaMemcpyToSymbol(c_PT, pt, sizeof(pt)));
        ZK_CUDA(cudaMemcpyToSymbol(c_RC, rc, sizeof(rc)));
    }
    ZK_CUDA(cudaMemcpyToSymbol(c_S, PG_S, sizeof(u64) * 507));
    // the kernels hard-code the circulant MDS; check it against the table the reference ships
    for (int j = 0; j < 12; j++)
        for (int i = 0; i < 12; i++)
        {
            static const u32 row0[12] = {17, 20, 34, 18, 39, 13, 13, 28, 2, 16, 41, 15};
            u32 e = row0[(i - j + 12) % 12] + ((i == 0 && j == 0) ? 8u : 0u);
            if (PG_M[j * 12 + i] != e)
            {
                set_error("poseidon: MDS table mismatch at [%d][%d]", j, i);
                return -1;
            }
        }
    C.consts_ready = true;
    return 0;
}

static inline unsigned nblocks(u64 n, unsigned per) { return (unsigned)((n + per - 1) / per); }

int poseidon_permute(u64 *out12, const u64 *in12, u64 n, cudaStream_t st)
{
    ZK_TRY(poseidon_init_constants());
    if (n == 0) return 0;
    permute_kernel<<<nblocks(n, 128), 128, 0, st>>>(out12, in12, n);
    ZK_LAUNCHED();
    ZK_CUDA(cudaGetLastError());
    return 0;
}

int merkle_leaves(u64 *leaves, const u64 *input, u64 row_words, u64 row_ld, u64 nrows, cudaStream_t st)
{
    ZK_TRY(poseidon_init_constants());
    if (nrows == 0) return 0;
    leaves_kernel<<<nblocks(nrows, 128), 128, 0, st>>>(leaves, input, row_words, row_ld, nrows);
    ZK_LAUNCHED();
    ZK_CUDA(cudaGetLastError());
    return 0;
}

int g_absorb_threads = 128;
int merkle_absorb(u64 *leaves, const u64 *rows, u64 strip_words, u64 row_ld, u64 nrows, bool first, cudaStream_t st, u64 *copy_to,
                  u64 copy_ld)
{
    ZK_TRY(poseidon_init_constants());
    if (nrows == 0 || strip_words == 0) return 0;
    if (copy_to)
    {
        absorb_kernel<128, true><<<nblocks(nrows, 128), 128, 0, st>>>(leaves, rows, strip_words, row_ld, nrows, first ? 1 : 0, copy_to, copy_ld);
        ZK_LAUNCHED();
        ZK_CUDA(cudaGetLastError());
        return 0;
    }
    // g_absorb_threads == 512: one 16-warp CTA per SM (two do not fit the register file), which leaves registers and
    // shared memory for a transform CTA to run beside it.  Measured (tools/overlap_probe.py, profiles/r1_overlap_probe.txt):
    // same sponge throughput as 128-thread CTAs, but running the LDE of the next strip beside the sponge gains only 3 %
    // over running them back to back, so the pipeline keeps them sequential.
    if (g_absorb_threads == 512)
        absorb_kernel<512><<<nblocks(nrows, 512), 512, 0, st>>>(leaves, rows, strip_words, row_ld, nrows, first ? 1 : 0);
    else
        absorb_kernel<128><<<nblocks(nrows, 128), 128, 0, st>>>(leaves, rows, strip_words, row_ld, nrows, first ? 1 : 0);
    ZK_LAUNCHED();
    ZK_CUDA(cudaGetLastError());
    return 0;
}

int merkle_leaves_batch(u64 *leaves, const u64 *input, u64 ncols, u64 dim, u64 batch, u64 nrows, cudaStream_t st)
{
    ZK_TRY(poseidon_init_constants());
    if (nrows == 0) return 0;
    if (batch == 0)
    {
        set_error("merkletree_batch: batch_size 0");
        return -1;
    }
    u64 nbatches = 1;
    if (ncols > 0) nbatches = (ncols + batch - 1) / batch;
    u64 *digests = nullptr;
    ZK_CUDA(cudaMallocAsync(&digests, nrows * nbatches * 4 * sizeof(u64), st));
    leaves_batch_kernel<<<nblocks(nrows, 128), 128, 0, st>>>(leaves, digests, input, ncols, dim, batch, nbatches, nrows);
    ZK_LAUNCHED();
    ZK_CUDA(cudaGetLastError());
    ZK_CUDA(cudaFreeAsync(digests, st));
    return 0;
}

int merkle_levels(u64 *tree, u64 nrows, cudaStream_t st)
{
    ZK_TRY(poseidon_init_constants());
    // reference loop (poseidon_goldilocks.cpp:283-300) for a power-of-two number of leaves.  Small levels are bound by the
    // latency of ONE permutation (~27 us: 30 rounds, each an S-box chain followed by an MDS chain), not by launches: a
    // cooperative 16-lanes-per-hash kernel measured 23-30 us per level in both rounds (profiles/r1_small_shape_coop_levels.txt,
    // profiles/r2_small_shapes.txt) and is not kept.
    u64 pending = nrows, next_index = 0;
    while (pending > 1)
    {
        const u64 next_n = pending / 2;
        level_kernel<<<nblocks(next_n, 128), 128, 0, st>>>(tree + next_index + pending * 4, tree + next_index, next_n);
        ZK_LAUNCHED();
        ZK_CUDA(cudaGetLastError());
        next_index += pending * 4;
        pending /= 2;
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------------
// zkgpu_merkletree_batch_noalloc: the batched Merkle tree for callers that must not allocate device memory.
//
// leaves_batch_kernel needs nrows * nbatches * 4 words of global memory for the per-batch digests of every row, beyond its
// input and its output.  merkle_leaves_batch takes them from the stream-ordered allocator on every call.  This variant
// keeps them in the module instead: a fixed array that the CUDA loader places in every device's memory when the library is
// loaded there and that stays until the process exits (8 MiB per GPU, used or not).
//
// g_static_digests is a DEVICE symbol.  Host code must not hand its name to a kernel: on the host the name stands for a
// shadow object of the runtime, not for device memory.  The device address comes from cudaGetSymbolAddress, asked once per
// device while that device is current (the array exists once per device, at a different address on each).
// ---------------------------------------------------------------------------------------------------
#define ZK_STATIC_DIGEST_WORDS ((size_t)1 << 20)
static __device__ __align__(256) u64 g_static_digests[ZK_STATIC_DIGEST_WORDS];
static u64 *s_static_digests_addr[64] = {}; // per device ordinal; written under that device's context lock

static int static_digests(Ctx &C, u64 **addr)
{
    if (C.device < 0 || C.device >= 64)
    {
        set_error("merkletree_batch_noalloc: device ordinal %d out of range", C.device);
        return -1;
    }
    if (!s_static_digests_addr[C.device])
    {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_digests); // the context's device is current (Enter)
        if (e != cudaSuccess || !p)
        {
            set_error("merkletree_batch_noalloc: cudaGetSymbolAddress failed on device %d: %s", C.device, cudaGetErrorString(e));
            cudaGetLastError();
            return -2;
        }
        s_static_digests_addr[C.device] = (u64 *)p;
    }
    *addr = s_static_digests_addr[C.device];
    return 0;
}

} // namespace zk

#ifndef ZK_NO_ENTRY_POINTS // tools/pos_bench.cu includes this file on its own, without the context code of ctx.cu
extern "C" int zkgpu_merkletree_batch_noalloc(uint64_t *tree, const uint64_t *input, uint64_t ncols, uint64_t nrows, uint64_t batch,
                                              uint64_t dim)
{
    using namespace zk;
    if (nrows == 0) return 0;
    if (nrows & (nrows - 1))
    {
        set_error("merkletree_batch_noalloc: num_rows %llu is not a power of two", (unsigned long long)nrows);
        return -1;
    }
    if (batch == 0)
    {
        set_error("merkletree_batch_noalloc: batch_size 0");
        return -1;
    }
    const u64 nbatches = ncols ? (ncols + batch - 1) / batch : 1;
    // nrows * nbatches * 4 without overflow: nrows <= 2^20 and nbatches <= 2^20 are both implied by the limit
    if (nrows > ZK_STATIC_DIGEST_WORDS || nbatches > ZK_STATIC_DIGEST_WORDS || nrows * nbatches * 4 > ZK_STATIC_DIGEST_WORDS)
    {
        set_error("merkletree_batch_noalloc: num_rows * ceil(num_cols / batch_size) * 4 = %llu * %llu * 4 words of digest scratch exceed "
                  "the built-in %llu words; use zkgpu_merkletree_batch (which allocates) for this shape",
                  (unsigned long long)nrows, (unsigned long long)nbatches, (unsigned long long)ZK_STATIC_DIGEST_WORDS);
        return -1;
    }
    ZK_ENTER;
    if (!is_device_ptr(tree) || (ncols * dim != 0 && !is_device_ptr(input)))
    {
        set_error("merkletree_batch_noalloc: tree and input must be device pointers (staging host data would need an allocation)");
        return -1;
    }
    if ((uintptr_t)tree & 15)
    {
        set_error("merkletree_batch_noalloc: the tree pointer must be 16-byte aligned");
        return -1;
    }
    u64 *digests = nullptr;
    ZK_TRY(static_digests(C, &digests));
    cudaStream_t st = C.cur;
    // every row writes its nbatches digests before it reads them back, so nothing in the array has to start at zero
    leaves_batch_kernel<<<nblocks(nrows, 128), 128, 0, st>>>(tree, digests, input, ncols, dim, batch, nbatches, nrows);
    ZK_LAUNCHED();
    ZK_CUDA(cudaGetLastError());
    return merkle_levels(tree, nrows, st);
}
#endif
