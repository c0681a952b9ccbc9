// ntt.cu — Goldilocks NTT / INTT / coset LDE for row-major column-interleaved matrices (sm_100a).
//
// Replaces NTT_Goldilocks::{NTT, INTT, NTT_iters, reversePermutation, extendPol}
// (reference src/goldilocks/src/ntt_goldilocks.cpp:12-164, 166-232, 243-305, 307-352).  The reference
// runs a radix-2 DIT network in <= 3 cache-blocked phases with a bit-reversal gather and row transposes
// between phases.  Field results are unique, so any exact algorithm gives the same canonical words; this
// file uses a different, GPU-shaped factorisation:
//
//   N = N_1 * N_2 * ... * N_P   (N_p = 2^l_p, l_p <= 8), one kernel launch ("pass") per factor.
//   input index  n = sum n_p * M_p      (M_p = N / (N_1..N_p), M_P = 1)
//   output index k = sum k_p * K_p      (K_p = N_1..N_{p-1},   K_1 = 1)
//   pass p: for every outer index o = (k_1..k_{p-1}) and every r_p < M_p, a size-N_p DFT over n_p of the
//           rows  o*N_p*M_p + n_p*M_p + r_p, whose inputs are first multiplied by the twiddle owed by the
//           previous level,  w_N^( (n_p*M_p + r_p) * k_{p-1} * K_{p-1} );  results are written back to the
//           same positions (k_p in place of n_p) except in the last pass, which scatters rows to natural
//           order k.  No bit-reversal pass and no transposes exist; every global access is a run of
//           consecutive words of one row (>= 128 B when the matrix is at least 16 columns wide).
//
//   Inside a pass a CTA owns a tile of N_p rows x LANES consecutive (r_p, column) words.  The size-N_p DFT
//   is done as R1 x R2 (R1*R2 = N_p, both <= 16): each thread runs a radix-R1 DIT network entirely in
//   registers, the R1 x R2 twiddle w_{N_p}^(j2*k1) is applied, values cross threads once through shared
//   memory, and a radix-R2 network in registers finishes.  The in-register networks only multiply by powers
//   of w_16 = 2^12, which the compiler strength-reduces.
//
//   The inverse transform uses the inverse roots (instead of the reference's forward network + index
//   reversal i -> (N-i) mod N, ntt_goldilocks.hpp:38-46): same values.
//
//   extendPol (blowup b): out[b*k + r] = sum_i c_i (7 w_{bN}^r)^i w_N^(ik)  — one size-N INTT and b size-N
//   coset NTTs (the coset factor and 1/N are a pre-scale of pass 1) written interleaved into the natural
//   order of the extended domain.  This never touches the zero-padded upper half the reference transforms
//   (ntt_goldilocks.cpp:259-277) and keeps every transform at size N.
#include "gl64.cuh"
#include "zk_internal.h"
#include <cstdlib>

namespace zk {

struct PassParams
{
    const u64 *src;
    u64 *dst;
    u64 src_ld, dst_ld;
    u64 totalQ;  // (N >> lgNp) * ncols : number of (o, r_p, column) lanes
    u32 ncols;
    u32 ncols_inv; // floor(2^32 / ncols), 0xFFFFFFFF for ncols == 1 (ntt_tile.cuh: fastdiv)
    u32 logN;
    u32 lgM;     // log2 M_p
    u32 lgNp;    // l_p
    u32 lgKp;    // log2 K_p
    u32 nprev;   // p - 1
    u32 lgNprev; // l_{p-1}
    u32 lgKprev; // log2 K_{p-1}
    u32 is_last;
    unsigned char lgs[8]; // l_1 .. l_{p-1}
    const u64 *T0, *T1;   // w_N^i (i < 4096), w_N^(4096 i)
    int pre_mode;
    u64 pre_const;
    const u64 *G0, *G1;
    int post_mode;
    u64 post_const;
    const u64 *H0, *H1;
    // last pass only, sharded commit: scatter over the peers' row shards (zk_internal.h: Scatter)
    const u64 *peer_tab;
    u32 sc_lg_b, sc_coset, sc_lg_rows;
    u64 sc_col0;
};

__device__ u64 g_w256[256]; // w_256^i, Goldilocks::W[8]^i (goldilocks_base_field.cpp:11-44)

// powers of w_16 = 2^12 and of its inverse (w_16^-e = -w_16^(8-e))
template <bool INV>
__device__ __forceinline__ constexpr u64 w16_pow(int e)
{
    constexpr u64 F[8] = {0x1ull, 0x1000ull, 0x1000000ull, 0x1000000000ull, 0x1000000000000ull, 0x1000000000000000ull,
                          0xffffffff00ull, 0xffffffff00000ull};
    constexpr u64 I[8] = {0x1ull, 0xffefffff00100001ull, 0xfffffeff00000101ull, 0xefffffff00000001ull,
                          0xfffeffff00000001ull, 0xffffffef00000001ull, 0xfffffffeff000001ull, 0xfffffffefffff001ull};
    return INV ? I[e] : F[e];
}

__host__ __device__ constexpr int bitrev_c(int x, int bits)
{
    int r = 0;
    for (int i = 0; i < bits; i++) r |= ((x >> i) & 1) << (bits - 1 - i);
    return r;
}

// t = hi * w_16^(+-e) with w_16 = 2^12 as a shift (gl::mul_pow2); the inverse root is -w_16^(8-e), i.e. the same shifts
// with the two outputs of the butterfly exchanged.
template <int E16, bool INV, bool HI_CANON>
__device__ __forceinline__ void butterfly16(u64 &lo, u64 &hi)
{
    const u64 u = lo;
    if constexpr (E16 == 0)
    {
        const u64 t = HI_CANON ? hi : gl::canon(hi); // gl::add / gl::sub need a canonical second operand
        lo = gl::add(u, t);
        hi = gl::sub(u, t);
    }
    else
    {
        const u64 t = gl::mul_pow2<12 * (INV ? 8 - E16 : E16)>(hi);
        if (INV)
        {
            lo = gl::sub(u, t);
            hi = gl::add(u, t);
        }
        else
        {
            lo = gl::add(u, t);
            hi = gl::sub(u, t);
        }
    }
}
// e16 is a compile-time constant after unrolling; dispatch it to the template
template <bool INV, bool HI_CANON>
__device__ __forceinline__ void constexpr_butterfly(u64 &lo, u64 &hi, const int e16)
{
    switch (e16)
    {
    case 0: butterfly16<0, INV, HI_CANON>(lo, hi); break;
    case 1: butterfly16<1, INV, false>(lo, hi); break;
    case 2: butterfly16<2, INV, false>(lo, hi); break;
    case 3: butterfly16<3, INV, false>(lo, hi); break;
    case 4: butterfly16<4, INV, false>(lo, hi); break;
    case 5: butterfly16<5, INV, false>(lo, hi); break;
    case 6: butterfly16<6, INV, false>(lo, hi); break;
    default: butterfly16<7, INV, false>(lo, hi); break;
    }
}

// size-2^LOGE DFT in registers, natural order in and out.  Inputs may be any words in [0, 2^64);
// outputs are in [0, 2^64) (not canonical).
// Written with one flat, constant-trip-count loop per stage (template recursion over the stages): with nested
// variable-bound loops nvcc stopped unrolling at E = 16 and the array went to local memory (LDL/STL in the butterfly
// loop, ncu source page of the first version).
template <int LOGE, bool INV, int S, bool CANON_IN>
struct DftStageGuard
{
    static __device__ __forceinline__ void run(u64 (&a)[1 << LOGE])
    {
        if constexpr (S <= LOGE)
        {
            constexpr int E = 1 << LOGE, m = 1 << S, half = m >> 1;
#pragma unroll
            for (int b = 0; b < E / 2; b++)
            {
                const int k = (b / half) * m, j = b % half;
                // the first stage reads the caller's values: canonical when CANON_IN
                constexpr_butterfly<INV, (CANON_IN && S == 1)>(a[k + j], a[k + j + half], j * (16 / m));
            }
            DftStageGuard<LOGE, INV, S + 1, CANON_IN>::run(a);
        }
    }
};

template <int LOGE, bool INV, bool CANON_IN = false>
__device__ __forceinline__ void dft_regs(u64 (&a)[1 << LOGE])
{
    constexpr int E = 1 << LOGE;
    if constexpr (LOGE > 0)
    {
        u64 b[E];
#pragma unroll
        for (int i = 0; i < E; i++) b[i] = a[bitrev_c(i, LOGE)];
#pragma unroll
        for (int i = 0; i < E; i++) a[i] = b[i];
        DftStageGuard<LOGE, INV, 1, CANON_IN>::run(a);
    }
}

template <int LR1, int LR2>
struct PassCfg
{
    static constexpr int R1 = 1 << LR1, R2 = 1 << LR2, R = R1 * R2;
    static constexpr int TG = (LR2 == 0) ? 1 : (R1 > R2 ? R1 : R2); // threads per lane
    static constexpr int LANES = (TG >= 8) ? 16 : 128 / TG;
    static constexpr int THREADS = LANES * TG;
};

__device__ __forceinline__ u64 pow2level(const u64 *__restrict__ A0, const u64 *__restrict__ A1, u64 i)
{
    return gl::mul(__ldg(A0 + (i & 4095)), __ldg(A1 + (i >> 12)));
}

template <int LR1, int LR2, bool INV>
__global__ void __launch_bounds__(PassCfg<LR1, LR2>::THREADS) ntt_pass_kernel(const PassParams P)
{
    using Cfg = PassCfg<LR1, LR2>;
    constexpr int R1 = Cfg::R1, R2 = Cfg::R2, R = Cfg::R, LANES = Cfg::LANES, THREADS = Cfg::THREADS;
    __shared__ u64 xch[LR2 ? R * LANES : 1];
    __shared__ u64 wr[LR2 ? R : 1];

    const int lane = threadIdx.x % LANES, g = threadIdx.x / LANES;
    if (LR2)
    {
        for (int e = threadIdx.x; e < R; e += THREADS)
        {
            int idx = e * (256 / R);
            if (INV) idx = (256 - idx) & 255;
            wr[e] = g_w256[idx];
        }
        __syncthreads();
    }

    const u64 Q = (u64)blockIdx.x * LANES + lane;
    const bool active = Q < P.totalQ;
    u64 o = 0;
    u32 rp = 0, c = 0;
    if (active)
    {
        const u64 F = (u64)P.ncols << P.lgM;
        o = Q / F;
        const u64 q = Q - o * F;
        rp = (u32)(q / P.ncols);
        c = (u32)(q - (u64)rp * P.ncols);
    }
    const u64 tile_row0 = (o << (P.lgNp + P.lgM)) + rp;
    const u32 nmask = (P.logN >= 32) ? 0xFFFFFFFFu : ((1u << P.logN) - 1u);

    u64 a[R1];
    if (g < R2 && active)
    {
        const u64 *s = P.src + c;
        const u32 kprev = (u32)o & ((1u << P.lgNprev) - 1u);
#pragma unroll
        for (int j1 = 0; j1 < R1; j1++)
        {
            const u32 np = (u32)(j1 * R2 + g);
            const u64 row = tile_row0 + ((u64)np << P.lgM);
            u64 v = s[row * P.src_ld];
            if (P.nprev)
            {
                const u32 rprev = (np << P.lgM) + rp;
                u32 x = (rprev * kprev) << P.lgKprev;
                if (INV) x = (0u - x) & nmask;
                u64 w = __ldg(P.T0 + (x & 4095));
                if (P.logN > 12) w = gl::mul(w, __ldg(P.T1 + (x >> 12)));
                v = gl::mul(v, w);
            }
            else if (P.pre_mode == SCALE_CONST)
                v = gl::mul(v, P.pre_const);
            else if (P.pre_mode == SCALE_POWER)
                v = gl::mul(v, pow2level(P.G0, P.G1, row));
            a[j1] = v;
        }
        dft_regs<LR1, INV>(a);
        if (LR2)
        {
#pragma unroll
            for (int k1 = 0; k1 < R1; k1++)
            {
                u64 v = a[k1];
                if (k1 > 0) v = gl::mul(v, wr[g * k1]);
                xch[(k1 * R2 + g) * LANES + lane] = v;
            }
        }
    }

    // ---- output side ----
    u64 rev = 0;
    if (P.is_last)
    {
        u64 oo = o;
        int sh = (int)P.lgKp;
        for (int q = (int)P.nprev - 1; q >= 0; --q)
        {
            const int lq = P.lgs[q];
            sh -= lq;
            rev |= (oo & ((1ull << lq) - 1ull)) << sh;
            oo >>= lq;
        }
    }
    auto emit = [&](u64 v, u32 kp) {
        u64 drow;
        if (P.is_last)
        {
            drow = rev + ((u64)kp << P.lgKp);
            if (P.post_mode == SCALE_CONST)
                v = gl::mul(v, P.post_const);
            else if (P.post_mode == SCALE_POWER)
                v = gl::mul(v, pow2level(P.H0, P.H1, drow));
            else
                v = gl::canon(v);
        }
        else
        {
            drow = (o << (P.lgNp + P.lgM)) + ((u64)kp << P.lgM) + rp;
            v = gl::canon(v);
        }
        if (P.peer_tab)
        {
            const u64 e = (drow << P.sc_lg_b) + P.sc_coset;
            u64 *base = reinterpret_cast<u64 *>(__ldg(P.peer_tab + (e >> P.sc_lg_rows)));
            base[(e & (((u64)1 << P.sc_lg_rows) - 1)) * P.dst_ld + P.sc_col0 + c] = v;
            return;
        }
        P.dst[drow * P.dst_ld + c] = v;
    };

    if (LR2)
    {
        __syncthreads();
        if (g < R1 && active)
        {
            u64 b[R2];
#pragma unroll
            for (int j2 = 0; j2 < R2; j2++) b[j2] = xch[(g * R2 + j2) * LANES + lane];
            dft_regs<LR2, INV>(b);
#pragma unroll
            for (int k2 = 0; k2 < R2; k2++) emit(b[k2], (u32)(g + R1 * k2));
        }
    }
    else
    {
        if (active)
        {
#pragma unroll
            for (int k = 0; k < R1; k++) emit(a[k], (u32)k);
        }
    }
}

} // namespace zk
#include "ntt_tile.cuh"
namespace zk {

// ---------------------------------------------------------------------------------------------------
// tables
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ u64 dev_pow(u64 b, u64 e)
{
    u64 r = 1;
    while (e)
    {
        if (e & 1) r = gl::mul(r, b);
        b = gl::mul(b, b);
        e >>= 1;
    }
    return r;
}
__global__ void build_pow_tables(u64 *G0, u64 *G1, u64 g, u64 scale, u32 n1)
{
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 4096) G0[i] = dev_pow(g, i);
    if (i < n1) G1[i] = gl::mul(scale, dev_pow(g, (u64)i << 12));
}

// Tables live in the calling thread's context (one set per device).  The build kernel is enqueued on the caller's stream and
// every table has its own "built" event: a user on another stream (zkgpu_set_stream, the sponge / exchange streams of the
// multi-GPU commit) waits for exactly the table it reads, whatever was built in between and wherever.
// The cache is bounded: past 96 entries (a blowup-16 LDE needs 18) it is flushed.
int get_power_tables(u64 g, u64 scale, unsigned log_n, const u64 **G0, const u64 **G1, cudaStream_t st)
{
    Ctx &C = *t_ctx;
    auto key = std::make_tuple(g, scale, log_n);
    auto it = C.tabs.find(key);
    if (it == C.tabs.end())
    {
        if (C.tabs.size() >= 96)
        {
            ZK_CUDA(cudaDeviceSynchronize());
            for (auto &kv : C.tabs)
            {
                cudaFree(kv.second.g0);
                cudaFree(kv.second.g1);
                if (kv.second.built) cudaEventDestroy(kv.second.built);
            }
            C.tabs.clear();
        }
        const u32 n1 = log_n > 12 ? (1u << (log_n - 12)) : 1u;
        Ctx::Tab t;
        ZK_CUDA(cudaMalloc(&t.g0, 4096 * sizeof(u64)));
        ZK_CUDA(cudaMalloc(&t.g1, (size_t)n1 * sizeof(u64)));
        ZK_CUDA(cudaEventCreateWithFlags(&t.built, cudaEventDisableTiming));
        const u32 n = n1 > 4096 ? n1 : 4096;
        build_pow_tables<<<(n + 255) / 256, 256, 0, st>>>(t.g0, t.g1, g, scale, n1);
        ZK_LAUNCHED();
        ZK_CUDA(cudaGetLastError());
        ZK_CUDA(cudaEventRecord(t.built, st));
        t.built_on = st;
        it = C.tabs.emplace(key, t).first;
    }
    else if (it->second.built_on != st)
        ZK_CUDA(cudaStreamWaitEvent(st, it->second.built, 0));
    *G0 = it->second.g0;
    *G1 = it->second.g1;
    return 0;
}

static int ensure_w256(cudaStream_t st)
{
    Ctx &C = *t_ctx;
    if (C.w256_ready) return 0;
    u64 h[256];
    const u64 w = gl::root_of_unity(8);
    h[0] = 1;
    for (int i = 1; i < 256; i++) h[i] = gl::mul_slow(h[i - 1], w);
    ZK_CUDA(cudaMemcpyToSymbol(g_w256, h, sizeof(h))); // synchronous: every later stream sees it
    C.w256_ready = true;
    return 0;
}

// ---------------------------------------------------------------------------------------------------
// host driver
// ---------------------------------------------------------------------------------------------------
template <int LR1, int LR2>
static int launch_pass(const PassParams &P, bool inv, cudaStream_t st)
{
    using Cfg = PassCfg<LR1, LR2>;
    const u64 blocks = (P.totalQ + Cfg::LANES - 1) / Cfg::LANES;
    if (blocks == 0) return 0;
    if (blocks > 0x7FFFFFFFull)
    {
        set_error("ntt pass: grid too large (%llu blocks)", (unsigned long long)blocks);
        return -1;
    }
    if (inv)
        ntt_pass_kernel<LR1, LR2, true><<<(unsigned)blocks, Cfg::THREADS, 0, st>>>(P);
    else
        ntt_pass_kernel<LR1, LR2, false><<<(unsigned)blocks, Cfg::THREADS, 0, st>>>(P);
    ZK_LAUNCHED();
    ZK_CUDA(cudaGetLastError());
    return 0;
}

int g_ntt_force_v1 = 0;           // test hook (ZKGPU_NTT_V1=1): always use the simple per-tile kernel
int g_ntt_threads_per_sm_cap = 0; // 0: fill the SM

// resident CTAs per SM of a kernel on the current device (also raises its dynamic shared-memory limit there, once)
template <typename K>
static int kernel_occupancy(K kern, int threads, size_t smem, int *out)
{
    Ctx &C = *t_ctx;
    auto it = C.occupancy.find((const void *)kern);
    if (it == C.occupancy.end())
    {
        if (smem > 48 * 1024) ZK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int nb = 0;
        ZK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, threads, smem));
        if (nb < 1) nb = 1;
        it = C.occupancy.emplace((const void *)kern, nb).first;
    }
    *out = it->second;
    return 0;
}

template <int LR1, int LR2>
static int launch_tile(const PassParams &P, bool inv, cudaStream_t st)
{
    using Cfg = TileCfg<LR1, LR2>;
    auto kf = ntt_tile_kernel<LR1, LR2, false>;
    auto ki = ntt_tile_kernel<LR1, LR2, true>;
    int bps = 1;
    ZK_TRY(kernel_occupancy(inv ? ki : kf, Cfg::THREADS, Cfg::SMEM, &bps));
    const u64 ntiles = (P.totalQ + 15) / 16;
    u64 grid = (u64)t_ctx->num_sms * bps;
    // experiment knob (zkgpu_tune key 1): leave most of each SM to another kernel
    if (g_ntt_threads_per_sm_cap > 0)
    {
        u64 b = (u64)g_ntt_threads_per_sm_cap / Cfg::THREADS;
        if (b < 1) b = 1;
        if (b < (u64)bps) grid = (u64)t_ctx->num_sms * b;
    }
    if (grid > ntiles) grid = ntiles;
    if (grid == 0) return 0;
    if (inv)
        ki<<<(unsigned)grid, Cfg::THREADS, Cfg::SMEM, st>>>(P, (u32)ntiles);
    else
        kf<<<(unsigned)grid, Cfg::THREADS, Cfg::SMEM, st>>>(P, (u32)ntiles);
    ZK_LAUNCHED();
    ZK_CUDA(cudaGetLastError());
    return 0;
}

static int dispatch_pass(const PassParams &P, bool inv, cudaStream_t st)
{
    static const bool env_v1 = getenv("ZKGPU_NTT_V1") != nullptr;
    const bool tile_ok = !env_v1 && !g_ntt_force_v1 && P.ncols >= 16 && P.totalQ < 0xFFFFFFF0ull &&
                         ((u64)P.ncols << P.lgM) < 0xFFFFFFFFull && P.logN <= 30;
    if (tile_ok)
    {
        switch (P.lgNp)
        {
        case 5: return launch_tile<3, 2>(P, inv, st);
        case 6: return launch_tile<3, 3>(P, inv, st);
        case 7: return launch_tile<4, 3>(P, inv, st);
        case 8: return launch_tile<4, 4>(P, inv, st);
        }
    }
    switch (P.lgNp)
    {
    case 0: return launch_pass<0, 0>(P, inv, st);
    case 1: return launch_pass<1, 0>(P, inv, st);
    case 2: return launch_pass<1, 1>(P, inv, st);
    case 3: return launch_pass<2, 1>(P, inv, st);
    case 4: return launch_pass<2, 2>(P, inv, st);
    case 5: return launch_pass<3, 2>(P, inv, st);
    case 6: return launch_pass<3, 3>(P, inv, st);
    case 7: return launch_pass<4, 3>(P, inv, st);
    case 8: return launch_pass<4, 4>(P, inv, st);
    }
    set_error("ntt pass: unsupported pass size 2^%u", P.lgNp);
    return -1;
}

int ntt_transform(u64 *dst, u64 dst_ld, const u64 *src, u64 src_ld, u64 *tmp, unsigned L, u64 ncols, bool inverse,
                  const Scale &pre, const Scale &post, cudaStream_t st, const Scatter *sc, unsigned sc_log_b, unsigned sc_coset)
{
    if (ncols == 0) return 0;
    if (L > 30)
    {
        set_error("ntt: domain 2^%u not supported by this build (max 2^30)", L);
        return -1;
    }
    if (ncols > 0xFFFFFFFFull)
    {
        set_error("ntt: too many columns");
        return -1;
    }
    ZK_TRY(ensure_w256(st));
    const int np = L == 0 ? 1 : (int)((L + 7) / 8);
    if (np > 1 && tmp == nullptr)
    {
        set_error("ntt: scratch buffer required for log2(size) > 8");
        return -1;
    }
    unsigned lg[8];
    for (int p = 0; p < np; p++) lg[p] = L / np + ((unsigned)p < L % np ? 1u : 0u);

    const u64 *T0 = nullptr, *T1 = nullptr;
    ZK_TRY(get_power_tables(gl::root_of_unity(L), 1, L, &T0, &T1, st));

    unsigned consumed = 0;
    for (int p = 0; p < np; p++)
    {
        PassParams P{};
        const bool first = p == 0, last = p == np - 1;
        P.src = first ? src : tmp;
        P.src_ld = first ? src_ld : ncols;
        P.dst = last ? dst : tmp;
        P.dst_ld = last ? dst_ld : ncols;
        P.ncols = (u32)ncols;
        P.ncols_inv = ncols == 1 ? 0xFFFFFFFFu : (u32)(((u64)1 << 32) / ncols);
        P.logN = L;
        P.lgNp = lg[p];
        P.lgKp = consumed;
        P.lgM = L - consumed - lg[p];
        P.nprev = (u32)p;
        P.lgNprev = p ? lg[p - 1] : 0;
        P.lgKprev = p ? consumed - lg[p - 1] : 0;
        P.is_last = last;
        for (int q = 0; q < p; q++) P.lgs[q] = (unsigned char)lg[q];
        P.T0 = T0;
        P.T1 = T1;
        P.pre_mode = first ? pre.mode : SCALE_NONE;
        P.pre_const = pre.konst;
        P.G0 = pre.G0;
        P.G1 = pre.G1;
        P.post_mode = last ? post.mode : SCALE_NONE;
        P.post_const = post.konst;
        P.H0 = post.G0;
        P.H1 = post.G1;
        P.totalQ = ((u64)1 << (L - lg[p])) * ncols;
        if (sc && last)
        {
            P.peer_tab = sc->peer_tab;
            P.sc_lg_b = sc_log_b;
            P.sc_coset = sc_coset;
            P.sc_lg_rows = sc->lg_rows;
            P.sc_col0 = sc->col0;
            P.dst = nullptr;
            P.dst_ld = sc->ld;
        }
        ZK_TRY(dispatch_pass(P, inverse, st));
        consumed += lg[p];
    }
    return 0;
}

// b = 2^log_blowup transforms of size N over the cosets g0 * w_ext^r of the extended domain, written interleaved into its
// natural order:  out[(k << log_blowup) + r] = scale * sum_i coef_i (g0 w_ext^r)^i w_N^(ik).
// This is the forward NTT of size N*b of the zero-padded coefficient rows (times the powers of g0) without ever touching
// the zero rows (the reference transforms them: ntt_goldilocks.cpp:259-277).
int coset_ntts(u64 *out, u64 out_ld, const u64 *coef, u64 coef_ld, u64 *tmp, unsigned log_n, unsigned log_blowup, u64 ncols, u64 g0,
               u64 scale, cudaStream_t st, const Scatter *sc)
{
    const u64 b = (u64)1 << log_blowup;
    const u64 wext = gl::root_of_unity(log_n + log_blowup);
    Scale none;
    u64 g = g0 % GL_P; // coset r: g_r = g0 * w_ext^r
    for (u64 r = 0; r < b; r++)
    {
        Scale pre;
        if (g != 1)
        {
            pre.mode = SCALE_POWER;
            ZK_TRY(get_power_tables(g, scale, log_n, &pre.G0, &pre.G1, st));
        }
        else if (scale != 1)
        {
            pre.mode = SCALE_CONST;
            pre.konst = scale;
        }
        if (sc)
            ZK_TRY(ntt_transform(nullptr, 0, coef, coef_ld, tmp, log_n, ncols, false, pre, none, st, sc, log_blowup, (unsigned)r));
        else
            ZK_TRY(ntt_transform(out + r * out_ld, out_ld * b, coef, coef_ld, tmp, log_n, ncols, false, pre, none, st));
        g = gl::mul_slow(g, wext);
    }
    return 0;
}

int lde_strip(u64 *out, u64 out_ld, const u64 *in, u64 in_ld, u64 *scratch, unsigned log_n, unsigned log_blowup,
              u64 ncols, cudaStream_t st, const Scatter *sc)
{
    if (ncols == 0) return 0;
    const u64 n = (u64)1 << log_n;
    u64 *coef = scratch;
    u64 *tmp = scratch + n * ncols;
    Scale none;
    // unscaled inverse transform: coef_i * N   (1/N is folded into the coset tables)
    ZK_TRY(ntt_transform(coef, ncols, in, in_ld, tmp, log_n, ncols, true, none, none, st));
    const u64 ninv = gl::pow_slow(n % GL_P, GL_P - 2);
    return coset_ntts(out, out_ld, coef, ncols, tmp, log_n, log_blowup, ncols, GL_SHIFT, ninv, st, sc);
}

} // namespace zk
