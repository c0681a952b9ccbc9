// ntt_tile.cuh — the streaming pass kernel used for the large shapes (included by ntt.cu after PassParams,
// dft_regs, g_w256 and pow2level are defined).  Same mathematics and same tile definition as ntt_pass_kernel
// (see the header of ntt.cu); what differs is how a CTA moves data and how its code is laid out:
//
//   * persistent CTAs loop over tiles; the NEXT tile is fetched with cp.async (LDGSTS, 8-byte granules: matrix
//     rows are only 8-byte aligned because ncols is odd in the headline shape) straight into shared memory while the
//     current tile is transformed, so no warp ever waits on a global load (ncu on the first version: 10-11 stalled
//     warps per issue on long_scoreboard, issue slots 39 % busy);
//   * the tile lives in shared memory; both halves of the R1 x R2 split are the SAME routine
//     (tile_step: read E words at a stride, multiply by a twiddle read from a shared-memory table, radix-E
//     network in registers, write to shared memory or to global memory), a real function called twice, so the
//     whole kernel is ~1.5k SASS instructions and stays inside the 32 KB instruction cache;
//   * the twiddle owed by the previous pass (or the coset factor of pass 1) is computed once per tile row by the
//     CTA (a 16-lane tile touches at most two distinct (outer index, r_p) pairs when ncols >= 16) instead of once
//     per element per thread;
//   * every thread always owns work in both halves (threads per lane = min(R1, R2)), which removes the barrier
//     stalls the l = 7 pass had.
#pragma once

namespace zk {

__device__ __forceinline__ void cp_async8(void *smem, const void *gmem)
{
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait()
{
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

struct StepOut
{
    u64 *dst;      // global destination (column already added), nullptr: write back to shared memory
    u64 row0;      // destination row of output k = 0
    u64 row_step;  // destination row increment per k
    u64 dst_ld;
    int post_mode; // last pass only
    u64 post_const;
    const u64 *H0, *H1;
    bool canon;    // canonicalise before the global store (last pass)
    // sharded commit (last pass of a coset transform): scatter over the peers' row shards, see zk_internal.h: Scatter
    const u64 *peer_tab;
    u32 sc_lg_b, sc_coset, sc_lg_rows;
    u64 sc_col; // sc_col0 + column of this lane
};

// E = 2^LOGE words at buf[(base + j*stride) * 16], j < E, each multiplied by tw[twbase + j*twstride] (if tw != nullptr),
// transformed, and written back in place (out.dst == nullptr) or to global rows out.row0 + k*out.row_step.
template <int LOGE, bool INV>
__device__ __noinline__ void tile_step(u64 *buf, int base, int stride, const u64 *tw, int twbase, int twstride, StepOut out)
{
    constexpr int E = 1 << LOGE;
    u64 a[E];
    // Only the upper half of the inputs (the "hi" operands of the first butterfly stage after the bit reversal) has to
    // be canonical; the lower half may stay un-normalised (3 instructions cheaper per product).  twbase < 0 marks the
    // w_R pass, whose twiddle for j = 0 is w^0 = 1.
    if (tw)
    {
        const bool j0_one = twbase < 0;
        if (j0_one) twbase = 0;
#pragma unroll
        for (int j = 0; j < E; j++)
        {
            const u64 v = buf[(base + j * stride) * 16], w = tw[twbase + j * twstride];
            if (j == 0)
                a[j] = j0_one ? v : gl::mul_nc(v, w);
            else
                a[j] = j < E / 2 ? gl::mul_nc(v, w) : gl::mul(v, w);
        }
    }
    else
    {
#pragma unroll
        for (int j = 0; j < E; j++) a[j] = j < E / 2 ? buf[(base + j * stride) * 16] : gl::canon(buf[(base + j * stride) * 16]);
    }
    dft_regs<LOGE, INV, true>(a); // the first butterfly stage skips its own canon()
    const bool scatter = out.peer_tab != nullptr;
    const bool scaled = !scatter && out.dst != nullptr && out.post_mode != SCALE_NONE;
    if (out.dst == nullptr || scaled)
    {
#pragma unroll
        for (int k = 0; k < E; k++) buf[(base + k * stride) * 16] = a[k];
        if (scatter)
        {
            // rolled like the post-scaled path: the destination GPU changes with the top bits of the output row
            u64 row = out.row0;
            const u64 mask = ((u64)1 << out.sc_lg_rows) - 1;
#pragma unroll 1
            for (int k = 0; k < E; k++)
            {
                const u64 v = gl::canon(buf[(base + k * stride) * 16]);
                const u64 e = (row << out.sc_lg_b) + out.sc_coset;
                u64 *pb = reinterpret_cast<u64 *>(__ldg(out.peer_tab + (e >> out.sc_lg_rows)));
                pb[(e & mask) * out.dst_ld + out.sc_col] = v;
                row += out.row_step;
            }
            return;
        }
        if (!scaled) return;
        // Post-scaled output (INTT's 1/N or 7^i/N: zkgpu_ntt only, never the LDE): a rolled loop over the values just
        // parked in shared memory.  Keeping the three output flavours inside the unrolled store loop tripled the code
        // of this function and pushed the kernel past the instruction cache (ncu: stall_no_instruction 1.4-1.9).
        u64 row = out.row0;
#pragma unroll 1
        for (int k = 0; k < E; k++)
        {
            u64 v = buf[(base + k * stride) * 16];
            v = out.post_mode == SCALE_CONST ? gl::mul(v, out.post_const) : gl::mul(v, pow2level(out.H0, out.H1, row));
            out.dst[row * out.dst_ld] = v;
            row += out.row_step;
        }
        return;
    }
    u64 *d = out.dst + out.row0 * out.dst_ld;
    const u64 dstep = out.row_step * out.dst_ld;
    if (out.canon)
    {
#pragma unroll
        for (int k = 0; k < E; k++) d[k * dstep] = gl::canon(a[k]);
    }
    else
    {
#pragma unroll
        for (int k = 0; k < E; k++) d[k * dstep] = a[k];
    }
}

template <int LR1, int LR2>
struct TileCfg
{
    static constexpr int R1 = 1 << LR1, R2 = 1 << LR2, R = R1 * R2;
    static constexpr int TG = R1 < R2 ? R1 : R2;   // threads per lane
    static constexpr int THREADS = 16 * TG;
    static constexpr int U0 = R2 / TG;             // radix-R1 networks per thread in the first half
    static constexpr int U1 = R1 / TG;             // radix-R2 networks per thread in the second half
    static constexpr int EPT = R / TG;             // words per thread per tile
    // shared memory: 2 tile buffers + pre-twiddle table (2 slots x R) + w_R table (R)
    static constexpr size_t SMEM = (size_t)(2 * R * 16 + 2 * R + R) * sizeof(u64);
};

struct LaneInfo
{
    u64 o;
    u32 rp, c;
    bool active;
};

// n / d with the host-computed M = floor(2^32 / d) (0xFFFFFFFF for d = 1): the estimate umulhi(n, M) is at most 2 short
__device__ __forceinline__ u32 fastdiv(u32 n, u32 d, u32 M)
{
    u32 q = __umulhi(n, M);
    u32 r = n - q * d;
    if (r >= d)
    {
        q++;
        r -= d;
    }
    if (r >= d) q++;
    return q;
}

__device__ __forceinline__ LaneInfo decode_lane(const PassParams &P, u64 Q)
{
    LaneInfo L;
    L.active = Q < P.totalQ;
    L.o = 0;
    L.rp = 0;
    L.c = 0;
    if (L.active)
    {
        // host guarantees totalQ < 2^32 for this kernel; lanes are numbered (o, r_p, column) with F = ncols << lgM
        // lanes per outer index o, so o = (Q >> lgM) / ncols
        const u32 q32 = (u32)Q;
        const u32 o = fastdiv(q32 >> P.lgM, P.ncols, P.ncols_inv);
        const u32 q = q32 - o * (P.ncols << P.lgM);
        L.rp = fastdiv(q, P.ncols, P.ncols_inv);
        L.c = q - L.rp * P.ncols;
        L.o = o;
    }
    return L;
}

// resident threads per SM the kernel is compiled for: 512 keeps 128 registers per thread (no spills); 768 spills 4-76 bytes
#ifndef ZK_TILE_OCC
#define ZK_TILE_OCC 512
#endif
template <int LR1, int LR2, bool INV>
__global__ void __launch_bounds__(TileCfg<LR1, LR2>::THREADS, ZK_TILE_OCC / TileCfg<LR1, LR2>::THREADS) ntt_tile_kernel(const PassParams P, const u32 ntiles)
{
    using Cfg = TileCfg<LR1, LR2>;
    constexpr int R1 = Cfg::R1, R2 = Cfg::R2, R = Cfg::R, TG = Cfg::TG, THREADS = Cfg::THREADS;
    extern __shared__ __align__(16) u64 smem[];
    u64 *bufs = smem;                   // [2][R][16]
    u64 *twt = smem + 2 * R * 16;       // [2][R]   pre-twiddle / pre-scale of tile row np, slot 0 / 1
    u64 *wr = twt + 2 * R;              // [R]      w_R^e

    const int lane = threadIdx.x & 15, g = threadIdx.x >> 4;
    for (int e = threadIdx.x; e < R; e += THREADS)
    {
        int idx = e * (256 / R);
        if (INV) idx = (256 - idx) & 255;
        wr[e] = g_w256[idx];
    }
    const u32 nmask = (1u << P.logN) - 1u;
    const bool has_pre = P.nprev != 0 || P.pre_mode != SCALE_NONE;

    auto prefetch = [&](u32 t, u64 *buf) -> LaneInfo {
        const LaneInfo L = decode_lane(P, (u64)t * 16 + lane);
        if (L.active)
        {
            const u64 row0 = (L.o << (P.lgNp + P.lgM)) + L.rp;
            // rows of this thread are TG << lgM apart: one 64-bit pointer increment per element
            const u64 *s = P.src + L.c + (row0 + ((u64)g << P.lgM)) * P.src_ld;
            const u64 sstep = ((u64)TG << P.lgM) * P.src_ld;
            u64 *b = buf + g * 16 + lane;
#pragma unroll 4
            for (int i = 0; i < Cfg::EPT; i++)
            {
                cp_async8(b, s);
                s += sstep;
                b += TG * 16;
            }
        }
        cp_async_commit();
        return L;
    };

    u32 t = blockIdx.x;
    LaneInfo cur = prefetch(t, bufs);
    for (int it = 0; t < ntiles; t += gridDim.x, it++)
    {
        u64 *buf = bufs + (size_t)(it & 1) * R * 16;
        const u32 tn = t + gridDim.x;
        LaneInfo nxt;
        if (tn < ntiles)
            nxt = prefetch(tn, bufs + (size_t)((it + 1) & 1) * R * 16);
        else
        {
            nxt.active = false;
            nxt.o = 0;
            nxt.rp = nxt.c = 0;
            cp_async_commit();
        }

        // (outer index, r_p) of lane 0 and of lane 15 of this tile: the (at most) two twiddle slots
        const LaneInfo A = decode_lane(P, (u64)t * 16);
        const u64 qlast = (u64)t * 16 + 15;
        const LaneInfo B = decode_lane(P, qlast < P.totalQ ? qlast : P.totalQ - 1);
        const bool two = (B.o != A.o) || (B.rp != A.rp);
        if (has_pre)
        {
            for (int e = threadIdx.x; e < (two ? 2 * R : R); e += THREADS)
            {
                const int np = e & (R - 1);
                const LaneInfo &S = (e < R) ? A : B;
                u64 w;
                if (P.nprev)
                {
                    const u32 kprev = (u32)S.o & ((1u << P.lgNprev) - 1u);
                    const u32 rprev = ((u32)np << P.lgM) + S.rp;
                    u32 x = (rprev * kprev) << P.lgKprev;
                    if (INV) x = (0u - x) & nmask;
                    w = __ldg(P.T0 + (x & 4095));
                    if (P.logN > 12) w = gl::mul(w, __ldg(P.T1 + (x >> 12)));
                }
                else if (P.pre_mode == SCALE_CONST)
                    w = P.pre_const;
                else
                    w = pow2level(P.G0, P.G1, ((u64)np << P.lgM) + S.rp);
                twt[e] = w;
            }
        }
        cp_async_wait<1>();
        __syncthreads();

        const int slot = (cur.o == A.o && cur.rp == A.rp) ? 0 : 1;
        // ---- first half: radix-R1 over j1 for j2 = g + TG*u ----
        {
            StepOut so{};
            so.dst = nullptr;
#pragma unroll 1
            for (int u = 0; u < Cfg::U0; u++)
            {
                const int j2 = g + TG * u;
                tile_step<LR1, INV>(buf + lane, j2, R2, has_pre ? twt : nullptr, slot * R + j2, R2, so);
            }
        }
        __syncthreads();
        // ---- second half: radix-R2 over j2 for k1 = g + TG*u, multiplied by w_R^(j2*k1) on the way in ----
        if (cur.active)
        {
            StepOut so;
            so.peer_tab = P.is_last ? P.peer_tab : nullptr;
            so.sc_lg_b = P.sc_lg_b;
            so.sc_coset = P.sc_coset;
            so.sc_lg_rows = P.sc_lg_rows;
            so.sc_col = P.sc_col0 + cur.c;
            so.dst = so.peer_tab ? nullptr : P.dst + cur.c;
            so.dst_ld = P.dst_ld;
            so.post_mode = P.is_last ? P.post_mode : SCALE_NONE;
            so.post_const = P.post_const;
            so.H0 = P.H0;
            so.H1 = P.H1;
            so.canon = P.is_last != 0;
            u64 rev = 0;
            if (P.is_last)
            {
                u64 oo = cur.o;
                int sh = (int)P.lgKp;
                for (int q = (int)P.nprev - 1; q >= 0; --q)
                {
                    const int lq = P.lgs[q];
                    sh -= lq;
                    rev |= (oo & ((1ull << lq) - 1ull)) << sh;
                    oo >>= lq;
                }
            }
#pragma unroll 1
            for (int u = 0; u < Cfg::U1; u++)
            {
                const int k1 = g + TG * u;
                if (P.is_last)
                {
                    so.row0 = rev + ((u64)k1 << P.lgKp);
                    so.row_step = (u64)R1 << P.lgKp;
                }
                else
                {
                    so.row0 = (cur.o << (P.lgNp + P.lgM)) + ((u64)k1 << P.lgM) + cur.rp;
                    so.row_step = (u64)R1 << P.lgM;
                }
                tile_step<LR2, INV>(buf + lane, k1 * R2, 1, wr, -1, k1, so);
            }
        }
        __syncthreads();
        cur = nxt;
    }
    cp_async_wait<0>();
}

} // namespace zk
