// api.cu — the C ABI of libzkgpu.so (include/zkgpu.h): context, host/device pointer handling, staging,
// stage timers, and the orchestration of the commit stage (LDE strips -> leaf hashing -> node levels).
#include "gl64.cuh"
#include "zk_internal.h"
#include "../../include/zkgpu.h"
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <atomic>
#include <chrono>
#include <thread>
#include <map>
#include <mutex>
#include <vector>

namespace zk {

static inline bool is_pow2(u64 x) { return x && !(x & (x - 1)); }
static inline unsigned ilog2(u64 x)
{
    unsigned r = 0;
    while (x > 1)
    {
        x >>= 1;
        r++;
    }
    return r;
}

// ---- small kernels ----
__device__ __forceinline__ u64 splitmix64(u64 x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__global__ void fill_kernel(u64 *dst, u64 n, u64 seed, u64 offset)
{
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x)
        dst[i] = splitmix64(seed + offset + i) % GL_P;
}
int fill_splitmix(u64 *dst, u64 n, u64 seed, u64 offset, cudaStream_t st)
{
    if (n == 0) return 0;
    u64 blocks = (n + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    fill_kernel<<<(unsigned)blocks, 256, 0, st>>>(dst, n, seed, offset);
    ZK_LAUNCHED();
    ZK_CUDA(cudaGetLastError());
    return 0;
}
__global__ void copy_block_kernel(u64 *dst, u64 dst_ld, const u64 *src, u64 src_ld, u64 nrows, u64 ncols)
{
    const u64 total = nrows * ncols;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (u64)gridDim.x * blockDim.x)
    {
        const u64 r = i / ncols, c = i - r * ncols;
        dst[r * dst_ld + c] = src[r * src_ld + c];
    }
}

// NTT_Goldilocks::reversePermutation: one warp-sized group of threads per destination row
__global__ void reverse_perm_kernel(u64 *dst, const u64 *src, u64 size, unsigned log_size, u64 offset_cols, u64 ncols, u64 ncols_all,
                                    u64 ext_limit)
{
    const u64 total = size * ncols;
    for (u64 e = (u64)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (u64)gridDim.x * blockDim.x)
    {
        const u64 i = e / ncols, c = e - i * ncols;
        const u64 r = log_size ? (__brevll(i) >> (64 - log_size)) : 0;
        const u64 off = r * ncols_all + offset_cols;
        dst[e] = (off < ext_limit) ? src[off + c] : 0;
    }
}
// in-place variant (dst == src, single block of columns): swap rows i and BR(i) for BR(i) < i
__global__ void reverse_perm_inplace_kernel(u64 *a, u64 size, unsigned log_size, u64 ncols)
{
    const u64 total = size * ncols;
    for (u64 e = (u64)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (u64)gridDim.x * blockDim.x)
    {
        const u64 i = e / ncols, c = e - i * ncols;
        const u64 r = log_size ? (__brevll(i) >> (64 - log_size)) : 0;
        if (r < i)
        {
            const u64 x = a[i * ncols + c], y = a[r * ncols + c];
            a[i * ncols + c] = y;
            a[r * ncols + c] = x;
        }
    }
}
// MerkleTreeGL::getGroupProof gather: row idx of source, then the sibling digest of every level
__global__ void group_proof_kernel(u64 *proof, const u64 *row, const u64 *nodes, u64 width, u64 height, u64 idx)
{
    for (u64 c = threadIdx.x; c < width; c += blockDim.x) proof[c] = row[c];
    if (nodes != nullptr && threadIdx.x < 4)
    {
        u64 off = 0, n = height, i = idx, lvl = 0;
        while (n > 1)
        {
            proof[width + 4 * lvl + threadIdx.x] = nodes[off + (i ^ 1) * 4 + threadIdx.x];
            off += n * 4;
            n >>= 1;
            i >>= 1;
            lvl++;
        }
    }
}

// the same for a batch of queries: block q serves idx[q] (FRIProve::prove asks for nQueries proofs per tree, friProve.cpp:144-160)
__global__ void group_proofs_kernel(u64 *proofs, const u64 *source, const u64 *nodes, u64 width, u64 height, const u64 *idx, u64 proof_words)
{
    const u64 i0 = idx[blockIdx.x];
    u64 *proof = proofs + (u64)blockIdx.x * proof_words;
    const u64 *row = source + i0 * width;
    for (u64 c = threadIdx.x; c < width; c += blockDim.x) proof[c] = row[c];
    if (threadIdx.x < 4)
    {
        u64 off = 0, n = height, i = i0, lvl = 0;
        while (n > 1)
        {
            proof[width + 4 * lvl + threadIdx.x] = nodes[off + (i ^ 1) * 4 + threadIdx.x];
            off += n * 4;
            n >>= 1;
            i >>= 1;
            lvl++;
        }
    }
}

__global__ void field_op_kernel(int op, u64 *out, const u64 *a, const u64 *b, u64 n)
{
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x)
    {
        const u64 x = a[i], y = b[i];
        u64 r;
        switch (op)
        {
        case 0: r = gl::canon(gl::add(x, gl::canon(y))); break; // gl::add needs one operand < p
        case 1: r = gl::canon(gl::sub(x, gl::canon(y))); break;
        case 2: r = gl::mul(x, y); break;
        case 3: r = gl::canon(gl::mul_nc(x, y)); break;
        default: r = gl::canon(x); break;
        }
        out[i] = r;
    }
}

// ---- LDE over all columns in strips that fit the scratch budget ----
static int lde_all(Ctx &C, u64 *out, u64 out_ld, const u64 *in, u64 in_ld, unsigned log_n, unsigned log_b, u64 ncols, cudaStream_t st)
{
    if (ncols == 0) return 0;
    const u64 n = (u64)1 << log_n;
    u64 w = ncols;
    // the whole width fits the scratch already held: no need to ask the driver (cudaMemGetInfo costs ~1 ms, which is most
    // of the host time of a small commit)
    if (2 * n * ncols > C.scratch_words)
    {
        size_t free_b = 0, total_b = 0;
        ZK_CUDA(cudaMemGetInfo(&free_b, &total_b));
        free_b += C.scratch_words * sizeof(u64);
        size_t budget = (size_t)(0.6 * (double)free_b);
        const size_t cap = (size_t)16 << 30;
        if (budget > cap) budget = cap;
        w = budget / (2 * n * sizeof(u64));
    }
    if (w >= ncols)
        w = ncols;
    else
    {
        w &= ~(u64)15;
        if (w == 0)
        {
            set_error("extendPol: not enough device memory for a 16-column strip of 2^%u rows", log_n);
            return -4;
        }
    }
    ZK_TRY(ctx_ensure_scratch(C, 2 * n * w));
    for (u64 c0 = 0; c0 < ncols; c0 += w)
    {
        const u64 cw = ncols - c0 < w ? ncols - c0 : w;
        ZK_TRY(lde_strip(out + c0, out_ld, in + c0, in_ld, C.scratch, log_n, log_b, cw, st));
    }
    return 0;
}

static int merkle_all(u64 *tree, const u64 *input, u64 row_words, u64 nrows, cudaStream_t st, cudaEvent_t mid)
{
    ZK_TRY(merkle_leaves(tree, input, row_words, row_words, nrows, st));
    if (mid) ZK_CUDA(cudaEventRecord(mid, st));
    ZK_TRY(merkle_levels(tree, nrows, st));
    return 0;
}

// staging helper: device mirror of a host buffer
struct Staged
{
    u64 *d = nullptr;
    bool owned = false;
    ~Staged()
    {
        if (owned && d) cudaFree(d);
    }
    int alloc(u64 words)
    {
        if (words == 0) words = 1;
        cudaError_t e = cudaMalloc(&d, words * sizeof(u64));
        if (e != cudaSuccess)
        {
            cudaGetLastError();
            set_error("zkgpu: cannot stage %llu MiB on the device (%s)", (unsigned long long)(words >> 17), cudaGetErrorString(e));
            return -4;
        }
        owned = true;
        return 0;
    }
};

// ---------------------------------------------------------------------------------------------------
// The commit stage as a strip pipeline.  Columns are independent in the LDE and the row sponge absorbs columns
// left to right, so the trace is processed in column strips (multiples of 16 columns):
//     [host->device copy of strip s+1]  ||  LDE of strip s  ->  leaf absorption of strip s  ||  [device->host of strip s-1]
// Host-resident input / output strips are double buffered on the device and moved with strided 2-D copies on their
// own streams; device-resident input / output is used in place.  The running leaf digests live in tree[0 .. 4*NExt).
// ---------------------------------------------------------------------------------------------------
// d_tree == nullptr: LDE only (extendPol with host pointers)
static int commit_pipeline(Ctx &C, u64 *out_ext, bool out_dev, u64 *d_tree, const u64 *input, bool in_dev, unsigned log_n, unsigned log_b,
                           u64 ncols, cudaStream_t st)
{
    const u64 n = (u64)1 << log_n, n_ext = n << log_b;
    ZK_TRY(ctx_pipe_streams(C));
    // ---- strip width from the memory budget: per column 2n words of LDE scratch, 2n per host-input buffer pair,
    //      2*NExt per host-output buffer pair
    const u64 per_col = 2 * n + (in_dev ? 0 : 2 * n) + (out_dev ? 0 : 2 * n_ext);
    u64 w = ncols;
    if (per_col * ncols > C.scratch_words) // see lde_all: ask the driver only when the scratch has to grow
    {
        size_t free_b = 0, total_b = 0;
        ZK_CUDA(cudaMemGetInfo(&free_b, &total_b));
        free_b += C.scratch_words * sizeof(u64);
        size_t budget = (size_t)(0.6 * (double)free_b);
        // with a device-resident extended trace most of the memory belongs to the caller: stay small.  With a host-resident
        // one the strips are all the library holds, and wide strips are what the copies want (rows of 384 B moved at
        // 29 GB/s up and 36 GB/s down, profiles/r2_e2e_full_trace.txt; the strided copies reach the link rate from ~900 B)
        const size_t cap = (size_t)(out_dev ? 24 : (getenv("ZKGPU_HOST_STRIP") ? 112 : 72)) << 30;
        if (budget > cap) budget = cap;
        w = budget / (per_col * sizeof(u64));
    }
    if (!in_dev || !out_dev)
    {
        // enough strips for the copies to overlap the compute
        u64 target = ((ncols + 5) / 6 + 15) & ~(u64)15;
        if (const char *e = getenv("ZKGPU_HOST_STRIP")) // development knob (tools/host_strip_sweep.py): strip width in columns
            if (atoi(e) >= 16) target = (u64)atoi(e) & ~(u64)15;
        if (target < 16) target = 16;
        if (w > target) w = target;
    }
    if (w >= ncols)
        w = ncols;
    else
    {
        w &= ~(u64)15;
        if (w == 0)
        {
            set_error("commit: not enough device memory for a 16-column strip of 2^%u rows", log_n);
            return -4;
        }
    }
    const u64 scratch_words = 2 * n * w, in_words = in_dev ? 0 : n * w, out_words = out_dev ? 0 : n_ext * w;
    ZK_TRY(ctx_ensure_scratch(C, scratch_words + 2 * in_words + 2 * out_words));
    u64 *lde_scratch = C.scratch;
    u64 *in_buf[2] = {C.scratch + scratch_words, C.scratch + scratch_words + in_words};
    u64 *out_buf[2] = {C.scratch + scratch_words + 2 * in_words, C.scratch + scratch_words + 2 * in_words + out_words};

    ctx_pipe_event(C, nullptr, st);
    C.segs.clear();
    // strip boundaries: full-width strips of w columns, except that with a host-resident trace the first strips ramp up
    // (16, 32, 48, 64 columns): only the copy of the first 16 columns is exposed and each later copy (1.33 ms per column)
    // nearly fits under the compute of the strip before it (2.0 ms per column).  Timelines (ZKGPU_PIPE_TRACE=1,
    // profiles/r1_e2e_pipeline_trace.txt) at H: 32,64 -> 1438 ms; 16,32,48,80 -> 1435 ms; 16,32,48,64 -> 1420 ms.
    std::vector<u64> start;
    for (u64 c = 0; c < ncols;)
    {
        start.push_back(c);
        u64 cw = w;
        // ZKGPU_STRIP_RAMP="a,b,c,d" overrides the ramp (development knob, multiples of 16)
        static u64 ramp[4] = {16, 32, 48, 64};
        static bool ramp_init = false;
        if (!ramp_init)
        {
            ramp_init = true;
            if (const char *e = getenv("ZKGPU_STRIP_RAMP"))
            {
                unsigned long long v[4] = {0, 0, 0, 0};
                const int k = sscanf(e, "%llu,%llu,%llu,%llu", &v[0], &v[1], &v[2], &v[3]);
                for (int i = 0; i < 4; i++) ramp[i] = (i < k && v[i] >= 16) ? (v[i] & ~15ull) : ~0ull;
            }
        }
        if (!in_dev && start.size() <= 4 && w > ramp[start.size() - 1]) cw = ramp[start.size() - 1];
        c += cw;
    }
    start.push_back(ncols);
    const u64 nstrips = start.size() - 1;
    auto strip_cols = [&](u64 s) { return start[s + 1] - start[s]; };
    // ZKGPU_PIPE_TRACE=1: print the timeline of the copies and kernels of this call (development aid; synchronises)
    static const bool trace = getenv("ZKGPU_PIPE_TRACE") != nullptr;
    std::vector<std::pair<int, int>> h2d_ev, d2h_ev;
    int t_origin = -1;
    if (trace) ZK_TRY(ctx_pipe_event(C, &t_origin, st));
    auto issue_h2d = [&](u64 s) -> int {
        const int b = (int)(s & 1);
        const u64 cw = strip_cols(s);
        if (s >= 2) ZK_CUDA(cudaStreamWaitEvent(C.h2d, C.in_free[b], 0));
        int t0 = -1, t1 = -1;
        if (trace) ZK_TRY(ctx_pipe_event(C, &t0, C.h2d));
        ZK_CUDA(cudaMemcpy2DAsync(in_buf[b], cw * sizeof(u64), input + start[s], ncols * sizeof(u64), cw * sizeof(u64), n,
                                  cudaMemcpyHostToDevice, C.h2d));
        if (trace) ZK_TRY(ctx_pipe_event(C, &t1, C.h2d));
        if (trace) h2d_ev.push_back({t0, t1});
        ZK_CUDA(cudaEventRecord(C.in_ready[b], C.h2d));
        return 0;
    };
    if (!in_dev)
    {
        // the copy stream must not start before work already queued on the compute stream (e.g. a previous call
        // still reading the staging buffers) has drained
        ZK_CUDA(cudaEventRecord(C.in_free[0], st));
        ZK_CUDA(cudaStreamWaitEvent(C.h2d, C.in_free[0], 0));
        ZK_TRY(issue_h2d(0));
    }
    if (!out_dev)
    {
        ZK_CUDA(cudaEventRecord(C.out_free[0], st));
        ZK_CUDA(cudaStreamWaitEvent(C.d2h, C.out_free[0], 0));
    }
    for (u64 s = 0; s < nstrips; s++)
    {
        const int b = (int)(s & 1);
        const u64 c0 = start[s], cw = strip_cols(s);
        const u64 *src = input + c0;
        u64 src_ld = ncols;
        if (!in_dev)
        {
            if (s + 1 < nstrips) ZK_TRY(issue_h2d(s + 1));
            ZK_CUDA(cudaStreamWaitEvent(st, C.in_ready[b], 0));
            src = in_buf[b];
            src_ld = cw;
        }
        u64 *dst = out_ext + c0;
        u64 dst_ld = ncols;
        if (!out_dev)
        {
            if (s >= 2) ZK_CUDA(cudaStreamWaitEvent(st, C.out_free[b], 0));
            dst = out_buf[b];
            dst_ld = cw;
        }
        int e0, e1, e2;
        ZK_TRY(ctx_pipe_event(C, &e0, st));
        ZK_TRY(lde_strip(dst, dst_ld, src, src_ld, lde_scratch, log_n, log_b, cw, st));
        ZK_TRY(ctx_pipe_event(C, &e1, st));
        if (!in_dev) ZK_CUDA(cudaEventRecord(C.in_free[b], st));
        if (d_tree) ZK_TRY(merkle_absorb(d_tree, dst, cw, dst_ld, n_ext, s == 0, st));
        ZK_TRY(ctx_pipe_event(C, &e2, st));
        C.segs.push_back({0, e0, e1});
        C.segs.push_back({1, e1, e2});
        if (!out_dev)
        {
            ZK_CUDA(cudaEventRecord(C.out_ready[b], st));
            ZK_CUDA(cudaStreamWaitEvent(C.d2h, C.out_ready[b], 0));
            int t0 = -1, t1 = -1;
            if (trace) ZK_TRY(ctx_pipe_event(C, &t0, C.d2h));
            ZK_CUDA(cudaMemcpy2DAsync(out_ext + c0, ncols * sizeof(u64), out_buf[b], cw * sizeof(u64), cw * sizeof(u64), n_ext,
                                      cudaMemcpyDeviceToHost, C.d2h));
            if (trace) ZK_TRY(ctx_pipe_event(C, &t1, C.d2h));
            if (trace) d2h_ev.push_back({t0, t1});
            ZK_CUDA(cudaEventRecord(C.out_free[b], C.d2h));
        }
    }
    if (!out_dev)
    {
        // the caller-visible completion point is the compute stream: make it wait for the last device->host strips
        ZK_CUDA(cudaStreamWaitEvent(st, C.out_free[0], 0));
        ZK_CUDA(cudaStreamWaitEvent(st, C.out_free[1], 0));
    }
    if (trace)
    {
        ZK_CUDA(cudaStreamSynchronize(st));
        if (C.h2d) ZK_CUDA(cudaStreamSynchronize(C.h2d));
        if (C.d2h) ZK_CUDA(cudaStreamSynchronize(C.d2h));
        auto at = [&](int e) {
            float ms = 0;
            cudaEventElapsedTime(&ms, C.pool[t_origin], C.pool[e]);
            return ms;
        };
        fprintf(stderr, "[zkgpu pipe] strip cols | h2d begin..end | lde begin..end | absorb end | d2h begin..end   (ms from call start)\n");
        for (u64 s = 0; s < nstrips; s++)
        {
            const auto &l = C.segs[2 * s], &a = C.segs[2 * s + 1];
            float hb = -1, he = -1;
            if (s < h2d_ev.size()) hb = at(h2d_ev[s].first), he = at(h2d_ev[s].second);
            float db = -1, de = -1;
            if (s < d2h_ev.size()) db = at(d2h_ev[s].first), de = at(d2h_ev[s].second);
            fprintf(stderr, "[zkgpu pipe] %2llu %4llu | %8.1f %8.1f | %8.1f %8.1f | %8.1f | %8.1f %8.1f\n", (unsigned long long)s,
                    (unsigned long long)strip_cols(s), hb, he, at(l.a), at(l.b), at(a.b), db, de);
        }
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------------
// FRI folding step (src/starkpil/fri/friProve.cpp:26-107, 192-217, 252-271).  Elements of the cubic extension
// F_p[x]/(x^3 - x - 1) are 3 consecutive words (Polinomial with dim 3).
// ---------------------------------------------------------------------------------------------------
// Polinomial::mulElement dim 3 x dim 3 (polinomial.hpp:193-206); a, b canonical, result canonical
__device__ __forceinline__ void gl3_mul(u64 (&o)[3], const u64 (&a)[3], const u64 (&b)[3])
{
    const u64 A = gl::mul(gl::add(a[0], a[1]), gl::add(b[0], b[1]));
    const u64 B = gl::mul(gl::add(a[0], a[2]), gl::add(b[0], b[2]));
    const u64 C = gl::mul(gl::add(a[1], a[2]), gl::add(b[1], b[2]));
    const u64 D = gl::mul(a[0], b[0]), E = gl::mul(a[1], b[1]), F = gl::mul(a[2], b[2]);
    const u64 G = gl::canon(gl::sub(D, E));
    o[0] = gl::canon(gl::sub(gl::add(C, G), F));
    o[1] = gl::canon(gl::sub(gl::sub(gl::sub(gl::add(A, C), E), E), D));
    o[2] = gl::canon(gl::sub(B, G));
}
// coef: the size-nX INTT of every group, [nX][pol2N][3] (group g = columns 3g..3g+2 of the nX x 3 pol2N matrix view
// of friPol).  polMulAxi scales coefficient i by sinv_g^i and evalPol evaluates at special_x: together the Horner
// evaluation at y_g = special_x * sinv_g, sinv_g = shiftInv * w(polBits)^-g.  One thread per group; reads are
// coalesced across g.
__global__ void __launch_bounds__(256) fri_fold_kernel(u64 *out, const u64 *coef, u64 nx, u64 pol2n, u64 shift_inv, u64 wi, u64 x0, u64 x1,
                                                        u64 x2)
{
    const u64 g = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= pol2n) return;
    u64 s = shift_inv, b = wi;
    for (u64 e = g; e; e >>= 1)
    {
        if (e & 1) s = gl::mul(s, b);
        b = gl::mul(b, b);
    }
    const u64 y[3] = {gl::mul(x0, s), gl::mul(x1, s), gl::mul(x2, s)};
    const u64 *c = coef + g * 3;
    const u64 stride = pol2n * 3;
    u64 res[3];
#pragma unroll
    for (int k = 0; k < 3; k++) res[k] = gl::canon(c[(nx - 1) * stride + k]);
    for (u64 i = nx - 1; i-- > 0;)
    {
        u64 t[3];
        gl3_mul(t, res, y);
#pragma unroll
        for (int k = 0; k < 3; k++) res[k] = gl::canon(gl::add(t[k], gl::canon(c[i * stride + k])));
    }
#pragma unroll
    for (int k = 0; k < 3; k++) out[g * 3 + k] = res[k];
}

// Stage 4 of Starks::genProof, the loop between the two transforms (src/starkpil/starks.cpp:264-281):
//   qq2[k*qDeg + p] = qq1[p*N + k] * shiftIn^p   (elements of qDim words, Goldilocks3::mul(Element&, Element&, Goldilocks::Element&)
//   multiplies every word, goldilocks_cubic_extension.hpp:1887-1892), k < N, p < qDeg, shiftIn = (1/7)^N.
// qq1 arrives UNSCALED from the inverse transform; sp[p] = shiftIn^p / NExt carries its 1/NExt.
// Output: the N non-zero rows of the reference's NExt x (qDim*qDeg) matrix qq2; one thread per word.
__global__ void __launch_bounds__(256) stage4_regroup_kernel(u64 *qq2, const u64 *qq1, u64 n, u32 q_dim, u32 q_deg, const u64 *sp)
{
    const u64 w = (u64)q_dim * q_deg, total = n * w;
    for (u64 t = (u64)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (u64)gridDim.x * blockDim.x)
    {
        const u64 k = t / w;
        const u32 j = (u32)(t - k * w), p = j / q_dim, d = j - p * q_dim;
        qq2[t] = gl::mul(qq1[((u64)p * n + k) * q_dim + d], __ldg(sp + p));
    }
}
static inline unsigned nblocks(u64 n, unsigned per) { return (unsigned)((n + per - 1) / per); }
// getTransposed: aux[i*h + j] = pol[j*w + i] (3-word elements); one thread per destination word
__global__ void __launch_bounds__(256) fri_transpose_kernel(u64 *aux, const u64 *pol, u64 w, u64 h)
{
    const u64 t = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= w * h * 3) return;
    const u64 e = t / 3, k = t - e * 3, i = e / h, j = e - i * h;
    aux[t] = pol[(j * w + i) * 3 + k];
}


// ---------------------------------------------------------------------------------------------------
// Commit cache: what lets UNCHANGED starks.cpp get the fused path.  The reference calls
//     ntt.extendPol(p_cm1_2ns, p_cm1_n, NExt, N, ncols, buffer)        (starks.cpp:53)
//     treesGL[0]->merkelize()  ->  merkletree(nodes, source = p_cm1_2ns) (starks.hpp:186, merkleTreeGL.cpp:37-44)
// on the same host buffer.  While extendPol has each extended strip on the device anyway, it also runs the row sponge over
// it (the commit pipeline; with host pointers the call is bound by the PCIe copies, so the hashing is hidden under them)
// and finishes the node levels; the tree is remembered under the key (host pointer, rows, columns).  The merkletree call
// that follows finds it and only copies 32*(2*NExt-1) bytes back instead of re-uploading the 89 GB extended trace.
// A few rows of the output are kept as a fingerprint: if the caller changed the buffer in between (or calls
// zkgpu_invalidate), the entry is dropped and the tree is computed from the data as usual.
// ---------------------------------------------------------------------------------------------------
struct Resident
{
    const u64 *host = nullptr; // key: the caller's extended-trace buffer
    u64 nrows = 0, ncols = 0;
    u64 *d_tree = nullptr;     // device copy of the finished tree (only when no pinned slot could be had)
    u64 *h_tree = nullptr;     // pinned host copy: an entry nobody claims must not hold device memory
    const void *handle = nullptr; // the NTT_Goldilocks object whose extendPol made the entry
    // fingerprint of the buffer as extendPol left it: the first and the last row word for word, and a keyed checksum of
    // CACHE_SAMPLES words at positions drawn from a per-entry seed (every word when the buffer is smaller than that, or when
    // ZKGPU_CACHE_VERIFY=full).  A caller cannot know which words are looked at.
    u64 seed = 0, sum = 0;
    bool full = false;
    std::vector<u64> rows_idx, rows_words;
};
static std::mutex g_cache_mu;
static std::vector<Resident> g_cache; // process-wide: the key is a host pointer
static u64 g_cache_stats[4] = {0, 0, 0, 0};               // hits, misses, stale, evicted (zkgpu_cache_stats)
static std::map<const void *, int> g_unclaimed;           // per NTT handle: entries dropped without a merkelize() asking for them
int g_fuse_commit = -1;               // -1: read ZKGPU_FUSE_COMMIT (default on)

static bool fuse_enabled()
{
    if (g_fuse_commit < 0)
    {
        const char *e = getenv("ZKGPU_FUSE_COMMIT");
        g_fuse_commit = (e && e[0] == '0') ? 0 : 1;
    }
    return g_fuse_commit != 0;
}
// pinned tree buffers of the multi-GPU path are recycled: pinning 1 GB costs a few hundred milliseconds
static std::mutex g_pin_mu;
static std::vector<std::pair<u64 *, size_t>> g_pin_pool; // (buffer, words)
static u64 *pinned_take(size_t words)
{
    {
        std::lock_guard<std::mutex> lk(g_pin_mu);
        for (size_t i = 0; i < g_pin_pool.size(); i++)
            if (g_pin_pool[i].second >= words)
            {
                u64 *p = g_pin_pool[i].first;
                g_pin_pool.erase(g_pin_pool.begin() + i);
                return p;
            }
    }
    u64 *p = nullptr;
    if (cudaMallocHost(&p, words * sizeof(u64)) != cudaSuccess)
    {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}
static void pinned_give(u64 *p, size_t words)
{
    std::lock_guard<std::mutex> lk(g_pin_mu);
    if (g_pin_pool.size() >= 2)
    {
        cudaFreeHost(g_pin_pool.front().first);
        g_pin_pool.erase(g_pin_pool.begin());
    }
    g_pin_pool.push_back({p, words});
}
static void cache_free(Resident &r)
{
    if (r.d_tree) cudaFree(r.d_tree);
    if (r.h_tree) pinned_give(r.h_tree, (2 * r.nrows - 1) * 4);
    r.d_tree = r.h_tree = nullptr;
}
// drop every entry whose buffer overlaps [p, p + bytes)  (bytes == 0: exact pointer; p == nullptr: everything)
static void cache_invalidate(const void *p, size_t bytes)
{
    std::lock_guard<std::mutex> lk(g_cache_mu);
    for (size_t i = 0; i < g_cache.size();)
    {
        Resident &r = g_cache[i];
        const char *a = (const char *)r.host, *b = a + r.nrows * r.ncols * sizeof(u64);
        const bool hit = p == nullptr || (bytes == 0 ? p == (const void *)a : ((const char *)p < b && (const char *)p + bytes > a));
        if (hit)
        {
            if (bytes != 0) // overwritten by a new extendPol before any merkelize() asked for it
            {
                g_unclaimed[r.handle]++;
                g_cache_stats[3]++;
            }
            cache_free(r);
            g_cache.erase(g_cache.begin() + i);
        }
        else
            i++;
    }
}
static const u64 CACHE_SAMPLES = (u64)1 << 20;
static const size_t CACHE_MAX_ENTRIES = 4;                // a genProof has at most a handful of commit stages in flight
static const size_t CACHE_MAX_BYTES = (size_t)5 << 30;    // of pinned host memory held for trees nobody has claimed yet
static inline u64 mix64(u64 x)
{
    x ^= x >> 30;
    x *= 0xBF58476D1CE4E5B9ull;
    x ^= x >> 27;
    x *= 0x94D049BB133111EBull;
    x ^= x >> 31;
    return x;
}
// order-independent keyed checksum of the host buffer: all words (full) or CACHE_SAMPLES pseudo-random positions
static u64 cache_checksum(const u64 *host, u64 total, u64 seed, bool full)
{
    const u64 count = full ? total : CACHE_SAMPLES;
    unsigned nt = std::thread::hardware_concurrency();
    if (nt > (full ? 32u : 8u)) nt = full ? 32 : 8;
    if (nt < 1 || count < ((u64)1 << 16)) nt = 1;
    std::vector<u64> part(nt, 0);
    auto work = [&](unsigned t) {
        const u64 lo = count * t / nt, hi = count * (t + 1) / nt;
        u64 acc = 0;
        for (u64 k = lo; k < hi; k++)
        {
            const u64 pos = full ? k : mix64(seed + k) % total;
            acc += mix64(host[pos] ^ (seed + pos * 0x9E3779B97F4A7C15ull));
        }
        part[t] = acc;
    };
    std::vector<std::thread> th;
    for (unsigned t = 1; t < nt; t++) th.emplace_back(work, t);
    work(0);
    for (auto &t : th) t.join();
    u64 acc = 0;
    for (u64 v : part) acc += v;
    return acc;
}
static void cache_fingerprint(Resident &r)
{
    static std::atomic<u64> counter{0};
    const char *e = getenv("ZKGPU_CACHE_VERIFY");
    const u64 total = r.nrows * r.ncols;
    r.full = (e && strcmp(e, "full") == 0) || total <= CACHE_SAMPLES;
    r.seed = mix64((u64)std::chrono::steady_clock::now().time_since_epoch().count() ^ mix64((u64)(uintptr_t)r.host + counter.fetch_add(1)));
    r.sum = cache_checksum(r.host, total, r.seed, r.full);
    const u64 picks[2] = {0, r.nrows - 1};
    for (u64 i : picks)
        if (i < r.nrows && (r.rows_idx.empty() || r.rows_idx.back() != i))
        {
            r.rows_idx.push_back(i);
            r.rows_words.insert(r.rows_words.end(), r.host + i * r.ncols, r.host + (i + 1) * r.ncols);
        }
}
static size_t cache_bytes_locked()
{
    size_t b = 0;
    for (auto &r : g_cache) b += (2 * r.nrows - 1) * 4 * sizeof(u64);
    return b;
}
static void cache_put(Resident &&r)
{
    std::lock_guard<std::mutex> lk(g_cache_mu);
    const size_t mine = (2 * r.nrows - 1) * 4 * sizeof(u64);
    while (!g_cache.empty() && (g_cache.size() >= CACHE_MAX_ENTRIES || cache_bytes_locked() + mine > CACHE_MAX_BYTES))
    {
        g_unclaimed[g_cache.front().handle]++;
        g_cache_stats[3]++;
        cache_free(g_cache.front());
        g_cache.erase(g_cache.begin());
    }
    g_cache.push_back(std::move(r));
}
// May extendPol of this NTT object hash on the way?  Not once two of its trees have gone unclaimed: the caller evidently does
// not build Goldilocks trees on these buffers (StarkRecursiveF extends and then builds a BN128 tree,
// starkRecursiveF.cpp:198-200), and the sponge would be wasted work.  In genProof every extendPol is followed by its
// merkelize() before the next extendPol, so entries of this handle that are still waiting when a new one arrives count too.
static bool cache_wanted(const void *handle)
{
    std::lock_guard<std::mutex> lk(g_cache_mu);
    int waiting = 0;
    for (auto &r : g_cache)
        if (r.handle == handle) waiting++;
    if (waiting >= 2 && g_unclaimed[handle] < 2)
    {
        g_unclaimed[handle] = 2;
        for (size_t i = 0; i < g_cache.size();)
            if (g_cache[i].handle == handle)
            {
                g_cache_stats[3]++;
                cache_free(g_cache[i]);
                g_cache.erase(g_cache.begin() + i);
            }
            else
                i++;
    }
    return g_unclaimed[handle] < 2;
}
static void cache_forget_handle(const void *handle)
{
    std::lock_guard<std::mutex> lk(g_cache_mu);
    g_unclaimed.erase(handle);
}
// the entry for this buffer if it is still what extendPol wrote; removed from the cache (the caller owns it now)
static bool cache_take(const u64 *host, u64 nrows, u64 ncols, Resident *out)
{
    Resident r;
    {
        std::lock_guard<std::mutex> lk(g_cache_mu);
        size_t i = 0;
        for (; i < g_cache.size(); i++)
            if (g_cache[i].host == host && g_cache[i].nrows == nrows && g_cache[i].ncols == ncols) break;
        if (i == g_cache.size())
        {
            g_cache_stats[1]++;
            return false;
        }
        r = std::move(g_cache[i]);
        g_cache.erase(g_cache.begin() + i);
    }
    bool same = true;
    for (size_t k = 0; k < r.rows_idx.size() && same; k++)
        same = memcmp(host + r.rows_idx[k] * ncols, r.rows_words.data() + k * ncols, ncols * sizeof(u64)) == 0;
    if (same) same = cache_checksum(host, nrows * ncols, r.seed, r.full) == r.sum; // outside the lock: may read the whole buffer
    std::lock_guard<std::mutex> lk(g_cache_mu);
    if (same)
    {
        g_cache_stats[0]++;
        g_unclaimed[r.handle] = 0;
        *out = std::move(r);
    }
    else
    {
        g_cache_stats[2]++;
        cache_free(r);
    }
    return same;
}

// "device pointers only" entry points: reject host pointers and mis-aligned digest arrays (the kernels access leaves and
// tree as 16-byte ulonglong2) with rc < 0 instead of a sticky device fault
static int need_dev(const char *fn, const void *p, const char *what, bool align16)
{
    if (!is_device_ptr(p))
    {
        set_error("%s: %s must be a device pointer", fn, what);
        return -1;
    }
    if (align16 && ((uintptr_t)p & 15))
    {
        set_error("%s: %s must be 16-byte aligned", fn, what);
        return -1;
    }
    return 0;
}

} // namespace zk

using namespace zk;

extern "C" {

// ZKGPU_NGPU=n in the environment makes a single-process host (zkevm-prover itself) use n GPUs without any code change:
// checked once, by the first commit-type call
static void multi_from_env()
{
    static std::once_flag once;
    std::call_once(once, []() {
        const char *e = getenv("ZKGPU_NGPU");
        if (e && atoi(e) > 1 && dist_world_of_process() == 0)
            if (zkgpu_init_multi(atoi(e)) < 0) fprintf(stderr, "[zkgpu] ZKGPU_NGPU=%s ignored: %s\n", e, zkgpu_last_error());
    });
}

int zkgpu_init(int device)
{
    // device >= 0 binds the calling host thread to that device's context; -1 keeps the thread's (or the process default) one
    zk::Enter en(device);
    return en.rc;
}
int zkgpu_set_stream(void *s)
{
    ZK_ENTER;
    C.cur = (s == (void *)-1) ? C.own : (cudaStream_t)s; // 0 is the legacy default stream, a valid choice
    return 0;
}
int zkgpu_sync(void)
{
    ZK_ENTER;
    ZK_CUDA(cudaStreamSynchronize(C.cur));
    return 0;
}
const char *zkgpu_last_error(void) { return last_error(); }
uint64_t zkgpu_launch_count(void) { return g_launches; }
int zkgpu_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess)
    {
        cudaGetLastError();
        return 0;
    }
    return n;
}
int zkgpu_mem_info(uint64_t *free_bytes, uint64_t *total_bytes)
{
    ZK_ENTER;
    size_t f = 0, t = 0;
    ZK_CUDA(cudaMemGetInfo(&f, &t));
    if (free_bytes) *free_bytes = f;
    if (total_bytes) *total_bytes = t;
    return 0;
}
int zkgpu_malloc(void **p, uint64_t bytes)
{
    ZK_ENTER;
    ZK_CUDA(cudaMalloc(p, bytes ? bytes : 8));
    return 0;
}
int zkgpu_free(void *p)
{
    ZK_ENTER;
    ZK_CUDA(cudaStreamSynchronize(C.cur));
    ZK_CUDA(cudaFree(p));
    return 0;
}
int zkgpu_malloc_host(void **p, uint64_t bytes)
{
    ZK_ENTER;
    ZK_CUDA(cudaMallocHost(p, bytes ? bytes : 8));
    return 0;
}
int zkgpu_free_host(void *p)
{
    ZK_ENTER;
    ZK_CUDA(cudaFreeHost(p));
    return 0;
}
int zkgpu_memcpy(void *dst, const void *src, uint64_t bytes)
{
    ZK_ENTER;
    ZK_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, C.cur));
    ZK_CUDA(cudaStreamSynchronize(C.cur));
    return 0;
}
int zkgpu_stage_ms(float out[5])
{
    ZK_ENTER;
    ZK_CUDA(cudaStreamSynchronize(C.cur));
    if (C.timed && !C.segs.empty())
    {
        C.stage[0] = C.stage[1] = 0;
        for (const auto &sg : C.segs)
        {
            float ms = 0;
            ZK_CUDA(cudaEventElapsedTime(&ms, C.pool[sg.a], C.pool[sg.b]));
            C.stage[sg.kind] += ms;
        }
        ZK_CUDA(cudaEventElapsedTime(&C.stage[2], C.ev[2], C.ev[3]));
        C.segs.clear();
        C.timed = false;
    }
    if (C.timed)
    {
        ZK_CUDA(cudaEventElapsedTime(&C.stage[0], C.ev[0], C.ev[1]));
        ZK_CUDA(cudaEventElapsedTime(&C.stage[1], C.ev[1], C.ev[2]));
        ZK_CUDA(cudaEventElapsedTime(&C.stage[2], C.ev[2], C.ev[3]));
        C.timed = false;
    }
    for (int i = 0; i < 5; i++) out[i] = C.stage[i];
    return 0;
}
int zkgpu_fill(uint64_t *dst, uint64_t n, uint64_t seed, uint64_t offset)
{
    ZK_ENTER;
    if (is_device_ptr(dst)) return fill_splitmix(dst, n, seed, offset, C.cur);
    Staged d;
    ZK_TRY(d.alloc(n));
    ZK_TRY(fill_splitmix(d.d, n, seed, offset, C.cur));
    ZK_CUDA(cudaMemcpyAsync(dst, d.d, n * sizeof(u64), cudaMemcpyDeviceToHost, C.cur));
    ZK_CUDA(cudaStreamSynchronize(C.cur));
    return 0;
}

int zkgpu_field_op(int op, uint64_t *out, const uint64_t *a, const uint64_t *b, uint64_t n)
{
    if (n == 0) return 0;
    ZK_ENTER;
    cudaStream_t st = C.cur;
    const bool dev = is_device_ptr(out);
    Staged da, db;
    const u64 *pa = a, *pb = b;
    u64 *po = out;
    if (!dev)
    {
        ZK_TRY(da.alloc(n));
        ZK_TRY(db.alloc(n));
        ZK_CUDA(cudaMemcpyAsync(da.d, a, n * sizeof(u64), cudaMemcpyHostToDevice, st));
        ZK_CUDA(cudaMemcpyAsync(db.d, b, n * sizeof(u64), cudaMemcpyHostToDevice, st));
        pa = da.d;
        pb = db.d;
        po = da.d;
    }
    u64 blocks = (n + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    field_op_kernel<<<(unsigned)blocks, 256, 0, st>>>(op, po, pa, pb, n);
    ZK_LAUNCHED();
    ZK_CUDA(cudaGetLastError());
    if (!dev)
    {
        ZK_CUDA(cudaMemcpyAsync(out, po, n * sizeof(u64), cudaMemcpyDeviceToHost, st));
        ZK_CUDA(cudaStreamSynchronize(st));
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------------
struct NttHandle
{
    u64 max_domain;
    int extension;
};
void *zkgpu_ntt_create(uint64_t max_domain_size, int extension)
{
    if (max_domain_size && ilog2(max_domain_size) > 32)
    {
        set_error("Domain size too big for the curve"); // ntt_goldilocks.hpp:94-97
        return nullptr;
    }
    return new NttHandle{max_domain_size, extension};
}
void zkgpu_ntt_destroy(void *h)
{
    cache_forget_handle(h); // the address may come back for another object
    delete (NttHandle *)h;
}

int zkgpu_ntt(void *, uint64_t *dst, const uint64_t *src, uint64_t size, uint64_t ncols, uint64_t *, int inverse, int extend)
{
    if (size == 0 || ncols == 0) return 0; // ntt_goldilocks.cpp:168-171
    if (!is_pow2(size))
    {
        set_error("NTT: size %llu is not a power of two", (unsigned long long)size); // reference asserts
        return -1;
    }
    if (!src)
    {
        set_error("NTT: src is NULL");
        return -1;
    }
    ZK_ENTER;
    if (!dst) dst = const_cast<uint64_t *>(src);
    const unsigned L = ilog2(size);
    cudaStream_t st = C.cur;
    Scale pre, post;
    if (inverse)
    {
        const u64 ninv = gl::pow_slow(size % GL_P, GL_P - 2);
        if (extend)
        {
            post.mode = SCALE_POWER;
            ZK_TRY(get_power_tables(GL_SHIFT, ninv, L, &post.G0, &post.G1, st));
        }
        else if (size > 1)
        {
            post.mode = SCALE_CONST;
            post.konst = ninv;
        }
    }
    const bool dev = is_device_ptr(src);
    if (dev != is_device_ptr(dst))
    {
        set_error("NTT: dst and src must both be host or both be device pointers");
        return -1;
    }
    const u64 words = size * ncols;
    if (dev)
    {
        if (L > 8) ZK_TRY(ctx_ensure_scratch(C, words));
        return ntt_transform(dst, ncols, src, ncols, C.scratch, L, ncols, inverse != 0, pre, post, st);
    }
    // host pointers: the device image and the transform's scratch both live in the context's scratch area (no allocation
    // per call: stage 4 and FRI call this in a loop, starks.cpp:262-285, friProve.cpp:100-102)
    ZK_TRY(ctx_ensure_scratch(C, 2 * words));
    u64 *d = C.scratch, *tmp = C.scratch + words;
    ZK_CUDA(cudaMemcpyAsync(d, src, words * sizeof(u64), cudaMemcpyHostToDevice, st));
    ZK_TRY(ntt_transform(d, ncols, d, ncols, tmp, L, ncols, inverse != 0, pre, post, st));
    ZK_CUDA(cudaMemcpyAsync(dst, d, words * sizeof(u64), cudaMemcpyDeviceToHost, st));
    ZK_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int zkgpu_reverse_permutation(void *h, uint64_t *dst, const uint64_t *src, uint64_t size, uint64_t offset_cols, uint64_t ncols,
                              uint64_t ncols_all)
{
    if (size == 0 || ncols == 0) return 0;
    if (!is_pow2(size))
    {
        set_error("reversePermutation: size %llu is not a power of two", (unsigned long long)size);
        return -1;
    }
    const int extension = h ? ((NttHandle *)h)->extension : 1;
    const bool inplace = dst == src || dst == nullptr;
    if (inplace && (extension > 1 || offset_cols != 0 || ncols != ncols_all))
    {
        set_error("reversePermutation: in place needs extension <= 1, offset_cols == 0 and ncols == ncols_all"); // :283, :302
        return -1;
    }
    ZK_ENTER;
    cudaStream_t st = C.cur;
    if (!dst) dst = const_cast<uint64_t *>(src);
    const unsigned L = ilog2(size);
    const bool dev = is_device_ptr(src);
    if (dev != is_device_ptr(dst))
    {
        set_error("reversePermutation: dst and src must both be host or both be device pointers");
        return -1;
    }
    // rows of the source the reference may touch: BR(i) < size, i.e. all `size` rows of ncols_all words
    const u64 src_words = size * ncols_all, dst_words = size * ncols;
    const u64 ext_limit = extension > 1 ? (size / (u64)extension) * ncols_all : ~(u64)0;
    Staged ds, dd;
    const u64 *ps = src;
    u64 *pd = dst;
    if (!dev)
    {
        const u64 need = extension > 1 ? (ext_limit < src_words ? ext_limit : src_words) : src_words;
        ZK_TRY(ds.alloc(src_words));
        ZK_CUDA(cudaMemcpyAsync(ds.d, src, need * sizeof(u64), cudaMemcpyHostToDevice, st));
        ps = ds.d;
        if (inplace)
            pd = ds.d;
        else
        {
            ZK_TRY(dd.alloc(dst_words));
            pd = dd.d;
        }
    }
    u64 blocks = (dst_words + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    if (inplace)
        reverse_perm_inplace_kernel<<<(unsigned)blocks, 256, 0, st>>>(pd, size, L, ncols);
    else
        reverse_perm_kernel<<<(unsigned)blocks, 256, 0, st>>>(pd, ps, size, L, offset_cols, ncols, ncols_all, ext_limit);
    ZK_LAUNCHED();
    ZK_CUDA(cudaGetLastError());
    if (!dev)
    {
        ZK_CUDA(cudaMemcpyAsync(dst, pd, dst_words * sizeof(u64), cudaMemcpyDeviceToHost, st));
        ZK_CUDA(cudaStreamSynchronize(st));
    }
    return 0;
}

static int check_lde_args(uint64_t n_ext, uint64_t n, unsigned *log_n, unsigned *log_b)
{
    if (!is_pow2(n) || !is_pow2(n_ext) || n_ext < n)
    {
        set_error("extendPol: N=%llu, N_Extended=%llu must be powers of two with N_Extended >= N", (unsigned long long)n,
                  (unsigned long long)n_ext);
        return -1;
    }
    *log_n = ilog2(n);
    *log_b = ilog2(n_ext) - *log_n;
    return 0;
}

int zkgpu_lde_columns(uint64_t *out, uint64_t out_ld, const uint64_t *in, uint64_t in_ld, uint64_t n_ext, uint64_t n,
                      uint64_t ncols)
{
    if (n == 0 || ncols == 0) return 0;
    unsigned log_n, log_b;
    ZK_TRY(check_lde_args(n_ext, n, &log_n, &log_b));
    ZK_ENTER;
    if (!is_device_ptr(out) || !is_device_ptr(in))
    {
        set_error("zkgpu_lde_columns: device pointers required");
        return -1;
    }
    return lde_all(C, out, out_ld, in, in_ld, log_n, log_b, ncols, C.cur);
}

int zkgpu_extend_pol(void *ntt_handle, uint64_t *output, const uint64_t *input, uint64_t n_ext, uint64_t n, uint64_t ncols, uint64_t *)
{
    if (n == 0 || ncols == 0 || n_ext == 0) return 0;
    unsigned log_n, log_b;
    ZK_TRY(check_lde_args(n_ext, n, &log_n, &log_b));
    multi_from_env();
    ZK_ENTER;
    cudaStream_t st = C.cur;
    const bool dev = is_device_ptr(input);
    if (dev != is_device_ptr(output))
    {
        set_error("extendPol: output and input must both be host or both be device pointers");
        return -1;
    }
    if (dev)
    {
        C.segs.clear();
        ZK_CUDA(cudaEventRecord(C.ev[0], st));
        ZK_TRY(lde_all(C, output, ncols, input, ncols, log_n, log_b, ncols, st));
        ZK_CUDA(cudaEventRecord(C.ev[1], st));
        ZK_CUDA(cudaEventRecord(C.ev[2], st));
        ZK_CUDA(cudaEventRecord(C.ev[3], st));
        C.timed = true;
        return 0;
    }
    // host pointers: column strips, host->device of strip s+1 and device->host of strip s-1 overlap the LDE of strip s.
    // With the commit cache on, the strips are hashed on the way (see Resident above) for the merkelize() that follows.
    cache_invalidate(output, n_ext * ncols * sizeof(u64));
    const bool fuse = fuse_enabled() && ncols > 4 && cache_wanted(ntt_handle);
    if (dist_world_of_process() > 1 && fuse && n_ext >= (u64)dist_world_of_process() && output != input)
    {
        // multi-GPU job (zkgpu_init_multi / ZKGPU_NGPU): all GPUs extend and hash; the tree waits in pinned host memory
        en__.lk.unlock(); // the worker threads lock their own contexts, device 0's included
        Resident r;
        r.host = output;
        r.nrows = n_ext;
        r.ncols = ncols;
        r.handle = ntt_handle;
        r.h_tree = pinned_take((2 * n_ext - 1) * 4);
        if (!r.h_tree)
        {
            set_error("extendPol: cannot allocate the pinned tree buffer");
            return -4;
        }
        const int rc = dist_multi_commit(output, r.h_tree, input, n_ext, n, ncols);
        if (rc < 0)
        {
            cache_free(r);
            return rc;
        }
        cache_fingerprint(r);
        cache_put(std::move(r));
        return 0;
    }
    u64 *d_tree = nullptr;
    if (fuse)
    {
        if (cudaMalloc(&d_tree, (2 * n_ext - 1) * 4 * sizeof(u64)) != cudaSuccess)
        {
            cudaGetLastError();
            d_tree = nullptr; // no room: plain extendPol
        }
    }
    int rc = commit_pipeline(C, output, false, d_tree, input, false, log_n, log_b, ncols, st);
    if (rc >= 0)
    {
        cudaEventRecord(C.ev[2], st);
        if (d_tree) rc = merkle_levels(d_tree, n_ext, st);
        cudaEventRecord(C.ev[3], st);
        C.timed = true;
    }
    cudaError_t e = cudaStreamSynchronize(st);
    if (rc < 0 || e != cudaSuccess)
    {
        if (d_tree) cudaFree(d_tree);
        if (rc >= 0) ZK_CUDA(e);
        return rc;
    }
    if (d_tree)
    {
        Resident r;
        r.host = output;
        r.nrows = n_ext;
        r.ncols = ncols;
        r.handle = ntt_handle;
        // the tree waits for its merkelize() in pinned host memory: an entry that is never claimed holds no device memory
        const size_t tree_words = (2 * n_ext - 1) * 4;
        r.h_tree = pinned_take(tree_words);
        if (r.h_tree && cudaMemcpy(r.h_tree, d_tree, tree_words * sizeof(u64), cudaMemcpyDeviceToHost) == cudaSuccess)
            cudaFree(d_tree);
        else
        {
            cudaGetLastError();
            if (r.h_tree) pinned_give(r.h_tree, tree_words);
            r.h_tree = nullptr;
            r.d_tree = d_tree;
        }
        cache_fingerprint(r);
        cache_put(std::move(r));
    }
    return 0;
}

int zkgpu_invalidate(const void *host_ptr)
{
    cache_invalidate(host_ptr, 0);
    return 0;
}
int zkgpu_cache_stats(uint64_t out[6])
{
    std::lock_guard<std::mutex> lk(g_cache_mu);
    for (int i = 0; i < 4; i++) out[i] = g_cache_stats[i];
    out[4] = g_cache.size();
    u64 dev = 0;
    for (auto &r : g_cache)
        if (r.d_tree) dev += (2 * r.nrows - 1) * 4 * sizeof(u64);
    out[5] = dev;
    return 0;
}

// ---------------------------------------------------------------------------------------------------
int zkgpu_hash_full_result(uint64_t *out12, const uint64_t *in12, uint64_t n)
{
    if (n == 0) return 0;
    ZK_ENTER;
    cudaStream_t st = C.cur;
    if (is_device_ptr(in12)) return poseidon_permute(out12, in12, n, st);
    Staged d;
    ZK_TRY(d.alloc(12 * n));
    ZK_CUDA(cudaMemcpyAsync(d.d, in12, 12 * n * sizeof(u64), cudaMemcpyHostToDevice, st));
    ZK_TRY(poseidon_permute(d.d, d.d, n, st));
    ZK_CUDA(cudaMemcpyAsync(out12, d.d, 12 * n * sizeof(u64), cudaMemcpyDeviceToHost, st));
    ZK_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int zkgpu_linear_hash(uint64_t *out4, const uint64_t *input, uint64_t size)
{
    ZK_ENTER;
    cudaStream_t st = C.cur;
    if (is_device_ptr(input)) return merkle_leaves(out4, input, size, size, 1, st);
    Staged d;
    ZK_TRY(d.alloc(size + 4));
    if (size) ZK_CUDA(cudaMemcpyAsync(d.d + 4, input, size * sizeof(u64), cudaMemcpyHostToDevice, st));
    ZK_TRY(merkle_leaves(d.d, d.d + 4, size, size, 1, st));
    ZK_CUDA(cudaMemcpyAsync(out4, d.d, 4 * sizeof(u64), cudaMemcpyDeviceToHost, st));
    ZK_CUDA(cudaStreamSynchronize(st));
    return 0;
}

static int merkletree_impl(uint64_t *tree, const uint64_t *input, uint64_t ncols, uint64_t nrows, uint64_t batch, uint64_t dim,
                           bool batched)
{
    if (nrows == 0) return 0; // poseidon_goldilocks.cpp:268-271
    if (!is_pow2(nrows))
    {
        set_error("merkletree: num_rows %llu is not a power of two", (unsigned long long)nrows);
        return -1;
    }
    const u64 row_words = ncols * dim;
    const u64 tree_words = (2 * nrows - 1) * 4;
    const bool in_dev = row_words == 0 || is_device_ptr(input);
    if (!in_dev && !batched)
    {
        // merkelize() of the buffer extendPol just produced: the tree is already there (commit cache)
        Resident r;
        if (cache_take(input, nrows, row_words, &r))
        {
            int rc = 0;
            if (r.h_tree)
            {
                if (is_device_ptr(tree))
                    rc = zkgpu_memcpy(tree, r.h_tree, tree_words * sizeof(u64));
                else
                    memcpy(tree, r.h_tree, tree_words * sizeof(u64));
            }
            else
            {
                Enter en(ptr_device(r.d_tree));
                rc = en.rc;
                if (rc >= 0 && cudaMemcpy(tree, r.d_tree, tree_words * sizeof(u64), cudaMemcpyDefault) != cudaSuccess)
                {
                    set_error("merkletree: copying the cached tree failed: %s", cudaGetErrorString(cudaGetLastError()));
                    rc = -2;
                }
            }
            cache_free(r);
            if (getenv("ZKGPU_PIPE_TRACE")) fprintf(stderr, "[zkgpu cache] merkletree(%p): tree taken from the commit cache, no upload\n", (const void *)input);
            return rc;
        }
    }
    ZK_ENTER;
    cudaStream_t st = C.cur;
    const bool tree_dev = is_device_ptr(tree);
    if (tree_dev && ((uintptr_t)tree & 15))
    {
        set_error("merkletree: a device tree pointer must be 16-byte aligned");
        return -1;
    }
    Staged dtree;
    u64 *d_tree = tree;
    if (!tree_dev)
    {
        ZK_TRY(dtree.alloc(tree_words));
        d_tree = dtree.d;
    }
    auto leaves_of = [&](u64 *leaves, const u64 *rows, u64 cnt) -> int {
        if (batched) return merkle_leaves_batch(leaves, rows, ncols, dim, batch, cnt, st);
        return merkle_leaves(leaves, rows, row_words, row_words, cnt, st);
    };
    ZK_CUDA(cudaEventRecord(C.ev[0], st));
    ZK_CUDA(cudaEventRecord(C.ev[1], st));
    C.segs.clear();
    if (in_dev)
        ZK_TRY(leaves_of(d_tree, input, nrows));
    else
    {
        // host rows: row chunks (~1 GiB), host->device copy of chunk k+1 overlaps the hashing of chunk k
        ZK_TRY(ctx_pipe_streams(C));
        u64 chunk = ((u64)1 << 27) / row_words;
        if (chunk == 0) chunk = 1;
        if (chunk > nrows) chunk = nrows;
        ZK_TRY(ctx_ensure_scratch(C, 2 * chunk * row_words));
        u64 *buf[2] = {C.scratch, C.scratch + chunk * row_words};
        const u64 nchunks = (nrows + chunk - 1) / chunk;
        auto issue = [&](u64 k) -> int {
            const int b = (int)(k & 1);
            const u64 r0 = k * chunk, cnt = nrows - r0 < chunk ? nrows - r0 : chunk;
            if (k >= 2) ZK_CUDA(cudaStreamWaitEvent(C.h2d, C.in_free[b], 0));
            ZK_CUDA(cudaMemcpyAsync(buf[b], input + r0 * row_words, cnt * row_words * sizeof(u64), cudaMemcpyHostToDevice, C.h2d));
            ZK_CUDA(cudaEventRecord(C.in_ready[b], C.h2d));
            return 0;
        };
        ZK_CUDA(cudaEventRecord(C.in_free[0], st));
        ZK_CUDA(cudaStreamWaitEvent(C.h2d, C.in_free[0], 0));
        ZK_TRY(issue(0));
        for (u64 k = 0; k < nchunks; k++)
        {
            const int b = (int)(k & 1);
            const u64 r0 = k * chunk, cnt = nrows - r0 < chunk ? nrows - r0 : chunk;
            if (k + 1 < nchunks) ZK_TRY(issue(k + 1));
            ZK_CUDA(cudaStreamWaitEvent(st, C.in_ready[b], 0));
            ZK_TRY(leaves_of(d_tree + 4 * r0, buf[b], cnt));
            ZK_CUDA(cudaEventRecord(C.in_free[b], st));
        }
    }
    ZK_CUDA(cudaEventRecord(C.ev[2], st));
    ZK_TRY(merkle_levels(d_tree, nrows, st));
    ZK_CUDA(cudaEventRecord(C.ev[3], st));
    C.timed = true;
    if (!tree_dev) ZK_CUDA(cudaMemcpyAsync(tree, d_tree, tree_words * sizeof(u64), cudaMemcpyDeviceToHost, st));
    if (!tree_dev || !in_dev) ZK_CUDA(cudaStreamSynchronize(st));
    return 0;
}
int zkgpu_merkletree(uint64_t *tree, const uint64_t *input, uint64_t ncols, uint64_t nrows, uint64_t dim)
{
    return merkletree_impl(tree, input, ncols, nrows, 0, dim, false);
}
int zkgpu_merkletree_batch(uint64_t *tree, const uint64_t *input, uint64_t ncols, uint64_t nrows, uint64_t batch, uint64_t dim)
{
    return merkletree_impl(tree, input, ncols, nrows, batch, dim, true);
}

int zkgpu_group_proof(uint64_t *proof, const uint64_t *source, const uint64_t *nodes, uint64_t width, uint64_t height, uint64_t idx)
{
    if (height == 0 || idx >= height)
    {
        set_error("getGroupProof: idx %llu out of range (height %llu)", (unsigned long long)idx, (unsigned long long)height);
        return -1;
    }
    ZK_ENTER;
    cudaStream_t st = C.cur;
    const unsigned levels = ilog2(height);
    const u64 proof_words = width + 4 * (u64)levels;
    const bool src_dev = width == 0 || is_device_ptr(source), nodes_dev = levels == 0 || is_device_ptr(nodes), proof_dev = is_device_ptr(proof);
    if (!src_dev && !nodes_dev && !proof_dev)
    {
        // everything on the host: a handful of words, no device work to do (what the reference does, merkleTreeGL.cpp:12-35)
        for (u64 c = 0; c < width; c++) proof[c] = source[idx * width + c];
        u64 off = 0, n = height, i = idx;
        for (unsigned l = 0; l < levels; l++, off += n * 4, n >>= 1, i >>= 1)
            for (int k = 0; k < 4; k++) proof[width + 4 * l + k] = nodes[off + (i ^ 1) * 4 + k];
        return 0;
    }
    // stage only what the proof needs from host-side buffers
    Staged drow, dproof;
    const u64 *d_row = source + idx * width;
    u64 *d_proof = proof;
    if (!proof_dev)
    {
        ZK_TRY(dproof.alloc(proof_words));
        d_proof = dproof.d;
    }
    if (!src_dev)
    {
        ZK_TRY(drow.alloc(width));
        ZK_CUDA(cudaMemcpyAsync(drow.d, source + idx * width, width * sizeof(u64), cudaMemcpyHostToDevice, st));
        d_row = drow.d;
    }
    if (!nodes_dev)
    {
        // host-resident nodes: copy the sibling digests straight into the device proof
        u64 off = 0, n = height, i = idx;
        for (unsigned l = 0; l < levels; l++, off += n * 4, n >>= 1, i >>= 1)
            ZK_CUDA(cudaMemcpyAsync(d_proof + width + 4 * l, nodes + off + (i ^ 1) * 4, 32, cudaMemcpyHostToDevice, st));
    }
    group_proof_kernel<<<1, 256, 0, st>>>(d_proof, d_row, nodes_dev ? nodes : nullptr, width, height, idx);
    ZK_LAUNCHED();
    ZK_CUDA(cudaGetLastError());
    if (!proof_dev)
    {
        ZK_CUDA(cudaMemcpyAsync(proof, d_proof, proof_words * sizeof(u64), cudaMemcpyDeviceToHost, st));
        ZK_CUDA(cudaStreamSynchronize(st));
    }
    return 0;
}

int zkgpu_group_proofs(uint64_t *proofs, const uint64_t *source, const uint64_t *nodes, uint64_t width, uint64_t height, const uint64_t *idx,
                       uint64_t nqueries)
{
    if (nqueries == 0) return 0;
    if (height == 0 || !idx || !proofs)
    {
        set_error("getGroupProof (batch): NULL argument or empty tree");
        return -1;
    }
    for (u64 q = 0; q < nqueries; q++)
        if (idx[q] >= height)
        {
            set_error("getGroupProof (batch): idx[%llu] = %llu out of range (height %llu)", (unsigned long long)q, (unsigned long long)idx[q],
                      (unsigned long long)height);
            return -1;
        }
    const unsigned levels = ilog2(height);
    const u64 proof_words = width + 4 * (u64)levels;
    const bool src_dev = width == 0 || is_device_ptr(source), nodes_dev = levels == 0 || is_device_ptr(nodes);
    if (!src_dev || !nodes_dev)
    {
        // buffers on the host (or mixed): the single-proof entry handles every combination
        for (u64 q = 0; q < nqueries; q++) ZK_TRY(zkgpu_group_proof(proofs + q * proof_words, source, nodes, width, height, idx[q]));
        return 0;
    }
    ZK_ENTER;
    cudaStream_t st = C.cur;
    const bool proof_dev = is_device_ptr(proofs);
    ZK_TRY(ctx_ensure_scratch(C, nqueries * (proof_words + 1)));
    u64 *d_idx = C.scratch, *d_proofs = proof_dev ? proofs : C.scratch + nqueries;
    ZK_CUDA(cudaMemcpyAsync(d_idx, idx, nqueries * sizeof(u64), cudaMemcpyHostToDevice, st));
    group_proofs_kernel<<<(unsigned)nqueries, 256, 0, st>>>(d_proofs, source, nodes, width, height, d_idx, proof_words);
    ZK_LAUNCHED();
    ZK_CUDA(cudaGetLastError());
    if (!proof_dev) ZK_CUDA(cudaMemcpyAsync(proofs, d_proofs, nqueries * proof_words * sizeof(u64), cudaMemcpyDeviceToHost, st));
    ZK_CUDA(cudaStreamSynchronize(st)); // idx is the caller's host buffer
    return 0;
}

int zkgpu_merkle_leaves(uint64_t *leaves, const uint64_t *rows, uint64_t row_words, uint64_t row_ld, uint64_t nrows)
{
    ZK_ENTER;
    if (nrows == 0) return 0;
    ZK_TRY(need_dev("merkle_leaves", leaves, "leaves", true));
    if (row_words) ZK_TRY(need_dev("merkle_leaves", rows, "rows", false));
    return merkle_leaves(leaves, rows, row_words, row_ld, nrows, C.cur);
}
int zkgpu_merkle_absorb(uint64_t *leaves, const uint64_t *rows, uint64_t strip_words, uint64_t row_ld, uint64_t nrows, int first)
{
    ZK_ENTER;
    if (nrows == 0 || strip_words == 0) return 0;
    ZK_TRY(need_dev("merkle_absorb", leaves, "leaves", true));
    ZK_TRY(need_dev("merkle_absorb", rows, "rows", false));
    return merkle_absorb(leaves, rows, strip_words, row_ld, nrows, first != 0, C.cur);
}
int zkgpu_merkle_absorb_copy(uint64_t *leaves, const uint64_t *rows, uint64_t strip_words, uint64_t row_ld, uint64_t nrows, int first,
                             uint64_t *copy_to, uint64_t copy_ld)
{
    ZK_ENTER;
    if (!copy_to)
    {
        set_error("merkle_absorb_copy: null destination");
        return -1;
    }
    if (nrows == 0 || strip_words == 0) return 0;
    ZK_TRY(need_dev("merkle_absorb_copy", leaves, "leaves", true));
    ZK_TRY(need_dev("merkle_absorb_copy", rows, "rows", false));
    ZK_TRY(need_dev("merkle_absorb_copy", copy_to, "copy_to", false));
    return merkle_absorb(leaves, rows, strip_words, row_ld, nrows, first != 0, C.cur, copy_to, copy_ld);
}
int zkgpu_merkle_levels(uint64_t *tree, uint64_t nrows)
{
    if (nrows == 0) return 0;
    if (!is_pow2(nrows))
    {
        set_error("merkle_levels: num_rows %llu is not a power of two", (unsigned long long)nrows);
        return -1;
    }
    ZK_ENTER;
    ZK_TRY(need_dev("merkle_levels", tree, "tree", true));
    return merkle_levels(tree, nrows, C.cur);
}
int zkgpu_copy_block(uint64_t *dst, uint64_t dst_ld, const uint64_t *src, uint64_t src_ld, uint64_t nrows, uint64_t ncols)
{
    ZK_ENTER;
    const u64 total = nrows * ncols;
    if (total == 0) return 0;
    ZK_TRY(need_dev("copy_block", dst, "dst", false));
    ZK_TRY(need_dev("copy_block", src, "src", false));
    u64 blocks = (total + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    copy_block_kernel<<<(unsigned)blocks, 256, 0, C.cur>>>(dst, dst_ld, src, src_ld, nrows, ncols);
    ZK_LAUNCHED();
    ZK_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------------
int zkgpu_extend_and_merkelize(void *, uint64_t *out_ext, uint64_t *tree, const uint64_t *input, uint64_t n_ext, uint64_t n,
                               uint64_t ncols)
{
    if (n == 0 || n_ext == 0) return 0;
    unsigned log_n, log_b;
    ZK_TRY(check_lde_args(n_ext, n, &log_n, &log_b));
    // after zkgpu_init_multi a host-resident trace is committed by all GPUs of the job (dist.cu)
    multi_from_env();
    if (dist_world_of_process() > 1 && ncols > 0 && !is_device_ptr(input) && n_ext >= (u64)dist_world_of_process())
        return dist_multi_commit(out_ext, tree, input, n_ext, n, ncols);
    ZK_ENTER;
    cudaStream_t st = C.cur;
    // each pointer may independently be host or device: e.g. host trace in, extended trace kept resident on the
    // device (what getGroupProof / stage 4 need later), tree back on the host.
    const bool in_dev = ncols == 0 || is_device_ptr(input);
    const bool out_dev = ncols == 0 || is_device_ptr(out_ext);
    const bool tree_dev = is_device_ptr(tree);
    if (tree_dev && ((uintptr_t)tree & 15))
    {
        set_error("commit: a device tree pointer must be 16-byte aligned");
        return -1;
    }
    const u64 tree_words = (2 * n_ext - 1) * 4;
    u64 leaves_sent = 0;
    Staged dtree;
    u64 *d_tree = tree;
    if (!tree_dev)
    {
        ZK_TRY(dtree.alloc(tree_words));
        d_tree = dtree.d;
    }
    ZK_CUDA(cudaEventRecord(C.ev[4], st));
    if (ncols <= 4)
    {
        // rows of <= 4 words are copied, not hashed (poseidon_goldilocks.cpp:234-239): no sponge to pipeline
        Staged din, dout;
        const u64 *d_in = input;
        u64 *d_out = out_ext;
        if (!in_dev)
        {
            ZK_TRY(din.alloc(n * ncols));
            ZK_CUDA(cudaMemcpyAsync(din.d, input, n * ncols * sizeof(u64), cudaMemcpyHostToDevice, st));
            d_in = din.d;
        }
        if (!out_dev)
        {
            ZK_TRY(dout.alloc(n_ext * ncols));
            d_out = dout.d;
        }
        ZK_CUDA(cudaEventRecord(C.ev[0], st));
        ZK_TRY(lde_all(C, d_out, ncols, d_in, ncols, log_n, log_b, ncols, st));
        ZK_CUDA(cudaEventRecord(C.ev[1], st));
        ZK_TRY(merkle_all(d_tree, d_out, ncols, n_ext, st, C.ev[2]));
        ZK_CUDA(cudaEventRecord(C.ev[3], st));
        C.segs.clear();
        if (!out_dev) ZK_CUDA(cudaMemcpyAsync(out_ext, d_out, n_ext * ncols * sizeof(u64), cudaMemcpyDeviceToHost, st));
        if (!out_dev || !in_dev) ZK_CUDA(cudaStreamSynchronize(st));
    }
    else
    {
        ZK_TRY(commit_pipeline(C, out_ext, out_dev, d_tree, input, in_dev, log_n, log_b, ncols, st));
        ZK_CUDA(cudaEventRecord(C.ev[2], st));
        if (!tree_dev)
        {
            // the leaf level (half of the tree) goes back to the host while the node levels are hashed
            ZK_CUDA(cudaStreamWaitEvent(C.d2h, C.ev[2], 0));
            ZK_CUDA(cudaMemcpyAsync(tree, d_tree, n_ext * 4 * sizeof(u64), cudaMemcpyDeviceToHost, C.d2h));
            ZK_CUDA(cudaEventRecord(C.out_free[0], C.d2h));
            leaves_sent = n_ext * 4;
        }
        ZK_TRY(merkle_levels(d_tree, n_ext, st));
        ZK_CUDA(cudaEventRecord(C.ev[3], st));
    }
    C.timed = true;
    if (!tree_dev)
    {
        ZK_CUDA(cudaMemcpyAsync(tree + leaves_sent, d_tree + leaves_sent, (tree_words - leaves_sent) * sizeof(u64), cudaMemcpyDeviceToHost, st));
        if (leaves_sent) ZK_CUDA(cudaStreamWaitEvent(st, C.out_free[0], 0));
    }
    ZK_CUDA(cudaEventRecord(C.ev[5], st));
    if (!in_dev || !out_dev || !tree_dev)
    {
        ZK_CUDA(cudaStreamSynchronize(st));
        C.stage[3] = 0;
        ZK_CUDA(cudaEventElapsedTime(&C.stage[4], C.ev[4], C.ev[5])); // whole call on the device clock
    }
    return 0;
}

int zkgpu_stage4_commit(void *, uint64_t *cm4_2ns, uint64_t *tree, const uint64_t *q_2ns, uint64_t n_ext, uint64_t n, uint64_t q_dim,
                        uint64_t q_deg)
{
    if (n == 0 || n_ext == 0 || q_dim == 0 || q_deg == 0) return 0;
    unsigned log_n, log_b;
    ZK_TRY(check_lde_args(n_ext, n, &log_n, &log_b));
    if (q_deg > ((u64)1 << log_b) || q_deg > 64 || q_dim > 1024)
    {
        set_error("stage4: qDeg = %llu exceeds the blowup factor %llu (qq2 rows k*qDeg + p would leave the first N rows)",
                  (unsigned long long)q_deg, (unsigned long long)1 << log_b);
        return -1;
    }
    if (!cm4_2ns || !q_2ns)
    {
        set_error("stage4: NULL argument");
        return -1;
    }
    ZK_ENTER;
    cudaStream_t st = C.cur;
    const unsigned log_e = log_n + log_b;
    const u64 w = q_dim * q_deg, in_words = n_ext * q_dim, out_words = n_ext * w, tree_words = (2 * n_ext - 1) * 4;
    const bool dev = is_device_ptr(q_2ns);
    if (dev != is_device_ptr(cm4_2ns))
    {
        set_error("stage4: cm4_2ns and q_2ns must both be host or both be device pointers");
        return -1;
    }
    const bool tree_dev = tree == nullptr || is_device_ptr(tree);
    if (tree && tree_dev && ((uintptr_t)tree & 15))
    {
        set_error("stage4: a device tree pointer must be 16-byte aligned");
        return -1;
    }
    Staged din, dout, dtree;
    const u64 *d_q = q_2ns;
    u64 *d_out = cm4_2ns, *d_tree = tree;
    if (!dev)
    {
        ZK_TRY(din.alloc(in_words));
        ZK_TRY(dout.alloc(out_words));
        ZK_CUDA(cudaMemcpyAsync(din.d, q_2ns, in_words * sizeof(u64), cudaMemcpyHostToDevice, st));
        d_q = din.d;
        d_out = dout.d;
    }
    if (tree && !tree_dev)
    {
        ZK_TRY(dtree.alloc(tree_words));
        d_tree = dtree.d;
    }
    // scratch: qq1 (NExt x qDim) | qq2 (N x w) | transform scratch max(NExt*qDim, N*w) | sp (qDeg)
    const u64 tmp_words = in_words > n * w ? in_words : n * w;
    ZK_TRY(ctx_ensure_scratch(C, in_words + n * w + tmp_words + 64));
    u64 *qq1 = C.scratch, *qq2 = qq1 + in_words, *tmp = qq2 + n * w, *sp = tmp + tmp_words;
    ZK_CUDA(cudaEventRecord(C.ev[0], st));
    Scale none;
    ZK_TRY(ntt_transform(qq1, q_dim, d_q, q_dim, tmp, log_e, q_dim, true, none, none, st)); // nttExtended.INTT(qq1, q_2ns, NExt, qDim), x NExt
    {
        u64 h[64];
        const u64 shift_in = gl::pow_slow(gl::pow_slow(GL_SHIFT, GL_P - 2), n); // Goldilocks::exp(inv(shift), N), starks.cpp:266
        h[0] = gl::pow_slow(n_ext % GL_P, GL_P - 2);
        for (u64 p = 1; p < q_deg; p++) h[p] = gl::mul_slow(h[p - 1], shift_in);
        ZK_CUDA(cudaMemcpyAsync(sp, h, q_deg * sizeof(u64), cudaMemcpyHostToDevice, st));
        ZK_CUDA(cudaStreamSynchronize(st)); // h[] is a stack buffer
    }
    u64 blocks = (n * w + 255) / 256;
    if (blocks > (u64)C.num_sms * 32) blocks = (u64)C.num_sms * 32;
    stage4_regroup_kernel<<<(unsigned)blocks, 256, 0, st>>>(qq2, qq1, n, (u32)q_dim, (u32)q_deg, sp);
    ZK_LAUNCHED();
    ZK_CUDA(cudaGetLastError());
    // nttExtended.NTT(cm4_2ns, qq2, NExt, qDim*qDeg) of the zero-padded rows = blowup coset transforms of size N
    ZK_TRY(coset_ntts(d_out, w, qq2, w, tmp, log_n, log_b, w, 1, 1, st));
    ZK_CUDA(cudaEventRecord(C.ev[1], st));
    C.segs.clear();
    if (tree)
        ZK_TRY(merkle_all(d_tree, d_out, w, n_ext, st, C.ev[2]));
    else
        ZK_CUDA(cudaEventRecord(C.ev[2], st));
    ZK_CUDA(cudaEventRecord(C.ev[3], st));
    C.timed = true;
    if (!dev) ZK_CUDA(cudaMemcpyAsync(cm4_2ns, d_out, out_words * sizeof(u64), cudaMemcpyDeviceToHost, st));
    if (tree && !tree_dev) ZK_CUDA(cudaMemcpyAsync(tree, d_tree, tree_words * sizeof(u64), cudaMemcpyDeviceToHost, st));
    if (!dev || (tree && !tree_dev)) ZK_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int zkgpu_tune(int key, int value)
{
    switch (key)
    {
    case 0: g_absorb_threads = value == 512 ? 512 : 128; return 0;
    case 1: g_ntt_threads_per_sm_cap = value < 0 ? 0 : value; return 0;
    case 2:
        g_fuse_commit = value ? 1 : 0;
        if (value)
        {
            std::lock_guard<std::mutex> lk(g_cache_mu);
            g_unclaimed.clear(); // switching the cache on again also forgives the handles it had given up on
        }
        return 0; // commit cache: extendPol with host pointers also hashes (default on)
    }
    set_error("zkgpu_tune: unknown key %d", key);
    return -1;
}

int zkgpu_build_const_tree(uint64_t *image, const uint64_t *const_pols, uint64_t n_bits, uint64_t n_bits_ext, uint64_t n_pols)
{
    if (n_bits > 32 || n_bits_ext > 32 || n_bits_ext < n_bits || !image)
    {
        set_error("buildConstTree: need nBits <= nBitsExt <= 32 and an image buffer");
        return -1;
    }
    const u64 n = (u64)1 << n_bits, n_ext = (u64)1 << n_bits_ext;
    {
        // The context is held for the header only.  With a multi-GPU job (zkgpu_init_multi / ZKGPU_NGPU) and host buffers the
        // commit below is done by one worker thread per device, and the worker of this thread's device must be able to
        // take the same context: holding it across the call would block that worker while this thread waits for it.
        ZK_ENTER;
        const u64 header[2] = {n_pols, n_ext}; // constTree[0] = nPols, constTree[1] = nExt (build_const_tree.cpp:576-577)
        ZK_CUDA(cudaMemcpyAsync(image, header, sizeof(header), cudaMemcpyDefault, C.cur));
        ZK_CUDA(cudaStreamSynchronize(C.cur)); // header[] is a stack buffer
    }
    return zkgpu_extend_and_merkelize(nullptr, image + 2, image + 2 + n_pols * n_ext, const_pols, n_ext, n, n_pols);
}

int zkgpu_fri_fold(uint64_t *pol2_e, const uint64_t *fri_pol, uint64_t pol_bits, uint64_t next_bits, uint64_t shift_inv,
                   const uint64_t *special_x)
{
    if (!pol2_e || !fri_pol || !special_x)
    {
        set_error("fri_fold: NULL argument");
        return -1;
    }
    if (next_bits > pol_bits || pol_bits > 32)
    {
        set_error("fri_fold: need next_bits <= pol_bits <= 32 (got %llu, %llu)", (unsigned long long)next_bits, (unsigned long long)pol_bits);
        return -1;
    }
    ZK_ENTER;
    cudaStream_t st = C.cur;
    const u64 pol2n = (u64)1 << next_bits, nx = (u64)1 << (pol_bits - next_bits), words_in = ((u64)3 << pol_bits), words_out = pol2n * 3;
    const bool dev = is_device_ptr(fri_pol);
    if (dev != is_device_ptr(pol2_e))
    {
        set_error("fri_fold: pol2_e and fri_pol must both be host or both be device pointers");
        return -1;
    }
    Staged in, out;
    const u64 *d_in = fri_pol;
    u64 *d_out = pol2_e;
    // every group's INTT at once: ppar[i] = friPol[i*pol2N + g] is row i, columns 3g..3g+2 of an nX x 3 pol2N matrix.
    // The coefficients live in the library's scratch area (followed by the transform's own scratch for nX > 256).
    const unsigned L = ilog2(nx);
    ZK_TRY(ctx_ensure_scratch(C, words_in * (L > 8 ? 2 : 1)));
    u64 *coef = C.scratch;
    if (!dev)
    {
        ZK_TRY(in.alloc(words_in));
        ZK_TRY(out.alloc(words_out));
        ZK_CUDA(cudaMemcpyAsync(in.d, fri_pol, words_in * sizeof(u64), cudaMemcpyHostToDevice, st));
        d_in = in.d;
        d_out = out.d;
    }
    Scale pre, post;
    if (nx > 1)
    {
        post.mode = SCALE_CONST;
        post.konst = gl::pow_slow(nx % GL_P, GL_P - 2);
    }
    ZK_TRY(ntt_transform(coef, words_out, d_in, words_out, C.scratch + words_in, L, words_out, true, pre, post, st));
    const u64 wi = gl::pow_slow(gl::root_of_unity((unsigned)pol_bits), GL_P - 2); // Goldilocks::inv(Goldilocks::w(polBits)), friProve.cpp:36
    fri_fold_kernel<<<nblocks(pol2n, 256), 256, 0, st>>>(d_out, coef, nx, pol2n, shift_inv % GL_P, wi, special_x[0] % GL_P,
                                                         special_x[1] % GL_P, special_x[2] % GL_P);
    ZK_LAUNCHED();
    ZK_CUDA(cudaGetLastError());
    if (!dev)
    {
        ZK_CUDA(cudaMemcpyAsync(pol2_e, d_out, words_out * sizeof(u64), cudaMemcpyDeviceToHost, st));
        ZK_CUDA(cudaStreamSynchronize(st)); // the staging buffers are freed on return
    }
    return 0; // device pointers: enqueued, like the other entry points
}

int zkgpu_fri_transpose(uint64_t *aux, const uint64_t *pol, uint64_t degree, uint64_t transpose_bits)
{
    if (degree == 0) return 0;
    if (!aux || !pol || aux == pol)
    {
        set_error("fri_transpose: NULL or aliased arguments");
        return -1;
    }
    const u64 w = (u64)1 << transpose_bits;
    if (transpose_bits > 32 || degree % w)
    {
        set_error("fri_transpose: degree %llu is not a multiple of 2^%llu", (unsigned long long)degree, (unsigned long long)transpose_bits);
        return -1;
    }
    ZK_ENTER;
    cudaStream_t st = C.cur;
    const u64 h = degree / w, words = degree * 3;
    const bool dev = is_device_ptr(pol);
    if (dev != is_device_ptr(aux))
    {
        set_error("fri_transpose: aux and pol must both be host or both be device pointers");
        return -1;
    }
    if (dev)
    {
        fri_transpose_kernel<<<nblocks(words, 256), 256, 0, st>>>(aux, pol, w, h);
        ZK_LAUNCHED();
        ZK_CUDA(cudaGetLastError());
        return 0;
    }
    Staged in, out;
    ZK_TRY(in.alloc(words));
    ZK_TRY(out.alloc(words));
    ZK_CUDA(cudaMemcpyAsync(in.d, pol, words * sizeof(u64), cudaMemcpyHostToDevice, st));
    fri_transpose_kernel<<<nblocks(words, 256), 256, 0, st>>>(out.d, in.d, w, h);
    ZK_LAUNCHED();
    ZK_CUDA(cudaGetLastError());
    ZK_CUDA(cudaMemcpyAsync(aux, out.d, words * sizeof(u64), cudaMemcpyDeviceToHost, st));
    ZK_CUDA(cudaStreamSynchronize(st));
    return 0;
}

} // extern "C"
