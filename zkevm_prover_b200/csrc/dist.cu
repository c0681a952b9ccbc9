// dist.cu — the commit stage sharded over the GPUs of one box, behind the C ABI (include/zkgpu.h: zkgpu_init_multi,
// zkgpu_dist_init, zkgpu_dist_commit; zkgpu_extend_and_merkelize dispatches here after zkgpu_init_multi).
//
// Reference call pair: ntt.extendPol(...) + treesGL[k]->merkelize() in Starks::genProof (src/starkpil/starks.cpp:51-58),
// one process, one proof in flight (src/prover/prover.cpp:182-221).  The split follows the data dependences of the
// reference code: the LDE is independent per column (ntt_goldilocks.cpp:103-110) -> column pieces; the leaf sponge
// needs whole rows in column order (poseidon_goldilocks.cpp:277-281) -> row shards; node levels are independent per
// subtree below the top log2(G) levels (:283-300).
//
//   rank g, round k:   [host->device of piece k+1]  ||  LDE of piece (k, g): the LAST PASS of every coset transform stores
//                      each extended row straight into the row shard of the rank that owns it — its own memory or a
//                      peer's over NVLink (ntt_tile.cuh: scatter output) — so the column-shard -> row-shard exchange is
//                      done by the transform's own stores, tile by tile, under the arithmetic of the next tile
//                      ||  sponge absorption of the columns of round k-1 on this rank's rows
//   then: subtree levels per rank, sub-roots all-gathered, top log2(G) levels on rank 0, levels stitched into the
//   reference's level-major node array.
//
// NCCL (dlopen'ed, no link-time dependency) provides the rendezvous, the cross-GPU stream barriers (1-word all-reduce),
// the exchange of CUDA IPC handles in the process-per-GPU form, the sub-root all-gather and the level stitch.  When
// peer access is unavailable (or ZKGPU_EXCHANGE=nccl) the exchange falls back to grouped ncclSend/ncclRecv per round.
#include "gl64.cuh"
#include "zk_internal.h"
#include "../../include/zkgpu.h"
#include <nccl.h>
#include <dlfcn.h>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>

namespace zk {

// ---------------------------------------------------------------------------------------------------
// NCCL through dlopen: the process may already hold a libnccl.so.2 (torch bundles one); share it
// ---------------------------------------------------------------------------------------------------
struct NcclApi
{
    void *h = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
};
static NcclApi g_nccl;
static std::mutex g_nccl_mu;

static int nccl_load()
{
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    if (g_nccl.h) return 0;
    // Order: an explicit path; the libnccl.so.2 the process already holds (a Python host that imported torch has torch's
    // bundled NCCL mapped: share it rather than map a second, possibly older, copy under the same soname); the system's.
    void *h = nullptr;
    if (const char *p = getenv("ZKGPU_NCCL_LIB")) h = dlopen(p, RTLD_NOW | RTLD_LOCAL);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL | RTLD_NOLOAD);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
    if (!h)
    {
        set_error("zkgpu: cannot load NCCL (libnccl.so.2): %s", dlerror());
        return -5;
    }
#define ZK_SYM(name)                                                        \
    *(void **)(&g_nccl.name) = dlsym(h, "nccl" #name);                      \
    if (!g_nccl.name)                                                       \
    {                                                                       \
        set_error("zkgpu: NCCL symbol nccl" #name " not found");            \
        return -5;                                                          \
    }
    ZK_SYM(GetUniqueId) ZK_SYM(CommInitRank) ZK_SYM(CommInitAll) ZK_SYM(CommDestroy) ZK_SYM(AllReduce) ZK_SYM(AllGather)
    ZK_SYM(Send) ZK_SYM(Recv) ZK_SYM(GroupStart) ZK_SYM(GroupEnd) ZK_SYM(GetErrorString) ZK_SYM(GetVersion)
#undef ZK_SYM
    g_nccl.h = h;
    return 0;
}
#define ZK_NCCL(call)                                                                                       \
    do {                                                                                                    \
        ncclResult_t r__ = (call);                                                                          \
        if (r__ != ncclSuccess) {                                                                           \
            zk::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(r__));         \
            return -5;                                                                                      \
        }                                                                                                   \
    } while (0)

// ---------------------------------------------------------------------------------------------------
// state
// ---------------------------------------------------------------------------------------------------
static const int MAX_WORLD = 16, MAX_ROUNDS = 16;

// How the extended rows travel from the rank that computed them (column pieces) to the rank that hashes them (row shards):
//   EX_STORES  the last pass of every coset transform stores each row straight into its owner's shard (peer memory over
//              NVLink): one kernel does the transform and the exchange, one round, no staging buffer;
//   EX_DMA     the transform writes a local staging buffer and the copy engines push the row blocks to the peers
//              (cudaMemcpy2DAsync device-to-device over NVLink) while the SMs extend the next round / hash the previous one;
//   EX_NCCL    grouped ncclSend/ncclRecv per round into a receive buffer that is then unpacked (no peer access needed).
// ZKGPU_EXCHANGE=stores|dma|nccl picks one; default dma when every pair of GPUs has peer access, else nccl.
// Measured at H (profiles/r2_bench_n{2,8}_exchange_*.json): 2 GPUs stores 703.8 / dma 703.0 / nccl 718.5 ms per commit;
// 8 GPUs stores 190.3 (the two scatter passes of every LDE run at the remote-store rate, 350 GB/s per GPU: LDE 64.0 ms
// instead of 46.6) / dma 174.6 (LDE 48.7, nothing exposed).
enum { EX_STORES = 0, EX_DMA = 1, EX_NCCL = 2 };
static int exchange_wanted()
{
    static const int w = []() {
        const char *e = getenv("ZKGPU_EXCHANGE");
        if (e && !strcmp(e, "nccl")) return (int)EX_NCCL;
        if (e && !strcmp(e, "stores")) return (int)EX_STORES;
        return (int)EX_DMA;
    }();
    return w;
}
static int g_exchange = -1; // what the job actually uses (set when peer access is known); before that: what is wanted
static int exchange_now() { return g_exchange >= 0 ? g_exchange : exchange_wanted(); }

struct Dist
{
    int rank = 0, world = 1;
    bool single_process = false;
    bool p2p = false, p2p_tried = false;
    ncclComm_t comm = nullptr;
    cudaStream_t cs = nullptr; // every NCCL operation of this rank is issued here
    cudaStream_t xs = nullptr; // EX_DMA: the peer copies
    cudaStream_t sa = nullptr; // the sponge: its own stream, so that the absorption of round k and the transform of round k+1
                               // start in whichever order their inputs arrive (exchange of k / host copy of k+1)
    cudaEvent_t ev_cp = nullptr, ev_sa = nullptr;
    cudaEvent_t ev_lde[MAX_ROUNDS] = {}, ev_bar[MAX_ROUNDS] = {}, ev_unp[MAX_ROUNDS] = {}, ev_a = nullptr, ev_b = nullptr;
    u64 *flag = nullptr;                                   // barrier word
    u64 *rows = nullptr;                                   // this rank's row shard of the extended trace, R x ncols
    size_t rows_cap = 0;                                   // words allocated
    u64 rows_n = 0, rows_c = 0;                            // shape of the last commit
    u64 *peer_rows[MAX_WORLD] = {};                        // every rank's shard as seen from this device
    void *ipc_open[MAX_WORLD] = {};                        // IPC mappings to close on reallocation
    u64 *peer_tab = nullptr;                               // device copy of peer_rows
    u64 *sub = nullptr;                                    // subtree over this rank's R leaves
    size_t sub_cap = 0;
    u64 *top = nullptr;                                    // (2*world - 1) * 4 words: levels above the sub-roots
    u64 *tree_dev = nullptr;                               // rank 0, host tree pointer, process-per-GPU: staging
    size_t tree_cap = 0;
    u64 *in_buf = nullptr;                                 // host-input staging, two slots
    size_t in_cap = 0;
    u64 *ext_buf = nullptr, *recv_buf = nullptr;           // fallback exchange (no peer access)
    size_t ext_cap = 0, recv_cap = 0;
    u64 *ipc_scratch = nullptr;                            // world * 64 bytes for the handle all-gather
};

// the ranks of a single-process job (zkgpu_init_multi): worker threads meet here
struct Group
{
    int n = 0;
    Ctx *ctx[MAX_WORLD] = {};
    std::mutex mu;
    std::condition_variable cv;
    int arrived = 0;
    unsigned long long phase = 0;
    u64 *shared_rows[MAX_WORLD] = {};
    int shared_rc[MAX_WORLD] = {};
    void barrier()
    {
        std::unique_lock<std::mutex> lk(mu);
        const unsigned long long ph = phase;
        if (++arrived == n)
        {
            arrived = 0;
            phase++;
            cv.notify_all();
        }
        else
            cv.wait(lk, [&] { return phase != ph; });
    }
    // every rank hands in its return code; all leave with the first failure (by rank) or 0.  Two barriers: the slots may be
    // rewritten by the next agreement only after everyone has read them.
    int agree(int rank, int rc, int *who = nullptr)
    {
        agree_rc[rank] = rc;
        barrier();
        int out = 0;
        for (int g = 0; g < n && out == 0; g++)
            if (agree_rc[g] < 0)
            {
                out = agree_rc[g];
                if (who) *who = g;
            }
        barrier();
        return out;
    }
    int agree_rc[MAX_WORLD] = {};
};
static Group g_group;
static std::mutex g_multi_mu; // one single-process multi-GPU commit at a time (the ranks' barriers and NCCL calls must not interleave)

static int dev_realloc(u64 **p, size_t *cap, size_t words)
{
    if (words <= *cap && *p) return 0;
    if (*p) ZK_CUDA(cudaFree(*p));
    *p = nullptr;
    *cap = 0;
    cudaError_t e = cudaMalloc(p, (words ? words : 1) * sizeof(u64));
    if (e != cudaSuccess)
    {
        cudaGetLastError();
        set_error("zkgpu dist: cannot allocate %llu MiB on the device (%s)", (unsigned long long)(words >> 17), cudaGetErrorString(e));
        return -4;
    }
    *cap = words;
    return 0;
}

static int dist_create(Ctx &C, int rank, int world, ncclComm_t comm, bool single)
{
    Dist *D = new Dist();
    D->rank = rank;
    D->world = world;
    D->comm = comm;
    D->single_process = single;
    // The exchange streams get the highest priority: a strided device-to-device copy and NCCL's barrier are kernels too, and
    // at equal priority the block scheduler would only start them once the sponge kernel queued before them has dispatched
    // its last CTA (measured at 8 GPUs, profiles/r2_dist_timeline_n8_dma.txt: every round's barrier arrived ~10 ms AFTER the
    // previous round's sponge had finished instead of during it).
    int prio_lo = 0, prio_hi = 0;
    ZK_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    ZK_CUDA(cudaStreamCreateWithPriority(&D->cs, cudaStreamNonBlocking, prio_hi));
    ZK_CUDA(cudaStreamCreateWithPriority(&D->xs, cudaStreamNonBlocking, prio_hi));
    ZK_CUDA(cudaEventCreateWithFlags(&D->ev_cp, cudaEventDisableTiming));
    ZK_CUDA(cudaEventCreateWithFlags(&D->ev_sa, cudaEventDisableTiming));
    ZK_CUDA(cudaStreamCreateWithFlags(&D->sa, cudaStreamNonBlocking));
    for (int i = 0; i < MAX_ROUNDS; i++)
    {
        ZK_CUDA(cudaEventCreateWithFlags(&D->ev_lde[i], cudaEventDisableTiming));
        ZK_CUDA(cudaEventCreateWithFlags(&D->ev_bar[i], cudaEventDisableTiming));
        ZK_CUDA(cudaEventCreateWithFlags(&D->ev_unp[i], cudaEventDisableTiming));
    }
    ZK_CUDA(cudaEventCreateWithFlags(&D->ev_a, cudaEventDisableTiming));
    ZK_CUDA(cudaEventCreateWithFlags(&D->ev_b, cudaEventDisableTiming));
    ZK_CUDA(cudaMalloc(&D->flag, 64));
    ZK_CUDA(cudaMemset(D->flag, 0, 64));
    ZK_CUDA(cudaMalloc(&D->peer_tab, MAX_WORLD * sizeof(u64)));
    ZK_CUDA(cudaMalloc(&D->top, (2 * (size_t)world - 1) * 4 * sizeof(u64)));
    ZK_CUDA(cudaMalloc(&D->ipc_scratch, (size_t)world * 64));
    C.dist = D;
    return 0;
}

// all ranks have finished everything enqueued on `after` before anything enqueued later on `then` starts
static int dist_barrier(Dist &D, cudaStream_t after, cudaStream_t then, cudaEvent_t e_in, cudaEvent_t e_out)
{
    ZK_CUDA(cudaEventRecord(e_in, after));
    ZK_CUDA(cudaStreamWaitEvent(D.cs, e_in, 0));
    ZK_NCCL(g_nccl.AllReduce(D.flag, D.flag + 1, 1, ncclUint64, ncclSum, D.comm, D.cs));
    ZK_CUDA(cudaEventRecord(e_out, D.cs));
    if (then) ZK_CUDA(cudaStreamWaitEvent(then, e_out, 0));
    return 0;
}

// ---------------------------------------------------------------------------------------------------
// column ownership: the row is cut into rounds x world consecutive pieces, piece (k, g) belonging to rank g in round k.
// Every round but the last covers a whole number of 8-column sponge blocks (so absorbing round by round chunks the row
// exactly as linear_hash does, poseidon_goldilocks.cpp:240-256); inside a round the columns are dealt out evenly.
// Device-resident input: one round (stores) or two (staged exchange).  Host input: uniform small rounds (below).
// ---------------------------------------------------------------------------------------------------
struct Plan
{
    int rounds = 0, world = 1;
    u64 c0[MAX_ROUNDS][MAX_WORLD], w[MAX_ROUNDS][MAX_WORLD];
    u64 round_c0[MAX_ROUNDS], round_w[MAX_ROUNDS];
};

static int make_plan(Plan &pl, u64 ncols, int world, bool host_input)
{
    pl.world = world;
    const u64 blocks = (ncols + 7) / 8;
    u64 per_round[MAX_ROUNDS];
    int nr = 0;
    if (blocks == 0 || (!host_input && exchange_now() == EX_STORES))
        per_round[nr++] = blocks;
    else if (!host_input)
    {
        // staged exchange (copy engines or NCCL): two rounds so that the transfer of one hides under the next one's transform
        // and the previous one's sponge
        static const u64 want = []() { const char *e = getenv("ZKGPU_DMA_ROUNDS"); return (u64)(e && atoi(e) > 0 && atoi(e) <= MAX_ROUNDS ? atoi(e) : 2); }();
        const u64 k = blocks >= want * (u64)world ? want : 1; // (8 GPUs, H: 2 rounds 184.7 ms, 3 rounds 185.9, 5 rounds 193.3)
        for (u64 i = 0; i < k; i++) per_round[nr++] = blocks / k + (i < blocks % k ? 1 : 0);
    }
    else
    {
        // Host input: uniform small rounds, two sponge blocks (16 columns) per rank and round, more when that would need more
        // than 14 rounds.  With 8 GPUs the host link is the bottleneck (20-24 GB/s per GPU): everything still to be computed
        // after the LAST copy lands is exposed, and a round's transform + sponge (32 ms) hides under the next round's copy
        // (49 ms).  Measured at 8 GPUs, H: ramp 16/32/35 columns per rank 390 ms, ramp up + down 372 ms
        // (profiles/r2_dist_timeline_n8_e2e_rampdown.txt: 100 ms of sponge left after the last copy), uniform rounds: see DESIGN.md.
        u64 per = 2 * (u64)world;
        if ((blocks + per - 1) / per > 14) per = (blocks + 13) / 14;
        u64 k = blocks / per; // floor: no round narrower than `per` blocks (pieces under 16 columns take the slow narrow kernel)
        if (k < 1) k = 1;
        for (u64 i = 0; i < k; i++) per_round[nr++] = blocks / k + (i < blocks % k ? 1 : 0);
    }
    pl.rounds = nr;
    u64 c = 0;
    int rot = 0;
    for (int k = 0; k < nr; k++)
    {
        u64 cols = per_round[k] * 8;
        if (cols > ncols - c) cols = ncols - c;
        pl.round_c0[k] = c;
        pl.round_w[k] = cols;
        const u64 base = cols / world, extra = cols % world;
        for (int g = 0; g < world; g++)
        {
            // the odd columns of a round go to a different group of ranks each round, so that column counts stay level
            const u64 wg = base + ((u64)((g - rot + world) % world) < extra ? 1 : 0);
            pl.c0[k][g] = c;
            pl.w[k][g] = wg;
            c += wg;
        }
        rot = (int)((rot + extra) % world);
    }
    return 0;
}

static int subtree_levels(u64 leaves, u64 *off, u64 *cnt)
{
    int l = 0;
    u64 o = 0, n = leaves;
    for (;;)
    {
        off[l] = o;
        cnt[l] = n;
        l++;
        if (n == 1) break;
        o += n * 4;
        n >>= 1;
    }
    return l;
}

// (re)allocate this rank's row shard and learn where every peer's shard is.  Collective: every rank calls it with the
// same shape.  grp != nullptr: ranks are threads of this process.
static int dist_setup_rows(Ctx &C, Dist &D, u64 need_words)
{
    const bool force_nccl = exchange_wanted() == EX_NCCL;
    if (D.rows && need_words <= D.rows_cap && D.p2p_tried) return 0;
    ZK_CUDA(cudaStreamSynchronize(C.cur));
    ZK_CUDA(cudaStreamSynchronize(D.cs));
    if (!D.single_process)
        for (int g = 0; g < D.world; g++)
            if (D.ipc_open[g])
            {
                cudaIpcCloseMemHandle(D.ipc_open[g]);
                D.ipc_open[g] = nullptr;
            }
    // the old shard is freed below: only after EVERY rank has drained its streams (nobody is still copying into a peer) and
    // closed its mapping of it.  Every rank takes this path together (same shapes on every rank, call after call).
    if (D.rows)
    {
        if (D.single_process)
            g_group.barrier();
        else
        {
            ZK_NCCL(g_nccl.AllReduce(D.flag, D.flag + 1, 1, ncclUint64, ncclSum, D.comm, D.cs));
            ZK_CUDA(cudaStreamSynchronize(D.cs));
        }
    }
    const int alloc_rc = dev_realloc(&D.rows, &D.rows_cap, need_words);
    if (D.single_process)
    {
        // the other worker threads are on their way to the barriers below: a rank that ran out of memory must meet them there
        // and everybody leaves with the error
        int who = -1;
        const int rc = g_group.agree(D.rank, alloc_rc, &who);
        if (rc < 0)
        {
            if (who != D.rank) set_error("zkgpu dist: rank %d could not allocate its row shard", who);
            return rc;
        }
    }
    else
        ZK_TRY(alloc_rc);
    for (int g = 0; g < D.world; g++) D.peer_rows[g] = nullptr;
    D.peer_rows[D.rank] = D.rows;
    bool ok = !force_nccl;
    if (D.single_process)
    {
        g_group.shared_rows[D.rank] = D.rows;
        g_group.barrier();
        for (int g = 0; g < D.world; g++) D.peer_rows[g] = g_group.shared_rows[g];
        if (ok)
            for (int g = 0; g < D.world && ok; g++)
                if (g != D.rank)
                {
                    int can = 0;
                    if (cudaDeviceCanAccessPeer(&can, C.device, g_group.ctx[g]->device) != cudaSuccess || !can) ok = false;
                }
        g_group.shared_rc[D.rank] = ok ? 1 : 0;
        g_group.barrier();
        for (int g = 0; g < D.world; g++) ok = ok && g_group.shared_rc[g];
        g_group.barrier(); // shared_* may be rewritten by the next call only after everyone has read them
    }
    else
    {
        cudaIpcMemHandle_t mine, all[MAX_WORLD];
        static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
        memset(&mine, 0, sizeof(mine));
        if (ok && cudaIpcGetMemHandle(&mine, D.rows) != cudaSuccess)
        {
            cudaGetLastError();
            ok = false;
        }
        ZK_CUDA(cudaMemcpyAsync((char *)D.ipc_scratch + 64 * D.rank, &mine, 64, cudaMemcpyHostToDevice, D.cs));
        ZK_NCCL(g_nccl.AllGather((char *)D.ipc_scratch + 64 * D.rank, D.ipc_scratch, 64, ncclChar, D.comm, D.cs));
        ZK_CUDA(cudaMemcpyAsync(all, D.ipc_scratch, 64 * (size_t)D.world, cudaMemcpyDeviceToHost, D.cs));
        ZK_CUDA(cudaStreamSynchronize(D.cs));
        for (int g = 0; g < D.world && ok; g++)
            if (g != D.rank)
            {
                void *p = nullptr;
                if (cudaIpcOpenMemHandle(&p, all[g], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess)
                {
                    cudaGetLastError();
                    ok = false;
                    break;
                }
                D.ipc_open[g] = p;
                D.peer_rows[g] = (u64 *)p;
            }
        // everyone must take the same path: sum of the failure flags
        const u64 bad = ok ? 0 : 1;
        ZK_CUDA(cudaMemcpyAsync(D.flag + 2, &bad, 8, cudaMemcpyHostToDevice, D.cs));
        ZK_NCCL(g_nccl.AllReduce(D.flag + 2, D.flag + 3, 1, ncclUint64, ncclSum, D.comm, D.cs));
        u64 total_bad = 0;
        ZK_CUDA(cudaMemcpyAsync(&total_bad, D.flag + 3, 8, cudaMemcpyDeviceToHost, D.cs));
        ZK_CUDA(cudaStreamSynchronize(D.cs));
        ok = total_bad == 0;
    }
    D.p2p = ok;
    D.p2p_tried = true;
    g_exchange = ok ? exchange_wanted() : (int)EX_NCCL; // the same on every rank: `ok` was agreed on above
    u64 tab[MAX_WORLD] = {};
    for (int g = 0; g < D.world; g++) tab[g] = (u64)(uintptr_t)D.peer_rows[g];
    ZK_CUDA(cudaMemcpy(D.peer_tab, tab, sizeof(u64) * D.world, cudaMemcpyHostToDevice));
    if (getenv("ZKGPU_DIST_TRACE")) fprintf(stderr, "[zkgpu dist] rank %d/%d: row shard %llu MiB, exchange = %s\n", D.rank, D.world,
                                            (unsigned long long)(need_words >> 17),
                                            g_exchange == EX_STORES ? "peer stores (fused into the LDE)" : g_exchange == EX_DMA ? "copy engines (peer cudaMemcpy2DAsync)" : "ncclSend/ncclRecv");
    return 0;
}

struct PieceSrc
{
    const u64 *p;
    u64 ld;
};

__global__ void dist_unpack_kernel(u64 *dst, u64 dst_ld, const u64 *src, u64 nrows, u64 w)
{
    const u64 total = nrows * w;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (u64)gridDim.x * blockDim.x)
    {
        const u64 r = i / w, c = i - r * w;
        dst[r * dst_ld + c] = src[i];
    }
}

// ---------------------------------------------------------------------------------------------------
// the program of one rank.  Runs with the rank's context locked and its device current.
//   pieces[k]   : this rank's N x w(k, rank) piece of round k (device memory, or host memory which is staged)
//   tree        : destination of the (2*NExt-1)*4 node words.  shared_host_tree: host memory every rank of this process
//                 writes its own level slices into (single-process form).  gather (same value on every rank): the level
//                 slices are shipped to rank 0, whose `tree` (device or host) receives the whole array; the other ranks
//                 pass nullptr.  Neither: the tree stays sharded (subtrees per rank, top levels on rank 0).
//   ext_host    : single-process form only: the caller's NExt x ncols host buffer; every rank copies out its rows
// ---------------------------------------------------------------------------------------------------
static int dist_commit_rank(Ctx &C, const PieceSrc *pieces, bool host_input, u64 n_ext, u64 n, u64 ncols, u64 *tree,
                            bool shared_host_tree, bool gather, u64 *ext_host)
{
    Dist &D = *C.dist;
    const int G = D.world, me = D.rank;
    unsigned log_n = 0, log_e = 0, lgR = 0;
    while (((u64)1 << log_n) < n) log_n++;
    while (((u64)1 << log_e) < n_ext) log_e++;
    const unsigned log_b = log_e - log_n;
    const u64 R = n_ext / (u64)G;
    while (((u64)1 << lgR) < R) lgR++;
    if (R == 0 || R * (u64)G != n_ext)
    {
        set_error("zkgpu dist: N_Extended = %llu must be a multiple of the number of GPUs (%d)", (unsigned long long)n_ext, G);
        return -1;
    }
    cudaStream_t st = C.cur;
    // Everything that can fail for one rank alone (stream creation, allocations) comes first, and in the single-process form
    // the worker threads then agree on the outcome: either all go on to the collectives below or all return the error.  A
    // rank that left on its own would leave the others in an NCCL wait that never ends.
    const int pipe_rc = ctx_pipe_streams(C);
    if (D.single_process)
    {
        int who = -1;
        const int rc = g_group.agree(me, pipe_rc, &who);
        if (rc < 0)
        {
            if (who != me) set_error("zkgpu dist: rank %d could not create its copy streams", who);
            return rc;
        }
    }
    else
        ZK_TRY(pipe_rc);
    ZK_TRY(dist_setup_rows(C, D, R * ncols)); // also settles the exchange mode, which the plan depends on (agrees among the ranks itself)
    Plan pl;
    ZK_TRY(make_plan(pl, ncols, G, host_input)); // same arguments on every rank: fails for all or for none
    D.rows_n = R;
    D.rows_c = ncols;
    const u64 sub_words = (2 * R - 1) * 4;
    u64 wmax = 0, rwmax = 0;
    for (int k = 0; k < pl.rounds; k++)
    {
        if (pl.w[k][me] > wmax) wmax = pl.w[k][me];
        if (pl.round_w[k] > rwmax) rwmax = pl.round_w[k];
    }
    const int ex = exchange_now();
    const bool tree_is_dev = tree != nullptr && is_device_ptr(tree);
    u64 *tree_d = tree_is_dev ? tree : nullptr; // rank 0's device image of the whole tree, when there is one
    auto allocate = [&]() -> int {
        ZK_TRY(dev_realloc(&D.sub, &D.sub_cap, sub_words));
        ZK_TRY(ctx_ensure_scratch(C, 2 * n * (wmax ? wmax : 1)));
        if (host_input) ZK_TRY(dev_realloc(&D.in_buf, &D.in_cap, 2 * n * wmax));
        if (ex != EX_STORES) ZK_TRY(dev_realloc(&D.ext_buf, &D.ext_cap, 2 * n_ext * wmax));
        if (ex == EX_NCCL) ZK_TRY(dev_realloc(&D.recv_buf, &D.recv_cap, 2 * R * rwmax));
        if (me == 0 && gather && !tree_is_dev)
        {
            ZK_TRY(dev_realloc(&D.tree_dev, &D.tree_cap, (2 * n_ext - 1) * 4));
            tree_d = D.tree_dev;
        }
        return 0;
    };
    int alloc_rc = allocate();
    if (D.single_process && alloc_rc == 0 && getenv("ZKGPU_TEST_FAIL_RANK") && atoi(getenv("ZKGPU_TEST_FAIL_RANK")) == me)
    {
        // test hook (tests/gpu_dist_worker.py): behave as if this rank had run out of device memory
        set_error("zkgpu dist: cannot allocate (ZKGPU_TEST_FAIL_RANK=%d)", me);
        alloc_rc = -4;
    }
    if (D.single_process)
    {
        int who = -1;
        const int rc = g_group.agree(me, alloc_rc, &who);
        if (rc < 0)
        {
            if (who != me) set_error("zkgpu dist: rank %d could not allocate its buffers", who);
            return rc;
        }
    }
    else
        ZK_TRY(alloc_rc);

    ctx_pipe_event(C, nullptr, st);
    C.segs.clear();
    // ZKGPU_DIST_TRACE=1: timeline of this rank's copies, transforms, barriers and sponge strips (development aid)
    static const bool trace = getenv("ZKGPU_DIST_TRACE") != nullptr;
    struct Mark { const char *what; int k, e; };
    std::vector<Mark> marks;
    auto mark = [&](const char *what, int k, cudaStream_t s) {
        if (!trace) return;
        int e = -1;
        if (ctx_pipe_event(C, &e, s) >= 0) marks.push_back({what, k, e});
    };
    ZK_CUDA(cudaEventRecord(C.ev[4], st));
    mark("call", 0, st);
    // nobody may still be reading its rows (previous commit) when a peer starts storing into them
    ZK_TRY(dist_barrier(D, st, st, D.ev_a, D.ev_b));
    // Host input: the sponge runs on its own stream, so that the absorption of round k and the transform of round k+1 start in
    // whichever order their inputs arrive (the exchange of k / the host copy of k+1): with the host link as the bottleneck the
    // sponge no longer queues behind a transform that is itself waiting for its copy (2 GPUs e2e 801 -> 759 ms).  Device input:
    // one stream — run side by side the two kernels only slow each other down (2 GPUs 691.8 -> 707.8 ms).
    cudaStream_t sa = host_input ? D.sa : st;
    ZK_CUDA(cudaStreamWaitEvent(sa, D.ev_b, 0)); // the sponge stream starts after everything the caller queued before this call

    auto issue_h2d = [&](int k) -> int {
        const int b = k & 1;
        const u64 w = pl.w[k][me];
        if (k >= 2) ZK_CUDA(cudaStreamWaitEvent(C.h2d, C.in_free[b], 0));
        mark("h2d+", k, C.h2d);
        if (w && pieces[k].ld == w)
            ZK_CUDA(cudaMemcpyAsync(D.in_buf + (size_t)b * n * wmax, pieces[k].p, n * w * sizeof(u64), cudaMemcpyHostToDevice, C.h2d));
        else if (w)
            ZK_CUDA(cudaMemcpy2DAsync(D.in_buf + (size_t)b * n * wmax, w * sizeof(u64), pieces[k].p, pieces[k].ld * sizeof(u64), w * sizeof(u64), n,
                                      cudaMemcpyHostToDevice, C.h2d));
        mark("h2d-", k, C.h2d);
        ZK_CUDA(cudaEventRecord(C.in_ready[b], C.h2d));
        return 0;
    };
    auto absorb_round = [&](int k) -> int {
        ZK_CUDA(cudaStreamWaitEvent(sa, D.ev_bar[k], 0));
        if (ex == EX_NCCL)
        {
            const u64 *rb = D.recv_buf + (size_t)(k & 1) * R * rwmax;
            for (int g = 0; g < G; g++)
            {
                const u64 w = pl.w[k][g];
                if (w)
                {
                    u64 blocks = (R * w + 255) / 256;
                    if (blocks > (u64)C.num_sms * 32) blocks = (u64)C.num_sms * 32;
                    dist_unpack_kernel<<<(unsigned)blocks, 256, 0, sa>>>(D.rows + pl.c0[k][g], ncols, rb, R, w);
                    ZK_LAUNCHED();
                    ZK_CUDA(cudaGetLastError());
                }
                rb += R * w;
            }
            ZK_CUDA(cudaEventRecord(D.ev_unp[k], sa));
        }
        if (ext_host && pl.round_w[k])
        {
            // this round's columns of my rows are complete: send them home while the sponge runs
            ZK_CUDA(cudaEventRecord(C.out_ready[0], sa));
            ZK_CUDA(cudaStreamWaitEvent(C.d2h, C.out_ready[0], 0));
            ZK_CUDA(cudaMemcpy2DAsync(ext_host + (size_t)me * R * ncols + pl.round_c0[k], ncols * sizeof(u64), D.rows + pl.round_c0[k],
                                      ncols * sizeof(u64), pl.round_w[k] * sizeof(u64), R, cudaMemcpyDeviceToHost, C.d2h));
        }
        if (ncols > 4 && pl.round_w[k])
        {
            int e0, e1;
            ZK_TRY(ctx_pipe_event(C, &e0, sa));
            ZK_TRY(merkle_absorb(D.sub, D.rows + pl.round_c0[k], pl.round_w[k], ncols, R, k == 0, sa));
            ZK_TRY(ctx_pipe_event(C, &e1, sa));
            C.segs.push_back({1, e0, e1});
            mark("abs-", k, sa);
        }
        return 0;
    };

    if (host_input)
    {
        ZK_CUDA(cudaEventRecord(C.in_free[0], st));
        ZK_CUDA(cudaStreamWaitEvent(C.h2d, C.in_free[0], 0));
        ZK_TRY(issue_h2d(0));
    }
    for (int k = 0; k < pl.rounds; k++)
    {
        const int b = k & 1;
        const u64 w = pl.w[k][me];
        const u64 *src = pieces[k].p;
        u64 src_ld = pieces[k].ld;
        if (host_input)
        {
            if (k + 1 < pl.rounds) ZK_TRY(issue_h2d(k + 1));
            ZK_CUDA(cudaStreamWaitEvent(st, C.in_ready[b], 0));
            src = D.in_buf + (size_t)b * n * wmax;
            src_ld = w;
        }
        if (k >= 2 && ex != EX_STORES) ZK_CUDA(cudaStreamWaitEvent(st, D.ev_bar[k - 2], 0)); // staging slot b has been shipped
        int e0, e1;
        mark("lde+", k, st);
        ZK_TRY(ctx_pipe_event(C, &e0, st));
        u64 *ext = ex == EX_STORES ? nullptr : D.ext_buf + (size_t)b * n_ext * wmax;
        if (w)
        {
            if (ex == EX_STORES)
            {
                Scatter sc;
                sc.peer_tab = D.peer_tab;
                sc.lg_rows = lgR;
                sc.ld = ncols;
                sc.col0 = pl.c0[k][me];
                ZK_TRY(lde_strip(nullptr, 0, src, src_ld, C.scratch, log_n, log_b, w, st, &sc));
            }
            else
                ZK_TRY(lde_strip(ext, w, src, src_ld, C.scratch, log_n, log_b, w, st));
        }
        ZK_TRY(ctx_pipe_event(C, &e1, st));
        C.segs.push_back({0, e0, e1});
        mark("lde-", k, st);
        if (host_input) ZK_CUDA(cudaEventRecord(C.in_free[b], st));
        // exchange point of round k
        ZK_CUDA(cudaEventRecord(D.ev_lde[k], st));
        if (ex == EX_DMA)
        {
            // the copy engines carry block h (rows of rank h, contiguous in the staging buffer) into rank h's shard, at this
            // piece's columns; the barrier below is entered once MY copies are done and left once EVERYBODY's are
            ZK_CUDA(cudaStreamWaitEvent(D.xs, D.ev_lde[k], 0));
            if (w)
                for (int i = 0; i < G; i++)
                {
                    const int h = (me + i) % G; // start with my own rows, spread the peers
                    ZK_CUDA(cudaMemcpy2DAsync(D.peer_rows[h] + pl.c0[k][me], ncols * sizeof(u64), ext + (size_t)h * R * w, w * sizeof(u64),
                                              w * sizeof(u64), R, cudaMemcpyDeviceToDevice, D.xs));
                }
            ZK_CUDA(cudaEventRecord(D.ev_cp, D.xs));
            ZK_CUDA(cudaStreamWaitEvent(D.cs, D.ev_cp, 0));
        }
        else
            ZK_CUDA(cudaStreamWaitEvent(D.cs, D.ev_lde[k], 0));
        if (ex != EX_NCCL)
            ZK_NCCL(g_nccl.AllReduce(D.flag, D.flag + 1, 1, ncclUint64, ncclSum, D.comm, D.cs)); // all rows of round k have landed
        else
        {
            if (k >= 2) ZK_CUDA(cudaStreamWaitEvent(D.cs, D.ev_unp[k - 2], 0)); // the receive slot has been unpacked
            u64 *rb = D.recv_buf + (size_t)b * R * rwmax;
            ZK_NCCL(g_nccl.GroupStart());
            for (int g = 0; g < G; g++)
            {
                if (w) ZK_NCCL(g_nccl.Send(ext + (size_t)g * R * w, R * w, ncclUint64, g, D.comm, D.cs));
                if (pl.w[k][g]) ZK_NCCL(g_nccl.Recv(rb, R * pl.w[k][g], ncclUint64, g, D.comm, D.cs));
                rb += R * pl.w[k][g];
            }
            ZK_NCCL(g_nccl.GroupEnd());
        }
        ZK_CUDA(cudaEventRecord(D.ev_bar[k], D.cs));
        mark("bar-", k, D.cs);
        if (k > 0) ZK_TRY(absorb_round(k - 1));
    }
    if (pl.rounds > 0) ZK_TRY(absorb_round(pl.rounds - 1));
    if (ncols <= 4)
    {
        // rows of <= 4 words are copied, not hashed (poseidon_goldilocks.cpp:234-239)
        int e0, e1;
        ZK_TRY(ctx_pipe_event(C, &e0, sa));
        ZK_TRY(merkle_leaves(D.sub, D.rows, ncols, ncols, R, sa));
        ZK_TRY(ctx_pipe_event(C, &e1, sa));
        C.segs.push_back({1, e0, e1});
    }
    ZK_CUDA(cudaEventRecord(C.ev[2], sa));
    // ---- stitch: level l of the global tree = the ranks' slices side by side (l <= log2 R), then the top over the sub-roots.
    // The leaf level is half of all the bytes: it travels to rank 0 while the node levels are being hashed.
    u64 soff[40], scnt[40], goff[40], gcnt[40];
    const int sl = subtree_levels(R, soff, scnt);
    subtree_levels(n_ext, goff, gcnt);
    const int LR = sl - 1; // level of the sub-roots
    auto ship_levels = [&](int l0, int l1) -> int {
        // process-per-GPU form, or a device tree on rank 0: every rank ships its level slices to rank 0
        ZK_NCCL(g_nccl.GroupStart());
        for (int l = l0; l < l1; l++)
        {
            if (me != 0)
                ZK_NCCL(g_nccl.Send(D.sub + soff[l], scnt[l] * 4, ncclUint64, 0, D.comm, D.cs));
            else
                for (int g = 1; g < G; g++) ZK_NCCL(g_nccl.Recv(tree_d + goff[l] + (size_t)g * scnt[l] * 4, scnt[l] * 4, ncclUint64, g, D.comm, D.cs));
        }
        ZK_NCCL(g_nccl.GroupEnd());
        return 0;
    };
    if (gather && LR > 0)
    {
        ZK_CUDA(cudaStreamWaitEvent(D.cs, C.ev[2], 0));
        ZK_TRY(ship_levels(0, 1));
    }
    ZK_TRY(merkle_levels(D.sub, R, sa));
    ZK_CUDA(cudaEventRecord(D.ev_a, sa));
    ZK_CUDA(cudaStreamWaitEvent(D.cs, D.ev_a, 0));
    u64 *top_dst = (me == 0 && tree_d) ? tree_d + goff[LR] : D.top;
    ZK_NCCL(g_nccl.AllGather(D.sub + soff[LR], top_dst, 4, ncclUint64, D.comm, D.cs));
    if (gather && LR > 1) ZK_TRY(ship_levels(1, LR));
    ZK_CUDA(cudaEventRecord(D.ev_b, D.cs));
    ZK_CUDA(cudaStreamWaitEvent(st, D.ev_b, 0)); // cs waited for the sponge stream (ev_a): st is now behind both
    ZK_CUDA(cudaStreamWaitEvent(st, D.ev_a, 0));
    if (gather && me == 0)
        for (int l = 0; l < LR; l++)
            ZK_CUDA(cudaMemcpyAsync(tree_d + goff[l], D.sub + soff[l], scnt[l] * 32, cudaMemcpyDeviceToDevice, st));
    if (me == 0) ZK_TRY(merkle_levels(top_dst, (u64)G, st));
    ZK_CUDA(cudaEventRecord(C.ev[3], st));
    C.timed = true;
    if (shared_host_tree && tree)
    {
        for (int l = 0; l < LR; l++)
            ZK_CUDA(cudaMemcpyAsync(tree + goff[l] + (size_t)me * scnt[l] * 4, D.sub + soff[l], scnt[l] * 32, cudaMemcpyDeviceToHost, st));
        if (me == 0) ZK_CUDA(cudaMemcpyAsync(tree + goff[LR], top_dst, (2 * (size_t)G - 1) * 32, cudaMemcpyDeviceToHost, st));
    }
    else if (me == 0 && tree && !tree_is_dev)
        ZK_CUDA(cudaMemcpyAsync(tree, tree_d, (2 * n_ext - 1) * 32, cudaMemcpyDeviceToHost, st));
    if (ext_host)
    {
        ZK_CUDA(cudaEventRecord(C.out_free[0], C.d2h));
        ZK_CUDA(cudaStreamWaitEvent(st, C.out_free[0], 0));
    }
    ZK_CUDA(cudaEventRecord(C.ev[5], st));
    mark("end", 0, st);
    if (host_input || ext_host || (tree && !tree_is_dev) || trace)
    {
        ZK_CUDA(cudaStreamSynchronize(st));
        ZK_CUDA(cudaEventElapsedTime(&C.stage[4], C.ev[4], C.ev[5]));
    }
    if (trace && !marks.empty())
    {
        ZK_CUDA(cudaStreamSynchronize(C.h2d));
        ZK_CUDA(cudaStreamSynchronize(D.cs));
        char line[4096];
        int o = snprintf(line, sizeof(line), "[zkgpu dist] rank %d timeline (ms):", me);
        for (auto &m : marks)
        {
            float ms = 0;
            cudaEventElapsedTime(&ms, C.pool[marks[0].e], C.pool[m.e]);
            if (o < (int)sizeof(line) - 32) o += snprintf(line + o, sizeof(line) - o, " %s%d=%.1f", m.what, m.k, ms);
        }
        fprintf(stderr, "%s\n", line);
    }
    return 0;
}

// the single-process form: one worker thread per device, each running the rank program on its own context
int dist_world_of_process() { return g_group.n; }

int dist_multi_commit(u64 *out_ext, u64 *tree, const u64 *input, u64 n_ext, u64 n, u64 ncols)
{
    const int G = g_group.n;
    if (is_device_ptr(input) || (out_ext && is_device_ptr(out_ext)))
    {
        set_error("zkgpu multi-GPU commit: the trace and the extended trace must be host buffers (or out_ext NULL: the extended "
                  "trace stays on the devices as row shards); the tree may be a host buffer or memory of device 0");
        return -1;
    }
    const bool tree_dev = tree && is_device_ptr(tree);
    if (tree_dev && ptr_device(tree) != g_group.ctx[0]->device)
    {
        set_error("zkgpu multi-GPU commit: a device tree must be memory of rank 0's device (%d), this one is on device %d",
                  g_group.ctx[0]->device, ptr_device(tree));
        return -1;
    }
    // Callers may be many host threads (the drop-in classes are used from OpenMP regions).  The worker threads of two commits
    // would meet at the same barriers and enqueue on the same communicators in arbitrary order: one commit at a time.
    std::lock_guard<std::mutex> one_at_a_time(g_multi_mu);
    Plan pl;
    ZK_TRY(make_plan(pl, ncols, G, true));
    int rcs[MAX_WORLD] = {};
    char errs[MAX_WORLD][512];
    std::vector<std::thread> th;
    for (int g = 0; g < G; g++)
        th.emplace_back([&, g]() {
            Enter en(g_group.ctx[g]->device);
            // the first agreement of the call: a rank that cannot take its context tells the others, and all of them return
            const int all_in = g_group.agree(g, en.rc);
            if (en.rc < 0 || all_in < 0)
            {
                rcs[g] = en.rc < 0 ? en.rc : all_in;
                snprintf(errs[g], sizeof(errs[g]), "%s", en.rc < 0 ? last_error() : "another rank could not enter its context");
                return;
            }
            PieceSrc ps[MAX_ROUNDS];
            for (int k = 0; k < pl.rounds; k++) ps[k] = {input + pl.c0[k][g], ncols};
            rcs[g] = dist_commit_rank(*en.c, ps, true, n_ext, n, ncols, (tree_dev && g != 0) ? nullptr : tree, tree && !tree_dev, tree_dev, out_ext);
            if (rcs[g] < 0) snprintf(errs[g], sizeof(errs[g]), "%s", last_error());
        });
    for (auto &t : th) t.join();
    for (int g = 0; g < G; g++)
        if (rcs[g] < 0)
        {
            set_error("rank %d: %s", g, errs[g]);
            return rcs[g];
        }
    return 0;
}

} // namespace zk

using namespace zk;
extern "C" {

int zkgpu_nccl_unique_id(void *out128)
{
    ZK_TRY(nccl_load());
    ncclUniqueId id;
    ZK_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(out128, &id, sizeof(id));
    return 0;
}

int zkgpu_dist_init(const void *nccl_unique_id, int rank, int world)
{
    if (world < 1 || world > MAX_WORLD || (world & (world - 1)) || rank < 0 || rank >= world || !nccl_unique_id)
    {
        set_error("zkgpu_dist_init: need a unique id, 0 <= rank < world, world a power of two <= %d", MAX_WORLD);
        return -1;
    }
    ZK_ENTER;
    if (C.dist)
    {
        set_error("zkgpu_dist_init: this device already belongs to a job");
        return -1;
    }
    ZK_TRY(nccl_load());
    ncclUniqueId id;
    memcpy(&id, nccl_unique_id, sizeof(id));
    ncclComm_t comm = nullptr;
    ZK_NCCL(g_nccl.CommInitRank(&comm, world, id, rank));
    ZK_TRY(dist_create(C, rank, world, comm, false));
    return dist_setup_rows(C, *C.dist, 1); // collective: finds out whether the peers' memory can be mapped (exchange mode)
}

int zkgpu_init_multi(int ngpu)
{
    int have = zkgpu_device_count();
    if (ngpu <= 0) ngpu = have;
    if (ngpu > have || ngpu > MAX_WORLD || (ngpu & (ngpu - 1)) || ngpu < 1)
    {
        set_error("zkgpu_init_multi: %d GPUs requested, %d visible; need a power of two <= %d", ngpu, have, MAX_WORLD);
        return -1;
    }
    if (g_group.n)
    {
        if (g_group.n == ngpu) return 0;
        set_error("zkgpu_init_multi: already initialised with %d GPUs", g_group.n);
        return -1;
    }
    if (ngpu == 1) return zkgpu_init(0);
    ZK_TRY(nccl_load());
    int devs[MAX_WORLD];
    for (int g = 0; g < ngpu; g++)
    {
        devs[g] = g;
        Enter en(g);
        if (en.rc < 0) return en.rc;
        g_group.ctx[g] = en.c;
        for (int h = 0; h < ngpu; h++)
            if (h != g)
            {
                int can = 0;
                if (cudaDeviceCanAccessPeer(&can, g, h) == cudaSuccess && can)
                {
                    cudaError_t e = cudaDeviceEnablePeerAccess(h, 0);
                    if (e != cudaSuccess) cudaGetLastError(); // already enabled is fine
                }
            }
    }
    ncclComm_t comms[MAX_WORLD];
    ZK_NCCL(g_nccl.CommInitAll(comms, ngpu, devs));
    for (int g = 0; g < ngpu; g++)
    {
        Enter en(g);
        if (en.rc < 0) return en.rc;
        ZK_TRY(dist_create(*en.c, g, ngpu, comms[g], true));
    }
    g_group.n = ngpu;
    {
        bool ok = exchange_wanted() != EX_NCCL;
        for (int g = 0; g < ngpu && ok; g++)
            for (int h = 0; h < ngpu && ok; h++)
                if (g != h)
                {
                    int can = 0;
                    if (cudaDeviceCanAccessPeer(&can, g, h) != cudaSuccess || !can) ok = false;
                }
        g_exchange = ok ? exchange_wanted() : (int)EX_NCCL;
    }
    zk::Enter back(0); // the calling thread stays bound to device 0
    return back.rc;
}

int zkgpu_dist_info(int *rank, int *world, int *peer_stores)
{
    ZK_ENTER;
    if (rank) *rank = C.dist ? C.dist->rank : 0;
    if (world) *world = g_group.n ? g_group.n : (C.dist ? C.dist->world : 1);
    if (peer_stores) *peer_stores = C.dist ? (exchange_now() == EX_STORES ? 1 : exchange_now() == EX_DMA ? 2 : 0) : 0;
    return 0;
}

int zkgpu_dist_plan(uint64_t ncols, int world, int host_input, int *rounds, uint64_t *c0, uint64_t *w)
{
    if (world < 1 || world > MAX_WORLD)
    {
        set_error("zkgpu_dist_plan: world out of range");
        return -1;
    }
    Plan pl;
    ZK_TRY(make_plan(pl, ncols, world, host_input != 0));
    if (rounds) *rounds = pl.rounds;
    for (int k = 0; k < pl.rounds; k++)
        for (int g = 0; g < world; g++)
        {
            if (c0) c0[k * world + g] = pl.c0[k][g];
            if (w) w[k * world + g] = pl.w[k][g];
        }
    return 0;
}

int zkgpu_dist_commit(void *, uint64_t *tree, const uint64_t *const *pieces, const uint64_t *piece_ld, int host_input, int gather_tree,
                      uint64_t n_ext, uint64_t n, uint64_t ncols)
{
    if (n == 0 || n_ext == 0) return 0;
    if ((n & (n - 1)) || (n_ext & (n_ext - 1)) || n_ext < n)
    {
        set_error("dist_commit: N and N_Extended must be powers of two with N_Extended >= N");
        return -1;
    }
    ZK_ENTER;
    if (!C.dist || C.dist->single_process)
    {
        set_error("dist_commit: call zkgpu_dist_init first (process-per-GPU form); after zkgpu_init_multi use zkgpu_extend_and_merkelize");
        return -1;
    }
    Plan pl;
    ZK_TRY(make_plan(pl, ncols, C.dist->world, host_input != 0));
    PieceSrc ps[MAX_ROUNDS];
    for (int k = 0; k < pl.rounds; k++)
    {
        ps[k] = {pieces[k], piece_ld[k]};
        if (pl.w[k][C.dist->rank] && (pieces[k] == nullptr || is_device_ptr(pieces[k]) == (host_input != 0)))
        {
            set_error("dist_commit: piece %d must be a %s pointer", k, host_input ? "host" : "device");
            return -1;
        }
    }
    return dist_commit_rank(C, ps, host_input != 0, n_ext, n, ncols, tree, false, gather_tree != 0, nullptr);
}

// Leaving a job takes two steps with a meeting of all ranks between them: first every rank drains its streams and closes
// its mappings of the peers' row shards, then (once EVERY rank has done so) the memory is freed.  Freeing a shard a peer
// still has open through cudaIpcOpenMemHandle, or is still copying into, is undefined behaviour.
static void dist_quiesce(Ctx &C)
{
    Dist *D = C.dist;
    if (!D) return;
    cudaStreamSynchronize(C.cur);
    cudaStreamSynchronize(D->cs);
    cudaStreamSynchronize(D->xs);
    cudaStreamSynchronize(D->sa);
    if (C.h2d) cudaStreamSynchronize(C.h2d);
    if (C.d2h) cudaStreamSynchronize(C.d2h);
    if (!D->single_process)
        for (int g = 0; g < D->world; g++)
            if (D->ipc_open[g])
            {
                cudaIpcCloseMemHandle(D->ipc_open[g]);
                D->ipc_open[g] = nullptr;
            }
}
static void dist_release(Ctx &C)
{
    Dist *D = C.dist;
    if (!D) return;
    u64 *bufs[] = {D->rows, D->sub, D->top, D->tree_dev, D->in_buf, D->ext_buf, D->recv_buf, D->ipc_scratch, D->flag, D->peer_tab};
    for (u64 *b : bufs)
        if (b) cudaFree(b);
    if (D->comm) g_nccl.CommDestroy(D->comm);
    cudaEvent_t evs[] = {D->ev_cp, D->ev_sa, D->ev_a, D->ev_b};
    for (cudaEvent_t e : evs)
        if (e) cudaEventDestroy(e);
    for (int i = 0; i < MAX_ROUNDS; i++)
    {
        if (D->ev_lde[i]) cudaEventDestroy(D->ev_lde[i]);
        if (D->ev_bar[i]) cudaEventDestroy(D->ev_bar[i]);
        if (D->ev_unp[i]) cudaEventDestroy(D->ev_unp[i]);
    }
    cudaStreamDestroy(D->cs);
    cudaStreamDestroy(D->xs);
    cudaStreamDestroy(D->sa);
    delete D;
    C.dist = nullptr;
}

int zkgpu_dist_finalize(void)
{
    if (g_group.n)
    {
        std::lock_guard<std::mutex> no_commit_running(g_multi_mu);
        for (int g = 0; g < g_group.n; g++)
        {
            Enter en(g_group.ctx[g]->device);
            if (en.rc >= 0) dist_quiesce(*en.c);
        }
        for (int g = 0; g < g_group.n; g++)
        {
            Enter en(g_group.ctx[g]->device);
            if (en.rc >= 0) dist_release(*en.c);
        }
        g_group.n = 0;
        g_exchange = -1;
        zk::Enter back(0);
        return back.rc;
    }
    ZK_ENTER;
    if (!C.dist) return 0;
    dist_quiesce(C);
    if (C.dist->world > 1)
    {
        // collective in the process-per-GPU form: every rank of the job calls zkgpu_dist_finalize
        Dist &D = *C.dist;
        ZK_NCCL(g_nccl.AllReduce(D.flag, D.flag + 1, 1, ncclUint64, ncclSum, D.comm, D.cs));
        ZK_CUDA(cudaStreamSynchronize(D.cs));
    }
    dist_release(C);
    g_exchange = -1;
    return 0;
}

int zkgpu_dist_rows(uint64_t **rows, uint64_t *nrows, uint64_t *ncols, uint64_t **subtree)
{
    ZK_ENTER;
    if (!C.dist)
    {
        set_error("dist_rows: no multi-GPU job on this device");
        return -1;
    }
    if (rows) *rows = C.dist->rows;
    if (nrows) *nrows = C.dist->rows_n;
    if (ncols) *ncols = C.dist->rows_c;
    if (subtree) *subtree = C.dist->sub;
    return 0;
}

} // extern "C"
