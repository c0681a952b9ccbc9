// zk_internal.h — declarations shared by the translation units of libzkgpu.so (not installed).
#pragma once
#include <cstdint>
#include <cstddef>
#include <cuda_runtime.h>
#include <atomic>
#include <map>
#include <mutex>
#include <tuple>
#include <vector>

typedef uint64_t u64;
typedef uint32_t u32;

namespace zk {

// ---- error plumbing (api.cu) ----
void set_error(const char *fmt, ...);
const char *last_error();
#define ZK_CUDA(call)                                                                              \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess) {                                                                  \
            zk::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));   \
            return -2;                                                                             \
        }                                                                                          \
    } while (0)
#define ZK_TRY(expr)                                                                               \
    do {                                                                                           \
        int r__ = (expr);                                                                          \
        if (r__ < 0) return r__;                                                                   \
    } while (0)

// every kernel launch of the library bumps this (zkgpu_launch_count): evidence for "gpu_launches" in bench.py
extern std::atomic<unsigned long long> g_launches;
#define ZK_LAUNCHED() (zk::g_launches.fetch_add(1, std::memory_order_relaxed))

// ---- per-device context (ctx.cu).  One Ctx per CUDA device, created on first use.  A host thread is bound to the
// context of the device it named in zkgpu_init(); threads that never called zkgpu_init share the process default
// context (the first one created).  Every entry point locks its context and makes its device current, so calls from
// starkpil's OpenMP threads are safe (they serialise per device) and the worker threads of the multi-GPU commit each
// drive their own device concurrently.
struct Dist; // dist.cu: NCCL communicator + peer row shards of the sharded commit
struct Resident; // ctx.cu: pointer-keyed residency cache (host buffer -> device copy the library kept)
struct Ctx
{
    int device = -1;
    bool ready = false;
    int num_sms = 0;
    std::recursive_mutex mu;
    cudaStream_t own = nullptr, cur = nullptr;
    u64 *scratch = nullptr;
    size_t scratch_words = 0;
    cudaEvent_t ev[8] = {};
    bool timed = false;
    float stage[5] = {0, 0, 0, 0, 0};
    // strip pipeline (commit_pipeline): copy streams, buffer hand-over events and per-strip timing segments
    cudaStream_t h2d = nullptr, d2h = nullptr;
    cudaEvent_t in_ready[2] = {}, in_free[2] = {}, out_ready[2] = {}, out_free[2] = {};
    std::vector<cudaEvent_t> pool; // timing events, grown on demand
    size_t pool_used = 0;
    struct Seg
    {
        int kind, a, b;
    };
    std::vector<Seg> segs; // segments of the last pipelined call: kind 0 = LDE, 1 = leaf absorption
    // per-device tables and one-time uploads (module globals such as __constant__ arrays exist once per device)
    bool consts_ready = false, w256_ready = false;
    struct Tab
    {
        u64 *g0 = nullptr, *g1 = nullptr;
        cudaEvent_t built = nullptr;     // recorded after this table's build kernel
        cudaStream_t built_on = nullptr; // users on that stream are ordered behind the build anyway
    };
    std::map<std::tuple<u64, u64, unsigned>, Tab> tabs; // get_power_tables cache
    std::map<const void *, int> occupancy;                                  // kernel -> resident CTAs per SM (smem attribute set)
    std::vector<Resident> *resident = nullptr;
    size_t resident_budget = 0;
    Dist *dist = nullptr;
};
extern thread_local Ctx *t_ctx; // the context of the entry point running on this thread (set by Enter)
Ctx *ctx_for_device(int device, bool create);
int ctx_ensure_scratch(Ctx &C, size_t words);
int ctx_pipe_streams(Ctx &C);
int ctx_pipe_event(Ctx &C, int *idx, cudaStream_t st); // idx == nullptr: reset the pool cursor
bool is_device_ptr(const void *p);
int ptr_device(const void *p); // device ordinal of a device pointer, -1 for host memory

// RAII guard of every entry point: picks the calling thread's context, locks it, makes its device current
struct Enter
{
    Ctx *c = nullptr;
    Ctx *prev = nullptr;
    int rc = 0;
    std::unique_lock<std::recursive_mutex> lk;
    explicit Enter(int device = -1);
    ~Enter();
};
#define ZK_ENTER              \
    zk::Enter en__;           \
    if (en__.rc < 0) return en__.rc; \
    zk::Ctx &C = *en__.c;     \
    (void)C

// ---- occupancy knobs of the overlapped commit pipeline (set by api.cu around its launches; zkgpu_tune for experiments)
extern int g_absorb_threads;         // 128 (fill the SM) or 512 (one CTA per SM)
extern int g_ntt_threads_per_sm_cap; // 0 = fill the SM

// ---- NTT (ntt.cu) ----
enum ScaleMode { SCALE_NONE = 0, SCALE_CONST = 1, SCALE_POWER = 2 };

struct Scale {
    int mode = SCALE_NONE;
    u64 konst = 1;          // SCALE_CONST: multiply every element by konst
    const u64 *G0 = nullptr; // SCALE_POWER: element of row i is multiplied by G0[i & 4095] * G1[i >> 12]
    const u64 *G1 = nullptr;
};

// Output scatter of the sharded commit (dist.cu): the last pass of a coset transform stores row k of the size-N result
// as row e = (k << log_blowup) + coset of the EXTENDED domain, and extended rows are row-sharded over the GPUs of the
// job: row e belongs to rank e >> lg_rows, which receives it at peer_tab[rank] + (e & (2^lg_rows - 1)) * ld + col0.
// peer_tab is a device array of `world` addresses (as u64): this rank's own shard and its peers' shards mapped over NVLink
// (cudaDeviceEnablePeerAccess / CUDA IPC), so the exchange is done by the transform's own stores.
struct Scatter {
    const u64 *peer_tab = nullptr;
    u32 lg_rows = 0;
    u64 ld = 0, col0 = 0;
};

// natural-order in, natural-order out transform of every column of a row-major matrix.
//   src: size x ncols words, row stride src_ld;  dst: row r of the result goes to dst + r*dst_ld
//   tmp: size*ncols words (row stride ncols), needed when log2(size) > 8; may alias neither src nor dst
//        (src == dst is allowed).
//   pre : applied to the input rows  (row index = input row)
//   post: applied to the output rows (row index = output row)
int ntt_transform(u64 *dst, u64 dst_ld, const u64 *src, u64 src_ld, u64 *tmp, unsigned log_size, u64 ncols,
                  bool inverse, const Scale &pre, const Scale &post, cudaStream_t st, const Scatter *sc = nullptr, unsigned sc_log_b = 0,
                  unsigned sc_coset = 0);

// two-level power table of g with a constant factor: G0[i] = g^i (i < 4096), G1[i] = scale * g^(4096 i), i < max(1, n >> 12).
// Cached for the life of the process.
int get_power_tables(u64 g, u64 scale, unsigned log_n, const u64 **G0, const u64 **G1, cudaStream_t st);

// LDE of a column strip: out rows are blowup*k + r (natural order of the extended domain).
//   in : n x ncols (row stride in_ld), out: (n << log_blowup) x ncols (row stride out_ld)
//   scratch: 2 * n * ncols words.  in may alias out (same ld) as the reference allows.
//   sc != nullptr: out / out_ld are ignored and the extended rows are scattered over the peers' row shards (see Scatter)
int lde_strip(u64 *out, u64 out_ld, const u64 *in, u64 in_ld, u64 *scratch, unsigned log_n, unsigned log_blowup,
              u64 ncols, cudaStream_t st, const Scatter *sc = nullptr);

int coset_ntts(u64 *out, u64 out_ld, const u64 *coef, u64 coef_ld, u64 *tmp, unsigned log_n, unsigned log_blowup, u64 ncols, u64 g0,
               u64 scale, cudaStream_t st, const Scatter *sc = nullptr);

// ---- Poseidon / Merkle (poseidon.cu) ----
int poseidon_init_constants();
// leaves: tree[4*i .. 4*i+3] = linear_hash(input row i), rows of row_words words with stride row_ld
int merkle_leaves(u64 *leaves, const u64 *input, u64 row_words, u64 row_ld, u64 nrows, cudaStream_t st);
// one column strip of the row sponge (strip_words a multiple of 8 except for the last strip of a row); the running
// digest lives in leaves[4*i .. 4*i+3]
int merkle_absorb(u64 *leaves, const u64 *rows, u64 strip_words, u64 row_ld, u64 nrows, bool first, cudaStream_t st, u64 *copy_to = nullptr,
                  u64 copy_ld = 0);
// two-level (batched) leaves: PoseidonGoldilocks::merkletree_batch semantics
int merkle_leaves_batch(u64 *leaves, const u64 *input, u64 ncols, u64 dim, u64 batch, u64 nrows, cudaStream_t st);
// all node levels above nrows leaves stored at tree[0 .. 4*nrows): reference level-major layout
int merkle_levels(u64 *tree, u64 nrows, cudaStream_t st);
// n independent permutations / hashes (KAT hooks)
int poseidon_permute(u64 *out12, const u64 *in12, u64 n, cudaStream_t st);

// ---- multi-GPU (dist.cu) ----
int dist_world_of_process(); // GPUs of the single-process job (zkgpu_init_multi), 0 when there is none
int dist_multi_commit(u64 *out_ext, u64 *tree, const u64 *input, u64 n_ext, u64 n, u64 ncols);

// ---- misc kernels (api.cu) ----
int fill_splitmix(u64 *dst, u64 n, u64 seed, u64 offset, cudaStream_t st);

} // namespace zk
