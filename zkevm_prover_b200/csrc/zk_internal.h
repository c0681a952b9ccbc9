// zk_internal.h — declarations shared by the translation units of libzkgpu.so (not installed).
#pragma once
#include <cstdint>
#include <cstddef>
#include <cuda_runtime.h>

typedef uint64_t u64;
typedef uint32_t u32;

namespace zk {

// ---- error plumbing (api.cu) ----
void set_error(const char *fmt, ...);
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
extern unsigned long long g_launches;
#define ZK_LAUNCHED() (++zk::g_launches)

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

// natural-order in, natural-order out transform of every column of a row-major matrix.
//   src: size x ncols words, row stride src_ld;  dst: row r of the result goes to dst + r*dst_ld
//   tmp: size*ncols words (row stride ncols), needed when log2(size) > 8; may alias neither src nor dst
//        (src == dst is allowed).
//   pre : applied to the input rows  (row index = input row)
//   post: applied to the output rows (row index = output row)
int ntt_transform(u64 *dst, u64 dst_ld, const u64 *src, u64 src_ld, u64 *tmp, unsigned log_size, u64 ncols,
                  bool inverse, const Scale &pre, const Scale &post, cudaStream_t st);

// two-level power table of g with a constant factor: G0[i] = g^i (i < 4096), G1[i] = scale * g^(4096 i), i < max(1, n >> 12).
// Cached for the life of the process.
int get_power_tables(u64 g, u64 scale, unsigned log_n, const u64 **G0, const u64 **G1, cudaStream_t st);

// LDE of a column strip: out rows are blowup*k + r (natural order of the extended domain).
//   in : n x ncols (row stride in_ld), out: (n << log_blowup) x ncols (row stride out_ld)
//   scratch: 2 * n * ncols words.  in may alias out (same ld) as the reference allows.
int lde_strip(u64 *out, u64 out_ld, const u64 *in, u64 in_ld, u64 *scratch, unsigned log_n, unsigned log_blowup,
              u64 ncols, cudaStream_t st);

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

// ---- misc kernels (api.cu) ----
int fill_splitmix(u64 *dst, u64 n, u64 seed, u64 offset, cudaStream_t st);

} // namespace zk
