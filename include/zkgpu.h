/* zkgpu.h — C ABI of libzkgpu.so: the B200 (sm_100a) implementation of zkevm-prover's STARK commit hot path
 * (Goldilocks NTT / INTT / extendPol + Poseidon-Goldilocks Merkle tree).
 *
 * The reference has no FFI layer for this path: starkpil links the C++ classes NTT_Goldilocks,
 * PoseidonGoldilocks and MerkleTreeGL directly.  include/ntt_goldilocks.hpp, include/poseidon_goldilocks.hpp
 * and include/merkleTreeGL.hpp re-declare those classes with identical signatures and forward to the
 * entry points below; each entry point cites the reference interface it replaces
 * (paths relative to the reference repository root).
 *
 * Conventions
 *   - Every data pointer may be a HOST pointer (pageable or pinned) or a DEVICE pointer of the current
 *     device; the library detects which (cudaPointerGetAttributes).  Host data is staged through device
 *     memory; device data is used in place.  All pointers of one call must be of the same kind.
 *   - Elements are uint64_t words of the Goldilocks field p = 2^64 - 2^32 + 1.  Inputs may be any 64-bit
 *     word (the reference accepts non-canonical words); every output word is canonical (< p), i.e. equal
 *     to Goldilocks::toU64 of the reference's output (goldilocks_base_field_tools.hpp:152-163).
 *   - Return value: 0 on success, < 0 on error; zkgpu_last_error() gives the message (thread local).
 *   - Calls are synchronous with respect to the host when given host pointers.  With device pointers the
 *     work is enqueued on the library stream (zkgpu_set_stream) and the call returns without waiting.
 *   - There is no CPU fallback: without a usable CUDA device every compute entry point fails.
 */
#ifndef ZKGPU_H
#define ZKGPU_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime ---- */
int zkgpu_init(int device);            /* bind the calling host thread to this device's context (created on first use);
                                          -1 = the context the thread already has, else the process default */
int zkgpu_set_stream(void *cuda_stream); /* cudaStream_t to enqueue on (0 = the legacy default stream);
                                            (void*)-1 = back to the library's own non-blocking stream */
int zkgpu_sync(void);                  /* wait for everything enqueued on the library stream */
const char *zkgpu_last_error(void);
int zkgpu_device_count(void);
uint64_t zkgpu_launch_count(void);     /* kernels this library has launched so far in this process */
int zkgpu_mem_info(uint64_t *free_bytes, uint64_t *total_bytes);
/* tuning / experiment hook: key 0 = threads per CTA of the sponge kernel (128 | 512), key 2 = commit cache on/off, key 1 = cap on transform threads per
 * SM (0 = none).  Used by tools/overlap_probe.py; not needed in normal use. */
int zkgpu_tune(int key, int value);
/* measurement hook (bench.py): sustained rates of this GPU's pipes, in thread-instructions (or operations) per second for the
 * whole GPU, from register-only loops of 8 independent chains at 32 warps per SM (csrc/probe.cu; ~30 ms):
 * out[0] IMAD.WIDE.U32, [1] IADD3 (ALU pipe), [2] DFMA, [3] DFMA and IMAD.WIDE interleaved 1:1 (both counted), [4] gl::mul_nc
 * (un-normalised 64-bit modular multiplies), [5] radix-2 butterflies (modular multiply + add + sub), [6] SM count, [7] 0.
 * The arithmetic roofs in the bench line are derived from these, measured in the same run.  No counterpart in the reference. */
int zkgpu_pipe_probe(double out[8]);
/* device-memory helpers so that C/C++ hosts need no CUDA headers */
int zkgpu_malloc(void **dev_ptr, uint64_t bytes);
int zkgpu_free(void *dev_ptr);
int zkgpu_malloc_host(void **host_ptr, uint64_t bytes); /* pinned */
int zkgpu_free_host(void *host_ptr);
int zkgpu_memcpy(void *dst, const void *src, uint64_t bytes); /* any direction, synchronous */
/* device timings (ms, CUDA events) of the stages of the last commit-type call on this thread's context:
 * [0] LDE (STARK_STEP_k_LDE), [1] Merkle leaves, [2] Merkle node levels (STARK_STEP_k_MERKLETREE = [1]+[2]),
 * [3] unused (0), [4] the whole call including host<->device copies when any pointer was a host pointer.  LDE and
 * leaf absorption run strip by strip, interleaved: [0] and [1] are sums over the strips.  Waits for the stream. */
int zkgpu_stage_ms(float out[5]);

/* synthetic input generator shared with the CPU oracle: dst[i] = splitmix64(seed + offset + i) mod p */
int zkgpu_fill(uint64_t *dst, uint64_t n, uint64_t seed, uint64_t offset);

/* element-wise field ops on n words (KAT hook for the device arithmetic that replaces
 * goldilocks_base_field_scalar.hpp:16-160): op 0 = add, 1 = sub, 2 = mul, 3 = mul (un-normalised path, then toU64),
 * 4 = toU64.  Operands may be any 64-bit words; results are canonical. */
int zkgpu_field_op(int op, uint64_t *out, const uint64_t *a, const uint64_t *b, uint64_t n);

/* ---- NTT_Goldilocks (src/goldilocks/src/ntt_goldilocks.hpp:49-136 ctor, :166-173 methods) ---- */
/* NTT_Goldilocks(maxDomainSize, nThreads, extension): returns NULL (and sets the error) if log2(maxDomainSize) > 32,
 * where the reference throws std::range_error (ntt_goldilocks.hpp:94-97). */
void *zkgpu_ntt_create(uint64_t max_domain_size, int extension);
void zkgpu_ntt_destroy(void *ntt);
/* NTT_Goldilocks::NTT / ::INTT (ntt_goldilocks.cpp:166-232, 307-324).  dst == NULL or dst == src: in place.
 * size must be a power of two (the reference asserts, ntt_goldilocks.cpp:28); size == 0 or ncols == 0: no-op.
 * inverse: INTT (scaled by 1/size); extend (inverse only): additionally row i is multiplied by 7^i, i.e. the
 * reference's r_[i] = 7^i / size (ntt_goldilocks.hpp:153-165).  buffer/nphase/nblock of the reference are CPU
 * blocking knobs: buffer is accepted and not used. */
int zkgpu_ntt(void *ntt, uint64_t *dst, const uint64_t *src, uint64_t size, uint64_t ncols, uint64_t *buffer, int inverse,
              int extend);
/* NTT_Goldilocks::extendPol (ntt_goldilocks.cpp:326-352): output (N_Extended x ncols) = evaluations on the coset
 * 7*<w_NExt> of the polynomials interpolating the N x ncols input columns.  output == input is allowed. */
int zkgpu_extend_pol(void *ntt, uint64_t *output, const uint64_t *input, uint64_t n_extended, uint64_t n, uint64_t ncols,
                     uint64_t *buffer);

/* NTT_Goldilocks::reversePermutation (ntt_goldilocks.cpp:243-305): dst[i*ncols + c] = src[BR(i)*ncols_all + offset_cols + c]
 * for i < size (BR = bit reversal over log2(size) bits).  When the object was created with extension > 1, rows whose
 * source offset lies at or beyond (size/extension)*ncols_all are zero-filled (:259-277).  dst == src is allowed only
 * for extension <= 1 with offset_cols == 0 and ncols == ncols_all (the reference asserts the same, :283, :302). */
int zkgpu_reverse_permutation(void *ntt, uint64_t *dst, const uint64_t *src, uint64_t size, uint64_t offset_cols, uint64_t ncols,
                              uint64_t ncols_all);

/* ---- PoseidonGoldilocks (src/goldilocks/src/poseidon_goldilocks.hpp:43-81) ---- */
/* hash_full_result (poseidon_goldilocks.cpp:166-228): n independent 12-word permutations */
int zkgpu_hash_full_result(uint64_t *out12, const uint64_t *in12, uint64_t n);
/* linear_hash (poseidon_goldilocks.cpp:229-265): 4-word digest of `size` words (size <= 4: copy, not hashed) */
int zkgpu_linear_hash(uint64_t *out4, const uint64_t *input, uint64_t size);
/* merkletree / merkletree_avx (poseidon_goldilocks.hpp:74-81, .cpp:266-301): tree has (2*num_rows - 1) * 4 words,
 * leaves first then each level (MerklehashGoldilocks::getTreeNumElements, merklehash_goldilocks.hpp:15-28).
 * num_rows must be a power of two (the reference's loop is only well defined then); num_rows == 0: no-op. */
int zkgpu_merkletree(uint64_t *tree, const uint64_t *input, uint64_t num_cols, uint64_t num_rows, uint64_t dim);
/* merkletree_batch (poseidon_goldilocks.cpp:302-352) */
int zkgpu_merkletree_batch(uint64_t *tree, const uint64_t *input, uint64_t num_cols, uint64_t num_rows, uint64_t batch_size,
                           uint64_t dim);
/* The same tree as zkgpu_merkletree_batch, for callers that cannot allocate device memory.  (This entry point has no
 * counterpart in the reference: its CPU code simply mallocs the per-batch digests, poseidon_goldilocks.cpp:302-352.)
 * Who needs it: a prover that reserves the whole GPU for its own arena when it starts (cudaMalloc then has nothing
 * left to give, and the stream-ordered pool behind cudaMallocAsync cannot grow), a service whose device-memory budget is
 * fixed by its operator so that no library may allocate behind its back, and a proving loop that must not stall on the
 * driver's allocator (a pool that has to grow or trim takes a device-wide lock and may synchronise).
 * zkgpu_merkletree_batch asks the stream-ordered allocator for num_rows * ceil(num_cols / batch_size) * 4 words of
 * digest scratch on every call; this variant keeps that scratch in a fixed array inside the library image instead,
 * 2^20 words = 8 MiB, which every GPU that loads libzkgpu.so holds until the process exits.  It calls no cudaMalloc / cudaFree / cudaMallocAsync and no routine that allocates.
 *   - tree and input must be DEVICE pointers (host data would have to be staged, which needs memory); tree 16-byte aligned.
 *   - Shapes whose scratch exceeds 2^20 words (e.g. more than 2^14 rows at 16 batches) are refused with an error (-1);
 *     nothing is launched.  Use zkgpu_merkletree_batch for those.
 *   - The first call on each device looks up the array's device address (and, if no other entry point ran there yet,
 *     creates the library's context and uploads the Poseidon constants).  Make that first call before any stream
 *     capture begins.
 *   - All calls on one device share the one array.  It still holds the previous call's digests when the next one starts;
 *     that is harmless (every row writes its digests before reading them), but two calls whose kernels overlap on the
 *     GPU (the library stream was switched with zkgpu_set_stream between them) would overwrite each other's digests:
 *     keep calls to this entry point ordered on one stream per device.
 * Work is enqueued on the library stream and the call returns without waiting, as for every device-pointer call. */
int zkgpu_merkletree_batch_noalloc(uint64_t *tree, const uint64_t *input, uint64_t num_cols, uint64_t num_rows, uint64_t batch_size,
                                   uint64_t dim);

/* MerkleTreeGL::getGroupProof (src/starkpil/merkleTree/merkleTreeGL.cpp:12-35): proof[0 .. width) = row idx of source,
 * then for every level below the root the 4-word sibling digest (nodes is the level-major array zkgpu_merkletree
 * writes).  proof has width + 4*log2(height) words.  source / nodes / proof may each be host or device pointers. */
int zkgpu_group_proof(uint64_t *proof, const uint64_t *source, const uint64_t *nodes, uint64_t width, uint64_t height, uint64_t idx);

/* the same for nqueries rows at once (FRIProve::prove asks every tree for nQueries proofs, friProve.cpp:144-160, 219-235): one launch
 * and one copy instead of one of each per query.  idx is a host array; proofs receives nqueries * (width + 4*log2(height)) words. */
int zkgpu_group_proofs(uint64_t *proofs, const uint64_t *source, const uint64_t *nodes, uint64_t width, uint64_t height, const uint64_t *idx,
                       uint64_t nqueries);

/* ---- the commit stage as one call: Starks::genProof steps 1-3 (src/starkpil/starks.cpp:51-58, 133-139, 214-220):
 * ntt.extendPol(out_ext, in, NExt, N, ncols) followed by MerkleTreeGL(NExt, ncols, out_ext).merkelize().
 * out_ext: n_extended x ncols, tree: (2*n_extended - 1)*4 words.  Here each of the three pointers may independently
 * be a host or a device pointer (e.g. host trace in, extended trace kept resident on the device for the later
 * getGroupProof / stage-4 calls, tree back on the host). */
int zkgpu_extend_and_merkelize(void *ntt, uint64_t *out_ext, uint64_t *tree, const uint64_t *input, uint64_t n_extended,
                               uint64_t n, uint64_t ncols);

/* The commit cache (csrc/api.cu: Resident).  Unchanged starks.cpp calls ntt.extendPol(p_cm_2ns, ...) (starks.cpp:53) and then
 * treesGL[k]->merkelize() on the same host buffer (starks.hpp:186).  zkgpu_extend_pol with host pointers therefore also
 * hashes the strips while they are on the device and remembers the finished tree under (output pointer, rows, columns); the
 * zkgpu_merkletree call on that buffer returns it without uploading the extended trace again.  A fingerprint of a few rows
 * guards against a buffer that was modified in between; zkgpu_invalidate(ptr) (NULL = everything) drops entries explicitly;
 * ZKGPU_FUSE_COMMIT=0 or zkgpu_tune(2, 0) turns the behaviour off. */
int zkgpu_invalidate(const void *host_ptr);
/* What the fingerprint covers, and the hazard that remains.  The first and the last row are compared word for word, and a
 * keyed 64-bit checksum is taken over 2^20 words at positions drawn from a per-entry random seed: every word when the buffer
 * has no more than that, and every word of any buffer with ZKGPU_CACHE_VERIFY=full in the environment (a threaded pass over
 * the host buffer: about a second at 89 GB, still cheaper than uploading and hashing it again).  In the default mode a caller
 * that rewrites a FEW words of a LARGE buffer between extendPol and merkelize() may go unnoticed and receive the tree of the
 * old contents; such a caller must call zkgpu_invalidate, set ZKGPU_CACHE_VERIFY=full, or switch the cache off.  Another
 * matrix at the same address (freed and re-allocated buffer) is caught in either mode.  The reference's genProof does not
 * write between the two calls (starks.cpp:53-57).  Entries wait in pinned host memory (no device memory is held for a tree
 * nobody has claimed), at most 4 of them and 5 GiB; an NTT_Goldilocks object two of whose trees went unclaimed stops hashing
 * in extendPol (StarkRecursiveF builds BN128 trees on its extended traces, starkRecursiveF.cpp:198-200).
 * zkgpu_cache_stats: out[0..3] = hits, misses, stale (fingerprint mismatch: tree recomputed), dropped unclaimed;
 * out[4] = entries waiting now, out[5] = bytes of DEVICE memory they hold (0 unless pinned memory could not be had). */
int zkgpu_cache_stats(uint64_t out[6]);

/* ---- stage 4 of Starks::genProof on the device (src/starkpil/starks.cpp:261-293):
 *   nttExtended.INTT(qq1, q_2ns, NExt, qDim);  qq2[k*qDeg + p] = qq1[p*N + k] * shiftIn^p (shiftIn = (1/7)^N, :266-281);
 *   nttExtended.NTT(cm4_2ns, qq2, NExt, qDim*qDeg);  treesGL[3]->merkelize().
 * q_2ns: NExt x qDim, cm4_2ns: NExt x (qDim*qDeg), both host or both device pointers; tree: (2*NExt-1)*4 words, host or
 * device, or NULL (no tree).  qDeg must not exceed NExt / N. */
int zkgpu_stage4_commit(void *ntt_extended, uint64_t *cm4_2ns, uint64_t *tree, const uint64_t *q_2ns, uint64_t n_extended, uint64_t n,
                        uint64_t q_dim, uint64_t q_deg);

/* ---- constant-tree builder: the Goldilocks branch of buildConstTree (tools/starkpil/bctree/build_const_tree.cpp:533-596;
 * interpolate :398-531 is the same LDE as extendPol: 1/N * shift^i pre-scale :361-395, coset shift 7).
 * image receives the reference's on-disk layout  [nPols, nExt, extended constant polynomials (nExt x nPols, row-major),
 * tree nodes ((2*nExt-1)*4 words)]  = 2 + nPols*nExt + (2*nExt-1)*4 words, which MerkleTreeGL(Goldilocks::Element *tree)
 * views directly (merkleTreeGL.hpp:24-32).  const_pols: (1<<n_bits) x n_pols row-major.  Host or device pointers. */
int zkgpu_build_const_tree(uint64_t *image, const uint64_t *const_pols, uint64_t n_bits, uint64_t n_bits_ext, uint64_t n_pols);

/* ---- FRI layer commit ----
 * One folding step for si > 0 (src/starkpil/fri/friProve.cpp:44-107): fri_pol holds 2^pol_bits elements of the cubic
 * extension (3 words each, Polinomial dim 3), pol2_e receives 2^next_bits elements:
 *   pol2_e[g] = evalPol( polMulAxi( INTT_nX( friPol[i * 2^next_bits + g], i < nX ), 1, shiftInv * w(polBits)^-g ), special_x )
 * with nX = 2^(pol_bits - next_bits) (polMulAxi :192-200, evalPol :201-217, the per-thread sinv walk :64-81, :105).
 * shift_inv is polShiftInv of the step (:34-35), special_x the transcript challenge (:30), a host pointer to 3 words.
 * pol2_e and fri_pol are both host or both device pointers.  (For si == 0 the reference copies friPol, :91-94.) */
int zkgpu_fri_fold(uint64_t *pol2_e, const uint64_t *fri_pol, uint64_t pol_bits, uint64_t next_bits, uint64_t shift_inv,
                   const uint64_t *special_x);
/* FRIProve::getTransposed (friProve.cpp:252-271) on 3-word elements: aux[i*h + j] = pol[j*w + i], w = 2^transpose_bits,
 * h = degree / w; aux is then the source of MerkleTreeGL(nGroups, groupSize * 3) (:118-122), i.e. of zkgpu_merkletree. */
int zkgpu_fri_transpose(uint64_t *aux, const uint64_t *pol, uint64_t degree, uint64_t transpose_bits);

/* ---- multi-GPU building blocks (device pointers only; one process per GPU, the host side does the exchange) ----
 * LDE of a column shard stored with an arbitrary row stride (columns are independent, ntt_goldilocks.cpp:103-110) */
int zkgpu_lde_columns(uint64_t *out, uint64_t out_ld, const uint64_t *in, uint64_t in_ld, uint64_t n_extended, uint64_t n,
                      uint64_t ncols);
/* leaf digests of a row shard (poseidon_goldilocks.cpp:277-281): leaves[4*i..] = linear_hash(row i) */
int zkgpu_merkle_leaves(uint64_t *leaves, const uint64_t *rows, uint64_t row_words, uint64_t row_ld, uint64_t num_rows);
/* one column strip of the row sponge (poseidon_goldilocks.cpp:244-256): row i continues from the running digest in
 * leaves[4*i..] (first != 0: starts from the zero capacity) and absorbs strip_words words found at rows + i*row_ld.
 * Every strip of a row but the last must be a multiple of 8 words; after the last strip leaves[] holds
 * linear_hash(row) for rows longer than 4 words. */
int zkgpu_merkle_absorb(uint64_t *leaves, const uint64_t *rows, uint64_t strip_words, uint64_t row_ld, uint64_t num_rows, int first);
/* the same, and the words absorbed are also stored at copy_to + i*copy_ld (assembles the row-major extended trace that
 * MerkleTreeGL::source points at, merkleTreeGL.hpp:18-28, from exchanged column pieces without a separate copy pass) */
int zkgpu_merkle_absorb_copy(uint64_t *leaves, const uint64_t *rows, uint64_t strip_words, uint64_t row_ld, uint64_t num_rows, int first,
                             uint64_t *copy_to, uint64_t copy_ld);
/* node levels above num_rows leaves already stored at tree[0 .. 4*num_rows) (poseidon_goldilocks.cpp:283-300) */
int zkgpu_merkle_levels(uint64_t *tree, uint64_t num_rows);
/* gather/scatter between a row-major matrix and a packed block: dst[r*ncols + c] = src[(row0 + r)*src_ld + col0 + c]
 * (pack) or the inverse (unpack) — used around the column-shard -> row-shard all-to-all */
int zkgpu_copy_block(uint64_t *dst, uint64_t dst_ld, const uint64_t *src, uint64_t src_ld, uint64_t nrows, uint64_t ncols);

/* ---- the commit stage over all GPUs of one box ----
 * The reference prover is one process with one proof in flight (src/prover/prover.cpp:182-221) whose commit stages are plain
 * method calls (src/starkpil/starks.cpp:51-58); so the multi-GPU commit has two forms behind this ABI:
 *
 * (1) single process (what the drop-in classes use): zkgpu_init_multi(ngpu) creates one context per device 0..ngpu-1,
 *     enables peer access and an NCCL communicator over them.  From then on zkgpu_extend_and_merkelize with a HOST trace is
 *     committed by all GPUs (one worker thread per device): out_ext may be a host buffer (every GPU copies its rows home) or
 *     NULL (the extended trace stays on the devices as row shards); tree may be a host buffer or memory of device 0.
 *     ngpu <= 0: all visible devices.  ngpu must be a power of two.
 * (2) one process per GPU (torchrun-style launchers): rank 0 calls zkgpu_nccl_unique_id and hands the 128 bytes to the other
 *     ranks by any means; every rank calls zkgpu_init(local_device) and zkgpu_dist_init(id, rank, world), then
 *     zkgpu_dist_commit with ITS column pieces of the trace.  zkgpu_dist_plan tells which columns those are:
 *     c0[k*world + g], w[k*world + g] = first column and width of the piece of rank g in round k (rounds <= 16).
 *     pieces[k] is an N x w matrix with row stride piece_ld[k]; all device pointers (host_input = 0) or all host pointers
 *     (host_input = 1: staged round by round under the compute; the plan differs, ask zkgpu_dist_plan with the same flag).
 *     gather_tree != 0 (same value on every rank): rank 0's `tree` (device or host, (2*NExt-1)*4 words) receives the whole
 *     level-major node array; other ranks pass NULL.  gather_tree == 0: the tree stays sharded.
 *     After the call zkgpu_dist_rows gives this rank's row shard of the extended trace (rows [rank*R, (rank+1)*R), R = NExt /
 *     world, row-major with ncols columns: the layout MerkleTreeGL::source has) and its subtree ((2R-1)*4 words).
 * The column-shard -> row-shard exchange goes into the peers' row shards over NVLink (peer
 * access / CUDA IPC); by default the rows are staged locally and the copy engines push them to the peers (ZKGPU_EXCHANGE=stores: the LDE stores them itself); without
 * peer access, or with ZKGPU_EXCHANGE=nccl, grouped ncclSend/ncclRecv per round do it. */
int zkgpu_init_multi(int ngpu);
int zkgpu_nccl_unique_id(void *out128);
int zkgpu_dist_init(const void *nccl_unique_id, int rank, int world);
int zkgpu_dist_info(int *rank, int *world, int *exchange); /* exchange: 1 = peer stores fused into the LDE, 2 = copy engines, 0 = NCCL */
int zkgpu_dist_plan(uint64_t ncols, int world, int host_input, int *rounds, uint64_t *c0, uint64_t *w);
int zkgpu_dist_commit(void *ntt, uint64_t *tree, const uint64_t *const *pieces, const uint64_t *piece_ld, int host_input, int gather_tree,
                      uint64_t n_extended, uint64_t n, uint64_t ncols);
int zkgpu_dist_rows(uint64_t **rows, uint64_t *num_rows, uint64_t *num_cols, uint64_t **subtree);
/* leave the job: waits for the device, frees the shards and staging buffers, destroys the communicator(s).  In the
 * process-per-GPU form this is COLLECTIVE: every rank calls it (the ranks meet between closing their mappings of the peers'
 * row shards and freeing their own, so that nobody frees memory a peer still has open). */
int zkgpu_dist_finalize(void);

#ifdef __cplusplus
}
#endif
#endif /* ZKGPU_H */
