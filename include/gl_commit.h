/* gl_commit.h -- C ABI of the B200 STARK-commit library (libglb200.so).
 *
 * Drop-in boundary for the commit hot path of MerlinLayer2/merlin-zkevm-prover.  The reference
 * has no FFI layer; its boundary is the C++ class API called by Starks::genProof
 * (src/starkpil/starks.cpp:52-56,133-137,214-218,261,284,289) and FRIProve::prove
 * (src/starkpil/fri/friProve.cpp:100-102,120-123).  Each entry point below names the reference
 * member it replaces; the header-compatible C++ shims in merlin-zkevm-prover_b200/host/ forward to
 * these functions.
 *
 * Conventions
 *   - plain pointers and sizes only; all matrices are row-major uint64_t words, element (row i,
 *     column j) of an n x ncols matrix at a[i*ncols + j] (ntt_goldilocks.cpp:95-110).
 *   - inputs may be any representative in [0, 2^64); every output word is CANONICAL (x mod p,
 *     p = 2^64 - 2^32 + 1), i.e. equal to Goldilocks::toU64 of the reference's result.
 *   - every function returns 0 on success, non-zero on error; gl_last_error() describes the last
 *     failure of the calling thread.  Nothing falls back to the CPU: without a CUDA device the
 *     calls fail.
 *   - *_host functions take HOST pointers, are synchronous and include the host<->device copies.
 *     *_dev functions take DEVICE pointers (cudaMalloc'ed, e.g. from gl_dev_alloc or a torch
 *     tensor's data_ptr) and run on the library's stream; they are synchronous on return unless
 *     noted.
 *   - aliasing follows the reference: dst == src or dst == NULL means in place; extendPol
 *     allows output == input.  When dst != src the source is not modified.
 */
#ifndef GL_COMMIT_H
#define GL_COMMIT_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GL_OK 0
#define GL_ERR_CUDA 1
#define GL_ERR_ARG 2
#define GL_ERR_RANGE 3 /* domain > 2^32: reference throws std::range_error (ntt_goldilocks.hpp:94-97) */
#define GL_ERR_NOMEM 4
#define GL_ERR_NCCL 5

/* ---- lifetime ------------------------------------------------------------------------- */
/* Bind the calling process to CUDA device `device`, upload constant tables.  Idempotent. */
int gl_init(int device);
void gl_shutdown(void);
const char *gl_last_error(void);
int gl_device_count(int *count);
/* Upper bound (bytes) for the library's internal scratch arena; 0 = default (32 GiB, and never more than half of the free device memory).
 * The call also hands back to the driver the device buffers that the *_host commit calls keep between calls (the extended matrix,
 * the tree and the upload staging: allocating and freeing 80 GB costs 50-120 ms per call); the library releases them by itself
 * whenever one of its own device allocations (gl_dev_alloc included) would otherwise fail. */
int gl_set_scratch_limit(size_t bytes);
/* Resident stage (off by default).  When on, gl_extend_pol_host -- NTT_Goldilocks::extendPol of the shims -- also hashes the
 * leaves and builds the tree while the extended matrix is being downloaded (the download hides it) and keeps the device copies
 * of both, keyed by the HOST pointer of `out`.  The gl_merkletree_host (MerkleTreeGL::merkelize, starks.cpp:52-56) that follows
 * on the same pointer and shape then only downloads the nodes: no 80 GB upload, no second hashing.  A sample of the host words is
 * compared with the device copy first; a buffer the host has written into since falls back to the normal path.  One stage is
 * resident at a time (the next gl_extend_pol_host / gl_commit_host replaces it); turning the mode off frees it. */
int gl_set_resident(int on);
/* Device pointers of the resident matrix / tree for queries (gl_group_proofs_dev, gl_rows_gather_dev); fails if `host_src` is
 * not the resident matrix. */
int gl_resident_query(const uint64_t *host_src, uint64_t ncols, uint64_t nrows, uint64_t **lde_dev, uint64_t **tree_dev);
/* Which NTT pass kernel runs the multi-pass transforms: 0 (default) ntt::k_pass -- direct global loads, two CTAs per SM alternate their
 * load and butterfly phases; 2 / 3 ntt::k_pass2 -- persistent CTAs fed by cp.async / TMA through a ring of 2 / 3 shared-memory tiles.
 * Same results; kept selectable because their ranking is shape dependent. */
int gl_set_ntt_variant(int variant);
/* Number of CUDA kernels launched by this library since the last call with reset != 0. */
uint64_t gl_kernel_launches(int reset);

/* ---- memory helpers ---------------------------------------------------------------------- */
int gl_dev_alloc(void **dptr, size_t bytes);
int gl_dev_free(void *dptr);
int gl_host_alloc(void **hptr, size_t bytes); /* pinned host memory */
int gl_host_free(void *hptr);
/* Page-lock an existing host allocation (e.g. the prover's big calloc/mmap arena, prover.cpp:106-119) so that
 * the *_host calls can overlap their copies with compute; pageable memory works but serialises. */
int gl_host_register(void *hptr, size_t bytes);
int gl_host_unregister(void *hptr);
int gl_h2d(void *dptr, const void *hptr, size_t bytes);
int gl_d2h(void *hptr, const void *dptr, size_t bytes);
int gl_sync(void);
/* Strided copy (cudaMemcpy2D semantics, pitches and width in bytes): kind 0 host->device, 1 device->host,
 * 2 device->device.  Used to move column blocks of row-major matrices (NTT nblock > 1, ntt_goldilocks.cpp:199-227). */
int gl_copy2d(void *dst, size_t dpitch, const void *src, size_t spitch, size_t width_bytes, size_t rows, int kind);
/* cudaStream_t of the library (as void*), for callers that want to order their own work. */
void *gl_stream(void);

/* ---- device-resident hot path -------------------------------------------------------------- */
/* NTT_Goldilocks::NTT / INTT (ntt_goldilocks.cpp:166-232, 307-324).
 * size must be a power of two <= 2^32 (0 or ncols == 0: no-op).  inverse != 0: INTT; extend != 0
 * (with inverse): coefficient i is additionally multiplied by 7^i (ntt_goldilocks.cpp:124-136). */
int gl_ntt_dev(uint64_t *dst, const uint64_t *src, uint64_t size, uint64_t ncols, int inverse, int extend);
/* NTT_Goldilocks::extendPol (ntt_goldilocks.cpp:326-352): out[k][c] = P_c(7 * w_E^k), k < n_ext,
 * where P_c interpolates in[.][c] over the size-n subgroup.  out holds n_ext*ncols words. */
int gl_extend_pol_dev(uint64_t *out, const uint64_t *in, uint64_t n_ext, uint64_t n, uint64_t ncols);
/* No counterpart in the reference.  gl_extend_pol_dev for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it): same arguments, same result, but the
 * twiddle / coset-scale tables and the column-block scratch live in one fixed file-scope device array (33 MiB, held by every GPU
 * that loads the library until the process exits) instead of cudaMalloc'ed buffers.  The call makes no cudaMalloc / cudaFree and
 * calls no allocating library routine (run gl_init once beforehand: it creates the library's stream).  LIMIT: n_ext <= 2^16 (any
 * ncols; columns go through the array in blocks of (2^22 / (n + n_ext)) columns); larger sizes fail with GL_ERR_RANGE and a message
 * naming the limit.  All calls share the array, which keeps the last call's contents (nothing in it has to start at zero: every word
 * is written before it is read); calls are serialised on the library's stream -- work that overlaps them on the GPU through a
 * second copy of the library's state would race on the array. */
int gl_extend_pol_dev_noalloc(uint64_t *out, const uint64_t *in, uint64_t n_ext, uint64_t n, uint64_t ncols);
/* PoseidonGoldilocks::merkletree_avx / MerkleTreeGL::merkelize (poseidon_goldilocks.cpp:266-301,
 * merkleTreeGL.cpp:37-44): tree holds (2*nrows-1)*4 words, leaf digests first, root last. */
int gl_merkletree_dev(uint64_t *tree, const uint64_t *src, uint64_t ncols, uint64_t nrows);
/* PoseidonGoldilocks::merkletree_batch_{seq,avx,avx512} (poseidon_goldilocks.cpp:113-164, 302-352, 521-577):
 * every row is hashed in slices of batch_size columns, the slice digests are hashed again (a different root
 * from gl_merkletree_*; not called by Starks/MerkleTreeGL, kept for the reference's KATs tests.cpp:2021-2223). */
int gl_merkletree_batch_dev(uint64_t *tree, const uint64_t *src, uint64_t ncols, uint64_t nrows, uint64_t batch_size);
int gl_merkletree_batch_host(uint64_t *tree, const uint64_t *src, uint64_t ncols, uint64_t nrows, uint64_t batch_size);
/* Starks::genProof stage commit (starks.cpp:52-56): extendPol followed by merkelize.
 * lde (n_ext*ncols words) and tree ((2*n_ext-1)*4 words) are device buffers; root is a HOST
 * pointer to 4 words. */
int gl_commit_dev(uint64_t *lde, uint64_t *tree, const uint64_t *in, uint64_t n_ext, uint64_t n,
                  uint64_t ncols, uint64_t root[4]);
/* MerkleTreeGL::getGroupProof (merkleTreeGL.cpp:12-35): proof (HOST, width + 4*log2(height) words)
 * = row idx of src followed by the sibling digests bottom-up; src/tree are device pointers. */
int gl_group_proof_dev(uint64_t *proof_host, const uint64_t *src, const uint64_t *tree,
                       uint64_t height, uint64_t width, uint64_t idx);
/* All queries of one tree in ONE launch and ONE download (FRIProve::queryPol loop, friProve.cpp:219-250, over
 * MerkleTreeGL::getGroupProof, merkleTreeGL.cpp:12-35): proofs_host receives n proofs of width + 4*log2(height) words
 * back to back; idx_host holds the n row indices (HOST).  No device allocation on the steady-state path. */
int gl_group_proofs_dev(uint64_t *proofs_host, const uint64_t *src, const uint64_t *tree, uint64_t height,
                        uint64_t width, const uint64_t *idx_host, uint64_t n);
/* n rows of a device-resident row-major matrix (the LDE row reads of Starks::evmap, starks.cpp:555-668):
 * rows_host[q*width + j] = src[idx_host[q]*width + j], canonical. */
int gl_rows_gather_dev(uint64_t *rows_host, const uint64_t *src, uint64_t width, const uint64_t *idx_host, uint64_t n);
/* Batched permutation PoseidonGoldilocks::hash_full_result (poseidon_goldilocks.cpp:166-228):
 * n states of 12 words. */
int gl_hash_full_result_dev(uint64_t *out, const uint64_t *in, uint64_t n);
/* PoseidonGoldilocks::linear_hash of every row (poseidon_goldilocks.cpp:229-265): out = nrows*4 words. */
int gl_linear_hash_rows_dev(uint64_t *out, const uint64_t *src, uint64_t ncols, uint64_t nrows);

/* ---- host-pointer convenience (what the C++ shims call) -------------------------------------- */
int gl_ntt_host(uint64_t *dst, const uint64_t *src, uint64_t size, uint64_t ncols, int inverse, int extend);
int gl_extend_pol_host(uint64_t *out, const uint64_t *in, uint64_t n_ext, uint64_t n, uint64_t ncols);
int gl_merkletree_host(uint64_t *tree, const uint64_t *src, uint64_t ncols, uint64_t nrows);
/* extendPol + merkelize with one upload: lde_out / tree_out may be NULL if not wanted. */
int gl_commit_host(uint64_t *lde_out, uint64_t *tree_out, const uint64_t *in, uint64_t n_ext, uint64_t n,
                   uint64_t ncols, uint64_t root[4]);
int gl_hash_full_result_host(uint64_t *out, const uint64_t *in, uint64_t n);
int gl_linear_hash_host(uint64_t out[4], const uint64_t *in, uint64_t size);

/* Batched field arithmetic, Goldilocks::add/sub/mul (goldilocks_base_field_scalar.hpp:16-160):
 * op 0 add(a,b), 1 sub(a,b), 2 mul(a,b), 3 a*b+c, 4 add with canonical b, 5 a^2, 6 canonicalise a,
 * 7 / 8 / 9: the add / sub / canonicalise forms used inside the NTT butterflies (same values as 0 / 1 / 6).
 * out[i] is canonical.  b / c may be NULL when the op does not use them. */
int gl_field_op_host(uint64_t *out, const uint64_t *a, const uint64_t *b, const uint64_t *c, uint64_t n, int op);

/* Benchmark helper (not on the commit path): dst[i] = splitmix64(seed + (first_index+i)*0x9E3779B97F4A7C15)
 * mod p, i < words; dst is a device pointer. */
int gl_fill_synthetic_dev(uint64_t *dst, uint64_t words, uint64_t seed, uint64_t first_index);

/* Measurement helper for bench.py's roofline (not on the commit path): issue rates measured now on this device, in
 * lane-operations per second: [0] IMAD.WIDE.U32 (32x32+64), [1] IADD3, [2] DFMA. */
int gl_measure_issue_peaks(double lane_ops_per_s[3]);

/* ---- layout helpers (merklehash_goldilocks.hpp:25-28, merkleTreeGL.hpp:71-78) ----------------- */
uint64_t gl_tree_num_elements(uint64_t height);
uint64_t gl_merkle_proof_size(uint64_t height);

/* ---- per-phase device timings of the last gl_commit_dev / gl_extend_pol_dev / gl_merkletree_dev
 *      (CUDA events on the library stream), in milliseconds: [0] LDE, [1] leaf hashing, [2] tree levels */
int gl_last_timings(float ms[3]);

/* ---- callers either side of the stage commit, device-resident -------------------------------- */
/* Starks::genProof step 4 (starks.cpp:261-284): q_2ns (n_ext x qdim evaluations on the coset) -> cm4_2ns
 * (n_ext x qdim*qdeg): INTT, coset un-shift by (7^-n)^p with the (p,k)->(k,p) regrouping, NTT.  qdeg*n <= n_ext. */
int gl_q_extend_dev(uint64_t *cm4, const uint64_t *q_2ns, uint64_t n, uint64_t n_ext, uint64_t qdim, uint64_t qdeg);
/* ... followed by treesGL[3]->merkelize() / getRoot (starks.cpp:289-290).  root: HOST pointer. */
int gl_q_commit_dev(uint64_t *cm4, uint64_t *tree, const uint64_t *q_2ns, uint64_t n, uint64_t n_ext, uint64_t qdim,
                    uint64_t qdeg, uint64_t root[4]);
/* One folding step of FRIProve::prove (friProve.cpp:36-107, si > 0): pol (2^pol_bits x 3) -> out (2^next_bits x 3)
 * evaluated at the F_p^3 challenge x (HOST pointer, 3 words); shift_inv is polShiftInv of that step.
 * Asynchronous on the library stream (gl_sync or a later synchronous call orders it). */
int gl_fri_fold_dev(uint64_t *out, const uint64_t *pol, uint32_t pol_bits, uint32_t next_bits, uint64_t shift_inv,
                    const uint64_t x[3]);
/* FRIProve::getTransposed (friProve.cpp:252-270) on 3-word elements; asynchronous. */
int gl_fri_transpose_dev(uint64_t *aux, const uint64_t *pol, uint64_t degree, uint32_t transpose_bits);
/* fold + getTransposed + MerkleTreeGL(2^tree_bits, 3 * 2^(next_bits-tree_bits)).merkelize() + getRoot
 * (friProve.cpp:36-126).  pol2 (2^next_bits x 3) and aux (same size, the tree's source) are outputs. */
int gl_fri_layer_commit_dev(uint64_t *tree, uint64_t *aux, uint64_t *pol2, const uint64_t *pol, uint32_t pol_bits,
                            uint32_t next_bits, uint32_t tree_bits, uint64_t shift_inv, const uint64_t x[3],
                            uint64_t root[4]);

/* Host-pointer forms of gl_q_commit_dev / gl_fri_layer_commit_dev: HOST buffers in and out (outputs may be NULL when not wanted),
 * synchronous, copies included -- the literal replacements of starks.cpp:261-290 and of one step of friProve.cpp:36-126 for a caller
 * that keeps its polynomials in host memory. */
int gl_q_commit_host(uint64_t *cm4_host, uint64_t *tree_host, const uint64_t *q_2ns_host, uint64_t n, uint64_t n_ext, uint64_t qdim,
                     uint64_t qdeg, uint64_t root[4]);
int gl_fri_layer_commit_host(uint64_t *tree_host, uint64_t *aux_host, uint64_t *pol2_host, const uint64_t *pol_host, uint32_t pol_bits,
                             uint32_t next_bits, uint32_t tree_bits, uint64_t shift_inv, const uint64_t x[3], uint64_t root[4]);

/* Offline constant tree, the `bctree` tool (tools/starkpil/bctree/build_const_tree.cpp:333-407, GL branch): the
 * .consttree image [nPols, nExt, nExt x nPols extended constants, (2 nExt - 1) x 4 nodes] that
 * MerkleTreeGL(Goldilocks::Element *tree) re-opens (merkleTreeGL.hpp:24-32, starks.hpp:190).  Its interpolate()
 * (:198-330) is the same coset extension as extendPol (coefficients * 7^i / n).  HOST buffers; const_tree holds
 * gl_consttree_num_elements(npols, n_ext) words; the verification key is its last 4 words (:386-392). */
uint64_t gl_consttree_num_elements(uint64_t npols, uint64_t n_ext);
int gl_build_consttree_host(uint64_t *const_tree, const uint64_t *const_pols, uint64_t n, uint64_t n_ext, uint64_t npols);

/* [0] LDE, [1] leaf hashing, [2] tree levels (incl. the gathered top), [3] multi-GPU exchange */
int gl_last_timings4(float ms[4]);

/* ---- multi-GPU: one process per GPU (the reference is single-process OpenMP and has no
 *      counterpart).  Columns are split over ranks for extendPol (ntt_goldilocks.cpp:326-352 is
 *      column-independent), an all-to-all over NVLink (peer copies through the copy engines into CUDA-IPC mappings of the
 *      peers' row_panels; grouped ncclSend/ncclRecv as the fallback) turns column blocks into row blocks, each
 *      rank hashes its E/world rows (poseidon_goldilocks.cpp:277-281) and builds its subtree
 *      (:283-300); the world subtree roots are gathered and the top log2(world) levels hashed on every
 *      rank.  Result: the same root and node values as the single-GPU call.
 *      Rank r owns columns [gl_mgpu_col_start(C, world, r), gl_mgpu_col_start(C, world, r+1)): a contiguous split balanced in
 *      units of 8 columns (the sponge rate), so that no 8-word chunk of a row straddles two ranks' panels. */
int gl_mgpu_unique_id(uint8_t id[128]);                      /* rank 0; broadcast the 128 bytes to all ranks */
int gl_mgpu_init(const uint8_t id[128], int rank, int world); /* after gl_init(device); world in {1,2,4,8} */
int gl_mgpu_shutdown(void);
uint64_t gl_mgpu_col_start(uint64_t ncols, int world, int rank);
/* How the last gl_commit_mgpu[_host] moved the column blocks into row panels: 1 = copy engines (every rank maps the peers' row_panels
 * buffers through CUDA IPC -- they must come from cudaMalloc / gl_dev_alloc -- and pushes its rows into them with peer copies over
 * NVLink, no SM involved; the default), 0 = grouped ncclSend/ncclRecv (GLB_MGPU_EXCHANGE=nccl, or when a buffer could not be mapped). */
int gl_mgpu_exchange_mode(void);
/* in_cols: n x w (this rank's w columns, row-major, pitch w); lde_cols: n_ext x w (out: this rank's
 * columns of the LDE); row_panels: (n_ext/world) * ncols_total words (out: this rank's rows, stored as
 * `world` panels -- panel p at word offset (n_ext/world)*col_start(p), row-major with pitch w_p);
 * subtree: (2*n_ext/world - 1)*4 words (out); root: HOST pointer, 4 words (same on every rank). */
int gl_commit_mgpu(uint64_t *lde_cols, uint64_t *row_panels, uint64_t *subtree, const uint64_t *in_cols,
                   uint64_t n_ext, uint64_t n, uint64_t ncols_total, uint64_t root[4]);
/* The same with HOST input and output (what a Starks caller that keeps the trace in host memory uses,
 * starks.cpp:52-56): in_cols_host is this rank's n x w column block (row-major, pitch w); row_panels_host
 * receives the rank's row panels and subtree_host its subtree (either may be NULL).  The upload of column
 * sub-block k+1 overlaps the LDE of sub-block k, and the download of the panels overlaps the leaf hashing.
 * lde_cols / row_panels / subtree are caller-owned DEVICE work buffers of the sizes gl_commit_mgpu needs. */
int gl_commit_mgpu_host(uint64_t *row_panels_host, uint64_t *subtree_host, const uint64_t *in_cols_host,
                        uint64_t *lde_cols, uint64_t *row_panels, uint64_t *subtree, uint64_t n_ext, uint64_t n,
                        uint64_t ncols_total, uint64_t root[4]);
/* Assemble the full (2*n_ext-1)*4-word node array (MerkleTreeGL::nodes layout) on every rank. */
int gl_mgpu_gather_tree(uint64_t *tree_full, const uint64_t *subtree, uint64_t n_ext);
/* Benchmark helper: the column range [col0, col0+ncols_local) of gl_fill_synthetic_dev's matrix. */
int gl_fill_synthetic_cols_dev(uint64_t *dst, uint64_t nrows, uint64_t col0, uint64_t ncols_local,
                               uint64_t ncols_total, uint64_t seed);

#ifdef __cplusplus
}
#endif
#endif /* GL_COMMIT_H */
