/* oracle/gl_oracle.h -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * CPU restatement, in plain C, of the reference's STARK-commit hot path.
 * Every function names the reference file:line whose behaviour it restates.
 * All outputs are CANONICAL field elements (x mod p); the reference stores
 * lazily reduced words and compares via Goldilocks::toU64
 * (src/goldilocks/src/goldilocks_base_field_tools.hpp:152-163).
 *
 * PARITY PINNED: tests/test_oracle.py checks this file against the golden
 * vectors of the reference's own tests (src/goldilocks/tests/tests.cpp,
 * benchs/bench.cpp) and against the unmodified reference compiled into
 * oracle/_ref/ (oracle/Makefile).
 */
#ifndef GLB200_GL_ORACLE_H
#define GLB200_GL_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define GLO_P 0xFFFFFFFF00000001ULL

/* field: goldilocks_base_field_scalar.hpp:16-160 (values, not representation) */
uint64_t glo_canon(uint64_t a);
uint64_t glo_add(uint64_t a, uint64_t b);
uint64_t glo_sub(uint64_t a, uint64_t b);
uint64_t glo_mul(uint64_t a, uint64_t b);
uint64_t glo_pow(uint64_t a, uint64_t e);
uint64_t glo_inv(uint64_t a);
/* Goldilocks::w(k), goldilocks_base_field.cpp:11-44; SHIFT=7 :51 */
uint64_t glo_root_of_unity(unsigned k);

/* NTT_Goldilocks::NTT / INTT, ntt_goldilocks.cpp:166-232, 307-324.
 * dst may equal src or be NULL (in place).  inverse!=0 -> INTT; extend!=0 (with inverse)
 * additionally scales coefficient i by 7^i (ntt_goldilocks.cpp:124-136). */
void glo_ntt(uint64_t *dst, const uint64_t *src, uint64_t size, uint64_t ncols, int inverse, int extend);
/* NTT_Goldilocks::extendPol, ntt_goldilocks.cpp:326-352 */
void glo_extend_pol(uint64_t *out, const uint64_t *in, uint64_t n_ext, uint64_t n, uint64_t ncols);

/* PoseidonGoldilocks::hash_full_result(_seq), poseidon_goldilocks.cpp:5-37 */
void glo_hash_full_result(uint64_t state[12], const uint64_t input[12]);
/* PoseidonGoldilocks::linear_hash(_seq), poseidon_goldilocks.cpp:38-74 */
void glo_linear_hash(uint64_t out[4], const uint64_t *in, uint64_t size);
/* PoseidonGoldilocks::merkletree_seq/_avx, poseidon_goldilocks.cpp:75-112, 266-301 */
void glo_merkletree(uint64_t *tree, const uint64_t *input, uint64_t ncols, uint64_t nrows);
/* PoseidonGoldilocks::merkletree_batch_seq, poseidon_goldilocks.cpp:113-164 */
void glo_merkletree_batch(uint64_t *tree, const uint64_t *input, uint64_t ncols, uint64_t nrows, uint64_t batch);
/* MerklehashGoldilocks::getTreeNumElements, merklehash_goldilocks.hpp:25-28 */
uint64_t glo_tree_num_elements(uint64_t height);
/* MerkleTreeGL::MerkleProofSize / getGroupProof, merkleTreeGL.hpp:71-78, merkleTreeGL.cpp:12-35 */
uint64_t glo_merkle_proof_size(uint64_t height);
void glo_group_proof(uint64_t *proof, const uint64_t *source, const uint64_t *nodes,
                     uint64_t height, uint64_t width, uint64_t idx);
/* Goldilocks3::mul, goldilocks_cubic_extension.hpp:1817-1830 */
void glo_gl3_mul(uint64_t r[3], const uint64_t a[3], const uint64_t b[3]);
/* Q-polynomial extension of Starks::genProof step 4, starks.cpp:261-284 (INTT, coset un-shift + (p,k)->(k,p)
 * transpose, NTT): q_2ns n_ext x qdim -> cm4 n_ext x qdim*qdeg */
void glo_q_extend(uint64_t *cm4, const uint64_t *q_2ns, uint64_t n, uint64_t n_ext, uint64_t qdim, uint64_t qdeg);
/* one FRI folding step (si > 0) of FRIProve::prove, friProve.cpp:36-107 */
void glo_fri_fold(uint64_t *out, const uint64_t *pol, unsigned pol_bits, unsigned next_bits, uint64_t shift_inv, const uint64_t x[3]);
/* FRIProve::getTransposed, friProve.cpp:252-270 */
void glo_fri_transpose(uint64_t *aux, const uint64_t *pol, uint64_t degree, unsigned transpose_bits);
/* xorshift64 synthetic input of SURVEY 8(c)/(d): seed 88172645463325252, value = s mod p */
void glo_fill_xorshift(uint64_t *dst, uint64_t count, uint64_t *state);
int glo_num_threads(void);
void glo_set_num_threads(int n);
#ifdef __cplusplus
}
#endif
#endif
