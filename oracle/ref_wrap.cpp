// oracle/ref_wrap.cpp -- TEST INFRASTRUCTURE ONLY.
//
// extern "C" veneer over the UNMODIFIED reference classes so that the Python
// test-suite and bench.py's cpu_baseline / --impl reference legs can call the
// reference's own CPU implementation of the commit path.  Compiled together
// with the reference sources where they lie under /root/reference (see
// oracle/Makefile); nothing from the reference is copied into this repo.
//
//   NTT_Goldilocks::NTT/INTT/extendPol   src/goldilocks/src/ntt_goldilocks.cpp:166,307,326
//   PoseidonGoldilocks::*                src/goldilocks/src/poseidon_goldilocks.cpp
//   MerkleTreeGL::getGroupProof          src/starkpil/merkleTree/merkleTreeGL.cpp:12
#include "goldilocks_base_field.hpp"
#include "ntt_goldilocks.hpp"
#include "poseidon_goldilocks.hpp"
#include "merkleTreeGL.hpp"
#include <omp.h>

typedef Goldilocks::Element E;

extern "C" {

int ref_num_threads(void) { return omp_get_max_threads(); }

int ref_has_avx512(void)
{
#ifdef __AVX512__
    return 1;
#else
    return 0;
#endif
}

void ref_ntt(uint64_t *dst, uint64_t *src, uint64_t size, uint64_t ncols, uint64_t nphase, uint64_t nblock)
{
    NTT_Goldilocks ntt(size);
    ntt.NTT((E *)dst, (E *)src, size, ncols, NULL, nphase, nblock);
}

void ref_intt(uint64_t *dst, uint64_t *src, uint64_t size, uint64_t ncols, uint64_t nphase, uint64_t nblock)
{
    NTT_Goldilocks ntt(size);
    ntt.INTT((E *)dst, (E *)src, size, ncols, NULL, nphase, nblock);
}

void ref_extend_pol(uint64_t *out, uint64_t *in, uint64_t n_ext, uint64_t n, uint64_t ncols, uint64_t *buffer)
{
    NTT_Goldilocks ntt(n);
    ntt.extendPol((E *)out, (E *)in, n_ext, n, ncols, (E *)buffer);
}

void ref_hash_full_result(uint64_t *state, const uint64_t *input, int variant)
{
    if (variant == 0)
        PoseidonGoldilocks::hash_full_result_seq((E *)state, (const E *)input);
    else
        PoseidonGoldilocks::hash_full_result((E *)state, (const E *)input);
}

void ref_linear_hash(uint64_t *out, uint64_t *in, uint64_t size, int variant)
{
    if (variant == 0)
        PoseidonGoldilocks::linear_hash_seq((E *)out, (E *)in, size);
    else
        PoseidonGoldilocks::linear_hash((E *)out, (E *)in, size);
}

// variant: 0 = _seq, 1 = _avx (the product default, merkleTreeGL.cpp:42), 2 = _avx512 (if built in)
int ref_merkletree(uint64_t *tree, uint64_t *input, uint64_t ncols, uint64_t nrows, int variant)
{
    if (variant == 0)
        PoseidonGoldilocks::merkletree_seq((E *)tree, (E *)input, ncols, nrows);
    else if (variant == 1)
        PoseidonGoldilocks::merkletree_avx((E *)tree, (E *)input, ncols, nrows);
    else
    {
#ifdef __AVX512__
        PoseidonGoldilocks::merkletree_avx512((E *)tree, (E *)input, ncols, nrows);
#else
        return -1;
#endif
    }
    return 0;
}

int ref_merkletree_batch(uint64_t *tree, uint64_t *input, uint64_t ncols, uint64_t nrows, uint64_t batch, int variant)
{
    if (variant == 0)
        PoseidonGoldilocks::merkletree_batch_seq((E *)tree, (E *)input, ncols, nrows, batch);
    else if (variant == 1)
        PoseidonGoldilocks::merkletree_batch_avx((E *)tree, (E *)input, ncols, nrows, batch);
    else
    {
#ifdef __AVX512__
        PoseidonGoldilocks::merkletree_batch_avx512((E *)tree, (E *)input, ncols, nrows, batch);
#else
        return -1;
#endif
    }
    return 0;
}

uint64_t ref_proof_size(uint64_t height, uint64_t width)
{
    MerkleTreeGL t;
    t.height = height;
    t.width = width;
    return width + t.MerkleProofSize() * HASH_SIZE;
}

void ref_group_proof(uint64_t *proof, uint64_t *source, uint64_t *nodes, uint64_t height, uint64_t width, uint64_t idx)
{
    MerkleTreeGL t;
    t.height = height;
    t.width = width;
    t.source = (E *)source;
    t.nodes = (E *)nodes;
    t.getGroupProof((E *)proof, idx);
}

// extendPol followed by MerkleTreeGL::merkelize, i.e. the stage-1 commit of
// Starks::genProof (src/starkpil/starks.cpp:52-56).  Returns the two timings in ms.
void ref_commit(uint64_t *lde, uint64_t *in, uint64_t *tree, uint64_t n_ext, uint64_t n, uint64_t ncols,
                uint64_t *buffer, double *ms_lde, double *ms_tree)
{
    double t0 = omp_get_wtime();
    {
        NTT_Goldilocks ntt(n);
        ntt.extendPol((E *)lde, (E *)in, n_ext, n, ncols, (E *)buffer);
    }
    double t1 = omp_get_wtime();
    MerkleTreeGL t;
    t.height = n_ext;
    t.width = ncols;
    t.source = (E *)lde;
    t.nodes = (E *)tree;
    t.merkelize();
    double t2 = omp_get_wtime();
    if (ms_lde) *ms_lde = (t1 - t0) * 1e3;
    if (ms_tree) *ms_tree = (t2 - t1) * 1e3;
}

} // extern "C"
