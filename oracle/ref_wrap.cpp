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
#include "goldilocks_cubic_extension.hpp"
#include <vector>
#include <omp.h>

typedef Goldilocks::Element E;

extern "C" {

int ref_num_threads(void) { return omp_get_max_threads(); }
// torch.distributed.run exports OMP_NUM_THREADS=1 to its workers; the CPU arms of bench.py set the count explicitly
void ref_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }

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

// ---- next rows.  The callers below live in Starks::genProof / FRIProve::prove, which cannot be
// compiled stand-alone (StarkInfo, Transcript, zklog ...).  Their loop bodies are restated here line by line
// on top of the reference's OWN arithmetic (NTT_Goldilocks, Goldilocks::*, Goldilocks3::mul), which is what
// pins the oracle restatement in gl_oracle.c.

void ref_gl3_mul(uint64_t *r, uint64_t *a, uint64_t *b)
{
    Goldilocks3::mul((Goldilocks3::Element &)*(E *)r, (Goldilocks3::Element &)*(E *)a, (Goldilocks3::Element &)*(E *)b);
}

// src/starkpil/starks.cpp:261-284
void ref_q_extend(uint64_t *cm4, uint64_t *q_2ns, uint64_t N, uint64_t NExtended, uint64_t qDim, uint64_t qDeg)
{
    NTT_Goldilocks nttExtended(NExtended);
    std::vector<E> qq1(NExtended * qDim), qq2(NExtended * qDeg * qDim, Goldilocks::zero());
    nttExtended.INTT(qq1.data(), (E *)q_2ns, NExtended, qDim, NULL, 2, 1);
    E shiftIn = Goldilocks::exp(Goldilocks::inv(Goldilocks::shift()), N);
    E curS = Goldilocks::one();
    for (uint64_t p = 0; p < qDeg; p++) {
        for (uint64_t k = 0; k < N; ++k)
            for (uint64_t d = 0; d < qDim; d++) qq2[(k * qDeg + p) * qDim + d] = qq1[(p * N + k) * qDim + d] * curS;
        curS = Goldilocks::mul(curS, shiftIn);
    }
    nttExtended.NTT((E *)cm4, qq2.data(), NExtended, qDim * qDeg);
}

// src/starkpil/fri/friProve.cpp:36-107 (si > 0 branch), polMulAxi :192-200, evalPol :201-217
void ref_fri_fold(uint64_t *out, uint64_t *pol, unsigned polBits, unsigned nextBits, uint64_t shiftInv, uint64_t *x)
{
    uint64_t pol2N = 1ULL << nextBits, nX = (1ULL << polBits) / pol2N;
    E wi = Goldilocks::inv(Goldilocks::w(polBits));
    E sinv_ = Goldilocks::fromU64(shiftInv);
    E *fp = (E *)pol, *o = (E *)out;
    for (uint64_t g = 0; g < pol2N; g++) {
        std::vector<E> ppar(nX * 3), ppar_c(nX * 3);
        for (uint64_t i = 0; i < nX; i++)
            for (int d = 0; d < 3; d++) ppar[i * 3 + d] = fp[((i * pol2N) + g) * 3 + d];
        NTT_Goldilocks ntt(nX, 1);
        ntt.INTT(ppar_c.data(), ppar.data(), nX, 3);
        E r = Goldilocks::one();
        for (uint64_t i = 0; i < nX; i++) {
            for (int d = 0; d < 3; d++) ppar_c[i * 3 + d] = ppar_c[i * 3 + d] * r;
            r = r * sinv_;
        }
        E res[3] = {ppar_c[(nX - 1) * 3], ppar_c[(nX - 1) * 3 + 1], ppar_c[(nX - 1) * 3 + 2]};
        for (int64_t i = (int64_t)nX - 2; i >= 0; i--) {
            E aux[3];
            Goldilocks3::mul((Goldilocks3::Element &)aux[0], (Goldilocks3::Element &)res[0], (Goldilocks3::Element &)*(E *)x);
            for (int d = 0; d < 3; d++) res[d] = aux[d] + ppar_c[i * 3 + d];
        }
        for (int d = 0; d < 3; d++) o[g * 3 + d] = res[d];
        sinv_ = sinv_ * wi;
    }
}

} // extern "C"
