// poseidon_goldilocks.hpp -- drop-in for the static interface of the reference's PoseidonGoldilocks
// (src/goldilocks/src/poseidon_goldilocks.hpp:43-68) on the B200.
//
// merkletree / merkletree_avx / merkletree_avx512 / merkletree_seq all produce the same tree in the
// reference (tests.cpp:1817-2019); here they are one GPU implementation.  nThreads is an OpenMP hint and is
// ignored.  hash_full_result / hash / linear_hash are provided for completeness (Transcript, tests);
// one-at-a-time calls pay a PCIe round trip -- batch through gl_hash_full_result_host where it matters.
#ifndef POSEIDON_GOLDILOCKS
#define POSEIDON_GOLDILOCKS

#include "goldilocks_base_field.hpp"
#include "glb200_shim.hpp"

#define RATE 8
#define CAPACITY 4
#define HASH_SIZE 4
#define SPONGE_WIDTH (RATE + CAPACITY)

class PoseidonGoldilocks
{
public:
    static void merkletree(Goldilocks::Element *tree, Goldilocks::Element *input, uint64_t num_cols, uint64_t num_rows, int nThreads = 0, uint64_t dim = 1)
    {
        (void)nThreads;
        if (num_rows == 0)
            return; // poseidon_goldilocks.cpp:268-271
        glb200::check(gl_merkletree_host((uint64_t *)tree, (const uint64_t *)input, num_cols * dim, num_rows), "PoseidonGoldilocks::merkletree");
    }
    static void merkletree_avx(Goldilocks::Element *tree, Goldilocks::Element *input, uint64_t num_cols, uint64_t num_rows, int nThreads = 0, uint64_t dim = 1) { merkletree(tree, input, num_cols, num_rows, nThreads, dim); }
    static void merkletree_avx512(Goldilocks::Element *tree, Goldilocks::Element *input, uint64_t num_cols, uint64_t num_rows, int nThreads = 0, uint64_t dim = 1) { merkletree(tree, input, num_cols, num_rows, nThreads, dim); }
    static void merkletree_seq(Goldilocks::Element *tree, Goldilocks::Element *input, uint64_t num_cols, uint64_t num_rows, int nThreads = 0, uint64_t dim = 1) { merkletree(tree, input, num_cols, num_rows, nThreads, dim); }

    static void hash_full_result(Goldilocks::Element *state, const Goldilocks::Element *input)
    {
        glb200::check(gl_hash_full_result_host((uint64_t *)state, (const uint64_t *)input, 1), "PoseidonGoldilocks::hash_full_result");
    }
    static void hash_full_result_seq(Goldilocks::Element *state, const Goldilocks::Element *input) { hash_full_result(state, input); }
    static void hash(Goldilocks::Element (&state)[CAPACITY], const Goldilocks::Element (&input)[SPONGE_WIDTH])
    {
        Goldilocks::Element aux[SPONGE_WIDTH];
        hash_full_result(aux, input);
        std::memcpy(state, aux, CAPACITY * sizeof(Goldilocks::Element)); // poseidon_goldilocks_avx.hpp:8-13
    }
    static void hash_seq(Goldilocks::Element (&state)[CAPACITY], const Goldilocks::Element (&input)[SPONGE_WIDTH]) { hash(state, input); }
    static void linear_hash(Goldilocks::Element *output, Goldilocks::Element *input, uint64_t size)
    {
        glb200::check(gl_linear_hash_host((uint64_t *)output, (const uint64_t *)input, size), "PoseidonGoldilocks::linear_hash");
    }
    static void linear_hash_seq(Goldilocks::Element *output, Goldilocks::Element *input, uint64_t size) { linear_hash(output, input, size); }
};

#endif
