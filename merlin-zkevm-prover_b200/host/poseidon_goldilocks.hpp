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

    // merkletree_batch family (poseidon_goldilocks.hpp:44,51,60,68; .cpp:113-164,302-352,521-577): rows hashed in slices of
    // batch_size columns, slice digests hashed again.  Not called by Starks / MerkleTreeGL; kept for the reference's tests.
    static void merkletree_batch(Goldilocks::Element *tree, Goldilocks::Element *input, uint64_t num_cols, uint64_t num_rows, uint64_t batch_size, int nThreads = 0, uint64_t dim = 1)
    {
        (void)nThreads;
        if (num_rows == 0)
            return;
        glb200::check(gl_merkletree_batch_host((uint64_t *)tree, (const uint64_t *)input, num_cols * dim, num_rows, batch_size * dim), "PoseidonGoldilocks::merkletree_batch");
    }
    static void merkletree_batch_seq(Goldilocks::Element *tree, Goldilocks::Element *input, uint64_t num_cols, uint64_t num_rows, uint64_t batch_size, int nThreads = 0, uint64_t dim = 1) { merkletree_batch(tree, input, num_cols, num_rows, batch_size, nThreads, dim); }
    static void merkletree_batch_avx(Goldilocks::Element *tree, Goldilocks::Element *input, uint64_t num_cols, uint64_t num_rows, uint64_t batch_size, int nThreads = 0, uint64_t dim = 1) { merkletree_batch(tree, input, num_cols, num_rows, batch_size, nThreads, dim); }
    static void merkletree_batch_avx512(Goldilocks::Element *tree, Goldilocks::Element *input, uint64_t num_cols, uint64_t num_rows, uint64_t batch_size, int nThreads = 0, uint64_t dim = 1) { merkletree_batch(tree, input, num_cols, num_rows, batch_size, nThreads, dim); }

    // The AVX-512 entry points work on TWO states / rows at once (poseidon_goldilocks.cpp:355-483).  Layout of the 24-word
    // buffer: 4-word groups alternate between state A and state B -- word w of A at 8*(w/4) + w%4, of B at 8*(w/4) + 4 + w%4
    // (see how linear_hash_avx512 fills it, :457-468).
    static void hash_full_result_avx512(Goldilocks::Element *state, const Goldilocks::Element *input)
    {
        Goldilocks::Element two[2 * SPONGE_WIDTH];
        for (int w = 0; w < SPONGE_WIDTH; w++)
        {
            two[w] = input[8 * (w / 4) + w % 4];
            two[SPONGE_WIDTH + w] = input[8 * (w / 4) + 4 + w % 4];
        }
        glb200::check(gl_hash_full_result_host((uint64_t *)two, (const uint64_t *)two, 2), "PoseidonGoldilocks::hash_full_result_avx512");
        for (int w = 0; w < SPONGE_WIDTH; w++)
        {
            state[8 * (w / 4) + w % 4] = two[w];
            state[8 * (w / 4) + 4 + w % 4] = two[SPONGE_WIDTH + w];
        }
    }
    static void hash_avx512(Goldilocks::Element (&state)[2 * CAPACITY], const Goldilocks::Element (&input)[2 * SPONGE_WIDTH])
    {
        Goldilocks::Element aux[2 * SPONGE_WIDTH];
        hash_full_result_avx512(aux, input);
        std::memcpy(state, aux, 2 * CAPACITY * sizeof(Goldilocks::Element)); // digest A | digest B
    }
    // input = row A (size words) followed by row B; output = digest A | digest B
    static void linear_hash_avx512(Goldilocks::Element *output, Goldilocks::Element *input, uint64_t size)
    {
        linear_hash(output, input, size);
        linear_hash(output + CAPACITY, input + size, size);
    }
};

#endif
