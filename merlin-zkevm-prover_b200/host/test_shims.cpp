// test_shims.cpp -- the reference's own commit-path tests, re-hosted on the drop-in classes
// (no gtest in this image).  Mirrors src/goldilocks/tests/tests.cpp: merkletree_seq (:1817-1882),
// ntt / ntt_block (:2225-2384), extendePol (:2554-2616), plus the stage-1 sequence of
// Starks::genProof (starks.cpp:52-56) through MerkleTreeGL.  Exit code 0 = all passed.
#include "merkleTreeGL.hpp"
#include "ntt_goldilocks.hpp"
#include <cstdio>
#include <thread>
#include <vector>

static const uint64_t P = 0xFFFFFFFF00000001ULL;
static int failures = 0;
#define EXPECT_EQ(a, b) do { uint64_t a_ = (a), b_ = (b); if (a_ != b_) { printf("FAIL %s:%d %s = %016lX, want %016lX\n", __FILE__, __LINE__, #a, a_, b_); failures++; } } while (0)

static Goldilocks::Element addm(Goldilocks::Element a, Goldilocks::Element b)
{
    unsigned __int128 s = (unsigned __int128)Goldilocks::toU64(a) + Goldilocks::toU64(b);
    return Goldilocks::fromU64((uint64_t)(s % P));
}

static void test_merkletree()
{
    uint64_t ncols = 128, nrows = 1 << 6;
    std::vector<Goldilocks::Element> cols(ncols * nrows);
    for (uint64_t i = 0; i < ncols; i++) cols[i] = cols[i + ncols] = Goldilocks::fromU64(i + 1);
    for (uint64_t j = 2; j < nrows; j++)
        for (uint64_t i = 0; i < ncols; i++) cols[j * ncols + i] = addm(cols[(j - 2) * ncols + i], cols[(j - 1) * ncols + i]);
    MerkleTreeGL mt(nrows, ncols, cols.data());
    mt.merkelize();
    Goldilocks::Element root[4];
    mt.getRoot(root);
    EXPECT_EQ(Goldilocks::toU64(root[0]), 0X918F7CD0C3E8701F);
    EXPECT_EQ(Goldilocks::toU64(root[1]), 0X83A130E00F961B02);
    EXPECT_EQ(Goldilocks::toU64(root[2]), 0X6921497B364123F8);
    EXPECT_EQ(Goldilocks::toU64(root[3]), 0XBD2B98A57B748BF4);
    // proof = row || siblings; recompute nothing, just check sizes and that sibling 0 is the neighbour leaf
    std::vector<Goldilocks::Element> proof(ncols + mt.MerkleProofSize() * HASH_SIZE);
    mt.getGroupProof(proof.data(), 5);
    EXPECT_EQ(mt.MerkleProofSize(), 6);
    for (int k = 0; k < 4; k++) EXPECT_EQ(Goldilocks::toU64(proof[ncols + k]), Goldilocks::toU64(mt.nodes[4 * 4 + k]));
    EXPECT_EQ(Goldilocks::toU64(proof[3]), Goldilocks::toU64(cols[5 * ncols + 3]));
    // edge: ncols = 0 (tests.cpp:1851-1865)
    MerkleTreeGL mt0(nrows, 0, NULL);
    mt0.merkelize();
    mt0.getRoot(root);
    EXPECT_EQ(Goldilocks::toU64(root[0]), 0X25225F1A5D49614A);
    EXPECT_EQ(Goldilocks::toU64(root[3]), 0X49561260080D30C3);
}

static void test_ntt_roundtrip()
{
    const uint64_t n = 1 << 4, ncols = 8;
    std::vector<Goldilocks::Element> a(n * ncols), init(n * ncols), b(n * ncols);
    for (uint64_t j = 0; j < ncols; j++) a[j] = a[ncols + j] = Goldilocks::fromU64(1 + j);
    for (uint64_t i = 2; i < n; i++)
        for (uint64_t j = 0; j < ncols; j++) a[i * ncols + j] = addm(a[(i - 1) * ncols + j], a[(i - 2) * ncols + j]);
    init = a;
    NTT_Goldilocks gntt(n);
    for (int rep = 0; rep < 3; rep++) { // dst == src
        gntt.NTT(a.data(), a.data(), n, ncols);
        gntt.INTT(a.data(), a.data(), n, ncols);
    }
    for (uint64_t i = 0; i < n * ncols; i++) EXPECT_EQ(Goldilocks::toU64(a[i]), Goldilocks::toU64(init[i]));
    gntt.NTT(b.data(), a.data(), n, ncols, NULL, 4, 3); // dst != src, odd nphase/nblock are only hints
    for (uint64_t i = 0; i < n * ncols; i++) EXPECT_EQ(Goldilocks::toU64(a[i]), Goldilocks::toU64(init[i]));
    gntt.INTT(NULL, b.data(), n, ncols);                // dst == NULL: in place
    for (uint64_t i = 0; i < n * ncols; i++) EXPECT_EQ(Goldilocks::toU64(b[i]), Goldilocks::toU64(init[i]));
    gntt.NTT(a.data(), a.data(), 0, ncols);             // size 0 / ncols 0: no-op
    gntt.NTT(a.data(), a.data(), n, 0);
    // root(domainPow, idx) = w(domainPow)^idx: w(1) = -1, w(4) = 2^12, w(6) = 2^3 (goldilocks_base_field.cpp:11-18)
    EXPECT_EQ(Goldilocks::toU64(gntt.root(1, 1)), P - 1);
    EXPECT_EQ(Goldilocks::toU64(gntt.root(4, 1)), 1ULL << 12);
    EXPECT_EQ(Goldilocks::toU64(gntt.root(6, 5)), 1ULL << 15);
    EXPECT_EQ(Goldilocks::toU64(gntt.root(4, 16)), 1);
    EXPECT_EQ(Goldilocks::toU64(gntt.root(32, 1)), 7277203076849721926ULL);
    bool thrown = false;
    try { NTT_Goldilocks big((uint64_t)1 << 33); } catch (const std::range_error &) { thrown = true; }
    EXPECT_EQ(thrown ? 1 : 0, 1);
}

static void test_extendPol_and_stage_commit()
{
    const uint64_t N = 1 << 4, NE = 1 << 5, ncols = 8;
    std::vector<Goldilocks::Element> a(NE * ncols), scratch(NE * ncols);
    for (uint64_t j = 0; j < ncols; j++) a[j] = a[ncols + j] = Goldilocks::fromU64(1 + j);
    for (uint64_t i = 2; i < N; i++)
        for (uint64_t j = 0; j < ncols; j++) a[i * ncols + j] = addm(a[(i - 1) * ncols + j], a[(i - 2) * ncols + j]);
    NTT_Goldilocks ntt(N);
    ntt.extendPol(a.data(), a.data(), NE, N, ncols, scratch.data()); // input == output, as tests.cpp:2578
    const uint64_t want[32] = {
        0X3E7CA26D67147C31, 0X1310720153E0ABE4, 0X20446D2EA50E8F96};
    for (int k = 0; k < 3; k++) EXPECT_EQ(Goldilocks::toU64(a[k * ncols]), want[k]);
    // stage commit as Starks::genProof does it (starks.cpp:52-56): tree over the extended matrix
    MerkleTreeGL tree(NE, ncols, a.data());
    tree.merkelize();
    Goldilocks::Element r1[4], r2[4];
    tree.getRoot(r1);
    std::vector<Goldilocks::Element> nodes(tree.getTreeNumElements());
    PoseidonGoldilocks::merkletree_seq(nodes.data(), a.data(), ncols, NE);
    std::memcpy(r2, &nodes[nodes.size() - 4], sizeof r2);
    for (int k = 0; k < 4; k++) EXPECT_EQ(Goldilocks::toU64(r1[k]), Goldilocks::toU64(r2[k]));
}

// FRIProve::prove builds NTT_Goldilocks(nX, 1) and calls INTT on (nX x 3) groups from inside an OpenMP parallel region
// (friProve.cpp:44-107): the drop-in must give every thread its own correct answer.  Each thread transforms a
// different matrix forth and back; a forward transform of a constant column must be (n*c, 0, 0, ...).
static void test_concurrent_callers()
{
    const int nthreads = 8, reps = 6;
    const uint64_t nX = 1 << 5, ncols = 3;
    std::vector<int> bad(nthreads, 0);
    std::vector<std::thread> pool;
    for (int t = 0; t < nthreads; t++)
        pool.emplace_back([&, t]() {
            try
            {
                for (int rep = 0; rep < reps; rep++)
                {
                    NTT_Goldilocks ntt(nX, 1);
                    std::vector<Goldilocks::Element> a(nX * ncols), init;
                    for (uint64_t i = 0; i < nX * ncols; i++) a[i] = Goldilocks::fromU64((0x9E3779B97F4A7C15ULL * (i + 1 + 1000 * t + 77 * rep)) % P);
                    init = a;
                    ntt.INTT(a.data(), a.data(), nX, ncols);
                    ntt.NTT(a.data(), a.data(), nX, ncols);
                    for (uint64_t i = 0; i < nX * ncols; i++)
                        if (Goldilocks::toU64(a[i]) != Goldilocks::toU64(init[i])) bad[t]++;
                    std::vector<Goldilocks::Element> c(nX * ncols, Goldilocks::fromU64(t + 1));
                    ntt.NTT(c.data(), c.data(), nX, ncols);
                    for (uint64_t i = 0; i < nX * ncols; i++)
                        if (Goldilocks::toU64(c[i]) != (i < ncols ? nX * (t + 1) : 0)) bad[t]++;
                }
            }
            catch (const std::exception &e)
            {
                printf("FAIL thread %d: %s\n", t, e.what());
                bad[t]++;
            }
        });
    for (auto &th : pool) th.join();
    for (int t = 0; t < nthreads; t++) EXPECT_EQ((uint64_t)bad[t], 0);
}

// merkletree_batch KAT (tests.cpp:2021-2052) and the two-state AVX-512 entry points against the one-state ones.
static void test_batch_and_avx512_entry_points()
{
    uint64_t ncols = 128, nrows = 1 << 6;
    std::vector<Goldilocks::Element> cols(ncols * nrows);
    for (uint64_t i = 0; i < ncols; i++) cols[i] = cols[i + ncols] = Goldilocks::fromU64(i + 1);
    for (uint64_t j = 2; j < nrows; j++)
        for (uint64_t i = 0; i < ncols; i++) cols[j * ncols + i] = addm(cols[(j - 2) * ncols + i], cols[(j - 1) * ncols + i]);
    std::vector<Goldilocks::Element> tree((2 * nrows - 1) * 4);
    PoseidonGoldilocks::merkletree_batch_avx(tree.data(), cols.data(), ncols, nrows, 32);
    EXPECT_EQ(Goldilocks::toU64(tree[tree.size() - 4]), 0XB2597514367E69FD);
    EXPECT_EQ(Goldilocks::toU64(tree[tree.size() - 1]), 0X3E8670A179011526);

    Goldilocks::Element a[12], b[12], ra[12], rb[12], two[24], out[24];
    for (int w = 0; w < 12; w++) { a[w] = Goldilocks::fromU64(w * 3 + 1); b[w] = Goldilocks::fromU64(1000003ULL * (w + 7)); }
    for (int w = 0; w < 12; w++) { two[8 * (w / 4) + w % 4] = a[w]; two[8 * (w / 4) + 4 + w % 4] = b[w]; }
    PoseidonGoldilocks::hash_full_result(ra, a);
    PoseidonGoldilocks::hash_full_result(rb, b);
    PoseidonGoldilocks::hash_full_result_avx512(out, two);
    for (int w = 0; w < 12; w++)
    {
        EXPECT_EQ(Goldilocks::toU64(out[8 * (w / 4) + w % 4]), Goldilocks::toU64(ra[w]));
        EXPECT_EQ(Goldilocks::toU64(out[8 * (w / 4) + 4 + w % 4]), Goldilocks::toU64(rb[w]));
    }
    Goldilocks::Element d2[8], da[4], db[4];
    PoseidonGoldilocks::linear_hash_avx512(d2, cols.data(), ncols); // rows 0 and 1
    PoseidonGoldilocks::linear_hash(da, cols.data(), ncols);
    PoseidonGoldilocks::linear_hash(db, cols.data() + ncols, ncols);
    for (int k = 0; k < 4; k++) { EXPECT_EQ(Goldilocks::toU64(d2[k]), Goldilocks::toU64(da[k])); EXPECT_EQ(Goldilocks::toU64(d2[4 + k]), Goldilocks::toU64(db[k])); }
}

// The literal stage-1 sequence of Starks::genProof (starks.cpp:52-59) with the resident stage switched on: extendPol keeps the
// extended matrix and the tree it hashed during the download on the device, merkelize() only fetches the nodes, and the queries
// (friProve.cpp:219-250) can be served from the device copies.  Everything must equal the non-resident path.
static void test_resident_stage()
{
    const uint64_t N = 1 << 10, NE = 1 << 11, ncols = 20;
    std::vector<Goldilocks::Element> in(N * ncols), out(NE * ncols), out2(NE * ncols);
    for (uint64_t i = 0; i < N * ncols; i++) in[i] = Goldilocks::fromU64((0x9E3779B97F4A7C15ULL * (i + 12345)) % P);
    NTT_Goldilocks ntt(N);
    EXPECT_EQ((uint64_t)gl_set_resident(1), 0);
    ntt.extendPol(out.data(), in.data(), NE, N, ncols);
    uint64_t *dl = nullptr, *dt = nullptr;
    EXPECT_EQ((uint64_t)gl_resident_query((const uint64_t *)out.data(), ncols, NE, &dl, &dt), 0);
    MerkleTreeGL tree(NE, ncols, out.data());
    tree.merkelize(); // served from the resident tree
    // queries from the device copies, all in one call
    const uint64_t idx[5] = {0, 1, 777, NE / 2, NE - 1};
    const uint64_t psize = ncols + tree.MerkleProofSize() * HASH_SIZE;
    std::vector<Goldilocks::Element> proofs(5 * psize), one(psize);
    EXPECT_EQ((uint64_t)gl_group_proofs_dev((uint64_t *)proofs.data(), dl, dt, NE, ncols, idx, 5), 0);
    for (int q = 0; q < 5; q++) {
        tree.getGroupProof(one.data(), idx[q]);
        for (uint64_t k = 0; k < psize; k++) EXPECT_EQ(Goldilocks::toU64(proofs[q * psize + k]), Goldilocks::toU64(one[k]));
    }
    // the same without the resident stage
    EXPECT_EQ((uint64_t)gl_set_resident(0), 0);
    EXPECT_EQ((uint64_t)(gl_resident_query((const uint64_t *)out.data(), ncols, NE, &dl, &dt) != 0), 1);
    ntt.extendPol(out2.data(), in.data(), NE, N, ncols);
    for (uint64_t i = 0; i < NE * ncols; i++) EXPECT_EQ(Goldilocks::toU64(out[i]), Goldilocks::toU64(out2[i]));
    MerkleTreeGL tree2(NE, ncols, out2.data());
    tree2.merkelize();
    for (uint64_t i = 0; i < tree.getTreeNumElements(); i++) EXPECT_EQ(Goldilocks::toU64(tree.nodes[i]), Goldilocks::toU64(tree2.nodes[i]));
    // a host buffer that changed after extendPol must not be served from the stale device copy: resident on, overwrite, merkelize
    EXPECT_EQ((uint64_t)gl_set_resident(1), 0);
    ntt.extendPol(out.data(), in.data(), NE, N, ncols);
    for (uint64_t i = 0; i < NE * ncols; i++) out[i] = Goldilocks::fromU64(Goldilocks::toU64(out[i]) ^ 1);
    MerkleTreeGL tree3(NE, ncols, out.data());
    tree3.merkelize();
    std::vector<Goldilocks::Element> nodes(tree3.getTreeNumElements());
    PoseidonGoldilocks::merkletree_seq(nodes.data(), out.data(), ncols, NE);
    for (int k = 0; k < 4; k++) EXPECT_EQ(Goldilocks::toU64(tree3.nodes[nodes.size() - 4 + k]), Goldilocks::toU64(nodes[nodes.size() - 4 + k]));
    gl_set_resident(0);
}

int main()
{
    try
    {
        test_merkletree();
        test_ntt_roundtrip();
        test_extendPol_and_stage_commit();
        test_concurrent_callers();
        test_batch_and_avx512_entry_points();
        test_resident_stage();
    }
    catch (const std::exception &e)
    {
        printf("FAIL exception: %s\n", e.what());
        return 2;
    }
    printf(failures ? "%d FAILURES\n" : "shim tests passed\n", failures);
    return failures ? 1 : 0;
}
