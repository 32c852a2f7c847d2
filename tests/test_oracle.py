"""CPU tests: pin the oracle (oracle/gl_oracle.c) against the reference's own known-answer vectors,
against outputs of the unmodified reference stored in tests/golden, and -- when oracle/_ref exists in
this checkout -- against the unmodified reference run live."""
import numpy as np
import pytest

from conftest import P, fib_rows, fib_vec, hexs


def test_field_basics(oracle):
    L = oracle.L
    assert L.glo_mul(P - 1, P - 1) == 1
    assert L.glo_add(P - 1, 1) == 0
    assert L.glo_sub(0, 1) == P - 1
    assert L.glo_canon(P) == 0 and L.glo_canon(2**64 - 1) == 2**32 - 2
    for k in range(1, 33):
        w = L.glo_root_of_unity(k)
        assert L.glo_pow(w, 1 << k) == 1 and L.glo_pow(w, 1 << (k - 1)) == P - 1
    # reference table: W[1]=-1, W[2]=2^48, W[3]=2^24, W[4]=2^12, W[5]=2^6, W[6]=2^3 (goldilocks_base_field.cpp:11-18)
    assert [L.glo_root_of_unity(k) for k in range(1, 7)] == [P - 1, 1 << 48, 1 << 24, 1 << 12, 1 << 6, 1 << 3]
    assert L.glo_root_of_unity(7) == 2198989700608
    assert L.glo_mul(L.glo_inv(12345), 12345) == 1


def test_poseidon_kats(oracle, kats):
    assert np.array_equal(oracle.hash_full_result(fib_vec(12)), kats["poseidon_fibonacci_state"])
    assert np.array_equal(oracle.hash_full_result(np.zeros(12, dtype=np.uint64)), kats["poseidon_zero_state"])


def test_linear_hash_kat(oracle, kats):
    assert np.array_equal(oracle.linear_hash(fib_vec(128)), kats["linear_hash_fibonacci_128"])


def test_linear_hash_short_inputs_are_not_hashed(oracle):
    for n in range(0, 5):
        v = np.arange(1, n + 1, dtype=np.uint64)
        want = np.zeros(4, dtype=np.uint64)
        want[:n] = v
        assert np.array_equal(oracle.linear_hash(v), want)


def test_merkle_kats(oracle, kats):
    t = oracle.merkletree(fib_rows(64, 128))
    assert np.array_equal(t[-4:], kats["merkle_root_64x128"]), hexs(t[-4:])
    t = oracle.merkletree(np.zeros(0, dtype=np.uint64), nrows=64, ncols=0)
    assert np.array_equal(t[-4:], kats["merkle_root_ncols0_64"])
    t = oracle.merkletree(np.zeros(0, dtype=np.uint64), nrows=1 << 17, ncols=0)
    assert np.array_equal(t[-4:], kats["merkle_root_ncols0_131072"])
    t = oracle.merkletree_batch(fib_rows(64, 128), 32)
    assert np.array_equal(t[-4:], kats["merkle_batch_root_64x128_b32"])


def test_lde_kats(oracle, kats):
    # tests.cpp:2385-2466: manual LDE with shift 49
    L = oracle.L
    a = fib_vec(16, 1, 1).reshape(16, 1)
    c = oracle.ntt(a, inverse=True)
    r = 1
    for i in range(16):
        c[i, 0] = L.glo_mul(int(c[i, 0]), r)
        r = L.glo_mul(r, 49)
    ext = np.zeros((32, 1), dtype=np.uint64)
    ext[:16] = c
    out = oracle.ntt(ext)
    assert np.array_equal(out[:, 0], kats["lde_shift49_16_to_32"])
    # tests.cpp:2554-2616: extendPol, 8 columns
    a = np.zeros((16, 8), dtype=np.uint64)
    for j in range(8):
        a[:, j] = fib_vec(16, 1 + j, 1 + j)
    out = oracle.extend_pol(a, 32)
    assert np.array_equal(out[:, 0], kats["extendpol_16_to_32_col0"])


def test_ntt_roundtrip_and_edge_sizes(oracle):
    rng = np.random.default_rng(7)
    for n, c in [(1, 1), (1, 5), (2, 1), (2, 3), (4, 2), (64, 8), (1024, 3)]:
        a = rng.integers(0, P, size=(n, c), dtype=np.uint64)
        assert np.array_equal(oracle.ntt(oracle.ntt(a), inverse=True), a)
    # size 1 is the identity, also for extend (reference tests.cpp:2339-2384 covers sizes 1, 2)
    a = rng.integers(0, P, size=(1, 4), dtype=np.uint64)
    assert np.array_equal(oracle.ntt(a), a) and np.array_equal(oracle.ntt(a, inverse=True, extend=True), a)
    # non-canonical input words are accepted
    b = np.array([[P], [P + 5]], dtype=np.uint64)
    assert np.array_equal(oracle.ntt(b), oracle.ntt(np.array([[0], [5]], dtype=np.uint64)))


def test_oracle_matches_reference_fixtures(oracle, fixtures):
    fin, fout, meta = fixtures["in"], fixtures["out"], fixtures["meta"]
    for (n, c) in meta["ntt_shapes"]:
        k = "%d_%d" % (n, c)
        assert np.array_equal(oracle.ntt(fin["ntt_" + k]), fout["ntt_" + k]), k
        assert np.array_equal(oracle.ntt(fin["ntt_" + k], inverse=True), fout["intt_" + k]), k
    for (n, e, c) in meta["lde"]:
        k = "lde_%d_%d_%d" % (n, e, c)
        assert np.array_equal(oracle.extend_pol(fin[k], e), fout[k]), k
    for (h, w) in meta["tree"]:
        k = "tree_%d_%d" % (h, w)
        t = oracle.merkletree(fin[k], nrows=h, ncols=w)
        assert np.array_equal(t, fout[k]), k
    for (h, w, idx) in meta["proof"]:
        src = fin["tree_%d_%d" % (h, w)]
        t = fout["tree_%d_%d" % (h, w)]
        assert np.array_equal(oracle.group_proof(src, t, idx), fout["proof_%d_%d_%d" % (h, w, idx)])
        assert oracle.proof_size(h, w) == fout["proof_%d_%d_%d" % (h, w, idx)].size


def test_config1_root(oracle, kats):
    """BASELINE.json configs[0]: 2^16 x 64, blowup 2 -- bit-exact root vs the reference's CPU AVX path."""
    x, _ = oracle.xorshift((1 << 16) * 64)
    lde = oracle.extend_pol(x.reshape(1 << 16, 64), 1 << 17)
    tree = oracle.merkletree(lde)
    assert np.array_equal(tree[-4:], kats["config1_root_2p16x64_blowup2"]), hexs(tree[-4:])


def test_oracle_vs_live_reference(oracle, ref):
    if ref is None:
        pytest.skip("oracle/_ref not built in this checkout (no /root/reference)")
    rng = np.random.default_rng(99)
    a = rng.integers(0, 2**64 - 1, size=(512, 11), dtype=np.uint64)  # includes non-canonical words
    assert np.array_equal(oracle.ntt(a), ref.ntt(a))
    assert np.array_equal(oracle.ntt(a, inverse=True), ref.ntt(a, inverse=True))
    assert np.array_equal(oracle.extend_pol(a, 2048), ref.extend_pol(a, 2048))
    for variant in (0, 1):
        assert np.array_equal(oracle.merkletree(a), ref.merkletree(a, variant=variant))
    assert np.array_equal(oracle.merkletree_batch(a, 4), ref.merkletree_batch(a, 4))
    for n in (0, 1, 4, 5, 8, 9, 16, 23):
        v = rng.integers(0, P, size=n, dtype=np.uint64)
        assert np.array_equal(oracle.linear_hash(v), ref.linear_hash(v)), n
    # nphase / nblock only change the reference's blocking, never the result (tests.cpp:2253-2384)
    for nphase, nblock in [(1, 1), (2, 3), (4, 11), (9, 2)]:
        assert np.array_equal(ref.ntt(a, nphase=nphase, nblock=nblock), oracle.ntt(a))


def test_next_rows_match_reference_fixtures(oracle, fixtures):
    """Q-polynomial extension (starks.cpp:261-284) and FRI folding (friProve.cpp:36-107) of the oracle against
    fixtures produced with the reference's own NTT / Goldilocks3 arithmetic (oracle/ref_wrap.cpp)."""
    for (n, e, qdeg) in fixtures["meta"]["q_extend"]:
        k = "qext_%d_%d_%d" % (n, e, qdeg)
        assert np.array_equal(oracle.q_extend(fixtures["in"][k], n, qdeg), fixtures["out"][k]), k
    for (pb, nb) in fixtures["meta"]["fri_fold"]:
        k = "frifold_%d_%d" % (pb, nb)
        pol, x, si = fixtures["in"][k]
        assert np.array_equal(oracle.fri_fold(pol, nb, si, x), fixtures["out"][k]), k


def test_fri_transpose_is_the_reference_index_map(oracle):
    # friProve.cpp:252-270: aux[i*h + j] = pol[j*w + i]
    pol = np.arange(3 * 64, dtype=np.uint64).reshape(64, 3)
    aux = oracle.fri_transpose(pol, 4)
    w, h = 16, 4
    for i in range(w):
        for j in range(h):
            assert np.array_equal(aux[i * h + j], pol[j * w + i])


def test_bctree_interpolate_pins_the_oracle(oracle, fixtures):
    """interpolate() of the reference's own bctree tool (tools/starkpil/bctree/build_const_tree.cpp:198-330,
    a second LDE implementation: own block FFT, interpolatePrepare with shift^i / N) -- stored outputs, and run live when
    oracle/_ref/libglref_bctree.so exists here -- equals the oracle's extendPol restatement word for word."""
    from oracle import pyoracle
    live = pyoracle.ref_bctree()
    for (n, e, c) in fixtures["meta"]["bctree"]:
        k = "bctree_%d_%d_%d" % (n, e, c)
        x = fixtures["in"][k]
        assert np.array_equal(oracle.extend_pol(x, e), fixtures["out"][k]), k
        if live is not None:
            assert np.array_equal(live.interpolate(x, e), fixtures["out"][k]), k
