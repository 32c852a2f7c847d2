"""GPU parity tests: the CUDA path, called through the C ABI (libglb200.so), against the reference's
golden vectors, the stored reference outputs and the CPU oracle on identical seeded inputs.
Bar: bit-exact on canonical values for every LDE word, every Merkle node and the root."""
import numpy as np
import pytest

from conftest import P, fib_rows, fib_vec, hexs

pytestmark = pytest.mark.gpu

EDGE = [0, 1, 2, 2**32 - 2, 2**32 - 1, 2**32, 2**32 + 1, P - 2, P - 1, P, P + 1, 2**63, 2**64 - 2**32, 2**64 - 2, 2**64 - 1,
        0xFFFFFFFE00000001, 0xFFFFFFFF00000000, 0x00000001FFFFFFFF]


def _pairs(rng, n):
    e = np.array(EDGE, dtype=np.uint64)
    a = np.concatenate([np.repeat(e, len(e)), rng.integers(0, 2**64 - 1, size=n, dtype=np.uint64, endpoint=True)])
    b = np.concatenate([np.tile(e, len(e)), rng.integers(0, 2**64 - 1, size=n, dtype=np.uint64, endpoint=True)])
    return a, b


def test_field_ops(glb):
    """F1: add/sub/mul on arbitrary (also non-canonical) representatives vs exact integer arithmetic."""
    rng = np.random.default_rng(1)
    a, b = _pairs(rng, 20000)
    c = np.roll(a, 7)
    ai, bi, ci = [int(x) for x in a], [int(x) for x in b], [int(x) for x in c]
    want = {
        "add": [(x + y) % P for x, y in zip(ai, bi)],
        "sub": [(x - y) % P for x, y in zip(ai, bi)],
        "mul": [(x * y) % P for x, y in zip(ai, bi)],
        "mad": [(x * y + z) % P for x, y, z in zip(ai, bi, ci)],
        "add_c": [(x + y) % P for x, y in zip(ai, bi)],
        "sqr": [(x * x) % P for x in ai],
        "canon": [x % P for x in ai],
        # the IMAD.WIDE-free butterfly forms of the NTT passes (gl64.cuh canon_sel / add_cx / sub_cx)
        "add_cx": [(x + y) % P for x, y in zip(ai, bi)],
        "sub_cx": [(x - y) % P for x, y in zip(ai, bi)],
        "canon_sel": [x % P for x in ai],
        "sqr3": [(x * x) % P for x in ai],
    }
    for op, w in want.items():
        got = glb.field_op_host(op, a, b, c)
        assert np.array_equal(got, np.array(w, dtype=np.uint64)), op


def test_poseidon_kats(glb, oracle, kats):
    inp = np.stack([fib_vec(12), np.zeros(12, dtype=np.uint64)])
    out = glb.hash_full_result_host(inp)
    assert np.array_equal(out[0], kats["poseidon_fibonacci_state"]), hexs(out[0])
    assert np.array_equal(out[1], kats["poseidon_zero_state"]), hexs(out[1])
    rng = np.random.default_rng(2)
    x = rng.integers(0, 2**64 - 1, size=(1000, 12), dtype=np.uint64, endpoint=True)  # non-canonical words too
    x[0, :] = P - 1
    x[1, :] = 2**64 - 1
    out = glb.hash_full_result_host(x)
    for i in range(0, 1000, 37):
        assert np.array_equal(out[i], oracle.hash_full_result(x[i])), i
    for i in (0, 1):
        assert np.array_equal(out[i], oracle.hash_full_result(x[i])), i


def test_linear_hash(glb, oracle, kats):
    assert np.array_equal(glb.linear_hash_host(fib_vec(128)), kats["linear_hash_fibonacci_128"])
    rng = np.random.default_rng(3)
    for n in list(range(0, 26)) + [63, 64, 65, 127, 600, 665]:
        v = rng.integers(0, P, size=n, dtype=np.uint64)
        assert np.array_equal(glb.linear_hash_host(v), oracle.linear_hash(v)), n


def test_merkletree_kats(glb, kats):
    t = glb.merkletree_host(fib_rows(64, 128))
    assert np.array_equal(t[-4:], kats["merkle_root_64x128"]), hexs(t[-4:])
    t = glb.merkletree_host(np.zeros(0, dtype=np.uint64), nrows=64, ncols=0)
    assert np.array_equal(t[-4:], kats["merkle_root_ncols0_64"])
    t = glb.merkletree_host(np.zeros(0, dtype=np.uint64), nrows=1 << 17, ncols=0)
    assert np.array_equal(t[-4:], kats["merkle_root_ncols0_131072"])


@pytest.mark.gpu
def test_merkletree_batch(glb, oracle, kats):
    """PoseidonGoldilocks::merkletree_batch_* (reference KAT tests.cpp:2049-2052) and every node vs the oracle."""
    t = glb.merkletree_batch_host(fib_rows(64, 128), 32)
    assert np.array_equal(t[-4:], kats["merkle_batch_root_64x128_b32"]), hexs(t[-4:])
    rng = np.random.default_rng(11)
    for (h, w, b) in [(16, 37, 8), (8, 9, 4), (32, 64, 64), (4, 3, 2), (2, 100, 7), (1, 5, 3), (64, 0, 4)]:
        x = rng.integers(0, P, size=(h, w), dtype=np.uint64)
        assert np.array_equal(glb.merkletree_batch_host(x, b), oracle.merkletree_batch(x, b)), (h, w, b)


def test_merkletree_every_node(glb, oracle, fixtures):
    fin, fout, meta = fixtures["in"], fixtures["out"], fixtures["meta"]
    for (h, w) in meta["tree"]:
        k = "tree_%d_%d" % (h, w)
        t = glb.merkletree_host(fin[k], nrows=h, ncols=w)
        assert np.array_equal(t, fout[k]), k
    rng = np.random.default_rng(4)
    for (h, w) in [(2048, 13), (4096, 600), (1 << 15, 8), (512, 665)]:
        a = rng.integers(0, 2**64 - 1, size=(h, w), dtype=np.uint64, endpoint=True)
        assert np.array_equal(glb.merkletree_host(a), oracle.merkletree(a)), (h, w)


def test_group_proofs(glb, oracle, fixtures):
    fin, fout, meta = fixtures["in"], fixtures["out"], fixtures["meta"]
    for (h, w, idx) in meta["proof"]:
        src = fin["tree_%d_%d" % (h, w)]
        ds = glb.DeviceBuffer(src.size).upload(src)
        dt = glb.DeviceBuffer(glb.tree_num_elements(h))
        glb.merkletree_dev(dt.ptr, ds.ptr, w, h)
        got = glb.group_proof_dev(ds.ptr, dt.ptr, h, w, idx)
        assert np.array_equal(got, fout["proof_%d_%d_%d" % (h, w, idx)]), (h, w, idx)


def test_ntt_fixtures(glb, fixtures):
    fin, fout, meta = fixtures["in"], fixtures["out"], fixtures["meta"]
    for (n, c) in meta["ntt_shapes"]:
        k = "%d_%d" % (n, c)
        assert np.array_equal(glb.ntt_host(fin["ntt_" + k]), fout["ntt_" + k]), k
        assert np.array_equal(glb.ntt_host(fin["ntt_" + k], inverse=True), fout["intt_" + k]), k


@pytest.mark.parametrize("logn", list(range(0, 15)) + [16, 17])
def test_ntt_vs_oracle_all_sizes(glb, oracle, logn):
    rng = np.random.default_rng(100 + logn)
    n = 1 << logn
    for c in ([1, 7, 32, 45] if logn <= 12 else [33]):
        a = rng.integers(0, 2**64 - 1, size=(n, c), dtype=np.uint64, endpoint=True)
        f = glb.ntt_host(a)
        assert np.array_equal(f, oracle.ntt(a)), (n, c, "fwd")
        assert np.array_equal(glb.ntt_host(a, inverse=True), oracle.ntt(a, inverse=True)), (n, c, "inv")
        assert np.array_equal(glb.ntt_host(a, inverse=True, extend=True), oracle.ntt(a, inverse=True, extend=True)), (n, c, "ext")
        # NTT o INTT = id (reference tests.cpp:2225-2384), on canonical values
        assert np.array_equal(glb.ntt_host(f, inverse=True), a % np.uint64(P)), (n, c, "roundtrip")


@pytest.mark.parametrize("logn,c", [(10, 17), (10, 24), (14, 55), (16, 88), (15, 20)])
def test_ntt_ragged_17_to_24_columns(glb, oracle, logn, c):
    """A ragged last column tile of 17..24 columns runs as a 16-column launch plus a <= 8-column launch (two / four row-tiles per
    warp) in the passes of >= 7 bits; every word against the oracle, all three transforms and an LDE."""
    rng = np.random.default_rng(300 + logn + c)
    n = 1 << logn
    a = rng.integers(0, 2**64 - 1, size=(n, c), dtype=np.uint64, endpoint=True)
    assert np.array_equal(glb.ntt_host(a), oracle.ntt(a)), "fwd"
    assert np.array_equal(glb.ntt_host(a, inverse=True), oracle.ntt(a, inverse=True)), "inv"
    assert np.array_equal(glb.ntt_host(a, inverse=True, extend=True), oracle.ntt(a, inverse=True, extend=True)), "ext"
    if logn <= 14:
        assert np.array_equal(glb.extend_pol_host(a, 4 * n), oracle.extend_pol(a, 4 * n)), "lde"


def test_ntt_empty_is_noop(glb):
    a = np.zeros((0, 4), dtype=np.uint64)
    assert glb.ntt_host(a).shape == (0, 4)
    assert glb.ntt_host(np.zeros((8, 0), dtype=np.uint64)).shape == (8, 0)


def test_ntt_errors(glb):
    with pytest.raises(glb.GLError):
        glb.ntt_host(np.ones((12, 2), dtype=np.uint64))  # not a power of two (reference: assert, ntt_goldilocks.cpp:28)



def test_concurrent_host_callers(glb, oracle):
    """The reference calls the tiny INTTs of the FRI fold from inside an OpenMP region (friProve.cpp:44-107) and
    ctypes releases the GIL, so these threads really overlap inside the C ABI: every call must still return its own
    answer (one stream and one scratch arena are shared; the public entry points serialise)."""
    import threading
    rng = np.random.default_rng(77)
    jobs = []
    for t in range(8):
        n = 1 << int(rng.integers(1, 11))
        c = int(rng.integers(1, 40))
        a = rng.integers(0, P, size=(n, c), dtype=np.uint64)
        jobs.append((a, bool(t & 1)))
    want = [oracle.ntt(a, inverse=inv) for a, inv in jobs]
    tree_src = rng.integers(0, P, size=(256, 20), dtype=np.uint64)
    want_tree = oracle.merkletree(tree_src)
    errors = []

    def worker(k):
        try:
            a, inv = jobs[k]
            for _ in range(10):
                if not np.array_equal(glb.ntt_host(a, inverse=inv), want[k]):
                    errors.append(("ntt", k))
                if k == 0 and not np.array_equal(glb.merkletree_host(tree_src), want_tree):
                    errors.append(("tree", k))
        except Exception as ex:  # noqa: BLE001 - reported below
            errors.append((k, repr(ex)))

    threads = [threading.Thread(target=worker, args=(k,)) for k in range(len(jobs))]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors[:4]


def test_ntt_dev_aliasing(glb, oracle):
    """dst == src (in place) and dst != src (source preserved), as ntt_goldilocks.cpp:166-232."""
    rng = np.random.default_rng(5)
    n, c = 1 << 11, 40
    a = rng.integers(0, P, size=(n, c), dtype=np.uint64)
    want = oracle.ntt(a)
    s = glb.DeviceBuffer(a.size).upload(a)
    d = glb.DeviceBuffer(a.size)
    glb.ntt_dev(d.ptr, s.ptr, n, c)
    assert np.array_equal(d.download().reshape(n, c), want)
    assert np.array_equal(s.download().reshape(n, c), a)
    glb.ntt_dev(s.ptr, s.ptr, n, c)
    assert np.array_equal(s.download().reshape(n, c), want)


def test_extend_pol_kats_and_fixtures(glb, kats, fixtures):
    a = np.zeros((16, 8), dtype=np.uint64)
    for j in range(8):
        a[:, j] = fib_vec(16, 1 + j, 1 + j)
    out = glb.extend_pol_host(a, 32)
    assert np.array_equal(out[:, 0], kats["extendpol_16_to_32_col0"]), hexs(out[:, 0])
    fin, fout, meta = fixtures["in"], fixtures["out"], fixtures["meta"]
    for (n, e, c) in meta["lde"]:
        k = "lde_%d_%d_%d" % (n, e, c)
        assert np.array_equal(glb.extend_pol_host(fin[k], e), fout[k]), k


def test_extend_pol_column_blocking(glb, oracle):
    """A scratch limit forces several column blocks; the result must not change."""
    rng = np.random.default_rng(6)
    n, e, c = 1 << 10, 1 << 12, 100
    a = rng.integers(0, P, size=(n, c), dtype=np.uint64)
    want = oracle.extend_pol(a, e)
    glb.lib().gl_set_scratch_limit((n + e) * 8 * 32)
    try:
        assert np.array_equal(glb.extend_pol_host(a, e), want)
    finally:
        glb.lib().gl_set_scratch_limit(0)
    assert np.array_equal(glb.extend_pol_host(a, e), want)


def test_config1_commit_every_word(glb, oracle, kats):
    """BASELINE.json configs[0]: 2^16 x 64 blowup 2.  Every extended element, every node, the root."""
    n, e, c = 1 << 16, 1 << 17, 64
    x, _ = oracle.xorshift(n * c)
    x = x.reshape(n, c)
    lde, tree, root = glb.commit_host(x, e)
    assert np.array_equal(root, kats["config1_root_2p16x64_blowup2"]), hexs(root)
    lde_o = oracle.extend_pol(x, e)
    assert np.array_equal(lde, lde_o)
    assert np.array_equal(tree, oracle.merkletree(lde_o))
    assert np.array_equal(tree[-4:], root)


def test_commit_host_blocks_and_buffer_reuse(glb, oracle):
    """The host-pointer commit walks the columns in blocks (32, 96, 128, 128 ... columns: upload of block k+1, LDE + leaf sponge of
    block k and download of block k-1 overlap) and keeps its device buffers between calls.  A ragged width that takes five blocks,
    twice in a row (the second call reuses the cached buffers), then a smaller and a larger shape through the same buffers, then
    again after the cache was handed back (gl_set_scratch_limit)."""
    rng = np.random.default_rng(81)
    for (n, e, c) in [(1 << 9, 1 << 11, 395), (1 << 9, 1 << 11, 395), (1 << 8, 1 << 9, 70), (1 << 10, 1 << 12, 403)]:
        a = rng.integers(0, P, size=(n, c), dtype=np.uint64)
        lde, tree, root = glb.commit_host(a, e)
        lde_o = oracle.extend_pol(a, e)
        assert np.array_equal(lde, lde_o), (n, e, c)
        assert np.array_equal(tree, oracle.merkletree(lde_o)), (n, e, c)
        assert np.array_equal(root, tree[-4:])
        assert np.array_equal(glb.extend_pol_host(a, e), lde_o), (n, e, c)
    glb.lib().gl_set_scratch_limit(0)
    lde, tree, root = glb.commit_host(a, e)
    assert np.array_equal(lde, lde_o) and np.array_equal(tree, oracle.merkletree(lde_o))


def test_recursive_shape_high_blowup(glb, oracle):
    """BASELINE.json configs[3] shape at reduced size: blowup 2^4."""
    rng = np.random.default_rng(8)
    n, e, c = 1 << 12, 1 << 16, 30
    a = rng.integers(0, P, size=(n, c), dtype=np.uint64)
    lde, tree, root = glb.commit_host(a, e)
    lde_o = oracle.extend_pol(a, e)
    assert np.array_equal(lde, lde_o)
    assert np.array_equal(tree, oracle.merkletree(lde_o))


def _fib_matrix_mod_p(nrows, ncols):
    """rows 0,1 = 1..ncols, row j = row j-1 + row j-2: element = (i+1) * Fib(j) mod p, vectorised."""
    f = np.empty(nrows, dtype=np.uint64)
    x, y = 1, 1
    for j in range(nrows):
        f[j] = x
        x, y = y, (x + y) % P
    k = np.arange(1, ncols + 1, dtype=np.uint64)
    lo = (f & np.uint64(0xFFFFFFFF))[:, None] * k[None, :]          # < 2^39
    hi = (f >> np.uint64(32))[:, None] * k[None, :]                 # < 2^39, weight 2^32
    # value = lo + hi*2^32 = lo + (hi_lo << 32) + hi_hi * 2^64, and 2^64 = 2^32 - 1 (mod p)
    hi_lo = hi & np.uint64(0xFFFFFFFF)
    hi_hi = hi >> np.uint64(32)                                     # < 2^7
    t = (hi_lo << np.uint64(32))                                    # < 2^64
    u = lo + hi_hi * np.uint64(0xFFFFFFFF)                          # < 2^40
    t = np.where(t >= np.uint64(P), t - np.uint64(P), t)
    s = t + u                                                       # may wrap
    wrapped = s < t
    s = np.where(wrapped, s + np.uint64(0xFFFFFFFF), s)             # 2^64 = 2^32 - 1
    return np.where(s >= np.uint64(P), s - np.uint64(P), s)


def test_fib_matrix_helper_small():
    assert np.array_equal(_fib_matrix_mod_p(200, 128), fib_rows(200, 128))


def test_merkle_root_production_scale_kat(glb, kats):
    """The reference's own full-size assertion (bench.cpp:440-443): 2^23 rows x 128 columns."""
    nrows, ncols = 1 << 23, 128
    dsrc = glb.DeviceBuffer(nrows * ncols)
    step = 1 << 20
    f = _fib_matrix_mod_p(nrows, ncols)
    for r0 in range(0, nrows, step):
        dsrc.upload(f[r0:r0 + step], word_offset=r0 * ncols)
    del f
    dtree = glb.DeviceBuffer(glb.tree_num_elements(nrows))
    glb.merkletree_dev(dtree.ptr, dsrc.ptr, ncols, nrows)
    root = dtree.download(4, glb.tree_num_elements(nrows) - 4)
    assert np.array_equal(root, kats["merkle_root_2p23x128"]), hexs(root)


def test_full_size_lde_sampled_columns_and_linearity(glb, oracle):
    """2^22 -> 2^23 rows (size-independent properties at a large size): the LDE is column-independent,
    so a few columns are checked against the oracle run on just those columns; plus linearity."""
    n, e, c = 1 << 20, 1 << 21, 40
    rng = np.random.default_rng(9)
    a = rng.integers(0, P, size=(n, c), dtype=np.uint64)
    got = glb.extend_pol_host(a, e)
    cols = [0, 17, 39]
    want = oracle.extend_pol(np.ascontiguousarray(a[:, cols]), e)
    assert np.array_equal(got[:, cols], want)
    b = rng.integers(0, P, size=(n, c), dtype=np.uint64)
    gb = glb.extend_pol_host(b, e)
    s = glb.field_op_host("add", a.reshape(-1), b.reshape(-1)).reshape(n, c)
    gs = glb.extend_pol_host(s, e)
    assert np.array_equal(gs.reshape(-1), glb.field_op_host("add", got.reshape(-1), gb.reshape(-1)))


@pytest.mark.gpu
def test_q_polynomial_commit(glb, oracle, fixtures):
    """Starks::genProof step 4 (starks.cpp:261-290): INTT, coset un-shift + regroup, NTT, merkelize."""
    shapes = [(n, e, qdeg, fixtures["in"]["qext_%d_%d_%d" % (n, e, qdeg)]) for (n, e, qdeg) in fixtures["meta"]["q_extend"]]
    rng = np.random.default_rng(21)
    shapes.append((1 << 10, 1 << 12, 3, rng.integers(0, P, size=(1 << 12, 3), dtype=np.uint64)))
    for (n, e, qdeg, q) in shapes:
        dq = glb.DeviceBuffer(e * 3).upload(q)
        dcm4 = glb.DeviceBuffer(e * 3 * qdeg)
        dtree = glb.DeviceBuffer(glb.tree_num_elements(e))
        root = glb.q_commit_dev(dcm4.ptr, dtree.ptr, dq.ptr, n, e, 3, qdeg)
        want = oracle.q_extend(q, n, qdeg)
        key = "qext_%d_%d_%d" % (n, e, qdeg)
        if key in fixtures["out"]:
            assert np.array_equal(want, fixtures["out"][key])
        assert np.array_equal(dcm4.download().reshape(e, 3 * qdeg), want), (n, e, qdeg)
        wt = oracle.merkletree(want)
        assert np.array_equal(dtree.download(), wt)
        assert np.array_equal(root, wt[-4:])
        # host-pointer form (gl_q_commit_host): same matrix, tree and root
        import ctypes as C
        q = np.ascontiguousarray(q, dtype=np.uint64)
        hcm4, htree, hroot = np.zeros((e, 3 * qdeg), dtype=np.uint64), np.zeros(wt.size, dtype=np.uint64), np.zeros(4, dtype=np.uint64)
        rc = glb.lib().gl_q_commit_host(hcm4.ctypes.data, htree.ctypes.data, q.ctypes.data, n, e, 3, qdeg, hroot.ctypes.data_as(C.POINTER(C.c_uint64)))
        assert rc == 0 and np.array_equal(hcm4, want) and np.array_equal(htree, wt) and np.array_equal(hroot, wt[-4:])


@pytest.mark.gpu
def test_fri_layer_commit(glb, oracle, fixtures):
    """FRIProve::prove one layer (friProve.cpp:36-126): fold in F_p^3, getTransposed, MerkleTreeGL merkelize."""
    cases = [(pb, nb, fixtures["in"]["frifold_%d_%d" % (pb, nb)]) for (pb, nb) in fixtures["meta"]["fri_fold"]]
    rng = np.random.default_rng(22)
    cases.append((14, 10, (rng.integers(0, P, size=(1 << 14, 3), dtype=np.uint64), rng.integers(0, P, size=3, dtype=np.uint64), 12345678901234567)))
    cases.append((11, 5, (rng.integers(0, P, size=(1 << 11, 3), dtype=np.uint64), rng.integers(0, P, size=3, dtype=np.uint64), 7)))
    for (pb, nb, (pol, x, si)) in cases:
        want = oracle.fri_fold(pol, nb, si, x)
        key = "frifold_%d_%d" % (pb, nb)
        if key in fixtures["out"]:
            assert np.array_equal(want, fixtures["out"][key])
        for tb in sorted({max(nb - 2, 0), nb}):
            dpol = glb.DeviceBuffer((1 << pb) * 3).upload(pol)
            dpol2 = glb.DeviceBuffer((1 << nb) * 3)
            daux = glb.DeviceBuffer((1 << nb) * 3)
            dtree = glb.DeviceBuffer(glb.tree_num_elements(1 << tb))
            root = glb.fri_layer_commit_dev(dtree.ptr, daux.ptr, dpol2.ptr, dpol.ptr, pb, nb, tb, si, x)
            assert np.array_equal(dpol2.download().reshape(-1, 3), want), (pb, nb)
            aux = oracle.fri_transpose(want, tb)
            assert np.array_equal(daux.download().reshape(-1, 3), aux)
            wt = oracle.merkletree(aux.reshape(1 << tb, -1))
            assert np.array_equal(dtree.download(), wt), (pb, nb, tb)
            assert np.array_equal(root, wt[-4:])
            # host-pointer form (gl_fri_layer_commit_host)
            import ctypes as C
            hp = np.ascontiguousarray(pol, dtype=np.uint64)
            hx = np.ascontiguousarray(x, dtype=np.uint64)
            hpol2, haux, htree, hroot = np.zeros_like(want), np.zeros_like(want), np.zeros(wt.size, dtype=np.uint64), np.zeros(4, dtype=np.uint64)
            u64p = C.POINTER(C.c_uint64)
            rc = glb.lib().gl_fri_layer_commit_host(htree.ctypes.data, haux.ctypes.data, hpol2.ctypes.data, hp.ctypes.data, pb, nb, tb, int(si),
                                                    hx.ctypes.data_as(u64p), hroot.ctypes.data_as(u64p))
            assert rc == 0 and np.array_equal(hpol2, want) and np.array_equal(haux, aux) and np.array_equal(htree, wt) and np.array_equal(hroot, wt[-4:])


@pytest.mark.gpu
def test_fri_fold_literal_shim_path(glb, oracle):
    """FRIProve::prove's fold loop as the reference writes it (friProve.cpp:36-107): per group an NTT_Goldilocks(nX, 1).INTT on
    an nX x 3 matrix -- here through gl_ntt_host, which is what the shim's INTT calls -- then polMulAxi (:192-200) and evalPol
    (:201-217).  2^14 -> 2^10.  Must equal the oracle's restatement and the fused device kernel (gl_fri_fold_dev)."""
    import ctypes as C
    rng = np.random.default_rng(33)
    pb, nb = 14, 10
    pol = rng.integers(0, P, size=(1 << pb, 3), dtype=np.uint64)
    x = rng.integers(0, P, size=3, dtype=np.uint64)
    shift_inv = int(rng.integers(1, P, dtype=np.uint64))
    pol2n, nx = 1 << nb, 1 << (pb - nb)
    L = oracle.L
    wi = L.glo_inv(L.glo_root_of_unity(pb))

    def gl3_mul(a, b):
        r = np.zeros(3, dtype=np.uint64)
        L.glo_gl3_mul(r.ctypes.data_as(C.POINTER(C.c_uint64)), np.ascontiguousarray(a).ctypes.data_as(C.POINTER(C.c_uint64)),
                      np.ascontiguousarray(b).ctypes.data_as(C.POINTER(C.c_uint64)))
        return r

    out = np.zeros((pol2n, 3), dtype=np.uint64)
    sinv = shift_inv
    for g in range(pol2n):
        ppar = np.ascontiguousarray(pol[g::pol2n])                      # rows i * pol2N + g, i < nX
        c = glb.ntt_host(ppar, inverse=True)                            # the shim's NTT_Goldilocks(nX, 1).INTT
        r = 1
        for i in range(nx):                                             # polMulAxi
            c[i] = [L.glo_mul(int(v), r) for v in c[i]]
            r = L.glo_mul(r, sinv)
        res = c[nx - 1].copy()
        for i in range(nx - 2, -1, -1):                                 # evalPol (Horner in F_p^3)
            res = gl3_mul(res, x)
            res = np.array([L.glo_add(int(res[d]), int(c[i][d])) for d in range(3)], dtype=np.uint64)
        out[g] = res
        sinv = L.glo_mul(sinv, wi)
    want = oracle.fri_fold(pol, nb, shift_inv, x)
    assert np.array_equal(out, want)
    dpol = glb.DeviceBuffer(pol.size).upload(pol)
    dout = glb.DeviceBuffer(pol2n * 3)
    assert glb.lib().gl_fri_fold_dev(C.c_void_p(dout.ptr), C.c_void_p(dpol.ptr), pb, nb, shift_inv, x.ctypes.data_as(C.POINTER(C.c_uint64))) == 0
    glb.lib().gl_sync()
    assert np.array_equal(dout.download().reshape(-1, 3), want)


@pytest.mark.gpu
def test_consttree_image(glb, oracle):
    """bctree (build_const_tree.cpp:333-407): [nPols, nExt, extended constants, nodes] re-opened as MerkleTreeGL."""
    rng = np.random.default_rng(23)
    for (n, e, c) in [(64, 128, 5), (256, 1024, 18), (16, 256, 1)]:
        x = rng.integers(0, P, size=(n, c), dtype=np.uint64)
        img = glb.build_consttree_host(x, e)
        lde = oracle.extend_pol(x, e)
        tree = oracle.merkletree(lde)
        assert img[0] == c and img[1] == e and img.size == 2 + c * e + tree.size
        assert np.array_equal(img[2:2 + c * e].reshape(e, c), lde)
        assert np.array_equal(img[2 + c * e:], tree)
        # the view MerkleTreeGL(Element *tree) takes (merkleTreeGL.hpp:24-32): width, height, source, nodes
        w, h = int(img[0]), int(img[1])
        proof = oracle.group_proof(img[2:2 + h * w].reshape(h, w), img[2 + h * w:], 3)
        assert np.array_equal(proof[:w], lde[3])


@pytest.mark.gpu
def test_consttree_image_vs_bctree_tool(glb, oracle, ref, fixtures):
    """gl_build_consttree_host against the reference tool itself: the extended constants must equal the stored outputs of
    interpolate() of tools/starkpil/bctree/build_const_tree.cpp:198-330 (compiled where it lies, tests/golden/make_golden.py),
    the nodes the reference's PoseidonGoldilocks::merkletree over them (:366) and the last 4 words the verification key (:386-392)."""
    cpu = ref if ref is not None else oracle
    for (n, e, c) in fixtures["meta"]["bctree"]:
        k = "bctree_%d_%d_%d" % (n, e, c)
        img = glb.build_consttree_host(fixtures["in"][k], e)
        want = fixtures["out"][k]
        assert img[0] == c and img[1] == e, k
        assert np.array_equal(img[2:2 + c * e].reshape(e, c), want), k
        tree = cpu.merkletree(want)
        assert np.array_equal(img[2 + c * e:], tree), k
        assert np.array_equal(img[-4:], tree[-4:]), k


@pytest.mark.gpu
def test_cpp_shims_run_reference_tests(glb):
    """host/test_shims: the reference's gtest cases (tests.cpp merkletree/ntt/extendePol + the starks.cpp:52-56 sequence)
    compiled against the drop-in C++ classes (NTT_Goldilocks, MerkleTreeGL, PoseidonGoldilocks) over the C ABI."""
    import os
    import subprocess
    from conftest import ROOT
    exe = os.path.join(ROOT, "merlin-zkevm-prover_b200", "host", "test_shims")
    if not os.path.exists(exe):
        subprocess.run(["make", "-C", os.path.dirname(exe)], check=True)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
