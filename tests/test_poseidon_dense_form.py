"""The CUDA permutation evaluates the 22 partial rounds in a different but equivalent form (csrc/poseidon.cuh): the same
small-entry MDS layer as the full rounds after an S-box on word 0, two rounds per layer via the squared matrix.  This test
proves the equivalence on the reference's own constants with exact integer arithmetic (no GPU, no oracle library):
the reference's optimised form (poseidon_goldilocks.cpp:5-37: dense matrix P, sparse matrices S) == dense form == paired form,
and the optimised form reproduces the reference's known-answer vector (tests.cpp:1631-1642)."""
import os
import random
import re

P = 0xFFFFFFFF00000001
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _arr(src, name):
    m = re.search(r"%s\[\d+\] = \{(.*?)\};" % name, src, re.S)
    return [int(x[:-3], 16) for x in re.findall(r"0x[0-9A-Fa-f]+ULL", m.group(1))]


SRC = open(os.path.join(ROOT, "oracle", "poseidon_constants.h")).read()
RC, PM, SP = _arr(SRC, "GLP_RC"), _arr(SRC, "GLP_PM"), _arr(SRC, "GLP_SP")
CIRC = [17, 20, 34, 18, 39, 13, 13, 28, 2, 16, 41, 15]


def M(j, i):
    return CIRC[(i - j) % 12] + (8 if i == 0 and j == 0 else 0)


def M2(j, i):
    return sum(M(j, k) * M(k, i) for k in range(12))


def mds(s):
    return [sum(M(j, i) * s[j] for j in range(12)) % P for i in range(12)]


def sb(x):
    return pow(x, 7, P)


def _head(s):
    s = [(s[i] + RC[i]) % P for i in range(12)]
    for r in range(3):
        s = mds([(sb(s[i]) + RC[(r + 1) * 12 + i]) % P for i in range(12)])
    return [(sb(s[i]) + RC[48 + i]) % P for i in range(12)]


def _tail(s):
    for r in range(3):
        s = mds([(sb(s[i]) + RC[82 + 12 * r + i]) % P for i in range(12)])
    return mds([sb(x) for x in s])


def perm_reference_form(s):
    s = _head(s)
    s = [sum(PM[j * 12 + i] * s[j] for j in range(12)) % P for i in range(12)]
    for r in range(22):
        t = (sb(s[0]) + RC[60 + r]) % P
        s0 = (t * SP[23 * r] + sum(s[i] * SP[23 * r + i] for i in range(1, 12))) % P
        s = [s0] + [(s[i] + t * SP[23 * r + 11 + i]) % P for i in range(1, 12)]
    return _tail(s)


def perm_dense_form(s):
    s = mds(_head(s))
    for r in range(22):
        s = mds([(sb(s[0]) + RC[60 + r]) % P] + s[1:])
    return _tail(s)


def perm_paired_form(s):
    s = mds(_head(s))
    for q in range(11):
        s = [(sb(s[0]) + RC[60 + 2 * q]) % P] + s[1:]
        x = sum(M(j, 0) * s[j] for j in range(12)) % P
        d = (sb(x) + RC[60 + 2 * q + 1] - x) % P
        s = [(sum(M2(j, i) * s[j] for j in range(12)) + M(0, i) * d) % P for i in range(12)]
    return _tail(s)


def test_constants_parsed():
    assert (len(RC), len(PM), len(SP)) == (118, 144, 506)
    assert all(SP[23 * r] == 25 for r in range(22))


def test_reference_form_reproduces_the_kat():
    out = perm_reference_form([0] * 12)
    assert out[0] == 0x3C18A9786CB0B359 and out[11] == 0x1792B1C4342109D7


def test_dense_and_paired_forms_equal_the_reference_form():
    rng = random.Random(7)
    cases = [[0] * 12, [P - 1] * 12, list(range(12))] + [[rng.randrange(P) for _ in range(12)] for _ in range(12)]
    for x in cases:
        want = perm_reference_form(x)
        assert perm_dense_form(x) == want
        assert perm_paired_form(x) == want


def test_fp64_exactness_bounds():
    """Every partial sum of the FP64 layers is an integer below 2^53: 2^52 bias + 32-bit offset + 12 (+1) products."""
    worst1 = max(sum(M(j, i) for j in range(12)) for i in range(12))
    worst2 = max(sum(M2(j, i) for j in range(12)) + M(0, i) for i in range(12))
    assert (1 << 52) + (1 << 32) + worst1 * (2**32 - 1) < (1 << 53)
    assert (1 << 52) + (1 << 32) + worst2 * (2**32 - 1) < (1 << 53)


# ---- the split-circulant evaluation of the MDS layers (csrc/poseidon.cuh: cyc12_pair) -----------------------------------
# c (*) v mod x^12 - 1 = CRT of a cyclic product mod x^6 - 1 on (v_lo + v_hi) and a negacyclic product mod x^6 + 1 on
# (v_lo - v_hi); the halved coefficient vectors are integers because c_lo +- c_hi is even.
def _cyc(c, v):
    n = len(c)
    return [sum(c[(k - a) % n] * v[a] for a in range(n)) for k in range(n)]


C2 = _cyc(CIRC, CIRC)  # circulant part of the squared matrix


def _split_cyc12(c, v):
    assert all((c[k] + c[k + 6]) % 2 == 0 for k in range(6))
    cp = [(c[k] + c[k + 6]) // 2 for k in range(6)]
    cm = [(c[k] - c[k + 6]) // 2 for k in range(6)]
    vp = [v[a] + v[a + 6] for a in range(6)]
    vm = [v[a] - v[a + 6] for a in range(6)]
    y = [0] * 12
    worst = 0
    for k in range(6):
        pk = sum(cp[(k - a) % 6] * vp[a] for a in range(6))
        nk = sum((1 if a <= k else -1) * cm[(k - a) % 6] * vm[a] for a in range(6))
        worst = max(worst, pk, abs(nk))
        y[k], y[k + 6] = pk + nk, pk - nk
    return y, worst


def test_split_circulant_layer_is_the_mds_layer():
    rng = random.Random(11)
    for _ in range(50):
        # 32-bit halves, as the kernel feeds them
        v = [rng.choice([0, 1, 2**32 - 1, rng.randrange(2**32)]) for _ in range(12)]
        y, worst = _split_cyc12(CIRC, v)
        y[0] += 8 * v[0]
        assert y == [sum(M(j, i) * v[j] for j in range(12)) for i in range(12)]
        assert (1 << 52) + (1 << 32) + worst < (1 << 53) and max(y) + (1 << 52) + (1 << 32) < (1 << 53)


def test_split_circulant_pair_layer():
    """(M M) a + d M[0][.]  ==  c2 (*) a + (8 a_0 + d) c + 8 (x + d) e_0  with x = <M[.][0], a>, on 32-bit halves."""
    rng = random.Random(12)
    assert all((C2[k] + C2[k + 6]) % 2 == 0 for k in range(6))
    for _ in range(50):
        a = [rng.choice([0, 2**32 - 1, rng.randrange(2**32)]) for _ in range(12)]
        d = rng.choice([0, 2**32 - 1, rng.randrange(2**32)])
        x = sum(M(j, 0) * a[j] for j in range(12))
        want = [sum(M2(j, i) * a[j] for j in range(12)) + M(0, i) * d for i in range(12)]
        y, worst = _split_cyc12(C2, a)
        w = 8 * a[0] + d
        got = [y[i] + CIRC[i] * w for i in range(12)]
        got[0] += 8 * (x + d)
        assert got == want
        assert (1 << 52) + (1 << 32) + worst + 64 * w + 8 * (x + d) < (1 << 53)


def test_split_circulant_pair_layer_mod_p_with_y():
    """What the kernel does: 8 (x + d) e_0 is replaced by 8 y e_0 (y = x^7 + c = x + d mod p), per 32-bit half."""
    rng = random.Random(13)
    for _ in range(40):
        a = [rng.randrange(2**64) for _ in range(12)]                       # any representatives
        cn = rng.randrange(P)
        x = sum(M(j, 0) * a[j] for j in range(12)) % P
        y = (sb(x) + cn) % P
        d = (y - x) % P
        want = [(sum(M2(j, i) * a[j] for j in range(12)) + M(0, i) * d) % P for i in range(12)]

        def half(sel):
            ah, dh, yh = [sel(v) for v in a], sel(d), sel(y)
            out, _ = _split_cyc12(C2, ah)
            w = 8 * ah[0] + dh
            out = [out[i] + CIRC[i] * w for i in range(12)]
            out[0] += 8 * yh
            assert 0 <= min(out) and max(out) + (1 << 52) + (1 << 32) < (1 << 53)
            return out
        lo, hi = half(lambda v: v & 0xFFFFFFFF), half(lambda v: v >> 32)
        assert [(lo[i] + (hi[i] << 32)) % P for i in range(12)] == want


# ---- second CRT level on the cyclic half (GLB_MDS_SPLIT=2 in csrc/poseidon.cuh): x^6 - 1 = (x^3 - 1)(x^3 + 1) ---------------
def _split2_cyc12(c, v):
    cp = [(c[k] + c[k + 6]) // 2 for k in range(6)]
    cm = [(c[k] - c[k + 6]) // 2 for k in range(6)]
    assert all((cp[k] + cp[k + 3]) % 2 == 0 for k in range(3))
    cpp = [(cp[k] + cp[k + 3]) // 2 for k in range(3)]
    cpm = [(cp[k] - cp[k + 3]) // 2 for k in range(3)]
    vp = [v[a] + v[a + 6] for a in range(6)]
    vm = [v[a] - v[a + 6] for a in range(6)]
    vpp = [vp[a] + vp[a + 3] for a in range(3)]
    vpm = [vp[a] - vp[a + 3] for a in range(3)]
    pk = [0] * 6
    worst = 0
    for k in range(3):
        pp = sum(cpp[(k - a) % 3] * vpp[a] for a in range(3))
        pn = sum((1 if a <= k else -1) * cpm[(k - a) % 3] * vpm[a] for a in range(3))
        pk[k], pk[k + 3] = pp + pn, pp - pn
        worst = max(worst, pp, abs(pn))
    y = [0] * 12
    for k in range(6):
        nk = sum((1 if a <= k else -1) * cm[(k - a) % 6] * vm[a] for a in range(6))
        y[k], y[k + 6] = pk[k] + nk, pk[k] - nk
        worst = max(worst, pk[k], abs(nk))
    return y, worst


def test_two_level_split_equals_one_level():
    rng = random.Random(14)
    for c in (CIRC, C2):
        for _ in range(40):
            v = [rng.choice([0, 2**32 - 1, rng.randrange(2**32)]) for _ in range(12)]
            y2, worst = _split2_cyc12(c, v)
            assert y2 == _cyc(c, v)
            assert (1 << 52) + (1 << 32) + worst < (1 << 53)
