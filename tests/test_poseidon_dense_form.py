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
