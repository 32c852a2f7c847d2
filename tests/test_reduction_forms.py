"""Carry logic of the Goldilocks reductions in csrc/gl64.cuh, simulated with exact 32-bit operations (no GPU):
reduce128 / reduce96 (one IMAD.WIDE for x2 * (2^32 - 1)), reduce96_one<1|2> (the MDS folds of the Poseidon layers: the second
step without a multiply-add) and reduce128_alu / reduce96_alu (the ALU-only forms of the round-2 experiment).  Each must return a value in [0, 2^64) congruent to its input mod p for
EVERY 128-bit (96-bit) input, with the single net correction k staying in {-1, 0, 1} ({0, 1})."""
import random

P = 0xFFFFFFFF00000001
M = 0xFFFFFFFF


def _addc(a, b, c=0):
    s = a + b + c
    return s & M, s >> 32


def _subc(a, b, c=0):
    d = a - b - c
    return d & M, 1 if d < 0 else 0


def _fix(r0, r1, k):
    assert k in (-1, 0, 1)
    r0, c = _addc(r0, (-k) & M)
    r1, c2 = _addc(r1, (k >> 31) & M, c)
    return r0 | (r1 << 32)


def reduce128(x0, x1, x2, x3):          # gl64.cuh reduce128
    r0, b = _subc(x0, x3)
    r1, b = _subc(x1, 0, b)
    t = (r0 | (r1 << 32)) + x2 * M      # mad.lo.cc / madc.hi.cc: one IMAD.WIDE with carry out
    return _fix(t & M, (t >> 32) & M, (t >> 64) - b)


def reduce128_alu(x0, x1, x2, x3):      # gl64.cuh reduce128_alu
    h, c1 = _addc(x1, x2)
    s0, s1 = _addc(x2, x3)
    r0, b = _subc(x0, s0)
    r1, b1 = _subc(h, s1, b)
    return _fix(r0, r1, c1 - b1)


def reduce96(x0, x1, x2):               # gl64.cuh reduce96
    t = (x0 | (x1 << 32)) + x2 * M
    k = t >> 64
    t = (t & ((1 << 64) - 1)) + k * M
    assert t < (1 << 64)
    return t


def reduce96_alu(x0, x1, x2):           # gl64.cuh reduce96_alu
    h, c1 = _addc(x1, x2)
    r0, b = _subc(x0, x2)
    r1, b1 = _subc(h, 0, b)
    k = c1 - b1
    assert k in (0, 1)
    return _fix(r0, r1, k)


def reduce96_one_1(x0, x1, x2):         # gl64.cuh reduce96_one<1>: second step as mask arithmetic (bfe.s32, add.cc, addc)
    t = (x0 | (x1 << 32)) + x2 * M
    k = t >> 64
    assert k in (0, 1)
    m = M if k else 0                   # sign extension of bit 0 of k
    r0, c = _addc(t & M, m)
    r1, c2 = _addc((t >> 32) & M, 0, c)
    assert c2 == 0                      # the added (2^32 - 1) never carries out again
    return r0 | (r1 << 32)


def reduce96_one_2(x0, x1, x2):         # gl64.cuh reduce96_one<2>: {r0 - k, r1 + k - borrow}
    t = (x0 | (x1 << 32)) + x2 * M
    k = t >> 64
    r0, b = _subc(t & M, k)
    r1 = ((t >> 32) & M) + k - b
    assert 0 <= r1 <= M
    return r0 | (r1 << 32)


EDGE = [0, 1, 2, M - 1, M, 0x80000000, 0x7FFFFFFF, 0xFFFF0000, 0x0000FFFF]


def _words(rng, n):
    return tuple(rng.choice(EDGE) if rng.random() < 0.4 else rng.randrange(1 << 32) for _ in range(n))


def test_reduce128_forms():
    rng = random.Random(1)
    cases = [(a, b, c, d) for a in EDGE for b in EDGE for c in EDGE for d in EDGE] + [_words(rng, 4) for _ in range(60000)]
    for x in cases:
        v = x[0] + (x[1] << 32) + (x[2] << 64) + (x[3] << 96)
        for f in (reduce128, reduce128_alu):
            r = f(*x)
            assert 0 <= r < (1 << 64) and r % P == v % P, (f.__name__, x)


def test_reduce96_forms():
    rng = random.Random(2)
    cases = [(a, b, c) for a in EDGE for b in EDGE for c in EDGE] + [_words(rng, 3) for _ in range(60000)]
    for x in cases:
        v = x[0] + (x[1] << 32) + (x[2] << 64)
        for f in (reduce96, reduce96_alu, reduce96_one_1, reduce96_one_2):
            r = f(*x)
            assert 0 <= r < (1 << 64) and r % P == v % P, (f.__name__, x)
