"""CPU mirror of the FP64 MDS layers of csrc/poseidon.cuh (mds_layer / mds_pair with GLB_MDS_SPLIT=2), operation by operation in
IEEE doubles: Python floats are binary64 and every intermediate is an exact integer below 2^53, so a * b + c here equals the
kernel's fused multiply-add.  Checks the 2^52 bias / raw-bit trick (fold_lh + reduce96), the two CRT levels and the pair layer
against plain integer arithmetic mod p, and asserts that no intermediate leaves the exact range."""
import random
import struct

from test_poseidon_dense_form import CIRC, C2, M, M2, P, sb

MASK = 0xFFFFFFFF
TWO52 = float(1 << 52)
TWO52_BITS = 0x4330000000000000
DM = (P - ((TWO52_BITS % P) * ((1 << 32) + 1)) % P) % P
BIAS_LO, BIAS_HI = TWO52 + float(DM & MASK), TWO52 + float(DM >> 32)


def _chk(x):
    assert x == int(x) and abs(x) < float(1 << 53), x
    return x


def fma(a, b, c):
    return _chk(a * b + c)


def bits(d):
    return struct.unpack("<Q", struct.pack("<d", d))[0]


def reduce96(x0, x1, x2):
    t = (x0 | (x1 << 32)) + x2 * MASK
    k = t >> 64
    return (t & ((1 << 64) - 1)) + k * MASK


def fold_lh(lo, hi):
    assert TWO52 <= lo < 2 * TWO52 and TWO52 <= hi < 2 * TWO52
    L, H = bits(lo), bits(hi)
    t = (L >> 32) + H
    assert t < (1 << 64)
    r = reduce96(L & MASK, t & MASK, t >> 32)
    assert r < (1 << 64)
    return r


def split_pm(v):
    return [_chk(v[a] + v[a + 6]) for a in range(6)], [_chk(v[a] - v[a + 6]) for a in range(6)]


def split_pm3(vp):
    return [_chk(vp[a] + vp[a + 3]) for a in range(3)], [_chk(vp[a] - vp[a + 3]) for a in range(3)]


def cp(c, i):
    return (c[i % 6] + c[i % 6 + 6]) // 2


def cyc_p2(c, k, vpp, vpm, bias):
    pp, pn = bias, 0.0
    for a in range(3):
        idx = (k - a + 3) % 3
        cpp = float((cp(c, idx) + cp(c, idx + 3)) // 2)
        cpm = float((1 if a <= k else -1) * ((cp(c, idx) - cp(c, idx + 3)) // 2))
        pp = fma(cpp, vpp[a], pp)
        pn = _chk(cpm * vpm[0]) if a == 0 else fma(cpm, vpm[a], pn)
    return _chk(pp + pn), _chk(pp - pn)


def cyc_n(c, k, vm):
    n = 0.0
    for a in range(6):
        idx = (k - a + 6) % 6
        cm = float((1 if a <= k else -1) * ((c[idx] - c[idx + 6]) // 2))
        n = _chk(cm * vm[0]) if a == 0 else fma(cm, vm[a], n)
    return n


def mds_layer(s):
    lo, hi = [float(x & MASK) for x in s], [float(x >> 32) for x in s]
    lp, lm = split_pm(lo)
    hp, hm = split_pm(hi)
    lpp, lpm = split_pm3(lp)
    hpp, hpm = split_pm3(hp)
    out = [0] * 12
    for k in range(3):
        pl, pl3 = cyc_p2(CIRC, k, lpp, lpm, BIAS_LO)
        ph, ph3 = cyc_p2(CIRC, k, hpp, hpm, BIAS_HI)
        nl, nh, nl3, nh3 = cyc_n(CIRC, k, lm), cyc_n(CIRC, k, hm), cyc_n(CIRC, k + 3, lm), cyc_n(CIRC, k + 3, hm)
        a_l, a_h = _chk(pl + nl), _chk(ph + nh)
        if k == 0:
            a_l, a_h = fma(8.0, lo[0], a_l), fma(8.0, hi[0], a_h)
        out[k] = fold_lh(a_l, a_h)
        out[k + 6] = fold_lh(_chk(pl - nl), _chk(ph - nh))
        out[k + 3] = fold_lh(_chk(pl3 + nl3), _chk(ph3 + nh3))
        out[k + 9] = fold_lh(_chk(pl3 - nl3), _chk(ph3 - nh3))
    return out


def mds_pair(s, c_next):
    lo, hi = [float(x & MASK) for x in s], [float(x >> 32) for x in s]
    dl, dh = BIAS_LO, BIAS_HI
    for j in range(12):
        dl, dh = fma(float(M(j, 0)), lo[j], dl), fma(float(M(j, 0)), hi[j], dh)
    x = fold_lh(dl, dh)
    y = (sb(x % P) + c_next) % P + (P if random.random() < 0.3 and (sb(x % P) + c_next) % P < (1 << 64) - P else 0)  # any representative
    d = (y - x) % P
    dlo, dhi = float(d & MASK), float(d >> 32)
    wl, wh = fma(8.0, lo[0], dlo), fma(8.0, hi[0], dhi)
    yl8, yh8 = _chk(8.0 * float(y & MASK)), _chk(8.0 * float(y >> 32))
    lp, lm = split_pm(lo)
    hp, hm = split_pm(hi)
    lpp, lpm = split_pm3(lp)
    hpp, hpm = split_pm3(hp)
    out = [0] * 12
    for k in range(3):
        pl, pl3 = cyc_p2(C2, k, lpp, lpm, BIAS_LO)
        ph, ph3 = cyc_p2(C2, k, hpp, hpm, BIAS_HI)
        nl, nh, nl3, nh3 = cyc_n(C2, k, lm), cyc_n(C2, k, hm), cyc_n(C2, k + 3, lm), cyc_n(C2, k + 3, hm)
        a_l, a_h = fma(float(CIRC[k]), wl, _chk(pl + nl)), fma(float(CIRC[k]), wh, _chk(ph + nh))
        if k == 0:
            a_l, a_h = _chk(a_l + yl8), _chk(a_h + yh8)
        out[k] = fold_lh(a_l, a_h)
        out[k + 6] = fold_lh(fma(float(CIRC[k + 6]), wl, _chk(pl - nl)), fma(float(CIRC[k + 6]), wh, _chk(ph - nh)))
        out[k + 3] = fold_lh(fma(float(CIRC[k + 3]), wl, _chk(pl3 + nl3)), fma(float(CIRC[k + 3]), wh, _chk(ph3 + nh3)))
        out[k + 9] = fold_lh(fma(float(CIRC[k + 9]), wl, _chk(pl3 - nl3)), fma(float(CIRC[k + 9]), wh, _chk(ph3 - nh3)))
    return out, d


def _states(rng, n):
    edge = [0, 1, P - 1, P, (1 << 64) - 1, MASK, MASK << 32, 1 << 63]
    yield [(1 << 64) - 1] * 12
    yield [0] * 12
    for _ in range(n):
        yield [rng.choice(edge) if rng.random() < 0.3 else rng.randrange(1 << 64) for _ in range(12)]


def test_mds_layer_emulation():
    rng = random.Random(21)
    for s in _states(rng, 300):
        got = mds_layer(s)
        want = [sum(M(j, i) * s[j] for j in range(12)) % P for i in range(12)]
        assert [g % P for g in got] == want


def test_mds_pair_emulation():
    rng = random.Random(22)
    random.seed(5)
    for s in _states(rng, 300):
        c_next = rng.randrange(P)
        got, d = mds_pair(s, c_next)
        want = [(sum(M2(j, i) * s[j] for j in range(12)) + M(0, i) * d) % P for i in range(12)]
        assert [g % P for g in got] == want
