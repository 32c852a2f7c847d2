"""ctypes bindings for the CPU oracle -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  It exposes

  * ``oracle``  -- oracle/libgloracle.so, our plain-C restatement (gl_oracle.c)
  * ``ref()``   -- oracle/_ref/libglref_{avx2,avx512}.so, the UNMODIFIED reference compiled by
                   oracle/Makefile from /root/reference (None when it was never built).

Nothing in the product package imports this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
P = 0xFFFFFFFF00000001
XORSHIFT_SEED = 88172645463325252

_u64p = C.POINTER(C.c_uint64)


def _p(a):
    if a is None:
        return None
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_u64p)


def build(ref=True, quiet=True):
    """(Re)build libgloracle.so and, when /root/reference exists, oracle/_ref/*.so."""
    targets = ["oracle"]
    if ref and os.path.isdir("/root/reference/src/goldilocks/src"):
        targets.append("ref")
    subprocess.run(["make", "-C", HERE, "-j2"] + targets, check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


class _Oracle:
    def __init__(self):
        path = os.path.join(HERE, "libgloracle.so")
        if not os.path.exists(path):
            build(ref=False)
        L = C.CDLL(path)
        u64 = C.c_uint64
        for name, res, args in [
            ("glo_canon", u64, [u64]), ("glo_add", u64, [u64, u64]), ("glo_sub", u64, [u64, u64]),
            ("glo_mul", u64, [u64, u64]), ("glo_pow", u64, [u64, u64]), ("glo_inv", u64, [u64]),
            ("glo_root_of_unity", u64, [C.c_uint]),
            ("glo_ntt", None, [_u64p, _u64p, u64, u64, C.c_int, C.c_int]),
            ("glo_extend_pol", None, [_u64p, _u64p, u64, u64, u64]),
            ("glo_hash_full_result", None, [_u64p, _u64p]),
            ("glo_linear_hash", None, [_u64p, _u64p, u64]),
            ("glo_merkletree", None, [_u64p, _u64p, u64, u64]),
            ("glo_merkletree_batch", None, [_u64p, _u64p, u64, u64, u64]),
            ("glo_tree_num_elements", u64, [u64]),
            ("glo_merkle_proof_size", u64, [u64]),
            ("glo_group_proof", None, [_u64p, _u64p, _u64p, u64, u64, u64]),
            ("glo_fill_xorshift", None, [_u64p, u64, _u64p]),
            ("glo_gl3_mul", None, [_u64p, _u64p, _u64p]),
            ("glo_q_extend", None, [_u64p, _u64p, u64, u64, u64, u64]),
            ("glo_fri_fold", None, [_u64p, _u64p, C.c_uint, C.c_uint, u64, _u64p]),
            ("glo_fri_transpose", None, [_u64p, _u64p, u64, C.c_uint]),
            ("glo_num_threads", C.c_int, []), ("glo_set_num_threads", None, [C.c_int]),
        ]:
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        self.L = L

    # numpy-level helpers -------------------------------------------------
    def ntt(self, src, inverse=False, extend=False):
        src = np.ascontiguousarray(src, dtype=np.uint64)
        size, ncols = src.shape
        dst = np.empty_like(src)
        self.L.glo_ntt(_p(dst), _p(src), size, ncols, int(inverse), int(extend))
        return dst if size and ncols else src.copy()

    def extend_pol(self, src, n_ext):
        src = np.ascontiguousarray(src, dtype=np.uint64)
        n, ncols = src.shape
        out = np.zeros((n_ext, ncols), dtype=np.uint64)
        self.L.glo_extend_pol(_p(out), _p(src), n_ext, n, ncols)
        return out

    def hash_full_result(self, inp):
        inp = np.ascontiguousarray(inp, dtype=np.uint64)
        out = np.empty(12, dtype=np.uint64)
        self.L.glo_hash_full_result(_p(out), _p(inp))
        return out

    def linear_hash(self, inp):
        inp = np.ascontiguousarray(inp, dtype=np.uint64).reshape(-1)
        out = np.empty(4, dtype=np.uint64)
        buf = inp if inp.size else np.zeros(1, dtype=np.uint64)
        self.L.glo_linear_hash(_p(out), _p(buf), inp.size)
        return out

    def merkletree(self, src, nrows=None, ncols=None):
        src = np.ascontiguousarray(src, dtype=np.uint64)
        if nrows is None:
            nrows, ncols = src.shape
        tree = np.zeros(self.L.glo_tree_num_elements(nrows), dtype=np.uint64)
        buf = src if src.size else np.zeros(1, dtype=np.uint64)
        self.L.glo_merkletree(_p(tree), _p(buf), ncols, nrows)
        return tree

    def merkletree_batch(self, src, batch):
        src = np.ascontiguousarray(src, dtype=np.uint64)
        nrows, ncols = src.shape
        tree = np.zeros(self.L.glo_tree_num_elements(nrows), dtype=np.uint64)
        self.L.glo_merkletree_batch(_p(tree), _p(src), ncols, nrows, batch)
        return tree

    def proof_size(self, height, width):
        return width + 4 * self.L.glo_merkle_proof_size(height)

    def q_extend(self, q_2ns, n, qdeg):
        q_2ns = np.ascontiguousarray(q_2ns, dtype=np.uint64)
        n_ext, qdim = q_2ns.shape
        out = np.zeros((n_ext, qdim * qdeg), dtype=np.uint64)
        self.L.glo_q_extend(_p(out), _p(q_2ns), n, n_ext, qdim, qdeg)
        return out

    def fri_fold(self, pol, next_bits, shift_inv, x):
        pol = np.ascontiguousarray(pol, dtype=np.uint64).reshape(-1, 3)
        pol_bits = int(pol.shape[0]).bit_length() - 1
        out = np.zeros((1 << next_bits, 3), dtype=np.uint64)
        x = np.ascontiguousarray(x, dtype=np.uint64)
        self.L.glo_fri_fold(_p(out), _p(pol), pol_bits, next_bits, int(shift_inv), _p(x))
        return out

    def fri_transpose(self, pol, transpose_bits):
        pol = np.ascontiguousarray(pol, dtype=np.uint64).reshape(-1, 3)
        aux = np.zeros_like(pol)
        self.L.glo_fri_transpose(_p(aux), _p(pol), pol.shape[0], transpose_bits)
        return aux

    def group_proof(self, src, nodes, idx):
        src = np.ascontiguousarray(src, dtype=np.uint64)
        h, w = src.shape
        proof = np.zeros(self.proof_size(h, w), dtype=np.uint64)
        self.L.glo_group_proof(_p(proof), _p(src), _p(nodes), h, w, idx)
        return proof

    def xorshift(self, count, seed=XORSHIFT_SEED):
        out = np.empty(count, dtype=np.uint64)
        st = np.array([seed], dtype=np.uint64)
        self.L.glo_fill_xorshift(_p(out), count, _p(st))
        return out, int(st[0])


class _Ref:
    """The unmodified reference (oracle/_ref).  Outputs are canonicalised here, because the
    reference stores lazily-reduced words and its own tests compare through toU64."""

    def __init__(self, path):
        L = C.CDLL(path)
        u64 = C.c_uint64
        for name, res, args in [
            ("ref_num_threads", C.c_int, []), ("ref_has_avx512", C.c_int, []), ("ref_set_num_threads", None, [C.c_int]),
            ("ref_ntt", None, [_u64p, _u64p, u64, u64, u64, u64]),
            ("ref_intt", None, [_u64p, _u64p, u64, u64, u64, u64]),
            ("ref_extend_pol", None, [_u64p, _u64p, u64, u64, u64, _u64p]),
            ("ref_hash_full_result", None, [_u64p, _u64p, C.c_int]),
            ("ref_linear_hash", None, [_u64p, _u64p, u64, C.c_int]),
            ("ref_merkletree", C.c_int, [_u64p, _u64p, u64, u64, C.c_int]),
            ("ref_merkletree_batch", C.c_int, [_u64p, _u64p, u64, u64, u64, C.c_int]),
            ("ref_proof_size", u64, [u64, u64]),
            ("ref_group_proof", None, [_u64p, _u64p, _u64p, u64, u64, u64]),
            ("ref_commit", None, [_u64p, _u64p, _u64p, u64, u64, u64, _u64p,
                                  C.POINTER(C.c_double), C.POINTER(C.c_double)]),
            ("ref_gl3_mul", None, [_u64p, _u64p, _u64p]),
            ("ref_q_extend", None, [_u64p, _u64p, u64, u64, u64, u64]),
            ("ref_fri_fold", None, [_u64p, _u64p, C.c_uint, C.c_uint, u64, _u64p]),
        ]:
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        self.L = L
        self.path = path
        self.avx512 = bool(L.ref_has_avx512())
        self.tree_variant = 2 if self.avx512 else 1

    @staticmethod
    def canon(a):
        a = a.copy()
        a[a >= np.uint64(P)] -= np.uint64(P)
        return a

    def ntt(self, src, inverse=False, nphase=3, nblock=1):
        src = np.ascontiguousarray(src, dtype=np.uint64).copy()
        size, ncols = src.shape
        dst = np.zeros_like(src)
        (self.L.ref_intt if inverse else self.L.ref_ntt)(_p(dst), _p(src), size, ncols, nphase, nblock)
        return self.canon(dst)

    def extend_pol(self, src, n_ext):
        src = np.ascontiguousarray(src, dtype=np.uint64).copy()
        n, ncols = src.shape
        out = np.zeros((n_ext, ncols), dtype=np.uint64)
        self.L.ref_extend_pol(_p(out), _p(src), n_ext, n, ncols, None)
        return self.canon(out)

    def hash_full_result(self, inp, variant=1):
        inp = np.ascontiguousarray(inp, dtype=np.uint64)
        out = np.empty(12, dtype=np.uint64)
        self.L.ref_hash_full_result(_p(out), _p(inp), variant)
        return self.canon(out)

    def linear_hash(self, inp, variant=1):
        inp = np.ascontiguousarray(inp, dtype=np.uint64).reshape(-1).copy()
        out = np.empty(4, dtype=np.uint64)
        self.L.ref_linear_hash(_p(out), _p(inp if inp.size else np.zeros(1, dtype=np.uint64)), inp.size, variant)
        return self.canon(out)

    def merkletree(self, src, variant=None, nrows=None, ncols=None):
        src = np.ascontiguousarray(src, dtype=np.uint64)
        if nrows is None:
            nrows, ncols = src.shape
        tree = np.zeros((2 * nrows - 1) * 4, dtype=np.uint64)
        buf = src if src.size else np.zeros(1, dtype=np.uint64)
        rc = self.L.ref_merkletree(_p(tree), _p(buf), ncols, nrows, self.tree_variant if variant is None else variant)
        assert rc == 0
        return self.canon(tree)

    def merkletree_batch(self, src, batch, variant=1):
        src = np.ascontiguousarray(src, dtype=np.uint64)
        nrows, ncols = src.shape
        tree = np.zeros((2 * nrows - 1) * 4, dtype=np.uint64)
        assert self.L.ref_merkletree_batch(_p(tree), _p(src), ncols, nrows, batch, variant) == 0
        return self.canon(tree)

    def group_proof(self, src, nodes, idx):
        src = np.ascontiguousarray(src, dtype=np.uint64)
        h, w = src.shape
        proof = np.zeros(self.L.ref_proof_size(h, w), dtype=np.uint64)
        self.L.ref_group_proof(_p(proof), _p(src), _p(nodes), h, w, idx)
        return self.canon(proof)

    def gl3_mul(self, a, b):
        a = np.ascontiguousarray(a, dtype=np.uint64).copy()
        b = np.ascontiguousarray(b, dtype=np.uint64).copy()
        r = np.zeros(3, dtype=np.uint64)
        self.L.ref_gl3_mul(_p(r), _p(a), _p(b))
        return self.canon(r)

    def q_extend(self, q_2ns, n, qdeg):
        q_2ns = np.ascontiguousarray(q_2ns, dtype=np.uint64).copy()
        n_ext, qdim = q_2ns.shape
        out = np.zeros((n_ext, qdim * qdeg), dtype=np.uint64)
        self.L.ref_q_extend(_p(out), _p(q_2ns), n, n_ext, qdim, qdeg)
        return self.canon(out)

    def fri_fold(self, pol, next_bits, shift_inv, x):
        pol = np.ascontiguousarray(pol, dtype=np.uint64).reshape(-1, 3).copy()
        pol_bits = int(pol.shape[0]).bit_length() - 1
        out = np.zeros((1 << next_bits, 3), dtype=np.uint64)
        x = np.ascontiguousarray(x, dtype=np.uint64).copy()
        self.L.ref_fri_fold(_p(out), _p(pol), pol_bits, next_bits, int(shift_inv), _p(x))
        return self.canon(out)

    def commit(self, src, n_ext, want_lde=True):
        """extendPol + merkelize (starks.cpp:52-56); returns (lde, tree, ms_lde, ms_tree)."""
        src = np.ascontiguousarray(src, dtype=np.uint64)
        n, ncols = src.shape
        lde = np.zeros((n_ext, ncols), dtype=np.uint64)
        scratch = np.zeros((n_ext, ncols), dtype=np.uint64)
        tree = np.zeros((2 * n_ext - 1) * 4, dtype=np.uint64)
        a, b = C.c_double(0), C.c_double(0)
        self.L.ref_commit(_p(lde), _p(src), _p(tree), n_ext, n, ncols, _p(scratch), C.byref(a), C.byref(b))
        return lde, tree, a.value, b.value


class _RefBctree:
    """interpolate() of the reference's offline constant-tree tool (tools/starkpil/bctree/build_const_tree.cpp:198-330),
    compiled where it lies by oracle/Makefile (oracle/_ref/libglref_bctree.so): a second, independent LDE implementation."""

    def __init__(self, path):
        L = C.CDLL(path)
        L.ref_bctree_interpolate.restype = None
        L.ref_bctree_interpolate.argtypes = [_u64p, _u64p, C.c_uint64, C.c_uint64, C.c_uint64]
        self.L = L

    def interpolate(self, src, n_ext):
        src = np.ascontiguousarray(src, dtype=np.uint64).copy()
        n, npols = src.shape
        out = np.zeros((n_ext, npols), dtype=np.uint64)
        self.L.ref_bctree_interpolate(_p(out), _p(src), npols, int(n).bit_length() - 1, int(n_ext).bit_length() - 1)
        return _Ref.canon(out)


def _cpu_has(flag):
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return flag in line.split()
    except OSError:
        pass
    return False


_oracle = None
_ref = {}


def oracle():
    global _oracle
    if _oracle is None:
        _oracle = _Oracle()
    return _oracle


def ref(prefer_avx512=False):
    """The compiled reference, or None if oracle/_ref was never built (e.g. no /root/reference)."""
    kind = "avx512" if (prefer_avx512 and _cpu_has("avx512f")) else "avx2"
    if kind not in _ref:
        path = os.path.join(HERE, "_ref", "libglref_%s.so" % kind)
        if not os.path.exists(path) or not _cpu_has("avx2"):
            _ref[kind] = None
        else:
            try:
                _ref[kind] = _Ref(path)
            except OSError:
                _ref[kind] = None
    return _ref[kind]


_bctree = []


def ref_bctree():
    """The reference bctree tool's interpolate(), or None when oracle/_ref/libglref_bctree.so was never built."""
    if not _bctree:
        path = os.path.join(HERE, "_ref", "libglref_bctree.so")
        try:
            _bctree.append(_RefBctree(path) if os.path.exists(path) and _cpu_has("avx2") else None)
        except OSError:
            _bctree.append(None)
    return _bctree[0]
