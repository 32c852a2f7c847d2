"""merlin-zkevm-prover_b200 -- B200 (sm_100a) implementation of the zkEVM prover's STARK commit path.

This Python module is a thin ctypes binding over the C ABI in ``include/gl_commit.h``
(``libglb200.so``, built in-tree from ``csrc/`` by ``__graft_entry__.build()``).  It exists for the
test-suite and ``bench.py``; the product boundary is the C ABI plus the header-compatible C++ shims
in ``host/`` (``NTT_Goldilocks``, ``MerkleTreeGL``, ``PoseidonGoldilocks``).

There is no CPU fallback: importing works without a GPU (so symbols can be checked), but every
compute call raises ``GLError`` unless the CUDA library is built and a B200 is visible.

The directory name contains a hyphen (it is the reference's name), so load it with
``importlib`` -- see ``load()`` in ``__graft_entry__.py`` -- under the module name
``merlin_zkevm_prover_b200``.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# GLB_LIB selects another build of the same C ABI (A/B runs of kernel variants under tools/); the product is libglb200.so
LIB_PATH = os.environ.get("GLB_LIB") or os.path.join(HERE, "libglb200.so")
P = 0xFFFFFFFF00000001

_u64p = C.POINTER(C.c_uint64)
_u64 = C.c_uint64

# name -> (restype, argtypes); must list every symbol of include/gl_commit.h (tests check this)
SIGNATURES = {
    "gl_init": (C.c_int, [C.c_int]),
    "gl_shutdown": (None, []),
    "gl_last_error": (C.c_char_p, []),
    "gl_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "gl_set_scratch_limit": (C.c_int, [C.c_size_t]),
    "gl_set_resident": (C.c_int, [C.c_int]),
    "gl_set_ntt_variant": (C.c_int, [C.c_int]),
    "gl_resident_query": (C.c_int, [_u64p, _u64, _u64, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]),
    "gl_kernel_launches": (_u64, [C.c_int]),
    "gl_dev_alloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t]),
    "gl_dev_free": (C.c_int, [C.c_void_p]),
    "gl_host_alloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t]),
    "gl_host_free": (C.c_int, [C.c_void_p]),
    "gl_host_register": (C.c_int, [C.c_void_p, C.c_size_t]),
    "gl_host_unregister": (C.c_int, [C.c_void_p]),
    "gl_h2d": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "gl_d2h": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "gl_sync": (C.c_int, []),
    "gl_stream": (C.c_void_p, []),
    "gl_ntt_dev": (C.c_int, [C.c_void_p, C.c_void_p, _u64, _u64, C.c_int, C.c_int]),
    "gl_extend_pol_dev": (C.c_int, [C.c_void_p, C.c_void_p, _u64, _u64, _u64]),
    "gl_extend_pol_dev_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, _u64, _u64, _u64]),
    "gl_merkletree_dev": (C.c_int, [C.c_void_p, C.c_void_p, _u64, _u64]),
    "gl_merkletree_batch_dev": (C.c_int, [C.c_void_p, C.c_void_p, _u64, _u64, _u64]),
    "gl_merkletree_batch_host": (C.c_int, [_u64p, _u64p, _u64, _u64, _u64]),
    "gl_commit_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, _u64, _u64, _u64, _u64p]),
    "gl_group_proof_dev": (C.c_int, [_u64p, C.c_void_p, C.c_void_p, _u64, _u64, _u64]),
    "gl_group_proofs_dev": (C.c_int, [_u64p, C.c_void_p, C.c_void_p, _u64, _u64, _u64p, _u64]),
    "gl_rows_gather_dev": (C.c_int, [_u64p, C.c_void_p, _u64, _u64p, _u64]),
    "gl_copy2d": (C.c_int, [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int]),
    "gl_measure_issue_peaks": (C.c_int, [C.POINTER(C.c_double)]),
    "gl_hash_full_result_dev": (C.c_int, [C.c_void_p, C.c_void_p, _u64]),
    "gl_linear_hash_rows_dev": (C.c_int, [C.c_void_p, C.c_void_p, _u64, _u64]),
    "gl_ntt_host": (C.c_int, [_u64p, _u64p, _u64, _u64, C.c_int, C.c_int]),
    "gl_extend_pol_host": (C.c_int, [_u64p, _u64p, _u64, _u64, _u64]),
    "gl_merkletree_host": (C.c_int, [_u64p, _u64p, _u64, _u64]),
    "gl_commit_host": (C.c_int, [_u64p, _u64p, _u64p, _u64, _u64, _u64, _u64p]),
    "gl_hash_full_result_host": (C.c_int, [_u64p, _u64p, _u64]),
    "gl_linear_hash_host": (C.c_int, [_u64p, _u64p, _u64]),
    "gl_field_op_host": (C.c_int, [_u64p, _u64p, _u64p, _u64p, _u64, C.c_int]),
    "gl_fill_synthetic_dev": (C.c_int, [C.c_void_p, _u64, _u64, _u64]),
    "gl_tree_num_elements": (_u64, [_u64]),
    "gl_merkle_proof_size": (_u64, [_u64]),
    "gl_last_timings": (C.c_int, [C.POINTER(C.c_float)]),
    "gl_q_extend_dev": (C.c_int, [C.c_void_p, C.c_void_p, _u64, _u64, _u64, _u64]),
    "gl_q_commit_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, _u64, _u64, _u64, _u64, _u64p]),
    "gl_fri_fold_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, _u64, _u64p]),
    "gl_fri_transpose_dev": (C.c_int, [C.c_void_p, C.c_void_p, _u64, C.c_uint32]),
    "gl_fri_layer_commit_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, _u64, _u64p, _u64p]),
    "gl_q_commit_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, _u64, _u64, _u64, _u64, _u64p]),
    "gl_fri_layer_commit_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, _u64, _u64p, _u64p]),
    "gl_consttree_num_elements": (_u64, [_u64, _u64]),
    "gl_build_consttree_host": (C.c_int, [_u64p, _u64p, _u64, _u64, _u64]),
    "gl_last_timings4": (C.c_int, [C.POINTER(C.c_float)]),
    "gl_mgpu_unique_id": (C.c_int, [C.POINTER(C.c_uint8)]),
    "gl_mgpu_init": (C.c_int, [C.POINTER(C.c_uint8), C.c_int, C.c_int]),
    "gl_mgpu_shutdown": (C.c_int, []),
    "gl_mgpu_col_start": (_u64, [_u64, C.c_int, C.c_int]),
    "gl_mgpu_exchange_mode": (C.c_int, []),
    "gl_commit_mgpu": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, _u64, _u64, _u64, _u64p]),
    "gl_commit_mgpu_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, _u64, _u64, _u64, _u64p]),
    "gl_mgpu_gather_tree": (C.c_int, [C.c_void_p, C.c_void_p, _u64]),
    "gl_fill_synthetic_cols_dev": (C.c_int, [C.c_void_p, _u64, _u64, _u64, _u64, _u64]),
}


class GLError(RuntimeError):
    pass


def build(verbose=False):
    """Compile csrc/ into libglb200.so for sm_100a (nvcc cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-C", os.path.join(HERE, "csrc")], capture_output=not verbose, text=True)
    if r.returncode != 0:
        raise GLError("building libglb200.so failed:\n%s\n%s" % (r.stdout, r.stderr))
    return LIB_PATH


_lib = None


def lib():
    """The loaded C-ABI library.  Raises GLError if it was never built -- there is no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GLError("libglb200.so is missing (%s): run __graft_entry__.build(); there is no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise GLError("glb200 error %d: %s" % (rc, lib().gl_last_error().decode()))


def _p(a):
    if a is None:
        return None
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_u64p)


def init(device=0):
    _check(lib().gl_init(device))


def device_count():
    n = C.c_int(0)
    rc = lib().gl_device_count(C.byref(n))
    return n.value if rc == 0 else 0


def kernel_launches(reset=False):
    return int(lib().gl_kernel_launches(1 if reset else 0))


def last_timings():
    ms = (C.c_float * 3)()
    lib().gl_last_timings(ms)
    return {"lde_ms": ms[0], "leaf_ms": ms[1], "tree_ms": ms[2]}


def last_timings4():
    ms = (C.c_float * 4)()
    lib().gl_last_timings4(ms)
    return {"lde_ms": ms[0], "leaf_ms": ms[1], "tree_ms": ms[2], "exchange_ms": ms[3]}


def tree_num_elements(height):
    return int(lib().gl_tree_num_elements(height))


def proof_size(height, width):
    return int(width + 4 * lib().gl_merkle_proof_size(height))


# ---- device buffers ---------------------------------------------------------------------------
class DeviceBuffer:
    """cudaMalloc'ed buffer of uint64 words (gl_dev_alloc)."""

    def __init__(self, words):
        self.words = int(words)
        p = C.c_void_p()
        _check(lib().gl_dev_alloc(C.byref(p), self.words * 8))
        self.ptr = p.value

    def upload(self, a, word_offset=0):
        a = np.ascontiguousarray(a, dtype=np.uint64)
        assert word_offset + a.size <= self.words
        if a.size:
            _check(lib().gl_h2d(C.c_void_p(self.ptr + 8 * word_offset), a.ctypes.data_as(C.c_void_p), a.size * 8))
        return self

    def download(self, words=None, word_offset=0):
        words = self.words - word_offset if words is None else int(words)
        out = np.empty(words, dtype=np.uint64)
        if words:
            _check(lib().gl_d2h(out.ctypes.data_as(C.c_void_p), C.c_void_p(self.ptr + 8 * word_offset), words * 8))
        return out

    def at(self, word_offset):
        return C.c_void_p(self.ptr + 8 * int(word_offset))

    def free(self):
        if self.ptr:
            lib().gl_dev_free(C.c_void_p(self.ptr))
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def ntt_dev(dst, src, size, ncols, inverse=False, extend=False):
    _check(lib().gl_ntt_dev(C.c_void_p(dst), C.c_void_p(src), size, ncols, int(inverse), int(extend)))


def extend_pol_dev(out, inp, n_ext, n, ncols):
    _check(lib().gl_extend_pol_dev(C.c_void_p(out), C.c_void_p(inp), n_ext, n, ncols))


def extend_pol_dev_noalloc(out, inp, n_ext, n, ncols):
    """gl_extend_pol_dev without any device allocation (static scratch, n_ext <= 2^16)."""
    _check(lib().gl_extend_pol_dev_noalloc(C.c_void_p(out), C.c_void_p(inp), n_ext, n, ncols))


def merkletree_dev(tree, src, ncols, nrows):
    _check(lib().gl_merkletree_dev(C.c_void_p(tree), C.c_void_p(src), ncols, nrows))


def commit_dev(lde, tree, inp, n_ext, n, ncols):
    root = np.zeros(4, dtype=np.uint64)
    _check(lib().gl_commit_dev(C.c_void_p(lde), C.c_void_p(tree), C.c_void_p(inp), n_ext, n, ncols, _p(root)))
    return root


def group_proof_dev(src, tree, height, width, idx):
    proof = np.zeros(proof_size(height, width), dtype=np.uint64)
    _check(lib().gl_group_proof_dev(_p(proof), C.c_void_p(src), C.c_void_p(tree), height, width, idx))
    return proof


def group_proofs_dev(src, tree, height, width, idxs):
    """All queries of one tree in one launch (FRIProve::queryPol, friProve.cpp:219-250)."""
    idxs = np.ascontiguousarray(idxs, dtype=np.uint64)
    proofs = np.zeros((idxs.size, proof_size(height, width)), dtype=np.uint64)
    _check(lib().gl_group_proofs_dev(_p(proofs), C.c_void_p(src), C.c_void_p(tree), height, width, _p(idxs), idxs.size))
    return proofs


def rows_gather_dev(src, width, idxs):
    idxs = np.ascontiguousarray(idxs, dtype=np.uint64)
    rows = np.zeros((idxs.size, width), dtype=np.uint64)
    _check(lib().gl_rows_gather_dev(_p(rows), C.c_void_p(src), width, _p(idxs), idxs.size))
    return rows


def download_cols(dptr, nrows, pitch, c0, w):
    """Columns [c0, c0+w) of a device-resident row-major nrows x pitch matrix as a packed (nrows, w) host array."""
    out = np.empty((nrows, w), dtype=np.uint64)
    _check(lib().gl_copy2d(out.ctypes.data_as(C.c_void_p), w * 8, C.c_void_p(dptr + 8 * c0), pitch * 8, w * 8, nrows, 1))
    return out


def measure_issue_peaks():
    v = (C.c_double * 3)()
    _check(lib().gl_measure_issue_peaks(v))
    return {"imad_wide_lane_ops_per_s": v[0], "iadd3_lane_ops_per_s": v[1], "dfma_lane_ops_per_s": v[2]}


def synthetic_host_cols(nrows, ncols_total, cols, seed=1):
    """Columns `cols` (array of column indices) of the nrows x ncols_total matrix gl_fill_synthetic_dev generates."""
    cols = np.asarray(cols, dtype=np.uint64)
    with np.errstate(over="ignore"):
        i = (np.arange(nrows, dtype=np.uint64)[:, None] * np.uint64(ncols_total) + cols[None, :])
        z = np.uint64(seed) + i * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return np.where(z >= np.uint64(P), z - np.uint64(P), z)


# ---- host-pointer calls (what the C++ shims use) -----------------------------------------------
def ntt_host(src, inverse=False, extend=False):
    src = np.ascontiguousarray(src, dtype=np.uint64)
    size, ncols = src.shape
    dst = np.empty_like(src)
    if size == 0 or ncols == 0:
        return src.copy()
    _check(lib().gl_ntt_host(_p(dst), _p(src), size, ncols, int(inverse), int(extend)))
    return dst


def extend_pol_host(src, n_ext):
    src = np.ascontiguousarray(src, dtype=np.uint64)
    n, ncols = src.shape
    out = np.zeros((n_ext, ncols), dtype=np.uint64)
    _check(lib().gl_extend_pol_host(_p(out), _p(src), n_ext, n, ncols))
    return out


def merkletree_host(src, nrows=None, ncols=None):
    src = np.ascontiguousarray(src, dtype=np.uint64)
    if nrows is None:
        nrows, ncols = src.shape
    tree = np.zeros(tree_num_elements(nrows), dtype=np.uint64)
    buf = src if src.size else np.zeros(1, dtype=np.uint64)
    _check(lib().gl_merkletree_host(_p(tree), _p(buf), ncols, nrows))
    return tree


def merkletree_batch_host(src, batch):
    src = np.ascontiguousarray(src, dtype=np.uint64)
    nrows, ncols = src.shape
    tree = np.zeros(tree_num_elements(nrows), dtype=np.uint64)
    buf = src if src.size else np.zeros(1, dtype=np.uint64)
    _check(lib().gl_merkletree_batch_host(_p(tree), _p(buf), ncols, nrows, batch))
    return tree


def commit_host(src, n_ext, want_lde=True, want_tree=True):
    src = np.ascontiguousarray(src, dtype=np.uint64)
    n, ncols = src.shape
    lde = np.zeros((n_ext, ncols), dtype=np.uint64) if want_lde else None
    tree = np.zeros(tree_num_elements(n_ext), dtype=np.uint64) if want_tree else None
    root = np.zeros(4, dtype=np.uint64)
    _check(lib().gl_commit_host(_p(lde), _p(tree), _p(src), n_ext, n, ncols, _p(root)))
    return lde, tree, root


def hash_full_result_host(inp):
    inp = np.ascontiguousarray(inp, dtype=np.uint64).reshape(-1, 12)
    out = np.empty_like(inp)
    _check(lib().gl_hash_full_result_host(_p(out), _p(inp), inp.shape[0]))
    return out


def linear_hash_host(inp):
    inp = np.ascontiguousarray(inp, dtype=np.uint64).reshape(-1)
    out = np.zeros(4, dtype=np.uint64)
    buf = inp if inp.size else np.zeros(1, dtype=np.uint64)
    _check(lib().gl_linear_hash_host(_p(out), _p(buf), inp.size))
    return out


FIELD_OPS = {"add": 0, "sub": 1, "mul": 2, "mad": 3, "add_c": 4, "sqr": 5, "canon": 6, "add_cx": 7, "sub_cx": 8, "canon_sel": 9, "sqr3": 10}


def field_op_host(op, a, b=None, c=None):
    a = np.ascontiguousarray(a, dtype=np.uint64)
    b = None if b is None else np.ascontiguousarray(b, dtype=np.uint64)
    c = None if c is None else np.ascontiguousarray(c, dtype=np.uint64)
    out = np.empty_like(a)
    _check(lib().gl_field_op_host(_p(out), _p(a), _p(b), _p(c), a.size, FIELD_OPS[op]))
    return out


def fill_synthetic_dev(ptr, words, seed=1, first_index=0):
    _check(lib().gl_fill_synthetic_dev(C.c_void_p(ptr), words, seed, first_index))


def synthetic_host(words, seed=1, first_index=0):
    """numpy twin of gl_fill_synthetic_dev (splitmix64 of the element index, mod p)."""
    with np.errstate(over="ignore"):
        i = np.arange(first_index, first_index + words, dtype=np.uint64)
        z = np.uint64(seed) + i * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return np.where(z >= np.uint64(P), z - np.uint64(P), z)


# ---- multi-GPU (one process per GPU) ---------------------------------------------------------------
def mgpu_unique_id():
    """128-byte NCCL unique id (call on rank 0, broadcast to every rank)."""
    buf = (C.c_uint8 * 128)()
    _check(lib().gl_mgpu_unique_id(buf))
    return bytes(buf)


def mgpu_init(unique_id, rank, world):
    buf = (C.c_uint8 * 128).from_buffer_copy(unique_id)
    _check(lib().gl_mgpu_init(buf, rank, world))


def mgpu_shutdown():
    lib().gl_mgpu_shutdown()


def commit_mgpu(lde_cols, row_panels, subtree, in_cols, n_ext, n, ncols_total):
    root = np.zeros(4, dtype=np.uint64)
    _check(lib().gl_commit_mgpu(C.c_void_p(lde_cols), C.c_void_p(row_panels), C.c_void_p(subtree), C.c_void_p(in_cols),
                                n_ext, n, ncols_total, _p(root)))
    return root


def commit_mgpu_host(row_panels_host, subtree_host, in_cols_host, lde_cols, row_panels, subtree, n_ext, n, ncols_total):
    """Host pointers (ints or None) in/out, device work buffers; pipelined upload / LDE and download / hashing."""
    root = np.zeros(4, dtype=np.uint64)
    _check(lib().gl_commit_mgpu_host(C.c_void_p(row_panels_host), C.c_void_p(subtree_host), C.c_void_p(in_cols_host), C.c_void_p(lde_cols),
                                     C.c_void_p(row_panels), C.c_void_p(subtree), n_ext, n, ncols_total, _p(root)))
    return root


def mgpu_gather_tree(tree_full, subtree, n_ext):
    _check(lib().gl_mgpu_gather_tree(C.c_void_p(tree_full), C.c_void_p(subtree), n_ext))


def fill_synthetic_cols_dev(ptr, nrows, col0, ncols_local, ncols_total, seed=1):
    _check(lib().gl_fill_synthetic_cols_dev(C.c_void_p(ptr), nrows, col0, ncols_local, ncols_total, seed))


# ---- next rows: Q-polynomial commit, FRI layer commit (device-resident) -----------------------------
def q_commit_dev(cm4, tree, q_2ns, n, n_ext, qdim, qdeg):
    root = np.zeros(4, dtype=np.uint64)
    _check(lib().gl_q_commit_dev(C.c_void_p(cm4), C.c_void_p(tree), C.c_void_p(q_2ns), n, n_ext, qdim, qdeg, _p(root)))
    return root


def fri_layer_commit_dev(tree, aux, pol2, pol, pol_bits, next_bits, tree_bits, shift_inv, x):
    root = np.zeros(4, dtype=np.uint64)
    x = np.ascontiguousarray(x, dtype=np.uint64)
    _check(lib().gl_fri_layer_commit_dev(C.c_void_p(tree), C.c_void_p(aux), C.c_void_p(pol2), C.c_void_p(pol), pol_bits, next_bits,
                                         tree_bits, int(shift_inv), _p(x), _p(root)))
    return root


def build_consttree_host(const_pols, n_ext):
    """The .consttree image of tools/starkpil/bctree: [nPols, nExt, extended constants, nodes]."""
    const_pols = np.ascontiguousarray(const_pols, dtype=np.uint64)
    n, npols = const_pols.shape
    out = np.zeros(int(lib().gl_consttree_num_elements(npols, n_ext)), dtype=np.uint64)
    _check(lib().gl_build_consttree_host(_p(out), _p(const_pols), n, n_ext, npols))
    return out
