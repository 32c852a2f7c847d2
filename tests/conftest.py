import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
P = 0xFFFFFFFF00000001


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle
    return pyoracle.oracle()


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference (oracle/_ref) or None when it was never built."""
    from oracle import pyoracle
    return pyoracle.ref()


@pytest.fixture(scope="session")
def kats():
    with open(os.path.join(GOLDEN, "reference_kats.json")) as f:
        raw = json.load(f)
    return {k: np.array([int(x, 16) for x in v["values"]], dtype=np.uint64) for k, v in raw.items()}


@pytest.fixture(scope="session")
def fixtures():
    fx = dict(np.load(os.path.join(GOLDEN, "ref_fixtures.npz")))
    with open(os.path.join(GOLDEN, "ref_fixtures.json")) as f:
        meta = json.load(f)
    # regenerate the inputs exactly as tests/golden/make_golden.py drew them
    rng = np.random.default_rng(meta["seed"])
    inputs = {}
    for (n, c) in meta["ntt_shapes"]:
        inputs["ntt_%d_%d" % (n, c)] = rng.integers(0, P, size=(n, c), dtype=np.uint64)
    for (n, e, c) in meta["lde"]:
        inputs["lde_%d_%d_%d" % (n, e, c)] = rng.integers(0, P, size=(n, c), dtype=np.uint64)
    for (h, w) in meta["tree"]:
        inputs["tree_%d_%d" % (h, w)] = rng.integers(0, P, size=(h, w), dtype=np.uint64)
    rng2 = np.random.default_rng(meta.get("seed2", 0))
    for (n, e, qdeg) in meta.get("q_extend", []):
        inputs["qext_%d_%d_%d" % (n, e, qdeg)] = rng2.integers(0, P, size=(e, 3), dtype=np.uint64)
    for (pb, nb) in meta.get("fri_fold", []):
        pol = rng2.integers(0, P, size=(1 << pb, 3), dtype=np.uint64)
        x = rng2.integers(0, P, size=3, dtype=np.uint64)
        si = int(rng2.integers(1, P, dtype=np.uint64))
        inputs["frifold_%d_%d" % (pb, nb)] = (pol, x, si)
    rng3 = np.random.default_rng(meta.get("seed3", 0))
    for (n, e, c) in meta.get("bctree", []):
        inputs["bctree_%d_%d_%d" % (n, e, c)] = rng3.integers(0, P, size=(n, c), dtype=np.uint64)
    return {"out": fx, "in": inputs, "meta": meta}


@pytest.fixture(scope="session")
def glb():
    """The product package (ctypes over libglb200.so), bound to cuda:0.  No fallback: if the CUDA
    library is missing or no device is usable this raises, it does not skip."""
    import __graft_entry__ as ge
    mod = ge.load()
    mod.init(0)
    return mod


def fib_rows(nrows, ncols):
    """rows 0,1 = 1..ncols; row j = row j-1 + row j-2 (reference tests.cpp:1822-1836)."""
    a = np.zeros((nrows, ncols), dtype=object)
    a[0, :] = a[1, :] = [i + 1 for i in range(ncols)]
    for j in range(2, nrows):
        a[j, :] = (a[j - 1, :] + a[j - 2, :]) % P
    return a.astype(np.uint64)


def fib_vec(n, a0=0, a1=1):
    v = [a0, a1]
    while len(v) < n:
        v.append((v[-1] + v[-2]) % P)
    return np.array(v[:n], dtype=np.uint64)


def hexs(a):
    return " ".join("%016X" % int(x) for x in np.asarray(a).reshape(-1))
