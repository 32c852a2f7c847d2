"""gl_extend_pol_dev_noalloc (allocation-free extendPol, static device scratch) against gl_extend_pol_dev, through the C ABI.
The entry point has no counterpart in the reference; its contract is "same result
as gl_extend_pol_dev", bit-exact like every other parity test here.  Collected last on purpose (file name).
Passes alone in a fresh process and after the other tests (it does not depend on the state they leave: the static
array keeps whatever the previous call wrote, and every call rewrites what it reads)."""
import numpy as np
import pytest

from conftest import P

pytestmark = pytest.mark.gpu


def _both(glb, x, n_ext):
    n, ncols = x.shape
    src = glb.DeviceBuffer(n * ncols).upload(x)
    a = glb.DeviceBuffer(n_ext * ncols)
    b = glb.DeviceBuffer(n_ext * ncols)
    glb.extend_pol_dev(a.ptr, src.ptr, n_ext, n, ncols)
    glb.extend_pol_dev_noalloc(b.ptr, src.ptr, n_ext, n, ncols)
    want = a.download().reshape(n_ext, ncols)
    got = b.download().reshape(n_ext, ncols)
    for buf in (src, a, b):
        buf.free()
    return want, got


def test_noalloc_extend_pol_matches_twice_in_a_row(glb, oracle):
    rng = np.random.default_rng(77)
    # two calls in a row on DIFFERENT shapes (the second one finds the first one's data in the shared array: wider blocks, another n,
    # another scale table), then the first shape again
    shapes = [(1 << 12, 1 << 13, 40), (1 << 10, 1 << 14, 37), (1 << 12, 1 << 13, 40)]
    first = None
    for k, (n, n_ext, ncols) in enumerate(shapes):
        x = rng.integers(0, P, size=(n, ncols), dtype=np.uint64) if k < 2 else first[0]
        want, got = _both(glb, x, n_ext)
        assert np.array_equal(want, got), (n, n_ext, ncols)
        if k == 0:
            first = (x, want)
            assert np.array_equal(got[:, :3], oracle.extend_pol(np.ascontiguousarray(x[:, :3]), n_ext))
        if k == 2:
            assert np.array_equal(got, first[1])


def test_noalloc_extend_pol_sizes_and_limit(glb):
    rng = np.random.default_rng(78)
    # tiny, one-pass, two-pass and the largest supported domain (several column blocks: 2^22 / (2^15 + 2^16) = 42 -> 32 columns per block)
    for (n, n_ext, ncols) in [(1, 1, 3), (2, 8, 5), (64, 256, 9), (1 << 9, 1 << 10, 1), (1 << 15, 1 << 16, 70), (1 << 16, 1 << 16, 33)]:
        x = rng.integers(0, 2**64 - 1, size=(n, ncols), dtype=np.uint64, endpoint=True)  # non-canonical words too
        want, got = _both(glb, x, n_ext)
        assert np.array_equal(want, got), (n, n_ext, ncols)
    # beyond the static scratch: a clear error, nothing written
    n, n_ext, ncols = 1 << 16, 1 << 17, 2
    src = glb.DeviceBuffer(n * ncols).upload(np.zeros(n * ncols, dtype=np.uint64))
    dst = glb.DeviceBuffer(n_ext * ncols)
    with pytest.raises(glb.GLError) as e:
        glb.extend_pol_dev_noalloc(dst.ptr, src.ptr, n_ext, n, ncols)
    assert "static scratch" in str(e.value) and "2^16" in str(e.value)
    src.free()
    dst.free()
