"""TEST SUPPORT (not part of the product package): the host-side sharding plan of the multi-GPU commit restated as
pure index arithmetic, no device code.

The same formulas are implemented in csrc/glb200.cu (col_start, level_offset, the panel layout of
gl_commit_mgpu); tests/test_shard_plan.py checks them on CPU with a world_size-2 gloo group by running
the exchange on oracle data and comparing the assembled tree with the single-process result.
"""
import numpy as np


def col_start(ncols, world, rank):
    """First column owned by `rank` (rank r owns [col_start(r), col_start(r+1))): contiguous, balanced in units of 8 columns (the sponge
    rate) so that no 8-word chunk of a row straddles two panels; the ncols % 8 left-over columns go to the last rank; balanced in
    single columns when there are fewer than 8 per rank.  Twin of col_start in csrc/glb200.cu."""
    if rank >= world:
        return ncols
    chunks = ncols // 8
    if chunks < world:
        return ncols * rank // world
    return 8 * (chunks * rank // world)


def col_range(ncols, world, rank):
    return col_start(ncols, world, rank), col_start(ncols, world, rank + 1)


def row_range(n_ext, world, rank):
    rows = n_ext // world
    return rank * rows, (rank + 1) * rows


def level_offset(height, level):
    """Word offset of tree level `level` (0 = leaves) in the MerkleTreeGL::nodes layout of a tree over
    `height` leaves (leaves first, level by level; merklehash_goldilocks.hpp:25-28)."""
    return sum((height >> q) * 4 for q in range(level))


def send_chunks(lde_cols, world):
    """Split this rank's column block (n_ext x w) into the `world` row chunks it sends (chunk h -> rank h)."""
    n_ext = lde_cols.shape[0]
    rows = n_ext // world
    return [np.ascontiguousarray(lde_cols[h * rows:(h + 1) * rows]) for h in range(world)]


def panel_offsets(n_ext, ncols, world):
    """Word offset and width of each received panel inside a rank's row_panels buffer."""
    rows = n_ext // world
    return [(rows * col_start(ncols, world, p), col_start(ncols, world, p + 1) - col_start(ncols, world, p)) for p in range(world)]


def rows_from_panels(panels):
    """Row-major (rows x ncols) view of what the panel-reading leaf kernel hashes."""
    return np.concatenate(panels, axis=1)


def assemble_tree(subtrees, n_ext):
    """Full node array from the per-rank subtrees (rank order), minus the top levels which the caller hashes:
    returns (tree, first_missing_level).  Mirrors gl_mgpu_gather_tree."""
    world = len(subtrees)
    rows = n_ext // world
    tree = np.zeros((2 * n_ext - 1) * 4, dtype=np.uint64)
    lv, cnt = 0, rows
    while True:
        for r, st in enumerate(subtrees):
            src = st[level_offset(rows, lv):level_offset(rows, lv) + cnt * 4]
            dst = level_offset(n_ext, lv) + r * cnt * 4
            tree[dst:dst + cnt * 4] = src
        if cnt == 1:
            break
        cnt >>= 1
        lv += 1
    return tree, lv + 1
