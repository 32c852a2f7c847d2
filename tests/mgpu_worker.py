"""Worker of the multi-rank tests (launched with torch.distributed.run by test_shard_plan.py / test_gpu_mgpu.py).

  mode "gloo": CPU only.  Every rank runs the ORACLE on its column block, the exchange goes through
               torch.distributed (gloo send/recv), the shard plan (tests/shard_plan.py)
               assembles the tree; checked against the single-process oracle.  Tests the host-side logic.
  mode "nccl": one GPU per rank.  gl_commit_mgpu / gl_mgpu_gather_tree through the C ABI, checked against the oracle.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402
from oracle import pyoracle  # noqa: E402

SHAPES = [(1 << 10, 1 << 11, 37), (1 << 8, 1 << 12, 9), (1 << 6, 1 << 7, 3), (1 << 9, 1 << 10, 64)]


def top_levels(o, tree, n_ext, first_level, sp):
    lv, cnt = first_level, n_ext >> first_level
    while cnt >= 1:
        src = tree[sp.level_offset(n_ext, lv - 1):sp.level_offset(n_ext, lv - 1) + cnt * 2 * 4].reshape(cnt, 8)
        for i in range(cnt):
            st = o.hash_full_result(np.concatenate([src[i], np.zeros(4, dtype=np.uint64)]))
            tree[sp.level_offset(n_ext, lv) + 4 * i:sp.level_offset(n_ext, lv) + 4 * i + 4] = st[:4]
        if cnt == 1:
            break
        cnt >>= 1
        lv += 1
    return tree


# root of the c2 stage commit (2^23 x 600 -> 2^24, gl_fill_synthetic seed 1): pinned on one GPU by
# tests/test_gpu_fullsize.py::test_c2_lde_sampled_columns_and_tree_paths (sampled LDE columns against the reference, sampled
# authentication paths re-hashed on the CPU up to this root)
C2_ROOT = [0xF68B21F12F746043, 0x1173C26EB33074DB, 0x5D94C8FA42CEBCC0, 0x2C3189EE0E59DD6B]


def full_size_check(glb, o, rank, world):
    """BASELINE configs[2]: the c2 commit sharded over `world` GPUs.  Every word of this rank's first 8 LDE columns against the CPU
    reference run on exactly those columns, this rank's received row panels against its own LDE columns of the same rows, and the
    root against the single-GPU one."""
    logn, loge, ncols = 23, 24, 600
    n, e = 1 << logn, 1 << loge
    L = glb.lib()
    c0, c1 = int(L.gl_mgpu_col_start(ncols, world, rank)), int(L.gl_mgpu_col_start(ncols, world, rank + 1))
    w, rows = c1 - c0, e // world
    din = glb.DeviceBuffer(n * w)
    dlde = glb.DeviceBuffer(e * w)
    dpan = glb.DeviceBuffer(rows * ncols)
    dsub = glb.DeviceBuffer(glb.tree_num_elements(rows))
    glb.fill_synthetic_cols_dev(din.ptr, n, c0, w, ncols, 1)
    root = glb.commit_mgpu(dlde.ptr, dpan.ptr, dsub.ptr, din.ptr, e, n, ncols)
    assert [int(v) for v in root] == C2_ROOT, "c2 root on %d GPUs: %s" % (world, " ".join("%016X" % int(v) for v in root))
    k = min(8, w)
    got = glb.download_cols(dlde.ptr, e, w, 0, k)
    cpu = pyoracle.ref() or o
    x = np.ascontiguousarray(glb.synthetic_host_cols(n, ncols, np.arange(c0, c0 + k), 1))
    assert np.array_equal(got, cpu.extend_pol(x, e)), "c2 LDE columns of rank %d" % rank
    # my rows of my own columns came back through the exchange unchanged (panel `rank`)
    off = rows * c0
    pan = dpan.download(rows * w, off).reshape(rows, w)
    mine = glb.download_cols(dlde.ptr + 8 * (rank * rows * w), rows, w, 0, w)
    assert np.array_equal(pan, mine), "row panel of rank %d" % rank


def main():
    mode = sys.argv[1]
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    glb = ge.load()
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import shard_plan as sp
    o = pyoracle.oracle()
    if mode == "gloo":
        dist.init_process_group("gloo")
    else:
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        glb.init(local_rank)
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(glb.mgpu_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        glb.mgpu_init(idt.cpu().numpy().tobytes(), rank, world)

    os.environ["GLB_MGPU_WAVE_ROWS"] = "16"   # small shapes: make the exchange run in several row chunks, as it does at full size
    for (n, e, ncols) in SHAPES:
        x, _ = o.xorshift(n * ncols)
        x = x.reshape(n, ncols)
        lde_full = o.extend_pol(x, e)
        tree_full = o.merkletree(lde_full)
        c0, c1 = sp.col_range(ncols, world, rank)
        r0, r1 = sp.row_range(e, world, rank)
        rows = e // world
        if mode == "gloo":
            mine = o.extend_pol(np.ascontiguousarray(x[:, c0:c1]), e) if c1 > c0 else np.zeros((e, 0), dtype=np.uint64)
            assert np.array_equal(mine, lde_full[:, c0:c1])
            chunks = sp.send_chunks(mine, world)
            panels = [None] * world
            reqs = []
            for p in range(world):
                pc0, pc1 = sp.col_range(ncols, world, p)
                if p == rank:
                    panels[p] = chunks[p]
                    continue
                panels[p] = np.zeros((rows, pc1 - pc0), dtype=np.uint64)
                if c1 > c0:
                    reqs.append(dist.isend(torch.from_numpy(chunks[p].view(np.int64)), p))
                if pc1 > pc0:
                    reqs.append(dist.irecv(torch.from_numpy(panels[p].view(np.int64)), p))
            for r in reqs:
                r.wait()
            myrows = sp.rows_from_panels(panels)
            assert np.array_equal(myrows, lde_full[r0:r1])
            sub = o.merkletree(myrows)
            subs = [torch.zeros(sub.size, dtype=torch.int64) for _ in range(world)]
            dist.all_gather(subs, torch.from_numpy(sub.view(np.int64)))
            subs = [s.numpy().view(np.uint64) for s in subs]
            tree, first = sp.assemble_tree(subs, e)
            tree = top_levels(o, tree, e, first, sp) if world > 1 else tree
            assert np.array_equal(tree, tree_full), "assembled tree differs (%d,%d,%d)" % (n, e, ncols)
        else:
            w = c1 - c0
            din = glb.DeviceBuffer(max(1, n * w)).upload(np.ascontiguousarray(x[:, c0:c1]))
            dlde = glb.DeviceBuffer(max(1, e * w))
            dpan = glb.DeviceBuffer(max(1, rows * ncols))
            dsub = glb.DeviceBuffer(glb.tree_num_elements(rows))
            dfull = glb.DeviceBuffer(glb.tree_num_elements(e))
            root = glb.commit_mgpu(dlde.ptr, dpan.ptr, dsub.ptr, din.ptr, e, n, ncols)
            assert np.array_equal(root, tree_full[-4:]), "root differs (%d,%d,%d)" % (n, e, ncols)
            # default transport: peer copies over NVLink through the copy engines (CUDA IPC mappings of the peers' panels)
            assert glb.lib().gl_mgpu_exchange_mode() == (1 if world > 1 and os.environ.get("GLB_MGPU_EXCHANGE") != "nccl" else 0), "exchange fell back to NCCL send/recv"
            assert np.array_equal(dlde.download(e * w).reshape(e, w), lde_full[:, c0:c1])
            pan = dpan.download(rows * ncols)
            for p, (off, pw) in enumerate(sp.panel_offsets(e, ncols, world)):
                pc0 = sp.col_start(ncols, world, p)
                assert np.array_equal(pan[off:off + rows * pw].reshape(rows, pw), lde_full[r0:r1, pc0:pc0 + pw])
            glb.mgpu_gather_tree(dfull.ptr, dsub.ptr, e)
            assert np.array_equal(dfull.download(), tree_full), "gathered tree differs"
            # the host-pointer (pipelined) entry point must produce the same panels, subtree and root
            sub_dev = dsub.download()
            hin = np.ascontiguousarray(x[:, c0:c1]) if w else np.zeros(1, dtype=np.uint64)
            hpan = np.zeros(max(1, rows * ncols), dtype=np.uint64)
            hsub = np.zeros(glb.tree_num_elements(rows), dtype=np.uint64)
            dlde2 = glb.DeviceBuffer(max(1, e * w))
            dpan2 = glb.DeviceBuffer(max(1, rows * ncols))
            dsub2 = glb.DeviceBuffer(glb.tree_num_elements(rows))
            root2 = glb.commit_mgpu_host(hpan.ctypes.data, hsub.ctypes.data, hin.ctypes.data, dlde2.ptr, dpan2.ptr, dsub2.ptr, e, n, ncols)
            assert np.array_equal(root2, tree_full[-4:]), "host-pointer root differs (%d,%d,%d)" % (n, e, ncols)
            assert np.array_equal(hpan[:rows * ncols], pan), "host-pointer panels differ"
            assert np.array_equal(hsub, sub_dev), "host-pointer subtree differs"
        dist.barrier()
    if mode == "nccl" and world > 1 and os.environ.get("GLB_MGPU_EXCHANGE") != "nccl":
        # the other transport (grouped ncclSend/ncclRecv, what a rank falls back to when a buffer cannot be mapped): new communicator
        glb.mgpu_shutdown()
        os.environ["GLB_MGPU_EXCHANGE"] = "nccl"
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(glb.mgpu_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        glb.mgpu_init(idt.cpu().numpy().tobytes(), rank, world)
        for (n, e, ncols) in SHAPES[:2]:
            x, _ = o.xorshift(n * ncols)
            x = x.reshape(n, ncols)
            tree_full = o.merkletree(o.extend_pol(x, e))
            c0, c1 = sp.col_range(ncols, world, rank)
            w, rows = c1 - c0, e // world
            din = glb.DeviceBuffer(max(1, n * w)).upload(np.ascontiguousarray(x[:, c0:c1]))
            dlde, dpan, dsub = glb.DeviceBuffer(max(1, e * w)), glb.DeviceBuffer(max(1, rows * ncols)), glb.DeviceBuffer(glb.tree_num_elements(rows))
            root = glb.commit_mgpu(dlde.ptr, dpan.ptr, dsub.ptr, din.ptr, e, n, ncols)
            assert glb.lib().gl_mgpu_exchange_mode() == 0
            assert np.array_equal(root, tree_full[-4:]), "NCCL-exchange root differs (%d,%d,%d)" % (n, e, ncols)
        dist.barrier()
        glb.mgpu_shutdown()
        os.environ.pop("GLB_MGPU_EXCHANGE")
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(glb.mgpu_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        glb.mgpu_init(idt.cpu().numpy().tobytes(), rank, world)
    os.environ.pop("GLB_MGPU_WAVE_ROWS", None)
    if mode == "nccl" and os.environ.get("GLB_MGPU_FULL") == "1":
        full_size_check(glb, o, rank, world)
        dist.barrier()
    if mode == "nccl":
        glb.mgpu_shutdown()
    dist.destroy_process_group()
    print("mgpu_worker %s rank %d/%d ok" % (mode, rank, world))


if __name__ == "__main__":
    main()
