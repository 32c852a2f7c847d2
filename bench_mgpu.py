"""bench_mgpu.py -- the N>1 leg of bench.py (one process per GPU, launched by torchrun).

The stage commit of the workload is sharded as SURVEY 8e / DESIGN.md "Multi-GPU": columns split over ranks
for the LDE, all-to-all over NVLink (peer copies through the copy engines; grouped ncclSend/ncclRecv as the fallback) to row blocks, per-rank leaf hashing +
subtree, gathered roots -> top levels.  Total work is fixed as N grows ("strong" scaling of BASELINE config 3);
value = E*C / (max over ranks of the device time of one step).
"""
import json
import time

import numpy as np


def run(args, glb, dist, rank, world, local_rank, config, workloads):
    import torch
    from bench import ClockSampler, SEED, leaf_roofline, peaks, permutations
    logn, loge, ncols = workloads[args.workload]
    n, e = 1 << logn, 1 << loge
    dev = torch.device("cuda", local_rank)

    # NCCL communicator of the library: rank 0 creates the id, torch.distributed broadcasts the 128 bytes
    idt = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        idt.copy_(torch.frombuffer(bytearray(glb.mgpu_unique_id()), dtype=torch.uint8))
    dist.broadcast(idt, 0)
    glb.mgpu_init(bytes(idt.cpu().numpy().tobytes()), rank, world)

    L = glb.lib()
    c0, c1 = int(L.gl_mgpu_col_start(ncols, world, rank)), int(L.gl_mgpu_col_start(ncols, world, rank + 1))
    w, rows = c1 - c0, e // world
    din = glb.DeviceBuffer(max(1, n * w))
    dlde = glb.DeviceBuffer(max(1, e * w))
    dpan = glb.DeviceBuffer(rows * ncols)
    dsub = glb.DeviceBuffer(glb.tree_num_elements(rows))
    glb.fill_synthetic_cols_dev(din.ptr, n, c0, w, ncols, SEED)

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        root = glb.commit_mgpu(dlde.ptr, dpan.ptr, dsub.ptr, din.ptr, e, n, ncols)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    glb.kernel_launches(reset=True)
    phase = {"lde_ms": 0.0, "leaf_ms": 0.0, "tree_ms": 0.0, "exchange_ms": 0.0}
    t0 = time.perf_counter()
    for _ in range(args.steps):
        root = glb.commit_mgpu(dlde.ptr, dpan.ptr, dsub.ptr, din.ptr, e, n, ncols)   # synchronous; CUDA events inside
        t = glb.last_timings4()
        for k in phase:
            phase[k] += t[k]
    barrier()
    wall = time.perf_counter() - t0
    launches = glb.kernel_launches()
    clocks = sampler.stop()
    dev_ms = sum(phase.values()) / args.steps
    # max over ranks of the device time, and of every phase
    tt = torch.tensor([dev_ms] + [phase[k] / args.steps for k in ("lde_ms", "exchange_ms", "leaf_ms", "tree_ms")], dtype=torch.float64, device=dev)
    per_rank = [torch.zeros_like(tt) for _ in range(world)]
    dist.all_gather(per_rank, tt.clone())
    per_rank = [[round(float(v), 3) for v in t.tolist()] for t in per_rank]   # [step, lde, exchange, leaf, tree] ms of every rank
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ln = torch.tensor([launches], dtype=torch.int64, device=dev)
    dist.all_reduce(ln, op=dist.ReduceOp.SUM)
    roots = [torch.zeros(4, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(roots, torch.from_numpy(root.view(np.int64)).to(dev))
    same_root = all(bool((r == roots[0]).all()) for r in roots)

    # ---- e2e: host-resident column block in, row block + subtree out (pinned), through the same C-ABI call
    e2e = None
    if not args.no_e2e:
        import ctypes as C
        hin, hrows, hsub = C.c_void_p(), C.c_void_p(), C.c_void_p()
        in_b, rows_b, sub_b = n * w * 8, rows * ncols * 8, glb.tree_num_elements(rows) * 8
        try:
            for hp, sz in ((hin, in_b), (hrows, rows_b), (hsub, sub_b)):
                if L.gl_host_alloc(C.byref(hp), max(sz, 8)) != 0:
                    raise glb.GLError(L.gl_last_error().decode())
            L.gl_d2h(hin, C.c_void_p(din.ptr), in_b)
            ts = []
            for i in range(1 + max(1, min(args.steps, 3))):
                barrier()
                t1 = time.perf_counter()
                r2 = glb.commit_mgpu_host(hrows.value, hsub.value, hin.value, dlde.ptr, dpan.ptr, dsub.ptr, e, n, ncols)
                barrier()
                if i > 0:
                    ts.append(time.perf_counter() - t1)
            assert np.array_equal(r2, root)
            te = torch.tensor([float(np.mean(ts))], dtype=torch.float64, device=dev)
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
            e2e = {"value": e * ncols / float(te.item()), "unit": "ext-elem/s", "ms_per_step": float(te.item()) * 1e3,
                   "h2d_bytes_per_step": n * ncols * 8, "d2h_bytes_per_step": e * ncols * 8 + world * sub_b + 32,
                   "call": "per rank: gl_commit_mgpu_host(row panels out, subtree out, column block in): sub-block upload overlaps the LDE, panel download overlaps the leaf hashing; pinned host buffers; bytes summed over ranks"}
        except Exception as ex:
            e2e = {"value": None, "unit": "ext-elem/s", "error": str(ex)[:200], "h2d_bytes_per_step": n * ncols * 8, "d2h_bytes_per_step": e * ncols * 8}
        finally:
            for hp in (hin, hrows, hsub):
                if hp.value:
                    L.gl_host_free(hp)

    exchange_mode = glb.lib().gl_mgpu_exchange_mode()
    glb.mgpu_shutdown()
    if rank != 0:
        return 0
    # (rank 0 measures the issue-rate probes for the roofline after the communicator is gone)
    dev_ms, lde_ms, ex_ms, leaf_ms, tree_ms = [float(x) for x in tt.tolist()]
    hbm_peak, peak_src = peaks()
    leaf_perm, tree_perm = permutations(loge, ncols)
    leaf_bytes = (e * ncols * 8 + e * 32) / world
    # mgpu_exchange_and_hash in csrc/glb200.cu: chunks of 1, 3, 9, ... waves of the leaf kernel (148 SMs x 4 CTAs x 128 rows); the last takes
    # the rest once that is no more than three times its predecessor
    wave, size, done, chunks = 148 * 4 * 128, 148 * 4 * 128, 0, 0
    while chunks < 7 and rows > wave * 4 and rows - done > size:
        done, size, chunks = done + size, size * 3, chunks + 1
    chunks += 1
    first_frac = wave / rows if chunks > 1 else 1.0
    config = dict(config)
    config["parallelism"] = "column-split LDE x%d -> all-to-all over NVLink (copy engines) -> row-block leaf hash + subtrees -> gathered root" % world
    line = {"metric": "commit_ext_elements_per_s", "value": e * ncols / (dev_ms / 1e3), "unit": "ext-elem/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms, "wall_ms_per_step": wall * 1e3 / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u64 (Goldilocks field, integer pipe)", "data": "synthetic (splitmix64 of the element index mod p, generated in HBM)",
            "config": config, "root": " ".join("%016X" % int(v) for v in root), "roots_agree": same_root, "clocks": clocks, "e2e": e2e,
            "gpu_launches": int(ln.item()),
            "roofline": leaf_roofline(glb, leaf_perm // world, leaf_ms, leaf_ms / dev_ms, leaf_bytes, clocks, hbm_peak, peak_src,
                                      "poseidon::k_linear_hash_rows<panels>, per GPU"),
            "ms_per_rank": {"columns": ["step", "lde", "exchange", "leaf", "tree"], "rows": per_rank},
            "phases": {"lde": {"ms": lde_ms},
                       "exchange": {"ms": ex_ms, "bytes_sent_per_rank": (world - 1) * rows * w * 8, "row_chunks": chunks,
                                    "transport": ("peer cudaMemcpyAsync over NVLink through the copy engines into CUDA-IPC mappings of the peers' panels, one-word ncclAllGather as the barrier"
                                                  if exchange_mode == 1 else "grouped ncclSend/ncclRecv"),
                                    "note": "exposed part only: the rows are exchanged in %d chunks of growing size (first %.4f of the rows) on a second stream; chunk j+1 moves while chunk j is hashed" % (chunks, first_frac),
                                    "GBps_per_rank_first_chunk": (world - 1) * rows * w * 8 * first_frac / max(ex_ms, 1e-9) / 1e6},
                       "leaf": {"ms": leaf_ms, "permutations_per_rank": leaf_perm // world}, "tree": {"ms": tree_ms}},
            "cpu_baseline": None}
    print(json.dumps(line))
    return 0
