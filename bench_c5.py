"""bench_c5.py -- BASELINE configs[4]: the full commit sequence of one synthetic batch proof on one B200.

Sequence (Starks::genProof, src/starkpil/starks.cpp:52-56, 133-137, 214-218, 261-290; FRIProve::prove,
src/starkpil/fri/friProve.cpp:20-149), every buffer resident in HBM:
  stage 1/2/3 : extendPol (in place: the trace sits in the first N rows of the extended buffer) + merkelize
  Q           : INTT(2^24 x 3), coset un-shift + regroup, NTT(2^24 x 3*qDeg), merkelize
  FRI         : per step fold in F_p^3 (challenge from a fixed synthetic transcript), getTransposed, merkelize
Column counts are synthetic (the real zkevm.starkinfo.json is an LFS blob missing from the reference tree):
stage-1 600, stage-2 48, stage-3 96, qDeg 2, FRI steps nBits [24, 19, 14, 10, 6].
Launched through bench.py --workload c5; prints the same JSON schema with per-stage phases.
"""
import json
import time

import numpy as np

STAGES = [("stage1", 600), ("stage2", 48), ("stage3", 96)]
QDIM, QDEG = 3, 2
FRI_STEPS = [24, 19, 14, 10, 6]
P = 0xFFFFFFFF00000001


def run(args, glb, config, peaks, ClockSampler, seed=1):
    logn, loge = 23, 24
    n, e = 1 << logn, 1 << loge
    bufs = {name: glb.DeviceBuffer(e * c) for name, c in STAGES}
    trees = {name: glb.DeviceBuffer(glb.tree_num_elements(e)) for name, _ in STAGES}
    dq = glb.DeviceBuffer(e * QDIM)
    dcm4 = glb.DeviceBuffer(e * QDIM * QDEG)
    tq = glb.DeviceBuffer(glb.tree_num_elements(e))
    fri_pol = glb.DeviceBuffer(e * 3)
    fri_tmp = [glb.DeviceBuffer((1 << b) * 3) for b in FRI_STEPS[1:]]
    fri_aux = [glb.DeviceBuffer((1 << b) * 3) for b in FRI_STEPS[1:]]
    fri_tree = [glb.DeviceBuffer(glb.tree_num_elements(1 << b)) for b in FRI_STEPS[2:]]
    x = np.array([0x1234567, 0x89ABCDEF01, 0x13579BDF], dtype=np.uint64)
    shift_inv = pow(7, P - 2, P)

    def fill():
        for i, (name, c) in enumerate(STAGES):
            glb.fill_synthetic_dev(bufs[name].ptr, n * c, seed + i)
        glb.fill_synthetic_dev(dq.ptr, e * QDIM, seed + 10)
        glb.fill_synthetic_dev(fri_pol.ptr, e * 3, seed + 11)

    def step():
        ph, roots = {}, []
        for name, c in STAGES:
            roots.append(glb.commit_dev(bufs[name].ptr, trees[name].ptr, bufs[name].ptr, e, n, c))
            t = glb.last_timings()
            ph[name] = {"cols": c, "lde_ms": t["lde_ms"], "leaf_ms": t["leaf_ms"], "tree_ms": t["tree_ms"]}
        roots.append(glb.q_commit_dev(dcm4.ptr, tq.ptr, dq.ptr, n, e, QDIM, QDEG))
        t = glb.last_timings()
        ph["q"] = {"cols": QDIM * QDEG, "lde_ms": t["lde_ms"], "leaf_ms": t["leaf_ms"], "tree_ms": t["tree_ms"]}
        # FRI: step 0 of the reference is a copy (reductionBits 0); steps 1.. fold the previous polynomial
        t0 = time.perf_counter()
        src, si = fri_pol, shift_inv
        for k in range(1, len(FRI_STEPS)):
            pb, nb = FRI_STEPS[k - 1], FRI_STEPS[k]
            if k < len(FRI_STEPS) - 1:
                roots.append(glb.fri_layer_commit_dev(fri_tree[k - 1].ptr, fri_aux[k - 1].ptr, fri_tmp[k - 1].ptr, src.ptr, pb, nb, FRI_STEPS[k + 1], si, x))
            else:
                rc = glb.lib().gl_fri_fold_dev(fri_tmp[k - 1].ptr, src.ptr, pb, nb, si, x.ctypes.data_as(glb._u64p))
                assert rc == 0
                glb.lib().gl_sync()
            src = fri_tmp[k - 1]
            for _ in range(pb - nb):
                si = si * si % P
        ph["fri"] = {"steps": FRI_STEPS, "wall_ms": (time.perf_counter() - t0) * 1e3}
        return ph, roots

    for _ in range(args.warmup):
        fill()
        step()
    glb.lib().gl_sync()
    sampler = ClockSampler(0)
    sampler.start()
    glb.kernel_launches(reset=True)
    tot, last = 0.0, None
    for _ in range(args.steps):
        fill()
        glb.lib().gl_sync()
        launches0 = glb.kernel_launches()
        t0 = time.perf_counter()
        last, roots = step()
        glb.lib().gl_sync()
        tot += time.perf_counter() - t0
    launches = glb.kernel_launches() - args.steps * (len(STAGES) + 2)   # minus the synthetic fills
    clocks = sampler.stop()
    ms = tot * 1e3 / args.steps
    elems = e * (sum(c for _, c in STAGES) + QDIM * QDEG) + sum((1 << b) * 3 for b in FRI_STEPS[1:])
    hbm_peak, peak_src = peaks()
    dev_ms = sum(v.get("lde_ms", 0) + v.get("leaf_ms", 0) + v.get("tree_ms", 0) for v in last.values()) + last["fri"]["wall_ms"]
    leaf_ms = sum(v.get("leaf_ms", 0) for v in last.values())
    leaf_bytes = e * 8 * (sum(c for _, c in STAGES) + QDIM * QDEG) + 4 * e * 32
    from bench import leaf_roofline
    leaf_perms = e * (sum((c + 7) // 8 for _, c in STAGES) + (QDIM * QDEG + 7) // 8)
    config = dict(config)
    config["workload"] = "c5: full commit sequence of one synthetic batch proof: stage-1/2/3 (600/48/96 cols) + Q (3 x qDeg 2) + FRI layers %s, 2^23 -> 2^24" % FRI_STEPS
    config["ncols"] = sum(c for _, c in STAGES) + QDIM * QDEG
    line = {"metric": "commit_ext_elements_per_s", "value": elems / (ms / 1e3), "unit": "ext-elem/s", "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "device_ms_per_step": dev_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64 (Goldilocks field, integer pipe)", "data": "synthetic (splitmix64 of the element index mod p, generated in HBM)",
            "config": config, "roots": [" ".join("%016X" % int(v) for v in r) for r in roots], "clocks": clocks, "e2e": None,
            "gpu_launches": launches,
            "roofline": leaf_roofline(glb, leaf_perms, leaf_ms, leaf_ms / dev_ms, leaf_bytes, clocks, hbm_peak, peak_src,
                                      "poseidon::k_linear_hash_rows over the 4 stage trees"),
            "phases": last, "cpu_baseline": None}
    print(json.dumps(line))
    return 0
