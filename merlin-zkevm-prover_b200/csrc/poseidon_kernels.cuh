// poseidon_kernels.cuh -- leaf (row sponge) and tree-level kernels + launchers.
//
//   k_linear_hash_rows : PoseidonGoldilocks::linear_hash over every row   (poseidon_goldilocks.cpp:229-265, loop :277-281)
//   k_tree_level       : parent = hash(left || right || 0,0,0,0)[0..4)     (poseidon_goldilocks.cpp:287-300)
//   k_hash_full_result : batched permutation (KAT entry point)             (poseidon_goldilocks.cpp:166-228)
#pragma once
#include "poseidon.cuh"
#include <cuda_runtime.h>

namespace poseidon {

constexpr int kThreads = 128;
// 4 CTAs of 128 threads per SM (<= 128 registers): the permutation with the FP64 MDS needs ~120 registers to stay spill-free;
// tools/ubench leaf: 1023 M perm/s against 1005 (5 CTAs, 96 registers, spills) and 992 (6 CTAs)
#ifndef GLB_LEAF_MINBLOCKS
#define GLB_LEAF_MINBLOCKS 4
#endif

__global__ void __launch_bounds__(kThreads) k_hash_full_result(u64 *__restrict__ out, const u64 *__restrict__ in, u64 n)
{
    u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n) return;
    u64 s[12];
#pragma unroll
    for (int i = 0; i < 12; i++) s[i] = in[idx * 12 + i];
    permute(s);
#pragma unroll
    for (int i = 0; i < 12; i++) out[idx * 12 + i] = gl::canon(s[i]);
}

// Where the words of a row live.  Single GPU: one panel = the row-major matrix.  Multi-GPU (after the
// column-split -> row-block exchange): panel g holds columns [start[g], start[g+1]) of this
// rank's rows, row-major with pitch start[g+1]-start[g], as received from rank g.
constexpr int kMaxPanels = 8;
struct RowPanels {
    const u64 *base[kMaxPanels];
    u32 start[kMaxPanels + 1];
    u32 npanels;
};

template <bool PANELS>
__device__ __forceinline__ u64 row_word(const RowPanels &rp, const u64 *p0, u64 row, u32 col)
{
    if (!PANELS) return p0[col];
    u32 g = 0;
    while (g + 1 < rp.npanels && col >= rp.start[g + 1]) g++; // warp-uniform
    const u32 w = rp.start[g + 1] - rp.start[g];
    return rp.base[g][row * w + (col - rp.start[g])];
}

// PoseidonGoldilocks::linear_hash of one row (poseidon_goldilocks.cpp:229-265): sponge, rate 8, capacity 4.
// ncols <= 4: the row is returned zero-padded WITHOUT hashing (reference :234-239).
template <bool PANELS>
__device__ __forceinline__ void linear_hash_row(u64 *d, const RowPanels &rp, const u64 *p, u64 row, u32 ncols)
{
    if (ncols <= 4) {
#pragma unroll
        for (u32 i = 0; i < 4; i++) d[i] = i < ncols ? gl::canon(row_word<PANELS>(rp, p, row, i)) : 0;
        return;
    }
    u64 s[12];
#pragma unroll
    for (int i = 8; i < 12; i++) s[i] = 0;
    const u32 full = ncols >> 3;
    const u32 rem = ncols & 7;
    if (PANELS) {
        // Stream the row panel by panel: `q` walks the current panel's copy of this row, `left` words remain in it.  A chunk
        // that lies inside one panel is eight independent loads (the common case); only chunks that straddle a panel
        // boundary take the word-by-word path.  All control flow is warp-uniform.
        u32 g = 0, left = rp.start[1] - rp.start[0];
        const u64 *q = rp.base[0] + row * left;
        auto next_word = [&]() {
            while (left == 0 && g + 1 < rp.npanels) {
                g++;
                left = rp.start[g + 1] - rp.start[g];
                q = rp.base[g] + row * left;
            }
            left--;
            return *q++;
        };
        for (u32 c = 0; c < full; c++) {
            if (left >= 8) {
#pragma unroll
                for (u32 i = 0; i < 8; i++) s[i] = q[i];
                q += 8;
                left -= 8;
            } else {
#pragma unroll
                for (u32 i = 0; i < 8; i++) s[i] = next_word();
            }
            permute<true>(s);
#pragma unroll
            for (int i = 0; i < 4; i++) s[8 + i] = s[i];
        }
        if (rem) {
#pragma unroll
            for (u32 i = 0; i < 8; i++) s[i] = i < rem ? next_word() : 0;
            permute<true>(s);
        }
    } else {
        for (u32 c = 0; c < full; c++) {
#pragma unroll
            for (u32 i = 0; i < 8; i++) s[i] = p[c * 8 + i];
            permute<true>(s);
#pragma unroll
            for (int i = 0; i < 4; i++) s[8 + i] = s[i];
        }
        if (rem) {
#pragma unroll
            for (u32 i = 0; i < 8; i++) s[i] = i < rem ? p[full * 8 + i] : 0;
            permute<true>(s);
        }
    }
#pragma unroll
    for (int i = 0; i < 4; i++) d[i] = gl::canon(s[i]);
}

// One thread per row: leaf digests of MerkleTreeGL::merkelize (poseidon_goldilocks.cpp:277-281).
template <bool PANELS>
__global__ void __launch_bounds__(kThreads, GLB_LEAF_MINBLOCKS) k_linear_hash_rows(u64 *__restrict__ digests, const __grid_constant__ RowPanels rp, u64 nrows, u32 ncols)
{
    u64 row = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
#ifdef GLB_LEAF_STAGGER
    // Round-2 experiment (off by default): every row costs the same, so the CTAs of one SM start together and their warps move
    // through the FP64 (MDS) and integer (S-box) stretches of the permutation in step.  Delay the first wave's CTA slots by
    // GLB_LEAF_STAGGER clocks each so that the four warps of a sub-partition sit in different stretches
    // (profiles/r1_27_ubench_fp64_phase.txt: -6 % on a synthetic two-phase loop); later CTAs inherit the offset of the slot they refill.
    {
        u32 nsm;
        asm("mov.u32 %0, %%nsmid;" : "=r"(nsm));
        if (blockIdx.x < GLB_LEAF_MINBLOCKS * nsm) {
            const long long wait = (long long)((blockIdx.x / nsm) % GLB_LEAF_MINBLOCKS) * (GLB_LEAF_STAGGER), t0 = clock64();
            while (clock64() - t0 < wait) { }
        }
    }
#endif
    linear_hash_row<PANELS>(digests + row * 4, rp, PANELS ? nullptr : rp.base[0] + row * ncols, row, ncols);
}

// The same sponge, one column block at a time: absorbs columns [c0, c0 + w) of every row (c0 a multiple of 8; w a
// multiple of 8 unless this is the last block) and keeps the running capacity words state[0..4) in the row's digest
// slot between blocks.  After the block with c0 + w == ncols_total the slot holds the leaf digest.  Lets the
// pipelined host-pointer commit hash block k while block k+1 is uploaded and transformed.  Needs ncols_total > 4.
__global__ void __launch_bounds__(kThreads, GLB_LEAF_MINBLOCKS) k_linear_hash_partial(u64 *__restrict__ digests, const u64 *__restrict__ src, u64 nrows, u64 pitch,
                                                                  u32 c0, u32 w)
{
    u64 row = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
    const u64 *p = src + row * pitch + c0;
    u64 *d = digests + row * 4;
    u64 s[12];
#pragma unroll
    for (int i = 0; i < 4; i++) s[8 + i] = c0 ? d[i] : 0;
    const u32 full = w >> 3;
    for (u32 c = 0; c < full; c++) {
#pragma unroll
        for (u32 i = 0; i < 8; i++) s[i] = p[c * 8 + i];
        permute<true>(s);
#pragma unroll
        for (int i = 0; i < 4; i++) s[8 + i] = s[i];
    }
    const u32 rem = w & 7;
    if (rem) {
#pragma unroll
        for (u32 i = 0; i < 8; i++) s[i] = i < rem ? p[full * 8 + i] : 0;
        permute<true>(s);
#pragma unroll
        for (int i = 0; i < 4; i++) s[8 + i] = s[i];
    }
#pragma unroll
    for (int i = 0; i < 4; i++) d[i] = gl::canon(s[8 + i]);
}

// merkletree_batch (poseidon_goldilocks.cpp:113-164, 302-352): digest of every `batch`-column slice of every
// row; thread v = (row, slice).  The per-row digest of these digests is a k_linear_hash_rows over the result.
__global__ void __launch_bounds__(kThreads) k_linear_hash_batches(u64 *__restrict__ out, const u64 *__restrict__ src, u64 nrows, u32 ncols,
                                                                  u32 batch, u32 nb)
{
    u64 v = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nrows * nb) return;
    const u64 row = v / nb;
    const u32 j = (u32)(v - row * nb);
    const u32 len = j == nb - 1 ? ncols - j * batch : batch;
    RowPanels rp = {};
    linear_hash_row<false>(out + v * 4, rp, src + row * ncols + (u64)j * batch, 0, len);
}

__global__ void __launch_bounds__(kThreads, GLB_LEAF_MINBLOCKS) k_tree_level(u64 *__restrict__ parents, const u64 *__restrict__ children, u64 nparents)
{
    u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= nparents) return;
    u64 s[12];
    const ulonglong2 *c = reinterpret_cast<const ulonglong2 *>(children + idx * 8);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        ulonglong2 v = c[i];
        s[2 * i] = v.x;
        s[2 * i + 1] = v.y;
    }
#pragma unroll
    for (int i = 8; i < 12; i++) s[i] = 0;
    permute<true>(s);
    ulonglong2 *o = reinterpret_cast<ulonglong2 *>(parents + idx * 4);
    o[0] = make_ulonglong2(gl::canon(s[0]), gl::canon(s[1]));
    o[1] = make_ulonglong2(gl::canon(s[2]), gl::canon(s[3]));
}

// The last levels of a tree in ONE launch: a single CTA walks the levels from `pending` <= 2 * kThreads nodes down to the root
// (one permutation per thread per level, a CTA barrier between levels).  A level this small is bound by the latency of one
// permutation per thread either way; fusing them removes log2(pending) - 1 launches from every tree, which is what the small FRI
// layer trees (friProve.cpp:120-123) and the gathered top of the multi-GPU commit consist of.
__global__ void __launch_bounds__(kThreads, GLB_LEAF_MINBLOCKS) k_tree_top(u64 *__restrict__ level, u64 pending)
{
    while (pending > 1) {
        const u64 nparents = (pending - 1) / 2 + 1;
        u64 *parents = level + pending * 4;
        for (u64 idx = threadIdx.x; idx < nparents; idx += blockDim.x) {
            u64 s[12];
            const ulonglong2 *c = reinterpret_cast<const ulonglong2 *>(level + idx * 8);
#pragma unroll
            for (int i = 0; i < 4; i++) {
                ulonglong2 v = c[i];
                s[2 * i] = v.x;
                s[2 * i + 1] = v.y;
            }
#pragma unroll
            for (int i = 8; i < 12; i++) s[i] = 0;
            permute<true>(s);
            ulonglong2 *o = reinterpret_cast<ulonglong2 *>(parents + idx * 4);
            o[0] = make_ulonglong2(gl::canon(s[0]), gl::canon(s[1]));
            o[1] = make_ulonglong2(gl::canon(s[2]), gl::canon(s[3]));
        }
        __syncthreads();
        level = parents;
        pending /= 2;
    }
}

// Derive the device tables from the generated parameter header (poseidon_constants.cuh).
inline cudaError_t upload_constants()
{
    static u64 rc0[12], rcf[96], rcp[22];
    for (int i = 0; i < 12; i++) rc0[i] = GLP_RC[i];
    for (int r = 0; r < 8; r++)
        for (int i = 0; i < 12; i++)
            rcf[r * 12 + i] = r < 4 ? GLP_RC[(r + 1) * 12 + i] : (r < 7 ? GLP_RC[82 + 12 * (r - 4) + i] : 0);
    for (int r = 0; r < 22; r++) rcp[r] = GLP_RC[60 + r];
    cudaError_t e;
    if ((e = cudaMemcpyToSymbol(c_RC0, rc0, sizeof rc0)) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(c_RCF, rcf, sizeof rcf)) != cudaSuccess) return e;
    if ((e = cudaMemcpyToSymbol(c_RCP, rcp, sizeof rcp)) != cudaSuccess) return e;
    return cudaSuccess;
}

inline unsigned grid_for(u64 n) { return (unsigned)((n + kThreads - 1) / kThreads); }

inline void launch_hash_full_result(u64 *out, const u64 *in, u64 n, cudaStream_t st)
{
    if (n) k_hash_full_result<<<grid_for(n), kThreads, 0, st>>>(out, in, n);
}

inline void launch_linear_hash_rows(u64 *digests, const u64 *src, u64 nrows, u64 ncols, cudaStream_t st)
{
    if (!nrows) return;
    RowPanels rp = {};
    rp.base[0] = src;
    rp.start[1] = (u32)ncols;
    rp.npanels = 1;
    k_linear_hash_rows<false><<<grid_for(nrows), kThreads, 0, st>>>(digests, rp, nrows, (u32)ncols);
}

inline void launch_linear_hash_panels(u64 *digests, const RowPanels &rp, u64 nrows, cudaStream_t st)
{
    if (nrows) k_linear_hash_rows<true><<<grid_for(nrows), kThreads, 0, st>>>(digests, rp, nrows, rp.start[rp.npanels]);
}

// Levels above the leaves.  tree holds (2*nrows-1)*4 words, leaves first (merklehash_goldilocks.hpp:25-28).
// Returns the number of kernels launched.
inline int launch_tree_levels(u64 *tree, u64 nrows, cudaStream_t st)
{
    int launches = 0;
    u64 pending = nrows, next_index = 0;
    while (pending > 1) {
        if (pending <= 2 * kThreads) { // the remaining levels in one CTA
            k_tree_top<<<1, kThreads, 0, st>>>(tree + next_index, pending);
            return launches + 1;
        }
        u64 next_n = (pending - 1) / 2 + 1; // reference :283-300 (power-of-two heights only are meaningful)
        k_tree_level<<<grid_for(next_n), kThreads, 0, st>>>(tree + next_index + pending * 4, tree + next_index, next_n);
        launches++;
        next_index += pending * 4;
        pending /= 2;
    }
    return launches;
}

} // namespace poseidon
