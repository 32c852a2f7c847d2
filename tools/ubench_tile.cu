// tools/ubench_tile.cu -- how fast can one NTT-pass tile pattern be moved through shared memory by TMA?
// A pass tile is R rows at stride 2^log_rest x W adjacent columns of a row-major u64 matrix (csrc/ntt.cuh).  Persistent CTAs
// (one or more per SM) walk the tiles with an NBUF-deep ring: cp.async.bulk.tensor load (mbarrier completion) -> [threads
// optionally read + rewrite the tile through shared memory] -> cp.async.bulk.tensor store.  No field arithmetic: this is the
// memory-side ceiling of the pass for a given tile shape, to be compared with cudaMemcpy D2D.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I merlin-zkevm-prover_b200/csrc -o tools/ubench_tile tools/ubench_tile.cu
#include "tma.cuh"
#include <cstdio>
#include <vector>
#include <cstdlib>
typedef unsigned long long u64;
typedef unsigned int u32;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

struct TileParams {
    u64 ntiles;     // row tiles x column tiles
    u32 coltiles;
    u32 log_rest;
    u32 R, W;
    int touch;      // 1: every thread reads 16 words of the tile and writes them back (+1) before the store
};

template <int NBUF>
__global__ void __launch_bounds__(512) k_tile_mover(const __grid_constant__ CUtensorMap src, const __grid_constant__ CUtensorMap dst, const TileParams p)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t full[NBUF];
    const u32 tile_bytes = p.R * p.W * 8;
    u64 *bufs = (u64 *)smem_raw;
    const u32 tid = threadIdx.x;
    if (tid == 0) {
        for (int i = 0; i < NBUF; i++) tma::mbar_init(&full[i], 1);
        tma::fence_barrier_init();
    }
    __syncthreads();
    const u64 first = blockIdx.x, step = gridDim.x;
    const u64 mine = p.ntiles > first ? (p.ntiles - first + step - 1) / step : 0;
    auto coords = [&](u64 k, int &c0, int &c1, int &c2) {
        const u64 id = first + k * step;
        const u64 tau = id / p.coltiles;
        const u32 ct = (u32)(id - tau * p.coltiles);
        c0 = (int)(ct * p.W);
        c1 = (int)(tau & ((1ULL << p.log_rest) - 1));
        c2 = (int)((tau >> p.log_rest) * p.R);
    };
    auto issue_load = [&](u64 k) {
        int c0, c1, c2;
        coords(k, c0, c1, c2);
        const int b = (int)(k % NBUF);
        tma::mbar_expect_tx(&full[b], tile_bytes);
        tma::load_3d((unsigned char *)bufs + (size_t)b * tile_bytes, &src, &full[b], c0, c1, c2);
    };
    if (tid == 0)
        for (u64 k = 0; k < NBUF - 1 && k < mine; k++) issue_load(k);
    for (u64 j = 0; j < mine; j++) {
        const int b = (int)(j % NBUF);
        if (tid == 0 && j + NBUF - 1 < mine) {
            if (j >= 1) tma::store_wait_read<0>(); // the buffer of tile j-1 is about to be overwritten
            issue_load(j + NBUF - 1);
        }
        u64 *t = (u64 *)((unsigned char *)bufs + (size_t)b * tile_bytes);
        if (p.touch) {
            tma::mbar_wait(&full[b], (u32)((j / NBUF) & 1));
            const u32 per = p.R * p.W / blockDim.x;
            u64 x[16];
#pragma unroll
            for (u32 i = 0; i < 16; i++) if (i < per) x[i] = t[i * blockDim.x + tid];
            __syncthreads();
#pragma unroll
            for (u32 i = 0; i < 16; i++) if (i < per) t[((i * 5 + 3) % per) * blockDim.x + tid] = x[i] + 1;
            tma::fence_proxy_async();
            __syncthreads();
        } else if (tid == 0) {
            tma::mbar_wait(&full[b], (u32)((j / NBUF) & 1));
        }
        if (tid == 0) {
            int c0, c1, c2;
            coords(j, c0, c1, c2);
            tma::store_3d(&dst, t, c0, c1, c2);
            tma::store_commit();
        }
    }
    if (tid == 0) tma::store_wait_all<0>();
}

static float time_ms(cudaEvent_t a, cudaEvent_t b) { float ms; cudaEventElapsedTime(&ms, a, b); return ms; }

int main()
{
    CK(cudaSetDevice(0));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    printf("device %s, %d SMs\n", prop.name, sms);
    const u64 logn = 22, n = 1ULL << logn, pitch = 64; // 2 GiB per matrix
    u64 *a, *b;
    const size_t bytes = n * pitch * 8;
    CK(cudaMalloc(&a, bytes)); CK(cudaMalloc(&b, bytes));
    CK(cudaMemset(a, 1, bytes)); CK(cudaMemset(b, 2, bytes));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int rep = 0; rep < 3; rep++) {
        CK(cudaEventRecord(e0));
        CK(cudaMemcpyAsync(b, a, bytes, cudaMemcpyDeviceToDevice));
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    }
    printf("tile   cudaMemcpy D2D %.2f GB          %8.3f ms  %8.1f GB/s (read+write)\n", bytes / 1e9, time_ms(e0, e1), 2.0 * bytes / time_ms(e0, e1) / 1e6);
    CK(cudaFuncSetAttribute(k_tile_mover<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(k_tile_mover<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(k_tile_mover<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    const u32 R = 256;
    for (u32 W : {16u, 32u})
        for (u32 log_rest : {14u, 7u, 0u})
            for (int nbuf : {2, 3, 4})
                for (int cps : {1, 2, 3})       // CTAs per SM
                    for (int touch : {0, 1}) {
                        const size_t smem = (size_t)nbuf * R * W * 8;
                        if (smem * cps > 220 * 1024) continue;
                        CUtensorMap ms, md;
                        if (tma::make_map_3d(&ms, a, pitch, pitch, n, log_rest, W, R) || tma::make_map_3d(&md, b, pitch, pitch, n, log_rest, W, R)) {
                            printf("tensor map encode failed\n");
                            return 1;
                        }
                        TileParams p;
                        p.coltiles = (u32)(pitch / W);
                        p.ntiles = (n / R) * p.coltiles;
                        p.log_rest = log_rest; p.R = R; p.W = W; p.touch = touch;
                        const int threads = touch ? (W == 32 ? 512 : 256) : 32;
                        float best = 1e30f;
                        for (int rep = 0; rep < 3; rep++) {
                            CK(cudaEventRecord(e0));
                            if (nbuf == 2) k_tile_mover<2><<<sms * cps, threads, smem>>>(ms, md, p);
                            if (nbuf == 3) k_tile_mover<3><<<sms * cps, threads, smem>>>(ms, md, p);
                            if (nbuf == 4) k_tile_mover<4><<<sms * cps, threads, smem>>>(ms, md, p);
                            CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
                            CK(cudaGetLastError());
                            if (rep && time_ms(e0, e1) < best) best = time_ms(e0, e1);
                        }
                        printf("tile   R=%u W=%2u (%3u B rows) log_rest=%2u nbuf=%d ctas/SM=%d touch=%d  %8.3f ms  %8.1f GB/s\n", R, W, W * 8, log_rest, nbuf, cps, touch, best,
                               2.0 * bytes / best / 1e6);
                    }
    // spot check of the copy (touch = 1 adds 1 to every word and permutes rows inside a thread's 16 words -- only check touch = 0)
    {
        CUtensorMap ms, md;
        tma::make_map_3d(&ms, a, pitch, pitch, n, 14, 32, R);
        tma::make_map_3d(&md, b, pitch, pitch, n, 14, 32, R);
        CK(cudaMemset(b, 0, bytes));
        std::vector<u64> h(1 << 16);
        for (size_t i = 0; i < h.size(); i++) h[i] = i * 0x9E3779B97F4A7C15ULL;
        CK(cudaMemcpy(a + 12345 * pitch, h.data(), h.size() * 8, cudaMemcpyHostToDevice));
        TileParams p;
        p.coltiles = (u32)(pitch / 32); p.ntiles = (n / R) * p.coltiles; p.log_rest = 14; p.R = R; p.W = 32; p.touch = 0;
        k_tile_mover<3><<<sms, 32, 3 * R * 32 * 8>>>(ms, md, p);
        CK(cudaDeviceSynchronize());
        std::vector<u64> g(h.size());
        CK(cudaMemcpy(g.data(), b + 12345 * pitch, g.size() * 8, cudaMemcpyDeviceToHost));
        size_t bad = 0;
        for (size_t i = 0; i < h.size(); i++) bad += g[i] != h[i];
        printf("tile   copy check: %zu mismatches of %zu\n", bad, h.size());
    }
    printf("ubench_tile done\n");
    return 0;
}
