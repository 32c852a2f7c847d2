// tools/pcie_all.cu -- aggregate host<->device bandwidth of G GPUs copying at the same time (pinned host memory, one stream per GPU
// and direction): where does the host side of gl_commit_mgpu_host saturate?  Build: nvcc -O3 -o tools/pcie_all tools/pcie_all.cu
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)
int main()
{
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    const size_t bytes = (size_t)4 << 30;
    std::vector<char *> h(ndev), h2(ndev), d(ndev), d2(ndev);
    std::vector<cudaStream_t> s(ndev), s2(ndev);
    for (int g = 0; g < ndev; g++) {
        CK(cudaSetDevice(g));
        CK(cudaHostAlloc(&h[g], bytes, cudaHostAllocPortable));
        CK(cudaHostAlloc(&h2[g], bytes, cudaHostAllocPortable));
        CK(cudaMalloc(&d[g], bytes)); CK(cudaMalloc(&d2[g], bytes));
        CK(cudaStreamCreate(&s[g])); CK(cudaStreamCreate(&s2[g]));
        for (size_t i = 0; i < bytes; i += 4096) { h[g][i] = 1; h2[g][i] = 2; }
    }
    for (int G = 1; G <= ndev; G *= 2)
        for (int mode = 0; mode < 3; mode++) { // 0 D2H, 1 H2D, 2 both directions at once
            double best = 1e30;
            for (int rep = 0; rep < 3; rep++) {
                for (int g = 0; g < G; g++) { CK(cudaSetDevice(g)); CK(cudaDeviceSynchronize()); }
                auto t0 = std::chrono::steady_clock::now();
                for (int g = 0; g < G; g++) {
                    CK(cudaSetDevice(g));
                    if (mode != 1) CK(cudaMemcpyAsync(h[g], d[g], bytes, cudaMemcpyDeviceToHost, s[g]));
                    if (mode != 0) CK(cudaMemcpyAsync(d2[g], h2[g], bytes, cudaMemcpyHostToDevice, s2[g]));
                }
                for (int g = 0; g < G; g++) { CK(cudaSetDevice(g)); CK(cudaDeviceSynchronize()); }
                double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
                if (sec < best) best = sec;
            }
            const double moved = (double)bytes * G * (mode == 2 ? 2 : 1);
            printf("pcie   %d GPU%s %-18s %6.1f GB in %7.1f ms = %6.1f GB/s aggregate (%5.1f per GPU%s)\n", G, G > 1 ? "s" : " ",
                   mode == 0 ? "D2H" : mode == 1 ? "H2D" : "D2H + H2D together", moved / 1e9, best * 1e3, moved / best / 1e9, moved / best / 1e9 / G,
                   mode == 2 ? ", both directions summed" : "");
        }
    return 0;
}
