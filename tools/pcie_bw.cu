// tools/pcie_bw.cu -- host<->device copy rates of the shapes the pipelined commit uses (pinned host memory):
// contiguous, and 2-D column-stripe copies (row pitch 600 words, stripe width w words) in both directions, alone
// and with the opposite direction running at the same time.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/pcie_bw tools/pcie_bw.cu
#include <cstdio>
#include <cstdlib>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)
int main()
{
    const size_t rows = 1 << 22, pitch = 600 * 8; // 20 GB matrix
    const size_t bytes = rows * pitch;
    char *h, *h2, *d, *d2;
    CK(cudaHostAlloc(&h, bytes, cudaHostAllocDefault));
    CK(cudaHostAlloc(&h2, bytes, cudaHostAllocDefault));
    CK(cudaMalloc(&d, bytes)); CK(cudaMalloc(&d2, bytes));
    CK(cudaMemset(d, 1, bytes)); CK(cudaMemset(d2, 1, bytes));
    for (size_t i = 0; i < bytes; i += 4096) { h[i] = 1; h2[i] = 2; }
    cudaStream_t s1, s2;
    CK(cudaStreamCreate(&s1)); CK(cudaStreamCreate(&s2));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float ms;
    const int widths[] = {0, 32, 128, 300};
    for (int dir = 0; dir < 2; dir++)          // 0 = D2H, 1 = H2D
        for (int both = 0; both < 2; both++)   // opposite direction running on another stream
            for (int wi = 0; wi < 4; wi++) {
                const size_t w = widths[wi] * 8;
                const size_t moved = w ? rows * w : bytes;
                for (int rep = 0; rep < 2; rep++) {
                    if (both) { // keep the other direction busy with a contiguous copy of the whole matrix
                        if (dir == 0) CK(cudaMemcpyAsync(d2, h2, bytes, cudaMemcpyHostToDevice, s2));
                        else CK(cudaMemcpyAsync(h2, d2, bytes, cudaMemcpyDeviceToHost, s2));
                    }
                    CK(cudaEventRecord(e0, s1));
                    if (w == 0) {
                        if (dir == 0) CK(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, s1));
                        else CK(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, s1));
                    } else {
                        if (dir == 0) CK(cudaMemcpy2DAsync(h, pitch, d, pitch, w, rows, cudaMemcpyDeviceToHost, s1));
                        else CK(cudaMemcpy2DAsync(d, pitch, h, pitch, w, rows, cudaMemcpyHostToDevice, s1));
                    }
                    CK(cudaEventRecord(e1, s1));
                    CK(cudaEventSynchronize(e1));
                    CK(cudaDeviceSynchronize());
                }
                CK(cudaEventElapsedTime(&ms, e0, e1));
                printf("%s %-22s width %4d words (pitch 600): %7.2f GB in %8.2f ms = %6.1f GB/s\n", dir ? "H2D" : "D2H",
                       both ? "(other direction busy)" : "(alone)", widths[wi], moved / 1e9, ms, moved / ms / 1e6);
            }
    // device-side repack cost: stripe -> contiguous (what a staged download would add)
    CK(cudaEventRecord(e0, s1));
    CK(cudaMemcpy2DAsync(d2, 128 * 8, d, pitch, 128 * 8, rows, cudaMemcpyDeviceToDevice, s1));
    CK(cudaEventRecord(e1, s1));
    CK(cudaEventSynchronize(e1));
    CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("D2D repack of a 128-word stripe to contiguous: %.2f GB in %.2f ms\n", rows * 1024 / 1e9, ms);
    return 0;
}
