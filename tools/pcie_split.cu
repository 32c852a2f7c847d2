// tools/pcie_split.cu -- does a column-stripe download go faster when its rows are split over several streams (copy engines)?
// D2H of a w-word stripe of a row-major matrix (pitch 600 words, 2^22 rows), as S row ranges on S streams, alone and with an
// upload running; and the same for the upload.  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/pcie_split tools/pcie_split.cu
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
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    printf("asyncEngineCount %d\n", prop.asyncEngineCount);
    cudaStream_t s[8], so;
    for (auto &x : s) CK(cudaStreamCreateWithFlags(&x, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&so, cudaStreamNonBlocking));
    cudaEvent_t e0, e1[8];
    CK(cudaEventCreate(&e0));
    for (auto &x : e1) CK(cudaEventCreate(&x));
    const int widths[] = {16, 32, 96, 128, 256};
    for (int dir = 0; dir < 2; dir++)
        for (int both = 0; both < 2; both++)
            for (int wi = 0; wi < 5; wi++)
                for (int S = 1; S <= 8; S *= 2) {
                    const size_t w = widths[wi] * 8, moved = rows * w;
                    float best = 1e30f;
                    for (int rep = 0; rep < 2; rep++) {
                        if (both) {
                            if (dir == 0) CK(cudaMemcpyAsync(d2, h2, bytes, cudaMemcpyHostToDevice, so));
                            else CK(cudaMemcpyAsync(h2, d2, bytes, cudaMemcpyDeviceToHost, so));
                        }
                        CK(cudaEventRecord(e0, s[0]));
                        for (int j = 1; j < S; j++) CK(cudaStreamWaitEvent(s[j], e0, 0));
                        const size_t rs = rows / S;
                        for (int j = 0; j < S; j++) {
                            if (dir == 0) CK(cudaMemcpy2DAsync(h + j * rs * pitch, pitch, d + j * rs * pitch, pitch, w, rs, cudaMemcpyDeviceToHost, s[j]));
                            else CK(cudaMemcpy2DAsync(d + j * rs * pitch, pitch, h + j * rs * pitch, pitch, w, rs, cudaMemcpyHostToDevice, s[j]));
                            CK(cudaEventRecord(e1[j], s[j]));
                        }
                        float ms = 0;
                        for (int j = 0; j < S; j++) {
                            CK(cudaEventSynchronize(e1[j]));
                            float t;
                            CK(cudaEventElapsedTime(&t, e0, e1[j]));
                            if (t > ms) ms = t;
                        }
                        CK(cudaDeviceSynchronize());
                        if (ms < best) best = ms;
                    }
                    printf("%s %-22s width %4d words, %d streams: %6.2f GB in %8.2f ms = %6.1f GB/s\n", dir ? "H2D" : "D2H", both ? "(other direction busy)" : "(alone)",
                           widths[wi], S, moved / 1e9, best, moved / best / 1e6);
                }
    return 0;
}
