// tools/ubench_fp64.cu -- does the FP64 pipe of the B200 SM issue beside the integer pipe?
// DFMA alone, IMAD.WIDE alone, and 1:1 / 1:2 mixes (8 independent chains per thread).
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/ubench_fp64 tools/ubench_fp64.cu
#include <cstdio>
#include <cstdlib>
typedef unsigned long long u64;
typedef unsigned int u32;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__constant__ double c_k[64];

template <int OP>
__global__ void k_issue(u64 *out, int iters, u32 seed)
{
    u32 a = seed + threadIdx.x, b = seed * 3 + blockIdx.x;
    u64 acc[8];
    double d[8];
    u32 r[8];
    u64 acc2[8];
    const double m = 1.0 + 1e-9 * seed, c = 1e-3 * threadIdx.x;
#pragma unroll
    for (int i = 0; i < 8; i++) { acc[i] = a + i; acc2[i] = b + i; d[i] = 1.0 + i; r[i] = a * (i + 1); }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int rep = 0; rep < 8; rep++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (OP == 0 || OP == 2 || OP == 3 || OP == 4) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(m), "d"(c));
                if (OP == 1 || OP == 2 || OP == 3 || OP == 4) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[i]) : "r"(a), "r"(b));
                if (OP == 3) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[i]) : "r"(b), "r"(a));
                if (OP == 4) asm volatile("add.cc.u32 %0, %0, %1;\n\taddc.u32 %0, %0, %2;" : "+r"(r[i]) : "r"(a), "r"(b));
                if (OP == 5) asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(d[i]) : "d"(c));
                if (OP >= 7 && OP <= 11) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(m), "d"(c));
                if (OP == 9) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[(i + 4) & 7]) : "d"(c), "d"(m));
                if (OP >= 7 && OP <= 11) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[i]) : "r"(a), "r"(b));
                if (OP == 7 || OP == 11) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc2[i]) : "r"(b), "r"(a));
                if (OP == 11) {
                    asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[(i + 4) & 7]) : "r"(a), "r"(a));
                    asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc2[(i + 4) & 7]) : "r"(b), "r"(b));
                }
                if (OP == 8) asm volatile("add.u32 %0, %0, %1;" : "+r"(r[i]) : "r"(b));
                if (OP == 10) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r[i]) : "r"(a), "r"(b));
                if (OP == 12 || OP == 13) d[i] = fma(d[i], c_k[(it * 8 + i) & 63], c); // constant through a uniform register (LDCU)
                if (OP == 13) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[i]) : "r"(a), "r"(b));
                if (OP == 6) { // DADD + LOP3
                    asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(d[i]) : "d"(c));
                    asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r[i]) : "r"(a), "r"(b));
                }
            }
        }
    }
    u64 t = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) t += acc[i] + acc2[i] + (u64)__double_as_longlong(d[i]) + r[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = t + a;
}

int main()
{
    CK(cudaSetDevice(0));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    { double hk[64]; for (int i = 0; i < 64; i++) hk[i] = 1.0 + 1e-12 * i; CK(cudaMemcpyToSymbol(c_k, hk, sizeof hk)); }
    u64 *out;
    CK(cudaMalloc(&out, (size_t)sms * 16 * 1024 * 8));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const char *names[] = {"DFMA", "IMAD.WIDE", "1 DFMA + 1 IMAD.WIDE", "1 DFMA + 2 IMAD.WIDE", "1 DFMA + 1 IMAD.WIDE + IADD3 pair", "DADD", "1 DADD + 1 LOP3",
                           "1 DFMA + 2 independent IMAD.WIDE", "1 DFMA + 1 IMAD.WIDE + 1 IADD3", "2 DFMA + 1 IMAD.WIDE", "1 DFMA + 1 IMAD.WIDE + 1 LOP3", "1 DFMA + 4 independent IMAD.WIDE", "DFMA with LDCU constant operand", "DFMA(LDCU const) + IMAD.WIDE"};
    const int per[] = {1, 1, 2, 3, 4, 1, 2, 3, 3, 3, 3, 5, 1, 2};
    for (int op = 0; op < 14; op++)
        for (int warps = 8; warps <= 32; warps *= 2) {
            int blocks = sms * 2, threads = warps * 32 / 2, iters = 2000;
            float ms = 0;
            for (int rep = 0; rep < 2; rep++) {
                CK(cudaEventRecord(e0));
                switch (op) {
                case 0: k_issue<0><<<blocks, threads>>>(out, iters, 12345); break;
                case 1: k_issue<1><<<blocks, threads>>>(out, iters, 12345); break;
                case 2: k_issue<2><<<blocks, threads>>>(out, iters, 12345); break;
                case 3: k_issue<3><<<blocks, threads>>>(out, iters, 12345); break;
                case 4: k_issue<4><<<blocks, threads>>>(out, iters, 12345); break;
                case 5: k_issue<5><<<blocks, threads>>>(out, iters, 12345); break;
                case 6: k_issue<6><<<blocks, threads>>>(out, iters, 12345); break;
                case 7: k_issue<7><<<blocks, threads>>>(out, iters, 12345); break;
                case 8: k_issue<8><<<blocks, threads>>>(out, iters, 12345); break;
                case 9: k_issue<9><<<blocks, threads>>>(out, iters, 12345); break;
                case 10: k_issue<10><<<blocks, threads>>>(out, iters, 12345); break;
                case 11: k_issue<11><<<blocks, threads>>>(out, iters, 12345); break;
                case 12: k_issue<12><<<blocks, threads>>>(out, iters, 12345); break;
                case 13: k_issue<13><<<blocks, threads>>>(out, iters, 12345); break;
                }
                CK(cudaEventRecord(e1));
                CK(cudaEventSynchronize(e1));
                CK(cudaEventElapsedTime(&ms, e0, e1));
            }
            // clocks per warp per group of `per` instructions, per SM sub-partition (4 per SM), at 1.9 GHz nominal
            double groups = (double)iters * 64.0;              // per thread
            double warps_per_smsp = warps / 4.0;
            double clk = ms * 1e-3 * 1.9e9 / (groups * warps_per_smsp);
            printf("fp64   %-36s warps/SM=%2d  %8.3f ms  %.2f clk per group of %d per SMSP (@1.9 GHz)\n", names[op], warps, ms, clk, per[op]);
        }
    printf("ubench_fp64 done\n");
    return 0;
}
