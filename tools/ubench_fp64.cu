// tools/ubench_fp64.cu -- does the FP64 pipe of the B200 SM issue beside the integer pipe?
// DFMA alone, IMAD.WIDE alone, and 1:1 / 1:2 mixes (8 independent chains per thread).
// CAUTION (found at the end of round 1 with cuobjdump): in ops 1-4, 7-11 and 13 the multiplicands of `mad.wide.u32` are loop
// invariant and ptxas strength-reduces them to an IADD3 + IADD3.X pair, so the lines labelled "IMAD.WIDE" there measure two
// ALU-pipe additions, not a multiply.  Ops 14-16 and the `phase ... (real IMAD.WIDE)` lines feed the low word of the
// accumulator back as the multiplicand, which keeps a real IMAD.WIDE.U32 in the SASS.
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
                if (OP == 15 || OP == 16) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(m), "d"(c));
                if (OP >= 14 && OP <= 16) asm volatile("{ .reg .u32 lo, hi; mov.b64 {lo, hi}, %0; mad.wide.u32 %0, lo, %1, %0; }" : "+l"(acc[i]) : "r"(b));
                if (OP == 16) asm volatile("{ .reg .u32 lo, hi; mov.b64 {lo, hi}, %0; mad.wide.u32 %0, lo, %1, %0; }" : "+l"(acc2[i]) : "r"(a));
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

// Cross-warp mixes: does the FP64 pipe overlap the integer pipes when DIFFERENT warps feed them?
//   MODE 0: even warps run DFMA only, odd warps IMAD.WIDE + IADD3-pair only (half the warps each).
//   MODE 1: every warp alternates a run of 288 DFMAs (one MDS layer) and a run of ~770 integer instructions (12 S-boxes),
//           all warps in phase.   MODE 2: the same, odd warps start with the integer run (half a period out of phase).
// Reported as milliseconds; compare MODE 0 with the two halves alone (MODE 3: DFMA half only, MODE 4: integer half only)
// and MODE 1 with MODE 2.
template <int MODE, bool REAL>
__global__ void k_phase(u64 *out, int iters, u32 seed)
{
    u32 a = seed + threadIdx.x, b = seed * 3 + blockIdx.x;
    u64 acc[8];
    double d[8];
    u32 r[8];
    const double m = 1.0 + 1e-9 * seed, c = 1e-3 * threadIdx.x;
#pragma unroll
    for (int i = 0; i < 8; i++) { acc[i] = a + i; d[i] = 1.0 + i; r[i] = a * (i + 1); }
    const bool odd = (threadIdx.x >> 5) & 1;
    auto fp_run = [&](int n8) { // n8 * 8 DFMAs
#pragma unroll 1
        for (int k = 0; k < n8; k++)
#pragma unroll
            for (int i = 0; i < 8; i++) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(m), "d"(c));
    };
    auto int_run = [&](int n8) { // n8 * 8 * (1 IMAD.WIDE + 2 IADD3-class)
#pragma unroll 1
        for (int k = 0; k < n8; k++)
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (REAL) asm volatile("{ .reg .u32 lo, hi; mov.b64 {lo, hi}, %0; mad.wide.u32 %0, lo, %1, %0; }" : "+l"(acc[i]) : "r"(b));
                else asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[i]) : "r"(a), "r"(b));
                asm volatile("add.cc.u32 %0, %0, %1;\n\taddc.u32 %0, %0, %2;" : "+r"(r[i]) : "r"(a), "r"(b));
            }
    };
    for (int it = 0; it < iters; it++) {
        if (MODE == 0) { if (odd) int_run(32); else fp_run(36); }
        if (MODE == 3) { if (!odd) fp_run(36); }
        if (MODE == 4) { if (odd) int_run(32); }
        if (MODE == 1) { fp_run(36); int_run(32); }
        if (MODE == 2) { if (odd) { int_run(32); fp_run(36); } else { fp_run(36); int_run(32); } }
    }
    u64 t = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) t += acc[i] + (u64)__double_as_longlong(d[i]) + r[i];
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
                           "1 DFMA + 2 independent IMAD.WIDE", "1 DFMA + 1 IMAD.WIDE + 1 IADD3", "2 DFMA + 1 IMAD.WIDE", "1 DFMA + 1 IMAD.WIDE + 1 LOP3", "1 DFMA + 4 independent IMAD.WIDE", "DFMA with LDCU constant operand", "DFMA(LDCU const) + IMAD.WIDE",
                           "real IMAD.WIDE (acc.lo fed back)", "1 DFMA + 1 real IMAD.WIDE", "1 DFMA + 2 real independent IMAD.WIDE"};
    const int per[] = {1, 1, 2, 3, 4, 1, 2, 3, 3, 3, 3, 5, 1, 2, 1, 2, 3};
    for (int op = 0; op < 17; op++)
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
                case 14: k_issue<14><<<blocks, threads>>>(out, iters, 12345); break;
                case 15: k_issue<15><<<blocks, threads>>>(out, iters, 12345); break;
                case 16: k_issue<16><<<blocks, threads>>>(out, iters, 12345); break;
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
    {
        const char *pn[] = {"xwarp: even warps DFMA, odd warps IMAD.WIDE+IADD3", "phases in step (288 DFMA | 768 int per warp)", "phases staggered (odd warps half a period off)",
                            "xwarp: DFMA half alone", "xwarp: integer half alone"};
        for (int real = 0; real < 2; real++)
        for (int mode = 0; mode < 5; mode++) {
            float ms = 0;
            for (int rep = 0; rep < 2; rep++) {
                CK(cudaEventRecord(e0));
                switch (mode + 5 * real) { // 4 CTAs of 128 threads per SM = the leaf kernel's residency (4 warps per sub-partition)
                case 0: k_phase<0, false><<<sms * 4, 128>>>(out, 200, 12345); break;
                case 1: k_phase<1, false><<<sms * 4, 128>>>(out, 200, 12345); break;
                case 2: k_phase<2, false><<<sms * 4, 128>>>(out, 200, 12345); break;
                case 3: k_phase<3, false><<<sms * 4, 128>>>(out, 200, 12345); break;
                case 4: k_phase<4, false><<<sms * 4, 128>>>(out, 200, 12345); break;
                case 5: k_phase<0, true><<<sms * 4, 128>>>(out, 200, 12345); break;
                case 6: k_phase<1, true><<<sms * 4, 128>>>(out, 200, 12345); break;
                case 7: k_phase<2, true><<<sms * 4, 128>>>(out, 200, 12345); break;
                case 8: k_phase<3, true><<<sms * 4, 128>>>(out, 200, 12345); break;
                case 9: k_phase<4, true><<<sms * 4, 128>>>(out, 200, 12345); break;
                }
                CK(cudaEventRecord(e1));
                CK(cudaEventSynchronize(e1));
                CK(cudaEventElapsedTime(&ms, e0, e1));
            }
            printf("phase  %-52s %s %8.3f ms\n", pn[mode], real ? "(real IMAD.WIDE)" : "(IADD3 pairs)    ", ms);
        }
    }
    printf("ubench_fp64 done\n");
    return 0;
}
