// tools/ubench_pipes.cu -- issue-cost model of the B200 SM sub-partition for the instructions the Goldilocks kernels use.
// Round 1's tools/ubench.cu labelled a line "IMAD.WIDE" whose multiply ptxas had strength-reduced to IADD3 + IADD3.X
// (loop-invariant multiplicands; cuobjdump).  Every kernel here keeps the intended opcode in the SASS -- the build step
// (tools/ubench_pipes.sh) prints the opcode histogram per kernel next to the expectation so that this can be checked
// BEFORE spending GPU time.  8 independent chains per thread; clocks are per warp instruction GROUP per SM sub-partition.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/ubench_pipes tools/ubench_pipes.cu
#include <cstdio>
#include <cstdlib>
typedef unsigned long long u64;
typedef unsigned int u32;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

// real 32x32+64 multiply-add with a 64-bit addend taken from the neighbouring chain: stays ONE IMAD.WIDE.U32 (with `mad.wide.u32 acc, lo, hi, acc`
// ptxas splits the accumulate into IMAD.WIDE + IADD3 + IMAD.X; with loop-invariant multiplicands it strength-reduces the multiply away)
#define WIDE(i)  asm volatile("{ .reg .u32 lo, hi, x, y; mov.b64 {lo, hi}, %0; mov.b64 {x, y}, %1; mad.lo.cc.u32 x, lo, hi, x; madc.hi.u32 y, lo, hi, y; mov.b64 %0, {x, y}; }" : "+l"(acc[i]) : "l"(acc[(i + 1) & 7]))
#define WIDE2(i) asm volatile("{ .reg .u32 lo, hi, x, y; mov.b64 {lo, hi}, %0; mov.b64 {x, y}, %1; mad.lo.cc.u32 x, lo, hi, x; madc.hi.u32 y, lo, hi, y; mov.b64 %0, {x, y}; }" : "+l"(acc2[i]) : "l"(acc2[(i + 1) & 7]))
#define WIDEP(i) asm volatile("{ .reg .u32 lo, hi; mov.b64 {lo, hi}, %0; mul.wide.u32 %0, lo, hi; }" : "+l"(acc[i]))
// additions whose other operands are neighbouring chains' values (ptxas folds r += b; r += b into r += 2b otherwise).  IADD: 2-input add
// (ptxas is free to emit IADD3 or IMAD.IADD); IADDT: 3-input add, which only IADD3 can do
#define IADD(i)  asm volatile("add.u32 %0, %0, %1;" : "+r"(r[i]) : "r"(r[(i + 1) & 7]))
#define IADD2(i) asm volatile("add.u32 %0, %0, %1;" : "+r"(s[i]) : "r"(s[(i + 1) & 7]))
#define IADD3_(i) asm volatile("add.u32 %0, %0, %1;" : "+r"(t[i]) : "r"(t[(i + 1) & 7]))
#define IADD4_(i) asm volatile("add.u32 %0, %0, %1;" : "+r"(u[i]) : "r"(u[(i + 1) & 7]))
#define IADDT(i) asm volatile("{ .reg .u32 z; add.u32 z, %0, %1; add.u32 %0, z, %2; }" : "+r"(r[i]) : "r"(r[(i + 1) & 7]), "r"(r[(i + 2) & 7]))
#define IADDT2(i) asm volatile("{ .reg .u32 z; add.u32 z, %0, %1; add.u32 %0, z, %2; }" : "+r"(s[i]) : "r"(s[(i + 1) & 7]), "r"(s[(i + 2) & 7]))
#define IADDT3(i) asm volatile("{ .reg .u32 z; add.u32 z, %0, %1; add.u32 %0, z, %2; }" : "+r"(t[i]) : "r"(t[(i + 1) & 7]), "r"(t[(i + 2) & 7]))
#define IADDT4(i) asm volatile("{ .reg .u32 z; add.u32 z, %0, %1; add.u32 %0, z, %2; }" : "+r"(u[i]) : "r"(u[(i + 1) & 7]), "r"(u[(i + 2) & 7]))
#define CARRY(i) asm volatile("add.cc.u32 %0, %0, %2;\n\taddc.u32 %1, %1, %3;" : "+r"(r[i]), "+r"(s[i]) : "r"(a), "r"(b))
#define LOP(i)   asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(t[i]) : "r"(a), "r"(b))
#define LOP2(i)  asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u[i]) : "r"(b), "r"(a))
#define SHF(i)   asm volatile("shf.l.wrap.b32 %0, %0, %1, 7;" : "+r"(t[i]) : "r"(u[i]))
#define SEL(i)   asm volatile("{ .reg .pred p; setp.lt.u32 p, %0, %1; selp.u32 %0, %2, %0, p; }" : "+r"(t[i]) : "r"(a), "r"(b))
#define IMAD(i)  asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(t[i]) : "r"(b), "r"(a))
#define IMADHI(i) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(t[i]) : "r"(b))
#define DFMA(i)  asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(m), "d"(c))
#define DFMA2(i) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(e[i]) : "d"(c), "d"(m))
#define DADD(i)  asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(d[i]) : "d"(c))
#define FFMA(i)  asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(fm), "f"(fc))
#define I2F(i)   asm volatile("cvt.rn.f64.u32 %0, %1;\n\tmov.b64 {%1, %2}, %0;" : "=d"(d[i]), "+r"(t[i]), "=r"(u[i]))
#define PRMT(i)  asm volatile("prmt.b32 %0, %0, %1, 0x2103;" : "+r"(t[i]) : "r"(a))

template <int OP>
__global__ void __launch_bounds__(512) k_issue(u64 *out, int iters, u32 seed)
{
    u32 a = seed + threadIdx.x, b = seed * 3 + blockIdx.x + threadIdx.x * 5; // both per-thread: a warp-uniform chain would move to the uniform datapath
    u64 acc[8], acc2[8];
    u32 r[8], s[8], t[8], u[8];
    double d[8], e[8];
    float f[8];
    const double m = 1.0 + 1e-9 * seed, c = 1e-3 * threadIdx.x;
    const float fm = 1.0f + 1e-6f * seed, fc = 1e-3f * threadIdx.x;
#pragma unroll
    for (int i = 0; i < 8; i++) { acc[i] = a + i; acc2[i] = b + i; r[i] = a * (i + 1); s[i] = b + i; t[i] = a ^ i; u[i] = b ^ (i * 7); d[i] = 1.0 + i; e[i] = 2.0 + i; f[i] = 1.0f + i; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int rep = 0; rep < 8; rep++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (OP == 0) { WIDE(i); }
                if (OP == 1) { IADD(i); }
                if (OP == 2) { CARRY(i); }
                if (OP == 3) { LOP(i); }
                if (OP == 4) { SHF(i); }
                if (OP == 5) { SEL(i); }
                if (OP == 6) { IMAD(i); }
                if (OP == 7) { IMADHI(i); }
                if (OP == 8) { WIDE(i); IADD(i); }
                if (OP == 9) { WIDE(i); IADD(i); IADD2(i); }
                if (OP == 10) { WIDE(i); IADD(i); IADD2(i); IADD3_(i); IADD4_(i); }
                if (OP == 11) { WIDE(i); LOP(i); }
                if (OP == 12) { WIDE(i); LOP(i); LOP2(i); }
                if (OP == 13) { LOP(i); IADD(i); }
                if (OP == 14) { LOP(i); IADD(i); IADD2(i); }
                if (OP == 15) { DFMA(i); }
                if (OP == 16) { DFMA(i); WIDE(i); }
                if (OP == 17) { DFMA(i); DFMA2(i); WIDE(i); }
                if (OP == 18) { DFMA(i); WIDE(i); WIDE2(i); }
                if (OP == 19) { DFMA(i); IADD(i); IADD2(i); }
                if (OP == 20) { DFMA(i); LOP(i); }
                if (OP == 21) { DFMA(i); WIDE(i); IADD(i); IADD2(i); }
                if (OP == 22) { DFMA(i); IMAD(i); }
                if (OP == 23) { FFMA(i); }
                if (OP == 24) { FFMA(i); WIDE(i); }
                if (OP == 25) { FFMA(i); LOP(i); }
                if (OP == 26) { FFMA(i); IADD(i); }
                if (OP == 27) { I2F(i); }
                if (OP == 28) { I2F(i); DFMA2(i); }
                if (OP == 29) { PRMT(i); }
                if (OP == 30) { IMAD(i); IADD(i); }
                if (OP == 31) { WIDE(i); WIDE2(i); CARRY(i); }
                if (OP == 32) { DADD(i); }
                if (OP == 33) { WIDE(i); SHF(i); }
                if (OP == 34) { IMAD(i); LOP2(i); }
                if (OP == 35) { WIDEP(i); }
                if (OP == 36) { IADDT(i); }
                if (OP == 37) { WIDE(i); IADDT(i); }
                if (OP == 38) { WIDE(i); IADDT(i); IADDT2(i); }
                if (OP == 39) { LOP(i); IADDT(i); }
                if (OP == 40) { DFMA(i); IADDT(i); IADDT2(i); }
                if (OP == 41) { DFMA(i); WIDE(i); IADDT(i); IADDT2(i); }
                if (OP == 42) { DFMA(i); WIDE(i); IADDT(i); }
                if (OP == 43) { WIDE(i); IADDT(i); IADDT2(i); IADDT3(i); IADDT4(i); }
                if (OP == 44) { DFMA(i); WIDE(i); IADDT(i); IADDT2(i); IADDT3(i); IADDT4(i); }
                if (OP == 45) { DFMA(i); DFMA2(i); WIDE(i); IADDT(i); IADDT2(i); IADDT3(i); IADDT4(i); }
                if (OP == 46) { IMAD(i); IADDT(i); IADDT2(i); }
                if (OP == 47) { WIDE(i); WIDE2(i); }
            }
        }
    }
    u64 x = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) x += acc[i] + acc2[i] + r[i] + s[i] + t[i] + u[i] + (u64)__double_as_longlong(d[i]) + (u64)__double_as_longlong(e[i]) + (u64)__float_as_uint(f[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = x + a;
}

struct Op { const char *name; int per; };
static const Op kOps[] = {
    {"IMAD.WIDE.U32 (real)", 1}, {"IADD3", 1}, {"IADD3 + IADD3.X carry pair", 2}, {"LOP3", 1}, {"SHF", 1}, {"ISETP + SEL", 2}, {"IMAD (32-bit, real)", 1},
    {"IMAD.HI.U32", 1}, {"1 WIDE + 1 IADD3", 2}, {"1 WIDE + 2 IADD3", 3}, {"1 WIDE + 4 IADD3", 5}, {"1 WIDE + 1 LOP3", 2}, {"1 WIDE + 2 LOP3", 3},
    {"1 LOP3 + 1 IADD3", 2}, {"1 LOP3 + 2 IADD3", 3}, {"DFMA", 1}, {"1 DFMA + 1 WIDE", 2}, {"2 DFMA + 1 WIDE", 3}, {"1 DFMA + 2 WIDE", 3},
    {"1 DFMA + 2 IADD3", 3}, {"1 DFMA + 1 LOP3", 2}, {"1 DFMA + 1 WIDE + 2 IADD3", 4}, {"1 DFMA + 1 IMAD", 2}, {"FFMA", 1}, {"1 FFMA + 1 WIDE", 2},
    {"1 FFMA + 1 LOP3", 2}, {"1 FFMA + 1 IADD3", 2}, {"I2F.F64.U32 (+mov)", 1}, {"1 I2F.F64.U32 + 1 DFMA", 2}, {"PRMT", 1}, {"1 IMAD + 1 IADD3", 2},
    {"2 WIDE + carry pair", 4}, {"DADD", 1}, {"1 WIDE + 1 SHF", 2}, {"1 IMAD + 1 LOP3", 2},
    {"IMAD.WIDE.U32 product only", 1}, {"IADD3 (3-input)", 1}, {"1 WIDE + 1 IADD3(3in)", 2}, {"1 WIDE + 2 IADD3(3in)", 3}, {"1 LOP3 + 1 IADD3(3in)", 2},
    {"1 DFMA + 2 IADD3(3in)", 3}, {"1 DFMA + 1 WIDE + 2 IADD3(3in)", 4}, {"1 DFMA + 1 WIDE + 1 IADD3(3in)", 3},
    {"1 WIDE + 4 IADD3(3in)", 5}, {"1 DFMA + 1 WIDE + 4 IADD3(3in)", 6}, {"2 DFMA + 1 WIDE + 4 IADD3(3in)", 7}, {"1 IMAD + 2 IADD3(3in)", 3}, {"2 WIDE (two chains)", 2},
};
constexpr int kNumOps = sizeof(kOps) / sizeof(kOps[0]);

template <int OP> void launch_one(u64 *out, int blocks, int threads, int iters) { k_issue<OP><<<blocks, threads>>>(out, iters, 12345); }
template <int... I> void launch(int op, u64 *out, int blocks, int threads, int iters, std::integer_sequence<int, I...>)
{
    ((op == I ? launch_one<I>(out, blocks, threads, iters) : (void)0), ...);
}

int main(int argc, char **argv)
{
    CK(cudaSetDevice(0));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    int khz = 0;
    CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0));
    printf("device %s, %d SMs, max clock %d kHz (clk figures assume the max clock; check clocks.csv)\n", prop.name, sms, khz);
    u64 *out;
    CK(cudaMalloc(&out, (size_t)sms * 4 * 1024 * 8));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int op = 0; op < kNumOps; op++)
        for (int warps = 8; warps <= 32; warps *= 2) {
            const int blocks = sms * 2, threads = warps * 32 / 2, iters = 1000;
            float ms = 0, best = 1e30f;
            for (int rep = 0; rep < 3; rep++) {
                CK(cudaEventRecord(e0));
                launch(op, out, blocks, threads, iters, std::make_integer_sequence<int, kNumOps>());
                CK(cudaEventRecord(e1));
                CK(cudaEventSynchronize(e1));
                CK(cudaEventElapsedTime(&ms, e0, e1));
                if (rep && ms < best) best = ms;
            }
            CK(cudaGetLastError());
            const double groups = (double)iters * 64.0, warps_per_smsp = warps / 4.0;
            const double clk = best * 1e-3 * (khz * 1e3) / (groups * warps_per_smsp);
            printf("pipe   %-30s warps/SM=%2d  %8.3f ms  %6.2f clk per group of %d per SMSP  (%.2f clk/instr, %.1f T lane-instr/s)\n", kOps[op].name, warps, best,
                   clk, kOps[op].per, clk / kOps[op].per, (double)kOps[op].per * groups * warps * 32 * sms / (best * 1e-3) / 1e12);
        }
    printf("ubench_pipes done\n");
    return 0;
}
