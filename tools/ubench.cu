// tools/ubench.cu -- B200 micro-benchmarks that size the design (not part of the product library).
//   1. integer-pipe issue rates: IMAD.WIDE.U32, IMAD, IADD3 carry chains, LOP3, mixes
//   2. gl::mul / gl::mad / S-box / full Poseidon permutation throughput from registers
//   3. HBM bandwidth of row-segment access (the NTT tile pattern): rows of `pitch` bytes, each CTA
//      touching `seg`-byte segments of strided rows
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -I merlin-zkevm-prover_b200/csrc -o tools/ubench tools/ubench.cu
#include "poseidon_kernels.cuh"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

typedef unsigned long long u64;
typedef unsigned int u32;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

template <int OP>
__global__ void k_issue(u64 *out, int iters, u32 seed)
{
    u32 a = seed + threadIdx.x, b = seed * 3 + blockIdx.x;
    u64 acc[8];
    u32 r[8], s[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { acc[i] = a + i; r[i] = a * (i + 1); s[i] = b + i; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int rep = 0; rep < 8; rep++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (OP == 0) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[i]) : "r"(a), "r"(b));
                if (OP == 1) asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(r[i]) : "r"(a), "r"(b));
                if (OP == 2) asm volatile("add.cc.u32 %0, %0, %2;\n\taddc.u32 %1, %1, %3;" : "+r"(r[i]), "+r"(s[i]) : "r"(a), "r"(b));
                if (OP == 3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r[i]) : "r"(a), "r"(b));
                if (OP == 4) { // 1 wide mad + 1 carry-chain add pair
                    asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[i]) : "r"(a), "r"(b));
                    asm volatile("add.cc.u32 %0, %0, %2;\n\taddc.u32 %1, %1, %3;" : "+r"(r[i]), "+r"(s[i]) : "r"(a), "r"(b));
                }
                if (OP == 5) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(r[i]) : "r"(b));
                if (OP == 6) asm volatile("shf.l.wrap.b32 %0, %0, %1, 7;" : "+r"(r[i]) : "r"(s[i]));
                if (OP == 8) { // 1 wide mad + 2 independent LOP3
                    asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[i]) : "r"(a), "r"(b));
                    asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r[i]) : "r"(a), "r"(b));
                    asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(s[i]) : "r"(a), "r"(b));
                }
                if (OP == 9) { // 1 wide mad + 1 LOP3
                    asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[i]) : "r"(a), "r"(b));
                    asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r[i]) : "r"(a), "r"(b));
                }
                if (OP == 10) { // 1 wide mad + 1 plain IADD3 (no carry)
                    asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[i]) : "r"(a), "r"(b));
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(r[i]) : "r"(b));
                }
                if (OP == 11) { // 2 plain IADD3
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(r[i]) : "r"(b));
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(s[i]) : "r"(a));
                }
                if (OP == 7) { // mad.lo.cc + madc.hi.cc pair (expected: one IMAD.WIDE with carry) + addc
                    asm volatile("mad.lo.cc.u32 %0, %3, %4, %0;\n\tmadc.hi.cc.u32 %1, %3, %4, %1;\n\taddc.u32 %2, %2, 0;"
                                 : "+r"(r[i]), "+r"(s[i]), "+r"(a) : "r"(b), "r"(seed));
                }
            }
        }
    }
    u64 t = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) t += acc[i] + r[i] + s[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = t + a;
}

template <int OP>
__global__ void k_field(u64 *out, int iters, u64 seed)
{
    u64 x[8];
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B97F4A7C15ULL + blockIdx.x;
    u64 c = seed | 1;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (OP == 0) x[i] = gl::mul(x[i], c);
            if (OP == 1) x[i] = gl::mad(x[i], c, x[(i + 1) & 7]);
            if (OP == 2) x[i] = poseidon::sbox_add(x[i], c);
            if (OP == 3) x[i] = gl::add_c(x[i], gl::canon(x[(i + 3) & 7]));
            if (OP == 4) x[i] = gl::sub_c(x[i], gl::canon(x[(i + 3) & 7]));
            if (OP == 5) x[i] = gl::sqr(x[i]);
        }
    }
    u64 t = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) t ^= x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
}

__global__ void __launch_bounds__(128) k_perm(u64 *out, int iters, u64 seed)
{
    u64 s[12];
#pragma unroll
    for (int i = 0; i < 12; i++) s[i] = seed * (threadIdx.x + 1) + i + blockIdx.x;
    for (int it = 0; it < iters; it++) poseidon::permute(s);
    u64 t = 0;
#pragma unroll
    for (int i = 0; i < 12; i++) t ^= s[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
}

// Row-segment copy: matrix rows x pitch words; CTA (bx, by) handles rows {bx + i*rstride, i < R} and the
// column segment [by*seg, (by+1)*seg) words.  One warp-wide access = `seg` contiguous words of one row
// (32/seg rows per warp instruction when seg < 32).
__global__ void k_segcopy(u64 *dst, const u64 *src, u64 pitch, u64 rstride, int R, int seg, int do_write, int do_read)
{
    int lane = threadIdx.x;                 // blockDim.x = 256
    int rows_per_pass = 256 / seg;
    int c = lane % seg, rr = lane / seg;
    u64 col = (u64)blockIdx.y * seg + c;
    u64 acc = 0;
    for (int i = rr; i < R; i += rows_per_pass) {
        u64 row = blockIdx.x + (u64)i * rstride;
        u64 v = lane;
        if (do_read) v = src[row * pitch + col];
        if (do_write) dst[row * pitch + col] = v + 1;
        else acc += v;
    }
    if (!do_write && acc == 0x123456789ULL) dst[0] = acc;
}

static float time_ms(cudaEvent_t a, cudaEvent_t b) { float ms; cudaEventElapsedTime(&ms, a, b); return ms; }

int main(int argc, char **argv)
{
    CK(cudaSetDevice(0));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    printf("device %s, %d SMs, clock %d kHz\n", prop.name, prop.multiProcessorCount, prop.clockRate);
    CK(poseidon::upload_constants());
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const int sms = prop.multiProcessorCount;
    u64 *out;
    CK(cudaMalloc(&out, (size_t)sms * 16 * 1024 * 8));

    const bool only_perm = argc > 1 && !strcmp(argv[1], "perm");
    if (only_perm) {
        k_perm<<<sms * 8, 128>>>(out, 40, 0x1234567ULL);
        CK(cudaDeviceSynchronize());
        k_perm<<<sms * 8, 128>>>(out, 40, 0x1234567ULL);
        CK(cudaDeviceSynchronize());
        printf("perm only done\n");
        return 0;
    }
    if (argc > 1 && !strcmp(argv[1], "leaf")) { // the leaf kernel of the library on a 2^20 x 600 matrix (75 perms per row)
        const u64 rows = 1ULL << 20, cols = 600;
        u64 *a, *dg;
        CK(cudaMalloc(&a, rows * cols * 8)); CK(cudaMalloc(&dg, rows * 32));
        CK(cudaMemset(a, 0x5A, rows * cols * 8));
        for (int rep = 0; rep < 3; rep++) {
            CK(cudaEventRecord(e0));
            poseidon::launch_linear_hash_rows(dg, a, rows, cols, 0);
            CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
            printf("leaf   2^20 x 600  %8.3f ms  %8.2f M perm/s\n", time_ms(e0, e1), rows * 75.0 / time_ms(e0, e1) / 1e3);
        }
        return 0;
    }
    const bool skip_mem = argc > 1 && !strcmp(argv[1], "nomem");
    // ---- 1. issue rates
    const char *names[] = {"IMAD.WIDE.U32 (mad.wide.u32)", "IMAD (mad.lo.u32)", "IADD3+IADD3.X pair (add.cc/addc)", "LOP3", "mix: 1 mad.wide + 1 add.cc/addc pair",
                           "IMAD.HI (mul.hi.u32)", "SHF", "mad.lo.cc+madc.hi.cc+addc",
                           "mix: 1 mad.wide + 2 LOP3", "mix: 1 mad.wide + 1 LOP3", "mix: 1 mad.wide + 1 IADD", "2 IADD (no carry)"};
    const int ops_per[] = {1, 1, 2, 1, 3, 1, 1, 3, 3, 2, 2, 2};
    for (int op = 0; op < 12; op++) {
        for (int warps = 8; warps <= 32; warps *= 2) {
            int blocks = sms * 2, threads = warps * 32 / 2, iters = 2000;
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
                }
                CK(cudaEventRecord(e1));
                CK(cudaEventSynchronize(e1));
            }
            double ms = time_ms(e0, e1);
            double ops = (double)blocks * threads * iters * 64.0 * ops_per[op];
            printf("issue  %-42s warps/SM=%2d  %8.3f ms  %8.2f T lane-ops/s  (%.1f per SM per clk @1.9GHz)\n", names[op], warps, ms,
                   ops / ms / 1e9, ops / ms / 1e9 * 1e12 / sms / 1.9e9);
        }
    }
    // ---- 2. field ops
    const char *fnames[] = {"gl::mul", "gl::mad", "sbox_add (x^7+c)", "add_c(canon)", "sub_c(canon)", "gl::sqr"};
    for (int op = 0; op < 6; op++) {
        int blocks = sms * 4, threads = 256, iters = 2000;
        for (int rep = 0; rep < 2; rep++) {
            CK(cudaEventRecord(e0));
            switch (op) {
            case 0: k_field<0><<<blocks, threads>>>(out, iters, 0x1234567ULL); break;
            case 1: k_field<1><<<blocks, threads>>>(out, iters, 0x1234567ULL); break;
            case 2: k_field<2><<<blocks, threads>>>(out, iters / 4, 0x1234567ULL); break;
            case 3: k_field<3><<<blocks, threads>>>(out, iters, 0x1234567ULL); break;
            case 4: k_field<4><<<blocks, threads>>>(out, iters, 0x1234567ULL); break;
            case 5: k_field<5><<<blocks, threads>>>(out, iters, 0x1234567ULL); break;
            }
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
        }
        double ms = time_ms(e0, e1);
        double n = (double)blocks * threads * (op == 2 ? iters / 4 : iters) * 8.0;
        printf("field  %-20s %8.3f ms  %8.2f G ops/s\n", fnames[op], ms, n / ms / 1e6);
    }
    for (int bps = 1; bps <= 8; bps *= 2) {
        int blocks = sms * bps, threads = 128, iters = 40;
        for (int rep = 0; rep < 2; rep++) {
            CK(cudaEventRecord(e0));
            k_perm<<<blocks, threads>>>(out, iters, 0x1234567ULL);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
        }
        double ms = time_ms(e0, e1);
        printf("perm   warps/SM=%2d  %8.3f ms  %8.2f M perm/s\n", bps * 4, ms, (double)blocks * threads * iters / ms / 1e3);
    }
    // ---- 3. row-segment bandwidth
    if (!skip_mem) {
        const u64 pitch = 600, rows = 1ULL << 22; // 20 GB matrix
        u64 *a, *b;
        size_t bytes = rows * pitch * 8;
        CK(cudaMalloc(&a, bytes)); CK(cudaMalloc(&b, bytes));
        CK(cudaMemset(a, 1, bytes)); CK(cudaMemset(b, 2, bytes));
        // plain copy baseline
        for (int rep = 0; rep < 3; rep++) {
            CK(cudaEventRecord(e0));
            CK(cudaMemcpyAsync(b, a, bytes, cudaMemcpyDeviceToDevice));
            CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        }
        printf("mem    cudaMemcpy D2D %.1f GB  %8.3f ms  %8.1f GB/s (read+write)\n", bytes / 1e9, time_ms(e0, e1), 2.0 * bytes / time_ms(e0, e1) / 1e6);
        int segs[] = {4, 8, 16, 32};
        int Rs[] = {256, 1024, 2048};
        for (int mode = 0; mode < 3; mode++) {       // 0 read only, 1 write only, 2 read+write
            for (int ri = 0; ri < 3; ri++) {
                for (int si = 0; si < 4; si++) {
                    int seg = segs[si], R = Rs[ri];
                    u64 rstride = rows / R;
                    dim3 grid((unsigned)rstride, (unsigned)(pitch / seg));
                    if (pitch % seg) grid.y = pitch / seg; // ignore the ragged tail
                    for (int rep = 0; rep < 2; rep++) {
                        CK(cudaEventRecord(e0));
                        k_segcopy<<<grid, 256>>>(b, a, pitch, rstride, R, seg, mode != 0, mode != 1);
                        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
                    }
                    CK(cudaGetLastError());
                    double moved = (double)rows * grid.y * seg * 8 * (mode == 2 ? 2 : 1);
                    printf("mem    %s R=%4d seg=%3dB  %8.3f ms  %8.1f GB/s\n", mode == 0 ? "read " : mode == 1 ? "write" : "r+w  ", R, seg * 8,
                           time_ms(e0, e1), moved / time_ms(e0, e1) / 1e6);
                }
            }
        }
        CK(cudaFree(a)); CK(cudaFree(b));
    }
    printf("ubench done\n");
    return 0;
}
