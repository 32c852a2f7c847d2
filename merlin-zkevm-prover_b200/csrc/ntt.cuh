// ntt.cuh -- multi-column Goldilocks NTT / INTT / LDE passes for sm_100a.
//
// Replaces NTT_Goldilocks::NTT_iters / reversePermutation / INTT / extendPol
// (reference src/goldilocks/src/ntt_goldilocks.cpp:12-164, 243-305, 307-352).
//
// Scheme (multi-pass "four-step", natural order in and out, no bit-reversal pass):
//   n = 2^k is split into m passes of b_p <= 8 bits, n_p = 2^b_p, rest_p = prod_{q>p} n_q,
//   prev_p = prod_{q<p} n_q.  A matrix row index is read as mixed-radix digits (d_1..d_m),
//   pos = sum d_q rest_q.  Pass p replaces digit j_p by k_p with an n_p-point DFT along that
//   digit (row stride rest_p) and multiplies by the inter-pass twiddle w_n^(prev_p*low*k_p),
//   low = pos mod rest_p.  Passes p < m work in place; the last pass writes row
//   sum k_q prev_q, i.e. natural order.
//
//   One CTA = one tile of R = 2^b rows x 32 adjacent columns (256-byte row segments, a warp is
//   always 32 columns of ONE row => coalesced global access, conflict-free shared memory,
//   warp-uniform twiddles).  A thread keeps 16 elements of one column in registers: a radix-16
//   DFT (twiddles are powers of w_16 = 2^12), the w_R^(t*k) twiddle, one exchange through
//   shared memory, a radix-R/16 DFT, the per-tile geometric twiddle/scale table, store.
//
// Roofline: HBM, 16 B per element per pass (one read + one write).
#pragma once
#include "gl64.cuh"
#include <cuda_runtime.h>

namespace ntt {

using gl::u32;
using gl::u64;

constexpr int kTileCols = 32;

// Power tables (uploaded by the host at init):
//   c_w32[j][i] = W32^(i * 2^(8j))   W32 = primitive 2^32-th root (reference Goldilocks::W[32])
//   c_p7 [j][i] = 7^(i * 2^(8j))     coset shift (Goldilocks::SHIFT, goldilocks_base_field.cpp:51)
//   c_inv2[k]   = 2^-k
//   c_w256[0][e] = w_256^e, c_w256[1][e] = w_256^-e
__constant__ u64 c_w32[4][256];
__constant__ u64 c_p7[4][256];
__constant__ u64 c_inv2[33];
__constant__ u64 c_w256[2][256];

__device__ __forceinline__ u64 pow_tab(const u64 (*tab)[256], u32 e)
{
    u64 r = tab[0][e & 255];
    r = gl::mul(r, tab[1][(e >> 8) & 255]);
    r = gl::mul(r, tab[2][(e >> 16) & 255]);
    r = gl::mul(r, tab[3][(e >> 24) & 255]);
    return r;
}

// 2^s mod p for 0 <= s < 192 (2^96 = -1 mod p), usable in constant folding.
__host__ __device__ constexpr u64 pow2mod(int s)
{
    s %= 192;
    if (s >= 96) return gl::P - pow2mod(s - 96);
    if (s < 64) return 1ULL << s;
    return (1ULL << (s - 32)) - (1ULL << (s - 64)); // 2^(s-64) * (2^32 - 1)
}

__host__ __device__ constexpr int bitrev(int x, int bits)
{
    int r = 0;
    for (int i = 0; i < bits; i++) r |= ((x >> i) & 1) << (bits - 1 - i);
    return r;
}

// In-register DFT of size 2^LOG on x[0 .. 2^LOG): decimation in frequency, so that on return
// register i holds output index bitrev(i).  Twiddles are powers of two: w_(2^LOG) = 2^(192/2^LOG).
template <int LOG, bool INV>
__device__ __forceinline__ void dft_regs(u64 *x)
{
    constexpr int N = 1 << LOG;
#pragma unroll
    for (int s = 0; s < LOG; s++) {
        const int half = N >> (s + 1);
#pragma unroll
        for (int blk = 0; blk < (1 << s); blk++) {
#pragma unroll
            for (int i = 0; i < half; i++) {
                const int ia = blk * 2 * half + i, ib = ia + half;
                u64 a = x[ia], b = gl::canon(x[ib]);
                x[ia] = gl::add_c(a, b);
                u64 d = gl::sub_c(a, b);
                const int e = i << s;               // exponent of w_N, 0 <= e < N/2
                if (e != 0) {
                    const int sh = (192 / N) * (INV ? (N - e) : e);
                    d = gl::mul(d, pow2mod(sh));
                }
                x[ib] = d;
            }
        }
    }
}

struct PassParams {
    const u64 *src;
    u64 *dst;
    u64 spitch, dpitch; // words per row
    u32 ncols;
    u32 logn;           // transform size
    u32 npass, pass;    // number of passes, this pass (0-based)
    u32 bits[4];        // bits per pass
    u32 log_rest;       // sum of bits of later passes
    u32 log_prev;       // sum of bits of earlier passes
    int inverse;        // use w^-1
    int scale_mode;     // last pass only: 0 none, 1 multiply by 1/n, 2 multiply row i by 7^i / n
    u64 nvalid;         // source rows >= nvalid are read as zeros (zero-padded LDE input)
};

template <int B2> struct Step2 {
    template <bool INV> static __device__ __forceinline__ void run(u64 *y)
    {
        constexpr int R2 = 1 << B2;
#pragma unroll
        for (int g = 0; g < 16 / R2; g++) dft_regs<B2, INV>(y + g * R2);
    }
};

// B2 = log2(R) - 4, blockDim = (32, 2^B2).
template <int B2, bool INV>
__global__ void __launch_bounds__(32 << B2) k_pass(PassParams p)
{
    constexpr int R2 = 1 << B2;
    constexpr int R = 16 * R2;
    extern __shared__ u64 smem[];
    u64 *tile = smem;                               // [R][32]   (only when B2 > 0)
    u64 *tw = smem + (B2 > 0 ? R * kTileCols : 0);  // [R] output multipliers
    u64 *pw = tw + R;                               // [32] two-level powers

    const u32 c = threadIdx.x, t = threadIdx.y;
    const u32 col = blockIdx.y * kTileCols + c;
    const bool colok = col < p.ncols;
    const u64 tau = blockIdx.x;
    const u64 hi = tau >> p.log_rest, low = tau & ((1ULL << p.log_rest) - 1);
    const u64 rest = 1ULL << p.log_rest;
    const u64 base_in = ((hi * R) << p.log_rest) + low;
    const bool last = (p.pass + 1 == p.npass);

    // ---- issue the global loads first; the twiddle table is built while they are in flight
    u64 x[16];
#pragma unroll
    for (int i1 = 0; i1 < 16; i1++) {
        u64 row = base_in + (((u64)i1 * R2 + t) << p.log_rest);
        x[i1] = (colok && row < p.nvalid) ? p.src[row * p.spitch + col] : 0;
    }

    // ---- per-tile output multipliers tw[k] = base * g^k
    u64 out_base = 0;
    bool use_tw;
    if (last) {
        // natural output row = digit-reversed(hi) + prev * k
        u64 rev = 0, h = hi;
        u32 shift = p.log_prev;
        for (int q = (int)p.npass - 2; q >= 0; q--) { // digits of hi, least significant = k_{m-1}
            shift -= p.bits[q];
            rev |= (h & ((1ULL << p.bits[q]) - 1)) << shift;
            h >>= p.bits[q];
        }
        out_base = rev;
        use_tw = (p.scale_mode == 2);
    } else {
        use_tw = true;
    }
    if (use_tw) {
        const u32 flat = t * 32 + c;
        if (flat < 32) {
            u64 g, b;
            if (last) { // 7^(rev + prev*k) / n
                g = pow_tab(c_p7, (u32)(1ULL << p.log_prev));
                b = gl::mul(pow_tab(c_p7, (u32)out_base), c_inv2[p.logn]);
            } else {    // w_n^(+-prev*low*k)
                u32 e = (u32)(((low << p.log_prev) & ((1ULL << p.logn) - 1)) << (32 - p.logn));
                if (p.inverse) e = 0u - e;
                g = pow_tab(c_w32, e);
                b = 1;
            }
            if (flat < 16) {
                pw[flat] = gl::pow(g, flat);
            } else {
                u64 g16 = g;
#pragma unroll
                for (int i = 0; i < 4; i++) g16 = gl::sqr(g16);
                pw[flat] = gl::mul(b, gl::pow(g16, flat - 16));
            }
        }
        __syncthreads();
        for (u32 k = t * 32 + c; k < (u32)R; k += 32 * R2) tw[k] = gl::mul(pw[16 + (k >> 4)], pw[k & 15]);
        // visibility of tw[] is guaranteed by the __syncthreads() below (B2 > 0) or the explicit one (B2 == 0)
    }

    // ---- step 1: radix-16 over i1
    dft_regs<4, INV>(x);

    u64 y[16];
    if (B2 > 0) {
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const int k1 = bitrev(i, 4);
            u64 v = x[i];
            if (k1 != 0) { // w_R^(t*k1) = w_256^(t*k1*256/R)
                u32 e = (t * k1) << (4 - B2);
                v = gl::mul(v, c_w256[INV ? 1 : 0][e]);
            }
            tile[(k1 * R2 + t) * kTileCols + c] = v;
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 16; i++) y[i] = tile[(16 * t + i) * kTileCols + c];
        Step2<B2>::template run<INV>(y);
    } else {
        if (use_tw) __syncthreads();
#pragma unroll
        for (int i = 0; i < 16; i++) y[i] = x[i];
    }

    // ---- store
    const u64 cscale = (last && p.scale_mode == 1) ? c_inv2[p.logn] : 0;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        u32 k;
        if (B2 > 0) {
            const int g = i / R2;
            const int k2 = bitrev(i % R2, B2);
            k = (16 * t) / R2 + g + 16 * k2;
        } else {
            k = bitrev(i, 4);
        }
        u64 v = y[i];
        if (use_tw) v = gl::mul(v, tw[k]);
        else if (cscale) v = gl::mul(v, cscale);
        u64 row = last ? out_base + ((u64)k << p.log_prev) : base_in + ((u64)k << p.log_rest);
        if (colok) p.dst[row * p.dpitch + col] = gl::canon(v);
    }
}

// n <= 8: one thread per column, direct O(n^2) evaluation (reference handles these sizes through
// the same radix-2 loop; ntt_goldilocks.cpp:28-36 clamps nphase).
__global__ void k_tiny(PassParams p)
{
    u32 col = blockIdx.x * blockDim.x + threadIdx.x;
    if (col >= p.ncols) return;
    const u32 n = 1u << p.logn;
    u64 x[8], w[8];
    u32 e = n > 1 ? (1u << (32 - p.logn)) : 0;
    if (p.inverse) e = 0u - e;
    u64 g = pow_tab(c_w32, e);
    w[0] = 1;
    for (u32 i = 1; i < n; i++) w[i] = gl::mul(w[i - 1], g);
    for (u32 i = 0; i < n; i++) x[i] = ((u64)i < p.nvalid) ? p.src[i * p.spitch + col] : 0;
    u64 sc = (p.scale_mode ? c_inv2[p.logn] : 1);
    for (u32 k = 0; k < n; k++) {
        u64 acc = 0;
        for (u32 j = 0; j < n; j++) acc = gl::mad(x[j], w[(j * k) & (n - 1)], acc);
        if (p.scale_mode) acc = gl::mul(acc, sc);
        p.dst[k * p.dpitch + col] = gl::canon(acc);
        if (p.scale_mode == 2) sc = gl::mul(sc, 7);
    }
}

} // namespace ntt
