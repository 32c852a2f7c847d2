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
// Roofline: 16 B of HBM traffic per element per pass (one read + one write), but the pass is bound by integer
// instruction issue first: 129 instructions per element per 8-bit pass (profiles/r1_17_ntt_ncu.txt).
#pragma once
#include "gl64.cuh"
#include <cuda_runtime.h>

namespace ntt {

using gl::u32;
using gl::u64;

constexpr int kTileCols = 32;
#ifndef GLB_NTT_CSEL
#define GLB_NTT_CSEL 1 // butterflies in the IMAD.WIDE-free form (gl::canon_sel / add_cx / sub_cx); 0: the round-1 form, kept for A/B builds
#endif
constexpr bool kCsel = GLB_NTT_CSEL != 0;

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

// d * 2^sh for the shifts that occur inside radix-16/8/4 butterflies (multiples of 12 below 96).  `sh` is a
// compile-time constant after unrolling; the switch folds away.
__device__ __forceinline__ u64 mul_pow2_sel(u64 d, int sh)
{
    switch (sh) {
    case 12: return gl::mul_pow2<12>(d);
    case 24: return gl::mul_pow2<24>(d);
    case 36: return gl::mul_pow2<36>(d);
    case 48: return gl::mul_pow2<48>(d);
    case 60: return gl::mul_pow2<60>(d);
    case 72: return gl::mul_pow2<72>(d);
    case 84: return gl::mul_pow2<84>(d);
    default: return d;
    }
}

// In-register forward DFT of size 2^LOG on x[0 .. 2^LOG): decimation in frequency, so that on return
// register i holds output index bitrev(i).  Twiddles are powers of two: w_(2^LOG) = 2^(192/2^LOG)
// (Goldilocks::W[1..6] = 2^96, 2^48, 2^24, 2^12, 2^6, 2^3; goldilocks_base_field.cpp:11-18), so a butterfly
// is canon (4) + add (4) + sub (5) instructions and, for 17 of the 32 butterflies of a radix-16, a
// 6-8 instruction shift-multiply.  Values stay lazily reduced; only the subtrahend is canonicalised.
// SBEG > 0: the caller has already done stages 0 .. SBEG-1 (dft16_ztop: trivial stages of a zero-padded input).
template <int LOG, bool CSEL = false, int SBEG = 0>
__device__ __forceinline__ void dft_regs(u64 *x)
{
    constexpr int N = 1 << LOG;
#pragma unroll
    for (int s = SBEG; s < LOG; s++) {
        const int half = N >> (s + 1);
#pragma unroll
        for (int blk = 0; blk < (1 << s); blk++) {
#pragma unroll
            for (int i = 0; i < half; i++) {
                const int ia = blk * 2 * half + i, ib = ia + half;
                const u64 a = x[ia], b = CSEL ? gl::canon_sel(x[ib]) : gl::canon(x[ib]);
                x[ia] = CSEL ? gl::add_cx(a, b) : gl::add_c(a, b);
                const u64 d = CSEL ? gl::sub_cx(a, b) : gl::sub_c(a, b);
                x[ib] = mul_pow2_sel(d, (192 / N) * (i << s)); // exponent of w_N: i << s < N/2
            }
        }
    }
}

// Radix-16 of a column whose registers 8..15 are zero (the first pass of an extended forward transform: rows >= nvalid of the source
// are padding, and with blowup >= 2 that is the upper half of every thread's 16 rows).  Stage 0 has nothing to add: x[i] stays and
// x[i + 8] = x[i] * w_16^i -- exactly the words the general butterflies produce from a zero operand (a + 0 and a - 0 return a's bits
// unchanged), so the results stay bit-identical.  ALL: only register 0 is non-zero (blowup >= 16): the transform of a delta is constant.
template <bool CSEL, bool ALL>
__device__ __forceinline__ void dft16_ztop(u64 *x)
{
    if (ALL) {
#pragma unroll
        for (int i = 1; i < 16; i++) x[i] = x[0];
        return;
    }
#pragma unroll
    for (int i = 0; i < 8; i++) x[i + 8] = mul_pow2_sel(x[i], 12 * i);
    dft_regs<4, CSEL, 1>(x);
}

template <int V> struct IntTag { static constexpr int value = V; };

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
    int inverse;        // last pass: output k goes to row (n-k) mod n (INTT = forward NTT + index reversal,
                        // exactly as the reference does it, ntt_goldilocks.cpp:139-148); k_tiny: use w^-1
    int scale_mode;     // last pass only: 0 none, 1 multiply by 1/n, 2 multiply row i by 7^i / n
    u64 nvalid;         // source rows >= nvalid are read as zeros (zero-padded LDE input)
    const u64 *roots;   // roots[i] = w_(2^log_roots)^i, i < 2^log_roots  (inter-pass twiddles)
    u32 log_roots;
    const u64 *rscale;  // scale_mode 2: rscale[i] = 7^i / n, i < n (nullptr: computed on the fly -- domains above 2^27)
    u64 cscale;         // scale_mode 1: 1/n
    // domains above 2^27 (the reference accepts up to 2^32, ntt_goldilocks.hpp:94-97): `roots` then holds only the first 2^log_lo powers
    // of w_(2^log_roots) and roots_hi[j] = w^(j * 2^log_lo); a twiddle is the product of one entry of each.  nullptr otherwise.
    const u64 *roots_hi;
    u32 log_lo;
    u32 pf_dist;        // k_pass: prefetch into L2 the tile of the CTA `pf_dist` row-tiles ahead (= the CTAs resident at once: the next one
                        // to get this SM slot), so that its loads hit L2 instead of HBM; 0 = off
};

template <int B2, bool CSEL = false> struct Step2 {
    static __device__ __forceinline__ void run(u64 *y)
    {
        constexpr int R2 = 1 << B2;
#pragma unroll
        for (int g = 0; g < 16 / R2; g++) dft_regs<B2, CSEL>(y + g * R2);
    }
};

// One pass: a radix-R DFT (R = 16 * 2^B2) along one mixed-radix digit of the row index, on a tile of
// R rows x 32 columns.  blockDim = (32, 2^B2): a warp is 32 adjacent columns of one row (256-byte
// coalesced segments, conflict-free shared memory, warp-uniform twiddles).
//   LAST = false: in-place positions, output k of the tile is multiplied by w_n^(prev*low*k) (table lookup).
//   LAST = true : natural-order (or index-reversed, for the inverse) output rows, optional scale, canonical words.
//   CW < 32 (ragged last column tile): the 32 lanes of a warp are 32/CW row-tiles x CW columns, so a block of
//   11 columns keeps 22 of 32 lanes busy instead of 11; tile index and inter-pass twiddle then differ between the
//   lane groups (per-lane instead of warp-uniform), everything that depends only on threadIdx.y stays uniform.
// BIG: two-level twiddle lookup / on-the-fly extend scale of the domains above 2^27 -- a separate instantiation, so that the kernel of
// the ordinary domains carries no second copy of its store loop (it cost 6 % of the c2 LDE as a run-time branch: instruction cache).
// PAD: the source may be zero padded (rows >= nvalid read as zeros: the first pass of the extended forward transform).  Every other
// pass is launched with PAD = false and carries no per-row validity tests (16 ISETP + 16 selects per thread on the ALU pipe, which
// bounds the pass).
template <int B2, bool LAST, int CW = kTileCols, bool BIG = false, bool PAD = true>
__global__ void __launch_bounds__(32 << B2, (B2 >= 4 ? 2 : (B2 == 3 ? 4 : 1))) k_pass(const PassParams p)
{
    constexpr int R2 = 1 << B2;
    constexpr int R = 16 * R2;
    extern __shared__ u64 tile[]; // [R][32]   (only when B2 > 0)

    const u32 c = threadIdx.x, t = threadIdx.y; // c = lane: the thread's slot in a shared-memory row
    const u32 col = blockIdx.y * CW + (c % CW);
    const bool colok = col < p.ncols;
    const u64 tau = (u64)blockIdx.x * (kTileCols / CW) + c / CW;
    const u64 hi = tau >> p.log_rest, low = tau & ((1ULL << p.log_rest) - 1);
    const u64 base_in = ((hi * R) << p.log_rest) + low;

    // rows row0 + i1 * rstep, i1 < 16; the first iv of them are below nvalid (the rest read as zeros)
    u64 x[16];
    {
        const u64 row0 = base_in + ((u64)t << p.log_rest);
        const u32 rshift = p.log_rest + B2; // rstep = R2 << log_rest
        u32 iv = colok ? 16u : 0u;
        if (PAD) {
            iv = 0;
            if (colok && p.nvalid > row0) {
                const u64 q = ((p.nvalid - row0 - 1) >> rshift) + 1;
                iv = q < 16 ? (u32)q : 16u;
            }
        }
        const u64 *sp = p.src + row0 * p.spitch + col;
        const u64 sstep = p.spitch << rshift;
#pragma unroll
        for (int i1 = 0; i1 < 16; i1++) {
            x[i1] = (u32)i1 < iv ? *sp : 0;
            sp += sstep;
        }
        if (CW == kTileCols && p.pf_dist && colok) { // the same rows of the tile pf_dist row-tiles ahead (same column tile)
            const u64 tau2 = tau + p.pf_dist;
            if (tau2 < gridDim.x) {
                const u64 row2 = ((((tau2 >> p.log_rest) * R) << p.log_rest) + (tau2 & ((1ULL << p.log_rest) - 1))) + ((u64)t << p.log_rest);
                u32 iv2 = 16u;
                if (PAD) {
                    iv2 = 0;
                    if (p.nvalid > row2) {
                        const u64 q = ((p.nvalid - row2 - 1) >> rshift) + 1;
                        iv2 = q < 16 ? (u32)q : 16u;
                    }
                }
                const u64 *pp = p.src + row2 * p.spitch + col;
#pragma unroll
                for (int i1 = 0; i1 < 16; i1++) {
                    if ((u32)i1 < iv2) asm volatile("prefetch.global.L2 [%0];" ::"l"(pp));
                    pp += sstep;
                }
            }
        }
    }

    // ---- step 1: radix-16 over i1 (row = (i1*R2 + t) within the tile).  With a zero-padded source no thread of the grid has more than
    // ivmax non-zero rows (grid-uniform): the trivial butterflies are skipped
    if (PAD) {
        const u64 ivmax = p.nvalid ? ((p.nvalid - 1) >> (p.log_rest + B2)) + 1 : 0;
        if (ivmax <= 1) dft16_ztop<kCsel, true>(x);
        else if (ivmax <= 8) dft16_ztop<kCsel, false>(x);
        else dft_regs<4, kCsel>(x);
    } else {
        dft_regs<4, kCsel>(x);
    }

    if (B2 > 0) {
        // ---- twiddle w_R^(t*k1), exchange through shared memory, step 2: radix-R2 over t
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const int k1 = bitrev(i, 4);
            u64 v = x[i];
            if (k1 != 0) v = gl::mul(v, c_w256[0][(t * k1) << (4 - B2)]); // w_R^(t*k1) = w_256^(t*k1*256/R)
            tile[(k1 * R2 + t) * kTileCols + c] = v;
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 16; i++) x[i] = tile[(16 * t + i) * kTileCols + c];
        Step2<B2, kCsel>::run(x);
    }

    // ---- store
    u64 out_base = 0;
    if (LAST) { // natural output row = digit-reversed(hi) + prev * k
        u64 h = hi;
        u32 shift = p.log_prev;
#pragma unroll
        for (int q = 2; q >= 0; q--) { // digits of hi, least significant = k_{m-1}; static indices keep p in the param bank
            if (q <= (int)p.npass - 2) {
                shift -= p.bits[q];
                out_base |= (h & ((1ULL << p.bits[q]) - 1)) << shift;
                h >>= p.bits[q];
            }
        }
    }
    const u32 tw_shift = p.log_prev + p.log_roots - p.logn;
    const u64 nmask = (1ULL << p.logn) - 1;
    // output index k = k0 + kc(i): k0 depends on the thread row t, kc on the register index only
    const u32 k0 = B2 > 0 ? (16 * t) / R2 : 0;
    const u64 dstep = p.dpitch << p.log_rest;                                 // non-last: one output index further
    u64 *dp0 = p.dst + (base_in + ((u64)k0 << p.log_rest)) * p.dpitch + col;
    const u32 e1 = (u32)(low << tw_shift);                                    // non-last: twiddle index step per output index (index < 2^log_roots <= 2^27)
    // last pass: the mode is warp-uniform; each mode gets its own copy of the store loop (LMODE 0 forward, 1 inverse
    // scaled by 1/n, 2 inverse scaled by 7^row/n) instead of re-testing it per element
    const u64 lstep = p.dpitch << p.log_prev;
    u64 *lp = p.dst + (out_base + ((u64)k0 << p.log_prev)) * p.dpitch + col;
    auto store_all = [&](auto lmode_tag) {
    constexpr int LMODE = decltype(lmode_tag)::value;
    // walk the outputs in ascending kc so that the store pointer advances by compile-time multiples of dstep
    u32 kprev = 0;
#pragma unroll
    for (int j = 0; j < 16; j++) {
        // j-th smallest kc: kc = g + 16*k2 with g = j % (16/R2), k2 = j / (16/R2); it lives in register i = g*R2 + bitrev(k2)
        constexpr int G = 16 / R2;
        const int g = B2 > 0 ? j % G : 0, k2 = B2 > 0 ? j / G : 0;
        const int i = B2 > 0 ? g * R2 + bitrev(k2, B2) : bitrev(j, 4);
        const u32 kc = B2 > 0 ? (u32)(g + 16 * k2) : (u32)j;
        u64 v = x[i];
        if (!LAST) {
            const u32 ti = e1 * (k0 + kc);
            if (BIG) v = gl::mul(gl::mul(v, p.roots[ti & ((1u << p.log_lo) - 1)]), p.roots_hi[ti >> p.log_lo]);
            else v = gl::mul(v, p.roots[ti]);
            dp0 += dstep * (kc - kprev);
            kprev = kc;
            if (colok) *dp0 = v;
        } else if (LMODE == 0) { // forward: natural rows, one running pointer
            lp += lstep * (kc - kprev);
            kprev = kc;
            if (colok) *lp = gl::canon(v);
        } else {
            const u64 row = (0 - (out_base + ((u64)(k0 + kc) << p.log_prev))) & nmask;
            v = gl::mul(v, LMODE == 1 ? p.cscale : (BIG ? gl::mul(pow_tab(c_p7, (u32)row), c_inv2[p.logn]) : p.rscale[row]));
            if (colok) p.dst[row * p.dpitch + col] = gl::canon(v);
        }
    }
    };
    if (!LAST || !p.inverse) store_all(IntTag<0>());
    else if (p.scale_mode == 2) store_all(IntTag<2>());
    else store_all(IntTag<1>()); // cscale is 1 when the caller asked for no scaling
}

// Table builders: roots[i] = w_(2^log_roots)^i ; rscale[i] = 7^i / 2^logn.
// roots[i] = w_(2^log_roots)^(i << shift), i < count  (shift = 0: the table itself; shift = log_lo: the second-level table)
__global__ void k_build_roots(u64 *roots, u32 log_roots, u64 count, u32 shift)
{
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (u64)gridDim.x * blockDim.x)
        roots[i] = gl::canon(pow_tab(c_w32, (u32)((i << shift) << (32 - log_roots))));
}
__global__ void k_build_rscale(u64 *rscale, u32 logn)
{
    const u64 n = 1ULL << logn;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x)
        rscale[i] = gl::canon(gl::mul(pow_tab(c_p7, (u32)i), c_inv2[logn]));
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
