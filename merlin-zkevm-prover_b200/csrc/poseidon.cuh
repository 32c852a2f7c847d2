// poseidon.cuh -- Poseidon-Goldilocks permutation for sm_100a.
//
// Replaces (reference, CPU AVX2/AVX-512):
//   PoseidonGoldilocks::hash_full_result   src/goldilocks/src/poseidon_goldilocks.cpp:166-228 (seq: 5-37)
//
// Mapping: one thread owns one 12-word state in registers; all lanes of a warp execute the same
// round, so round constants / matrices are warp-uniform operands read from __constant__ memory.
//
// What the measured machine dictates (profiles/r1_01_ubench.txt, profiles/r1_01_leaf_ncu.txt):
//   * IMAD.WIDE issues at 16 lanes/clk/SMSP (2 clk per warp instruction), IADD3-class at ~22
//     lanes/clk/SMSP, and the two do NOT overlap in practice: cost ~ 2*#IMAD + 1.5*#ALU clocks.
//     So every instruction counts; a modular multiply is 5 IMAD.WIDE + 11 ALU (~26 clk).
//   * The first version (88 KB of straight-line SASS) stalled mostly on instruction fetch
//     ("no_instruction" was the top stall).  The whole permutation is now ONE loop of 19 stages
//     (8 full rounds, 11 pairs of partial rounds) and ~30 KB of SASS.
//   * The 22 partial rounds are evaluated in the DENSE form (x^7 + c on word 0, then the same small-entry MDS layer
//     as a full round), two rounds per layer through the squared matrix (mds_pair) -- not in the reference's sparse
//     form, whose 64-bit coefficients cost 23 full modular multiplications per round on the integer pipe; see permute().
//   * The circulant MDS layers use the small entries (< 64) directly: 24 multiply-adds per output on
//     32-bit halves and one 96-bit reduction -- issued as exact DFMAs on the otherwise idle FP64 pipe
//     (mds_layer below; profiles/r1_09_ubench_fp64.txt, r1_10_perm_mds_fp64_vs_int.txt).
#pragma once
#include "gl64.cuh"
#include "poseidon_constants.cuh"

namespace poseidon {

using gl::u32;
using gl::u64;

// Tables derived from GLP_RC by upload_constants() (poseidon_kernels.cuh):
__constant__ u64 c_RC0[12];          // pre-round constants C[0..12)
__constant__ u64 c_RCF[8 * 12];      // constants added after the S-box of full round r (row 7 = 0)
__constant__ u64 c_RCP[22];          // partial-round constants C[60..82)

// MDS circulant (reference M[j][i] = circ[(i-j) mod 12] + 8*[i==j==0]) as immediates.
__device__ __forceinline__ constexpr u32 mds_entry(int j, int i)
{
    constexpr u32 circ[12] = {17, 20, 34, 18, 39, 13, 13, 28, 2, 16, 41, 15};
    return circ[(i - j + 12) % 12] + ((i == 0 && j == 0) ? 8u : 0u);
}

#ifndef GLB_POSEIDON_SQR3
#define GLB_POSEIDON_SQR3 1 // both squarings of the S-box with three 32 x 32 products (gl::sqr3): 1584 -> 1601 M perm/s (profiles/r2_08); 0: a * a
#endif
#ifndef GLB_POSEIDON_RED_ALU
#define GLB_POSEIDON_RED_ALU 0 // round-2 experiment: 1 = S-box reductions, 2 = also the 96-bit reductions of the MDS layers, in ALU-only form
#endif
// x^7 + c   (c any representative)
__device__ __forceinline__ u64 sbox_add(u64 x, u64 c)
{
#if GLB_POSEIDON_RED_ALU
    u64 x2 = gl::sqr_alu(x);
    u64 x3 = gl::mul_alu(x, x2);
    u64 x4 = gl::sqr_alu(x2);
    return gl::mad_alu(x3, x4, c);
#elif defined(GLB_POSEIDON_SQR3) && GLB_POSEIDON_SQR3
    // A/B: squarings with three 32 x 32 products (gl::sqr3); 2: only the first of the two
    u64 x2 = gl::sqr3(x);
    u64 x3 = gl::mul(x, x2);
    u64 x4 = GLB_POSEIDON_SQR3 == 2 ? gl::sqr(x2) : gl::sqr3(x2);
    return gl::mad(x3, x4, c);
#else
    u64 x2 = gl::sqr(x);
    u64 x3 = gl::mul(x, x2);
    u64 x4 = gl::sqr(x2);
    return gl::mad(x3, x4, c);
#endif
}
#ifndef GLB_POSEIDON_FOLD
#define GLB_POSEIDON_FOLD 1 // 1 / 2: the 96-bit reduction of the MDS folds with one IMAD.WIDE instead of two (gl::reduce96_one<1 / 2>); 0: gl::reduce96
#endif
// Three places that spent an IMAD.WIDE (4 issue clocks, nothing beside it) on a conditional +-(2^32 - 1): the second step of the MDS
// folds (GLB_POSEIDON_FOLD), the canonicalisation inside mds_pair's y - x (GLB_POSEIDON_SUB_SEL) and the 12 additions of the
// pre-round constants (GLB_POSEIDON_RC0_CX).  Leaf kernel on 2^20 x 600: 1601.5 -> 1613.9 (fold) -> 1621.2 M perm/s (all three),
// profiles/r2_37_leaf_wide_saving_ab.txt; 0 restores the round-1 forms for A/B builds.
#ifndef GLB_POSEIDON_SUB_SEL
#define GLB_POSEIDON_SUB_SEL 1
#endif
#ifndef GLB_POSEIDON_RC0_CX
#define GLB_POSEIDON_RC0_CX 1
#endif
__device__ __forceinline__ u64 red96(u32 x0, u32 x1, u32 x2)
{
#if GLB_POSEIDON_FOLD
    return gl::reduce96_one<GLB_POSEIDON_FOLD>(x0, x1, x2);
#elif GLB_POSEIDON_RED_ALU >= 2
    return gl::reduce96_alu(x0, x1, x2);
#else
    return gl::reduce96(x0, x1, x2);
#endif
}

// new[i] = sum_j M[j][i] old[j]  (poseidon_goldilocks.hpp:148-162).  Entries < 64, so the sums of
// the low and of the high 32-bit halves each fit 64 bits (< 2^42); one 96-bit reduction per output.
//
// The 24 multiply-adds per output run on the FP64 pipe, which issues beside the integer pipe on this SM
// (profiles/r1_09_ubench_fp64.txt: DFMA + IMAD.WIDE pairs issue in the time of one): a 32-bit half h becomes a
// double with ONE I2F.F64.U32 (exact; it issues on the conversion pipe, which is otherwise idle -- the earlier
// (2^52 + h) - 2^52 form cost a DADD on the FP64 pipe plus ~2.3 register moves per half to build the pair:
// 1359 -> 1398 M perm/s, profiles/r1_21_leaf_variants.txt), every partial sum 2^52 + sum m*h is an integer
// below 2^53 and therefore exact, and the low 52 bits of the final double ARE the integer sum.
__device__ __forceinline__ double u32_to_f64(u32 h) { return __uint2double_rn(h); }

constexpr double kTwo52 = 4503599627370496.0;
constexpr u64 kTwo52Bits = 0x4330000000000000ULL;
// -(x) mod p for a 128-bit x (compile time)
__host__ __device__ constexpr u64 neg_mod_p(unsigned __int128 x) { return (u64)((gl::P - (u64)(x % gl::P)) % gl::P); }

#ifndef GLB_MDS_SPLIT
#define GLB_MDS_SPLIT 2 // 2: two CRT levels (default); 1: one level; 0: plain 12 x 12 form (1554 / 1398 M perm/s against 1584: kept for A/B runs of tools/ubench)
#endif
// ---- split-circulant evaluation -----------------------------------------------------------------------------------
// new = c (*) old + 8 old[0] e_0 with c the MDS circulant (cyclic convolution of length 12).  By the CRT on
// x^12 - 1 = (x^6 - 1)(x^6 + 1) the product is a cyclic product of length 6 on (v_lo + v_hi) and a negacyclic one on
// (v_lo - v_hi); c_lo +- c_hi is even, so the halved coefficients are integers and
//     y[k] = P_k + N_k,  y[k+6] = P_k - N_k,   P_k = sum_a cp[(k-a) mod 6] vp[a],  N_k = sum_a +-cm[(k-a) mod 6] vm[a]:
// 72 multiply-adds + 24 additions per 32-bit half instead of 144 (tests/test_poseidon_dense_form.py proves the identity
// and the < 2^53 bounds with exact integers).  The same holds for the circulant part c (*) c of the squared matrix.
__host__ __device__ constexpr int circ12(int k)
{
    constexpr int circ[12] = {17, 20, 34, 18, 39, 13, 13, 28, 2, 16, 41, 15};
    return circ[k % 12];
}
__host__ __device__ constexpr int circ12sq(int k) // (c (*) c)[k]
{
    int acc = 0;
    for (int a = 0; a < 12; a++) acc += circ12(a) * circ12((k - a + 12) % 12);
    return acc;
}
struct CircM {  // the MDS circulant
    static __host__ __device__ constexpr int c(int k) { return circ12(k); }
};
struct CircM2 { // circulant part of the squared matrix
    static __host__ __device__ constexpr int c(int k) { return circ12sq(k); }
};
template <class K> __host__ __device__ constexpr bool circ_halves_even()
{
    for (int k = 0; k < 6; k++)
        if ((K::c(k) + K::c(k + 6)) % 2 != 0) return false;
    return true;
}
static_assert(circ_halves_even<CircM>() && circ_halves_even<CircM2>(), "halved circulant coefficients must be integers");

// vp/vm of one 32-bit half plane
__device__ __forceinline__ void split_pm(const double (&v)[12], double (&vp)[6], double (&vm)[6])
{
#pragma unroll
    for (int a = 0; a < 6; a++) {
        vp[a] = v[a] + v[a + 6];
        vm[a] = v[a] - v[a + 6];
    }
}
// P_k (started at `bias`) and N_k of output pair (k, k+6)
template <class K, int k>
__device__ __forceinline__ void cyc_pair(const double (&vp)[6], const double (&vm)[6], double bias, double &P, double &N)
{
    P = bias;
#pragma unroll
    for (int a = 0; a < 6; a++) {
        const int idx = (k - a + 6) % 6;
        const double cp = (double)((K::c(idx) + K::c(idx + 6)) / 2);
        const double cm = (double)((a <= k ? 1 : -1) * ((K::c(idx) - K::c(idx + 6)) / 2));
        P = fma(cp, vp[a], P);
        N = a == 0 ? cm * vm[0] : fma(cm, vm[a], N);
    }
}

// ---- GLB_MDS_SPLIT >= 2 (default): a second CRT level on the cyclic half, x^6 - 1 = (x^3 - 1)(x^3 + 1):
// P_k = PP_k + PN_k, P_{k+3} = PP_k - PN_k with a cyclic / negacyclic product of length 3 on vp_lo +- vp_hi -- 30 instead
// of 36 operations per half plane (tests/test_poseidon_dense_form.py::test_two_level_split_equals_one_level).
template <class K> __host__ __device__ constexpr int circ_cp(int i) { return (K::c(i % 6) + K::c(i % 6 + 6)) / 2; }
template <class K> __host__ __device__ constexpr bool circ_quarters_even()
{
    for (int k = 0; k < 3; k++)
        if ((circ_cp<K>(k) + circ_cp<K>(k + 3)) % 2 != 0) return false;
    return true;
}
static_assert(circ_quarters_even<CircM>() && circ_quarters_even<CircM2>(), "quartered circulant coefficients must be integers");
__device__ __forceinline__ void split_pm3(const double (&vp)[6], double (&vpp)[3], double (&vpm)[3])
{
#pragma unroll
    for (int a = 0; a < 3; a++) {
        vpp[a] = vp[a] + vp[a + 3];
        vpm[a] = vp[a] - vp[a + 3];
    }
}
template <class K, int k> // k < 3
__device__ __forceinline__ void cyc_p2(const double (&vpp)[3], const double (&vpm)[3], double bias, double &Pk, double &Pk3)
{
    double pp = bias, pn = 0.0;
#pragma unroll
    for (int a = 0; a < 3; a++) {
        const int idx = (k - a + 3) % 3;
        const double cpp = (double)((circ_cp<K>(idx) + circ_cp<K>(idx + 3)) / 2);
        const double cpm = (double)((a <= k ? 1 : -1) * ((circ_cp<K>(idx) - circ_cp<K>(idx + 3)) / 2));
        pp = fma(cpp, vpp[a], pp);
        pn = a == 0 ? cpm * vpm[0] : fma(cpm, vpm[a], pn);
    }
    Pk = pp + pn;
    Pk3 = pp - pn;
}
template <class K, int k> // k < 6
__device__ __forceinline__ double cyc_n(const double (&vm)[6])
{
    double N = 0.0;
#pragma unroll
    for (int a = 0; a < 6; a++) {
        const int idx = (k - a + 6) % 6;
        const double cm = (double)((a <= k ? 1 : -1) * ((K::c(idx) - K::c(idx + 6)) / 2));
        N = a == 0 ? cm * vm[0] : fma(cm, vm[a], N);
    }
    return N;
}
// (L, H) raw biased sums -> word
__device__ __forceinline__ u64 fold_lh(double l, double h)
{
    const u64 L = (u64)__double_as_longlong(l), H = (u64)__double_as_longlong(h);
    const u64 t = (L >> 32) + H;
    return red96((u32)L, (u32)t, (u32)(t >> 32));
}

template <int V> struct KTag { static constexpr int value = V; };

template <int NOUT = 12>
__device__ __forceinline__ void mds_layer(u64 (&s)[12])
{
    // bias of the two sums, value L + H 2^32: kTwo52Bits (1 + 2^32); cancelled by 32-bit offsets in the starting values
    constexpr u64 DM = neg_mod_p((unsigned __int128)(kTwo52Bits % gl::P) * (((u64)1 << 32) + 1));
    double lo[12], hi[12];
#pragma unroll
    for (int j = 0; j < 12; j++) {
        lo[j] = u32_to_f64((u32)s[j]);
        hi[j] = u32_to_f64((u32)(s[j] >> 32));
    }
    if constexpr (NOUT == 12 && GLB_MDS_SPLIT >= 2) {
        const double lo0 = lo[0], hi0 = hi[0];
        double lp[6], lm[6], hp[6], hm[6], lpp[3], lpm[3], hpp[3], hpm[3];
        split_pm(lo, lp, lm);
        split_pm(hi, hp, hm);
        split_pm3(lp, lpp, lpm);
        split_pm3(hp, hpp, hpm);
        auto emit = [&](auto ktag) { // outputs k, k+3, k+6, k+9
            constexpr int k = decltype(ktag)::value;
            double pl, pl3, ph, ph3;
            cyc_p2<CircM, k>(lpp, lpm, kTwo52 + (double)(DM & 0xFFFFFFFFu), pl, pl3);
            cyc_p2<CircM, k>(hpp, hpm, kTwo52 + (double)(DM >> 32), ph, ph3);
            const double nl = cyc_n<CircM, k>(lm), nh = cyc_n<CircM, k>(hm), nl3 = cyc_n<CircM, k + 3>(lm), nh3 = cyc_n<CircM, k + 3>(hm);
            double a_l = pl + nl, a_h = ph + nh;
            if (k == 0) { a_l = fma(8.0, lo0, a_l); a_h = fma(8.0, hi0, a_h); }
            s[k] = fold_lh(a_l, a_h);
            s[k + 6] = fold_lh(pl - nl, ph - nh);
            s[k + 3] = fold_lh(pl3 + nl3, ph3 + nh3);
            s[k + 9] = fold_lh(pl3 - nl3, ph3 - nh3);
        };
        emit(KTag<0>()); emit(KTag<1>()); emit(KTag<2>());
    } else if constexpr (NOUT == 12 && GLB_MDS_SPLIT != 0) {
        const double lo0 = lo[0], hi0 = hi[0];
        double lp[6], lm[6], hp[6], hm[6];
        split_pm(lo, lp, lm);
        split_pm(hi, hp, hm);
        auto emit = [&](auto ktag) {
            constexpr int k = decltype(ktag)::value;
            double pl, nl, ph, nh;
            cyc_pair<CircM, k>(lp, lm, kTwo52 + (double)(DM & 0xFFFFFFFFu), pl, nl);
            cyc_pair<CircM, k>(hp, hm, kTwo52 + (double)(DM >> 32), ph, nh);
            double a_l = pl + nl, a_h = ph + nh;
            const double b_l = pl - nl, b_h = ph - nh;
            if (k == 0) { a_l = fma(8.0, lo0, a_l); a_h = fma(8.0, hi0, a_h); } // the +8 on M[0][0]
            {
                const u64 L = (u64)__double_as_longlong(a_l), H = (u64)__double_as_longlong(a_h);
                const u64 t = (L >> 32) + H;
                s[k] = red96((u32)L, (u32)t, (u32)(t >> 32));
            }
            {
                const u64 L = (u64)__double_as_longlong(b_l), H = (u64)__double_as_longlong(b_h);
                const u64 t = (L >> 32) + H;
                s[k + 6] = red96((u32)L, (u32)t, (u32)(t >> 32));
            }
        };
        emit(KTag<0>()); emit(KTag<1>()); emit(KTag<2>()); emit(KTag<3>()); emit(KTag<4>()); emit(KTag<5>());
    } else {
#pragma unroll
        for (int i = 0; i < NOUT; i++) {
            double dl = kTwo52 + (double)(DM & 0xFFFFFFFFu), dh = kTwo52 + (double)(DM >> 32);
#pragma unroll
            for (int j = 0; j < 12; j++) {
                dl = fma((double)mds_entry(j, i), lo[j], dl);
                dh = fma((double)mds_entry(j, i), hi[j], dh);
            }
            const u64 L = (u64)__double_as_longlong(dl), H = (u64)__double_as_longlong(dh); // raw: bias included
            const u64 t = (L >> 32) + H; // < 2^64
            s[i] = red96((u32)L, (u32)t, (u32)(t >> 32));
        }
    }
}

// (M*M)[j][i] = sum_k M[j][k] M[k][i]: entries <= 6818, column sums <= 67784 < 2^17 (so a 12-term sum of 32-bit halves
// times these entries stays below 2^49 and is exact in a double next to the 2^52 bias).
__device__ __forceinline__ constexpr u32 mds2_entry(int j, int i)
{
    u32 acc = 0;
    for (int k = 0; k < 12; k++) acc += mds_entry(j, k) * mds_entry(k, i);
    return acc;
}

// TWO consecutive partial rounds in one layer.  On entry s = a_r (the state of round r with its S-box already applied to
// word 0).  Round r+1 only needs word 0 of M a_r:  x = <M[.][0], a_r>,  y = x^7 + c_next,  d = y - x;  then
//     v_{r+2} = M (M a_r + d e_0) = (M M) a_r + d M[0][.],
// so the second matrix product collapses into the first with the squared matrix plus a rank-1 term: 12 + 1 outputs of
// 24 (+2) DFMAs instead of 24 outputs, and one reduction per word instead of two.
__device__ __forceinline__ void mds_pair(u64 (&s)[12], u64 c_next)
{
    constexpr u64 DM = neg_mod_p((unsigned __int128)(kTwo52Bits % gl::P) * (((u64)1 << 32) + 1));
    double lo[12], hi[12];
#pragma unroll
    for (int j = 0; j < 12; j++) {
        lo[j] = u32_to_f64((u32)s[j]);
        hi[j] = u32_to_f64((u32)(s[j] >> 32));
    }
    u64 x;
    {
        double dl = kTwo52 + (double)(DM & 0xFFFFFFFFu), dh = kTwo52 + (double)(DM >> 32);
#pragma unroll
        for (int j = 0; j < 12; j++) {
            dl = fma((double)mds_entry(j, 0), lo[j], dl);
            dh = fma((double)mds_entry(j, 0), hi[j], dh);
        }
        const u64 L = (u64)__double_as_longlong(dl), H = (u64)__double_as_longlong(dh);
        const u64 t = (L >> 32) + H;
        x = red96((u32)L, (u32)t, (u32)(t >> 32));
    }
    const u64 y = sbox_add(x, c_next);
#if defined(GLB_POSEIDON_SUB_SEL) && GLB_POSEIDON_SUB_SEL
    const u64 d = gl::sub_c(y, gl::canon_sel(x)); // canonicalisation without an IMAD.WIDE
#else
    const u64 d = gl::sub(y, x);
#endif
    const double dlo = u32_to_f64((u32)d), dhi = u32_to_f64((u32)(d >> 32));
#if GLB_MDS_SPLIT >= 2
    {
        const double wl = fma(8.0, lo[0], dlo), wh = fma(8.0, hi[0], dhi);
        const double yl8 = 8.0 * u32_to_f64((u32)y), yh8 = 8.0 * u32_to_f64((u32)(y >> 32));
        double lp[6], lm[6], hp[6], hm[6], lpp[3], lpm[3], hpp[3], hpm[3];
        split_pm(lo, lp, lm);
        split_pm(hi, hp, hm);
        split_pm3(lp, lpp, lpm);
        split_pm3(hp, hpp, hpm);
        auto emit = [&](auto ktag) { // outputs k, k+3, k+6, k+9
            constexpr int k = decltype(ktag)::value;
            double pl, pl3, ph, ph3;
            cyc_p2<CircM2, k>(lpp, lpm, kTwo52 + (double)(DM & 0xFFFFFFFFu), pl, pl3);
            cyc_p2<CircM2, k>(hpp, hpm, kTwo52 + (double)(DM >> 32), ph, ph3);
            const double nl = cyc_n<CircM2, k>(lm), nh = cyc_n<CircM2, k>(hm), nl3 = cyc_n<CircM2, k + 3>(lm), nh3 = cyc_n<CircM2, k + 3>(hm);
            double a_l = fma((double)circ12(k), wl, pl + nl), a_h = fma((double)circ12(k), wh, ph + nh);
            if (k == 0) { a_l += yl8; a_h += yh8; }
            s[k] = fold_lh(a_l, a_h);
            s[k + 6] = fold_lh(fma((double)circ12(k + 6), wl, pl - nl), fma((double)circ12(k + 6), wh, ph - nh));
            s[k + 3] = fold_lh(fma((double)circ12(k + 3), wl, pl3 + nl3), fma((double)circ12(k + 3), wh, ph3 + nh3));
            s[k + 9] = fold_lh(fma((double)circ12(k + 9), wl, pl3 - nl3), fma((double)circ12(k + 9), wh, ph3 - nh3));
        };
        emit(KTag<0>()); emit(KTag<1>()); emit(KTag<2>());
    }
#elif GLB_MDS_SPLIT
    {
        // (M M) a + d M[0][.]  =  (c (*) c) (*) a + (8 a_0 + d) c + 8 (x + d) e_0, and x + d = y (mod p): per 32-bit half
        const double wl = fma(8.0, lo[0], dlo), wh = fma(8.0, hi[0], dhi);
        const double yl8 = 8.0 * u32_to_f64((u32)y), yh8 = 8.0 * u32_to_f64((u32)(y >> 32));
        double lp[6], lm[6], hp[6], hm[6];
        split_pm(lo, lp, lm);
        split_pm(hi, hp, hm);
        auto emit = [&](auto ktag) {
            constexpr int k = decltype(ktag)::value;
            double pl, nl, ph, nh;
            cyc_pair<CircM2, k>(lp, lm, kTwo52 + (double)(DM & 0xFFFFFFFFu), pl, nl);
            cyc_pair<CircM2, k>(hp, hm, kTwo52 + (double)(DM >> 32), ph, nh);
            double a_l = fma((double)circ12(k), wl, pl + nl), a_h = fma((double)circ12(k), wh, ph + nh);
            const double b_l = fma((double)circ12(k + 6), wl, pl - nl), b_h = fma((double)circ12(k + 6), wh, ph - nh);
            if (k == 0) { a_l += yl8; a_h += yh8; }
            {
                const u64 L = (u64)__double_as_longlong(a_l), H = (u64)__double_as_longlong(a_h);
                const u64 t = (L >> 32) + H;
                s[k] = red96((u32)L, (u32)t, (u32)(t >> 32));
            }
            {
                const u64 L = (u64)__double_as_longlong(b_l), H = (u64)__double_as_longlong(b_h);
                const u64 t = (L >> 32) + H;
                s[k + 6] = red96((u32)L, (u32)t, (u32)(t >> 32));
            }
        };
        emit(KTag<0>()); emit(KTag<1>()); emit(KTag<2>()); emit(KTag<3>()); emit(KTag<4>()); emit(KTag<5>());
    }
#else
#pragma unroll
    for (int i = 0; i < 12; i++) {
        double dl = kTwo52 + (double)(DM & 0xFFFFFFFFu), dh = kTwo52 + (double)(DM >> 32);
#pragma unroll
        for (int j = 0; j < 12; j++) {
            dl = fma((double)mds2_entry(j, i), lo[j], dl);
            dh = fma((double)mds2_entry(j, i), hi[j], dh);
        }
        dl = fma((double)mds_entry(0, i), dlo, dl);
        dh = fma((double)mds_entry(0, i), dhi, dh);
        const u64 L = (u64)__double_as_longlong(dl), H = (u64)__double_as_longlong(dh); // raw: bias included
        const u64 t = (L >> 32) + H; // < 2^64
        s[i] = red96((u32)L, (u32)t, (u32)(t >> 32));
    }
#endif
}

// Full permutation.  Input: any representatives; output: lazily reduced.  OUT4: only s[0..4) of the result
// are computed (all the sponge and the tree ever read: poseidon_goldilocks.cpp:242-249, 292-295) -- the last
// MDS layer then costs a third; s[4..12) are left undefined.
//
// DENSE FORM (default).  The reference evaluates the 22 partial rounds in the "optimised" form: a dense 12x12
// matrix P, then per round a 12-term dot product and a rank-1 update with 64-bit coefficients S
// (poseidon_goldilocks.cpp:17-28).  Those matrices are a re-factorisation of the MDS matrix M: with
// w_r = T_r v_r, T_r = diag(1, T^_r) acting on coordinates 1..11 only (so it commutes with the S-box on coordinate
// 0 and with adding a constant there), S_r = T_{r+1} M T_r^-1, P = T_0 M and T_22 = I, the same state sequence is
//     v_0 = M (u + C[48..60)),     v_{r+1} = M (sigma_0(v_r) + C[60+r] e_0),   r = 0..21,
// i.e. every partial round is "x^7 + c on word 0, then the SAME small-entry MDS layer as a full round".  On this
// machine that form is far cheaper: the MDS layer runs as exact DFMAs on the FP64 pipe beside the integer pipe
// (mds_layer), whereas the sparse form costs 23 full 64-bit modular multiplications per round on the integer pipe.
// The two forms are equal as maps on F_p^12 (tests/test_poseidon_dense_form.py checks it on the reference's constants;
// every GPU parity test checks it against the reference itself).  The sparse form and the variants that were measured
// on the way (integer MDS, FP64 dot products, one layer per partial round) are in the history and in DESIGN.md 4.2.
template <bool OUT4 = false>
__device__ __forceinline__ void permute(u64 (&s)[12])
{
#pragma unroll
#if defined(GLB_POSEIDON_RC0_CX) && GLB_POSEIDON_RC0_CX
    for (int i = 0; i < 12; i++) s[i] = gl::add_cx(s[i], c_RC0[i]); // the IMAD.WIDE-free addition of the NTT butterflies (constants are canonical)
#else
    for (int i = 0; i < 12; i++) s[i] = gl::add_c(s[i], c_RC0[i]);
#endif
    // 19 stages, one loop body: stages 0..3 and 15..18 are full rounds (S-box on all 12 words, MDS layer), stages 4..14
    // are PAIRS of partial rounds (mds_pair)
#pragma unroll 1
    for (int st = 0; st < 19; st++) {
        if (st < 4 || st >= 15) {
            const int f = st < 4 ? st : st - 11;
#pragma unroll
            for (int i = 0; i < 12; i++) s[i] = sbox_add(s[i], c_RCF[f * 12 + i]);
            if (OUT4 && st == 18) mds_layer<4>(s);
            else mds_layer(s);
        } else {
            const int q = st - 4;
            s[0] = sbox_add(s[0], c_RCP[2 * q]);
            mds_pair(s, c_RCP[2 * q + 1]);
        }
    }
}

} // namespace poseidon
