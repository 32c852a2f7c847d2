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
//     ("no_instruction" was the top stall).  All 8 full rounds now share ONE loop body and the 22
//     partial rounds another; the whole permutation is ~30 KB of SASS.
//   * The 22 partial rounds are evaluated in the DENSE form (x^7 + c on word 0, then the same small-entry MDS layer
//     as a full round), two rounds per layer through the squared matrix (mds_pair) -- not in the reference's sparse
//     form, whose 64-bit coefficients cost 23 full modular multiplications per round on the integer pipe; see permute().
//   * The circulant MDS layers use the small entries (< 64) directly: 24 multiply-adds per output on
//     32-bit halves and one 96-bit reduction -- issued as exact DFMAs on the otherwise idle FP64 pipe
//     (mds_layer below; profiles/r1_09_ubench_fp64.txt, r1_10_perm_mds_fp64_vs_int.txt).
#pragma once
#include "gl64.cuh"
#include "poseidon_constants.cuh"

// loop unrolling of the dense layer / the partial rounds (tools/ubench leaf: 1 is best or equal)
#ifndef GLB_PARTIAL_UNROLL
#define GLB_PARTIAL_UNROLL 1
#endif
#ifndef GLB_DENSE_UNROLL
#define GLB_DENSE_UNROLL 1
#endif

namespace poseidon {

using gl::u32;
using gl::u64;

// Tables derived from GLP_RC / GLP_PM / GLP_SP by upload_constants() (poseidon_kernels.cuh):
__constant__ u64 c_RC0[12];          // pre-round constants C[0..12)
__constant__ u64 c_RCF[8 * 12];      // constants added after the S-box of full round r (row 7 = 0)
__constant__ u64 c_RCP[22];          // partial-round constants C[60..82)
#ifdef GLB_POSEIDON_SPARSE // tables of the reference's sparse form (dense matrix P, sparse matrices S); see permute()
__constant__ uint4 c_PL[12 * 12];    // dense layer: limbs of P[j][i] at [i*12 + j] (x,y,z; w unused)
__constant__ uint4 c_SL[22 * 12];    // partial round r: limbs of the dot row S[23r + i] at [r*12 + i]
__constant__ u64 c_SV[22 * 11];      // partial round r: rank-1 column S[23r + 12 + i], i < 11
__constant__ double4 c_PLd[12 * 12]; // the same dense-layer / dot-row constants as four 16-bit limbs held as doubles
__constant__ double4 c_SLd[22 * 12];
#endif

// MDS circulant (reference M[j][i] = circ[(i-j) mod 12] + 8*[i==j==0]) as immediates.
__device__ __forceinline__ constexpr u32 mds_entry(int j, int i)
{
    constexpr u32 circ[12] = {17, 20, 34, 18, 39, 13, 13, 28, 2, 16, 41, 15};
    return circ[(i - j + 12) % 12] + ((i == 0 && j == 0) ? 8u : 0u);
}

// x^7 + c   (c any representative)
__device__ __forceinline__ u64 sbox_add(u64 x, u64 c)
{
    u64 x2 = gl::sqr(x);
    u64 x3 = gl::mul(x, x2);
    u64 x4 = gl::sqr(x2);
    return gl::mad(x3, x4, c);
}

// new[i] = sum_j M[j][i] old[j]  (poseidon_goldilocks.hpp:148-162).  Entries < 64, so the sums of
// the low and of the high 32-bit halves each fit 64 bits (< 2^42); one 96-bit reduction per output.
//
// The 24 multiply-adds per output run on the FP64 pipe, which issues beside the integer pipe on this SM
// (profiles/r1_09_ubench_fp64.txt: DFMA + IMAD.WIDE pairs issue in the time of one): a 32-bit half h is the
// double (2^52 + h) - 2^52 (register pairing + one DADD), every partial sum 2^52 + sum m*h is an integer
// below 2^53 and therefore exact, and the low 52 bits of the final double ARE the integer sum.
__device__ __forceinline__ double u32_to_f64(u32 h) { return __hiloint2double(0x43300000, (int)h) - 4503599627370496.0; }

constexpr double kTwo52 = 4503599627370496.0;
constexpr u64 kTwo52Bits = 0x4330000000000000ULL;
// -(x) mod p for a 128-bit x (compile time)
__host__ __device__ constexpr u64 neg_mod_p(unsigned __int128 x) { return (u64)((gl::P - (u64)(x % gl::P)) % gl::P); }

// Lazy dot product on the FP64 pipe: sum of up to 12 products a*c, a any u64 (as two exact doubles a0, a1 < 2^32),
// c a constant given as four 16-bit limbs held as doubles.  Eight columns col[h][l] of weight 2^(32h+16l); every
// partial sum is 2^52 + (an integer < 2^52) and therefore exact.  The 2^52 biases are never masked off: the
// starting values of the h = 0 columns carry 16-bit offsets d_l with  sum_l d_l 2^(16l) = -(bias total) mod p,
// so the RAW bit patterns of the eight doubles already sum to the dot product mod p.
struct DotF {
    double c0[4], c1[4];
    __device__ __forceinline__ void init()
    {
        // bias total = kTwo52Bits * sum_{h,l} 2^(32h+16l)
        constexpr unsigned __int128 W = ((unsigned __int128)1 + ((unsigned __int128)1 << 16) + ((unsigned __int128)1 << 32) + ((unsigned __int128)1 << 48)) *
                                        (((unsigned __int128)1 << 32) + 1);
        constexpr u64 D = neg_mod_p((unsigned __int128)(kTwo52Bits % gl::P) * (W % gl::P));
#pragma unroll
        for (int l = 0; l < 4; l++) {
            c0[l] = kTwo52 + (double)((D >> (16 * l)) & 0xFFFF);
            c1[l] = kTwo52;
        }
    }
    __device__ __forceinline__ void mad(double a0, double a1, const double4 &k)
    {
        c0[0] = fma(a0, k.x, c0[0]); c0[1] = fma(a0, k.y, c0[1]); c0[2] = fma(a0, k.z, c0[2]); c0[3] = fma(a0, k.w, c0[3]);
        c1[0] = fma(a1, k.x, c1[0]); c1[1] = fma(a1, k.y, c1[1]); c1[2] = fma(a1, k.z, c1[2]); c1[3] = fma(a1, k.w, c1[3]);
    }
    __device__ __forceinline__ u64 fold() const
    {
        typedef unsigned __int128 w128;
        const w128 i0 = (w128)(u64)__double_as_longlong(c0[0]) + ((w128)(u64)__double_as_longlong(c0[1]) << 16) +
                        ((w128)(u64)__double_as_longlong(c0[2]) << 32) + ((w128)(u64)__double_as_longlong(c0[3]) << 48);
        const w128 i1 = (w128)(u64)__double_as_longlong(c1[0]) + ((w128)(u64)__double_as_longlong(c1[1]) << 16) +
                        ((w128)(u64)__double_as_longlong(c1[2]) << 32) + ((w128)(u64)__double_as_longlong(c1[3]) << 48);
        // i0, i1 < 2^112; i0 + i1 2^32 < 2^145: five words, the top one < 2^17
        const u64 lo = (u64)i0, hi = (u64)(i0 >> 64), l1 = (u64)i1, h1 = (u64)(i1 >> 64);
        u32 x0 = (u32)lo, x1 = (u32)(lo >> 32), x2 = (u32)hi, x3 = (u32)(hi >> 32), x4;
        asm("add.cc.u32 %0, %0, %4;\n\taddc.cc.u32 %1, %1, %5;\n\taddc.cc.u32 %2, %2, %6;\n\taddc.u32 %3, %7, 0;"
            : "+r"(x1), "+r"(x2), "+r"(x3), "=r"(x4)
            : "r"((u32)l1), "r"((u32)(l1 >> 32)), "r"((u32)h1), "r"((u32)(h1 >> 32)));
        return gl::reduce160(x0, x1, x2, x3, x4);
    }
};

template <int NOUT = 12>
__device__ __forceinline__ void mds_layer(u64 (&s)[12])
{
#ifndef GLB_MDS_INT
    // bias of the two sums, value L + H 2^32: kTwo52Bits (1 + 2^32); cancelled by 32-bit offsets in the starting values
    constexpr u64 DM = neg_mod_p((unsigned __int128)(kTwo52Bits % gl::P) * (((u64)1 << 32) + 1));
    double lo[12], hi[12];
#pragma unroll
    for (int j = 0; j < 12; j++) {
        lo[j] = u32_to_f64((u32)s[j]);
        hi[j] = u32_to_f64((u32)(s[j] >> 32));
    }
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
        s[i] = gl::reduce96((u32)L, (u32)t, (u32)(t >> 32));
    }
#else
    u32 lo[12], hi[12];
#pragma unroll
    for (int j = 0; j < 12; j++) {
        lo[j] = (u32)s[j];
        hi[j] = (u32)(s[j] >> 32);
    }
#pragma unroll
    for (int i = 0; i < NOUT; i++) {
        u64 L = 0, H = 0;
#pragma unroll
        for (int j = 0; j < 12; j++) {
            gl::madw(L, lo[j], mds_entry(j, i));
            gl::madw(H, hi[j], mds_entry(j, i));
        }
        u64 t = (L >> 32) + H; // < 2^43
        s[i] = gl::reduce96((u32)L, (u32)t, (u32)(t >> 32));
    }
#endif
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
        x = gl::reduce96((u32)L, (u32)t, (u32)(t >> 32));
    }
    const u64 y = sbox_add(x, c_next);
    const u64 d = gl::sub(y, x);
    const double dlo = u32_to_f64((u32)d), dhi = u32_to_f64((u32)(d >> 32));
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
        s[i] = gl::reduce96((u32)L, (u32)t, (u32)(t >> 32));
    }
}

#ifdef GLB_POSEIDON_SPARSE
// Dense layer before the partial rounds: new[i] = sum_j P[j][i] old[j]   (poseidon_goldilocks.cpp:17)
__device__ __forceinline__ void dense_layer(u64 (&s)[12])
{
#ifdef GLB_DENSE_F64
    double o0[12], o1[12];
#pragma unroll
    for (int j = 0; j < 12; j++) {
        o0[j] = u32_to_f64((u32)s[j]);
        o1[j] = u32_to_f64((u32)(s[j] >> 32));
    }
#else
    u64 o[12];
#pragma unroll
    for (int j = 0; j < 12; j++) o[j] = s[j];
#endif
    constexpr int kUnroll = GLB_DENSE_UNROLL;
#pragma unroll kUnroll
    for (int i = 0; i < 12; i++) {
#ifdef GLB_DENSE_F64
        DotF d;
        d.init();
#pragma unroll
        for (int j = 0; j < 12; j++) d.mad(o0[j], o1[j], c_PLd[i * 12 + j]);
#else
        gl::Dot d;
        d.zero();
#pragma unroll
        for (int j = 0; j < 12; j++) {
            const uint4 l = c_PL[i * 12 + j];
            d.mad(o[j], l.x, l.y, l.z);
        }
#endif
        const u64 acc = d.fold();
        // rotate instead of indexing registers dynamically: after 12 steps s[] holds new[0..12)
        s[0] = s[1]; s[1] = s[2]; s[2] = s[3]; s[3] = s[4]; s[4] = s[5]; s[5] = s[6];
        s[6] = s[7]; s[7] = s[8]; s[8] = s[9]; s[9] = s[10]; s[10] = s[11]; s[11] = acc;
    }
}

// 22 partial rounds (poseidon_goldilocks.cpp:19-28): t = s0^7 + C; s0' = <(t, s1..s11), S_row>;
// s_i += t * S_col[i].
__device__ __forceinline__ void partial_rounds(u64 (&s)[12])
{
    constexpr int kUnroll = GLB_PARTIAL_UNROLL;
#pragma unroll kUnroll
    for (int r = 0; r < 22; r++) {
        const u64 t = sbox_add(s[0], c_RCP[r]);
#ifdef GLB_PARTIAL_F64
        DotF d;
        d.init();
        d.c0[0] = fma(25.0, u32_to_f64((u32)t), d.c0[0]); // S[23r] == 25 in every round (checked when the tables are built)
        d.c1[0] = fma(25.0, u32_to_f64((u32)(t >> 32)), d.c1[0]);
#pragma unroll
        for (int i = 1; i < 12; i++) d.mad(u32_to_f64((u32)s[i]), u32_to_f64((u32)(s[i] >> 32)), c_SLd[r * 12 + i]);
#else
        gl::Dot d;
        d.zero();
        gl::madw(d.c00, (u32)t, 25u); // S[23r] == 25 in every round (checked when the tables are built)
        gl::madw(d.c10, (u32)(t >> 32), 25u);
#pragma unroll
        for (int i = 1; i < 12; i++) {
            const uint4 l = c_SL[r * 12 + i];
            d.mad(s[i], l.x, l.y, l.z);
        }
#endif
#pragma unroll
        for (int i = 1; i < 12; i++) s[i] = gl::mad(t, c_SV[r * 11 + i - 1], s[i]);
        s[0] = d.fold();
    }
}

#endif // GLB_POSEIDON_SPARSE

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
// every GPU parity test checks it against the reference itself).  -DGLB_POSEIDON_SPARSE selects the sparse form.
template <bool OUT4 = false>
__device__ __forceinline__ void permute(u64 (&s)[12])
{
#pragma unroll
    for (int i = 0; i < 12; i++) s[i] = gl::add_c(s[i], c_RC0[i]);
#ifdef GLB_POSEIDON_SPARSE
#pragma unroll 1
    for (int r = 0; r < 8; r++) {
#pragma unroll
        for (int i = 0; i < 12; i++) s[i] = sbox_add(s[i], c_RCF[r * 12 + i]);
        if (r == 3) {
            dense_layer(s);
            partial_rounds(s);
        } else if (OUT4 && r == 7) {
            mds_layer<4>(s);
        } else {
            mds_layer(s);
        }
    }
#else
    // 19 stages, one loop body: stages 0..3 and 15..18 are full rounds (S-box on all 12 words, MDS layer), stages 4..14
    // are PAIRS of partial rounds (mds_pair); -DGLB_POSEIDON_SINGLE keeps one MDS layer per partial round
#ifndef GLB_POSEIDON_SINGLE
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
#else
#pragma unroll 1
    for (int r = 0; r < 30; r++) {
        if (r < 4 || r >= 26) {
            const int f = r < 4 ? r : r - 22;
#pragma unroll
            for (int i = 0; i < 12; i++) s[i] = sbox_add(s[i], c_RCF[f * 12 + i]);
        } else {
            s[0] = sbox_add(s[0], c_RCP[r - 4]);
        }
        if (OUT4 && r == 29) mds_layer<4>(s);
        else mds_layer(s);
    }
#endif
#endif
}

} // namespace poseidon
