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
//   * Dot products against 64-bit constants (the dense pre-partial matrix and the 12-term row of
//     every partial round) are accumulated lazily in six carry-free 64-bit columns (gl::Dot: the
//     constant is pre-split into 22/21/21-bit limbs, 6 IMAD.WIDE per product, no ALU work) and
//     reduced once per dot product instead of once per term.
//   * The circulant MDS layers use the small entries (< 64) directly: 24 IMAD.WIDE per output on
//     32-bit halves and one 96-bit reduction.
#pragma once
#include "gl64.cuh"
#include "poseidon_constants.cuh"

namespace poseidon {

using gl::u32;
using gl::u64;

// Tables derived from GLP_RC / GLP_PM / GLP_SP by upload_constants() (poseidon_kernels.cuh):
__constant__ u64 c_RC0[12];          // pre-round constants C[0..12)
__constant__ u64 c_RCF[8 * 12];      // constants added after the S-box of full round r (row 7 = 0)
__constant__ u64 c_RCP[22];          // partial-round constants C[60..82)
__constant__ uint4 c_PL[12 * 12];    // dense layer: limbs of P[j][i] at [i*12 + j] (x,y,z; w unused)
__constant__ uint4 c_SL[22 * 12];    // partial round r: limbs of the dot row S[23r + i] at [r*12 + i]
__constant__ u64 c_SV[22 * 11];      // partial round r: rank-1 column S[23r + 12 + i], i < 11

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
template <int NOUT = 12>
__device__ __forceinline__ void mds_layer(u64 (&s)[12])
{
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
}

// Dense layer before the partial rounds: new[i] = sum_j P[j][i] old[j]   (poseidon_goldilocks.cpp:17)
__device__ __forceinline__ void dense_layer(u64 (&s)[12])
{
    u64 o[12];
#pragma unroll
    for (int j = 0; j < 12; j++) o[j] = s[j];
#pragma unroll 1
    for (int i = 0; i < 12; i++) {
        gl::Dot d;
        d.zero();
#pragma unroll
        for (int j = 0; j < 12; j++) {
            const uint4 l = c_PL[i * 12 + j];
            d.mad(o[j], l.x, l.y, l.z);
        }
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
#pragma unroll 1
    for (int r = 0; r < 22; r++) {
        const u64 t = sbox_add(s[0], c_RCP[r]);
        gl::Dot d;
        d.zero();
        gl::madw(d.c00, (u32)t, 25u); // S[23r] == 25 in every round (checked when the tables are built)
        gl::madw(d.c10, (u32)(t >> 32), 25u);
#pragma unroll
        for (int i = 1; i < 12; i++) {
            const uint4 l = c_SL[r * 12 + i];
            d.mad(s[i], l.x, l.y, l.z);
        }
#pragma unroll
        for (int i = 1; i < 12; i++) s[i] = gl::mad(t, c_SV[r * 11 + i - 1], s[i]);
        s[0] = d.fold();
    }
}

// Full permutation.  Input: any representatives; output: lazily reduced.  OUT4: only s[0..4) of the result
// are computed (all the sponge and the tree ever read: poseidon_goldilocks.cpp:242-249, 292-295) -- the last
// MDS layer then costs a third; s[4..12) are left undefined.
template <bool OUT4 = false>
__device__ __forceinline__ void permute(u64 (&s)[12])
{
#pragma unroll
    for (int i = 0; i < 12; i++) s[i] = gl::add_c(s[i], c_RC0[i]);
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
}

} // namespace poseidon
