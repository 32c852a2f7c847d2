// poseidon.cuh -- Poseidon-Goldilocks permutation, sponge and Merkle kernels for sm_100a.
//
// Replaces (reference, CPU AVX2/AVX-512):
//   PoseidonGoldilocks::hash_full_result   src/goldilocks/src/poseidon_goldilocks.cpp:166-228 (seq: 5-37)
//   PoseidonGoldilocks::linear_hash        poseidon_goldilocks.cpp:229-265
//   PoseidonGoldilocks::merkletree_avx     poseidon_goldilocks.cpp:266-301
//
// Mapping: one thread owns one 12-word state in registers (24 x 32-bit); all lanes of a warp
// execute the same round, so round constants / sparse matrices are read from __constant__
// memory as warp-uniform operands.  Work per permutation: 118 S-boxes (x^7 = 4 modmuls, the
// round constant rides on the last multiply as a fused multiply-add), 7 circulant MDS layers
// done on 32-bit halves with one lazy reduction per output, one dense 12x12 layer and 22
// sparse partial rounds.  Integer-pipe bound; HBM traffic is 8 B per absorbed element.
#pragma once
#include "gl64.cuh"
#include "poseidon_constants.cuh"

namespace poseidon {

using gl::u32;
using gl::u64;

__constant__ u64 c_RC[118];
__constant__ u64 c_PM[144];
__constant__ u64 c_SP[506];

// MDS circulant (reference M[j][i] = circ[(i-j) mod 12] + 8*[i==j==0]) as immediates.
__device__ __forceinline__ constexpr u32 mds_entry(int j, int i)
{
    constexpr u32 circ[12] = {17, 20, 34, 18, 39, 13, 13, 28, 2, 16, 41, 15};
    return circ[(i - j + 12) % 12] + ((i == 0 && j == 0) ? 8u : 0u);
}

// x^7 + c   (c canonical)
__device__ __forceinline__ u64 sbox_add(u64 x, u64 c)
{
    u64 x2 = gl::sqr(x);
    u64 x3 = gl::mul(x, x2);
    u64 x4 = gl::sqr(x2);
    return gl::mad(x3, x4, c);
}

__device__ __forceinline__ u64 sbox(u64 x)
{
    u64 x2 = gl::sqr(x);
    u64 x3 = gl::mul(x, x2);
    u64 x4 = gl::sqr(x2);
    return gl::mul(x3, x4);
}

// new[i] = sum_j M[j][i] old[j]  (poseidon_goldilocks.hpp:148-162).  Entries < 64, so the sums of
// the low and of the high 32-bit halves each fit 64 bits (< 2^41); one 96-bit reduction per output.
__device__ __forceinline__ void mds_layer(u64 (&s)[12])
{
    u32 lo[12], hi[12];
#pragma unroll
    for (int j = 0; j < 12; j++) {
        lo[j] = (u32)s[j];
        hi[j] = (u32)(s[j] >> 32);
    }
#pragma unroll
    for (int i = 0; i < 12; i++) {
        u64 L = 0, H = 0;
#pragma unroll
        for (int j = 0; j < 12; j++) {
            L += (u64)lo[j] * mds_entry(j, i);
            H += (u64)hi[j] * mds_entry(j, i);
        }
        u64 t = (L >> 32) + H; // < 2^42
        s[i] = gl::reduce96((u32)L, (u32)t, (u32)(t >> 32));
    }
}

// Dense layer before the partial rounds: new[i] = sum_j PM[j*12+i] old[j].
__device__ __forceinline__ void dense_layer(u64 (&s)[12])
{
    u64 o[12];
#pragma unroll
    for (int j = 0; j < 12; j++) o[j] = s[j];
#pragma unroll 1
    for (int i = 0; i < 12; i++) {
        u64 acc = gl::mul(o[0], c_PM[i]);
#pragma unroll
        for (int j = 1; j < 12; j++) acc = gl::mad(o[j], c_PM[j * 12 + i], acc);
        // dynamic register indexing is avoided by a rotating write below
        s[0] = s[1]; s[1] = s[2]; s[2] = s[3]; s[3] = s[4]; s[4] = s[5]; s[5] = s[6];
        s[6] = s[7]; s[7] = s[8]; s[8] = s[9]; s[9] = s[10]; s[10] = s[11]; s[11] = acc;
    }
}

// Full permutation, all 12 output words.  Input: any representatives; output: lazily reduced.
__device__ __forceinline__ void permute(u64 (&s)[12])
{
#pragma unroll
    for (int i = 0; i < 12; i++) s[i] = gl::add_c(s[i], c_RC[i]);
#pragma unroll 1
    for (int r = 0; r < 3; r++) {
#pragma unroll
        for (int i = 0; i < 12; i++) s[i] = sbox_add(s[i], c_RC[(r + 1) * 12 + i]);
        mds_layer(s);
    }
#pragma unroll
    for (int i = 0; i < 12; i++) s[i] = sbox_add(s[i], c_RC[48 + i]);
    dense_layer(s);
#pragma unroll 1
    for (int r = 0; r < 22; r++) {
        const u64 *S = &c_SP[23 * r];
        u64 t = sbox_add(s[0], c_RC[60 + r]);
        u64 d = gl::mul(t, S[0]);
#pragma unroll
        for (int i = 1; i < 12; i++) d = gl::mad(s[i], S[i], d);
#pragma unroll
        for (int i = 1; i < 12; i++) s[i] = gl::mad(t, S[11 + i], s[i]);
        s[0] = d;
    }
#pragma unroll 1
    for (int r = 0; r < 3; r++) {
#pragma unroll
        for (int i = 0; i < 12; i++) s[i] = sbox_add(s[i], c_RC[82 + 12 * r + i]);
        mds_layer(s);
    }
#pragma unroll
    for (int i = 0; i < 12; i++) s[i] = sbox(s[i]);
    mds_layer(s);
}

} // namespace poseidon
