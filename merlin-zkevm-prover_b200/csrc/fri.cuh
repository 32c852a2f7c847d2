// fri.cuh -- the callers either side of the commit path that SURVEY 8(f) ranks next:
//   * Q-polynomial commit of Starks::genProof step 4 (src/starkpil/starks.cpp:261-289): the coset un-shift +
//     (p,k) -> (k,p) regrouping between the INTT and the NTT;
//   * one FRI layer of FRIProve::prove (src/starkpil/fri/friProve.cpp:36-126): folding in F_p^3, getTransposed.
// Small streaming kernels; the transforms and the trees around them are ntt.cuh / poseidon_kernels.cuh.
#pragma once
#include "gl64.cuh"
#include <cuda_runtime.h>

namespace fri {

using gl::u32;
using gl::u64;

constexpr int kMaxQDeg = 16;
struct QScale {
    u64 s[kMaxQDeg]; // (7^-n)^p
};

// qq2[k][p*qdim + d] = qq1[(p*n + k)][d] * S^p for k < n   (starks.cpp:265-280; Goldilocks3::mul by a base-field
// scalar is component-wise, goldilocks_cubic_extension.hpp:1831-1836).  Rows >= n of qq2 are never written: the
// forward NTT that follows reads them as zeros (nvalid).
__global__ void k_q_regroup(u64 *__restrict__ qq2, const u64 *__restrict__ qq1, u64 n, u32 qdim, u32 qdeg, const QScale sc)
{
    const u64 w = (u64)qdim * qdeg, total = n * w;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (u64)gridDim.x * blockDim.x) {
        const u64 k = i / w;
        const u32 c = (u32)(i - k * w), p = c / qdim, d = c - p * qdim;
        qq2[i] = gl::mul(qq1[((u64)p * n + k) * qdim + d], sc.s[p]);
    }
}

struct F3 {
    u64 a, b, c;
};

// Goldilocks3::mul, goldilocks_cubic_extension.hpp:1817-1830 (F_p[x]/(x^3 - x - 1)); any representatives in and out.
__device__ __forceinline__ F3 mul3(const F3 &x, const F3 &y)
{
    const u64 A = gl::mul(gl::add(x.a, x.b), gl::add(y.a, y.b));
    const u64 B = gl::mul(gl::add(x.a, x.c), gl::add(y.a, y.c));
    const u64 C = gl::mul(gl::add(x.b, x.c), gl::add(y.b, y.c));
    const u64 D = gl::mul(x.a, y.a), E = gl::mul(x.b, y.b), F = gl::mul(x.c, y.c);
    const u64 G = gl::sub(D, E);
    F3 r;
    r.a = gl::sub(gl::add(C, G), F);
    r.b = gl::sub(gl::sub(gl::sub(gl::add(A, C), E), E), D);
    r.c = gl::sub(B, G);
    return r;
}

// One FRI folding step (friProve.cpp:85-104): thread g takes the nX = 2^lognx elements pol[i*pol2n + g], and
// evaluates their interpolant (INTT of size nX, coefficient i scaled by sinv_g^i, sinv_g = shift_inv * w^-g)
// at the challenge x in F_p^3.  Since sum_i c_i sinv^i x^i = sum_i c_i (sinv*x)^i this is a Horner evaluation at
// y = sinv_g * x; coefficient c_i = nX^-1 sum_j v_j w_nX^(-ij) is recomputed on the fly (nX <= 64).
// roots[i] = w_(2^log_roots)^i.
__global__ void __launch_bounds__(128) k_fri_fold(u64 *__restrict__ out, const u64 *__restrict__ pol, u32 pol_bits, u32 next_bits, u64 shift_inv,
                                                  F3 x, const u64 *__restrict__ roots, u32 log_roots, u64 nx_inv)
{
    const u64 pol2n = 1ULL << next_bits, g = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= pol2n) return;
    const u32 lognx = pol_bits - next_bits, nx = 1u << lognx;
    const u64 rmask = (1ULL << log_roots) - 1;
    // w(pol_bits)^-g
    const u64 wig = roots[(((1ULL << pol_bits) - g) << (log_roots - pol_bits)) & rmask];
    const u64 sinv = gl::mul(shift_inv, wig);
    F3 y = {gl::mul(x.a, sinv), gl::mul(x.b, sinv), gl::mul(x.c, sinv)};
    F3 acc = {0, 0, 0};
    for (int i = (int)nx - 1; i >= 0; i--) {
        u64 c0 = 0, c1 = 0, c2 = 0;
        for (u32 j = 0; j < nx; j++) {
            // w_nX^(-i*j) = roots[(nX - (i*j mod nX)) * 2^(L - lognx)]
            const u32 e = (nx - ((i * j) & (nx - 1))) & (nx - 1);
            const u64 tw = roots[(u64)e << (log_roots - lognx)];
            const u64 *v = pol + ((u64)j * pol2n + g) * 3;
            c0 = gl::mad(v[0], tw, c0);
            c1 = gl::mad(v[1], tw, c1);
            c2 = gl::mad(v[2], tw, c2);
        }
        acc = mul3(acc, y);
        acc.a = gl::add(acc.a, gl::mul(c0, nx_inv));
        acc.b = gl::add(acc.b, gl::mul(c1, nx_inv));
        acc.c = gl::add(acc.c, gl::mul(c2, nx_inv));
    }
    out[g * 3 + 0] = gl::canon(acc.a);
    out[g * 3 + 1] = gl::canon(acc.b);
    out[g * 3 + 2] = gl::canon(acc.c);
}

// FRIProve::getTransposed (friProve.cpp:252-270): aux[i*h + j] = pol[j*w + i], w = 2^transpose_bits, 3-word elements.
__global__ void k_fri_transpose(u64 *__restrict__ aux, const u64 *__restrict__ pol, u64 degree, u32 transpose_bits)
{
    const u64 w = 1ULL << transpose_bits, h = degree >> transpose_bits;
    for (u64 o = (u64)blockIdx.x * blockDim.x + threadIdx.x; o < degree; o += (u64)gridDim.x * blockDim.x) {
        const u64 i = o / h, j = o - i * h, f = j * w + i; // o = i*h + j: consecutive threads write consecutive elements
        aux[o * 3 + 0] = pol[f * 3 + 0];
        aux[o * 3 + 1] = pol[f * 3 + 1];
        aux[o * 3 + 2] = pol[f * 3 + 2];
    }
}

} // namespace fri
