// tools/poseidon_parts.cu -- the pieces of one Poseidon permutation as separate kernels, so that their SASS can be costed
// statically with tools/sass_cost.py (no GPU needed): a full round = 12 S-boxes + mds_layer, a pair of partial rounds =
// 1 S-box + mds_pair.  One permutation = 8 full rounds + 11 pairs (+ 12 additions).
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I merlin-zkevm-prover_b200/csrc -cubin -o /tmp/parts.cubin tools/poseidon_parts.cu
#include "poseidon.cuh"
typedef unsigned long long u64;
using namespace poseidon;
__global__ void __launch_bounds__(128, 4) part_full_round(u64 *o, const u64 *a, int f)
{
    u64 s[12];
    for (int i = 0; i < 12; i++) s[i] = a[threadIdx.x * 12 + i];
#pragma unroll
    for (int i = 0; i < 12; i++) s[i] = sbox_add(s[i], c_RCF[f * 12 + i]);
    mds_layer(s);
    for (int i = 0; i < 12; i++) o[threadIdx.x * 12 + i] = s[i];
}
__global__ void __launch_bounds__(128, 4) part_pair(u64 *o, const u64 *a, int q)
{
    u64 s[12];
    for (int i = 0; i < 12; i++) s[i] = a[threadIdx.x * 12 + i];
    s[0] = sbox_add(s[0], c_RCP[2 * q]);
    mds_pair(s, c_RCP[2 * q + 1]);
    for (int i = 0; i < 12; i++) o[threadIdx.x * 12 + i] = s[i];
}
__global__ void __launch_bounds__(128, 4) part_sbox12(u64 *o, const u64 *a, int f)
{
    u64 s[12];
    for (int i = 0; i < 12; i++) s[i] = a[threadIdx.x * 12 + i];
#pragma unroll
    for (int i = 0; i < 12; i++) s[i] = sbox_add(s[i], c_RCF[f * 12 + i]);
    for (int i = 0; i < 12; i++) o[threadIdx.x * 12 + i] = s[i];
}
__global__ void __launch_bounds__(128, 4) part_mds(u64 *o, const u64 *a)
{
    u64 s[12];
    for (int i = 0; i < 12; i++) s[i] = a[threadIdx.x * 12 + i];
    mds_layer(s);
    for (int i = 0; i < 12; i++) o[threadIdx.x * 12 + i] = s[i];
}
__global__ void __launch_bounds__(128, 4) part_perm(u64 *o, const u64 *a)
{
    u64 s[12];
    for (int i = 0; i < 12; i++) s[i] = a[threadIdx.x * 12 + i];
    permute<true>(s);
    for (int i = 0; i < 4; i++) o[threadIdx.x * 4 + i] = s[i];
}
