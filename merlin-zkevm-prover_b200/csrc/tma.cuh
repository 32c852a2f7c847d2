// tma.cuh -- thin PTX wrappers for the sm_100a asynchronous copy machinery used by the NTT passes:
// mbarrier (transaction-count barriers in shared memory), cp.async.bulk.tensor (TMA tiled loads / stores through a
// CUtensorMap), the async-proxy fence, and a host helper that encodes a 3-D tensor map without linking libcuda
// (the driver entry point is fetched through the runtime).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace tma {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
// make the mbarrier initialisation visible to the async proxy (TMA unit)
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// generic-proxy writes to shared memory (st.shared by threads) -> visible to the async proxy (TMA store reads)
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok)
                 : "r"(smem_u32(bar)), "r"(parity)
                 : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    while (!mbar_try_wait(bar, parity)) { }
}

// TMA tiled load of one box (coordinates innermost first) into shared memory; completion is signalled on `bar` as
// box-bytes of transaction count.  Elements of the box that lie outside the tensor are filled with zeros.
__device__ __forceinline__ void load_3d(void *smem_dst, const CUtensorMap *map, uint64_t *bar, int c0, int c1, int c2)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
// TMA tiled store of one box from shared memory (bulk async-group completion); out-of-tensor elements are not written.
__device__ __forceinline__ void store_3d(const CUtensorMap *map, const void *smem_src, int c0, int c1, int c2)
{
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4}], [%1];"
                 ::"l"(map), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
// 4-D variants (the pass output view splits the row index inside a tile into two digits, see ntt2.cuh)
__device__ __forceinline__ void load_4d(void *smem_dst, const CUtensorMap *map, uint64_t *bar, int c0, int c1, int c2, int c3)
{
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                 ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
                 : "memory");
}
__device__ __forceinline__ void store_4d(const CUtensorMap *map, const void *smem_src, int c0, int c1, int c2, int c3)
{
    asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
                 ::"l"(map), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
                 : "memory");
}
// Ampere-style 8-byte asynchronous copy global -> shared (LDGSTS); src_bytes = 0 writes zeros instead of reading.
__device__ __forceinline__ void cp_async_8(void *smem_dst, const void *gmem_src, uint32_t src_bytes)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(src_bytes) : "memory");
}
// arrive on `bar` once all cp.async issued so far by this thread have landed (the arrival is part of the barrier's expected count)
__device__ __forceinline__ void cp_async_arrive_noinc(uint64_t *bar)
{
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// wait until at most N of this thread's committed bulk groups still READ their shared-memory source
template <int N> __device__ __forceinline__ void store_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
// ... until at most N are not yet complete (writes visible in global memory)
template <int N> __device__ __forceinline__ void store_wait_all() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void prefetch_map(const CUtensorMap *map) { asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory"); }

// ---- host side ---------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_fn()
{
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return nullptr;
        fn = (EncodeTiledFn)p;
    }
    return fn;
}

// Rows of a row-major u64 matrix (`pitch` words per row, `ncols` of them addressable) seen as
//     [hi = row >> log_rest][low = row & (2^log_rest - 1)][col]:      dims (ncols, 2^log_rest, nrows >> log_rest),
// box = (box_cols, 1, box_rows): `box_rows` rows at stride 2^log_rest -- one column tile of one NTT pass tile.
// nrows_valid <= nrows limits the third dimension so that rows beyond it read as zeros (zero-padded LDE input): it must be a
// multiple of 2^log_rest.  Requires base 16-byte aligned and pitch * 8 a multiple of 16.  Returns 0 on success.
inline int make_map_3d(CUtensorMap *map, const void *base, uint64_t pitch, uint64_t ncols, uint64_t nrows_valid, uint32_t log_rest, uint32_t box_cols,
                       uint32_t box_rows)
{
    EncodeTiledFn fn = encode_fn();
    if (!fn) return -1;
    const cuuint64_t dims[3] = {ncols, 1ULL << log_rest, nrows_valid >> log_rest};
    const cuuint64_t strides[2] = {pitch * 8, (pitch * 8) << log_rest};
    const cuuint32_t box[3] = {box_cols, 1, box_rows};
    const cuuint32_t estr[3] = {1, 1, 1};
    // 64-bit words travel as pairs of 32-bit lanes for the copy engine: there is no 8-byte unsigned type in every driver, and
    // a plain copy does not care -- use UINT64 where available.
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : (int)r;
}

// Output view of a non-last pass tile: the position index k inside a tile of R = 16 * R2 rows is written k = k1 + 16 * k2, so the
// matrix row ((hi * R + k) << log_rest) + low is seen as [k2 + R2 * hi][k1][low][col]:
//     dims (ncols, 2^log_rest, 16, nrows >> (log_rest + 4)),   box = (box_cols, 1, 1, R2):
// the R2 outputs of one first-stage frequency k1, which a thread group holds in R2 consecutive shared-memory rows.
inline int make_map_4d(CUtensorMap *map, const void *base, uint64_t pitch, uint64_t ncols, uint64_t nrows, uint32_t log_rest, uint32_t box_cols, uint32_t r2)
{
    EncodeTiledFn fn = encode_fn();
    if (!fn) return -1;
    const cuuint64_t dims[4] = {ncols, 1ULL << log_rest, 16, nrows >> (log_rest + 4)};
    const cuuint64_t strides[3] = {pitch * 8, (pitch * 8) << log_rest, (pitch * 8) << (log_rest + 4)};
    const cuuint32_t box[4] = {box_cols, 1, 1, r2};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT64, 4, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : (int)r;
}

} // namespace tma
