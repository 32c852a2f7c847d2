// ntt2.cuh -- the NTT pass as a persistent, asynchronously fed kernel for sm_100a (round 2).
//
// Same transform and the same pass semantics as ntt::k_pass (ntt.cuh; reference NTT_Goldilocks::NTT_iters,
// src/goldilocks/src/ntt_goldilocks.cpp:12-164): a pass replaces one mixed-radix digit of the row index by a radix-R DFT
// (R = 16 * 2^B2) and multiplies by the inter-pass twiddle.  What changed is how a tile moves:
//
//   * tile = R rows at stride 2^log_rest x 16 adjacent columns (128-byte row segments, 32 KB for R = 256), held in shared memory;
//   * persistent CTAs (a few per SM) walk the tiles of whole STRIPS (one row-tile, all its column tiles) with a ring of NBUF tile
//     buffers: the next tile is in flight while the current one is in the butterflies --
//       first pass : cp.async (LDGSTS, 8-byte, zero-fill) from the caller's row-major matrix, whose pitch may be odd
//                    (600 / 665 columns: a tensor map needs 16-byte strides), completion on an mbarrier;
//       other passes: ONE cp.async.bulk.tensor (TMA, 3-D box: 16 columns x 1 x R rows) per tile from the scratch block, issued by one
//                    thread, completion as transaction bytes on the same mbarrier;
//   * non-last passes write their outputs back into the thread's OWN shared-memory rows and hand them to TMA stores (4-D box: the
//     R2 outputs of one first-stage frequency) -- no global address arithmetic and no second CTA barrier; the last pass stores
//     straight from registers to the caller's row-major matrix (natural order, optional scale, canonical words);
//   * the 16 inter-pass twiddles a thread needs are staged once per strip in shared memory (one coalesced-by-strip gather of R
//     words) and reused by every column tile of the strip, instead of one root-table sector per output row per column tile;
//   * the w_R^(t k1) table is read from shared memory (two row groups per warp would serialise the constant cache).
//
// Roofline: 16 B of HBM traffic per element per pass.  The pass is bound by the integer pipes (profiles/r2_01_ubench_pipes.txt:
// IMAD.WIDE 4 clk exclusive, every other integer instruction 2 clk); the memory side alone moves 4.6 TB/s with these boxes
// (profiles/r2_02_ubench_tile_tma.txt), so the asynchronous ring only has to keep the pipes fed.
#pragma once
#include "ntt.cuh"
#include "tma.cuh"

namespace ntt {

constexpr int kW2 = 16; // columns per tile

struct Pass2Params : PassParams {
    u64 nstrips;  // row tiles: 2^logn >> bits[pass]
    u32 coltiles; // ceil(ncols / 16)
};

enum { kFirst = 0, kMid = 1, kLast = 2 };

template <int B2, int MODE, int NBUF, bool CSEL>
__global__ void __launch_bounds__(16 << B2, (B2 >= 4 ? (NBUF >= 3 ? 2 : 3) : (B2 == 3 ? (NBUF >= 3 ? 4 : 6) : 8)))
k_pass2(const __grid_constant__ CUtensorMap in_map, const __grid_constant__ CUtensorMap out_map, const Pass2Params p)
{
    constexpr int R2 = 1 << B2, R = 16 * R2, NT = 16 * R2, NW = NT / 32;
    static_assert(B2 >= 1 && B2 <= 4, "tile of 32..256 rows");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    u64 *tiles = (u64 *)smem_raw;              // NBUF x [R][16]
    u64 *tw = tiles + NBUF * R * kW2;          // 2 x [R]   inter-pass twiddles of the current / next strip
    u64 *w256s = tw + 2 * R;                   // [256]     w_256^e
    uint64_t *full = (uint64_t *)(w256s + 256); // tile landed
    uint64_t *empty = full + NBUF;              // tile buffer may be overwritten

    const u32 tid = threadIdx.x, c = tid & 15, t = tid >> 4, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
#pragma unroll
        for (int b = 0; b < NBUF; b++) {
            tma::mbar_init(&full[b], MODE == kFirst ? NT : 1);
            tma::mbar_init(&empty[b], NW);
        }
        tma::fence_barrier_init();
    }
    for (u32 i = tid; i < 256; i += NT) w256s[i] = c_w256[0][i];
    __syncthreads();

    const u64 first = blockIdx.x, step = gridDim.x;
    const u64 nstrip_mine = p.nstrips > first ? (p.nstrips - first + step - 1) / step : 0;
    const u64 ntile = nstrip_mine * p.coltiles;
    const u64 restmask = (1ULL << p.log_rest) - 1;
    const u32 tw_shift = p.log_prev + p.log_roots - p.logn;
    const u64 nmask = (1ULL << p.logn) - 1;

    auto advance = [&](u64 &tau, u32 &ct) {
        if (++ct == p.coltiles) {
            ct = 0;
            tau += step;
        }
    };
    // producer side of the ring: tile q of this CTA goes to buffer q % NBUF once its previous user has released it
    auto issue_load = [&](u64 q, u64 tau, u32 ct) {
        const u32 b = (u32)(q % NBUF);
        const u64 u = q / NBUF;
        u64 *buf = tiles + (size_t)b * R * kW2;
        const u64 hi = tau >> p.log_rest, low = tau & restmask;
        if (MODE == kFirst) {
            if (u > 0) tma::mbar_wait(&empty[b], (u32)((u - 1) & 1));
            const u64 row0 = (((hi * R) << p.log_rest) + low) + ((u64)t << p.log_rest);
            const u32 rshift = p.log_rest + B2; // rows row0 + i1 * (R2 << log_rest), i1 < 16
            const u32 col = ct * kW2 + c;
            u32 iv = 0; // the first iv of them are below nvalid (the rest read as zeros: zero-padded LDE input)
            if (col < p.ncols && p.nvalid > row0) {
                const u64 nq = ((p.nvalid - row0 - 1) >> rshift) + 1;
                iv = nq < 16 ? (u32)nq : 16u;
            }
            const u64 *sp = p.src + row0 * p.spitch + col;
            const u64 sstep = p.spitch << rshift;
#pragma unroll
            for (int i1 = 0; i1 < 16; i1++) {
                const bool ok = (u32)i1 < iv;
                tma::cp_async_8(&buf[(i1 * R2 + t) * kW2 + c], ok ? sp : p.src, ok ? 8u : 0u);
                sp += sstep;
            }
            tma::cp_async_arrive_noinc(&full[b]);
        } else if (tid == 0) {
            if (u > 0) tma::mbar_wait(&empty[b], (u32)((u - 1) & 1));
            tma::mbar_expect_tx(&full[b], R * kW2 * 8);
            tma::load_3d(buf, &in_map, &full[b], (int)(ct * kW2), (int)low, (int)(hi * R));
        }
    };

    u64 lq = 0, ltau = first;
    u32 lct = 0;
    for (; lq < NBUF - 1 && lq < ntile; lq++) {
        issue_load(lq, ltau, lct);
        advance(ltau, lct);
    }
    u64 tau = first;
    u32 ct = 0, spar = 0;
    for (u64 q = 0; q < ntile; q++) {
        if (lq < ntile) {
            issue_load(lq, ltau, lct);
            advance(ltau, lct);
            lq++;
        }
        const u32 b = (u32)(q % NBUF);
        u64 *buf = tiles + (size_t)b * R * kW2;
        const u64 hi = tau >> p.log_rest, low = tau & restmask;
        // inter-pass twiddles w_n^(prev * low * k), k < R, of this strip: thread tid fetches the one for k = tid (R == NT)
        const bool newstrip = MODE != kLast && ct == 0;
        u64 twv = 0;
        if (newstrip) {
            spar ^= 1;
            const u32 ti = (u32)(low << tw_shift) * tid;
            twv = p.roots_hi ? gl::mul(p.roots[ti & ((1u << p.log_lo) - 1)], p.roots_hi[ti >> p.log_lo]) : p.roots[ti];
        }
        tma::mbar_wait(&full[b], (u32)((q / NBUF) & 1));

        // ---- step 1: radix-16 over i1 (tile row i1 * R2 + t), twiddle w_R^(t k1), back to the same rows
        u64 x[16];
#pragma unroll
        for (int i1 = 0; i1 < 16; i1++) x[i1] = buf[(i1 * R2 + t) * kW2 + c];
        dft_regs<4, CSEL>(x);
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const int k1 = bitrev(i, 4);
            u64 v = x[i];
            if (k1 != 0) v = gl::mul(v, w256s[(t * k1) << (4 - B2)]); // w_R^(t*k1) = w_256^(t*k1*256/R)
            buf[(k1 * R2 + t) * kW2 + c] = v;
        }
        if (newstrip) tw[spar * R + tid] = twv;
        __syncthreads();
        // ---- step 2: radix-R2 over t for 16 / R2 values of k1: tile rows 16 t .. 16 t + 15 = R2 * k1 + t', k1 = k0 + g
#pragma unroll
        for (int i = 0; i < 16; i++) x[i] = buf[(16 * t + i) * kW2 + c];
        Step2<B2, CSEL>::run(x);
        constexpr int G = 16 / R2;
        const u32 k0 = (16 * t) / R2;

        if (MODE != kLast) {
            // output k = k1 + 16 k2 (times its twiddle) goes back to row R2 * k1 + k2 -- one of this thread's own rows -- and from there
            // to matrix position k of the tile through the 4-D store view
            const u64 *twp = tw + spar * R;
#pragma unroll
            for (int j = 0; j < 16; j++) {
                const int g = j / R2, k2 = j % R2;
                const int i = g * R2 + bitrev(k2, B2);
                buf[(16 * t + j) * kW2 + c] = gl::mul(x[i], twp[k0 + g + 16 * k2]);
            }
            tma::fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
                constexpr int KPW = 32 / R2; // first-stage frequencies held by one warp (rows 32 warp .. 32 warp + 31)
#pragma unroll
                for (int j = 0; j < KPW; j++) {
                    const u32 k1 = warp * KPW + j;
                    tma::store_4d(&out_map, buf + (size_t)(R2 * k1) * kW2, (int)(ct * kW2), (int)low, (int)k1, (int)(hi * R2));
                }
                tma::store_commit();
                tma::store_wait_read<0>();
                tma::mbar_arrive(&empty[b]);
            }
        } else {
            __syncwarp();
            if (lane == 0) tma::mbar_arrive(&empty[b]); // every lane's tile words are in registers (consumed by step 2)
            // natural-order (or index-reversed, for the inverse) rows of the caller's matrix: row = digit-reversed(hi) + prev * k
            u64 out_base = 0;
            {
                u64 h = hi;
                u32 shift = p.log_prev;
#pragma unroll
                for (int qq = 2; qq >= 0; qq--) {
                    if (qq <= (int)p.npass - 2) {
                        shift -= p.bits[qq];
                        out_base |= (h & ((1ULL << p.bits[qq]) - 1)) << shift;
                        h >>= p.bits[qq];
                    }
                }
            }
            const u32 col = ct * kW2 + c;
            const bool colok = col < p.ncols;
            const u64 lstep = p.dpitch << p.log_prev;
            u64 *lp = p.dst + (out_base + ((u64)k0 << p.log_prev)) * p.dpitch + col;
            auto store_all = [&](auto lmode_tag) {
                constexpr int LMODE = decltype(lmode_tag)::value;
                u32 kprev = 0;
#pragma unroll
                for (int j = 0; j < 16; j++) { // ascending kc = g + 16 k2
                    const int g = j % G, k2 = j / G;
                    const int i = g * R2 + bitrev(k2, B2);
                    const u32 kc = (u32)(g + 16 * k2);
                    u64 v = x[i];
                    if (LMODE == 0) {
                        lp += lstep * (kc - kprev);
                        kprev = kc;
                        if (colok) *lp = gl::canon(v);
                    } else {
                        const u64 row = (0 - (out_base + ((u64)(k0 + kc) << p.log_prev))) & nmask;
                        v = gl::mul(v, LMODE == 1 ? p.cscale : (p.rscale ? p.rscale[row] : gl::mul(pow_tab(c_p7, (u32)row), c_inv2[p.logn])));
                        if (colok) p.dst[row * p.dpitch + col] = gl::canon(v);
                    }
                }
            };
            if (!p.inverse) store_all(IntTag<0>());
            else if (p.scale_mode == 2) store_all(IntTag<2>());
            else store_all(IntTag<1>());
        }
        advance(tau, ct);
    }
    if (MODE != kLast && lane == 0) tma::store_wait_all<0>();
}

template <int B2, int NBUF> constexpr size_t pass2_smem() { return (size_t)NBUF * (16 << B2) * kW2 * 8 + 2 * (16 << B2) * 8 + 256 * 8 + 2 * NBUF * 8; }

} // namespace ntt
