// glb200.cu -- C ABI (include/gl_commit.h) over the sm_100a kernels.
//
// One process drives one GPU.  The library owns a stream, a grow-only scratch arena and the
// constant tables; all kernels are launched here.  There is NO CPU fallback: every compute entry
// point fails with GL_ERR_CUDA when no device is usable.
#include "../../include/gl_commit.h"
#include "ntt.cuh"
#include "ntt2.cuh"
#include "poseidon_kernels.cuh"
#include "fri.cuh"

#include <dlfcn.h>
#include <nccl.h> // types and prototypes only: libnccl.so.2 is dlopen'ed by gl_mgpu_init (single-GPU users need no NCCL)

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

namespace {

typedef unsigned long long u64;
typedef unsigned int u32;
typedef unsigned __int128 u128_t;

thread_local std::string t_err;

int fail(int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    t_err = buf;
    return code;
}

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess)                                                                    \
            return fail(e__ == cudaErrorMemoryAllocation ? GL_ERR_NOMEM : GL_ERR_CUDA, "%s: %s (%s:%d)", #call, \
                        cudaGetErrorString(e__), __FILE__, __LINE__);                              \
    } while (0)

// inside the copy / compute pipelines: remember the first failing call (the loop condition stops on it; clean-up follows the loop)
#define CUP(call)                                                                                          \
    do {                                                                                                   \
        cudaError_t e__ = (call);                                                                          \
        if (e__ != cudaSuccess && rc == GL_OK) rc = fail(GL_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

#define TRY(call)                   \
    do {                            \
        int rc__ = (call);          \
        if (rc__ != GL_OK) return rc__; \
    } while (0)

struct Ctx {
    bool ready = false;
    int device = -1;
    cudaStream_t stream = nullptr;
    cudaStream_t s_h2d = nullptr, s_d2h = nullptr; // copy streams of the pipelined host-pointer calls
    cudaStream_t s_comm = nullptr;                 // multi-GPU exchange stream (row chunks overlap the leaf hashing)
    cudaEvent_t ev_chunk[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    void *arena = nullptr;
    size_t arena_bytes = 0;
    u64 *roots = nullptr, *rscale = nullptr; // twiddle / coset-scale tables (device)
    u64 *roots_hi = nullptr;                 // second-level twiddle table of the domains above 2^27
    u32 log_roots = 0, log_rscale = 0;
    size_t scratch_limit = (size_t)32 << 30; // upper bound; pick_col_block also stays within half of the free device memory
    std::atomic<uint64_t> launches{0};
    float ms[4] = {0, 0, 0, 0}; // LDE, leaf hashing, tree levels, multi-GPU exchange
    // Resident stage (opt-in, gl_set_resident): the device copies of the last gl_extend_pol_host output and of the tree hashed
    // over it while it was being downloaded, keyed by the HOST pointer the caller got the matrix in.
    bool resident_on = false;
    const void *res_host = nullptr;
    u64 res_rows = 0, res_cols = 0;
    void *res_lde = nullptr, *res_tree = nullptr;
    size_t res_lde_bytes = 0, res_tree_bytes = 0;
    int ntt_prefetch = 0;       // k_pass: L2 prefetch distance in units of the resident CTAs (0 off); env GLB_NTT_PREFETCH
    int ntt_variant = 0;        // 0: k_pass (direct loads, two CTAs per SM alternate load and butterfly phases -- the faster one),
                                // 2 / 3: k_pass2 (ntt2.cuh: TMA / cp.async ring of 2 / 3 tile buffers); env GLB_NTT or gl_set_ntt_variant
    int sms = 148;
    void *small_dev = nullptr;  // 1 MiB device staging of the small host-pointer transforms (no cudaMalloc per call: FRIProve's fold calls
                                // INTT on nX x 3 matrices from an OpenMP loop, friProve.cpp:44-107)
    u64 *qbuf = nullptr;        // grow-only staging of the batched query calls: [indices | gathered words] (device)
    size_t qbuf_words = 0;
    // Device buffers of the host-pointer commit (extended matrix, tree, upload staging) kept between calls: cudaMalloc + cudaFree of
    // the 80 GB matrix cost 50-120 ms of every c2 call (profiles/r2_29_alloc_cost.txt).  One buffer per slot, reused when large
    // enough; given back to the driver by gl_shutdown, gl_set_resident(0), gl_set_scratch_limit and whenever a device allocation
    // of the library fails (it then retries once).
    struct PoolSlot { void *p = nullptr; size_t bytes = 0; } pool[3];
    std::mutex mu;
} g;

void pool_trim()
{
    for (auto &sl : g.pool) {
        if (sl.p) cudaFree(sl.p);
        sl.p = nullptr;
        sl.bytes = 0;
    }
}

// cudaMalloc that gives the cached buffers back to the driver and retries once when the device is full
cudaError_t dev_malloc(void **p, size_t bytes)
{
    cudaError_t e = cudaMalloc(p, bytes);
    if (e == cudaErrorMemoryAllocation && (g.pool[0].p || g.pool[1].p || g.pool[2].p)) {
        cudaGetLastError(); // clear the sticky-free allocation error
        pool_trim();
        e = cudaMalloc(p, bytes);
    }
    return e;
}

// ---- host-side field helpers (table generation only) -------------------------------------------
const u64 P = 0xFFFFFFFF00000001ULL;
u64 hmul(u64 a, u64 b) { return (u64)(((u128_t)a * b) % P); }
u64 hpow(u64 a, u64 e)
{
    u64 r = 1;
    while (e) {
        if (e & 1) r = hmul(r, a);
        a = hmul(a, a);
        e >>= 1;
    }
    return r;
}
const u64 W32 = 7277203076849721926ULL; // Goldilocks::W[32], goldilocks_base_field.cpp:44

int upload_tables()
{
    static u64 w32[4][256], p7[4][256], inv2[33], w256[2][256];
    for (int j = 0; j < 4; j++) {
        u64 gw = hpow(W32, 1ULL << (8 * j)), g7 = hpow(7, 1ULL << (8 * j));
        w32[j][0] = p7[j][0] = 1;
        for (int i = 1; i < 256; i++) {
            w32[j][i] = hmul(w32[j][i - 1], gw);
            p7[j][i] = hmul(p7[j][i - 1], g7);
        }
    }
    u64 half = hpow(2, P - 2);
    inv2[0] = 1;
    for (int i = 1; i <= 32; i++) inv2[i] = hmul(inv2[i - 1], half);
    u64 w = hpow(W32, 1ULL << 24), wi = hpow(w, P - 2);
    if (hpow(w, 16) != 4096 || hpow(w, 4) != 8) return fail(GL_ERR_ARG, "root-of-unity chain mismatch");
    w256[0][0] = w256[1][0] = 1;
    for (int i = 1; i < 256; i++) {
        w256[0][i] = hmul(w256[0][i - 1], w);
        w256[1][i] = hmul(w256[1][i - 1], wi);
    }
    CU(cudaMemcpyToSymbol(ntt::c_w32, w32, sizeof w32));
    CU(cudaMemcpyToSymbol(ntt::c_p7, p7, sizeof p7));
    CU(cudaMemcpyToSymbol(ntt::c_inv2, inv2, sizeof inv2));
    CU(cudaMemcpyToSymbol(ntt::c_w256, w256, sizeof w256));
    CU(poseidon::upload_constants());
    return GL_OK;
}

int need_ctx()
{
    if (!g.ready) return gl_init(0);
    return GL_OK;
}

// The library owns ONE stream, ONE scratch arena and one set of timing events, so the public calls are serialised:
// the reference's callers are single-threaded per prover (prover.cpp:187-258) except FRIProve::prove, which builds
// NTT_Goldilocks(nX, 1) and calls INTT from inside an OpenMP region (friProve.cpp:44-107).  Recursive because the
// *_host entry points are written on top of the *_dev ones.  The CUDA "current device" is per host thread, so every
// entry also binds the calling thread to the library's device.
std::recursive_mutex g_api_mu;
int enter_api()
{
    TRY(need_ctx());
    int cur = -1;
    if (cudaGetDevice(&cur) != cudaSuccess || cur != g.device) CU(cudaSetDevice(g.device));
    return GL_OK;
}
#define API_ENTER()                                         \
    std::lock_guard<std::recursive_mutex> api_lock_(g_api_mu); \
    TRY(enter_api())

int arena_reserve(size_t bytes)
{
    if (bytes <= g.arena_bytes) return GL_OK;
    if (g.arena) {
        CU(cudaStreamSynchronize(g.stream));
        CU(cudaFree(g.arena));
        g.arena = nullptr;
        g.arena_bytes = 0;
    }
    CU(dev_malloc(&g.arena, bytes));
    g.arena_bytes = bytes;
    return GL_OK;
}

u32 ilog2(u64 n)
{
    u32 k = 0;
    while ((1ULL << k) < n) k++;
    return k;
}

bool is_pow2(u64 n) { return n && !(n & (n - 1)); }

// ---- NTT tables / plan / launch ---------------------------------------------------------------------
// roots[i] = w_(2^L)^i for the largest transform seen so far (the reference keeps the same table per
// NTT_Goldilocks object and REBUILDS it serially on every extendPol, ntt_goldilocks.cpp:328; here it is
// built once on the device and reused); rscale[i] = 7^i / n for the most recent extend size (computeR,
// ntt_goldilocks.hpp:153-165).
// Domains above 2^27 (the reference accepts up to 2^32, ntt_goldilocks.hpp:94-97) keep only the first 2^27 powers of w_(2^L) plus a
// second-level table of the 2^(L-27) powers w^(j 2^27); the extend scale 7^i / n is then computed on the fly instead of tabulated.
constexpr u32 kLogRootsLo = 27;

int ensure_roots(u32 logn)
{
    const bool big = logn > kLogRootsLo;
    if (g.roots && (big ? g.log_roots == logn : (g.log_roots >= logn && g.log_roots <= kLogRootsLo))) return GL_OK;
    u32 L = logn < 16 ? 16 : logn;
    if (g.roots) {
        CU(cudaStreamSynchronize(g.stream));
        CU(cudaFree(g.roots));
        g.roots = nullptr;
    }
    if (g.roots_hi) {
        CU(cudaFree(g.roots_hi));
        g.roots_hi = nullptr;
    }
    const u64 count = 1ULL << (big ? kLogRootsLo : L);
    CU(dev_malloc((void **)&g.roots, count * 8));
    ntt::k_build_roots<<<148 * 4, 256, 0, g.stream>>>(g.roots, L, count, 0);
    g.launches++;
    if (big) {
        CU(dev_malloc((void **)&g.roots_hi, (size_t)8 << (L - kLogRootsLo)));
        ntt::k_build_roots<<<1, 32, 0, g.stream>>>(g.roots_hi, L, 1ULL << (L - kLogRootsLo), kLogRootsLo);
        g.launches++;
    }
    CU(cudaGetLastError());
    g.log_roots = L;
    return GL_OK;
}

int ensure_rscale(u32 logn)
{
    if (logn > kLogRootsLo) { // computed on the fly by the last pass
        if (g.rscale) {
            CU(cudaStreamSynchronize(g.stream));
            CU(cudaFree(g.rscale));
        }
        g.rscale = nullptr;
        g.log_rscale = 0;
        return GL_OK;
    }
    if (g.rscale && g.log_rscale == logn) return GL_OK;
    if (g.rscale) {
        CU(cudaStreamSynchronize(g.stream));
        CU(cudaFree(g.rscale));
        g.rscale = nullptr;
    }
    CU(dev_malloc((void **)&g.rscale, (size_t)8 << logn));
    ntt::k_build_rscale<<<148 * 4, 256, 0, g.stream>>>(g.rscale, logn);
    g.launches++;
    CU(cudaGetLastError());
    g.log_rscale = logn;
    return GL_OK;
}

template <int B2, bool LAST, int CW, bool BIG, bool PAD>
int launch_pass_pad(const ntt::PassParams &p, u64 ntiles, u32 coltiles)
{
    constexpr int R = 16 << B2;
    size_t smem = B2 > 0 ? (size_t)R * ntt::kTileCols * 8 : 0;
    static bool attr_done[16] = {false}; // per device: the attribute belongs to the device's copy of the function
    if (!attr_done[g.device & 15] && smem > 48 * 1024) {
        CU(cudaFuncSetAttribute(ntt::k_pass<B2, LAST, CW, BIG, PAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_done[g.device & 15] = true;
    }
    dim3 grid((unsigned)(ntiles / (ntt::kTileCols / CW)), coltiles), block(32, 1 << B2);
    ntt::PassParams q = p;
    q.pf_dist = (u32)(g.ntt_prefetch * g.sms * (B2 >= 4 ? 2 : (B2 == 3 ? 4 : 1))); // resident CTAs per SM as in __launch_bounds__
    ntt::k_pass<B2, LAST, CW, BIG, PAD><<<grid, block, smem, g.stream>>>(q);
    g.launches++;
    CU(cudaGetLastError());
    return GL_OK;
}
template <int B2, bool LAST, int CW, bool BIG>
int launch_pass_inst(const ntt::PassParams &p, u64 ntiles, u32 coltiles)
{
    // only the first pass of an extended forward transform reads a zero-padded source; every other pass skips the validity tests
    if (p.nvalid < (1ULL << p.logn)) return launch_pass_pad<B2, LAST, CW, BIG, true>(p, ntiles, coltiles);
    return launch_pass_pad<B2, LAST, CW, BIG, false>(p, ntiles, coltiles);
}
template <int B2, bool LAST, int CW>
int launch_pass_cw(const ntt::PassParams &p, u64 ntiles, u32 coltiles)
{
    // domains above 2^27: only the full-width tiles of 7- and 8-bit passes exist there (2^28 = 7+7+7+7 ... 2^32 = 8+8+8+8)
    if (p.roots_hi || (p.scale_mode == 2 && !p.rscale)) {
        if constexpr (B2 >= 3 && CW == ntt::kTileCols) return launch_pass_inst<B2, LAST, CW, true>(p, ntiles, coltiles);
        else return fail(GL_ERR_ARG, "NTT above 2^27: unexpected pass shape");
    }
    return launch_pass_inst<B2, LAST, CW, false>(p, ntiles, coltiles);
}

// Full 32-column tiles in one launch; a ragged remainder of <= 24 columns in one or two more launches that pack two (four)
// row-tiles into each warp: 16 columns (CW = 16), then what is left as <= 8 (CW = 8) or <= 16 -- a remainder of 17..24 columns used
// to cost a whole 32-column tile (c2 on 4 GPUs: 152 = 4 x 32 + 24 columns per rank ran as long as 160).  Only for the wide passes of
// large transforms, where it matters.
template <int B2, bool LAST>
int launch_pass_t(const ntt::PassParams &p, u64 ntiles, u32 coltiles)
{
    u32 rag = p.ncols % ntt::kTileCols;
    const bool big = p.roots_hi || (p.scale_mode == 2 && !p.rscale); // domains above 2^27: full-width tiles only (launch_pass_cw)
    if (B2 >= 3 && rag != 0 && rag <= 24 && ntiles >= 4 && !big) {
        u32 done = p.ncols - rag;
        if (done) {
            ntt::PassParams q = p;
            q.ncols = done;
            TRY((launch_pass_cw<B2, LAST, ntt::kTileCols>(q, ntiles, coltiles - 1)));
        }
        if (rag > 16) {
            ntt::PassParams q = p;
            q.src = p.src + done;
            q.dst = p.dst + done;
            q.ncols = 16;
            TRY((launch_pass_cw<(B2 >= 3 ? B2 : 3), LAST, 16>(q, ntiles, 1)));
            done += 16;
            rag -= 16;
        }
        ntt::PassParams q = p;
        q.src = p.src + done;
        q.dst = p.dst + done;
        q.ncols = rag;
        if (rag <= 8) return launch_pass_cw<(B2 >= 3 ? B2 : 3), LAST, 8>(q, ntiles, 1);
        return launch_pass_cw<(B2 >= 3 ? B2 : 3), LAST, 16>(q, ntiles, 1);
    }
    return launch_pass_cw<B2, LAST, ntt::kTileCols>(p, ntiles, coltiles);
}

int launch_pass(const ntt::PassParams &p)
{
    u32 b = p.bits[p.pass];
    u64 ntiles = (1ULL << p.logn) >> b;
    u32 coltiles = (p.ncols + ntt::kTileCols - 1) / ntt::kTileCols;
    if (ntiles > 0x7FFFFFFFULL || coltiles > 65535) return fail(GL_ERR_ARG, "grid too large");
    const bool last = p.pass + 1 == p.npass;
    switch ((b - 4) * 2 + (last ? 1 : 0)) {
    case 0: return launch_pass_t<0, false>(p, ntiles, coltiles);
    case 1: return launch_pass_t<0, true>(p, ntiles, coltiles);
    case 2: return launch_pass_t<1, false>(p, ntiles, coltiles);
    case 3: return launch_pass_t<1, true>(p, ntiles, coltiles);
    case 4: return launch_pass_t<2, false>(p, ntiles, coltiles);
    case 5: return launch_pass_t<2, true>(p, ntiles, coltiles);
    case 6: return launch_pass_t<3, false>(p, ntiles, coltiles);
    case 7: return launch_pass_t<3, true>(p, ntiles, coltiles);
    case 8: return launch_pass_t<4, false>(p, ntiles, coltiles);
    case 9: return launch_pass_t<4, true>(p, ntiles, coltiles);
    }
    return fail(GL_ERR_ARG, "bad pass width %u", b);
}

// ---- k_pass2 (ntt2.cuh): persistent CTAs, tiles fed by cp.async (first pass) / TMA (other passes) ----------------------------
template <int B2, int MODE, int NBUF>
int launch_pass2_inst(const ntt::Pass2Params &p, const CUtensorMap &in_map, const CUtensorMap &out_map)
{
    auto kern = ntt::k_pass2<B2, MODE, NBUF, GLB_NTT_CSEL != 0>;
    const size_t smem = ntt::pass2_smem<B2, NBUF>();
    static int occ[16] = {0}; // per device: resident CTAs per SM
    int &o = occ[g.device & 15];
    if (o == 0) {
        CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, kern, 16 << B2, smem));
        if (o < 1) return fail(GL_ERR_CUDA, "k_pass2<%d,%d,%d> does not fit an SM", B2, MODE, NBUF);
    }
    u64 grid = (u64)g.sms * o;
    if (grid > p.nstrips) grid = p.nstrips;
    kern<<<(unsigned)grid, 16 << B2, smem, g.stream>>>(in_map, out_map, p);
    g.launches++;
    CU(cudaGetLastError());
    return GL_OK;
}
template <int B2, int MODE>
int launch_pass2_b(const ntt::Pass2Params &p, const CUtensorMap &in_map, const CUtensorMap &out_map, int nbuf)
{
    if (nbuf == 3) return launch_pass2_inst<B2, MODE, 3>(p, in_map, out_map);
    return launch_pass2_inst<B2, MODE, 2>(p, in_map, out_map);
}
template <int MODE>
int launch_pass2_m(const ntt::Pass2Params &p, const CUtensorMap &in_map, const CUtensorMap &out_map, int nbuf)
{
    switch (p.bits[p.pass]) {
    case 5: return launch_pass2_b<1, MODE>(p, in_map, out_map, nbuf);
    case 6: return launch_pass2_b<2, MODE>(p, in_map, out_map, nbuf);
    case 7: return launch_pass2_b<3, MODE>(p, in_map, out_map, nbuf);
    case 8: return launch_pass2_b<4, MODE>(p, in_map, out_map, nbuf);
    }
    return fail(GL_ERR_ARG, "k_pass2: bad pass width %u", p.bits[p.pass]);
}

// The new passes need: >= 2 passes, every pass 5..8 bits wide, and a scratch block a tensor map can describe (16-byte aligned base,
// even pitch, padded to whole 16-column tiles).
bool pass2_usable(u32 logn, u32 m, const u32 *bits, const u64 *tmp, u64 tpitch, u32 ncols)
{
    if (g.ntt_variant == 0 || m < 2 || !tmp) return false;
    for (u32 i = 0; i < m; i++)
        if (bits[i] < 5 || bits[i] > 8) return false;
    if (((uintptr_t)tmp & 15) || (tpitch & 1) || ((ncols + 15) & ~15u) > tpitch) return false;
    (void)logn;
    return true;
}

int launch_pass2(const ntt::PassParams &pp)
{
    ntt::Pass2Params p;
    memcpy(static_cast<ntt::PassParams *>(&p), &pp, sizeof(ntt::PassParams));
    const u32 b = p.bits[p.pass];
    p.nstrips = (1ULL << p.logn) >> b;
    p.coltiles = (p.ncols + ntt::kW2 - 1) / ntt::kW2;
    const bool first = p.pass == 0, last = p.pass + 1 == p.npass;
    const u64 n = 1ULL << p.logn, wpad = (u64)p.coltiles * ntt::kW2;
    CUtensorMap in_map, out_map;
    memset(&in_map, 0, sizeof in_map);
    memset(&out_map, 0, sizeof out_map);
    if (!first && tma::make_map_3d(&in_map, p.src, p.spitch, wpad, n, p.log_rest, ntt::kW2, 1u << b)) return fail(GL_ERR_CUDA, "cuTensorMapEncodeTiled failed (pass input)");
    if (!last && tma::make_map_4d(&out_map, p.dst, p.dpitch, wpad, n, p.log_rest, ntt::kW2, 1u << (b - 4))) return fail(GL_ERR_CUDA, "cuTensorMapEncodeTiled failed (pass output)");
    const int nbuf = g.ntt_variant == 3 ? 3 : 2;
    if (first) return launch_pass2_m<ntt::kFirst>(p, in_map, out_map, nbuf);
    if (last) return launch_pass2_m<ntt::kLast>(p, in_map, out_map, nbuf);
    return launch_pass2_m<ntt::kMid>(p, in_map, out_map, nbuf);
}

// One transform of `ncols` columns (ncols small enough for tmp).  src/dst may alias.
// scale_mode: 0 none, 1: 1/n, 2: 7^i/n.  nvalid: rows >= nvalid of src are zeros.
int run_ntt(u64 *dst, u64 dpitch, const u64 *src, u64 spitch, u64 *tmp, u64 tpitch, u32 logn, u32 ncols,
            bool inverse, int scale_mode, u64 nvalid)
{
    ntt::PassParams p;
    memset(&p, 0, sizeof p);
    p.ncols = ncols;
    p.logn = logn;
    p.inverse = inverse ? 1 : 0;
    if (logn <= 3) {
        p.src = src; p.dst = dst; p.spitch = spitch; p.dpitch = dpitch;
        p.scale_mode = scale_mode; p.nvalid = nvalid; p.npass = 1;
        ntt::k_tiny<<<(ncols + 127) / 128, 128, 0, g.stream>>>(p);
        g.launches++;
        CU(cudaGetLastError());
        return GL_OK;
    }
    u32 m = (logn + 7) / 8;
    p.npass = m;
    for (u32 i = 0; i < m; i++) p.bits[i] = logn / m + (i < logn % m ? 1 : 0);
    if (m > 1) TRY(ensure_roots(logn));
    if (scale_mode == 2) TRY(ensure_rscale(logn));
    p.roots = g.roots;
    p.log_roots = g.log_roots;
    p.rscale = g.rscale;
    p.roots_hi = logn > kLogRootsLo ? g.roots_hi : nullptr; // a one-pass transform after a big one still finds the two-level table in g: it reads no roots
    p.log_lo = kLogRootsLo;
    p.cscale = scale_mode == 1 ? hpow((P + 1) / 2, logn) : 1; // 2^-logn (1: unscaled)
    const bool v2 = pass2_usable(logn, m, p.bits, tmp, tpitch, ncols);
    u32 prev = 0;
    for (u32 i = 0; i < m; i++) {
        p.pass = i;
        p.log_prev = prev;
        p.log_rest = logn - prev - p.bits[i];
        bool first = (i == 0), last = (i + 1 == m);
        p.src = first ? src : tmp;
        p.spitch = first ? spitch : tpitch;
        p.dst = last ? dst : tmp;
        p.dpitch = last ? dpitch : tpitch;
        p.nvalid = first ? nvalid : (1ULL << logn);
        p.scale_mode = last ? scale_mode : 0;
        TRY(v2 ? launch_pass2(p) : launch_pass(p));
        prev += p.bits[i];
    }
    return GL_OK;
}

// Columns per scratch block: as many as the scratch limit and half of the free device memory allow (wide blocks amortise the
// launch tails), a multiple of 32, at least 32.  The free-memory query is a driver call that takes milliseconds when several
// processes share a node (it showed up as 20 ms of idle stream per commit at 8 GPUs), so the answer is remembered per
// (ncols, bytes per column, limit) and reused as long as the arena it was sized for is still there.
u32 pick_col_block(u64 ncols, size_t bytes_per_col)
{
    static u64 k_ncols = 0, k_cb = 0;
    static size_t k_bpc = 0, k_limit = 0;
    if (k_cb && k_ncols == ncols && k_bpc == bytes_per_col && k_limit == g.scratch_limit && g.arena_bytes >= bytes_per_col * k_cb) return (u32)k_cb;
    size_t budget = g.scratch_limit, free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) {
        free_b += g.arena_bytes; // the arena is ours to reuse
        for (const auto &sl : g.pool) free_b += sl.bytes; // and so are the cached host-path buffers (dev_malloc gives them back on demand)
        if (free_b / 2 < budget) budget = free_b / 2;
    }
    u64 cb = budget / bytes_per_col;
    cb = (cb / 32) * 32;
    if (cb < 32) cb = 32;
    if (cb > ncols) cb = ncols;
    k_ncols = ncols; k_bpc = bytes_per_col; k_limit = g.scratch_limit; k_cb = cb;
    return (u32)cb;
}
// scratch pitch of a block of cb columns: whole 16-column tiles (tensor maps need an even pitch; k_pass2 reads / writes whole tiles)
u64 pad16(u64 cb) { return g.ntt_variant == 0 ? cb : ((cb + 15) & ~(u64)15); }

int ntt_dev(u64 *dst, const u64 *src, u64 size, u64 ncols, int inverse, int extend)
{
    if (size == 0 || ncols == 0) return GL_OK; // ntt_goldilocks.cpp:168-171
    if (!is_pow2(size)) return fail(GL_ERR_ARG, "NTT size %llu is not a power of two", size);
    if (size > (1ULL << 32)) return fail(GL_ERR_RANGE, "Domain size too big for the curve");
    if (!dst) dst = (u64 *)src;
    u32 logn = ilog2(size);
    int mode = inverse ? (extend ? 2 : 1) : 0;
    u32 cb = pick_col_block(ncols, size * 8);
    const u64 cp = pad16(cb);
    if (logn > 8) TRY(arena_reserve((size_t)size * cp * 8));
    for (u64 c0 = 0; c0 < ncols; c0 += cb) {
        u32 w = (u32)((ncols - c0 < cb) ? ncols - c0 : cb);
        TRY(run_ntt(dst + c0, ncols, src + c0, ncols, (u64 *)g.arena, cp, logn, w, inverse != 0, mode, size));
    }
    return GL_OK;
}

int extend_pol_dev(u64 *out, const u64 *in, u64 n_ext, u64 n, u64 ncols)
{
    if (n == 0 || ncols == 0) return GL_OK;
    if (!is_pow2(n) || !is_pow2(n_ext) || n_ext < n) return fail(GL_ERR_ARG, "extendPol: bad sizes %llu -> %llu", n, n_ext);
    if (n_ext > (1ULL << 32)) return fail(GL_ERR_RANGE, "Domain size too big for the curve");
    u32 logn = ilog2(n), loge = ilog2(n_ext);
    u32 cb = pick_col_block(ncols, (n + n_ext) * 8);
    const u64 cp = pad16(cb);
    TRY(arena_reserve((size_t)(n + n_ext) * cp * 8));
    u64 *coef = (u64 *)g.arena;          // n x cp      INTT result, scaled by 7^i / n
    u64 *tmp = coef + (size_t)n * cp;    // n_ext x cp  pass scratch
    for (u64 c0 = 0; c0 < ncols; c0 += cb) {
        u32 w = (u32)((ncols - c0 < cb) ? ncols - c0 : cb);
        TRY(run_ntt(coef, cp, in + c0, ncols, tmp, cp, logn, w, true, 2, n));
        TRY(run_ntt(out + c0, ncols, coef, cp, tmp, cp, loge, w, false, 0, n));
    }
    return GL_OK;
}

int merkletree_dev(u64 *tree, const u64 *src, u64 ncols, u64 nrows, bool timed)
{
    if (nrows == 0) return GL_OK; // poseidon_goldilocks.cpp:268-271
    if (timed) CU(cudaEventRecord(g.ev[1], g.stream));
    poseidon::launch_linear_hash_rows(tree, src, nrows, ncols, g.stream);
    g.launches++;
    CU(cudaGetLastError());
    if (timed) CU(cudaEventRecord(g.ev[2], g.stream));
    g.launches += poseidon::launch_tree_levels(tree, nrows, g.stream);
    CU(cudaGetLastError());
    if (timed) CU(cudaEventRecord(g.ev[3], g.stream));
    return GL_OK;
}

int collect_timings(bool lde, bool tree)
{
    CU(cudaStreamSynchronize(g.stream));
    if (lde) CU(cudaEventElapsedTime(&g.ms[0], g.ev[0], g.ev[1])); else g.ms[0] = 0;
    if (tree) {
        CU(cudaEventElapsedTime(&g.ms[1], g.ev[1], g.ev[2]));
        CU(cudaEventElapsedTime(&g.ms[2], g.ev[2], g.ev[3]));
    } else {
        g.ms[1] = g.ms[2] = 0;
    }
    g.ms[3] = 0;
    return GL_OK;
}

__global__ void k_gather_proof(u64 *proof, const u64 *src, const u64 *tree, u64 height, u64 width, u64 idx)
{
    u64 tid = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (tid < width) {
        proof[tid] = gl::canon(src[idx * width + tid]);
        return;
    }
    u64 k = tid - width; // sibling word k: level k/4, word k%4   (merkleTreeGL.cpp:24-35)
    u64 level = k >> 2, off = 0, n = height;
    for (u64 l = 0; l < level; l++) {
        off += n * 4;
        n >>= 1;
    }
    if (n <= 1) return;
    u64 sib = (idx >> level) ^ 1;
    proof[tid] = gl::canon(tree[off + sib * 4 + (k & 3)]);
}

// n proofs in one launch: blockIdx.y = query, proof q at proofs + q * words (MerkleTreeGL::getGroupProof for every query of
// one tree, friProve.cpp:219-250).  An index >= height yields an all-zero proof (the host side rejects it beforehand).
__global__ void k_gather_proofs(u64 *proofs, const u64 *src, const u64 *tree, u64 height, u64 width, u64 words, const u64 *idx_list)
{
    const u64 tid = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= words) return;
    const u64 idx = idx_list[blockIdx.y];
    u64 *proof = proofs + (u64)blockIdx.y * words;
    if (tid < width) {
        proof[tid] = gl::canon(src[idx * width + tid]);
        return;
    }
    const u64 k = tid - width, level = k >> 2;
    u64 off = 0, n = height;
    for (u64 l = 0; l < level; l++) {
        off += n * 4;
        n >>= 1;
    }
    proof[tid] = n > 1 ? gl::canon(tree[off + ((idx >> level) ^ 1) * 4 + (k & 3)]) : 0;
}

// n rows of a row-major matrix in one launch (the row reads of Starks::evmap, starks.cpp:555-668): out[q][.] = src[idx[q]][.]
__global__ void k_gather_rows(u64 *out, const u64 *src, u64 width, const u64 *idx_list)
{
    const u64 tid = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= width) return;
    out[(u64)blockIdx.y * width + tid] = gl::canon(src[idx_list[blockIdx.y] * width + tid]);
}

// Issue-rate probes for bench.py's roofline (not on the commit path): one real IMAD.WIDE.U32 / IADD3 / DFMA per chain step, 8
// independent chains per thread, the forms whose SASS tools/ubench_pipes.cu verifies (nothing is strength-reduced away).
template <int OP>
__global__ void __launch_bounds__(512) k_issue_probe(u64 *out, int iters, u32 seed)
{
    u32 a = seed + threadIdx.x, b = seed * 3 + blockIdx.x + threadIdx.x * 5;
    u64 acc[8];
    u32 r[8];
    double d[8];
    const double m = 1.0 + 1e-9 * seed, c = 1e-3 * threadIdx.x;
#pragma unroll
    for (int i = 0; i < 8; i++) { acc[i] = ((u64)b << 32) + a + i; r[i] = a * (i + 1); d[i] = 1.0 + i; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int rep = 0; rep < 8; rep++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (OP == 0) asm volatile("{ .reg .u32 lo, hi, x, y; mov.b64 {lo, hi}, %0; mov.b64 {x, y}, %1; mad.lo.cc.u32 x, lo, hi, x; madc.hi.u32 y, lo, hi, y; mov.b64 %0, {x, y}; }" : "+l"(acc[i]) : "l"(acc[(i + 1) & 7]));
                if (OP == 1) asm volatile("{ .reg .u32 z; add.u32 z, %0, %1; add.u32 %0, z, %2; }" : "+r"(r[i]) : "r"(r[(i + 1) & 7]), "r"(r[(i + 2) & 7]));
                if (OP == 2) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(m), "d"(c));
            }
        }
    }
    u64 x = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) x += acc[i] + r[i] + (u64)__double_as_longlong(d[i]);
    out[(u64)blockIdx.x * blockDim.x + threadIdx.x] = x;
}

// Batched Goldilocks::add/sub/mul (goldilocks_base_field_scalar.hpp:16-160) -- F1 parity entry point.
__global__ void k_field_op(u64 *out, const u64 *a, const u64 *b, const u64 *c, u64 n, int op)
{
    u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    u64 r;
    switch (op) {
    case 0: r = gl::add(a[i], b[i]); break;
    case 1: r = gl::sub(a[i], b[i]); break;
    case 2: r = gl::mul(a[i], b[i]); break;
    case 3: r = gl::mad(a[i], b[i], c[i]); break;
    case 4: r = gl::add_c(a[i], gl::canon(b[i])); break;
    case 5: r = gl::sqr(a[i]); break;
    case 7: r = gl::add_cx(a[i], gl::canon_sel(b[i])); break;
    case 8: r = gl::sub_cx(a[i], gl::canon_sel(b[i])); break;
    case 9: r = gl::canon_sel(a[i]); break;
    case 10: r = gl::sqr3(a[i]); break;
    default: r = a[i];
    }
    out[i] = gl::canon(r);
}

// Synthetic benchmark input: word i = splitmix64(seed + i) mod p (counter based, so the CPU side can
// regenerate any slice).  Not part of the commit path.
__global__ void k_fill_synthetic(u64 *dst, u64 n, u64 seed, u64 first)
{
    u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        u64 z = seed + (first + i) * 0x9E3779B97F4A7C15ULL;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
        z = z ^ (z >> 31);
        dst[i] = gl::canon(z);
    }
}

// ---- multi-GPU: column-split LDE -> all-to-all to row blocks -> per-GPU subtrees -> gathered root
struct Nccl {
    void *h = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclSend) Send = nullptr;
    decltype(&ncclRecv) Recv = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    u64 *top = nullptr; // (2*world-1)*4 words: gathered subtree roots + the levels above them
    // Copy-engine exchange (default; GLB_MGPU_EXCHANGE=nccl selects grouped ncclSend/ncclRecv): every rank maps the peers' row-panel
    // buffers (CUDA IPC) and PUSHES its column block's rows into them with peer cudaMemcpyAsync over NVLink -- no SM is used, whereas
    // NCCL's send/recv kernels took ~3 % of the leaf hashing they run under (profiles/r2_33, r2_34).  A one-word ncclAllGather is the barrier.
    int ce_mode = 1;          // 0: NCCL send/recv, 1: copy engines when every rank could map every peer
    bool ce_ok = false;       // outcome of the last mapping round (the same on every rank)
    bool ce_used = false;     // the last commit went through the copy engines
    u64 *ipc_dev = nullptr;   // device staging: kMaxPanels records of kIpcWords words, then the barrier words
    struct Peer {
        u64 rec[10] = {0};    // last record seen from this rank: 64-byte cudaIpcMemHandle_t, byte offset of row_panels in its allocation, valid flag
        void *base = nullptr; // mapping of the peer's allocation in this process (cudaIpcOpenMemHandle)
    } peer[8];
} nc;
constexpr int kIpcWords = 10;

#define NC(call)                                                                                                 \
    do {                                                                                                         \
        ncclResult_t r__ = (call);                                                                               \
        if (r__ != ncclSuccess)                                                                                  \
            return fail(GL_ERR_NCCL, "%s: %s (%s:%d)", #call, nc.GetErrorString ? nc.GetErrorString(r__) : "?", __FILE__, __LINE__); \
    } while (0)

int nccl_load()
{
    if (nc.h) return GL_OK;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *n : names)
        if ((nc.h = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
    if (!nc.h) return fail(GL_ERR_NCCL, "cannot dlopen libnccl.so.2: %s", dlerror());
#define SYM(field, name)                                                              \
    nc.field = (decltype(nc.field))dlsym(nc.h, name);                                 \
    if (!nc.field) return fail(GL_ERR_NCCL, "libnccl lacks %s", name)
    SYM(GetUniqueId, "ncclGetUniqueId");
    SYM(CommInitRank, "ncclCommInitRank");
    SYM(CommDestroy, "ncclCommDestroy");
    SYM(GetErrorString, "ncclGetErrorString");
    SYM(GroupStart, "ncclGroupStart");
    SYM(GroupEnd, "ncclGroupEnd");
    SYM(Send, "ncclSend");
    SYM(Recv, "ncclRecv");
    SYM(AllGather, "ncclAllGather");
#undef SYM
    return GL_OK;
}

// First column of rank r when ncols columns are split over `world` ranks: contiguous, balanced in units of 8 columns (the rate of the
// Poseidon sponge, poseidon_goldilocks.cpp:229-265), so that after the exchange no 8-word chunk of a row straddles two panels and
// the leaf kernel reads every chunk with eight independent loads; the ncols % 8 columns left over go to the last rank.  With fewer
// than 8 columns per rank the split is balanced in single columns.
u64 col_start(u64 ncols, int world, int r)
{
    if (r >= world) return ncols;
    const u64 chunks = ncols / 8;
    if (chunks < (u64)world) return ncols * (u64)r / (u64)world;
    return 8 * (chunks * (u64)r / (u64)world);
}

// word offset of level l (0 = leaves) in a tree over h leaves (merklehash_goldilocks.hpp:25-28 layout)
u64 level_offset(u64 h, u32 l)
{
    u64 off = 0;
    for (u32 q = 0; q < l; q++) off += (h >> q) * 4;
    return off;
}

// gather the `world` subtree roots and hash the log2(world) levels above them; result root on host
int mgpu_top(const u64 *subtree_root, uint64_t root[4])
{
    NC(nc.AllGather(subtree_root, nc.top, 4, ncclUint64, nc.comm, g.stream));
    g.launches += poseidon::launch_tree_levels(nc.top, (u64)nc.world, g.stream);
    CU(cudaGetLastError());
    if (root) CU(cudaMemcpyAsync(root, nc.top + gl_tree_num_elements(nc.world) - 4, 32, cudaMemcpyDeviceToHost, g.stream));
    return GL_OK;
}

// Synthetic input restricted to a column range: dst[r*w + c] = word (r*ncols_total + col0 + c) of gl_fill_synthetic_dev.
__global__ void k_fill_synthetic_cols(u64 *dst, u64 nrows, u64 col0, u64 w, u64 ncols_total, u64 seed)
{
    u64 stride = (u64)gridDim.x * blockDim.x, n = nrows * w;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        u64 r = i / w, c = i - r * w;
        u64 z = seed + (r * ncols_total + col0 + c) * 0x9E3779B97F4A7C15ULL;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
        z = z ^ (z >> 31);
        dst[i] = gl::canon(z);
    }
}

struct DevBuf {
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    void *release() { void *q = p; p = nullptr; return q; }
    int alloc(size_t bytes)
    {
        CU(dev_malloc(&p, bytes ? bytes : 8));
        return GL_OK;
    }
    uint64_t *w() { return (uint64_t *)p; }
};

// A buffer taken from pool slot `slot` (or freshly allocated) and put back there when the call ends.
struct PoolBuf {
    void *p = nullptr;
    size_t bytes = 0;
    int slot;
    explicit PoolBuf(int s) : slot(s) {}
    ~PoolBuf() { put_back(p, bytes); }
    void put_back(void *q, size_t b)
    {
        if (!q) return;
        Ctx::PoolSlot &sl = g.pool[slot];
        if (sl.p && sl.bytes >= b) { cudaFree(q); return; } // the slot already holds a buffer at least as large
        if (sl.p) cudaFree(sl.p);
        sl.p = q;
        sl.bytes = b;
    }
    int alloc(size_t want)
    {
        if (!want) want = 8;
        Ctx::PoolSlot &sl = g.pool[slot];
        if (sl.p && sl.bytes >= want) {
            p = sl.p; bytes = sl.bytes;
            sl.p = nullptr; sl.bytes = 0;
            return GL_OK;
        }
        if (sl.p) { cudaFree(sl.p); sl.p = nullptr; sl.bytes = 0; }
        CU(dev_malloc(&p, want));
        bytes = want;
        return GL_OK;
    }
    void *release(size_t *b) { void *q = p; *b = bytes; p = nullptr; bytes = 0; return q; }
};

// Column blocks of the pipelined host-pointer calls: a narrow first block (32 columns) so that the download -- the bottleneck: 81.6 GB
// on c2 -- starts after a short upload, a second block that completes the first `cb` columns, then blocks of `cb` columns.  Wide
// blocks matter: a column-stripe copy moves w * 8 bytes per row and the copy engine sustains 55 GB/s at 128 columns but only 28 GB/s
// at 32 (16 at 16) while the upload runs (profiles/r2_31_pcie_split_streams.txt; splitting a stripe over several streams does not
// help, the copies of one direction share one engine).  A schedule that starts at 16 columns and grows by 1.75x per block was
// measured 8 % SLOWER end to end (profiles/r2_30_e2e_block_schedule_ab.txt); GLB_HOST_BLOCKS=1 selects it for tuning runs.
std::vector<u64> block_starts(u64 ncols, u64 cb)
{
    std::vector<u64> starts;
    static const int sched = getenv("GLB_HOST_BLOCKS") ? atoi(getenv("GLB_HOST_BLOCKS")) : 0;
    if (sched == 0) {
        const bool graded = cb > 32 && ncols > 2 * cb;
        for (u64 c0 = 0; c0 < ncols;) {
            starts.push_back(c0);
            c0 += (graded && c0 == 0) ? 32 : ((graded && c0 == 32) ? cb - 32 : cb);
        }
        starts.push_back(ncols);
        return starts;
    }
    const bool graded = cb > 16 && ncols >= 64;
    u64 w = graded ? (sched >= 8 ? (u64)sched : 16) : cb;
    for (u64 c0 = 0; c0 < ncols;) {
        starts.push_back(c0);
        c0 += w;
        if (graded && w < cb) {
            w = ((w * 7 / 4) / 8) * 8;
            if (w > cb) w = cb;
        }
    }
    starts.push_back(ncols);
    return starts;
}

void resident_drop()
{
    if (g.res_lde) { PoolBuf b(0); b.put_back(g.res_lde, g.res_lde_bytes); }   // back to the pool: the next host-pointer commit reuses them
    if (g.res_tree) { PoolBuf b(1); b.put_back(g.res_tree, g.res_tree_bytes); }
    g.res_lde = g.res_tree = nullptr;
    g.res_lde_bytes = g.res_tree_bytes = 0;
    g.res_host = nullptr;
    g.res_rows = g.res_cols = 0;
}

// Host-pointer stage commit / extendPol, pipelined over column blocks: while block k is transformed on the
// compute stream, block k+1 is uploaded (strided 2-D copy out of the row-major host matrix) and the finished
// block k-1 of the extended matrix is downloaded; the leaf hashing then overlaps the tail of the downloads.
// PCIe is full duplex, so the wall time approaches max(download time, upload + hashing) instead of the sum.
// Host buffers should be pinned (gl_host_alloc / gl_host_register) -- pageable memory still works but the
// copies then serialise.
int commit_host_pipelined(uint64_t *lde_out, uint64_t *tree_out, const uint64_t *in, u64 n_ext, u64 n, u64 ncols, uint64_t root[4],
                          bool do_tree, bool keep_resident = false)
{
    resident_drop(); // at most one resident stage: its 80 GB would not fit twice
    if (!is_pow2(n) || !is_pow2(n_ext) || n_ext < n) return fail(GL_ERR_ARG, "extendPol: bad sizes %llu -> %llu", n, n_ext);
    if (n_ext > (1ULL << 32)) return fail(GL_ERR_RANGE, "Domain size too big for the curve");
    const u32 logn = ilog2(n), loge = ilog2(n_ext);
    PoolBuf l(0), t(1), stage(2);
    TRY(l.alloc((size_t)n_ext * ncols * 8));
    if (do_tree) TRY(t.alloc(gl_tree_num_elements(n_ext) * 8));
    u64 *lde = (u64 *)l.p;
    if (ncols) {
        // block width: >= 4 blocks when there are enough columns, <= 128 columns, bounded by free memory
        u64 cb = ((ncols + 3) / 4 + 31) / 32 * 32;
        if (cb > 128) cb = 128;
        size_t free_b = 0, total_b = 0;
        CU(cudaMemGetInfo(&free_b, &total_b));
        free_b += g.pool[2].bytes; // ours to reuse: the cached staging block ...
        const size_t per_col = (size_t)(3 * n + n_ext) * 8; // coef + pass scratch + two staging buffers
        // ... and the arena, but only as far as this call needs it: an arena left larger by an earlier call is kept, not handed back
        auto fits = [&](u64 c) { return per_col * c + ((size_t)1 << 30) <= free_b + std::min<size_t>(g.arena_bytes, (size_t)((n + n_ext) * 8 * c)); };
        while (cb > 32 && !fits(cb)) cb -= 32;
        if (cb > ncols) cb = ncols;
        const u64 cp = pad16(cb);
        TRY(arena_reserve((size_t)(n + n_ext) * cp * 8));
        TRY(stage.alloc((size_t)2 * n * cb * 8));
        u64 *coef = (u64 *)g.arena, *tmp = coef + (size_t)n * cp;
        const std::vector<u64> starts = block_starts(ncols, cb);
        const u64 nb = starts.size() - 1;
        // GLB_HOST_TRACE=1 (tuning runs): timed events, the per-block timeline goes to stderr after the call
        static const bool trace = getenv("GLB_HOST_TRACE") && atoi(getenv("GLB_HOST_TRACE")) != 0;
        const unsigned evflags = trace ? cudaEventDefault : cudaEventDisableTiming;
        std::vector<cudaEvent_t> ev_in(nb), ev_done(nb), ev_hash, ev_dn;
        cudaEvent_t ev_t0 = nullptr;
        for (u64 k = 0; k < nb; k++) {
            CU(cudaEventCreateWithFlags(&ev_in[k], evflags));
            CU(cudaEventCreateWithFlags(&ev_done[k], evflags));
        }
        if (trace) {
            ev_hash.resize(nb);
            ev_dn.resize(nb);
            for (u64 k = 0; k < nb; k++) { CU(cudaEventCreate(&ev_hash[k])); CU(cudaEventCreate(&ev_dn[k])); }
            CU(cudaEventCreate(&ev_t0));
            CU(cudaEventRecord(ev_t0, g.s_h2d));
        }
        int rc = GL_OK;
        for (u64 k = 0; k < nb && rc == GL_OK; k++) {
            const u64 c0 = starts[k];
            const u32 w = (u32)(starts[k + 1] - c0);
            u64 *st = (u64 *)stage.p + (k & 1) * (size_t)n * cb;
            if (k >= 2) CUP(cudaStreamWaitEvent(g.s_h2d, ev_done[k - 2], 0)); // staging buffer free again
            CUP(cudaMemcpy2DAsync(st, cb * 8, in + c0, ncols * 8, (size_t)w * 8, n, cudaMemcpyHostToDevice, g.s_h2d));
            CUP(cudaEventRecord(ev_in[k], g.s_h2d));
            CUP(cudaStreamWaitEvent(g.stream, ev_in[k], 0));
            if (rc == GL_OK) rc = run_ntt(coef, cp, st, cb, tmp, cp, logn, w, true, 2, n);
            if (rc == GL_OK) rc = run_ntt(lde + c0, ncols, coef, cp, tmp, cp, loge, w, false, 0, n);
            CUP(cudaEventRecord(ev_done[k], g.stream));
            if (do_tree && ncols > 4 && rc == GL_OK) { // hash this block's columns now; the capacity words wait in the leaf slots
                poseidon::k_linear_hash_partial<<<poseidon::grid_for(n_ext), poseidon::kThreads, 0, g.stream>>>((u64 *)t.p, lde, n_ext, ncols, (u32)c0, w);
                g.launches++;
            }
            if (trace) CUP(cudaEventRecord(ev_hash[k], g.stream));
            if (lde_out) {
                CUP(cudaStreamWaitEvent(g.s_d2h, ev_done[k], 0));
                CUP(cudaMemcpy2DAsync(lde_out + c0, ncols * 8, lde + c0, ncols * 8, (size_t)w * 8, n_ext, cudaMemcpyDeviceToHost, g.s_d2h));
            }
            if (trace) CUP(cudaEventRecord(ev_dn[k], g.s_d2h));
        }
        if (trace && rc == GL_OK) {
            cudaDeviceSynchronize();
            fprintf(stderr, "[glb200 host trace] %llu blocks of %llu x %llu -> %llu (ms since the first upload started)\n", (u64)nb, (u64)n, (u64)ncols, (u64)n_ext);
            for (u64 k = 0; k < nb; k++) {
                float a = 0, b = 0, c = 0, d = 0;
                cudaEventElapsedTime(&a, ev_t0, ev_in[k]);
                cudaEventElapsedTime(&b, ev_t0, ev_done[k]);
                cudaEventElapsedTime(&c, ev_t0, ev_hash[k]);
                cudaEventElapsedTime(&d, ev_t0, ev_dn[k]);
                fprintf(stderr, "  block %2llu cols %4llu..%4llu  uploaded %8.1f  extended %8.1f  hashed %8.1f  downloaded %8.1f\n", (u64)k, (u64)starts[k],
                        (u64)starts[k + 1], a, b, c, d);
            }
            for (u64 k = 0; k < nb; k++) { cudaEventDestroy(ev_hash[k]); cudaEventDestroy(ev_dn[k]); }
            cudaEventDestroy(ev_t0);
        }
        cudaError_t e = cudaGetLastError();
        if (rc == GL_OK && e != cudaSuccess) rc = fail(GL_ERR_CUDA, "pipelined commit: %s", cudaGetErrorString(e));
        if (rc != GL_OK) {
            cudaDeviceSynchronize();
            for (u64 k = 0; k < nb; k++) { cudaEventDestroy(ev_in[k]); cudaEventDestroy(ev_done[k]); }
            return rc;
        }
        for (u64 k = 0; k < nb; k++) { cudaEventDestroy(ev_in[k]); cudaEventDestroy(ev_done[k]); } // released when complete
    }
    if (do_tree) {
        if (ncols > 4) { // leaves were hashed block by block above
            g.launches += poseidon::launch_tree_levels((u64 *)t.p, n_ext, g.stream);
            CU(cudaGetLastError());
        } else {
            TRY(merkletree_dev((u64 *)t.p, lde, ncols, n_ext, false));
        }
        const u64 tw = gl_tree_num_elements(n_ext);
        if (tree_out) CU(cudaMemcpyAsync(tree_out, t.p, tw * 8, cudaMemcpyDeviceToHost, g.stream));
        if (root) CU(cudaMemcpyAsync(root, (u64 *)t.p + tw - 4, 32, cudaMemcpyDeviceToHost, g.stream));
    }
    CU(cudaStreamSynchronize(g.stream));
    CU(cudaStreamSynchronize(g.s_d2h));
    CU(cudaStreamSynchronize(g.s_h2d));
    if (keep_resident && lde_out && do_tree) {
        g.res_host = lde_out;
        g.res_rows = n_ext;
        g.res_cols = ncols;
        g.res_lde = l.release(&g.res_lde_bytes);
        g.res_tree = t.release(&g.res_tree_bytes);
    }
    return GL_OK;
}

// Is `src` (nrows x ncols, host) the matrix whose device copy is resident?  The pointer and the shape must match and a sample
// of its words must still equal the device copy (the host side may have written into the buffer since: then the copy is stale).
bool resident_matches(const uint64_t *src, u64 ncols, u64 nrows)
{
    if (!g.resident_on || !g.res_lde || src != g.res_host || nrows != g.res_rows || ncols != g.res_cols) return false;
    const u64 words = nrows * ncols;
    if (words == 0) return true;
    uint64_t probe[64];
    for (int i = 0; i < 64; i++) {
        const u64 at = i == 63 ? words - 1 : (words / 64) * i + (i * 7919ULL) % (words / 64 ? words / 64 : 1);
        if (cudaMemcpyAsync(&probe[i], (const u64 *)g.res_lde + at, 8, cudaMemcpyDeviceToHost, g.stream) != cudaSuccess) return false;
    }
    if (cudaStreamSynchronize(g.stream) != cudaSuccess) return false;
    for (int i = 0; i < 64; i++) {
        const u64 at = i == 63 ? words - 1 : (words / 64) * i + (i * 7919ULL) % (words / 64 ? words / 64 : 1);
        u64 h = src[at];
        if (h >= P) h -= P; // the device copy is canonical
        if (h != probe[i]) return false;
    }
    return true;
}

} // namespace

// =================================================================================================
extern "C" {

const char *gl_last_error(void) { return t_err.c_str(); }

int gl_device_count(int *count)
{
    CU(cudaGetDeviceCount(count));
    return GL_OK;
}

int gl_init(int device)
{
    std::lock_guard<std::mutex> lk(g.mu);
    if (g.ready && g.device == device) return GL_OK;
    if (g.ready) return fail(GL_ERR_ARG, "already bound to device %d", g.device);
    int n = 0;
    CU(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n) return fail(GL_ERR_CUDA, "no CUDA device %d (found %d)", device, n);
    CU(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail(GL_ERR_CUDA, "device %s is sm_%d%d; this library is built for sm_100a only", prop.name, prop.major, prop.minor);
    g.sms = prop.multiProcessorCount;
    if (const char *v = getenv("GLB_NTT")) g.ntt_variant = atoi(v);
    if (const char *v = getenv("GLB_NTT_PREFETCH")) g.ntt_prefetch = atoi(v);
    if (const char *v = getenv("GLB_SCRATCH_MIB")) g.scratch_limit = (size_t)atoll(v) << 20; // tuning runs (tools/): same as gl_set_scratch_limit
    CU(cudaStreamCreateWithFlags(&g.stream, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&g.s_h2d, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&g.s_d2h, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&g.s_comm, cudaStreamNonBlocking));
    for (auto &e : g.ev_chunk) CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto &e : g.ev) CU(cudaEventCreate(&e));
    TRY(upload_tables());
    g.device = device;
    g.ready = true;
    return GL_OK;
}

void gl_shutdown(void)
{
    std::lock_guard<std::recursive_mutex> api_lock_(g_api_mu);
    std::lock_guard<std::mutex> lk(g.mu);
    if (!g.ready) return;
    cudaStreamSynchronize(g.stream);
    if (g.arena) cudaFree(g.arena);
    g.arena = nullptr;
    g.arena_bytes = 0;
    if (g.roots) cudaFree(g.roots);
    if (g.rscale) cudaFree(g.rscale);
    if (g.roots_hi) cudaFree(g.roots_hi);
    g.roots_hi = nullptr;
    resident_drop();
    pool_trim();
    if (g.small_dev) cudaFree(g.small_dev);
    g.small_dev = nullptr;
    if (g.qbuf) cudaFree(g.qbuf);
    g.qbuf = nullptr;
    g.qbuf_words = 0;
    g.roots = g.rscale = nullptr;
    g.log_roots = g.log_rscale = 0;
    for (auto &e : g.ev) { cudaEventDestroy(e); e = nullptr; }
    cudaStreamDestroy(g.stream);
    cudaStreamDestroy(g.s_h2d);
    cudaStreamDestroy(g.s_d2h);
    cudaStreamDestroy(g.s_comm);
    for (auto &e : g.ev_chunk) { cudaEventDestroy(e); e = nullptr; }
    g.stream = g.s_h2d = g.s_d2h = g.s_comm = nullptr;
    g.ready = false;
    g.device = -1;
}

int gl_set_resident(int on)
{
    std::lock_guard<std::recursive_mutex> api_lock_(g_api_mu);
    g.resident_on = on != 0;
    if (!g.resident_on && g.ready) {
        resident_drop();
        pool_trim(); // turning the mode off hands the memory back
    }
    return GL_OK;
}

int gl_resident_query(const uint64_t *host_src, uint64_t ncols, uint64_t nrows, uint64_t **lde_dev, uint64_t **tree_dev)
{
    API_ENTER();
    const bool ok = resident_matches(host_src, ncols, nrows);
    if (lde_dev) *lde_dev = ok ? (uint64_t *)g.res_lde : nullptr;
    if (tree_dev) *tree_dev = ok ? (uint64_t *)g.res_tree : nullptr;
    return ok ? GL_OK : fail(GL_ERR_ARG, "no resident copy of this matrix");
}

int gl_set_ntt_variant(int v)
{
    if (v != 0 && v != 2 && v != 3) return fail(GL_ERR_ARG, "NTT pass variant %d (0: k_pass, 2 / 3: k_pass2 with 2 / 3 tile buffers)", v);
    g.ntt_variant = v;
    return GL_OK;
}

int gl_set_scratch_limit(size_t bytes)
{
    std::lock_guard<std::recursive_mutex> api_lock_(g_api_mu);
    g.scratch_limit = bytes ? bytes : ((size_t)32 << 30);
    if (g.ready) pool_trim(); // also gives the buffers cached by the host-pointer commit back to the driver
    return GL_OK;
}

uint64_t gl_kernel_launches(int reset)
{
    uint64_t v = g.launches.load();
    if (reset) g.launches = 0;
    return v;
}

int gl_dev_alloc(void **dptr, size_t bytes)
{
    API_ENTER();
    CU(dev_malloc(dptr, bytes ? bytes : 8));
    return GL_OK;
}
int gl_dev_free(void *dptr)
{
    if (dptr) CU(cudaFree(dptr));
    return GL_OK;
}
int gl_host_alloc(void **hptr, size_t bytes)
{
    API_ENTER();
    CU(cudaHostAlloc(hptr, bytes ? bytes : 8, cudaHostAllocDefault));
    return GL_OK;
}
int gl_host_free(void *hptr)
{
    if (hptr) CU(cudaFreeHost(hptr));
    return GL_OK;
}
int gl_h2d(void *dptr, const void *hptr, size_t bytes)
{
    API_ENTER();
    CU(cudaMemcpyAsync(dptr, hptr, bytes, cudaMemcpyHostToDevice, g.stream));
    CU(cudaStreamSynchronize(g.stream));
    return GL_OK;
}
int gl_d2h(void *hptr, const void *dptr, size_t bytes)
{
    API_ENTER();
    CU(cudaMemcpyAsync(hptr, dptr, bytes, cudaMemcpyDeviceToHost, g.stream));
    CU(cudaStreamSynchronize(g.stream));
    return GL_OK;
}
int gl_sync(void)
{
    API_ENTER();
    CU(cudaStreamSynchronize(g.stream));
    return GL_OK;
}
void *gl_stream(void) { return (void *)g.stream; }

uint64_t gl_tree_num_elements(uint64_t height) { return height ? height * 4 + (height - 1) * 4 : 0; }

uint64_t gl_merkle_proof_size(uint64_t height)
{
    if (height > 1) return (uint64_t)ceil(log10((double)height) / log10(2.0)); // merkleTreeGL.hpp:71-78
    return 0;
}

int gl_last_timings(float ms[3])
{
    ms[0] = g.ms[0]; ms[1] = g.ms[1]; ms[2] = g.ms[2];
    return GL_OK;
}

// ---- device-resident ----------------------------------------------------------------------------
int gl_ntt_dev(uint64_t *dst, const uint64_t *src, uint64_t size, uint64_t ncols, int inverse, int extend)
{
    API_ENTER();
    CU(cudaEventRecord(g.ev[0], g.stream));
    TRY(ntt_dev((u64 *)dst, (const u64 *)src, size, ncols, inverse, extend));
    CU(cudaEventRecord(g.ev[1], g.stream));
    return collect_timings(true, false);
}

int gl_extend_pol_dev(uint64_t *out, const uint64_t *in, uint64_t n_ext, uint64_t n, uint64_t ncols)
{
    API_ENTER();
    CU(cudaEventRecord(g.ev[0], g.stream));
    TRY(extend_pol_dev((u64 *)out, (const u64 *)in, n_ext, n, ncols));
    CU(cudaEventRecord(g.ev[1], g.stream));
    return collect_timings(true, false);
}

int gl_merkletree_dev(uint64_t *tree, const uint64_t *src, uint64_t ncols, uint64_t nrows)
{
    API_ENTER();
    if (nrows == 0) return GL_OK;
    TRY(merkletree_dev((u64 *)tree, (const u64 *)src, ncols, nrows, true));
    return collect_timings(false, true);
}

int gl_commit_dev(uint64_t *lde, uint64_t *tree, const uint64_t *in, uint64_t n_ext, uint64_t n, uint64_t ncols,
                  uint64_t root[4])
{
    API_ENTER();
    if (n == 0 || n_ext == 0) return fail(GL_ERR_ARG, "commit: empty domain");
    CU(cudaEventRecord(g.ev[0], g.stream));
    TRY(extend_pol_dev((u64 *)lde, (const u64 *)in, n_ext, n, ncols));
    TRY(merkletree_dev((u64 *)tree, (const u64 *)lde, ncols, n_ext, true));
    if (root)
        CU(cudaMemcpyAsync(root, tree + gl_tree_num_elements(n_ext) - 4, 32, cudaMemcpyDeviceToHost, g.stream));
    return collect_timings(true, true);
}

int gl_group_proof_dev(uint64_t *proof_host, const uint64_t *src, const uint64_t *tree, uint64_t height,
                       uint64_t width, uint64_t idx)
{
    API_ENTER();
    if (height == 0 || idx >= height) return fail(GL_ERR_ARG, "getGroupProof: idx %llu out of range", (u64)idx);
    u64 levels = 0;
    while ((height >> levels) > 1) levels++;
    u64 words = width + 4 * levels;
    if (words == 0) return GL_OK;
    DevBuf d;
    TRY(d.alloc(words * 8));
    k_gather_proof<<<(unsigned)((words + 127) / 128), 128, 0, g.stream>>>((u64 *)d.p, (const u64 *)src, (const u64 *)tree,
                                                                          height, width, idx);
    g.launches++;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(proof_host, d.p, words * 8, cudaMemcpyDeviceToHost, g.stream));
    CU(cudaStreamSynchronize(g.stream));
    return GL_OK;
}

// grow-only device staging for the batched query calls (no cudaMalloc on the steady-state path)
static int qbuf_reserve(size_t words)
{
    if (words <= g.qbuf_words) return GL_OK;
    if (g.qbuf) {
        CU(cudaStreamSynchronize(g.stream));
        CU(cudaFree(g.qbuf));
        g.qbuf = nullptr;
        g.qbuf_words = 0;
    }
    size_t cap = words < 4096 ? 4096 : words + words / 2;
    CU(dev_malloc((void **)&g.qbuf, cap * 8));
    g.qbuf_words = cap;
    return GL_OK;
}

int gl_group_proofs_dev(uint64_t *proofs_host, const uint64_t *src, const uint64_t *tree, uint64_t height, uint64_t width,
                        const uint64_t *idx_host, uint64_t n)
{
    API_ENTER();
    if (n == 0) return GL_OK;
    if (height == 0) return fail(GL_ERR_ARG, "getGroupProof: empty tree");
    for (u64 q = 0; q < n; q++)
        if (idx_host[q] >= height) return fail(GL_ERR_ARG, "getGroupProof: idx %llu out of range (query %llu)", (u64)idx_host[q], q);
    if (n > 65535) return fail(GL_ERR_ARG, "getGroupProof: at most 65535 queries per call");
    u64 levels = 0;
    while ((height >> levels) > 1) levels++;
    const u64 words = width + 4 * levels;
    if (words == 0) return GL_OK;
    TRY(qbuf_reserve(n + n * words));
    u64 *didx = g.qbuf, *dproofs = g.qbuf + n;
    CU(cudaMemcpyAsync(didx, idx_host, n * 8, cudaMemcpyHostToDevice, g.stream));
    k_gather_proofs<<<dim3((unsigned)((words + 127) / 128), (unsigned)n), 128, 0, g.stream>>>(dproofs, (const u64 *)src, (const u64 *)tree, height, width,
                                                                                             words, didx);
    g.launches++;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(proofs_host, dproofs, n * words * 8, cudaMemcpyDeviceToHost, g.stream));
    CU(cudaStreamSynchronize(g.stream));
    return GL_OK;
}

int gl_rows_gather_dev(uint64_t *rows_host, const uint64_t *src, uint64_t width, const uint64_t *idx_host, uint64_t n)
{
    API_ENTER();
    if (n == 0 || width == 0) return GL_OK;
    if (n > 65535) return fail(GL_ERR_ARG, "rows_gather: at most 65535 rows per call");
    TRY(qbuf_reserve(n + n * width));
    u64 *didx = g.qbuf, *drows = g.qbuf + n;
    CU(cudaMemcpyAsync(didx, idx_host, n * 8, cudaMemcpyHostToDevice, g.stream));
    k_gather_rows<<<dim3((unsigned)((width + 127) / 128), (unsigned)n), 128, 0, g.stream>>>(drows, (const u64 *)src, width, didx);
    g.launches++;
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(rows_host, drows, n * width * 8, cudaMemcpyDeviceToHost, g.stream));
    CU(cudaStreamSynchronize(g.stream));
    return GL_OK;
}

int gl_copy2d(void *dst, size_t dpitch, const void *src, size_t spitch, size_t width_bytes, size_t rows, int kind)
{
    API_ENTER();
    if (width_bytes == 0 || rows == 0) return GL_OK;
    if (kind < 0 || kind > 2) return fail(GL_ERR_ARG, "copy2d: kind %d (0 H2D, 1 D2H, 2 D2D)", kind);
    const cudaMemcpyKind k = kind == 0 ? cudaMemcpyHostToDevice : (kind == 1 ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice);
    CU(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width_bytes, rows, k, g.stream));
    CU(cudaStreamSynchronize(g.stream));
    return GL_OK;
}

int gl_measure_issue_peaks(double lane_ops_per_s[3])
{
    API_ENTER();
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, g.device));
    const int blocks = prop.multiProcessorCount * 2, threads = 512, iters = 2000;
    DevBuf out;
    TRY(out.alloc((size_t)blocks * threads * 8));
    for (int op = 0; op < 3; op++) {
        float best = 1e30f;
        for (int rep = 0; rep < 3; rep++) {
            CU(cudaEventRecord(g.ev[4], g.stream));
            if (op == 0) k_issue_probe<0><<<blocks, threads, 0, g.stream>>>((u64 *)out.p, iters, 12345);
            if (op == 1) k_issue_probe<1><<<blocks, threads, 0, g.stream>>>((u64 *)out.p, iters, 12345);
            if (op == 2) k_issue_probe<2><<<blocks, threads, 0, g.stream>>>((u64 *)out.p, iters, 12345);
            g.launches++;
            CU(cudaEventRecord(g.ev[5], g.stream));
            CU(cudaStreamSynchronize(g.stream));
            float ms = 0;
            CU(cudaEventElapsedTime(&ms, g.ev[4], g.ev[5]));
            if (rep && ms < best) best = ms;
        }
        lane_ops_per_s[op] = (double)iters * 64.0 * blocks * threads / (best * 1e-3);
    }
    return GL_OK;
}

int gl_hash_full_result_dev(uint64_t *out, const uint64_t *in, uint64_t n)
{
    API_ENTER();
    poseidon::launch_hash_full_result((u64 *)out, (const u64 *)in, n, g.stream);
    if (n) g.launches++;
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(g.stream));
    return GL_OK;
}

int gl_linear_hash_rows_dev(uint64_t *out, const uint64_t *src, uint64_t ncols, uint64_t nrows)
{
    API_ENTER();
    poseidon::launch_linear_hash_rows((u64 *)out, (const u64 *)src, nrows, ncols, g.stream);
    if (nrows) g.launches++;
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(g.stream));
    return GL_OK;
}

// ---- host-pointer convenience --------------------------------------------------------------------
int gl_ntt_host(uint64_t *dst, const uint64_t *src, uint64_t size, uint64_t ncols, int inverse, int extend)
{
    if (size == 0 || ncols == 0) return GL_OK;
    API_ENTER();
    if (!dst) dst = (uint64_t *)src;
    size_t bytes = (size_t)size * ncols * 8;
    constexpr size_t kSmall = (size_t)1 << 20;
    if (bytes <= kSmall) { // persistent staging: H2D, one or two launches, D2H on the library stream, one synchronisation
        if (!g.small_dev) CU(cudaMalloc(&g.small_dev, kSmall));
        CU(cudaMemcpyAsync(g.small_dev, src, bytes, cudaMemcpyHostToDevice, g.stream));
        TRY(ntt_dev((u64 *)g.small_dev, (const u64 *)g.small_dev, size, ncols, inverse, extend));
        CU(cudaMemcpyAsync(dst, g.small_dev, bytes, cudaMemcpyDeviceToHost, g.stream));
        CU(cudaStreamSynchronize(g.stream));
        return GL_OK;
    }
    DevBuf d;
    TRY(d.alloc(bytes));
    TRY(gl_h2d(d.p, src, bytes));
    TRY(gl_ntt_dev(d.w(), d.w(), size, ncols, inverse, extend));
    return gl_d2h(dst, d.p, bytes);
}

int gl_extend_pol_host(uint64_t *out, const uint64_t *in, uint64_t n_ext, uint64_t n, uint64_t ncols)
{
    if (n == 0 || ncols == 0) return GL_OK;
    API_ENTER();
    // resident mode: hash the leaves while the extended matrix is on its way down (the download hides it) and keep both on the
    // device, so that the MerkleTreeGL::merkelize() that follows (starks.cpp:52-56) neither uploads 80 GB nor hashes again
    if (g.resident_on) return commit_host_pipelined(out, nullptr, in, n_ext, n, ncols, nullptr, true, true);
    return commit_host_pipelined(out, nullptr, in, n_ext, n, ncols, nullptr, false);
}

int gl_merkletree_host(uint64_t *tree, const uint64_t *src, uint64_t ncols, uint64_t nrows)
{
    if (nrows == 0) return GL_OK;
    API_ENTER();
    if (resident_matches(src, ncols, nrows)) // the tree was built when this matrix was extended: only the nodes travel
        return gl_d2h(tree, g.res_tree, gl_tree_num_elements(nrows) * 8);
    DevBuf s, t;
    TRY(s.alloc((size_t)nrows * ncols * 8));
    TRY(t.alloc(gl_tree_num_elements(nrows) * 8));
    if (ncols) TRY(gl_h2d(s.p, src, (size_t)nrows * ncols * 8));
    TRY(gl_merkletree_dev(t.w(), s.w(), ncols, nrows));
    return gl_d2h(tree, t.p, gl_tree_num_elements(nrows) * 8);
}

int gl_commit_host(uint64_t *lde_out, uint64_t *tree_out, const uint64_t *in, uint64_t n_ext, uint64_t n,
                   uint64_t ncols, uint64_t root[4])
{
    API_ENTER();
    if (n == 0 || n_ext == 0) return fail(GL_ERR_ARG, "commit: empty domain");
    return commit_host_pipelined(lde_out, tree_out, in, n_ext, n, ncols, root, true);
}

int gl_host_register(void *hptr, size_t bytes)
{
    API_ENTER();
    CU(cudaHostRegister(hptr, bytes, cudaHostRegisterDefault));
    return GL_OK;
}
int gl_host_unregister(void *hptr)
{
    CU(cudaHostUnregister(hptr));
    return GL_OK;
}

int gl_merkletree_batch_dev(uint64_t *tree, const uint64_t *src, uint64_t ncols, uint64_t nrows, uint64_t batch)
{
    API_ENTER();
    if (nrows == 0) return GL_OK;
    if (batch == 0) return fail(GL_ERR_ARG, "merkletree_batch: batch size 0");
    const u64 nb = ncols > 0 ? (ncols + batch - 1) / batch : 1; // poseidon_goldilocks.cpp:122-127
    DevBuf tmp;
    TRY(tmp.alloc(nrows * nb * 32));
    poseidon::k_linear_hash_batches<<<poseidon::grid_for(nrows * nb), poseidon::kThreads, 0, g.stream>>>((u64 *)tmp.p, (const u64 *)src, nrows,
                                                                                                         (u32)ncols, (u32)batch, (u32)nb);
    poseidon::launch_linear_hash_rows((u64 *)tree, (const u64 *)tmp.p, nrows, nb * 4, g.stream);
    g.launches += 2;
    CU(cudaGetLastError());
    g.launches += poseidon::launch_tree_levels((u64 *)tree, nrows, g.stream);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(g.stream));
    return GL_OK;
}

int gl_merkletree_batch_host(uint64_t *tree, const uint64_t *src, uint64_t ncols, uint64_t nrows, uint64_t batch)
{
    if (nrows == 0) return GL_OK;
    API_ENTER();
    DevBuf s, t;
    TRY(s.alloc((size_t)nrows * ncols * 8));
    TRY(t.alloc(gl_tree_num_elements(nrows) * 8));
    if (ncols) TRY(gl_h2d(s.p, src, (size_t)nrows * ncols * 8));
    TRY(gl_merkletree_batch_dev(t.w(), s.w(), ncols, nrows, batch));
    return gl_d2h(tree, t.p, gl_tree_num_elements(nrows) * 8);
}

int gl_hash_full_result_host(uint64_t *out, const uint64_t *in, uint64_t n)
{
    if (n == 0) return GL_OK;
    API_ENTER();
    DevBuf d;
    TRY(d.alloc((size_t)n * 96));
    TRY(gl_h2d(d.p, in, (size_t)n * 96));
    TRY(gl_hash_full_result_dev(d.w(), d.w(), n));
    return gl_d2h(out, d.p, (size_t)n * 96);
}

int gl_linear_hash_host(uint64_t out[4], const uint64_t *in, uint64_t size)
{
    API_ENTER();
    DevBuf s, o;
    TRY(s.alloc((size_t)size * 8));
    TRY(o.alloc(32));
    if (size) TRY(gl_h2d(s.p, in, (size_t)size * 8));
    TRY(gl_linear_hash_rows_dev(o.w(), s.w(), size, 1));
    return gl_d2h(out, o.p, 32);
}

int gl_field_op_host(uint64_t *out, const uint64_t *a, const uint64_t *b, const uint64_t *c, uint64_t n, int op)
{
    if (n == 0) return GL_OK;
    API_ENTER();
    if (op < 0 || op > 10) return fail(GL_ERR_ARG, "unknown field op %d", op);
    DevBuf d;
    TRY(d.alloc((size_t)n * 8 * 4));
    u64 *base = (u64 *)d.p;
    TRY(gl_h2d(base, a, n * 8));
    TRY(gl_h2d(base + n, b ? b : a, n * 8));
    TRY(gl_h2d(base + 2 * n, c ? c : a, n * 8));
    k_field_op<<<(unsigned)((n + 255) / 256), 256, 0, g.stream>>>(base + 3 * n, base, base + n, base + 2 * n, n, op);
    g.launches++;
    CU(cudaGetLastError());
    return gl_d2h(out, base + 3 * n, n * 8);
}

int gl_fill_synthetic_dev(uint64_t *dst, uint64_t words, uint64_t seed, uint64_t first_index)
{
    API_ENTER();
    if (words == 0) return GL_OK;
    k_fill_synthetic<<<148 * 8, 256, 0, g.stream>>>((u64 *)dst, words, seed, first_index);
    g.launches++;
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(g.stream));
    return GL_OK;
}

} // extern "C" (closed for the helper below)

namespace {
// base address of the cudaMalloc allocation that contains `ptr` (cuMemGetAddressRange through the runtime's driver entry point: the
// library links cudart only).  An IPC handle always names a whole allocation; the caller's row_panels may sit inside one.
bool alloc_base(const void *ptr, void **base)
{
    typedef int (*RangeFn)(unsigned long long *, size_t *, unsigned long long);
    static RangeFn fn = nullptr;
    if (!fn) {
        void *q = nullptr;
        cudaDriverEntryPointQueryResult r;
        if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &q, cudaEnableDefault, &r) != cudaSuccess || r != cudaDriverEntryPointSuccess) return false;
        fn = (RangeFn)q;
    }
    unsigned long long b = 0;
    size_t sz = 0;
    if (fn(&b, &sz, (unsigned long long)(uintptr_t)ptr) != 0) return false;
    *base = (void *)(uintptr_t)b;
    return true;
}

// one-word all-gather on the exchange stream: completes on a rank only after every rank's stream has reached it
int mgpu_barrier()
{
    u64 *bar = nc.ipc_dev + (size_t)poseidon::kMaxPanels * kIpcWords;
    NC(nc.AllGather(bar + nc.rank, bar, 1, ncclUint64, nc.comm, g.s_comm));
    return GL_OK;
}

// Collective (every rank calls it at the start of every multi-GPU commit): tell the peers where this rank's row_panels buffer is and
// map theirs.  Steady state = one 80-byte all-gather and one stream synchronisation (~30 us); the mappings are redone only when some
// rank's record changed.  All ranks see the same records, so all take the same branches; when any rank cannot export or map a
// buffer every rank falls back to the NCCL exchange.
int mgpu_map_peers(const u64 *row_panels, bool *use_ce)
{
    *use_ce = false;
    const int G = nc.world, me = nc.rank;
    if (G == 1 || nc.ce_mode == 0) return GL_OK;
    u64 all[poseidon::kMaxPanels * kIpcWords], mine[kIpcWords];
    memset(mine, 0, sizeof mine);
    {
        void *base = nullptr;
        cudaIpcMemHandle_t h;
        static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t size");
        if (alloc_base(row_panels, &base) && cudaIpcGetMemHandle(&h, base) == cudaSuccess) {
            memcpy(mine, &h, 64);
            mine[8] = (u64)((const char *)row_panels - (const char *)base);
            mine[9] = 1;
        } else {
            cudaGetLastError();
        }
    }
    CU(cudaMemcpyAsync(nc.ipc_dev + (size_t)me * kIpcWords, mine, sizeof mine, cudaMemcpyHostToDevice, g.s_comm));
    NC(nc.AllGather(nc.ipc_dev + (size_t)me * kIpcWords, nc.ipc_dev, kIpcWords, ncclUint64, nc.comm, g.s_comm));
    CU(cudaMemcpyAsync(all, nc.ipc_dev, (size_t)G * kIpcWords * 8, cudaMemcpyDeviceToHost, g.s_comm));
    CU(cudaStreamSynchronize(g.s_comm));
    bool changed = false, exported = true;
    for (int p = 0; p < G; p++) {
        if (memcmp(all + p * kIpcWords, nc.peer[p].rec, sizeof mine) != 0) changed = true;
        if (all[p * kIpcWords + 9] != 1) exported = false;
    }
    if (changed) {
        u64 ok = exported ? 1 : 0;
        for (int p = 0; p < G; p++) {
            const bool differs = memcmp(all + p * kIpcWords, nc.peer[p].rec, sizeof mine) != 0;
            memcpy(nc.peer[p].rec, all + p * kIpcWords, sizeof mine);
            if (p == me || !differs) continue;
            if (nc.peer[p].base) {
                cudaStreamSynchronize(g.stream); // nothing of an earlier commit still writes through the old mapping
                cudaIpcCloseMemHandle(nc.peer[p].base);
                nc.peer[p].base = nullptr;
            }
            if (!exported) continue;
            cudaIpcMemHandle_t h;
            memcpy(&h, all + p * kIpcWords, 64);
            if (cudaIpcOpenMemHandle(&nc.peer[p].base, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                cudaGetLastError();
                nc.peer[p].base = nullptr;
                ok = 0;
            }
        }
        // second round: did every rank map every peer?
        u64 *flags = nc.ipc_dev + (size_t)poseidon::kMaxPanels * kIpcWords + poseidon::kMaxPanels;
        u64 oks[poseidon::kMaxPanels];
        CU(cudaMemcpyAsync(flags + me, &ok, 8, cudaMemcpyHostToDevice, g.s_comm));
        NC(nc.AllGather(flags + me, flags, 1, ncclUint64, nc.comm, g.s_comm));
        CU(cudaMemcpyAsync(oks, flags, (size_t)G * 8, cudaMemcpyDeviceToHost, g.s_comm));
        CU(cudaStreamSynchronize(g.s_comm));
        nc.ce_ok = true;
        for (int p = 0; p < G; p++)
            if (oks[p] != 1) nc.ce_ok = false;
    }
    *use_ce = nc.ce_ok;
    return GL_OK;
}

u64 *peer_panels(int p) { return (u64 *)((char *)nc.peer[p].base + nc.peer[p].rec[8]); }

// Exchange + leaf hashing of the multi-GPU commit, in row chunks: chunk j of this rank's rows is exchanged on the
// communication stream (grouped ncclSend/ncclRecv: rows [h*rows + r0, +rc) of my column block go to rank h; I receive
// rank p's columns of my rows r0..r0+rc into panel p) while chunk j-1 is being hashed on the compute stream, so only the
// first chunk's transfer is exposed.  Both sides of every message are contiguous (the column block and the panels are
// row-major), so chunking needs no re-pack.  With row_panels_host the received chunk also starts its way to the host.
// Records g.ev[1] when the first chunk has arrived and g.ev[2] after the last leaf kernel.
int mgpu_exchange_and_hash(const u64 *lde_cols, u64 *row_panels, u64 *subtree, u64 rows, u64 w, u64 ncols_total, u64 *row_panels_host, bool ce)
{
    const int G = nc.world, me = nc.rank;
    nc.ce_used = ce;
    poseidon::RowPanels rp = {};
    rp.npanels = (u32)G;
    for (int p = 0; p <= G; p++) rp.start[p] = (u32)col_start(ncols_total, G, p);
    for (int p = 0; p < G; p++) rp.base[p] = (const u64 *)row_panels + rows * rp.start[p];
    // Row chunks of GROWING size: only the first chunk's transfer is exposed (the hashing of chunk j hides the transfer of chunk
    // j+1, and hashing a row takes ~9x as long as moving it).  Every row costs the same, so a leaf launch runs in whole WAVES of
    // resident CTAs and a chunk that is not a multiple of a wave wastes the rest of its last wave (five uneven chunks cost 30
    // waves for 27.7 waves of work at 8 GPUs: +6 % on the leaf phase).  Chunks are therefore 1, 3, 9, ... waves, and the last one
    // takes what is left once that is no more than three times its predecessor: the transfer of a chunk only has the hashing time of
    // the chunk before it (with "1, 3, 9, rest" the rest was 97.7 of 110.7 waves on 2 GPUs -- 17.8 GB to move in 32 ms).
    u64 bounds[9];
    u64 J = 0;
    bounds[0] = 0;
    {
        u64 wave = (u64)g.sms * GLB_LEAF_MINBLOCKS * poseidon::kThreads; // rows hashed by one wave of the leaf kernel
        if (const char *v = getenv("GLB_MGPU_WAVE_ROWS")) wave = strtoull(v, nullptr, 10) ? strtoull(v, nullptr, 10) : wave; // tests: chunked path on small shapes
        u64 size = wave, done = 0;
        while (J < 7 && rows > wave * 4 && rows - done > size) {
            done += size;
            bounds[++J] = done;
            size *= 3;
        }
        bounds[++J] = rows;
    }
    CU(cudaEventRecord(g.ev[5], g.stream));
    CU(cudaStreamWaitEvent(g.s_comm, g.ev[5], 0)); // the column block is complete
    // copy engines: every rank has finished its LDE -- and with it, in stream order, the hashing of its previous commit -- before
    // anyone writes into its panels
    if (ce) TRY(mgpu_barrier());
    for (u64 j = 0; j < J; j++) {
        const u64 r0 = bounds[j], rc = bounds[j + 1] - bounds[j];
        if (ce) {
            // push rows [p * rows + r0, + rc) of my column block into panel `me` of rank p (both sides contiguous), then the barrier:
            // when it completes here, every peer's pushes of this chunk -- ordered before its barrier on its stream -- have landed
            for (int q = 1; q < G && w; q++) {
                const int p = (me + q) % G; // staggered: at any moment the ranks write to different peers
                CU(cudaMemcpyAsync(peer_panels(p) + rows * rp.start[me] + r0 * w, lde_cols + ((u64)p * rows + r0) * w, rc * w * 8, cudaMemcpyDeviceToDevice, g.s_comm));
            }
            TRY(mgpu_barrier());
        } else {
            NC(nc.GroupStart());
            for (int p = 0; p < G; p++) {
                if (p == me) continue;
                const u64 wp = rp.start[p + 1] - rp.start[p];
                if (w) NC(nc.Send(lde_cols + ((u64)p * rows + r0) * w, rc * w, ncclUint64, p, nc.comm, g.s_comm));
                if (wp) NC(nc.Recv((void *)(rp.base[p] + r0 * wp), rc * wp, ncclUint64, p, nc.comm, g.s_comm));
            }
            NC(nc.GroupEnd());
        }
        if (w) CU(cudaMemcpyAsync((void *)(rp.base[me] + r0 * w), lde_cols + ((u64)me * rows + r0) * w, rc * w * 8, cudaMemcpyDeviceToDevice, g.s_comm));
        CU(cudaEventRecord(g.ev_chunk[j], g.s_comm));
        CU(cudaStreamWaitEvent(g.stream, g.ev_chunk[j], 0));
        if (j == 0) CU(cudaEventRecord(g.ev[1], g.stream));
        poseidon::RowPanels rj = rp;
        for (int p = 0; p < G; p++) rj.base[p] = rp.base[p] + r0 * (rp.start[p + 1] - rp.start[p]);
        poseidon::launch_linear_hash_panels(subtree + r0 * 4, rj, rc, g.stream);
        g.launches++;
        CU(cudaGetLastError());
        if (row_panels_host) {
            CU(cudaStreamWaitEvent(g.s_d2h, g.ev_chunk[j], 0));
            for (int p = 0; p < G; p++) {
                const u64 wp = rp.start[p + 1] - rp.start[p];
                if (wp) CU(cudaMemcpyAsync(row_panels_host + rows * rp.start[p] + r0 * wp, rp.base[p] + r0 * wp, rc * wp * 8, cudaMemcpyDeviceToHost, g.s_d2h));
            }
        }
    }
    CU(cudaEventRecord(g.ev[2], g.stream));
    return GL_OK;
}

} // namespace

extern "C" {

// ---- multi-GPU (one process per GPU) ----------------------------------------------------
int gl_mgpu_unique_id(uint8_t id[128])
{
    TRY(nccl_load());
    ncclUniqueId u;
    NC(nc.GetUniqueId(&u));
    static_assert(sizeof(u) == 128, "ncclUniqueId size");
    memcpy(id, &u, 128);
    return GL_OK;
}

int gl_mgpu_init(const uint8_t id[128], int rank, int world)
{
    API_ENTER();
    TRY(nccl_load());
    if (nc.comm) return fail(GL_ERR_ARG, "multi-GPU communicator already initialised");
    if (world < 1 || world > poseidon::kMaxPanels || (world & (world - 1)) || rank < 0 || rank >= world)
        return fail(GL_ERR_ARG, "world size %d (rank %d): need a power of two <= %d", world, rank, poseidon::kMaxPanels);
    ncclUniqueId u;
    memcpy(&u, id, 128);
    NC(nc.CommInitRank(&nc.comm, world, u, rank));
    nc.rank = rank;
    nc.world = world;
    CU(cudaMalloc(&nc.top, gl_tree_num_elements(world) * 8));
    CU(cudaMalloc(&nc.ipc_dev, ((size_t)poseidon::kMaxPanels * kIpcWords + 2 * poseidon::kMaxPanels) * 8));
    CU(cudaMemset(nc.ipc_dev, 0, ((size_t)poseidon::kMaxPanels * kIpcWords + 2 * poseidon::kMaxPanels) * 8));
    nc.ce_mode = 1;
    if (const char *v = getenv("GLB_MGPU_EXCHANGE")) nc.ce_mode = strcmp(v, "nccl") == 0 ? 0 : 1;
    nc.ce_ok = nc.ce_used = false;
    for (auto &pr : nc.peer) pr = Nccl::Peer();
    return GL_OK;
}

int gl_mgpu_exchange_mode(void) { return nc.ce_used ? 1 : 0; }

int gl_mgpu_shutdown(void)
{
    std::lock_guard<std::recursive_mutex> api_lock_(g_api_mu);
    if (nc.comm) {
        cudaStreamSynchronize(g.stream);
        nc.CommDestroy(nc.comm);
        nc.comm = nullptr;
    }
    for (auto &pr : nc.peer) {
        if (pr.base) cudaIpcCloseMemHandle(pr.base);
        pr = Nccl::Peer();
    }
    if (nc.ipc_dev) cudaFree(nc.ipc_dev);
    nc.ipc_dev = nullptr;
    nc.ce_ok = nc.ce_used = false;
    if (nc.top) cudaFree(nc.top);
    nc.top = nullptr;
    nc.rank = 0;
    nc.world = 1;
    return GL_OK;
}

uint64_t gl_mgpu_col_start(uint64_t ncols, int world, int rank) { return col_start(ncols, world, rank); }

int gl_commit_mgpu(uint64_t *lde_cols, uint64_t *row_panels, uint64_t *subtree, const uint64_t *in_cols, uint64_t n_ext,
                   uint64_t n, uint64_t ncols_total, uint64_t root[4])
{
    API_ENTER();
    if (!nc.comm) return fail(GL_ERR_ARG, "gl_commit_mgpu: call gl_mgpu_init first");
    const int G = nc.world, me = nc.rank;
    if (n == 0 || n_ext == 0 || !is_pow2(n) || !is_pow2(n_ext) || n_ext < n) return fail(GL_ERR_ARG, "commit: bad domain %llu -> %llu", (u64)n, (u64)n_ext);
    if (n_ext < (u64)G) return fail(GL_ERR_ARG, "commit: %llu extended rows cannot be split over %d ranks", (u64)n_ext, G);
    const u64 rows = n_ext / G;
    const u64 c0 = col_start(ncols_total, G, me), w = col_start(ncols_total, G, me + 1) - c0;
    bool use_ce = false;
    TRY(mgpu_map_peers((const u64 *)row_panels, &use_ce));

    CU(cudaEventRecord(g.ev[0], g.stream));
    TRY(extend_pol_dev((u64 *)lde_cols, (const u64 *)in_cols, n_ext, n, w));
    CU(cudaEventRecord(g.ev[4], g.stream));

    TRY(mgpu_exchange_and_hash((const u64 *)lde_cols, (u64 *)row_panels, (u64 *)subtree, rows, w, ncols_total, nullptr, use_ce));
    g.launches += poseidon::launch_tree_levels((u64 *)subtree, rows, g.stream);
    CU(cudaGetLastError());
    TRY(mgpu_top((const u64 *)subtree + gl_tree_num_elements(rows) - 4, root));
    CU(cudaEventRecord(g.ev[3], g.stream));
    CU(cudaStreamSynchronize(g.stream));
    CU(cudaEventElapsedTime(&g.ms[0], g.ev[0], g.ev[4]));
    CU(cudaEventElapsedTime(&g.ms[3], g.ev[4], g.ev[1]));
    CU(cudaEventElapsedTime(&g.ms[1], g.ev[1], g.ev[2]));
    CU(cudaEventElapsedTime(&g.ms[2], g.ev[2], g.ev[3]));
    return GL_OK;
}

// Host-pointer variant of gl_commit_mgpu, pipelined: the rank's column block is uploaded one column sub-block at a
// time while the previous sub-block is being extended; after the exchange the received row panels are downloaded on
// the copy stream WHILE the leaves are hashed and the subtree is built; the subtree follows.  Device buffers as for
// gl_commit_mgpu (caller-owned work space); host buffers should be pinned.
int gl_commit_mgpu_host(uint64_t *row_panels_host, uint64_t *subtree_host, const uint64_t *in_cols_host, uint64_t *lde_cols,
                        uint64_t *row_panels, uint64_t *subtree, uint64_t n_ext, uint64_t n, uint64_t ncols_total, uint64_t root[4])
{
    API_ENTER();
    if (!nc.comm) return fail(GL_ERR_ARG, "gl_commit_mgpu_host: call gl_mgpu_init first");
    const int G = nc.world, me = nc.rank;
    if (n == 0 || n_ext == 0 || !is_pow2(n) || !is_pow2(n_ext) || n_ext < n) return fail(GL_ERR_ARG, "commit: bad domain %llu -> %llu", (u64)n, (u64)n_ext);
    if (n_ext < (u64)G) return fail(GL_ERR_ARG, "commit: %llu extended rows cannot be split over %d ranks", (u64)n_ext, G);
    const u64 rows = n_ext / G;
    const u64 c0r = col_start(ncols_total, G, me), w = col_start(ncols_total, G, me + 1) - c0r;
    const u32 logn = ilog2(n), loge = ilog2(n_ext);
    bool use_ce = false;
    TRY(mgpu_map_peers((const u64 *)row_panels, &use_ce));

    CU(cudaEventRecord(g.ev[0], g.stream));
    DevBuf stage;
    if (w) {
        u64 cb = ((w + 3) / 4 + 31) / 32 * 32; // >= 4 sub-blocks when there are enough columns, <= 128 columns each
        if (cb > 128) cb = 128;
        size_t free_b = 0, total_b = 0;
        CU(cudaMemGetInfo(&free_b, &total_b));
        const size_t per_col = (size_t)(3 * n + n_ext) * 8; // coef + pass scratch + two staging buffers
        auto fits = [&](u64 c) { return per_col * c + ((size_t)1 << 30) <= free_b + std::min<size_t>(g.arena_bytes, (size_t)((n + n_ext) * 8 * c)); };
        while (cb > 32 && !fits(cb)) cb -= 32;
        if (cb > w) cb = w;
        const u64 cp = pad16(cb);
        TRY(arena_reserve((size_t)(n + n_ext) * cp * 8));
        TRY(stage.alloc((size_t)2 * n * cb * 8));
        u64 *coef = (u64 *)g.arena, *tmp = coef + (size_t)n * cp;
        const u64 nb = (w + cb - 1) / cb;
        std::vector<cudaEvent_t> ev_in(nb), ev_done(nb);
        for (u64 k = 0; k < nb; k++) {
            CU(cudaEventCreateWithFlags(&ev_in[k], cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&ev_done[k], cudaEventDisableTiming));
        }
        int rc = GL_OK;
        for (u64 k = 0; k < nb && rc == GL_OK; k++) {
            const u64 c0 = k * cb;
            const u32 ww = (u32)((w - c0 < cb) ? w - c0 : cb);
            u64 *st = (u64 *)stage.p + (k & 1) * (size_t)n * cb;
            if (k >= 2) CUP(cudaStreamWaitEvent(g.s_h2d, ev_done[k - 2], 0)); // staging buffer free again
            CUP(cudaMemcpy2DAsync(st, cb * 8, in_cols_host + c0, w * 8, (size_t)ww * 8, n, cudaMemcpyHostToDevice, g.s_h2d));
            CUP(cudaEventRecord(ev_in[k], g.s_h2d));
            CUP(cudaStreamWaitEvent(g.stream, ev_in[k], 0));
            if (rc == GL_OK) rc = run_ntt(coef, cp, st, cb, tmp, cp, logn, ww, true, 2, n);
            if (rc == GL_OK) rc = run_ntt((u64 *)lde_cols + c0, w, coef, cp, tmp, cp, loge, ww, false, 0, n);
            CUP(cudaEventRecord(ev_done[k], g.stream));
        }
        cudaError_t e = cudaGetLastError();
        for (u64 k = 0; k < nb; k++) { cudaEventDestroy(ev_in[k]); cudaEventDestroy(ev_done[k]); } // released when complete
        if (rc == GL_OK && e != cudaSuccess) rc = fail(GL_ERR_CUDA, "pipelined multi-GPU commit: %s", cudaGetErrorString(e));
        if (rc != GL_OK) {
            cudaDeviceSynchronize();
            return rc;
        }
    }
    CU(cudaEventRecord(g.ev[4], g.stream));

    TRY(mgpu_exchange_and_hash((const u64 *)lde_cols, (u64 *)row_panels, (u64 *)subtree, rows, w, ncols_total, (u64 *)row_panels_host, use_ce));
    g.launches += poseidon::launch_tree_levels((u64 *)subtree, rows, g.stream);
    CU(cudaGetLastError());
    TRY(mgpu_top((const u64 *)subtree + gl_tree_num_elements(rows) - 4, root));
    CU(cudaEventRecord(g.ev[3], g.stream));
    if (subtree_host) CU(cudaMemcpyAsync(subtree_host, subtree, gl_tree_num_elements(rows) * 8, cudaMemcpyDeviceToHost, g.stream));
    CU(cudaStreamSynchronize(g.stream));
    CU(cudaStreamSynchronize(g.s_d2h));
    CU(cudaStreamSynchronize(g.s_h2d));
    CU(cudaEventElapsedTime(&g.ms[0], g.ev[0], g.ev[4]));
    CU(cudaEventElapsedTime(&g.ms[3], g.ev[4], g.ev[1]));
    CU(cudaEventElapsedTime(&g.ms[1], g.ev[1], g.ev[2]));
    CU(cudaEventElapsedTime(&g.ms[2], g.ev[2], g.ev[3]));
    return GL_OK;
}

int gl_mgpu_gather_tree(uint64_t *tree_full, const uint64_t *subtree, uint64_t n_ext)
{
    API_ENTER();
    if (!nc.comm) return fail(GL_ERR_ARG, "gl_mgpu_gather_tree: call gl_mgpu_init first");
    const u64 G = (u64)nc.world, rows = n_ext / G;
    u32 lv = 0;
    for (u64 cnt = rows; cnt >= 1; cnt >>= 1, lv++) {
        NC(nc.AllGather((const u64 *)subtree + level_offset(rows, lv), (u64 *)tree_full + level_offset(n_ext, lv), cnt * 4, ncclUint64,
                        nc.comm, g.stream));
        if (cnt == 1) break;
    }
    // levels above the subtree roots (G nodes at level lv)
    u64 pending = G, next_index = level_offset(n_ext, lv);
    while (pending > 1) {
        poseidon::k_tree_level<<<poseidon::grid_for(pending / 2), poseidon::kThreads, 0, g.stream>>>((u64 *)tree_full + next_index + pending * 4,
                                                                                                       (u64 *)tree_full + next_index, pending / 2);
        g.launches++;
        next_index += pending * 4;
        pending /= 2;
    }
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(g.stream));
    return GL_OK;
}

int gl_fill_synthetic_cols_dev(uint64_t *dst, uint64_t nrows, uint64_t col0, uint64_t ncols_local, uint64_t ncols_total, uint64_t seed)
{
    API_ENTER();
    if (nrows == 0 || ncols_local == 0) return GL_OK;
    k_fill_synthetic_cols<<<148 * 8, 256, 0, g.stream>>>((u64 *)dst, nrows, col0, ncols_local, ncols_total, seed);
    g.launches++;
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(g.stream));
    return GL_OK;
}

// ---- next rows: Q-polynomial commit, FRI layer commit -----------------------------------
int gl_q_extend_dev(uint64_t *cm4, const uint64_t *q_2ns, uint64_t n, uint64_t n_ext, uint64_t qdim, uint64_t qdeg)
{
    API_ENTER();
    if (n == 0 || qdim == 0 || qdeg == 0) return GL_OK;
    if (!is_pow2(n) || !is_pow2(n_ext) || n_ext < n) return fail(GL_ERR_ARG, "q_extend: bad sizes %llu -> %llu", (u64)n, (u64)n_ext);
    if (qdeg > (u64)fri::kMaxQDeg || qdeg * n > n_ext) return fail(GL_ERR_ARG, "q_extend: qDeg %llu out of range", (u64)qdeg);
    const u64 w = qdim * qdeg;
    const u32 loge = ilog2(n_ext);
    DevBuf qq1, qq2;
    TRY(qq1.alloc(n_ext * qdim * 8));
    TRY(qq2.alloc(n * w * 8));
    CU(cudaEventRecord(g.ev[0], g.stream));
    TRY(ntt_dev((u64 *)qq1.p, (const u64 *)q_2ns, n_ext, qdim, 1, 0));                    // starks.cpp:261
    fri::QScale sc;
    const u64 shift_in = hpow(hpow(7, P - 2), n);                                        // (1/7)^N, starks.cpp:265
    sc.s[0] = 1;
    for (int p = 1; p < fri::kMaxQDeg; p++) sc.s[p] = hmul(sc.s[p - 1], shift_in);
    fri::k_q_regroup<<<148 * 8, 256, 0, g.stream>>>((u64 *)qq2.p, (const u64 *)qq1.p, n, (u32)qdim, (u32)qdeg, sc); // :268-280
    g.launches++;
    CU(cudaGetLastError());
    // forward NTT over the extended domain; rows >= n of qq2 are zeros (:284)
    u32 cb = pick_col_block(w, n_ext * 8);
    const u64 cp = pad16(cb);
    if (loge > 8) TRY(arena_reserve((size_t)n_ext * cp * 8));
    for (u64 c0 = 0; c0 < w; c0 += cb) {
        u32 ww = (u32)((w - c0 < cb) ? w - c0 : cb);
        TRY(run_ntt((u64 *)cm4 + c0, w, (const u64 *)qq2.p + c0, w, (u64 *)g.arena, cp, loge, ww, false, 0, n));
    }
    CU(cudaEventRecord(g.ev[1], g.stream));
    return collect_timings(true, false);
}

int gl_q_commit_dev(uint64_t *cm4, uint64_t *tree, const uint64_t *q_2ns, uint64_t n, uint64_t n_ext, uint64_t qdim, uint64_t qdeg,
                    uint64_t root[4])
{
    API_ENTER();
    TRY(gl_q_extend_dev(cm4, q_2ns, n, n_ext, qdim, qdeg));
    const float lde_ms = g.ms[0];
    TRY(merkletree_dev((u64 *)tree, (const u64 *)cm4, qdim * qdeg, n_ext, true));          // starks.cpp:289
    if (root) CU(cudaMemcpyAsync(root, tree + gl_tree_num_elements(n_ext) - 4, 32, cudaMemcpyDeviceToHost, g.stream));
    TRY(collect_timings(false, true));
    g.ms[0] = lde_ms;
    return GL_OK;
}

int gl_fri_fold_dev(uint64_t *out, const uint64_t *pol, uint32_t pol_bits, uint32_t next_bits, uint64_t shift_inv, const uint64_t x[3])
{
    API_ENTER();
    if (next_bits > pol_bits || pol_bits > 32 || pol_bits - next_bits > 6) return fail(GL_ERR_ARG, "fri_fold: bits %u -> %u unsupported", pol_bits, next_bits);
    if (pol_bits > kLogRootsLo) return fail(GL_ERR_ARG, "fri_fold: polynomials above 2^%u are not supported", kLogRootsLo);
    TRY(ensure_roots(pol_bits));
    fri::F3 xx = {x[0], x[1], x[2]};
    const u64 pol2n = 1ULL << next_bits;
    const u64 nx_inv = hpow((P + 1) / 2, pol_bits - next_bits);
    fri::k_fri_fold<<<(unsigned)((pol2n + 127) / 128), 128, 0, g.stream>>>((u64 *)out, (const u64 *)pol, pol_bits, next_bits, shift_inv, xx, g.roots,
                                                                           g.log_roots, nx_inv);
    g.launches++;
    CU(cudaGetLastError());
    return GL_OK;
}

int gl_fri_transpose_dev(uint64_t *aux, const uint64_t *pol, uint64_t degree, uint32_t transpose_bits)
{
    API_ENTER();
    if (degree == 0) return GL_OK;
    if ((degree >> transpose_bits) == 0) return fail(GL_ERR_ARG, "fri_transpose: 2^%u columns exceed the degree", transpose_bits);
    fri::k_fri_transpose<<<148 * 8, 256, 0, g.stream>>>((u64 *)aux, (const u64 *)pol, degree, transpose_bits);
    g.launches++;
    CU(cudaGetLastError());
    return GL_OK;
}

int gl_fri_layer_commit_dev(uint64_t *tree, uint64_t *aux, uint64_t *pol2, const uint64_t *pol, uint32_t pol_bits, uint32_t next_bits,
                            uint32_t tree_bits, uint64_t shift_inv, const uint64_t x[3], uint64_t root[4])
{
    API_ENTER();
    if (tree_bits > next_bits) return fail(GL_ERR_ARG, "fri_layer: tree of 2^%u groups from 2^%u elements", tree_bits, next_bits);
    TRY(gl_fri_fold_dev(pol2, pol, pol_bits, next_bits, shift_inv, x));                    // friProve.cpp:36-107
    const u64 pol2n = 1ULL << next_bits, ngroups = 1ULL << tree_bits, group = pol2n / ngroups;
    TRY(gl_fri_transpose_dev(aux, pol2, pol2n, tree_bits));                                // :112-117
    TRY(merkletree_dev((u64 *)tree, (const u64 *)aux, group * 3, ngroups, true));          // :120-123
    if (root) CU(cudaMemcpyAsync(root, tree + gl_tree_num_elements(ngroups) - 4, 32, cudaMemcpyDeviceToHost, g.stream));
    return collect_timings(false, true);
}

// Host-pointer forms of the two callers above (what a Starks / FRIProve that keeps its polynomials in host memory calls): upload,
// the device-resident call, download of whatever the caller asked for (NULL = not wanted).
int gl_q_commit_host(uint64_t *cm4_host, uint64_t *tree_host, const uint64_t *q_2ns_host, uint64_t n, uint64_t n_ext, uint64_t qdim,
                     uint64_t qdeg, uint64_t root[4])
{
    API_ENTER();
    if (n == 0 || n_ext == 0 || qdim == 0 || qdeg == 0) return fail(GL_ERR_ARG, "q_commit: empty input");
    DevBuf q, cm4, tree;
    const size_t qb = (size_t)n_ext * qdim * 8, cb = qb * qdeg, tb = gl_tree_num_elements(n_ext) * 8;
    TRY(q.alloc(qb));
    TRY(cm4.alloc(cb));
    TRY(tree.alloc(tb));
    CU(cudaMemcpyAsync(q.p, q_2ns_host, qb, cudaMemcpyHostToDevice, g.stream));
    TRY(gl_q_commit_dev(cm4.w(), tree.w(), q.w(), n, n_ext, qdim, qdeg, root));
    if (cm4_host) CU(cudaMemcpyAsync(cm4_host, cm4.p, cb, cudaMemcpyDeviceToHost, g.stream));
    if (tree_host) CU(cudaMemcpyAsync(tree_host, tree.p, tb, cudaMemcpyDeviceToHost, g.stream));
    CU(cudaStreamSynchronize(g.stream));
    return GL_OK;
}

int gl_fri_layer_commit_host(uint64_t *tree_host, uint64_t *aux_host, uint64_t *pol2_host, const uint64_t *pol_host, uint32_t pol_bits,
                             uint32_t next_bits, uint32_t tree_bits, uint64_t shift_inv, const uint64_t x[3], uint64_t root[4])
{
    API_ENTER();
    if (next_bits > pol_bits || tree_bits > next_bits) return fail(GL_ERR_ARG, "fri_layer: bits %u -> %u, tree 2^%u", pol_bits, next_bits, tree_bits);
    DevBuf pol, pol2, aux, tree;
    const size_t pb = (size_t)24 << pol_bits, nb = (size_t)24 << next_bits, tb = gl_tree_num_elements(1ULL << tree_bits) * 8;
    TRY(pol.alloc(pb));
    TRY(pol2.alloc(nb));
    TRY(aux.alloc(nb));
    TRY(tree.alloc(tb));
    CU(cudaMemcpyAsync(pol.p, pol_host, pb, cudaMemcpyHostToDevice, g.stream));
    TRY(gl_fri_layer_commit_dev(tree.w(), aux.w(), pol2.w(), pol.w(), pol_bits, next_bits, tree_bits, shift_inv, x, root));
    if (pol2_host) CU(cudaMemcpyAsync(pol2_host, pol2.p, nb, cudaMemcpyDeviceToHost, g.stream));
    if (aux_host) CU(cudaMemcpyAsync(aux_host, aux.p, nb, cudaMemcpyDeviceToHost, g.stream));
    if (tree_host) CU(cudaMemcpyAsync(tree_host, tree.p, tb, cudaMemcpyDeviceToHost, g.stream));
    CU(cudaStreamSynchronize(g.stream));
    return GL_OK;
}

// ---- offline constant tree (tools/starkpil/bctree/build_const_tree.cpp:333-407) -------------------------
uint64_t gl_consttree_num_elements(uint64_t npols, uint64_t n_ext) { return 2 + npols * n_ext + gl_tree_num_elements(n_ext); }

int gl_build_consttree_host(uint64_t *const_tree, const uint64_t *const_pols, uint64_t n, uint64_t n_ext, uint64_t npols)
{
    API_ENTER();
    if (n == 0 || n_ext == 0 || npols == 0) return fail(GL_ERR_ARG, "buildConstTree: empty input");
    const_tree[0] = npols; // header read back by MerkleTreeGL(Goldilocks::Element *tree), merkleTreeGL.hpp:24-32
    const_tree[1] = n_ext;
    return commit_host_pipelined(const_tree + 2, const_tree + 2 + npols * n_ext, const_pols, n_ext, n, npols, nullptr, true);
}

int gl_last_timings4(float ms[4])
{
    for (int i = 0; i < 4; i++) ms[i] = g.ms[i];
    return GL_OK;
}

} // extern "C"

// ---- allocation-free extendPol (no counterpart in the reference) ---------------
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it): the
// twiddle table, the coset-scale table and the column-block scratch of extend_pol_dev live in ONE file-scope device array of fixed
// size instead of cudaMalloc'ed buffers.  Every GPU that loads the library holds the array (33.0 MiB) until the process exits.
// Layout (words): [0, 2^16) roots w_(2^16)^i | [2^16, 2^17) 7^i / n of the last n | [2^17, 2^17 + 2^22) INTT result + pass scratch.
// All calls share the array and it keeps the previous call's contents.  Nothing in it has to start at zero: a call writes every
// word before reading it (the INTT writes all n x w coefficients, the first forward pass reads only rows < n and writes all
// n_ext x w words of the pass scratch), the roots are built once per device and the scale table is rebuilt when n changes.
// Calls are serialised on the library stream; two calls overlapping on the GPU (another process-wide user of the symbol) would race.
constexpr unsigned kNoAllocLogMax = 16;                                     // n_ext <= 2^16
constexpr unsigned long long kNoAllocRootWords = 1ULL << kNoAllocLogMax;
constexpr unsigned long long kNoAllocScaleWords = 1ULL << kNoAllocLogMax;
constexpr unsigned long long kNoAllocWorkWords = 1ULL << 22;                // (n + n_ext) * (columns per block)
static __device__ __align__(256) unsigned long long s_noalloc[kNoAllocRootWords + kNoAllocScaleWords + kNoAllocWorkWords];

namespace {
struct NoAllocDev {
    u64 *base = nullptr;   // DEVICE address of s_noalloc on this device (the symbol's name is only a host-side stand-in)
    bool roots_built = false;
    u32 log_scale = ~0u;   // n of the scale table currently in the array
} na_dev[16];

// once per device, with that device current
int noalloc_setup(NoAllocDev **out)
{
    if (g.device < 0 || g.device >= 16) return fail(GL_ERR_ARG, "no-alloc extendPol: device %d out of range", g.device);
    NoAllocDev &d = na_dev[g.device];
    if (!d.base) {
        void *ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, s_noalloc);
        if (e != cudaSuccess || !ptr) return fail(GL_ERR_CUDA, "cudaGetSymbolAddress(s_noalloc): %s", cudaGetErrorString(e));
        d.base = (u64 *)ptr;
        d.roots_built = false;
        d.log_scale = ~0u;
    }
    if (!d.roots_built) {
        ntt::k_build_roots<<<148, 256, 0, g.stream>>>(d.base, kNoAllocLogMax, kNoAllocRootWords, 0);
        g.launches++;
        CU(cudaGetLastError());
        d.roots_built = true;
    }
    *out = &d;
    return GL_OK;
}

// run_ntt with caller-supplied tables, always through ntt::k_pass (no tensor maps)
int run_ntt_tables(u64 *dst, u64 dpitch, const u64 *src, u64 spitch, u64 *tmp, u64 tpitch, u32 logn, u32 ncols, bool inverse,
                   int scale_mode, u64 nvalid, const u64 *roots, u32 log_roots, const u64 *rscale)
{
    ntt::PassParams p;
    memset(&p, 0, sizeof p);
    p.ncols = ncols;
    p.logn = logn;
    p.inverse = inverse ? 1 : 0;
    if (logn <= 3) {
        p.src = src; p.dst = dst; p.spitch = spitch; p.dpitch = dpitch;
        p.scale_mode = scale_mode; p.nvalid = nvalid; p.npass = 1;
        ntt::k_tiny<<<(ncols + 127) / 128, 128, 0, g.stream>>>(p);
        g.launches++;
        CU(cudaGetLastError());
        return GL_OK;
    }
    const u32 m = (logn + 7) / 8;
    p.npass = m;
    for (u32 i = 0; i < m; i++) p.bits[i] = logn / m + (i < logn % m ? 1 : 0);
    p.roots = roots;
    p.log_roots = log_roots;
    p.rscale = rscale;
    p.roots_hi = nullptr;
    p.log_lo = kLogRootsLo;
    p.cscale = scale_mode == 1 ? hpow((P + 1) / 2, logn) : 1;
    u32 prev = 0;
    for (u32 i = 0; i < m; i++) {
        const bool first = i == 0, last = i + 1 == m;
        p.pass = i;
        p.log_prev = prev;
        p.log_rest = logn - prev - p.bits[i];
        p.src = first ? src : tmp;
        p.spitch = first ? spitch : tpitch;
        p.dst = last ? dst : tmp;
        p.dpitch = last ? dpitch : tpitch;
        p.nvalid = first ? nvalid : (1ULL << logn);
        p.scale_mode = last ? scale_mode : 0;
        TRY(launch_pass(p));
        prev += p.bits[i];
    }
    return GL_OK;
}
} // namespace

extern "C" int gl_extend_pol_dev_noalloc(uint64_t *out, const uint64_t *in, uint64_t n_ext, uint64_t n, uint64_t ncols)
{
    API_ENTER();
    if (n == 0 || ncols == 0) return GL_OK;
    if (!is_pow2(n) || !is_pow2(n_ext) || n_ext < n) return fail(GL_ERR_ARG, "extendPol: bad sizes %llu -> %llu", (u64)n, (u64)n_ext);
    if (n_ext > (1ULL << kNoAllocLogMax))
        return fail(GL_ERR_RANGE, "gl_extend_pol_dev_noalloc: n_ext = %llu exceeds the static scratch (at most 2^%u extended rows); use gl_extend_pol_dev",
                    (u64)n_ext, kNoAllocLogMax);
    if (ncols > 0xFFFFFFFFULL) return fail(GL_ERR_ARG, "extendPol: %llu columns", (u64)ncols);
    NoAllocDev *d = nullptr;
    TRY(noalloc_setup(&d));
    const u32 logn = ilog2(n), loge = ilog2(n_ext);
    u64 *roots = d->base, *rscale = d->base + kNoAllocRootWords, *work = rscale + kNoAllocScaleWords;
    if (d->log_scale != logn) { // stream order keeps the rebuild behind the previous call's readers
        ntt::k_build_rscale<<<148, 256, 0, g.stream>>>(rscale, logn);
        g.launches++;
        CU(cudaGetLastError());
        d->log_scale = logn;
    }
    u64 cb = kNoAllocWorkWords / (n + n_ext); // >= 32 at the largest size
    if (cb > 32) cb = (cb / 32) * 32;
    if (cb > ncols) cb = ncols;
    u64 *coef = work, *tmp = work + n * cb;
    for (u64 c0 = 0; c0 < ncols; c0 += cb) {
        const u32 w = (u32)((ncols - c0 < cb) ? ncols - c0 : cb);
        TRY(run_ntt_tables(coef, cb, (const u64 *)in + c0, ncols, tmp, cb, logn, w, true, 2, n, roots, kNoAllocLogMax, rscale));
        TRY(run_ntt_tables((u64 *)out + c0, ncols, coef, cb, tmp, cb, loge, w, false, 0, n, roots, kNoAllocLogMax, rscale));
    }
    CU(cudaStreamSynchronize(g.stream));
    return GL_OK;
}
