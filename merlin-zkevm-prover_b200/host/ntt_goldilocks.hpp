// ntt_goldilocks.hpp -- drop-in for the reference's NTT_Goldilocks (src/goldilocks/src/ntt_goldilocks.hpp:12-173)
// that runs the transforms on the B200 through the C ABI (include/gl_commit.h).
//
// Same class name, constructor and member signatures/defaults, so Starks::genProof (starks.cpp:52,133,214,
// 261,284,325-326), StarkRecursiveF::genProof and FRIProve::prove (friProve.cpp:100-102) compile unchanged.
// Differences that are invisible to those callers:
//   * no root table is built (the reference rebuilds 2^24 roots on every extendPol, ntt_goldilocks.cpp:328);
//   * nThreads, nphase, nblock and the scratch `buffer` are CPU blocking hints and are ignored;
//   * results are canonical words (the reference returns lazily reduced words; equal under toU64).
#ifndef NTT_GOLDILOCKS
#define NTT_GOLDILOCKS

#include "goldilocks_base_field.hpp"
#include "glb200_shim.hpp"
#include <cassert>

#define NUM_PHASES 3
#define NUM_BLOCKS 1

class NTT_Goldilocks
{
private:
    u_int32_t s = 0;
    u_int32_t nThreads = 0;
    int extension = 1;

    static u_int32_t log2(u_int64_t size)
    {
        assert(size != 0);
        u_int32_t res = 0;
        while (size != 1)
        {
            size >>= 1;
            res++;
        }
        return res;
    }

public:
    NTT_Goldilocks(u_int64_t maxDomainSize, u_int32_t _nThreads = 0, int extension_ = 1)
    {
        if (maxDomainSize == 0)
            return;
        nThreads = _nThreads;
        extension = extension_;
        u_int32_t domainPow = NTT_Goldilocks::log2(maxDomainSize);
        if (domainPow > 32)
        {
            throw std::range_error("Domain size too big for the curve"); // ntt_goldilocks.hpp:94-97
        }
        s = domainPow;
    }
    ~NTT_Goldilocks() {}

    // ntt_goldilocks.hpp:169-172: w(domainPow)^idx, i.e. roots[idx << (s - domainPow)] of the table the reference builds in its
    // constructor.  No table is kept here (the transforms use the device table), so the value is computed on demand:
    // w(k) = W[32]^(2^(32-k)) with W[32] = 7277203076849721926 (goldilocks_base_field.cpp:44).  The reference returned refers
    // to a per-thread slot that the next call on the same thread overwrites.
    Goldilocks::Element &root(u_int32_t domainPow, u_int64_t idx)
    {
        static thread_local Goldilocks::Element slot;
        const uint64_t p = 0xFFFFFFFF00000001ULL;
        auto mulm = [p](uint64_t a, uint64_t b) { return (uint64_t)(((unsigned __int128)a * b) % p); };
        uint64_t w = 7277203076849721926ULL;
        for (u_int32_t i = domainPow; i < 32; i++) w = mulm(w, w);
        uint64_t r = 1;
        for (uint64_t e = idx; e; e >>= 1, w = mulm(w, w))
            if (e & 1) r = mulm(r, w);
        slot = Goldilocks::fromU64(r);
        return slot;
    }

    void NTT(Goldilocks::Element *dst, Goldilocks::Element *src, u_int64_t size, u_int64_t ncols = 1, Goldilocks::Element *buffer = NULL, u_int64_t nphase = NUM_PHASES, u_int64_t nblock = NUM_BLOCKS, bool inverse = false, bool extend = false)
    {
        (void)buffer; (void)nphase; (void)nblock;
        if (ncols == 0 || size == 0)
            return; // ntt_goldilocks.cpp:168-171
        glb200::check(gl_ntt_host((uint64_t *)(dst ? dst : src), (const uint64_t *)src, size, ncols, inverse ? 1 : 0, extend ? 1 : 0), "NTT_Goldilocks::NTT");
    }

    void INTT(Goldilocks::Element *dst, Goldilocks::Element *src, u_int64_t size, u_int64_t ncols = 1, Goldilocks::Element *buffer = NULL, u_int64_t nphase = NUM_PHASES, u_int64_t nblock = NUM_BLOCKS, bool extend = false)
    {
        NTT(dst, src, size, ncols, buffer, nphase, nblock, true, extend); // ntt_goldilocks.cpp:307-324
    }

    void extendPol(Goldilocks::Element *output, Goldilocks::Element *input, uint64_t N_Extended, uint64_t N, uint64_t ncols, Goldilocks::Element *buffer = NULL, u_int64_t nphase = NUM_PHASES, u_int64_t nblock = NUM_BLOCKS)
    {
        (void)buffer; (void)nphase; (void)nblock;
        glb200::check(gl_extend_pol_host((uint64_t *)output, (const uint64_t *)input, N_Extended, N, ncols), "NTT_Goldilocks::extendPol");
    }
};

#endif
