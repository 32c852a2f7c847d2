// oracle/ref_bctree_wrap.cpp -- TEST INFRASTRUCTURE ONLY.
// extern "C" veneer over interpolate() of the reference's offline constant-tree tool
// (tools/starkpil/bctree/build_const_tree.cpp:198-330: its own block FFT, interpolateBitReverse :115-124,
// interpolatePrepare :160-196 with shift^i / N), compiled where it lies under /root/reference by oracle/Makefile with the
// compile-only stubs of oracle/shim/bctree/.  This is a SECOND implementation of the coset extension, independent of
// NTT_Goldilocks::extendPol (the tool's implicit zero padding is made explicit by oracle/shim/bctree/zalloc.h); tests/golden/make_golden.py stores its outputs as fixtures for the gl_build_consttree_host parity test.
#include "build_const_tree.hpp"
#include "utils.hpp"
#include <cstdlib>
#include <sstream>

void interpolate(Goldilocks::Element *buffSrc, uint64_t nPols, uint64_t nBits, Goldilocks::Element *buffDst, uint64_t nBitsExt);

// helpers of src/utils that buildConstTree() (never called here) links against
std::string DateAndTime(struct timeval &) { return std::string(); }
void json2file(const nlohmann::json &, const std::string &) { abort(); }
void file2json(const std::string &, nlohmann::json &) { abort(); }
void *copyFile(const std::string &, uint64_t) { abort(); }

extern "C" void ref_bctree_interpolate(uint64_t *dst, uint64_t *src, uint64_t nPols, uint64_t nBits, uint64_t nBitsExt)
{
    // the tool sizes its blocks from omp_get_max_threads() / 2 and divides by it (:170, :208): with an OpenMP team of one
    // (e.g. after NTT_Goldilocks(nX, 1) of the FRI fold called omp_set_num_threads(1), ntt_goldilocks.cpp:59) it divides by zero
    if (omp_get_max_threads() < 2) omp_set_num_threads(2);
    std::ostringstream sink; // interpolate() logs its phases to std::cout
    std::streambuf *old = std::cout.rdbuf(sink.rdbuf());
    interpolate((Goldilocks::Element *)src, nPols, nBits, (Goldilocks::Element *)dst, nBitsExt);
    std::cout.rdbuf(old);
}
