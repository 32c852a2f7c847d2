// glb200_shim.hpp -- shared plumbing of the header-compatible shims: C-ABI declarations + error mapping.
#ifndef GLB200_SHIM_HPP
#define GLB200_SHIM_HPP
#include "gl_commit.h"
#include <stdexcept>
#include <string>

namespace glb200
{
    // C-ABI status -> the exception the reference would have raised at the same point:
    // std::range_error for a domain > 2^32 (ntt_goldilocks.hpp:94-97), std::runtime_error otherwise
    // (the reference asserts / exits the process; a drop-in must at least not continue silently).
    inline void check(int rc, const char *where)
    {
        if (rc == GL_OK)
            return;
        std::string msg = std::string(where) + ": " + gl_last_error();
        if (rc == GL_ERR_RANGE)
            throw std::range_error("Domain size too big for the curve");
        throw std::runtime_error(msg);
    }
}
#endif
