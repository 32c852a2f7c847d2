// Stand-in for the reference's goldilocks_base_field.hpp, used ONLY when the shims are built outside
// the reference tree (this repo's tests).  Inside the reference tree the real header
// (src/goldilocks/src/goldilocks_base_field.hpp) is found first on the include path and this file is
// not used.  It declares just what the shim headers need: the Element type and U64 conversion.
#ifndef GOLDILOCKS_BASE
#define GOLDILOCKS_BASE
#include <cstdint>
#include <cstring>
#include <sys/types.h>

#define GOLDILOCKS_PRIME 0xFFFFFFFF00000001ULL

class Goldilocks
{
public:
    typedef struct
    {
        uint64_t fe;
    } Element;

    static inline Element fromU64(uint64_t v)
    {
        Element e;
        e.fe = v >= GOLDILOCKS_PRIME ? v - GOLDILOCKS_PRIME : v;
        return e;
    }
    static inline uint64_t toU64(const Element &e) { return e.fe >= GOLDILOCKS_PRIME ? e.fe - GOLDILOCKS_PRIME : e.fe; }
    static inline Element zero() { return fromU64(0); }
    static inline Element one() { return fromU64(1); }
};
#endif
