/* Minimal mpz_class -- TEST INFRASTRUCTURE ONLY (oracle/_ref build).
 * Only the operations the reference's goldilocks_base_field_tools.hpp:118-240
 * spells out are provided (string ctor, +, -, %, comparisons, get_ui/si/str).
 * Also pulls in the std headers the real gmpxx.h leaks and the reference
 * relies on.
 */
#ifndef GLB200_ORACLE_GMPXX_SHIM_H
#define GLB200_ORACLE_GMPXX_SHIM_H
#include <gmp.h>
#include <cstring>
#include <cstdlib>
#include <cstdint>
#include <cmath>
#include <iostream>
#include <string>
#include <vector>
#include <stdexcept>

class mpz_class
{
    mpz_t v;
public:
    mpz_class() { mpz_init(v); }
    mpz_class(const mpz_class &o) { mpz_init_set(v, o.v); }
    mpz_class(unsigned long x) { mpz_init_set_ui(v, x); }
    mpz_class(unsigned long long x) { mpz_init_set_ui(v, (unsigned long)x); }
    mpz_class(long x) { mpz_init_set_ui(v, (unsigned long)(x < 0 ? -x : x)); if (x < 0) { mpz_t z; mpz_init(z); mpz_sub(v, z, v); mpz_clear(z); } }
    mpz_class(int x) : mpz_class((long)x) {}
    mpz_class(unsigned int x) : mpz_class((unsigned long)x) {}
    mpz_class(const std::string &s, int radix = 10)
    {
        if (mpz_init_set_str(v, s.c_str(), radix) != 0) { mpz_clear(v); throw std::invalid_argument("mpz_class: bad string"); }
    }
    ~mpz_class() { mpz_clear(v); }
    mpz_class &operator=(const mpz_class &o) { if (this != &o) mpz_set(v, o.v); return *this; }
    unsigned long get_ui() const { return mpz_get_ui(v); }
    long get_si() const { return mpz_get_si(v); }
    std::string get_str(int radix = 10) const
    {
        char *c = mpz_get_str(NULL, radix, v);
        std::string s(c);
        free(c);
        return s;
    }
    friend mpz_class operator+(const mpz_class &a, const mpz_class &b) { mpz_class r; mpz_add(r.v, a.v, b.v); return r; }
    friend mpz_class operator-(const mpz_class &a, const mpz_class &b) { mpz_class r; mpz_sub(r.v, a.v, b.v); return r; }
    friend mpz_class operator%(const mpz_class &a, const mpz_class &b) { mpz_class r; mpz_mod(r.v, a.v, b.v); return r; }
    friend bool operator>(const mpz_class &a, const mpz_class &b) { return mpz_cmp(a.v, b.v) > 0; }
    friend bool operator<(const mpz_class &a, const mpz_class &b) { return mpz_cmp(a.v, b.v) < 0; }
    friend bool operator==(const mpz_class &a, const mpz_class &b) { return mpz_cmp(a.v, b.v) == 0; }
};
#endif
