// oracle/shim/bctree/nlohmann/json.hpp -- TEST INFRASTRUCTURE ONLY.
// Compile-only stand-in for nlohmann::json (absent from this image) so that the reference's
// tools/starkpil/bctree/build_const_tree.cpp compiles where it lies; only its interpolate() (:198-330) is ever
// called (through oracle/ref_bctree_wrap.cpp).  buildConstTree(), the one function that touches json, is never run.
#pragma once
#include <cstdint>
#include <ostream>
#include <string>
namespace nlohmann {
struct json {
    json() {}
    template <class T> json(const T &) {}
    json &operator[](const std::string &) { return *this; }
    json &operator[](const char *) { return *this; }
    json &operator[](int) { return *this; }
    template <class T> json &operator=(const T &) { return *this; }
    operator uint64_t() const { return 0; }
    bool operator==(const char *) const { return false; }
};
inline std::ostream &operator<<(std::ostream &o, const json &) { return o; }
typedef json ordered_json;
} // namespace nlohmann
