// oracle/shim/bctree/merkleTreeBN128.hpp -- TEST INFRASTRUCTURE ONLY: compile-only stand-in for the BN128 tree (out of scope,
// SURVEY 2) named by buildConstTree()'s BN128 branch, which is never run here.
#pragma once
#include "goldilocks_base_field.hpp"
#include <string>
struct RawFr {
    struct Element { uint64_t v[4]; };
    std::string toString(const Element &) { return std::string(); }
};
struct MerkleTreeBN128 {
    uint64_t source_width = 0, height = 0, numNodes = 0;
    Goldilocks::Element *source = nullptr;
    RawFr::Element *nodes = nullptr;
    MerkleTreeBN128(uint64_t, uint64_t, Goldilocks::Element *) {}
    void merkelize() {}
    void getRoot(RawFr::Element *) {}
};
