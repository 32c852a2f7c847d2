// merkleTreeGL.hpp -- drop-in for the reference's MerkleTreeGL (src/starkpil/merkleTree/merkleTreeGL.hpp:9-79,
// merkleTreeGL.cpp:5-44): same public fields (height, width, source, nodes), constructors, ownership rules
// and methods.  merkelize() builds the tree on the B200; the proof getters read the host copies exactly as
// the reference does.
#ifndef MERKLETREEGL
#define MERKLETREEGL

#include "goldilocks_base_field.hpp"
#include "poseidon_goldilocks.hpp"
#include <cstdlib>
#include <math.h>

#define MERKLEHASHGL_ARITY 2
class MerkleTreeGL
{
private:
    void genMerkleProof(Goldilocks::Element *proof, uint64_t idx, uint64_t offset, uint64_t n)
    {
        // iterative form of merkleTreeGL.cpp:24-35
        while (n > HASH_SIZE)
        {
            uint64_t si = (idx ^ 1) * HASH_SIZE;
            std::memcpy(proof, &nodes[offset + si], HASH_SIZE * sizeof(Goldilocks::Element));
            proof += HASH_SIZE;
            idx >>= 1;
            uint64_t nextN = ((n - 1) / 8 + 1) * HASH_SIZE;
            offset += nextN * 2;
            n = nextN;
        }
    }

public:
    uint64_t height;
    uint64_t width;
    Goldilocks::Element *source;
    Goldilocks::Element *nodes;
    bool isSourceAllocated = false;
    bool isNodesAllocated = false;
    MerkleTreeGL(){};
    MerkleTreeGL(Goldilocks::Element *tree)
    {
        // view over a [width, height, source..., nodes...] buffer (.consttree layout, merkleTreeGL.hpp:24-32)
        width = Goldilocks::toU64(tree[0]);
        height = Goldilocks::toU64(tree[1]);
        source = &tree[2];
        nodes = &tree[2 + height * width];
    };
    MerkleTreeGL(uint64_t _height, uint64_t _width, Goldilocks::Element *_source) : height(_height), width(_width), source(_source)
    {
        if (source == NULL)
        {
            source = (Goldilocks::Element *)calloc(height * width, sizeof(Goldilocks::Element));
            isSourceAllocated = true;
        }
        nodes = (Goldilocks::Element *)calloc(getTreeNumElements(), sizeof(Goldilocks::Element));
        isNodesAllocated = true;
    };
    ~MerkleTreeGL()
    {
        if (isSourceAllocated)
            free(source);
        if (isNodesAllocated)
            free(nodes);
    };
    void copySource(Goldilocks::Element *_source)
    {
        std::memcpy(source, _source, height * width * sizeof(Goldilocks::Element));
    }
    void merkelize()
    {
        PoseidonGoldilocks::merkletree(nodes, source, width, height); // merkleTreeGL.cpp:37-44
    }
    uint64_t getTreeNumElements()
    {
        return height * HASH_SIZE + (height - 1) * HASH_SIZE;
    }
    void getRoot(Goldilocks::Element *root)
    {
        std::memcpy(root, &nodes[getTreeNumElements() - HASH_SIZE], HASH_SIZE * sizeof(Goldilocks::Element));
    }
    void getGroupProof(Goldilocks::Element *proof, uint64_t idx)
    {
        std::memcpy(proof, &source[idx * width], width * sizeof(Goldilocks::Element)); // merkleTreeGL.cpp:12-22
        genMerkleProof(&proof[width], idx, 0, height * HASH_SIZE);
    }
    uint64_t MerkleProofSize()
    {
        if (height > 1)
            return (uint64_t)ceil(log10(height) / log10(MERKLEHASHGL_ARITY));
        return 0;
    }
};

#endif
