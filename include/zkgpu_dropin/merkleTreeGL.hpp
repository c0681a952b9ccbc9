// merkleTreeGL.hpp — drop-in for the reference's src/starkpil/merkleTree/merkleTreeGL.{hpp,cpp}.
//
// Same class name, public fields (height, width, source, nodes — FRI reads `width` directly, friProve.cpp:225),
// constructors and methods as reference merkleTreeGL.hpp:17-78 / merkleTreeGL.cpp:5-44.  merkelize() runs on the GPU
// (zkgpu_merkletree); getGroupProof() is served by zkgpu_group_proof, which reads the row and the sibling path from
// wherever `source` / `nodes` live (host memory, or device memory when the commit stage left them resident —
// MerkleTreeGL(height, width, source, nodes) below wraps such device buffers without copying).
#ifndef MERKLETREEGL
#define MERKLETREEGL
#include "goldilocks_element.hpp"
#include "poseidon_goldilocks.hpp"
#include <cstdlib>
#include <cstring>

#define MERKLEHASHGL_ARITY 2
class MerkleTreeGL
{
public:
    uint64_t height;
    uint64_t width;
    Goldilocks::Element *source;
    Goldilocks::Element *nodes;
    bool isSourceAllocated = false;
    bool isNodesAllocated = false;
    MerkleTreeGL(){};
    // view of a file image [width, height, source..., nodes...] (merkleTreeGL.hpp:24-32)
    MerkleTreeGL(Goldilocks::Element *tree)
    {
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
    // extension: wrap caller-owned buffers (host or device) without allocating — used with the resident commit
    MerkleTreeGL(uint64_t _height, uint64_t _width, Goldilocks::Element *_source, Goldilocks::Element *_nodes)
        : height(_height), width(_width), source(_source), nodes(_nodes){};
    MerkleTreeGL(const MerkleTreeGL &) = delete;
    MerkleTreeGL &operator=(const MerkleTreeGL &) = delete;
    ~MerkleTreeGL()
    {
        if (isSourceAllocated) free(source);
        if (isNodesAllocated) free(nodes);
    };
    void copySource(Goldilocks::Element *_source)
    {
        zkgpu_dropin::check(zkgpu_memcpy(source, _source, height * width * sizeof(Goldilocks::Element)), "MerkleTreeGL::copySource");
    }
    void merkelize()
    {
        PoseidonGoldilocks::merkletree(nodes, source, width, height); // merkleTreeGL.cpp:37-44
    }
    uint64_t getTreeNumElements() { return height * HASH_SIZE + (height - 1) * HASH_SIZE; }
    void getRoot(Goldilocks::Element *root)
    {
        zkgpu_dropin::check(zkgpu_memcpy(root, &nodes[getTreeNumElements() - HASH_SIZE], HASH_SIZE * sizeof(Goldilocks::Element)), "MerkleTreeGL::getRoot");
    }
    // proof = row idx of source (width words) followed by the sibling digest of every level (merkleTreeGL.cpp:12-35)
    void getGroupProof(Goldilocks::Element *proof, uint64_t idx)
    {
        zkgpu_dropin::check(zkgpu_group_proof(zkgpu_dropin::words(proof), zkgpu_dropin::words(source), zkgpu_dropin::words(nodes), width, height, idx),
                            "MerkleTreeGL::getGroupProof");
    }
    uint64_t MerkleProofSize()
    {
        uint64_t n = 0;
        for (uint64_t h = height; h > 1; h >>= 1) n++; // ceil(log2(height)) for the power-of-two heights in use
        return n;
    }
};
#endif
