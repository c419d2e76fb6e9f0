// Shim (see FaissAssert.h): the packing interface pq4_fast_scan.h derives CodePackerPQ4 from.
#pragma once
#include <cstddef>
#include <cstdint>
namespace faiss {
struct CodePacker {
    size_t code_size = 0;    // bytes of one flat code
    size_t nvec = 0;         // vectors per block
    size_t block_size = 0;   // bytes of one block
    virtual void pack_1(const uint8_t* flat_code, size_t offset, uint8_t* block) const = 0;
    virtual void unpack_1(const uint8_t* block, size_t offset, uint8_t* flat_code) const = 0;
    virtual ~CodePacker() {}
};
}  // namespace faiss
