// Shim (see FaissAssert.h): the result-handler plumbing pq4_fast_scan_search_1.cpp calls, reduced to the one handler the
// oracle needs -- a store-to-memory handler (faiss's StoreResultHandler: uint16 distances, row q at data + q * ld).
#pragma once
#include <cstddef>
#include <cstdint>
#include <simdlib_avx2.h>
namespace faiss {
inline bool is_aligned_pointer(const void* p) { return ((uintptr_t)p & 31) == 0; }
struct SIMDResultHandler {
    virtual void handle(size_t q, size_t b, simd16uint16 d0, simd16uint16 d1) = 0;
    virtual void set_block_origin(size_t i0, size_t j0) = 0;
    virtual ~SIMDResultHandler() {}
};
namespace simd_result_handlers {
// distances of NQ queries x BB groups of 16 vectors held until the block is done
template <int NQ, int BB>
struct FixedStorageHandler : SIMDResultHandler {
    simd16uint16 dis[NQ][BB];
    int i0 = 0;
    void handle(size_t q, size_t b, simd16uint16 d0, simd16uint16 d1) final { dis[q + i0][2 * b] = d0; dis[q + i0][2 * b + 1] = d1; }
    void set_block_origin(size_t i0_in, size_t) final { i0 = (int)i0_in; }
    template <class Other>
    void to_other_handler(Other& other) const {
        for (int q = 0; q < NQ; q++)
            for (int b = 0; b < BB; b += 2) other.handle(q, b / 2, dis[q][b], dis[q][b + 1]);
    }
};
struct StoreResultHandler : SIMDResultHandler {
    uint16_t* data;
    size_t ld, i0 = 0, j0 = 0;
    StoreResultHandler(uint16_t* data, size_t ld) : data(data), ld(ld) {}
    void handle(size_t q, size_t b, simd16uint16 d0, simd16uint16 d1) final {
        size_t ofs = (q + i0) * ld + j0 + b * 32;
        d0.storeu(data + ofs);
        d1.storeu(data + ofs + 16);
    }
    void set_block_origin(size_t i0_in, size_t j0_in) final { i0 = i0_in; j0 = j0_in; }
};
template <class Consumer, class... Types>
void dispatch_SIMDResultHandler(SIMDResultHandler& res, Consumer& consumer, Types... args) {
    consumer.template f<StoreResultHandler>(static_cast<StoreResultHandler&>(res), args...);
}
}  // namespace simd_result_handlers
}  // namespace faiss
