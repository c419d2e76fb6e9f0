// Shim (see FaissAssert.h): the two scalers pq4_fast_scan_search_1.cpp is instantiated with, restated from faiss's
// published interface (faiss/impl/LookupTableScaler.h, v1.7.4): DummyScaler has nscale = 0 and is never looked up;
// NormTableScaler multiplies the last nscale = 2 sub-quantizers' table values by an integer, in uint16 lanes.
#pragma once
#include <faiss/impl/FaissAssert.h>
#include <simdlib_avx2.h>
namespace faiss {
struct DummyScaler {
    static constexpr int nscale = 0;
    simd32uint8 lookup(const simd32uint8&, const simd32uint8&) const { FAISS_THROW_MSG("DummyScaler::lookup"); }
    simd16uint16 scale_lo(const simd32uint8&) const { FAISS_THROW_MSG("DummyScaler::scale_lo"); }
    simd16uint16 scale_hi(const simd32uint8&) const { FAISS_THROW_MSG("DummyScaler::scale_hi"); }
};
struct NormTableScaler {
    static constexpr int nscale = 2;
    int scale_int;
    simd16uint16 scale_simd;
    explicit NormTableScaler(int scale) : scale_int(scale), scale_simd(scale) {}
    simd32uint8 lookup(const simd32uint8& lut, const simd32uint8& c) const { return lut.lookup_2_lanes(c); }
    simd16uint16 scale_lo(const simd32uint8& res) const { return simd16uint16(res) * scale_simd; }
    simd16uint16 scale_hi(const simd32uint8& res) const { return (simd16uint16(res) >> 8) * scale_simd; }
};
}  // namespace faiss
