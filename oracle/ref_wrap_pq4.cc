// TEST INFRASTRUCTURE.  Compiles the reference's loose faiss PQ4 fast-scan copies IN PLACE
// (benchmark/kernel/faiss-simd/pq4_fast_scan.cpp + pq4_fast_scan_search_1.cpp + simdlib_avx2.h, found through -I) against
// the small header shims in oracle/faiss_shim/ (the reference vendors no faiss tree), and exports plain-C entry points
// for oracle/gen_golden_pq4.py.  Output: oracle/_ref/libpq4_ref.so (git-ignored).  `make -C oracle ref`.
#include <pq4_fast_scan.cpp>
#include <pq4_fast_scan_search_1.cpp>

extern "C" {
int ref_pq4_pack_codes(const uint8_t* codes, size_t ntotal, size_t M, size_t nb, size_t bbs, size_t nsq, uint8_t* blocks) {
    try { faiss::pq4_pack_codes(codes, ntotal, M, nb, bbs, nsq, blocks); } catch (const std::exception&) { return 1; }
    return 0;
}
int ref_pq4_pack_codes_range(const uint8_t* codes, size_t M, size_t i0, size_t i1, size_t bbs, size_t nsq, uint8_t* blocks) {
    try { faiss::pq4_pack_codes_range(codes, M, i0, i1, bbs, nsq, blocks); } catch (const std::exception&) { return 1; }
    return 0;
}
int ref_pq4_get_packed_element(const uint8_t* data, size_t bbs, size_t nsq, size_t vector_id, size_t sq) {
    return faiss::pq4_get_packed_element(data, bbs, nsq, vector_id, sq);
}
void ref_pq4_set_packed_element(uint8_t* data, int code, size_t bbs, size_t nsq, size_t vector_id, size_t sq) {
    faiss::pq4_set_packed_element(data, (uint8_t)code, bbs, nsq, vector_id, sq);
}
void ref_pq4_pack_LUT(int nq, int nsq, const uint8_t* src, uint8_t* dest) { faiss::pq4_pack_LUT(nq, nsq, src, dest); }
int ref_pq4_pack_LUT_qbs(int qbs, int nsq, const uint8_t* src, uint8_t* dest) {
    try { return faiss::pq4_pack_LUT_qbs(qbs, nsq, src, dest); } catch (const std::exception&) { return -1; }
}
int ref_pq4_pack_LUT_qbs_q_map(int qbs, int nsq, const uint8_t* src, const int* q_map, uint8_t* dest) {
    try { return faiss::pq4_pack_LUT_qbs_q_map(qbs, nsq, src, q_map, dest); } catch (const std::exception&) { return -1; }
}
void ref_pq4_code_packer(size_t nsq, size_t bbs, const uint8_t* flat, size_t offset, uint8_t* block, uint8_t* flat_back, size_t* sizes) {
    faiss::CodePackerPQ4 p(nsq, bbs);
    sizes[0] = p.code_size; sizes[1] = p.nvec; sizes[2] = p.block_size;
    p.pack_1(flat, offset, block);
    p.unpack_1(block, offset, flat_back);
}
// pq4_accumulate_loop into a uint16 [nq][nb] matrix; scale < 0: no scaler (DummyScaler), else NormTableScaler(scale).
// codes / LUT must be 32-byte aligned (the reference checks).  Returns 1 when the reference throws.
int ref_pq4_accumulate_loop(int nq, size_t nb, int bbs, int nsq, const uint8_t* codes, const uint8_t* LUT, uint16_t* accu, int scale) {
    try {
        faiss::simd_result_handlers::StoreResultHandler res(accu, nb);
        if (scale >= 0) {
            faiss::NormTableScaler sc(scale);
            faiss::pq4_accumulate_loop(nq, nb, bbs, nsq, codes, LUT, res, &sc);
        } else {
            faiss::pq4_accumulate_loop(nq, nb, bbs, nsq, codes, LUT, res, nullptr);
        }
    } catch (const std::exception&) { return 1; }
    return 0;
}
}
