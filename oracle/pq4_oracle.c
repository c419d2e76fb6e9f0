/* pq4_oracle.c -- CPU restatement of the reference's PQ4 fast-scan path.  TEST INFRASTRUCTURE ONLY (see lutvq_oracle.c):
 * nothing under rvv-lut-vq_b200/ links or calls this file.
 *
 * Follows /root/reference/benchmark/kernel/faiss-simd (loose copies of faiss, MIT):
 *   pq4_fast_scan.cpp:45-91     pq4_pack_codes            pq4_fast_scan.cpp:93-137   pq4_pack_codes_range
 *   pq4_fast_scan.cpp:144-204   get/set packed element    pq4_fast_scan.cpp:253-281  pq4_pack_LUT / pq4_pack_LUT_qbs
 *   pq4_fast_scan_search_1.cpp:30-117  kernel_accumulate_block  (uint16 lanes, wrap-around arithmetic, optional scaler)
 *   pq4_fast_scan_search_1.cpp:119-171 accumulate_fixed_blocks / the (nq, bbs/32) dispatch table
 * The 256-bit lanes are written out as uint16_t[16] arrays; every step of the kernel (the ">> 8" accumulators, the
 * "-= << 8" fix-up, combine2x2) is kept so that overflow behaviour is the reference's.
 * Parity: PINNED by tests/golden/pq4_*.npz, produced by the reference's own compiled code (oracle/_ref/libpq4_ref.so,
 * built from the files above by `make -C oracle ref`; oracle/gen_golden_pq4.py). */
#include <stdint.h>
#include <stddef.h>
#include <string.h>

static const uint8_t kPerm0[16] = {0, 8, 1, 9, 2, 10, 3, 11, 4, 12, 5, 13, 6, 14, 7, 15};

/* column (i.., j) of the (m, n) byte matrix, zero outside: pq4_fast_scan.cpp:26-41 */
static void matrix_column32(const uint8_t* src, size_t m, size_t n, int64_t i, int64_t j, uint8_t* dest) {
    for (int64_t k = 0; k < 32; k++) dest[k] = (k + i >= 0 && (size_t)(k + i) < m) ? src[(k + i) * n + j] : 0;
}

static void pack_chunk(const uint8_t* c, uint8_t* codes2, int accumulate) {
    uint8_t c0[32], c1[32];
    for (int j = 0; j < 32; j++) { c0[j] = c[j] & 15; c1[j] = c[j] >> 4; }
    for (int j = 0; j < 16; j++) {
        uint8_t d0 = c0[kPerm0[j]] | (uint8_t)(c0[kPerm0[j] + 16] << 4);
        uint8_t d1 = c1[kPerm0[j]] | (uint8_t)(c1[kPerm0[j] + 16] << 4);
        if (accumulate) { codes2[j] |= d0; codes2[j + 16] |= d1; } else { codes2[j] = d0; codes2[j + 16] = d1; }
    }
}

/* returns 1 where the reference throws (bbs % 32, nb % bbs, nsq % 2) */
int orc_pq4_pack_codes(const uint8_t* codes, size_t ntotal, size_t M, size_t nb, size_t bbs, size_t nsq, uint8_t* blocks) {
    if (bbs % 32 != 0 || nb % bbs != 0 || nsq % 2 != 0) return 1;
    if (nb == 0) return 0;
    memset(blocks, 0, nb * nsq / 2);
    uint8_t* codes2 = blocks;
    for (size_t i0 = 0; i0 < nb; i0 += bbs)
        for (size_t sq = 0; sq < nsq; sq += 2)
            for (size_t i = 0; i < bbs; i += 32) {
                uint8_t c[32];
                matrix_column32(codes, ntotal, (M + 1) / 2, (int64_t)(i0 + i), (int64_t)(sq / 2), c);
                pack_chunk(c, codes2, 0);
                codes2 += 32;
            }
    return 0;
}

void orc_pq4_pack_codes_range(const uint8_t* codes, size_t M, size_t i0, size_t i1, size_t bbs, size_t nsq, uint8_t* blocks) {
    size_t block0 = i0 / bbs, block1 = ((i1 - 1) / bbs) + 1;
    for (size_t b = block0; b < block1; b++) {
        uint8_t* codes2 = blocks + b * bbs * nsq / 2;
        int64_t i_base = (int64_t)(b * bbs) - (int64_t)i0;
        for (size_t sq = 0; sq < nsq; sq += 2)
            for (size_t i = 0; i < bbs; i += 32) {
                uint8_t c[32];
                matrix_column32(codes, i1 - i0, (M + 1) / 2, i_base + (int64_t)i, (int64_t)(sq / 2), c);
                pack_chunk(c, codes2, 1);
                codes2 += 32;
            }
    }
}

static size_t vector_address(size_t bbs, size_t vector_id, size_t sq, int* shift) {
    vector_id %= bbs;
    *shift = vector_id > 15;
    vector_id &= 15;
    size_t address = vector_id < 8 ? vector_id << 1 : ((vector_id - 8) << 1) + 1;
    if (sq & 1) address += 16;
    return (sq >> 1) * bbs + address;
}

int orc_pq4_get_packed_element(const uint8_t* data, size_t bbs, size_t nsq, size_t vector_id, size_t sq) {
    data += (vector_id / bbs) * (((nsq + 1) / 2) * bbs);
    int shift;
    size_t a = vector_address(bbs, vector_id, sq, &shift);
    return shift ? data[a] >> 4 : data[a] & 15;
}

void orc_pq4_set_packed_element(uint8_t* data, int code, size_t bbs, size_t nsq, size_t vector_id, size_t sq) {
    data += (vector_id / bbs) * (((nsq + 1) / 2) * bbs);
    int shift;
    size_t a = vector_address(bbs, vector_id, sq, &shift);
    if (shift) data[a] = (uint8_t)((code << 4) | (data[a] & 15));
    else data[a] = (uint8_t)(code | (data[a] & ~15));
}

void orc_pq4_pack_LUT(int nq, int nsq, const uint8_t* src, uint8_t* dest) {
    for (int q = 0; q < nq; q++)
        for (int sq = 0; sq < nsq; sq += 2) {
            memcpy(dest + (size_t)(sq / 2 * nq + q) * 32, src + (size_t)(q * nsq + sq) * 16, 16);
            memcpy(dest + (size_t)(sq / 2 * nq + q) * 32 + 16, src + (size_t)(q * nsq + sq + 1) * 16, 16);
        }
}

int orc_pq4_pack_LUT_qbs(int qbs, int nsq, const uint8_t* src, uint8_t* dest) {
    if (nsq % 2 != 0) return -1;
    size_t dim12 = 16 * (size_t)nsq;
    int i0 = 0;
    for (int qi = qbs; qi; qi >>= 4) {
        int nq = qi & 15;
        orc_pq4_pack_LUT(nq, nsq, src + i0 * dim12, dest + i0 * dim12);
        i0 += nq;
    }
    return i0;
}

int orc_pq4_pack_LUT_qbs_q_map(int qbs, int nsq, const uint8_t* src, const int* q_map, uint8_t* dest) {
    if (nsq % 2 != 0) return -1;
    size_t dim12 = 16 * (size_t)nsq;
    int i0 = 0;
    for (int qi = qbs; qi; qi >>= 4) {
        int nq = qi & 15;
        uint8_t* d = dest + i0 * dim12;
        for (int k = 0; k < nq; k++) {
            int q = q_map[i0 + k];
            for (int sq = 0; sq < nsq; sq += 2) {
                memcpy(d + (size_t)(sq / 2 * nq + k) * 32, src + (size_t)(q * nsq + sq) * 16, 16);
                memcpy(d + (size_t)(sq / 2 * nq + k) * 32 + 16, src + (size_t)(q * nsq + sq + 1) * 16, 16);
            }
        }
        i0 += nq;
    }
    return i0;
}

/* one (query, 32-vector group) of kernel_accumulate_block: accu[4][16] in uint16 lanes */
typedef struct { uint16_t a[4][16]; } Accu;

static void lookup_add(Accu* A, const uint8_t* lut32, const uint8_t* c32, int scale) {
    uint8_t res0[32], res1[32];
    for (int j = 0; j < 32; j++) {                       /* vpshufb works per 128-bit half: bytes 16.. index the second table */
        const uint8_t* t = lut32 + (j / 16) * 16;
        res0[j] = t[c32[j] & 15];
        res1[j] = t[(c32[j] >> 4) & 15];
    }
    for (int l = 0; l < 16; l++) {
        uint16_t w0 = (uint16_t)(res0[2 * l] | (res0[2 * l + 1] << 8)), w1 = (uint16_t)(res1[2 * l] | (res1[2 * l + 1] << 8));
        if (scale < 0) {
            A->a[0][l] += w0; A->a[1][l] += (uint16_t)(w0 >> 8);
            A->a[2][l] += w1; A->a[3][l] += (uint16_t)(w1 >> 8);
        } else {                                          /* NormTableScaler::scale_lo / scale_hi */
            A->a[0][l] += (uint16_t)(w0 * (uint16_t)scale); A->a[1][l] += (uint16_t)((uint16_t)(w0 >> 8) * (uint16_t)scale);
            A->a[2][l] += (uint16_t)(w1 * (uint16_t)scale); A->a[3][l] += (uint16_t)((uint16_t)(w1 >> 8) * (uint16_t)scale);
        }
    }
}

/* returns 1 where the reference throws: misaligned pointers, bbs % 32, nb % bbs, (nq, bbs/32) not in the dispatch table */
int orc_pq4_accumulate_loop(int nq, size_t nb, int bbs, int nsq, const uint8_t* codes, const uint8_t* LUT, uint16_t* accu, int scale) {
    if (((uintptr_t)codes & 31) || ((uintptr_t)LUT & 31) || bbs % 32 != 0 || nb % (size_t)bbs != 0) return 1;
    const int BB = bbs / 32, key = nq * 1000 + BB;
    if (!(key == 1001 || key == 1002 || key == 1003 || key == 1004 || key == 1005 || key == 2001 || key == 2002 || key == 3001 || key == 4001)) return 1;
    const int nscale = scale < 0 ? 0 : 2;
    for (size_t j0 = 0; j0 < nb; j0 += (size_t)bbs) {
        const uint8_t* cb = codes + j0 * (size_t)nsq / 2;
        for (int q = 0; q < nq; q++)
            for (int b = 0; b < BB; b++) {
                Accu A;
                memset(&A, 0, sizeof A);
                for (int sq = 0; sq < nsq; sq += 2)
                    lookup_add(&A, LUT + (size_t)(sq / 2 * nq + q) * 32, cb + (size_t)(sq / 2 * BB + b) * 32, sq < nsq - nscale ? -1 : scale);
                uint16_t dis[32];
                for (int l = 0; l < 16; l++) { A.a[0][l] -= (uint16_t)(A.a[1][l] << 8); A.a[2][l] -= (uint16_t)(A.a[3][l] << 8); }
                for (int l = 0; l < 8; l++) {                /* combine2x2: low half = a.lo + a.hi, high half = b.lo + b.hi */
                    dis[l] = (uint16_t)(A.a[0][l] + A.a[0][l + 8]);
                    dis[8 + l] = (uint16_t)(A.a[1][l] + A.a[1][l + 8]);
                    dis[16 + l] = (uint16_t)(A.a[2][l] + A.a[2][l + 8]);
                    dis[24 + l] = (uint16_t)(A.a[3][l] + A.a[3][l + 8]);
                }
                memcpy(accu + (size_t)q * nb + j0 + (size_t)b * 32, dis, sizeof dis);
            }
    }
    return 0;
}
