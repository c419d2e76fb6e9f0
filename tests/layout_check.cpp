// layout_check.cpp -- host-side proof of the shared-memory LUT layout used by gemv_lut.cu (round-2 layout: 128 u16 tables).
// Built and run by tests/test_host_logic.py (g++, no GPU).  Exhaustively checks, with the SAME inline
// functions the kernel uses (rvv-lut-vq_b200/csrc/lut_layout.h):
//   1. the PRMT-composed LDS address of every (lane, half, image byte, idx) is the entry of the table that byte of the
//      prepacked image belongs to (prepack permutation + kernel addressing agree), every lane visits each of its 32
//      code bytes exactly once, and the 4 lanes of a row cover its 128 tables exactly once;
//   2. the 32 lanes of every lookup step hit 32 distinct banks (conflict-free for any index pattern);
//   3. the build side (mma.sync C fragments of both codebooks -> packed u16 pairs -> 128-bit stores) writes every
//      (table, idx) entry exactly once, to the address the lookups read, and each quarter-warp phase of every store
//      touches 8 distinct 16-byte columns;
//   4. the transposed warp reduction leaves row `lane` of the 32-row set in lane `lane`;
//   5. the split-k word: 63 slices of maximal partials fit the 29-bit fields, the arrival count survives.
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <set>
#include <vector>
#include "../rvv-lut-vq_b200/csrc/lut_layout.h"

using namespace lutvq;

static uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {   // PTX prmt.b32 default mode, no sign replication used
    uint8_t src[8];
    for (int i = 0; i < 4; ++i) { src[i] = (a >> (8 * i)) & 0xFF; src[4 + i] = (b >> (8 * i)) & 0xFF; }
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) r |= (uint32_t)src[(sel >> (4 * i)) & 7] << (8 * i);
    return r;
}

#define CHECK(cond, ...) do { if (!(cond)) { printf("FAIL %s:%d: ", __FILE__, __LINE__); printf(__VA_ARGS__); printf("\n"); return 1; } } while (0)

int main() {
    // ---- 1 + 2: streaming side.  A tile row = 128 code bytes with distinct values so that idx identifies the table. ----
    for (int p = 0; p < 2; ++p)
        for (int k = 0; k < 16; ++k) {
            std::set<int> banks;
            for (int lane = 0; lane < 32; ++lane) {
                // the row's slice bytes in memory order: code(table) = (7 * table + 3 * row + 1) & 255
                const int row = stream_row_in_tile(lane);
                auto code_of = [&](int table) { return (7 * table + 3 * row + 1) & 255; };
                const int chunk = stream_chunk(lane, p);
                uint32_t in[4], img[4];
                for (int w = 0; w < 4; ++w) {
                    in[w] = 0;
                    for (int b = 0; b < 4; ++b) in[w] |= (uint32_t)code_of(16 * chunk + 4 * w + b) << (8 * b);
                }
                pack_lane_image(in, lane, img);
                const int j = k >> 2, b = k & 3, u = k >> 1, h = k & 1;
                const uint32_t addr = prmt(img[j], stream_lm(lane, u), stream_sel(b)) + stream_imm(p, h);
                const int table = stream_table(lane, p, k);
                CHECK(table >= 16 * chunk && table < 16 * chunk + 16, "table out of chunk lane %d", lane);
                const int idx = code_of(table);
                CHECK(addr == lut_entry_offset(table, idx), "addr mismatch lane %d p %d k %d: %u vs %u", lane, p, k, addr, lut_entry_offset(table, idx));
                CHECK(addr < (uint32_t)kSliceBytes, "addr out of the slice");
                banks.insert((addr >> 2) & 31);
                CHECK(((addr >> 2) & 31) == (uint32_t)lut_bank(table), "bank function");
            }
            CHECK(banks.size() == 32, "bank conflict at step p=%d k=%d: %zu banks", p, k, banks.size());
        }
    for (int row = 0; row < 8; ++row) {
        std::set<int> seen;
        for (int a = 0; a < 4; ++a) {
            const int lane = 8 * a + row;
            std::set<int> mine;
            for (int p = 0; p < 2; ++p) for (int k = 0; k < 16; ++k) mine.insert(stream_table(lane, p, k));
            CHECK(mine.size() == 32, "lane %d covers %zu tables", lane, mine.size());
            seen.insert(mine.begin(), mine.end());
        }
        CHECK(seen.size() == 128 && *seen.begin() == 0 && *seen.rbegin() == 127, "row %d: %zu tables", row, seen.size());
    }
    // packed offsets of one item are a bijection onto [0, 4096)
    {
        std::set<uint32_t> offs;
        for (int t = 0; t < 4; ++t) for (int p = 0; p < 2; ++p) for (int lane = 0; lane < 32; ++lane) for (int k = 0; k < 16; ++k) offs.insert(packed_offset(t, p, lane, k));
        CHECK(offs.size() == (size_t)kSetBytes && *offs.rbegin() == (uint32_t)kSetBytes - 1, "packed offsets");
    }

    // ---- 3: build side ----
    std::vector<int> written(kSliceBytes / 2, 0);                  // per u16 entry
    for (int m = 0; m < 16; ++m)                                   // idx block
        for (int hb = 0; hb < 2; ++hb)
            for (int hh = 0; hh < 2; ++hh)
                for (int which = 0; which < 2; ++which) {          // first / second store instruction
                    for (int phase = 0; phase < 4; ++phase) {      // quarter-warp phases of a 128-bit store
                        std::set<uint32_t> cols;
                        for (int l8 = 0; l8 < 8; ++l8) {
                            const int lane = 8 * phase + l8, r = lane >> 2, q = lane & 3;
                            const bool odd = r & 1;
                            const int pair_local = which == 0 ? (odd ? 1 : 0) : (odd ? 0 : 1);
                            const int idx = 16 * m + r + 8 * hh;
                            const uint32_t off = build_store_offset(idx, 2 * hb + pair_local, q);
                            CHECK(off % 16 == 0 && off + 16 <= (uint32_t)kSliceBytes, "store offset");
                            cols.insert((off % 128) / 16);
                            for (int e = 0; e < 2; ++e)
                                for (int f = 0; f < 2; ++f) {
                                    // word 2e+f of the store = C fragment element (row r + 8hh, column 2q+f) of N-tile 4hb + 2*pair_local + e
                                    const int group = build_group_of_col(4 * hb + 2 * pair_local + e, 2 * q + f);
                                    CHECK(group >= 0 && group < kSliceGroups, "group range");
                                    const uint32_t word_off = off + 4 * (2 * e + f);
                                    for (int n = 0; n < 2; ++n) {                      // codebook n -> half n of the word
                                        CHECK(word_off + 2 * n == lut_entry_offset(2 * group + n, idx), "build/lookup disagree idx %d group %d n %d", idx, group, n);
                                        written[(word_off + 2 * n) / 2] += 1;
                                    }
                                }
                        }
                        CHECK(cols.size() == 8, "store phase conflict m=%d hb=%d hh=%d which=%d phase=%d: %zu columns", m, hb, hh, which, phase, cols.size());
                    }
                }
    for (size_t i = 0; i < written.size(); ++i) CHECK(written[i] == 1, "entry %zu written %d times", i, written[i]);
    {   // the B fragment columns cover the slice's 64 groups exactly once
        std::set<int> groups;
        for (int j = 0; j < 8; ++j) for (int col = 0; col < 8; ++col) groups.insert(build_group_of_col(j, col));
        CHECK(groups.size() == 64 && *groups.begin() == 0 && *groups.rbegin() == 63, "B fragment columns");
    }

    // ---- 4: transposed reduction.  acc[t] of lane (a, r) = partial of tile t, row r over chunks a, a+4 ----
    {
        uint32_t acc[32][4], k0[32], k1[32], g0[32], g1[32], keep[32], give[32];
        for (int lane = 0; lane < 32; ++lane)
            for (int t = 0; t < 4; ++t) acc[lane][t] = 1000u * t + 10u * (lane & 7) + (1u << (16 + (lane >> 3)));   // distinct per (t, r), a marks the bit
        for (int lane = 0; lane < 32; ++lane) {
            const bool b4 = (lane >> 4) & 1;
            k0[lane] = b4 ? acc[lane][2] : acc[lane][0]; k1[lane] = b4 ? acc[lane][3] : acc[lane][1];
            g0[lane] = b4 ? acc[lane][0] : acc[lane][2]; g1[lane] = b4 ? acc[lane][1] : acc[lane][3];
        }
        uint32_t k0b[32], k1b[32];
        for (int lane = 0; lane < 32; ++lane) { k0b[lane] = k0[lane] + g0[lane ^ 16]; k1b[lane] = k1[lane] + g1[lane ^ 16]; }
        for (int lane = 0; lane < 32; ++lane) {
            const bool b3 = (lane >> 3) & 1;
            keep[lane] = b3 ? k1b[lane] : k0b[lane]; give[lane] = b3 ? k0b[lane] : k1b[lane];
        }
        for (int lane = 0; lane < 32; ++lane) {
            const uint32_t total = keep[lane] + give[lane ^ 8];
            const int t = lane >> 3, r = lane & 7;                  // row `lane` of the set = tile t, row r
            const uint32_t want = 4u * (1000u * t + 10u * r) + (1u << 16) + (1u << 17) + (1u << 18) + (1u << 19);
            CHECK(total == want, "reduction lane %d: %u vs %u", lane, total, want);
        }
    }

    // ---- 5: split-k word ----
    {
        const unsigned long long maxpart = 128ull * 65535ull;       // a row's partial over one slice
        unsigned long long word = 0;
        for (int s = 0; s < 63; ++s) word += maxpart + (maxpart << 29) + (1ull << 58);
        CHECK((word >> 58) == 63, "arrival count");
        CHECK((word & ((1ull << 29) - 1)) == 63 * maxpart && ((word >> 29) & ((1ull << 29) - 1)) == 63 * maxpart, "fields overflow");
    }
    printf("layout OK\n");
    return 0;
}
