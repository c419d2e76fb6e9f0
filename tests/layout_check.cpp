// layout_check.cpp -- host-side proof of the shared-memory LUT layout used by gemv_lut.cu.
// Built and run by tests/test_host_logic.py (g++, no GPU).  Exhaustively checks, with the SAME inline
// functions the kernel uses (rvv-lut-vq_b200/csrc/lut_layout.h):
//   1. the PRMT-composed LDS address of every (lane, step, idx) is the entry of the table the lane's code
//      byte belongs to;
//   2. the 32 lanes of every lookup step hit 32 distinct banks (conflict-free for any index pattern);
//   3. the build side (mma.sync C fragments -> 128-bit stores) writes every (table, idx) entry exactly once,
//      to the address the lookups read, and each quarter-warp phase of every store touches 8 distinct
//      16-byte columns;
//   4. the transposed warp reduction leaves row `lane` of the 32-row set in lane `lane`.
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

// ---- 3b: build side for NCB codebooks per group (AQLM Kx8).  Simulates the kernel's 8 MMAs per idx block symbolically:
// MMA i = n * NT + t yields, in thread (r, q), element (h, u) = entry of (idx 16m + r + 8h, codebook n, group of column 2q+u of
// tile t); the store code picks build_val_mma / build_val_col.  Every (position, idx) must be written once, with the entry of
// (group p / NCB, codebook p % NCB).
template <int NCB>
static int check_build() {
    constexpr int NT = 8 / NCB;
    std::vector<int> written(kSliceBytes / 4, -1);
    for (int m = 0; m < 16; ++m)
        for (int h = 0; h < 2; ++h)
            for (int jp = 0; jp < 2; ++jp)
                for (int which = 0; which < 2; ++which)
                    for (int lane = 0; lane < 32; ++lane) {
                        const int r = lane >> 2, q = lane & 3, odd = r & 1;
                        const int j0 = 2 * jp, j1 = 2 * jp + 1;
                        const int j = (which == 0) ? (odd ? j1 : j0) : (odd ? j0 : j1);
                        const int idx = 16 * m + r + 8 * h;
                        const uint32_t off = build_store_offset(idx, j, q);
                        for (int beta = 0; beta < 4; ++beta) {
                            const int i = build_val_mma<NCB>(j, beta), u = build_val_col<NCB>(j, beta);
                            CHECK(i >= 0 && i < 8 && (u == 0 || u == 1), "NCB=%d: bad MMA index", NCB);
                            const int n = i / NT, t = i % NT;                               // codebook, N-tile of MMA i
                            const int group = build_group<NCB>(t, 2 * q + u);
                            CHECK(group >= 0 && group < 64 / NCB, "NCB=%d: group %d out of the slice", NCB, group);
                            const int p = group * NCB + n;
                            CHECK(off + 4 * beta == lut_entry_offset(p, idx), "NCB=%d: build offset mismatch (j=%d beta=%d q=%d)", NCB, j, beta, q);
                            CHECK(written[(off >> 2) + beta] == -1, "NCB=%d: entry written twice", NCB);
                            written[(off >> 2) + beta] = p * 256 + idx;
                        }
                    }
    for (int p = 0; p < 64; ++p)
        for (int idx = 0; idx < 256; ++idx)
            CHECK(written[lut_entry_offset(p, idx) >> 2] == p * 256 + idx, "NCB=%d: entry (p=%d, idx=%d) not built", NCB, p, idx);
    std::set<int> groups;                                           // the NT tiles cover the slice's groups exactly once
    for (int t = 0; t < NT; ++t) for (int col = 0; col < 8; ++col) groups.insert(build_group<NCB>(t, col));
    CHECK((int)groups.size() == 64 / NCB && *groups.rbegin() == 64 / NCB - 1, "NCB=%d: N-tiles do not cover the slice groups", NCB);
    return 0;
}

int main() {
    if (check_build<1>() || check_build<2>() || check_build<4>() || check_build<8>()) return 1;
    // ---- 1 + 2: streaming side ----
    for (int c = 0; c < 4; ++c)
        for (int d = 0; d < 4; ++d) {
            std::set<int> banks;
            for (int lane = 0; lane < 32; ++lane) {
                const int e = stream_first_half(lane);
                const int p = stream_pos(lane, c, d);
                // the lane's 16 bytes: memory chunk stream_chunk(lane); registers w0,w1 = first loaded half (half e), w2,w3 = other half
                uint8_t mem[16];
                for (int i = 0; i < 16; ++i) mem[i] = (uint8_t)(17 * i + 3 * lane + 1);     // arbitrary distinct indices
                uint32_t w[4];
                for (int rw = 0; rw < 4; ++rw) {
                    const int mw = rw ^ (2 * e);                                          // register word rw holds memory word rw^2e
                    w[rw] = mem[4 * mw] | (mem[4 * mw + 1] << 8) | (mem[4 * mw + 2] << 16) | ((uint32_t)mem[4 * mw + 3] << 24);
                }
                const uint32_t addr = prmt(w[c], stream_lm(lane, c >> 1, d), stream_sel(lane, d)) + stream_imm(c);
                const int byte_in_chunk = p - 16 * stream_chunk(lane);
                CHECK(byte_in_chunk >= 0 && byte_in_chunk < 16, "pos out of chunk lane %d", lane);
                const int idx = mem[byte_in_chunk];
                CHECK(addr == lut_entry_offset(p, idx), "addr mismatch lane %d c %d d %d: %u vs %u", lane, c, d, addr, lut_entry_offset(p, idx));
                banks.insert((addr >> 2) & 31);
            }
            CHECK(banks.size() == 32, "bank conflict at step c=%d d=%d: %zu banks", c, d, banks.size());
        }
    // every lane visits each of its 16 bytes exactly once
    for (int lane = 0; lane < 32; ++lane) {
        std::set<int> seen;
        for (int c = 0; c < 4; ++c) for (int d = 0; d < 4; ++d) seen.insert(stream_pos(lane, c, d));
        CHECK(seen.size() == 16, "lane %d covers %zu bytes", lane, seen.size());
        CHECK(*seen.begin() == 16 * stream_chunk(lane) && *seen.rbegin() == 16 * stream_chunk(lane) + 15, "lane %d range", lane);
    }

    // ---- 3: build side.  written[offset/4] = (p, idx) tag ----
    std::vector<int> written(kSliceBytes / 4, -1);
    for (int m = 0; m < 16; ++m)                 // idx blocks of 16
        for (int h = 0; h < 2; ++h)
            for (int jp = 0; jp < 2; ++jp)
                for (int which = 0; which < 2; ++which) {            // first / second store of the pair
                    for (int phase = 0; phase < 4; ++phase) {
                        std::set<uint32_t> cols;
                        for (int lane = 8 * phase; lane < 8 * phase + 8; ++lane) {
                            const int r = lane >> 2, q = lane & 3, odd = r & 1;
                            const int j0 = 2 * jp, j1 = 2 * jp + 1;
                            const int j = (which == 0) ? (odd ? j1 : j0) : (odd ? j0 : j1);
                            const int idx = 16 * m + r + 8 * h;
                            const uint32_t off = build_store_offset(idx, j, q);
                            cols.insert((off >> 4) & 7);
                            for (int beta = 0; beta < 4; ++beta) {
                                const int u = beta >> 1, n = beta & 1;
                                // C fragment: thread (r,q) holds columns 2q+u of N-tile j  -> local group
                                const int group = build_group_of_col(j, 2 * q + u);
                                const int p = 2 * group + n;                 // byte position of table (group, codebook n)
                                CHECK(off + 4 * beta == lut_entry_offset(p, idx), "build offset mismatch");
                                CHECK(written[(off >> 2) + beta] == -1, "entry written twice");
                                written[(off >> 2) + beta] = p * 256 + idx;
                            }
                        }
                        CHECK(cols.size() == 8, "store phase conflict: %zu columns", cols.size());
                    }
                }
    for (int p = 0; p < 64; ++p)
        for (int idx = 0; idx < 256; ++idx)
            CHECK(written[lut_entry_offset(p, idx) >> 2] == p * 256 + idx, "entry (p=%d, idx=%d) not built", p, idx);
    // B-fragment columns of the 4 N-tiles cover the 32 local groups exactly once
    {
        std::set<int> groups;
        for (int j = 0; j < 4; ++j) for (int col = 0; col < 8; ++col) groups.insert(build_group_of_col(j, col));
        CHECK(groups.size() == 32 && *groups.rbegin() == 31, "N-tiles do not cover the slice groups");
    }

    // ---- 4: transposed reduction ----
    {
        // value held by lane L for tile t before the reduction: partial sum of row 8t + (L&7) over chunk L>>3
        // encode as: contribution id = (row, chunk); result must be the set of 4 chunks of row == lane
        std::vector<std::vector<std::set<int>>> acc(32, std::vector<std::set<int>>(4));
        for (int L = 0; L < 32; ++L) for (int t = 0; t < 4; ++t) acc[L][t].insert((8 * t + (L & 7)) * 4 + (L >> 3));
        std::vector<std::set<int>> k0(32), k1(32), g0(32), g1(32), keep(32), give(32);
        for (int L = 0; L < 32; ++L) {
            const bool b4 = (L >> 4) & 1;
            k0[L] = b4 ? acc[L][2] : acc[L][0]; k1[L] = b4 ? acc[L][3] : acc[L][1];
            g0[L] = b4 ? acc[L][0] : acc[L][2]; g1[L] = b4 ? acc[L][1] : acc[L][3];
        }
        for (int L = 0; L < 32; ++L) { auto a = g0[L ^ 16], b = g1[L ^ 16]; k0[L].insert(a.begin(), a.end()); k1[L].insert(b.begin(), b.end()); }
        for (int L = 0; L < 32; ++L) { const bool b3 = (L >> 3) & 1; keep[L] = b3 ? k1[L] : k0[L]; give[L] = b3 ? k0[L] : k1[L]; }
        std::vector<std::set<int>> fin(32);
        for (int L = 0; L < 32; ++L) { fin[L] = keep[L]; auto a = give[L ^ 8]; fin[L].insert(a.begin(), a.end()); }
        for (int L = 0; L < 32; ++L) {
            CHECK(fin[L].size() == 4, "lane %d reduced %zu terms", L, fin[L].size());
            for (int v : fin[L]) CHECK(v / 4 == L, "lane %d holds a term of row %d", L, v / 4);
        }
    }
    printf("layout OK\n");
    return 0;
}
