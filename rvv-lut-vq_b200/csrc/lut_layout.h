// lut_layout.h -- shared-memory LUT slice layout of the fused LUT-GEMV kernel (host + device), round-2 design.
//
// One k-slice covers 128 consecutive code bytes of every row (= 128 lookup tables = 64 groups x 2 codebooks for
// AQLM 2x8).  Its LUT holds 16-bit fixed-point entries, one 256-byte row per index value:
//
//     byte_offset(t, idx) = idx*256 + 2*t          t = table = byte position in the slice (0..127), 64 KiB in all
//
// so that  bank(t) = (t >> 1) & 31  depends on the TABLE only (never on idx): a warp whose 32 lanes look up 32 tables with
// distinct banks is conflict-free for ANY index pattern, and the gather runs at one shared-memory wavefront per 32 code
// bytes.  Entries are u16 = round(L * S) + 32768 (S: one scale per GEMV, gemv_lut.cu); the 32-bit word at
// idx*256 + 4*g holds {codebook 0, codebook 1} of local group g.
//
// Streaming side: a warp tile is 8 rows x 128 B.  lane = 8*a + r handles row r of the tile and the two 16-byte chunks
// c = a and c = a + 4 of that row ("halves" p = 0, 1).  The prepacked code image stores, for each lane and half, the 16
// code bytes in the order the lane walks them:
//     image byte k = 2u + h (u = 0..7, h = 0..1)  <->  table  16c + 2*(u ^ r) + h
// => at step (p, u, h) the 32 lanes hit banks 8a + (u ^ r): pairwise distinct.  The full LDS address is formed by ONE
// PRMT: byte0 <- lane constant Lm[u] = 32a + 4*(u ^ r), byte1 <- the code byte, plus the immediate 128p + 2h.
// (The XOR swizzle lives in the packed image, so every PRMT selector is a compile-time constant.)
//
// Packed image:  [k-slice][32-row set][tile 0..3][half p][lane][16 B]   (tile = 1 KiB, set = 4 KiB, contiguous per worker)
//
// Build side (mma.sync m16n8k8, M = idx, N = groups): column `col` of N-tile j is local group
//     16*(j>>1) + 4*(col>>1) + 2*(j&1) + (col&1)
// so thread (r = lane>>2, q = lane&3) ends up with, for each tile pair i = j>>1 and idx row, the four consecutive groups
// 16i + 4q + {0,1,2,3} -> one 16-byte store at idx*256 + 64i + 16q.  Odd-r lanes store pair i^1 first so that each
// quarter-warp phase of the 128-bit store touches 8 distinct 16-byte columns.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define LUTVQ_HD __host__ __device__ __forceinline__
#else
#define LUTVQ_HD inline
#endif

namespace lutvq {

constexpr int kSliceTables = 128;           // code bytes per row per k-slice
constexpr int kSliceGroups = 64;            // groups of 8 activations per k-slice (2 code bytes each)
constexpr int kIdxStride   = 256;           // bytes between consecutive idx rows of the slice LUT
constexpr int kSliceBytes  = 256 * 256;     // 64 KiB of u16 entries
constexpr int kTileRows    = 8;             // rows per warp tile
constexpr int kSetRows     = 32;            // rows per warp iteration (4 tiles, transposed-reduced to 1 row/lane)
constexpr int kTileBytes   = 1024;          // packed bytes of one tile ([half][lane][16])
constexpr int kSetBytes    = 4096;          // packed bytes of one (k-slice, 32-row set) item

LUTVQ_HD uint32_t lut_entry_offset(int t, int idx) { return (uint32_t)(idx * kIdxStride + 2 * t); }
LUTVQ_HD int lut_bank(int t) { return (t >> 1) & 31; }

// ---- streaming side ----
LUTVQ_HD int      stream_row_in_tile(int lane)            { return lane & 7; }
LUTVQ_HD int      stream_chunk(int lane, int p)           { return (lane >> 3) + 4 * p; }
// table looked up by `lane` with byte k (0..15) of its half-p image
LUTVQ_HD int      stream_table(int lane, int p, int k)    { return 16 * stream_chunk(lane, p) + 2 * ((k >> 1) ^ (lane & 7)) + (k & 1); }
LUTVQ_HD uint32_t stream_lm(int lane, int u)              { return (uint32_t)(32 * (lane >> 3) + 4 * (u ^ (lane & 7))); }
LUTVQ_HD uint32_t stream_imm(int p, int h)                { return (uint32_t)(128 * p + 2 * h); }
LUTVQ_HD uint32_t stream_sel(int byte_in_word)            { return 0x6504u | (uint32_t)(byte_in_word << 4); }   // byte0 <- Lm, byte1 <- code byte
// byte offset inside the packed image of (tile t, half p, lane, byte k) of an item
LUTVQ_HD uint32_t packed_offset(int t, int p, int lane, int k) { return (uint32_t)(t * kTileBytes + p * 512 + lane * 16 + k); }

// prepack: the 16 chunk bytes of (lane, half) in memory order (4 words) -> the lane's image (4 words): image 16-bit pair u
// holds chunk pair u ^ r
LUTVQ_HD void     pack_lane_image(const uint32_t (&in)[4], int lane, uint32_t (&out)[4]) {
    const int rr = lane & 7;
    for (int j = 0; j < 4; ++j) {
        const int u0 = (2 * j) ^ rr, u1 = (2 * j + 1) ^ rr;
        const uint32_t lo16 = (in[u0 >> 1] >> (16 * (u0 & 1))) & 0xFFFFu;
        const uint32_t hi16 = (in[u1 >> 1] >> (16 * (u1 & 1))) & 0xFFFFu;
        out[j] = lo16 | (hi16 << 16);
    }
}

// ---- build side ----
// local group (0..63) held by column `col` (0..7) of N-tile j (0..7)
LUTVQ_HD int      build_group_of_col(int j, int col)      { return 16 * (j >> 1) + 4 * (col >> 1) + 2 * (j & 1) + (col & 1); }
// 16-byte store of tile pair i (tiles 2i, 2i+1) by thread q for index row idx: groups 16i + 4q .. +3
LUTVQ_HD uint32_t build_store_offset(int idx, int i, int q){ return (uint32_t)(idx * kIdxStride + 64 * i + 16 * q); }

}  // namespace lutvq
