// lut_layout.h -- shared-memory LUT slice layout of the fused LUT-GEMV kernel (host + device).
//
// One k-slice covers 64 consecutive code bytes of every row (= 64 lookup tables = 32 groups x 2
// codebooks for AQLM 2x8).  Its fp32 LUT is a 256 x 64 array of words, 64 KiB:
//
//     byte_offset(p, idx) = idx*256 + (w&1)*128 + a*32 + (w>>1)*16 + beta*4
//         p = byte position in the slice (0..63) = 16*a + 4*w + beta     (a: 16-byte chunk, w: word, beta: byte)
//
// so that  bank(p) = 8*a + 4*(w>>1) + beta  depends on the TABLE only (never on idx): a warp whose 32
// lanes look up 32 tables with distinct banks is conflict-free for ANY index pattern, which is what
// lets the gather run at one shared-memory wavefront per 32 code bytes.
//
// Streaming side: a warp tile is 8 rows x 64 B.  lane = 8*a + 4*e + b loads chunk a of row (lane&7)
// as two 8-byte halves, swapped when e=1, and walks its 16 bytes in the order
//     step (c,d), c,d in 0..3:  register word c  (= memory word c^2e),  byte d^b
// => position p(lane,c,d) = 16a + 4(c^2e) + (d^b), whose banks are pairwise distinct across the warp
// for every (c,d).  The full LDS address is formed by ONE PRMT: byte0 <- lane constant
// Lm[c>>1][d] = 4*lane ^ (16*(c>>1) + 4*d), byte1 <- the code byte, plus immediate (c&1)*128.
//
// Build side (mma.sync m16n8k8, M = idx, N = groups): thread (r = lane>>2, q = lane&3) ends up with, for
// each N-tile j in 0..3 and idx row, the four words beta = 2u+n of chunk a=q, word w_j = 2*(j&1) + (j>>1);
// they are one 16-byte store at idx*256 + (j>>1)*128 + 32q + (j&1)*16.  Odd-r lanes store tile j^1 first
// so each quarter-warp phase of the 128-bit store touches 8 distinct 16-byte columns.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define LUTVQ_HD __host__ __device__ __forceinline__
#else
#define LUTVQ_HD inline
#endif

namespace lutvq {

constexpr int kSliceTables = 64;            // code bytes per row per k-slice
constexpr int kIdxStride   = 256;           // bytes between consecutive idx rows of the slice LUT
constexpr int kSliceBytes  = 256 * 256;     // 64 KiB of fp32 entries
constexpr int kTileRows    = 8;             // rows per warp tile
constexpr int kSetRows     = 32;            // rows per warp iteration (4 tiles, transposed-reduced to 1 row/lane)

LUTVQ_HD uint32_t lut_entry_offset(int p, int idx) {
    const int a = p >> 4, w = (p >> 2) & 3, beta = p & 3;
    return (uint32_t)(idx * kIdxStride + (w & 1) * 128 + a * 32 + (w >> 1) * 16 + beta * 4);
}
LUTVQ_HD int lut_bank(int p) { return (int)((lut_entry_offset(p, 0) >> 2) & 31); }

// ---- streaming side ----
LUTVQ_HD int      stream_row_in_tile(int lane)            { return lane & 7; }
LUTVQ_HD int      stream_chunk(int lane)                  { return lane >> 3; }
LUTVQ_HD int      stream_first_half(int lane)             { return (lane >> 2) & 1; }          // e: which 8-byte half is loaded first
LUTVQ_HD int      stream_pos(int lane, int c, int d)      { return 16 * (lane >> 3) + 4 * (c ^ (2 * ((lane >> 2) & 1))) + (d ^ (lane & 3)); }
LUTVQ_HD uint32_t stream_sel(int lane, int d)             { return 0x7604u | (uint32_t)((d ^ (lane & 3)) << 4); }
LUTVQ_HD uint32_t stream_lm(int lane, int h, int d)       { return (uint32_t)((4 * lane) ^ (16 * h + 4 * d)); }
LUTVQ_HD uint32_t stream_imm(int c)                       { return (uint32_t)((c & 1) * 128); }

// ---- build side ----
LUTVQ_HD int      build_word_of_tile(int j)               { return 2 * (j & 1) + (j >> 1); }   // w_j
// local group (0..31) held by column `col` (0..7) of N-tile j, for a format with 2 code bytes per group
LUTVQ_HD int      build_group_of_col(int j, int col)      { return 8 * (col >> 1) + 2 * build_word_of_tile(j) + (col & 1); }
// ---- build side, any number of codebooks per group NCB in {1, 2, 4, 8} (AQLM Kx8) ----
// Byte position p of a slice means (group p / NCB, codebook p % NCB).  The 8 MMAs of an idx block are i = n * NT + t
// (codebook n, N-tile t of 8 groups, NT = 8 / NCB).  The 16-byte store (chunk a = q, word w_j) covers positions
// 16q + 4w_j + beta; the thread owns MMA columns 2q + u.  Which x-group sits in column `col` of tile t, and which MMA
// output is byte beta of store j:
//   NCB = 1: t = 2j + n', group 16(col>>1) + 4w_j + 2n' + (col&1);  beta = 2n' + u   -> MMA 2j + (beta>>1),  u = beta & 1
//   NCB = 2: t = j,       group  8(col>>1) + 2w_j + (col&1);        beta = 2u + n    -> MMA 4(beta&1) + j,   u = beta >> 1
//   NCB = 4: t = k, j = 2k + u, group 4(col>>1) + w_{2k + (col&1)}; beta = n         -> MMA 2beta + (j>>1),  u = j & 1
//   NCB = 8: t = 0,       group col;                  codebook 4(j>>1) + beta        -> MMA 4(j>>1) + beta,  u = j & 1
template <int NCB> LUTVQ_HD int build_group(int t, int col) {
    return NCB == 1 ? 16 * (col >> 1) + 4 * build_word_of_tile(t >> 1) + 2 * (t & 1) + (col & 1)
         : NCB == 2 ? build_group_of_col(t, col)
         : NCB == 4 ? 4 * (col >> 1) + build_word_of_tile(2 * t + (col & 1))
                    : col;
}
template <int NCB> LUTVQ_HD int build_val_mma(int j, int beta) {
    return NCB == 1 ? 2 * j + (beta >> 1) : NCB == 2 ? 4 * (beta & 1) + j : NCB == 4 ? 2 * beta + (j >> 1) : 4 * (j >> 1) + beta;
}
template <int NCB> LUTVQ_HD int build_val_col(int j, int beta) {
    return NCB == 1 ? (beta & 1) : NCB == 2 ? (beta >> 1) : (j & 1);
}
LUTVQ_HD uint32_t build_store_offset(int idx, int j, int q){ return (uint32_t)(idx * kIdxStride + (j >> 1) * 128 + 32 * q + (j & 1) * 16); }

}  // namespace lutvq
