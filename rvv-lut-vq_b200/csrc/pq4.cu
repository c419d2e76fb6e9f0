// pq4.cu -- PQ4 fast-scan distance accumulation: 4-bit codes, 16-entry uint8 tables, uint16 sums.
// Replaces benchmark/kernel/faiss-simd/pq4_fast_scan_search_1.cpp:30-117 (kernel_accumulate_block: two 16-entry vpshufb
// tables per 32-byte code chunk, uint16 accumulators, optional NormTableScaler on the last two sub-quantizers) and the
// packing helpers of pq4_fast_scan.cpp:45-137,253-322.  Same packed layouts, so images packed by either side are
// interchangeable; results are the reference's bit for bit (sums modulo 2^16).
//
// Kernel idea: PRMT is the GPU's vpshufb.  A 16-entry byte table lives in 4 registers; one PRMT picks 4 bytes out of 8 by
// four 3-bit selectors, so  lo = prmt(T0, T1, s & 0x7777), hi = prmt(T2, T3, s & 0x7777), r = prmt(lo, hi, 0x3210 | bit3 << 2)
// does FOUR table lookups with 16 bits of packed codes as the selector -- no shared-memory lookups, no bank conflicts.
// The four result bytes are accumulated the way the reference accumulates its uint16 lanes: R += r (as one integer) and
// H += the odd bytes; R - (H << 8) then holds the even-byte sums, H the odd-byte sums, two uint16 per register.
// Thread mapping: 4 lanes per group of 32 database vectors; lane s loads 8 of the chunk's 32 bytes (lanes 0,1: the half that
// indexes sub-quantizer 2p, lanes 2,3: the half of 2p+1), so every 32-byte sector is read exactly once.
// Bound: HBM (nb * nsq / 2 code bytes in, nq * nb * 2 bytes out); ~2.2 issue slots per lookup.
#include "common.cuh"

#include <type_traits>

namespace lutvq {

constexpr int kPq4Threads = 256;
constexpr int kPq4Chunk = 8;                 // pairs per lane and software-pipeline stage
constexpr int kPq4SegPairs = 256;            // pairs per segment: 256 * 255 < 2^16, so the packed sums cannot carry

__device__ __forceinline__ uint32_t pq4_lut4(const uint4& T, uint32_t s, uint32_t t) {       // t = s >> 1
    const uint32_t m = s & 0x7777u;
    return __byte_perm(__byte_perm(T.x, T.y, m), __byte_perm(T.z, T.w, m), (t & 0x4444u) | 0x3210u);
}
__device__ __forceinline__ uint32_t add16x2(uint32_t a, uint32_t b) {                       // two independent uint16 adds
    return (((a & 0xffffu) + (b & 0xffffu)) & 0xffffu) | (((a >> 16) + (b >> 16)) << 16);
}
__device__ __forceinline__ uint32_t mul16x2(uint32_t a, uint32_t k) {
    return (((a & 0xffffu) * k) & 0xffffu) | (((a >> 16) * k) << 16);
}

// accu[q * ld + j] = sum_sq LUT[q][sq][code(j, sq)] (mod 2^16), the last two sub-quantizers times `scale` when scale >= 0.
// codes: blocks of bbs vectors, chunk (p, b) of a block at (p * BB + b) * 32; LUT: (p * NQ + q) * 32 (pq4_pack_LUT).
template <int NQ>
__global__ void __launch_bounds__(kPq4Threads) pq4_accumulate_kernel(size_t n_groups, int BB, int nsq, const uint8_t* __restrict__ codes,
                                                                       const uint8_t* __restrict__ LUT, uint16_t* __restrict__ accu,
                                                                       size_t ld, int scale) {
    extern __shared__ __align__(16) uint8_t lut_s[];
    const int npairs = nsq / 2;
    for (int i = threadIdx.x; i < npairs * NQ * 2; i += kPq4Threads)
        reinterpret_cast<uint4*>(lut_s)[i] = __ldg(reinterpret_cast<const uint4*>(LUT) + i);
    __syncthreads();
    const int sub = threadIdx.x & 3, half = sub >> 1;                 // half 0: table of sq 2p, half 1: table of sq 2p + 1
    const int n_plain = scale >= 0 ? npairs - 1 : npairs;
    const int n_chunks = (n_plain + kPq4Chunk - 1) / kPq4Chunk;
    const size_t pstride = (size_t)BB * 32, gstride = (size_t)gridDim.x * (kPq4Threads / 4);
    auto base_of = [&](size_t g_) {
        const unsigned gi = (unsigned)g_, blk = BB == 1 ? gi : gi / (unsigned)BB, b = gi - blk * (unsigned)BB;   // n_groups < 2^32 (host check)
        return codes + (size_t)blk * ((size_t)BB * 32 * npairs) + b * 32 + sub * 8;
    };
    // One chunk = kPq4Chunk pairs of one group = kPq4Chunk 8-byte loads per lane.  The NEXT chunk (of this group or of the
    // lane's next group) is requested before the current one is looked up, so a lane always has 64..128 bytes in flight:
    // with load-wait-compute the kernel ran at 0.42 of the HBM peak, bound by latency (6 warps x 32 B per SMSP per round trip).
    auto load_chunk = [&](uint2 (&buf)[kPq4Chunk], const uint8_t* base_, int c_) {
        #pragma unroll
        for (int u = 0; u < kPq4Chunk; ++u) {
            const int p = c_ * kPq4Chunk + u;
            buf[u] = p < n_plain ? __ldg(reinterpret_cast<const uint2*>(base_ + p * pstride)) : make_uint2(0u, 0u);
        }
    };
    size_t g = (size_t)blockIdx.x * (kPq4Threads / 4) + (threadIdx.x >> 2);
    if (g >= n_groups) return;
    const uint8_t* base = base_of(g);
    uint2 cur[kPq4Chunk], nxt[kPq4Chunk];
    if (n_chunks > 0) load_chunk(cur, base, 0);
    for (;;) {
        const size_t g_next = g + gstride;
        const uint8_t* base_next = g_next < n_groups ? base_of(g_next) : nullptr;
        // F[q][k]: finished uint16 pairs; k = 2 * word + piece for the even result bytes, 4 + that for the odd ones
        uint32_t F[NQ][8];
        #pragma unroll
        for (int q = 0; q < NQ; ++q)
            #pragma unroll
            for (int k = 0; k < 8; ++k) F[q][k] = 0u;
        uint32_t R[NQ][4], H[NQ][4];
        #pragma unroll
        for (int q = 0; q < NQ; ++q)
            #pragma unroll
            for (int k = 0; k < 4; ++k) R[q][k] = H[q][k] = 0u;
        for (int c = 0; c < n_chunks; ++c) {
            const bool last = c + 1 == n_chunks;
            if (!last) load_chunk(nxt, base, c + 1);
            else if (base_next != nullptr) load_chunk(nxt, base_next, 0);
            #pragma unroll
            for (int u = 0; u < kPq4Chunk; ++u) {
                const int p = c * kPq4Chunk + u;
                if (p < n_plain) {                                     // uniform over the CTA
                    const uint32_t s[4] = {cur[u].x, cur[u].x >> 16, cur[u].y, cur[u].y >> 16};
                    const uint32_t t[4] = {cur[u].x >> 1, cur[u].x >> 17, cur[u].y >> 1, cur[u].y >> 17};
                    #pragma unroll
                    for (int q = 0; q < NQ; ++q) {
                        const uint4 T = *reinterpret_cast<const uint4*>(lut_s + ((size_t)(p * NQ + q) * 2 + half) * 16);
                        #pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const uint32_t r = pq4_lut4(T, s[k], t[k]);
                            R[q][k] += r;
                            H[q][k] += __byte_perm(r, 0u, 0x4341);    // bytes 1 and 3 as two uint16
                        }
                    }
                }
            }
            if (last || (c + 1) % (kPq4SegPairs / kPq4Chunk) == 0) {   // close the segment before the packed sums can carry
                #pragma unroll
                for (int q = 0; q < NQ; ++q)
                    #pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        F[q][k] = add16x2(F[q][k], R[q][k] - (H[q][k] << 8));
                        F[q][4 + k] = add16x2(F[q][4 + k], H[q][k]);
                        R[q][k] = H[q][k] = 0u;
                    }
            }
            #pragma unroll
            for (int u = 0; u < kPq4Chunk; ++u) cur[u] = nxt[u];
        }
        if (scale >= 0) {                                               // NormTableScaler: (value * scale) in uint16 lanes
            const int p = npairs - 1;
            const uint2 c = __ldg(reinterpret_cast<const uint2*>(base + p * pstride));
            const uint32_t s[4] = {c.x, c.x >> 16, c.y, c.y >> 16};
            #pragma unroll
            for (int q = 0; q < NQ; ++q) {
                const uint4 T = *reinterpret_cast<const uint4*>(lut_s + ((size_t)(p * NQ + q) * 2 + half) * 16);
                #pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const uint32_t r = pq4_lut4(T, s[k], s[k] >> 1);
                    F[q][k] = add16x2(F[q][k], mul16x2(r & 0x00ff00ffu, (uint32_t)scale));
                    F[q][4 + k] = add16x2(F[q][4 + k], mul16x2((r >> 8) & 0x00ff00ffu, (uint32_t)scale));
                }
            }
        }
        // lanes s and s ^ 2 hold the two sub-quantizers of every pair for the same vectors (the four lanes of a group run the
        // same trip count, other groups of the warp may already have left the loop: sync only the group)
        const unsigned quad = 0xfu << (threadIdx.x & 28);
        #pragma unroll
        for (int q = 0; q < NQ; ++q)
            #pragma unroll
            for (int k = 0; k < 8; ++k) F[q][k] = add16x2(F[q][k], __shfl_xor_sync(quad, F[q][k], 2));
        // word w = 2 * (s & 1) + j of the half: piece A = vectors {2w, 2w+16 | 2w+8, 2w+24}, piece B = the same + 1.
        // lanes 0,1 store word j = 0, lanes 2,3 word j = 1.
        const int j = half, w = 2 * (sub & 1) + j;
        #pragma unroll
        for (int q = 0; q < NQ; ++q) {
            const uint32_t eA = j ? F[q][2] : F[q][0], eB = j ? F[q][3] : F[q][1], oA = j ? F[q][6] : F[q][4], oB = j ? F[q][7] : F[q][5];
            uint32_t* out = reinterpret_cast<uint32_t*>(accu + (size_t)q * ld + g * 32);
            out[w] = __byte_perm(eA, eB, 0x5410);                       // vectors 2w, 2w+1
            out[w + 4] = __byte_perm(eA, eB, 0x7632);                   // 2w+8, 2w+9
            out[w + 8] = __byte_perm(oA, oB, 0x5410);                   // 2w+16, 2w+17
            out[w + 12] = __byte_perm(oA, oB, 0x7632);                  // 2w+24, 2w+25
        }
        if (base_next == nullptr) break;
        g = g_next;
        base = base_next;
    }
}

// ---- second kernel: byte-indexed pair tables in shared memory (the fused LUT-GEMV's gather idea applied to PQ4) ----
// A code BYTE holds the codes of vectors v (low nibble) and v + 16 (high nibble) for one sub-quantizer, so a 256-entry table
// tab[c] = {T[c & 15], T[c >> 4]} returns BOTH lookups of a byte, already widened to two uint16, with one LDS.32; they are
// accumulated with one integer add.  That is 1 PRMT (address) + 1 LDS + 1 add per byte = 0.5 ALU-pipe slots per lookup where
// the PRMT kernel above needs 1.9 (it is bound by the ALU pipe, 64 lanes / clk / SM, at 0.45 of the HBM peak).
// Layout: entry (sq, c) = NQP words (one per query) at plane(sq / SPP) * 64 KiB + c * 256 + (sq % SPP) * 4 NQP, SPP = 64 / NQP --
// the bank depends on the TABLE only, never on the code.  2 lanes per group of 32 vectors (lane = 16 groups x {sq 2p half, sq 2p+1 half}); lane (grp, h)
// walks the pairs in the rotated order p + grp, so at every step the 32 lanes of a warp use 32 different tables = 32 different
// banks (needs >= 16 pairs): conflict-free for any codes.  A NormTableScaler is baked into the last two tables (value * scale
// mod 2^16); their adds, and every add once nsq * 255 can reach 2^16, are done per uint16 half.
constexpr int kPq4BThreads = 256;
constexpr int kPq4BChunk = 4;                // pairs (16-byte loads) per lane and pipeline stage
constexpr int kPq4PlaneBytes = 65536;

template <int NQ>
__global__ void __launch_bounds__(kPq4BThreads) pq4_bytelut_kernel(size_t n_groups, int BB, int nsq, const uint8_t* __restrict__ codes,
                                                                     const uint8_t* __restrict__ LUT, uint16_t* __restrict__ accu,
                                                                     size_t ld, int scale) {
    extern __shared__ __align__(1024) uint8_t tab[];
    const int npairs = nsq / 2, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // An entry holds the pairs of ALL queries of the call side by side (NQP = 1, 2 or 4 words), so one LDS.32 / .64 / .128
    // serves every query: the lookups are bound by the LDS instruction rate, not by its bytes.
    constexpr int NQP = NQ == 3 ? 4 : NQ, SPP = 64 / NQP;            // words per entry; tables (sub-quantizers) per 64 KiB plane
    const int n_planes = (nsq + SPP - 1) / SPP;
    // ---- build: warp w fills rows c = 32w .. 32w+31, lane = word of the 256-byte row (conflict-free stores) ----
    for (int pw = 0; pw < n_planes * 2; ++pw) {
        const int word = (pw & 1) * 32 + lane, sq = (pw >> 1) * SPP + word / NQP, q = word % NQP;
        if (sq < nsq && q < NQ) {
            const uint8_t* T = LUT + ((size_t)(sq / 2 * NQ + q) * 32 + 16 * (sq & 1));
            const uint32_t k = (scale >= 0 && sq >= nsq - 2) ? (uint32_t)scale : 1u;
            uint8_t* dst = tab + (size_t)(pw >> 1) * kPq4PlaneBytes + word * 4;
            for (int c = warp * 32; c < warp * 32 + 32; ++c) {
                const uint32_t lo = (__ldg(T + (c & 15)) * k) & 0xffffu, hi = (__ldg(T + (c >> 4)) * k) & 0xffffu;
                *reinterpret_cast<uint32_t*>(dst + c * 256) = lo | (hi << 16);
            }
        }
    }
    __syncthreads();
    const int h = lane & 1, grp = lane >> 1;
    const bool safe_adds = (size_t)nsq * 255 >= 65536;                // per-lane sums can carry between the uint16 halves
    const size_t pstride = (size_t)BB * 32, gstride = (size_t)gridDim.x * (kPq4BThreads / 2);
    auto base_of = [&](size_t g_) {
        const unsigned gi = (unsigned)g_, blk = BB == 1 ? gi : gi / (unsigned)BB, b = gi - blk * (unsigned)BB;
        return codes + (size_t)blk * ((size_t)BB * 32 * npairs) + b * 32 + h * 16;
    };
    auto pair_of = [&](int i) { int p = i + grp; return p >= npairs ? p - npairs : p; };      // rotated walk (npairs >= 16 > grp)
    auto load_chunk = [&](uint4 (&buf)[kPq4BChunk], const uint8_t* base_, int c_) {
        #pragma unroll
        for (int u = 0; u < kPq4BChunk; ++u) {
            const int i = c_ * kPq4BChunk + u;
            buf[u] = i < npairs ? __ldg(reinterpret_cast<const uint4*>(base_ + pair_of(i) * pstride)) : make_uint4(0u, 0u, 0u, 0u);
        }
    };
    const int n_chunks = (npairs + kPq4BChunk - 1) / kPq4BChunk;
    size_t g = (size_t)blockIdx.x * (kPq4BThreads / 2) + (threadIdx.x >> 1);
    if (g >= n_groups) return;
    const uint8_t* base = base_of(g);
    uint4 cur[kPq4BChunk], nxt[kPq4BChunk];
    load_chunk(cur, base, 0);
    for (;;) {
        const size_t g_next = g + gstride;
        const uint8_t* base_next = g_next < n_groups ? base_of(g_next) : nullptr;
        uint32_t acc[NQ][16];
        #pragma unroll
        for (int q = 0; q < NQ; ++q)
            #pragma unroll
            for (int j = 0; j < 16; ++j) acc[q][j] = 0u;
        for (int c = 0; c < n_chunks; ++c) {
            if (c + 1 < n_chunks) load_chunk(nxt, base, c + 1);
            else if (base_next != nullptr) load_chunk(nxt, base_next, 0);
            #pragma unroll
            for (int u = 0; u < kPq4BChunk; ++u) {
                const int i = c * kPq4BChunk + u;
                if (i < npairs) {                                      // uniform over the CTA
                    const int sq = 2 * pair_of(i) + h;
                    const bool safe = safe_adds || (scale >= 0 && sq >= nsq - 2);
                    const uint32_t w[4] = {cur[u].x, cur[u].y, cur[u].z, cur[u].w};
                    const uint32_t lm = (uint32_t)((sq % SPP) * 4 * NQP) | ((uint32_t)(sq / SPP) << 16);   // byte 0: table column, byte 2: plane
                    auto gather = [&](auto safe_c) {
                        #pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const uint8_t* e = tab + __byte_perm(w[j >> 2], lm, 0x7604u | ((j & 3) << 4));
                            uint32_t v[NQP];
                            if (NQP == 1) v[0] = *reinterpret_cast<const uint32_t*>(e);
                            else if (NQP == 2) { const uint2 t2 = *reinterpret_cast<const uint2*>(e); v[0] = t2.x; v[NQP - 1] = t2.y; }
                            else { const uint4 t4 = *reinterpret_cast<const uint4*>(e); v[0] = t4.x; v[1 % NQP] = t4.y; v[2 % NQP] = t4.z; v[3 % NQP] = t4.w; }
                            #pragma unroll
                            for (int q = 0; q < NQ; ++q) acc[q][j] = decltype(safe_c)::value ? add16x2(acc[q][j], v[q]) : acc[q][j] + v[q];
                        }
                    };
                    if (!safe) gather(std::false_type{}); else gather(std::true_type{});
                }
            }
            #pragma unroll
            for (int u = 0; u < kPq4BChunk; ++u) cur[u] = nxt[u];
        }
        // the two lanes of a group hold the even / odd sub-quantizers of every pair: combine (per uint16 half), then lane h
        // stores vectors 16h .. 16h+15.  Byte j of a half is vector perm0[j] = j/2 + 8 (j & 1) (+16 in the high uint16).
        const unsigned duo = 0x3u << (lane & 30);
        #pragma unroll
        for (int q = 0; q < NQ; ++q) {
            #pragma unroll
            for (int j = 0; j < 16; ++j) acc[q][j] = add16x2(acc[q][j], __shfl_xor_sync(duo, acc[q][j], 1));
            const uint32_t sel = h ? 0x7632u : 0x5410u;
            uint4 o0, o1;                                              // vectors 16h + 0..7 and 16h + 8..15
            o0.x = __byte_perm(acc[q][0], acc[q][2], sel);   o0.y = __byte_perm(acc[q][4], acc[q][6], sel);
            o0.z = __byte_perm(acc[q][8], acc[q][10], sel);  o0.w = __byte_perm(acc[q][12], acc[q][14], sel);
            o1.x = __byte_perm(acc[q][1], acc[q][3], sel);   o1.y = __byte_perm(acc[q][5], acc[q][7], sel);
            o1.z = __byte_perm(acc[q][9], acc[q][11], sel);  o1.w = __byte_perm(acc[q][13], acc[q][15], sel);
            uint4* out = reinterpret_cast<uint4*>(accu + (size_t)q * ld + g * 32 + h * 16);
            out[0] = o0;
            out[1] = o1;
        }
        if (base_next == nullptr) break;
        g = g_next;
        base = base_next;
    }
}

// pq4_fast_scan.cpp:45-137: one thread per byte of the packed image.  v_base/v_count: which global vector ids `codes` holds
// (rows v_base .. v_base + v_count); accumulate: OR into the image (pack_codes_range) instead of overwriting it.
__global__ void pq4_pack_codes_kernel(const uint8_t* __restrict__ codes, size_t v_base, size_t v_count, size_t ncol, size_t block0, size_t n_bytes,
                                      size_t bbs, size_t nsq, uint8_t* __restrict__ blocks, int accumulate) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_bytes) return;
    const size_t per_block = bbs * nsq / 2, blk = block0 + t / per_block, r = t % per_block;
    const size_t p = r / bbs, i = (r % bbs) / 32, jb = r % 32, jj = jb % 16;
    const size_t perm = (jj >> 1) + ((jj & 1) << 3);                    // perm0 = {0, 8, 1, 9, ...}
    const size_t v0 = blk * bbs + i * 32 + perm, v1 = v0 + 16;
    const uint8_t c0 = (v0 >= v_base && v0 - v_base < v_count) ? codes[(v0 - v_base) * ncol + p] : 0;
    const uint8_t c1 = (v1 >= v_base && v1 - v_base < v_count) ? codes[(v1 - v_base) * ncol + p] : 0;
    const uint8_t d = jb < 16 ? (uint8_t)((c0 & 15) | ((c1 & 15) << 4)) : (uint8_t)((c0 >> 4) | ((c1 >> 4) << 4));
    uint8_t* o = blocks + blk * per_block + r;
    *o = accumulate ? (uint8_t)(*o | d) : d;
}

// pq4_fast_scan.cpp:253-264 / 283-303: dest[(sq/2 * nq + qi) * 32 + 16 * (sq & 1) + e] = src[(q * nsq + sq) * 16 + e]
__global__ void pq4_pack_lut_kernel(int nq, int nsq, const uint8_t* __restrict__ src, const int* __restrict__ q_map, uint8_t* __restrict__ dest) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nq * nsq * 16) return;
    const int e = t & 15, sq = (t >> 4) % nsq, qi = (t >> 4) / nsq;
    const int q = q_map ? q_map[qi] : qi;
    dest[(size_t)(sq / 2 * nq + qi) * 32 + 16 * (sq & 1) + e] = src[((size_t)q * nsq + sq) * 16 + e];
}

static int g_pq4_mode = 0;                   // 0: choose by size; 1: PRMT kernel; 2: byte-table kernel whenever it applies (tests)
void pq4_set_mode(int m) { g_pq4_mode = m; }

static int pq4_accumulate_launch(int nq, size_t nb, int bbs, int nsq, const uint8_t* codes, const uint8_t* LUT, uint16_t* accu, size_t ld,
                                 int scale, cudaStream_t st) {
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    const size_t smem = (size_t)nq * nsq * 16;
    LUTVQ_REQUIRE(smem <= di.smem_optin - 1024, "packed LUT of %zu bytes does not fit in shared memory", smem);
    const size_t n_groups = nb / 32;
    LUTVQ_REQUIRE(n_groups < (1ull << 32), "nb=%zu: at most 2^37 database vectors per call", nb);
    // Large scans with >= 16 pairs take the byte-table kernel (HBM-bound); it needs 64 KiB of shared memory per 64 tables,
    // 16-byte aligned outputs and enough groups to amortise the per-CTA table build.
    const int nqp = nq == 3 ? 4 : nq, planes = (nsq + 64 / nqp - 1) / (64 / nqp);
    const bool bytelut = g_pq4_mode != 1 && nsq >= 32 && nsq <= 512 && (size_t)planes * kPq4PlaneBytes <= di.smem_optin - 1024 &&
                         (n_groups >= 8192 || g_pq4_mode == 2) && ld % 8 == 0 && (reinterpret_cast<uintptr_t>(accu) & 15u) == 0;
    if (bytelut) {
        const size_t smem2 = (size_t)planes * kPq4PlaneBytes;
        const int per_sm = planes == 1 ? (nq == 1 ? 3 : 2) : 1;        // registers: 73 / 128 / 139 / 141 for nq = 1..4
        const size_t want2 = (n_groups + kPq4BThreads / 2 - 1) / (kPq4BThreads / 2), cap2 = (size_t)di.sm_count * per_sm;
        const unsigned grid2 = (unsigned)(want2 < cap2 ? want2 : cap2);
#define LUTVQ_PQ4B_LAUNCH(NQ)                                                                                                      \
    case NQ: {                                                                                                                     \
        LUTVQ_CUDA_OK(cudaFuncSetAttribute(pq4_bytelut_kernel<NQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));      \
        pq4_bytelut_kernel<NQ><<<grid2, kPq4BThreads, smem2, st>>>(n_groups, bbs / 32, nsq, codes, LUT, accu, ld, scale);          \
        break;                                                                                                                     \
    }
        switch (nq) {
            LUTVQ_PQ4B_LAUNCH(1) LUTVQ_PQ4B_LAUNCH(2) LUTVQ_PQ4B_LAUNCH(3) LUTVQ_PQ4B_LAUNCH(4)
            default: return fail(LUTVQ_EINVAL, "nq=%d: the kernel takes 1..4 queries per call", nq);
        }
#undef LUTVQ_PQ4B_LAUNCH
        LUTVQ_CUDA_OK(cudaGetLastError());
        return LUTVQ_OK;
    }
    const size_t want = (n_groups + kPq4Threads / 4 - 1) / (kPq4Threads / 4), cap = (size_t)di.sm_count * 8;
    const unsigned grid = (unsigned)(want < cap ? want : cap);
#define LUTVQ_PQ4_LAUNCH(NQ)                                                                                                       \
    case NQ: {                                                                                                                     \
        if (smem > 48 * 1024) LUTVQ_CUDA_OK(cudaFuncSetAttribute(pq4_accumulate_kernel<NQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        pq4_accumulate_kernel<NQ><<<grid, kPq4Threads, smem, st>>>(n_groups, bbs / 32, nsq, codes, LUT, accu, ld, scale);          \
        break;                                                                                                                     \
    }
    switch (nq) {
        LUTVQ_PQ4_LAUNCH(1) LUTVQ_PQ4_LAUNCH(2) LUTVQ_PQ4_LAUNCH(3) LUTVQ_PQ4_LAUNCH(4)
        default: return fail(LUTVQ_EINVAL, "nq=%d: the kernel takes 1..4 queries per call", nq);
    }
#undef LUTVQ_PQ4_LAUNCH
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

}  // namespace lutvq

using namespace lutvq;

extern "C" {

// tests / tools: 0 = choose the accumulation kernel by size (default), 1 = PRMT kernel, 2 = byte-table kernel whenever it applies
void lutvq_debug_set_pq4_mode(int mode) { pq4_set_mode(mode); }

int lutvq_pq4_pack_codes(const void* codes, size_t ntotal, size_t M, size_t nb, size_t bbs, size_t nsq, void* blocks, void* stream) {
    LUTVQ_REQUIRE(bbs > 0 && bbs % 32 == 0 && nb % bbs == 0 && nsq % 2 == 0, "bbs %% 32, nb %% bbs and nsq %% 2 must be 0 (pq4_fast_scan.cpp:53-55)");
    LUTVQ_REQUIRE(nsq == (M + 1) / 2 * 2, "nsq=%zu must be M=%zu rounded up to a multiple of 2", nsq, M);
    if (nb == 0) return LUTVQ_OK;
    LUTVQ_REQUIRE(blocks && (codes || ntotal == 0), "null pointer");
    const size_t n_bytes = nb * nsq / 2;
    pq4_pack_codes_kernel<<<(unsigned)((n_bytes + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const uint8_t*>(codes), 0, ntotal, (M + 1) / 2, 0, n_bytes, bbs, nsq, static_cast<uint8_t*>(blocks), 0);
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

int lutvq_pq4_pack_codes_range(const void* codes, size_t M, size_t i0, size_t i1, size_t bbs, size_t nsq, void* blocks, void* stream) {
    LUTVQ_REQUIRE(bbs > 0 && bbs % 32 == 0 && nsq % 2 == 0 && nsq == (M + 1) / 2 * 2, "bbs %% 32 and nsq = M rounded up to 2");
    if (i1 <= i0) return LUTVQ_OK;
    LUTVQ_REQUIRE(codes && blocks, "null pointer");
    const size_t block0 = i0 / bbs, block1 = (i1 - 1) / bbs + 1, n_bytes = (block1 - block0) * bbs * nsq / 2;
    pq4_pack_codes_kernel<<<(unsigned)((n_bytes + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const uint8_t*>(codes), i0, i1 - i0, (M + 1) / 2, block0, n_bytes, bbs, nsq, static_cast<uint8_t*>(blocks), 1);
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

int lutvq_pq4_pack_LUT(int nq, int nsq, const void* src, void* dest, void* stream) {
    LUTVQ_REQUIRE(nq >= 0 && nsq >= 0 && nsq % 2 == 0, "nsq must be a multiple of 2");
    if (nq == 0 || nsq == 0) return LUTVQ_OK;
    LUTVQ_REQUIRE(src && dest, "null pointer");
    pq4_pack_lut_kernel<<<(nq * nsq * 16 + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(nq, nsq, static_cast<const uint8_t*>(src), nullptr,
                                                                                                   static_cast<uint8_t*>(dest));
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

int lutvq_pq4_qbs_to_nq(int qbs) {
    int nq = 0;
    for (unsigned q = (unsigned)qbs; q; q >>= 4) nq += (int)(q & 15u);
    return nq;
}

// q_map may be null (pq4_pack_LUT_qbs) or a DEVICE array of nq source-row indices (pq4_pack_LUT_qbs_q_map)
int lutvq_pq4_pack_LUT_qbs(int qbs, int nsq, const void* src, const int* q_map, void* dest, void* stream) {
    LUTVQ_REQUIRE(nsq > 0 && nsq % 2 == 0, "nsq must be a multiple of 2 (pq4_fast_scan.cpp:267)");
    LUTVQ_REQUIRE(src && dest, "null pointer");
    const size_t dim12 = 16 * (size_t)nsq;
    int i0 = 0;
    for (unsigned qi = (unsigned)qbs; qi; qi >>= 4) {
        const int nq = (int)(qi & 15u);
        if (nq == 0) continue;
        // without a map the block's rows are src rows i0..; with one they are q_map[i0..] (the kernel indexes src from row 0)
        pq4_pack_lut_kernel<<<(nq * nsq * 16 + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
            nq, nsq, static_cast<const uint8_t*>(src) + (q_map ? 0 : i0 * dim12), q_map ? q_map + i0 : nullptr, static_cast<uint8_t*>(dest) + i0 * dim12);
        LUTVQ_CUDA_OK(cudaGetLastError());
        i0 += nq;
    }
    return LUTVQ_OK;
}

int lutvq_pq4_accumulate_loop(int nq, size_t nb, int bbs, int nsq, const void* codes, const void* LUT, void* accu, size_t ld, int scale, void* stream) {
    LUTVQ_REQUIRE(bbs > 0 && bbs % 32 == 0 && nb % (size_t)bbs == 0, "bbs %% 32 and nb %% bbs must be 0 (pq4_fast_scan_search_1.cpp:149-150)");
    LUTVQ_REQUIRE(nsq > 0 && nsq % 2 == 0, "nsq must be a positive multiple of 2");
    LUTVQ_REQUIRE(nq >= 1 && nq <= 4, "nq=%d: 1..4 queries per call (the reference instantiates (nq, bbs/32) in {1x1..5, 2x1, 2x2, 3x1, 4x1})", nq);
    LUTVQ_REQUIRE(scale < 65536, "scale must fit 16 bits");
    if (nb == 0) return LUTVQ_OK;
    LUTVQ_REQUIRE(codes && LUT && accu, "null pointer");
    LUTVQ_REQUIRE((reinterpret_cast<uintptr_t>(codes) & 31u) == 0 && (reinterpret_cast<uintptr_t>(LUT) & 31u) == 0,
                  "codes and LUT must be 32-byte aligned (pq4_fast_scan_search_1.cpp:147-148)");
    LUTVQ_REQUIRE(ld >= nb && ld % 2 == 0 && (reinterpret_cast<uintptr_t>(accu) & 3u) == 0, "accu: ld >= nb, even, 4-byte aligned");
    return pq4_accumulate_launch(nq, nb, bbs, nsq, static_cast<const uint8_t*>(codes), static_cast<const uint8_t*>(LUT), static_cast<uint16_t*>(accu), ld,
                                 scale, static_cast<cudaStream_t>(stream));
}

int lutvq_pq4_accumulate_loop_qbs(int qbs, size_t nb, int nsq, const void* codes, const void* LUT, void* accu, size_t ld, int scale, void* stream) {
    LUTVQ_REQUIRE(nb % 32 == 0, "nb must be a multiple of 32 (the qbs kernels run with bbs = 32)");
    LUTVQ_REQUIRE(nsq > 0 && nsq % 2 == 0, "nsq must be a positive multiple of 2");
    for (unsigned qi = (unsigned)qbs; qi; qi >>= 4) LUTVQ_REQUIRE((qi & 15u) >= 1 && (qi & 15u) <= 4, "qbs=0x%x: every digit must be 1..4", qbs);
    LUTVQ_REQUIRE(scale < 65536, "scale must fit 16 bits");
    if (nb == 0 || qbs == 0) return LUTVQ_OK;
    LUTVQ_REQUIRE(codes && LUT && accu, "null pointer");
    LUTVQ_REQUIRE((reinterpret_cast<uintptr_t>(codes) & 31u) == 0 && (reinterpret_cast<uintptr_t>(LUT) & 31u) == 0, "codes and LUT must be 32-byte aligned");
    LUTVQ_REQUIRE(ld >= nb && ld % 2 == 0 && (reinterpret_cast<uintptr_t>(accu) & 3u) == 0, "accu: ld >= nb, even, 4-byte aligned");
    int i0 = 0;
    for (unsigned qi = (unsigned)qbs; qi; qi >>= 4) {
        const int nq = (int)(qi & 15u);
        if (int rc = pq4_accumulate_launch(nq, nb, 32, nsq, static_cast<const uint8_t*>(codes), static_cast<const uint8_t*>(LUT) + (size_t)i0 * 16 * nsq,
                                           static_cast<uint16_t*>(accu) + (size_t)i0 * ld, ld, scale, static_cast<cudaStream_t>(stream)))
            return rc;
        i0 += nq;
    }
    return LUTVQ_OK;
}

}  // extern "C"
