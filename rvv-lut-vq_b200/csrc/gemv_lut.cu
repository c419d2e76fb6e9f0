// gemv_lut.cu -- fused LUT-build + code-indexed gather-accumulate GEMV for AQLM 2x8 codes (sm_100a).
//
// Replaces Code2x8MatVec (reference benchmark/AQLM/inference_lib/src/aqlm/inference_kernels/
// cuda_kernel.cu:144-233) with the LUT formulation of the golden model (vq_dataflow_sim/vq.py:280-298):
//
//   K1  L[g][n][k] = sum_j C[n][k][j] * x[8g+j]      built per CTA for ITS 32-group k-slice with
//                                                     mma.sync.m16n8k8 (fp32 accumulate) into 64 KiB of smem
//   K2  y[o] = sum_g sum_n L[g][n][code(o,g,n)]       codes streamed HBM -> L2 (bulk prefetch) -> registers
//                                                     (one 128-bit no-allocate load per lane per 8-row tile),
//                                                     one PRMT + one LDS + one FADD per code byte,
//                                                     conflict-free for any index pattern (lut_layout.h)
//
// Work decomposition.  The work items of one GEMV are (k-slice, 32-row set) pairs, linearised slice-major.
// The grid is EXACTLY one CTA per SM (a "worker"); worker w owns a contiguous range of items:
//   aligned  (n_k <= #SMs): the workers are dealt to the slices (floor or ceil(#SMs/n_k) each) and split that
//            slice's row sets evenly -> one LUT build per worker;
//   linear   (otherwise, or when cheaper): range w*T/W .. (w+1)*T/W, which may straddle slices (<= 2 builds
//            when T/W >= n_sets, more for very long K).
// One CTA per SM per launch matters under programmatic dependent launch: consecutive launches then pair up
// 1:1 on every SM (occupancy 2), instead of a 144-CTA grid doubling up on the 4 idle SMs and straggling.
//
// Cross-CTA split-k reduction: one 64-bit accumulator word per output row, (fixed-point sum << 8) | arrivals.  Each
// (row, k-slice) partial is ONE integer atomic add executed at the row's L2 home slice; the atomic's return value
// tells the contributor whether it was the last of the n_k, in which case it holds the complete sum, applies the
// reference epilogue order (round to dtype, * scale, + bias; cuda_kernel.cpp:95-111), writes y and zeroes the word.
// Integer addition commutes => the result is bit-identical whatever the arrival order; nothing polls, nothing
// needs the CTAs of a launch to be co-resident, and the workspace is self-cleaning (all zero between launches).
// The return value is examined PF sets after the atomic was issued, so its L2 round trip hides behind lookups;
// workers of different k-slices walk their rows in rotated order so the n_k atomics of a row do not collide.
// (Measured alternatives on B200, 4096x11008: {value, tag} words polled through L2 cost 1.7-2.7 us per launch
// after the last store -- one poll round of 16 loads took 2.4 us -- against ~1 us for the atomic scheme.)
//
// Stream overlap (programmatic dependent launch).  Weights (codes, codebooks, scales, bias) never depend on the
// previous kernel, so the first tile sets are requested into registers and the rest of the worker's share is pushed
// HBM -> L2 BEFORE griddepcontrol.wait.  With LUTVQ_FLAG_X_READY the caller additionally promises that x was not
// written by the kernel(s) immediately preceding this call in the stream and that the workspace is not shared with
// them (include/lutvq.h); then the LUT build, the gather and the split-k atomics all run before the wait and only
// the y stores (and this kernel's completion) are ordered behind the previous kernel.
//
// Two code layouts feed the same gather loop:
//   reference  int8 [O][G][2] exactly as QuantizedLinear.codes: a warp's 128-bit load touches 8 rows x 64 B;
//   packed     lutvq_code2x8_prepack(): [k-slice][32-row set][8-row tile][lane] x 16 B, the lane's register
//              image, so a warp's load is 512 contiguous bytes and a worker's share is one contiguous range.
//
// Algorithmic HBM bytes per launch: O*(I/8)*2 codes + 8192 codebooks + 2I x + 2O scales + 2O y.
#include "common.cuh"
#include "lut_layout.h"

#ifndef LUTVQ_DBG
#define LUTVQ_DBG 0          // 1: compile the experiment knobs (flags 0x100 no lookups, 0x200 no streaming, 0x400 no atomics)
#endif
#if LUTVQ_DBG
#define LUTVQ_DBG_AND(x) && (x)
#else
#define LUTVQ_DBG_AND(x)
#endif

namespace lutvq {

constexpr size_t kWsHeaderBytes = 256;        // reserved

constexpr int kMaxPeers = 8;                  // one NVSwitch domain
constexpr int kRangeTable = 320;              // workers whose item ranges ride in the kernel parameters (B200: 148 or 2 x 148)

constexpr int kMaxProblems = 8;               // linears that share x (q/k/v, gate/up) or rows of a small batch that share the codes: ONE launch

// One linear layer of a (possibly grouped) launch.  Every worker (CTA) belongs to exactly one problem.
// two fp16 lanes of silu(g) * u, rounded once: the arithmetic of decode_glue.cu silu_mul_kernel, bit for bit
__device__ __forceinline__ uint32_t silu_mul_halves(uint32_t g, uint32_t u) {
    const __half2 gh = *reinterpret_cast<const __half2*>(&g), uh = *reinterpret_cast<const __half2*>(&u);
    const float a0 = __low2float(gh), a1 = __high2float(gh);
    const __half r0 = __float2half_rn(a0 / (1.f + __expf(-a0)) * __low2float(uh));
    const __half r1 = __float2half_rn(a1 / (1.f + __expf(-a1)) * __high2float(uh));
    const __half2 r = __halves2half2(r0, r1);
    return *reinterpret_cast<const uint32_t*>(&r);
}

struct GemvProblem {
    const uint8_t* codes;     // reference layout [O][G][2]  or packed layout (see above)
    const void* codebook;     // [2][256][8]
    void* y;                  // [O]
    const void* scales;       // [O] or null
    const void* bias;         // [O] or null
    unsigned long long* acc;  // [O_pad] split-k accumulator words, zero between launches
    int O, n_sets;
    unsigned x_off;           // byte offset of this problem's activation row from GemvParams::x / x_gate (rows of a multi-row call)
};

struct GemvParams {
    GemvProblem prob[kMaxProblems];
    int n_prob;
    const void* x;            // [I], shared by the problems of a group
    const void* x_gate;       // fp16 only, or null: the activations are silu(x_gate[i]) * x[i] (SwiGLU fused into the x load)
    int G, n_k;
    unsigned flags;
    // ---- output-channel-sharded decode (shard.cu): besides y, every finished row is stored into each peer's gather
    // buffer over NVLink (peer pointers already offset to this rank's row range), and the LAST CTA of the launch to
    // finish publishes the slot's epoch to every peer.  n_push == 0: plain single-GPU GEMV, none of this is touched.
    int n_push;
    void* push_y[kMaxPeers];                    // peer r's gather buffer + this rank's row offset
    unsigned* push_flag[kMaxPeers];             // peer r's epoch flag for (slot, this rank)
    unsigned* push_sent;                        // local: pushes completed on this slot so far (epoch counter)
    unsigned* cta_arrivals;                     // local: workspace header word counting finished CTAs (self-resetting)
    int has_table;                              // 1: the host tabulated every worker's range (W <= kRangeTable)
    int tab_begin[kRangeTable];                 // worker w owns linear items [tab_begin[w], tab_end[w]) of ITS problem
    int tab_end[kRangeTable];
    unsigned short tab_ks[kRangeTable];         // k-slice of item tab_begin[w]
    unsigned char tab_prob[kRangeTable];        // problem of worker w (ranges are per problem)
    unsigned long long* tl;   // debug timeline [256 CTAs][16] of %globaltimer stamps for this launch, or null
    GemvLL ll;                // sharded decode: outputs as LL words into every rank's gather slot / inputs from LL words (common.cuh)
};

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#if LUTVQ_DBG
#define TL_STAMP(i)                                                                                              \
    do {                                                                                                         \
        if (p.tl != nullptr && threadIdx.x == 0) p.tl[(size_t)blockIdx.x * 16 + (i)] = globaltimer_ns();          \
    } while (0)
#else
#define TL_STAMP(i) do { } while (0)
#endif

template <int DT> struct MmaK8;
template <> struct MmaK8<LUTVQ_F16> {
    static __device__ __forceinline__ void run(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(b0));
    }
};
template <> struct MmaK8<LUTVQ_BF16> {
    static __device__ __forceinline__ void run(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(b0));
    }
};

__device__ __forceinline__ void sts_v4(uint32_t addr, float a, float b, float c, float d) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" :: "r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
// Split-k accumulator word (one per output row, zero between launches):  (fixed-point sum << 8) | arrivals.
// Every (row, k-slice) partial is ONE 64-bit integer atomic add performed at the row's L2 home slice; integer
// addition commutes, so the result is bit-identical whatever the arrival order.  The atomic returns the old word:
// the contributor that sees arrivals == n_k - 1 holds the complete sum (old + its own), writes y and zeroes the word.
// Fixed point: 2^-kFxBits per unit in a 56-bit signed field.  fp16 (|y| <= 65504): 2^-30 -> +-3.4e7 of range with
// 9.3e-10 resolution (below half an ulp of every normal fp16 output).  bf16: 2^-20 -> +-3.4e10.  Partials are
// saturated to the field (a non-finite partial saturates too: the output becomes +-inf in fp16, not NaN).
template <int DT> struct Fx;
template <> struct Fx<LUTVQ_F16>  { static constexpr float scale = 1073741824.f, inv = 1.f / 1073741824.f; };   // 2^30
template <> struct Fx<LUTVQ_BF16> { static constexpr float scale = 1048576.f,    inv = 1.f / 1048576.f; };      // 2^20

template <int DT>
__device__ __forceinline__ unsigned long long fx_contribution(float v) {
    // clamp so that the sum of n_k <= 255 contributions cannot leave the signed 56-bit field: |fx| < 2^47
    float sv = v * Fx<DT>::scale;
    sv = fminf(fmaxf(sv, -1.4e14f), 1.4e14f);                   // NaN -> -1.4e14 (fmaxf/fminf drop NaN), 1.4e14 < 2^47
    const long long fx = __float2ll_rn(sv);
    return ((unsigned long long)fx << 8) + 1ull;
}
__device__ __forceinline__ unsigned long long atom_add_u64(unsigned long long* p, unsigned long long v) {
    unsigned long long old;
    asm volatile("atom.relaxed.gpu.global.add.u64 %0, [%1], %2;" : "=l"(old) : "l"(p), "l"(v) : "memory");
    return old;
}
template <int DT>
__device__ __forceinline__ float fx_total(unsigned long long word) {
    return (float)((long long)word >> 8) * Fx<DT>::inv;          // arithmetic shift drops the arrival count
}
// asynchronous HBM -> L2 prefetch of a contiguous range (no registers, no shared memory)
// L2 residency: the code stream is read exactly once per GEMV (evict_first), while the few KB of codebook / scales / bias
// of every layer are re-read every token (evict_last): without the hints the 11 MB stream of each launch pushes them out
// and every launch pays HBM latency for its codebook in the prologue and for its scales in the tail
// (round 2: 5.13 -> 4.50 us per independent launch, 8.1 -> 7.7 dependent, on the 16-bit-LUT kernel this was measured on).
__device__ __forceinline__ uint4 ldcg_u128(const void* p) {            // L2-coherent: x may be the previous kernel's output
    uint4 v;
    asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ uint4 ldg_stream_u128_pol(const void* p, uint64_t pol) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p), "l"(pol));
    return r;
}
__device__ __forceinline__ uint32_t ldg_u32_pol(const void* p, uint64_t pol) {
    uint32_t r;
    asm volatile("ld.global.nc.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(r) : "l"(p), "l"(pol));
    return r;
}
__device__ __forceinline__ uint16_t ldg_u16_pol(const void* p, uint64_t pol) {
    uint16_t r;
    asm volatile("ld.global.nc.L2::cache_hint.u16 %0, [%1], %2;" : "=h"(r) : "l"(p), "l"(pol));
    return r;
}
__device__ __forceinline__ void l2_prefetch_bulk_pol(const void* p, uint32_t bytes, uint64_t pol) {
    asm volatile("cp.async.bulk.prefetch.L2.global.L2::cache_hint [%0], %1, %2;" :: "l"(p), "r"(bytes), "l"(pol) : "memory");
}
__device__ __forceinline__ void l2_prefetch_bulk(const void* p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(p), "r"(bytes) : "memory");
}
// per-thread (non-uniform address) L2 prefetch of the line holding p
__device__ __forceinline__ void l2_prefetch_line(const void* p) {
    asm volatile("prefetch.global.L2 [%0];" :: "l"(p));
}

template <int DT>
__device__ __forceinline__ void epilogue_store(void* y, const void* scales, const void* bias, int row, float acc) {
    using D = Dtype<DT>;
    using T = typename D::T;
    T r = D::from_f(acc);                                                  // Code2x8MatVec writes round(acc)   (cuda_kernel.cu:221-231)
    if (scales) r = D::mul(r, static_cast<const T*>(scales)[row]);         // flat_output *= scales             (cuda_kernel.cpp:101)
    if (bias)   r = D::add(r, static_cast<const T*>(bias)[row]);           // flat_output += bias               (cuda_kernel.cpp:103)
    static_cast<T*>(y)[row] = r;
}

// ---- fixed-point LUT entries --------------------------------------------------------------------------------
// The LUT holds INTEGERS: entry = round(L * 2^e) with one exponent e per (CTA, k-slice), chosen from the bound
// |L| <= 8 * max|codebook| * max|x_slice| so that |entry| <= 2^22 (the whole signed range of the magic-number trick).  The conversion is free: the build adds the
// magic constant 1.5 * 2^23 with the scaling FMA, which leaves 0x4B400000 + entry in the float's bits, and the
// gather adds those words as plain 32-bit integers (one IADD3 per TWO lookups instead of one FADD per lookup); the
// 64 magic offsets of a row partial (64 tables) are removed once, mod 2^32.  Integer sums are exact and order-free;
// resolution is 2^-22 of the slice bound, i.e. >= 2^11 times finer than the fp16 LUT of the golden model (vq.py:285).
constexpr float    kFxMagic     = 12582912.f;                       // 1.5 * 2^23
constexpr uint32_t kFxMagicBits = 0x4B400000u;
constexpr uint32_t kFxMagic64   = (uint32_t)(64ull * kFxMagicBits); // 64 lookups per row partial, mod 2^32

// |v| bits of the two halves of a packed fp16x2 / bf16x2 word, as an integer that orders like |v| (NaN highest)
__device__ __forceinline__ uint32_t absmax_halves(uint32_t w) { return max(w & 0x7FFFu, (w >> 16) & 0x7FFFu); }
template <int DT> __device__ __forceinline__ float abs_half_bits_to_float(uint32_t h);
template <> __device__ __forceinline__ float abs_half_bits_to_float<LUTVQ_F16>(uint32_t h)  { return __half2float(__ushort_as_half((unsigned short)h)); }
template <> __device__ __forceinline__ float abs_half_bits_to_float<LUTVQ_BF16>(uint32_t h) { return __uint_as_float(h << 16); }
// scale 2^e and its inverse for a slice whose entries are bounded by `bound` (>= 0): |L * 2^e| <= 2^22, so a row
// partial (64 entries) stays below 2^28.
// Degenerate bounds: < 2^-107 -> inverse flushes to 0 (outputs 0); inf/NaN -> inverse NaN (the partial saturates).
__device__ __forceinline__ void fx_scales(float bound, float& scale, float& inv) {
    uint32_t ex = (__float_as_uint(bound) >> 23) & 0xFFu;            // bound < 2^(ex - 126)
    const bool bad = ex == 255u;
    ex = max(ex, 21u);
    scale = __uint_as_float((275u - ex) << 23);                      // 2^(148 - ex): bound * scale < 2^22
    inv   = bad ? __uint_as_float(0x7FC00000u) : __uint_as_float((ex - 21u) << 23);
}

// store one finished output element locally and into every peer's gather buffer (sharded decode)
template <int DT>
__device__ __forceinline__ void store_y(const GemvParams& p, const GemvProblem& PB, int row, typename Dtype<DT>::T v) {
    using T = typename Dtype<DT>::T;
    if (p.ll.nranks > 0) {
        // LL output: {value | epoch16 << 16} into every rank's gather slot (one atomic 32-bit store each, no flag, no fence)
        const int pi = (int)(&PB - p.prob);
        const unsigned ep = ll_epoch16(*p.ll.tick, p.ll.out_seq[pi]);       // after the dependency wait (every y store is)
        const uint32_t word = (uint32_t)*reinterpret_cast<const unsigned short*>(&v) | (ep << 16);
        const size_t off = p.ll.out_off[pi] + (size_t)row * 4;
        for (int r = 0; r < p.ll.nranks; ++r) ll_st(p.ll.peer_base[r], off, word);
        return;
    }
    static_cast<T*>(PB.y)[row] = v;
    for (int r = 0; r < p.n_push; ++r) static_cast<T*>(p.push_y[r])[row] = v;       // NVLink stores (peer memory)
}

// One warp tile set = 32 rows x 64 bytes: 4 tiles of 8 rows; per lane 4 x 16 bytes.
struct CodeRegs { uint32_t w[4][4]; };

// Fallback when the grid is larger than the tabulated ranges (never on a 148-SM B200): the linear plan, one problem.
__device__ __forceinline__ void worker_range(const GemvParams& p, int w, int W, int& begin, int& end) {
    const long long T = (long long)p.n_k * p.prob[0].n_sets;
    begin = (int)((T * w) / W);
    end   = (int)((T * (w + 1)) / W);
}

// NCB = codebooks per group of 8 inputs (AQLM Kx8: 1, 2, 4 or 8).  A k-slice is 64 code bytes of every row whatever NCB is
// (= 64 / NCB groups); only the LUT build knows which (group, codebook) a byte position means -- the gather, the split-k
// protocol and the epilogue are format-blind.
template <int DT, bool PACKED, int NW, int PF, int PA, int MINB, int NCB = 2>
__global__ void __launch_bounds__(NW * 32, MINB)
lut_gemv_2x8_kernel(const GemvParams p) {
    static_assert(NCB == 1 || NCB == 2 || NCB == 4 || NCB == 8, "codebooks per group");
    static_assert(!PACKED || NCB == 2, "the packed code image is built for 2x8");
    constexpr int kSliceGroups = kSliceTables / NCB;                  // groups of 8 activations per k-slice
    constexpr int kSliceXBytes = kSliceGroups * 16;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t lut_base = smem_u32(smem_raw);
    const GemvProblem& PB = p.prob[p.n_prob > 1 ? (int)p.tab_prob[blockIdx.x] : 0];       // indexed constant-bank loads

    TL_STAMP(0);
#if LUTVQ_DBG
    if (p.tl != nullptr && threadIdx.x == 0) {
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        p.tl[(size_t)blockIdx.x * 16 + 15] = smid + 1;
    }
#endif
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int W = gridDim.x;
    int begin, end;
    if (p.has_table) { begin = p.tab_begin[blockIdx.x]; end = p.tab_end[blockIdx.x]; }          // two constant-bank loads
    else worker_range(p, blockIdx.x, W, begin, end);
    const size_t row_bytes = (size_t)p.G * NCB;
    const int last_row = PB.O - 1;

    // ---------------- streaming-side lane constants ----------------
    // Reference layout: no predicated loads.  Rows past the end are clamped to the last row (result discarded)
    // and lanes whose 16-byte chunk lies past the end of the row (last, partial k-slice) re-read chunk 0 --
    // their tables were built from x = 0, so whatever index they fetch contributes exactly 0.
    const int e = stream_first_half(lane);
    const uint64_t pol_stream = l2_policy_evict_first();
    auto load_set = [&](CodeRegs& r, int ks, int set) {
        if (PACKED) {
            const uint4* base = reinterpret_cast<const uint4*>(PB.codes) + (((size_t)ks * PB.n_sets + set) * 4) * 32 + lane;
            #pragma unroll
            for (int t = 0; t < 4; ++t) {
                const uint4 v = ldg_stream_u128_pol(base + t * 32, pol_stream);
                r.w[t][0] = v.x; r.w[t][1] = v.y; r.w[t][2] = v.z; r.w[t][3] = v.w;
            }
        } else {
            const int slice_byte0 = ks * kSliceTables;
            const bool chunk_valid = (size_t)(slice_byte0 + 16 * stream_chunk(lane)) < row_bytes;
            const uint8_t* lane_base = PB.codes + slice_byte0 + (chunk_valid ? 16 * stream_chunk(lane) : 0);
            const int rbase = set * kSetRows + stream_row_in_tile(lane);
            #pragma unroll
            for (int t = 0; t < 4; ++t) {
                const uint32_t row = (uint32_t)min(rbase + t * kTileRows, last_row);
                const uint4 v = ldg_stream_u128_pol(lane_base + (uint64_t)row * (uint32_t)row_bytes, pol_stream);
                // register word c holds memory word c ^ 2e (lut_layout.h)
                r.w[t][0] = e ? v.z : v.x; r.w[t][1] = e ? v.w : v.y;
                r.w[t][2] = e ? v.x : v.z; r.w[t][3] = e ? v.y : v.w;
            }
        }
    };

    TL_STAMP(8);
    // The workers of different k-slices walk the same rows; each starts at a slice-dependent rotation of its own
    // range so that the n_k atomics of one row are spread over the whole gather instead of colliding at one instant.
    auto rot_of = [&](int ks_, int n) { return (int)(((unsigned)ks_ * (unsigned)n) / (unsigned)p.n_k); };
    auto wrap = [](int v, int n) { while (v >= n) v -= n; return v; };
    // first segment of this worker's range: slice ks, sets [lo, hi)
    int pos = begin;
    int ks = p.has_table ? (int)p.tab_ks[blockIdx.x] : (begin < end) ? (int)((unsigned)begin / (unsigned)PB.n_sets) : 0;
    int lo = begin - ks * PB.n_sets;
    int hi = min(PB.n_sets, lo + (end - begin));

    int seg_rot = rot_of(ks, hi - lo);
    // ---------------- prologue that does not depend on the previous kernel in the stream ----------------
    // Codes and codebooks are weights.  (1) the first PF tile sets are requested into registers, (2) the rest
    // of this worker's share is pushed HBM -> L2, (3) x / the workspace header are pulled into L2 (L2 is the
    // coherence point, so this is safe even while the previous kernel is still writing them), (4) the codebook
    // MMA fragments are loaded -- all BEFORE the programmatic-dependency wait.
    CodeRegs buf[PF];
    #pragma unroll
    for (int i = 0; i < PF; ++i)
        if (lo + warp + i * NW < hi) load_set(buf[i], ks, lo + wrap(warp + i * NW + seg_rot, hi - lo));
    TL_STAMP(9);
    if (PACKED) {
        // the worker's whole share is ONE contiguous range of the packed image (linear item index * 2048 B)
        if (lane == 0) {
            const uint8_t* base = PB.codes + (size_t)begin * 2048;
            const int n = end - begin;
#if LUTVQ_DBG
            if (!(p.flags & 0x800u))                               // DEBUG 0x800: no L2 prefetch of the share
#endif
            for (int c = warp * 4 + PF * NW; c < n; c += NW * 4)
                l2_prefetch_bulk_pol(base + (size_t)c * 2048, (uint32_t)min(4, n - c) * 2048u, pol_stream);
        }
    } else {
        int ppos = begin;
        while (ppos < end) {
            const int s = ppos / PB.n_sets, l = ppos - s * PB.n_sets, h = min(PB.n_sets, l + (end - ppos));
            const size_t off = (size_t)s * kSliceTables;
            const int r1 = min(h * kSetRows, PB.O);
            for (int r = l * kSetRows + threadIdx.x; r < r1; r += NW * 32)
                l2_prefetch_line(PB.codes + (size_t)r * row_bytes + off);
            ppos += h - l;
        }
    }
    if (p.n_k > 1 && begin < end) {
        // scale / bias of this worker's first-segment rows: 64 B per set each, pulled into L2 for the in-loop fetch
        const int nl = (hi - lo + 1) / 2;                             // 128-byte lines
        if ((int)threadIdx.x < nl) {
            const size_t off = (size_t)lo * 64 + (size_t)threadIdx.x * 128;
            if (off < (size_t)PB.O * 2) {
                if (PB.scales) l2_prefetch_line(static_cast<const uint8_t*>(PB.scales) + off);
                if (PB.bias)   l2_prefetch_line(static_cast<const uint8_t*>(PB.bias) + off);
            }
        }
    }
    if (warp == NW - 1) {
        if (p.x != nullptr && lane < kSliceXBytes / 128 && (size_t)ks * kSliceXBytes + lane * 128 < (size_t)p.G * 16)
        {
            l2_prefetch_line(static_cast<const uint8_t*>(p.x) + PB.x_off + (size_t)ks * kSliceXBytes + lane * 128);
            if (p.x_gate) l2_prefetch_line(static_cast<const uint8_t*>(p.x_gate) + PB.x_off + (size_t)ks * kSliceXBytes + lane * 128);
        }
    }
    TL_STAMP(10);
    const int r = lane >> 2, q = lane & 3;
    constexpr int MB_PER_WARP = 16 / NW;                                   // 16 idx blocks of 16 per codebook
    uint32_t afrag[MB_PER_WARP][NCB][2];
    {
        const uint32_t* cb32 = static_cast<const uint32_t*>(PB.codebook);  // [NCB][256][4] words (2 values each)
        const uint64_t pol_keep = l2_policy_evict_last();
        // A fragments for this warp's idx blocks: rows 16m + r and 16m + r + 8, k = 2q, 2q+1
        #pragma unroll
        for (int mi = 0; mi < MB_PER_WARP; ++mi) {
            const int m = warp * MB_PER_WARP + mi;
            #pragma unroll
            for (int n = 0; n < NCB; ++n) {
                afrag[mi][n][0] = ldg_u32_pol(cb32 + (n * 256 + 16 * m + r) * 4 + q, pol_keep);
                afrag[mi][n][1] = ldg_u32_pol(cb32 + (n * 256 + 16 * m + r + 8) * 4 + q, pol_keep);
            }
        }
    }
    // max |codebook| over the CTA (every warp holds 1/NW of it in its A fragments): one REDUX + one smem word per warp
    __shared__ uint32_t cmax_red[NW];
    {
        uint32_t m = 0;
        #pragma unroll
        for (int mi = 0; mi < MB_PER_WARP; ++mi)
            #pragma unroll
            for (int n = 0; n < NCB; ++n) m = max(m, max(absmax_halves(afrag[mi][n][0]), absmax_halves(afrag[mi][n][1])));
        m = __reduce_max_sync(0xffffffffu, m);
        if (lane == 0) cmax_red[warp] = m;
    }
    TL_STAMP(1);
    // The next launch may become resident now that OUR weight requests are in flight (it still waits for us below).
    // Measured on B200 (4096x11008, 3 CTAs/SM): triggering at kernel entry instead costs 5.4 us per launch vs 4.0 us.
    pdl_launch_dependents();
    // X_READY: x is not the previous kernel's output and the workspace is exclusively ours (include/lutvq.h), so
    // nothing but the final y stores has to be ordered behind the previous kernel.
    const bool xready = (p.flags & LUTVQ_FLAG_X_READY) != 0;
    bool waited = !xready;
    if (!xready) pdl_wait();    // x and the workspace belong to the previous kernel until here
    TL_STAMP(2);
    // The slice's 32 activation groups (512 B = 4 lines) are requested ONCE per CTA and shared through shared memory: every
    // CTA of a slice reads the same few lines at the same moment, and when all 8 warps fetched their B fragments themselves
    // (4 scattered 4-byte loads per lane) each hot line took 16 requests per CTA at its L2 slice.
    __shared__ __align__(16) uint4 xs[kSliceGroups];
    auto load_x_slice = [&](int ks_) {
        if (threadIdx.x < kSliceGroups) {
            if (NCB == 2 && p.ll.x_ll != nullptr && (p.flags & LUTVQ_FLAG_X_READY)) {
                // dataflow launch: the words may be several microseconds away and EVERY CTA of the launch (and of the launches
                // queued behind it) waits for the same few lines.  One lane polls one word, with a pause between polls, the rest
                // of the warp waits at the barrier.  (All 32 lanes of every CTA polling at full rate saturate the L2 slice that
                // holds the lines and hold up the very stores they wait for: 2.0 ms per 7B token instead of 1.2.)
                if (threadIdx.x == 0) {
                    const unsigned ep0 = ll_epoch16(*p.ll.tick, p.ll.x_seq);
                    const uint32_t* w0 = p.ll.x_ll + (size_t)min(ks_ * kSliceGroups, p.G - 1) * 8;
                    for (unsigned spin = 0; spin < (1u << 18); ++spin) {
                        uint32_t v0;
                        asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v0) : "l"(w0) : "memory");
                        if ((v0 >> 16) == ep0) break;
                        __nanosleep(250u);
                    }
                }
                __syncwarp();
            }
            const int g = ks_ * kSliceGroups + (int)threadIdx.x;
            uint4 v = make_uint4(0u, 0u, 0u, 0u);
            if (g < p.G) {
                if (p.ll.x_ll != nullptr) {
                    // the activations are an earlier exchange's LL words in the local gather slot: poll this group's 8 words
                    const unsigned tick = *p.ll.tick;
                    v = ll_load8(p.ll.x_ll + (size_t)g * 8, ll_epoch16(tick, p.ll.x_seq), p.ll.error);
                    if (p.ll.gate_ll != nullptr) {
                        const uint4 gt = ll_load8(p.ll.gate_ll + (size_t)g * 8, ll_epoch16(tick, p.ll.gate_seq), p.ll.error);
                        v.x = silu_mul_halves(gt.x, v.x); v.y = silu_mul_halves(gt.y, v.y);
                        v.z = silu_mul_halves(gt.z, v.z); v.w = silu_mul_halves(gt.w, v.w);
                    }
                } else {
                    v = ldcg_u128(reinterpret_cast<const uint4*>(static_cast<const uint8_t*>(p.x) + PB.x_off) + g);   // L2-coherent: x may be the previous kernel's output
                    if (DT == LUTVQ_F16 && p.x_gate != nullptr) {
                        const uint4 gt = ldcg_u128(reinterpret_cast<const uint4*>(static_cast<const uint8_t*>(p.x_gate) + PB.x_off) + g);
                        v.x = silu_mul_halves(gt.x, v.x); v.y = silu_mul_halves(gt.y, v.y);
                        v.z = silu_mul_halves(gt.z, v.z); v.w = silu_mul_halves(gt.w, v.w);
                    }
                }
            }
            xs[threadIdx.x] = v;
        }
    };
    load_x_slice(ks);

    // PRMT operands (lut_layout.h): byte0 <- lm, byte1 <- code byte, bytes 2,3 <- 0.  The slice base and the
    // (c&1)*128 half-select are compile-time constants for ptxas and ride in the LDS immediate.
    uint32_t sel[4], lm[2][4];
    #pragma unroll
    for (int d = 0; d < 4; ++d) {
        sel[d] = stream_sel(lane, d);
        lm[0][d] = stream_lm(lane, 0, d);
        lm[1][d] = stream_lm(lane, 1, d);
    }
    __syncthreads();
    float cmax8;
    {
        uint32_t m = 0;
        #pragma unroll
        for (int i = 0; i < NW; ++i) m = max(m, cmax_red[i]);
        cmax8 = 8.f * abs_half_bits_to_float<DT>(m);
    }
    const bool b3 = (lane >> 3) & 1, b4 = (lane >> 4) & 1;
    const bool odd = r & 1;
    const bool round16 = (p.flags & LUTVQ_FLAG_LUT_FP16) != 0;
    bool first_seg = true;
    float fx_scale = 1.f, fx_inv = 1.f;                                  // this segment's 2^e and 2^-e

    // ---- split-k completion: see the accumulator word comment above ----
    using D = Dtype<DT>;
    using T = typename D::T;
    // A contribution's atomic is issued at the end of its set and its return value is examined PF sets later, so the
    // L2 round trip hides behind lookups.  Pending state per slot: the returned word, the partial and the row.
    static_assert(PA % PF == 0, "the pending-atomic ring must be a whole number of code rings");
    unsigned long long pend_old[PA];
    float pend_keep[PA];
    int pend_row[PA];
    bool pend_valid[PA];
    #pragma unroll
    for (int i = 0; i < PA; ++i) { pend_old[i] = 0; pend_keep[i] = 0.f; pend_row[i] = 0; pend_valid[i] = false; }
    const unsigned last_count = (unsigned)p.n_k - 1u;
    auto finish_row = [&](unsigned long long old, float keep, int row) {
        if ((unsigned)(old & 255ull) == last_count) {              // this lane delivered the last partial of `row`
            if (!waited) { pdl_wait(); waited = true; }            // y stores are ordered behind the previous kernel
            const float total = fx_total<DT>(old + fx_contribution<DT>(keep));
            asm volatile("st.global.cg.b64 [%0], %1;" :: "l"(PB.acc + row), "l"(0ull) : "memory");   // hand the word back
            if (row < PB.O) {
                T rr = D::from_f(total);                                               // Code2x8MatVec writes round(acc)   (cuda_kernel.cu:221-231)
                const uint64_t pol_keep = l2_policy_evict_last();
                // scale / bias were pulled into L2 by the prologue; only the last arriver (1 in n_k) pays this round trip
                if (PB.scales) { const uint16_t b = ldg_u16_pol(static_cast<const T*>(PB.scales) + row, pol_keep); rr = D::mul(rr, *reinterpret_cast<const T*>(&b)); }   // flat_output *= scales (cuda_kernel.cpp:101)
                if (PB.bias)   { const uint16_t b = ldg_u16_pol(static_cast<const T*>(PB.bias) + row, pol_keep);   rr = D::add(rr, *reinterpret_cast<const T*>(&b)); }   // flat_output += bias   (cuda_kernel.cpp:103)
                store_y<DT>(p, PB, row, rr);
            }
        }
    };

    while (pos < end) {
        // ---------------- K1: build this worker's 64 KiB LUT for slice ks ----------------
        {
            // B fragments: column r of N-tile t is local group build_group<NCB>(t, r); k = 2q, 2q+1
            constexpr int NT = 8 / NCB;                                // N-tiles of 8 groups per slice
            uint32_t bfrag[NT];
            #pragma unroll
            for (int t = 0; t < NT; ++t) bfrag[t] = reinterpret_cast<const uint32_t*>(xs)[build_group<NCB>(t, r) * 4 + q];
            {   // the warp's 32 lanes x NT words are exactly the slice's activations
                uint32_t m = 0;
                #pragma unroll
                for (int t = 0; t < NT; ++t) m = max(m, absmax_halves(bfrag[t]));
                m = __reduce_max_sync(0xffffffffu, m);
                fx_scales(cmax8 * abs_half_bits_to_float<DT>(m), fx_scale, fx_inv);
            }
            #pragma unroll
            for (int mi = 0; mi < MB_PER_WARP; ++mi) {
                const int m = warp * MB_PER_WARP + mi;
                float c[8][4];                                         // MMA i = n * NT + t: codebook n x N-tile t
                #pragma unroll
                for (int n = 0; n < NCB; ++n)
                    #pragma unroll
                    for (int t = 0; t < NT; ++t) {
                        float (&ci)[4] = c[n * NT + t];
                        ci[0] = ci[1] = ci[2] = ci[3] = 0.f;
                        MmaK8<DT>::run(ci, afrag[mi][n][0], afrag[mi][n][1], bfrag[t]);
                    }
                #pragma unroll
                for (int i = 0; i < 8; ++i)
                    #pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        float v = c[i][k];
                        if (round16) v = __half2float(__float2half_rn(v));
                        c[i][k] = fmaf(v, fx_scale, kFxMagic);             // float bits = 0x4B400000 + round(v * 2^e)
                    }
                // The 16-byte store (idx = 16m + r + 8h, chunk q, word w_j) holds byte positions 16q + 4w_j + beta, beta = 0..3;
                // build_val<NCB>(j, beta, h) names the MMA output that belongs there (lut_layout.h).
                auto val = [&](int j, int beta, int h) -> float { return c[build_val_mma<NCB>(j, beta)][2 * h + build_val_col<NCB>(j, beta)]; };
                #pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int idx = 16 * m + r + 8 * h;
                    #pragma unroll
                    for (int jp = 0; jp < 2; ++jp) {
                        const int j0 = 2 * jp, j1 = 2 * jp + 1;
                        // first store: odd lanes take tile j1, even lanes tile j0 (8 distinct 16-byte columns per phase)
                        float f0 = odd ? val(j1, 0, h) : val(j0, 0, h);
                        float f1 = odd ? val(j1, 1, h) : val(j0, 1, h);
                        float f2 = odd ? val(j1, 2, h) : val(j0, 2, h);
                        float f3 = odd ? val(j1, 3, h) : val(j0, 3, h);
                        float s0 = odd ? val(j0, 0, h) : val(j1, 0, h);
                        float s1 = odd ? val(j0, 1, h) : val(j1, 1, h);
                        float s2 = odd ? val(j0, 2, h) : val(j1, 2, h);
                        float s3 = odd ? val(j0, 3, h) : val(j1, 3, h);
                        const uint32_t first  = lut_base + build_store_offset(idx, odd ? j1 : j0, q);
                        const uint32_t second = lut_base + build_store_offset(idx, odd ? j0 : j1, q);
                        sts_v4(first, f0, f1, f2, f3);
                        sts_v4(second, s0, s1, s2, s3);
                    }
                }
            }
        }
        __syncthreads();
        if (first_seg) TL_STAMP(3);
        first_seg = false;

        // ---------------- K2: stream codes, gather-accumulate ----------------
        {
            const int n_seg = hi - lo;
            int v = warp + seg_rot;                                  // rotated local index of this warp's next set
            if (v >= n_seg) v -= n_seg;
            int vnext = v;                                           // ... and of the set PF iterations ahead
            #pragma unroll
            for (int i = 0; i < PF; ++i) { vnext += NW; if (vnext >= n_seg) vnext -= n_seg; }
            for (int li = warp; li < n_seg; li += PA * NW) {
                #pragma unroll
                for (int sa = 0; sa < PA; ++sa) {
                    constexpr int kDummy = 0; (void)kDummy;
                    const int st = sa % PF;                          // code-ring slot (compile-time after unrolling)
                    if (li + sa * NW < n_seg) {                      // warp-uniform
                        const int row = (lo + v) * kSetRows + lane;
                        uint32_t acc[4];
#if LUTVQ_DBG
                        if (p.flags & 0x100u) {                      // DEBUG: no lookups (pure code streaming)
                            #pragma unroll
                            for (int t = 0; t < 4; ++t) acc[t] = buf[st].w[t][0] ^ buf[st].w[t][1] ^ buf[st].w[t][2] ^ buf[st].w[t][3];
                        } else
#endif
                        {
                            #pragma unroll
                            for (int c = 0; c < 4; ++c)
                                #pragma unroll
                                for (int d = 0; d < 4; d += 2)
                                    #pragma unroll
                                    for (int t = 0; t < 4; ++t) {
                                        const uint32_t ad0 = __byte_perm(buf[st].w[t][c], lm[c >> 1][d], sel[d]);
                                        const uint32_t ad1 = __byte_perm(buf[st].w[t][c], lm[c >> 1][d + 1], sel[d + 1]);
                                        const uint32_t v0 = *reinterpret_cast<const uint32_t*>(smem_raw + ad0 + stream_imm(c));
                                        const uint32_t v1 = *reinterpret_cast<const uint32_t*>(smem_raw + ad1 + stream_imm(c));
                                        if (c == 0 && d == 0) acc[t] = v0 + v1;
                                        else                  acc[t] = acc[t] + v0 + v1;       // IADD3
                                    }
                        }
                        // refill this stage with the set PF iterations ahead before the (short) reduction
                        if (li + (sa + PF) * NW < n_seg LUTVQ_DBG_AND(!(p.flags & 0x200u))) load_set(buf[st], ks, lo + vnext);
                        // transposed reduce over the 4 chunk lanes (lane bits 3,4): lane ends with tile (2*b4+b3), row lane&7
                        uint32_t k0 = b4 ? acc[2] : acc[0], k1 = b4 ? acc[3] : acc[1];
                        uint32_t g0 = b4 ? acc[0] : acc[2], g1 = b4 ? acc[1] : acc[3];
                        k0 += __shfl_xor_sync(0xffffffffu, g0, 16);
                        k1 += __shfl_xor_sync(0xffffffffu, g1, 16);
                        uint32_t keep_u = b3 ? k1 : k0, give = b3 ? k0 : k1;
                        keep_u += __shfl_xor_sync(0xffffffffu, give, 8);
                        const float keep = (float)(int)(keep_u - kFxMagic64) * fx_inv;     // 64 lookups' magic offsets, mod 2^32
                        if (p.n_k == 1) {
                            if (!waited) { pdl_wait(); waited = true; }
                            if (row < PB.O) {
                                T rr = D::from_f(keep);
                                if (PB.scales) rr = D::mul(rr, static_cast<const T*>(PB.scales)[row]);
                                if (PB.bias)   rr = D::add(rr, static_cast<const T*>(PB.bias)[row]);
                                store_y<DT>(p, PB, row, rr);
                            }
                        } else {
                            // finish the atomic this slot issued PA sets ago (its L2 round trip, ~1.3 us under load, has
                            // long completed), then issue this set's
                            if (pend_valid[sa]) finish_row(pend_old[sa], pend_keep[sa], pend_row[sa]);
                            pend_row[sa] = row; pend_valid[sa] = true; pend_keep[sa] = keep;
#if LUTVQ_DBG
                            if (p.flags & 0x400u) { pend_old[sa] = 0; if (row < PB.O) static_cast<T*>(PB.y)[row] = D::from_f(keep); }   // DEBUG: no atomics
                            else
#endif
                            pend_old[sa] = atom_add_u64(PB.acc + row, fx_contribution<DT>(keep));
                        }
                        v += NW;     if (v >= n_seg) v -= n_seg;
                        vnext += NW; if (vnext >= n_seg) vnext -= n_seg;
                    }
                }
            }
        }
        pos += hi - lo;
        if (pos < end) {                                           // next segment (linear plan only)
            ks = pos / PB.n_sets;
            lo = pos - ks * PB.n_sets;
            hi = min(PB.n_sets, lo + (end - pos));
            seg_rot = rot_of(ks, hi - lo);
            #pragma unroll
            for (int i = 0; i < PF; ++i)
                if (lo + warp + i * NW < hi) load_set(buf[i], ks, lo + wrap(warp + i * NW + seg_rot, hi - lo));
            load_x_slice(ks);                                      // the old slice's xs was last read before the post-build barrier
            __syncthreads();                                       // everyone is done reading this slice's LUT; xs is complete
        }
    }
    TL_STAMP(4);

    #pragma unroll
    for (int i = 0; i < PA; ++i)
        if (pend_valid[i]) finish_row(pend_old[i], pend_keep[i], pend_row[i]);
    TL_STAMP(5);
    if (!waited) pdl_wait();                                       // completion stays stream-ordered
    TL_STAMP(6);
    if (p.n_push > 0) {
        // All of this CTA's y / peer stores are issued; make them visible system-wide, then count the CTA in.  The CTA
        // that completes the count has (by fence cumulativity) every CTA's stores ordered before its own flag writes.
        __syncthreads();
        __shared__ unsigned last_flag;
        if (threadIdx.x == 0) {
            __threadfence_system();
            const unsigned prev = atomicAdd(p.cta_arrivals, 1u);
            last_flag = (prev == gridDim.x - 1) ? 1u : 0u;
        }
        __syncthreads();
        if (last_flag) {
            if (threadIdx.x == 0) *p.cta_arrivals = 0u;                               // self-cleaning, like the accumulators
            __threadfence_system();
            const unsigned epoch = *p.push_sent + 1u;
            if ((int)threadIdx.x < p.n_push)
                asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p.push_flag[threadIdx.x]), "r"(epoch) : "memory");
            __syncthreads();
            if (threadIdx.x == 0) *p.push_sent = epoch;
        }
    }
}

// [k-slice][set][tile][lane] x 16 B register images from reference-layout codes (see file header)
__global__ void prepack_2x8_kernel(const uint8_t* __restrict__ codes, uint4* __restrict__ packed, int O, int G,
                                   int n_k, int n_sets_total) {
    const size_t total = (size_t)n_k * n_sets_total * 128;
    const size_t row_bytes = (size_t)G * 2;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int lane = (int)(i & 31), t = (int)((i >> 5) & 3);
        const size_t ss = i >> 7;
        const int set = (int)(ss % n_sets_total), ks = (int)(ss / n_sets_total);
        const int row = set * kSetRows + t * kTileRows + stream_row_in_tile(lane);
        const size_t off = (size_t)ks * kSliceTables + 16 * stream_chunk(lane);
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (row < O && off < row_bytes) v = *reinterpret_cast<const uint4*>(codes + (size_t)row * row_bytes + off);
        if (stream_first_half(lane)) v = make_uint4(v.z, v.w, v.x, v.y);
        packed[i] = v;
    }
}

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------
struct GemvTuning { int nw = 0, pf = 0, pa = 0; };   // 0 = choose per shape (launch_tuned)
static GemvTuning g_tuning;
static int g_dep_workers_per_sm = 2;   // workers (CTAs) per SM of a DEPENDENT launch, see gemv_2x8_launch_group
void gemv_set_dependent_workers(int m) { g_dep_workers_per_sm = m < 1 ? 1 : (m > 2 ? 2 : m); }
static unsigned long long* g_timeline = nullptr;   // debug: [64 launches][256 CTAs][8] stamps
static unsigned g_timeline_seq = 0;
void gemv_set_timeline(void* buf) { g_timeline = static_cast<unsigned long long*>(buf); g_timeline_seq = 0; }

struct GemvPlan { int n_k, n_sets, W, aligned, q, r, max_range; };

// One worker (CTA) per SM.  Cost model in units of one 32-row set of gather work; a LUT build costs about 10.
static GemvPlan plan_gemv(int O, int G, int sm_count, int ncb = 2) {
    GemvPlan pl;
    pl.n_k = ceil_div(G * ncb, kSliceTables);
    pl.n_sets = ceil_div(O, kSetRows);
    const long long T = (long long)pl.n_k * pl.n_sets;
    pl.W = (int)(T < sm_count ? T : sm_count);
    if (pl.W > 1024) pl.W = 1024;                                  // worker_range's magic division is exact up to 1024
    const int kBuildCost = 10;
    const long long per = (T + pl.W - 1) / pl.W;
    const long long lin_builds = 1 + (per + pl.n_sets - 1) / pl.n_sets;            // a range of `per` items touches at most this many slices
    const long long cost_linear = per + kBuildCost * (lin_builds > pl.n_k ? pl.n_k : lin_builds);
    pl.aligned = 0; pl.q = 0; pl.r = 0;
    pl.max_range = (int)per;
    if (pl.n_k <= pl.W) {
        const int q = pl.W / pl.n_k;
        const long long cost_aligned = (pl.n_sets + q - 1) / q + kBuildCost;
        if (cost_aligned <= cost_linear) {
            pl.aligned = 1; pl.q = q; pl.r = pl.W % pl.n_k;
            pl.max_range = (pl.n_sets + q - 1) / q;
        }
    }
    return pl;
}

size_t gemv_workspace_bytes(int O, int G) {
    (void)G;
    const size_t O_pad = (size_t)ceil_div(O, kSetRows) * kSetRows;
    return kWsHeaderBytes + O_pad * sizeof(unsigned long long);
}

size_t gemv_packed_bytes(int O, int G) {
    const int n_k = ceil_div(G * 2, kSliceTables);
    return (size_t)n_k * ceil_div(O, kSetRows) * 2048;
}

int gemv_prepack_launch(const void* codes, void* packed, int O, int I, cudaStream_t stream, void* zero, size_t zero_bytes) {
    (void)zero; (void)zero_bytes;
    const int G = I / 8;
    const int n_k = ceil_div(G * 2, kSliceTables), n_sets_total = ceil_div(O, kSetRows);
    const size_t total = (size_t)n_k * n_sets_total * 128;
    int grid = (int)((total + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    prepack_2x8_kernel<<<grid, 256, 0, stream>>>(static_cast<const uint8_t*>(codes), static_cast<uint4*>(packed), O, G, n_k, n_sets_total);
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

template <int DT, bool PACKED, int NW, int PF, int PA, int MINB = 2, int NCB = 2>
static int launch_variant(const GemvParams& p, int grid, size_t smem, cudaStream_t stream, bool pdl) {
    auto kern = lut_gemv_2x8_kernel<DT, PACKED, NW, PF, PA, MINB, NCB>;
    static thread_local int attr_device = -1;
    int dev = 0;
    LUTVQ_CUDA_OK(cudaGetDevice(&dev));
    if (attr_device != dev) {
        LUTVQ_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSliceBytes));
        attr_device = dev;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(NW * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    LUTVQ_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, p));
    return LUTVQ_OK;
}

template <int DT, bool PACKED>
static int launch_tuned(const GemvParams& p, int grid, size_t smem, cudaStream_t stream, bool pdl) {
    GemvTuning t = g_tuning;
    if (t.nw == 0) {
        // Measured on B200 (X_READY ring, packed codes): 3 CTAs/SM with a 1-deep code ring wins while a row has few
        // split-k arrivals (4096x11008: 3.75 vs 4.39 us; 4096x4096: 2.39 vs 2.51 us); with many k-slices the exposed
        // atomic round trip of the shallow ring loses to 2 CTAs/SM with 2-deep rings (11008x4096: 6.08 vs 5.07 us).
        // That is the INDEPENDENT (X_READY) ring.  A dependent chain (decode) runs one launch at a time: there the third
        // resident CTA buys nothing, the deeper ring wins on every shape, and alternating between the two variants
        // (different register / occupancy footprints) costs the PDL overlap: Llama-2-7B layer, 4 GEMV launches chained:
        // 51.7 us mixed, 36.6 all 3-CTA, 31.0 all 2-CTA (tools/decode_parts.py).
        t.nw = 8;
        if (p.n_k <= 24 && (p.flags & LUTVQ_FLAG_X_READY) && p.ll.x_ll == nullptr) { t.pf = 1; t.pa = 1; } else { t.pf = 2; t.pa = 2; }
        // dependent launches: the split-k atomic's return is examined 4 sets later instead of 2 (ncu on a batch-8 launch: a quarter
        // of the stall samples sat on that return): Llama-2-7B decode 812 -> 825 tok/s, batch-1 dependent 6.63 -> 6.42 us
        if (!(p.flags & LUTVQ_FLAG_X_READY)) t.pa = 4;
    }
    if (t.nw == 8 && t.pf == 1 && t.pa == 1)  return launch_variant<DT, PACKED, 8, 1, 1, 3>(p, grid, smem, stream, pdl);   // 3 CTAs / SM
    if (t.nw == 8 && t.pf == 2 && t.pa == 2)  return launch_variant<DT, PACKED, 8, 2, 2, 2>(p, grid, smem, stream, pdl);   // 2 CTAs / SM, deeper ring
    if (t.nw == 8 && t.pf == 2 && t.pa == 4)  return launch_variant<DT, PACKED, 8, 2, 4>(p, grid, smem, stream, pdl);   // atomic returns examined 4 sets later
#if LUTVQ_DBG
    if (t.nw == 8 && t.pf == 2 && t.pa == 6)  return launch_variant<DT, PACKED, 8, 2, 6>(p, grid, smem, stream, pdl);
    if (t.nw == 8 && t.pf == 1 && t.pa == 4)  return launch_variant<DT, PACKED, 8, 1, 4>(p, grid, smem, stream, pdl);
    if (t.nw == 8 && t.pf == 1 && t.pa == 6)  return launch_variant<DT, PACKED, 8, 1, 6>(p, grid, smem, stream, pdl);
    if (t.nw == 16 && t.pf == 1 && t.pa == 3) return launch_variant<DT, PACKED, 16, 1, 3>(p, grid, smem, stream, pdl);
    if (t.nw == 4 && t.pf == 2 && t.pa == 8)  return launch_variant<DT, PACKED, 4, 2, 8, 3>(p, grid, smem, stream, pdl);
    if (t.nw == 8 && t.pf == 1 && t.pa == 2)  return launch_variant<DT, PACKED, 8, 1, 2, 3>(p, grid, smem, stream, pdl);
    if (t.nw == 8 && t.pf == 2 && t.pa == 8)  return launch_variant<DT, PACKED, 8, 2, 2, 3>(p, grid, smem, stream, pdl);   // "8,2,8" = (8,2,2) at 3 CTAs/SM
#endif
    return fail(LUTVQ_EINVAL, "no kernel variant for tuning nw=%d pf=%d pa=%d", t.nw, t.pf, t.pa);
}

// AQLM Kx8 with K = NCB != 2 (reference layout only): the 2-CTA / 2-deep-ring variant, whatever the tuning knobs say
template <int DT>
static int launch_kx8(const GemvParams& p, int ncb, int grid, size_t smem, cudaStream_t stream, bool pdl) {
    if (ncb == 1) return launch_variant<DT, false, 8, 2, 2, 2, 1>(p, grid, smem, stream, pdl);
    if (ncb == 4) return launch_variant<DT, false, 8, 2, 2, 2, 4>(p, grid, smem, stream, pdl);
    if (ncb == 8) return launch_variant<DT, false, 8, 2, 2, 1, 8>(p, grid, smem, stream, pdl);
    return fail(LUTVQ_EUNSUPPORTED, "the fused LUT-GEMV takes 1, 2, 4 or 8 codebooks per group, got %d", ncb);
}

// worker w of a problem planned as `pl` -> its item range [begin, end) over (slice, set) pairs, slice-major.
// Aligned plan: slices < r have q+1 workers, the others q; a slice's n_sets sets are dealt base (+1 for the first rem).
static void host_worker_range(const GemvPlan& pl, int w, int& begin, int& end) {
    if (pl.aligned) {
        const int big = pl.r * (pl.q + 1);
        const bool hi_part = w >= big;
        const int nw = hi_part ? pl.q : pl.q + 1;
        const int w2 = hi_part ? w - big : w;
        const int s2 = w2 / nw, j = w2 - s2 * nw;
        const int s = (hi_part ? pl.r : 0) + s2;
        const int base = pl.n_sets / nw, rem = pl.n_sets % nw;
        begin = s * pl.n_sets + j * base + (j < rem ? j : rem);
        end   = begin + base + (j < rem ? 1 : 0);
    } else {
        const long long T = (long long)pl.n_k * pl.n_sets;
        begin = (int)((T * w) / pl.W);
        end   = (int)((T * (w + 1)) / pl.W);
    }
}

size_t gemv_group_workspace_bytes(int n, const int* O) {
    size_t b = kWsHeaderBytes;
    for (int i = 0; i < n; ++i) b += (size_t)ceil_div(O[i], kSetRows) * kSetRows * sizeof(unsigned long long);
    return b;
}

// n linears that share x (and prob_k) in ONE launch: y_i = epilogue_i(W_i(codes_i) . x).  The workers (one CTA per SM)
// are dealt to the problems in proportion to their rows; each problem is then planned exactly like a single GEMV on
// its share of the SMs.  Validation is done by the caller (api.cu / shard.cu).
// x_gate != null (fp16): the activations are silu(x_gate) * x, evaluated in the kernel's x loads.
static int gemv_launch_impl(int n, const GemvItem* items, int packed, const void* x, const void* x_gate, int I, int dtype, unsigned flags,
                            void* ws, size_t ws_bytes, cudaStream_t stream, const GemvPush* push, const GemvLL* ll = nullptr, int ncb = 2) {
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    LUTVQ_REQUIRE(n >= 1 && n <= kMaxProblems, "a group holds 1..%d linears, got %d", kMaxProblems, n);
    // An independent (X_READY) ring fills the SMs by overlapping consecutive launches, one worker per SM each.  A dependent
    // launch runs alone: it may be planned over g_dep_workers_per_sm workers per SM (2 fit with the 2-CTA variant).
    // A dataflow launch (X_READY with LL activations, shard.cu) is a link of a dependent chain too: only one of the co-resident
    // launches has its data at any time, so it is planned -- and tuned -- like a dependent launch.
    const bool independent = (flags & LUTVQ_FLAG_X_READY) && !(ll && ll->x_ll);
    const int workers = di.sm_count * (independent ? 1 : g_dep_workers_per_sm);
    LUTVQ_REQUIRE(n == 1 || workers <= kRangeTable, "grouped launches need the tabulated plan (<= %d workers)", kRangeTable);
    LUTVQ_REQUIRE(n == 1 || push == nullptr, "the fused push epilogue takes one linear per launch");
    const int G = I / 8;
    int Os[kMaxProblems];
    long long sets_total = 0;
    for (int i = 0; i < n; ++i) { Os[i] = items[i].O; sets_total += ceil_div(items[i].O, kSetRows); }
    const size_t need = gemv_group_workspace_bytes(n, Os);
    LUTVQ_REQUIRE(ws != nullptr && ws_bytes >= need, "workspace too small: need %zu bytes, got %zu", need, ws_bytes);
    LUTVQ_REQUIRE(aligned16(ws), "workspace must be 16-byte aligned");
    GemvParams p;
    p.n_prob = n; p.x = x; p.G = G;
    p.x_gate = x_gate;
    p.ll = ll ? *ll : GemvLL{};
    LUTVQ_REQUIRE(x != nullptr || p.ll.x_ll != nullptr, "null x pointer");
    LUTVQ_REQUIRE(p.ll.gate_ll == nullptr || dtype == LUTVQ_F16, "the fused SwiGLU input is fp16 only");
    LUTVQ_REQUIRE(p.x_gate == nullptr || (dtype == LUTVQ_F16 && !(flags & LUTVQ_FLAG_X_READY)), "the fused SwiGLU input is fp16 only and never X_READY");
    p.has_table = 0;
    int W_total = 0, left = workers;
    uint8_t* accp = static_cast<uint8_t*>(ws) + kWsHeaderBytes;
    for (int i = 0; i < kMaxProblems; ++i) p.prob[i] = GemvProblem{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, 0, 0u};
    for (int i = 0; i < n; ++i) {
        // SM share of problem i: proportional to its row sets, at least one, the last problem takes the remainder
        const int n_sets_i = ceil_div(items[i].O, kSetRows);
        int share = (i == n - 1) ? left : (int)((long long)workers * n_sets_i / sets_total);
        if (share < 1) share = 1;
        if (share > left - (n - 1 - i)) share = left - (n - 1 - i);
        left -= share;
        GemvPlan pl = plan_gemv(items[i].O, G, share, ncb);
        if (pl.n_k > 255) return fail(LUTVQ_EUNSUPPORTED, "prob_k=%d needs %d k-slices; the split-k arrival counter holds 255 (prob_k <= 65280)", I, pl.n_k);
        p.n_k = pl.n_k;
        GemvProblem& pb = p.prob[i];
        pb.codes = static_cast<const uint8_t*>(items[i].codes); pb.codebook = items[i].codebook; pb.y = items[i].y;
        pb.scales = items[i].scales; pb.bias = items[i].bias;
        pb.acc = reinterpret_cast<unsigned long long*>(accp);
        pb.O = items[i].O; pb.n_sets = pl.n_sets;
        LUTVQ_REQUIRE(items[i].x_off < (1ull << 32) && items[i].x_off % 16 == 0, "activation row offset must be 16-byte aligned and below 4 GiB");
        pb.x_off = (unsigned)items[i].x_off;
        accp += (size_t)pl.n_sets * kSetRows * sizeof(unsigned long long);
        if (W_total + pl.W <= kRangeTable LUTVQ_DBG_AND(!(flags & 0x1000u))) {
            for (int w = 0; w < pl.W; ++w) {
                int b, e;
                host_worker_range(pl, w, b, e);
                p.tab_begin[W_total + w] = b; p.tab_end[W_total + w] = e;
                p.tab_ks[W_total + w] = (unsigned short)(b < e ? b / pl.n_sets : 0);
                p.tab_prob[W_total + w] = (unsigned char)i;
            }
            p.has_table = 1;
        } else {
            LUTVQ_REQUIRE(n == 1, "grouped launches need the tabulated plan");
            p.has_table = 0;                                        // device fallback: linear plan over W CTAs
        }
        W_total += pl.W;
    }
    p.n_push = 0; p.push_sent = nullptr;
    p.cta_arrivals = reinterpret_cast<unsigned*>(ws);                // first word of the reserved workspace header
    for (int r = 0; r < kMaxPeers; ++r) { p.push_y[r] = nullptr; p.push_flag[r] = nullptr; }
    if (push != nullptr && push->n > 0) {
        LUTVQ_REQUIRE(push->n <= kMaxPeers, "at most %d push targets", kMaxPeers);
        p.n_push = push->n; p.push_sent = push->sent;
        for (int r = 0; r < push->n; ++r) { p.push_y[r] = push->y[r]; p.push_flag[r] = push->flag[r]; }
    }
    const bool pdl = !(flags & LUTVQ_FLAG_NO_PDL);
    p.flags = pdl ? flags : (flags & ~LUTVQ_FLAG_X_READY);        // the X_READY overlap only exists under PDL
    p.tl = (g_timeline && W_total <= 256) ? g_timeline + (size_t)(g_timeline_seq++ % 64) * 256 * 16 : nullptr;
    const size_t smem = kSliceBytes;
    if (ncb != 2) {
        LUTVQ_REQUIRE(!packed, "Kx8 formats other than 2x8 run on reference-layout codes");
        if (dtype == LUTVQ_F16)  return launch_kx8<LUTVQ_F16>(p, ncb, W_total, smem, stream, pdl);
        if (dtype == LUTVQ_BF16) return launch_kx8<LUTVQ_BF16>(p, ncb, W_total, smem, stream, pdl);
        return fail(LUTVQ_EUNSUPPORTED, "dtype %d not supported (fp16=0, bf16=1)", dtype);
    }
    if (dtype == LUTVQ_F16)  return packed ? launch_tuned<LUTVQ_F16, true>(p, W_total, smem, stream, pdl)  : launch_tuned<LUTVQ_F16, false>(p, W_total, smem, stream, pdl);
    if (dtype == LUTVQ_BF16) return packed ? launch_tuned<LUTVQ_BF16, true>(p, W_total, smem, stream, pdl) : launch_tuned<LUTVQ_BF16, false>(p, W_total, smem, stream, pdl);
    return fail(LUTVQ_EUNSUPPORTED, "dtype %d not supported (fp16=0, bf16=1)", dtype);
}

int gemv_2x8_launch_group(int n, const GemvItem* items, const void* x, const void* x_gate, int I, int dtype, unsigned flags,
                          void* ws, size_t ws_bytes, cudaStream_t stream, const GemvPush* push, const GemvLL* ll) {
    return gemv_launch_impl(n, items, 1, x, x_gate, I, dtype, flags, ws, ws_bytes, stream, push, ll);
}

// One GEMV on a packed code image: y = epilogue(W(codes) . x).
int gemv_2x8_launch(const void* packed, const void* x, void* y, const void* codebook, const void* scales,
                    const void* bias, int O, int I, int dtype, unsigned flags, void* ws, size_t ws_bytes,
                    cudaStream_t stream, const GemvPush* push) {
    const GemvItem it{packed, codebook, y, scales, bias, O};
    return gemv_launch_impl(1, &it, 1, x, nullptr, I, dtype, flags, ws, ws_bytes, stream, push);
}

// One GEMV on reference-layout codes int8 [O][G][2] (the kernel's PACKED = false build: a warp's 128-bit load touches
// 8 rows x 64 B; no re-layout pass, no staging area).
int gemv_2x8_launch_ref(const void* codes, const void* x, void* y, const void* codebook, const void* scales,
                        const void* bias, int O, int I, int dtype, unsigned flags, void* ws, size_t ws_bytes,
                        cudaStream_t stream) {
    const GemvItem it{codes, codebook, y, scales, bias, O};
    return gemv_launch_impl(1, &it, 0, x, nullptr, I, dtype, flags, ws, ws_bytes, stream, nullptr);
}

// One GEMV of an AQLM Kx8 layer (K = ncb codebooks of 256 entries per group of 8 inputs) on reference-layout codes
// int8 [O][I/8][ncb]; rows must be whole 16-byte chunks ((I/8) * ncb % 16 == 0).  Same kernel as 2x8, other LUT build.
int gemv_kx8_launch_ref(const void* codes, const void* x, void* y, const void* codebook, const void* scales,
                        const void* bias, int O, int I, int ncb, int dtype, unsigned flags, void* ws, size_t ws_bytes,
                        cudaStream_t stream) {
    const GemvItem it{codes, codebook, y, scales, bias, O};
    return gemv_launch_impl(1, &it, 0, x, nullptr, I, dtype, flags, ws, ws_bytes, stream, nullptr, nullptr, ncb);
}

// `rows` (1..kMaxProblems) activation rows against ONE layer in one launch: the rows are the "problems" of a grouped launch -- each
// gets a share of the SMs, its own split-k words and its own x row (x_off); they stream the same codes at the same pace, so the
// second..last readers of a code line hit L2.  The kernel's fixed part (prologue, dependency wait, LUT build latency, tail) is
// paid once per launch instead of once per row.  Workspace: gemv_rows_workspace_bytes(rows, O).  Results are bit-identical to
// `rows` single-row launches (integer split-k sums).
int gemv_2x8_launch_rows(const void* codes, int packed, const void* x, void* y, const void* codebook, const void* scales, const void* bias,
                         int rows, int O, int I, int dtype, unsigned flags, void* ws, size_t ws_bytes, cudaStream_t stream, int ncb) {
    LUTVQ_REQUIRE(rows >= 1 && rows <= kMaxProblems, "1..%d rows per launch, got %d", kMaxProblems, rows);
    GemvItem items[kMaxProblems];
    for (int b = 0; b < rows; ++b)
        items[b] = GemvItem{codes, codebook, static_cast<uint8_t*>(y) + (size_t)b * O * 2, scales, bias, O, (size_t)b * I * 2};
    return gemv_launch_impl(rows, items, packed, x, nullptr, I, dtype, flags, ws, ws_bytes, stream, nullptr, nullptr, ncb);
}
int gemv_max_rows_per_launch() { return kMaxProblems; }
size_t gemv_rows_workspace_bytes(int rows, int O) {
    return kWsHeaderBytes + (size_t)rows * ceil_div(O, kSetRows) * kSetRows * sizeof(unsigned long long);
}

size_t gemv_packed_workspace_bytes(int O, int G) { return gemv_workspace_bytes(O, G); }

void gemv_set_tuning(int nw, int pf, int pa) {
    if (nw > 0) g_tuning.nw = nw;
    if (pf > 0) g_tuning.pf = pf;
    if (pa > 0) g_tuning.pa = pa;
}
void gemv_get_plan(int O, int I, int* n_k, int* n_chunks, int* rows_per_chunk) {
    DeviceInfo di;
    int sm = 148;
    if (get_device_info(&di) == 0) sm = di.sm_count;
    GemvPlan pl = plan_gemv(O, I / 8, sm);
    if (n_k) *n_k = pl.n_k;
    if (n_chunks) *n_chunks = pl.W;
    if (rows_per_chunk) *rows_per_chunk = pl.max_range * kSetRows;
}

}  // namespace lutvq
