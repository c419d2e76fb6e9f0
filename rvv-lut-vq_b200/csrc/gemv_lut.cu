// gemv_lut.cu -- fused LUT-build + code-indexed gather-accumulate GEMV for AQLM 2x8 codes (sm_100a), round-2 design.
//
// Replaces Code2x8MatVec (reference benchmark/AQLM/inference_lib/src/aqlm/inference_kernels/
// cuda_kernel.cu:144-233) with the LUT formulation of the golden model (vq_dataflow_sim/vq.py:280-298):
//
//   K1  L[g][n][k] = sum_j C[n][k][j] * x[8g+j]      built per CTA for ITS 64-group k-slice with mma.sync.m16n8k8
//                                                     (fp32 accumulate), stored as 16-bit fixed point: 128 tables in
//                                                     64 KiB of shared memory (lut_layout.h)
//   K2  y[o] = sum_g sum_n L[g][n][code(o,g,n)]       packed codes streamed HBM -> shared memory by 1-D bulk async copies
//                                                     (cp.async.bulk + mbarrier, a private ring per warp), then per code
//                                                     byte one PRMT + one LDS.U16 + half an IADD3, conflict-free for any
//                                                     index pattern; warp-shuffle transposed reduction
//
// Fixed-point LUT.  One scale per GEMV: S = 32704 / bound with bound = max_k |C[n][k][:]|_2 * max_g |x[8g..8g+7]|_2 >= |L|
// (Cauchy-Schwarz), computed identically by every CTA from (x, codebook) alone -- so the result does not depend on the
// work decomposition.  entry = round(L * S) + 32768 as u16 (the conversion is one FFMA with the magic constant
// 1.5 * 2^23 + 32768); sums of entries are exact integers, order-free => bit-reproducible.  Quantisation step
// bound / 32704: about 4x finer than the fp16 LUT of the golden model (vq.py:285) for N(0,1) data.
//
// Work decomposition.  Items = (k-slice, 32-row set), slice-major; one CTA ("worker") per SM owns a contiguous item
// range: aligned (workers dealt to slices, one LUT build each) or linear (balanced, may straddle slices).  The ranges
// ride in the kernel parameters.
//
// Cross-CTA split-k reduction: one 64-bit word per PAIR of output rows:  [arrivals:6][row 2i+1 : 29][row 2i : 29].
// Each (row pair, k-slice) partial is ONE integer atomic add at the word's L2 home; the contributor that sees
// arrivals == n_k - 1 in the returned word holds both complete sums, applies the reference epilogue order (round to
// dtype, * scale, + bias; cuda_kernel.cpp:95-111), writes y and zeroes the word.  Nothing polls; the workspace is
// self-cleaning.  (Round 1 used one word per row and 64 tables per slice: 4x the atomics.)
//
// Stream overlap (programmatic dependent launch).  Codes and codebooks are weights: ring fills, L2 prefetch of the
// worker's whole share, codebook fragments and codebook norms all happen BEFORE griddepcontrol.wait.  With
// LUTVQ_FLAG_X_READY (include/lutvq.h) the LUT build, the gather and the atomics run before the wait too.
//
// Algorithmic HBM bytes per launch: O*(I/8)*2 codes + 8192 codebooks + 2I x + 2O scales + 2O y.
#include "common.cuh"
#include "lut_layout.h"

#ifndef LUTVQ_DBG
#define LUTVQ_DBG 0          // 1: compile the experiment knobs (flags 0x100 no lookups, 0x200 no ring refills, 0x400 no atomics)
#endif
#if LUTVQ_DBG
#define LUTVQ_DBG_AND(x) && (x)
#else
#define LUTVQ_DBG_AND(x)
#endif

namespace lutvq {

constexpr size_t kWsHeaderBytes = 256;        // reserved
constexpr unsigned kFlagL2Prefetch = 0x10000u; // internal: push the worker's share HBM -> L2 in the prologue (dependent launches)

constexpr int kMaxPeers = 8;                  // one NVSwitch domain
constexpr int kRangeTable = 320;              // workers whose item ranges ride in the kernel parameters (B200: 148 or 2 x 148)
constexpr int kMaxProblems = 3;               // linears that share x and go out in ONE launch (q/k/v, gate/up)
constexpr int kMaxSlices = 63;                // the arrival counter of the split-k word has 6 bits

// two fp16 lanes of silu(g) * u, rounded once: the arithmetic of decode_glue.cu silu_mul_kernel, bit for bit
__device__ __forceinline__ uint32_t silu_mul_halves(uint32_t g, uint32_t u) {
    const __half2 gh = *reinterpret_cast<const __half2*>(&g), uh = *reinterpret_cast<const __half2*>(&u);
    const float a0 = __low2float(gh), a1 = __high2float(gh);
    const __half r0 = __float2half_rn(a0 / (1.f + __expf(-a0)) * __low2float(uh));
    const __half r1 = __float2half_rn(a1 / (1.f + __expf(-a1)) * __high2float(uh));
    const __half2 r = __halves2half2(r0, r1);
    return *reinterpret_cast<const uint32_t*>(&r);
}

// One linear layer of a (possibly grouped) launch.  Every worker (CTA) belongs to exactly one problem.
struct GemvProblem {
    const uint8_t* codes;     // packed image (lutvq_code2x8_prepack)
    const void* codebook;     // [2][256][8]
    void* y;                  // [O]
    const void* scales;       // [O] or null
    const void* bias;         // [O] or null
    unsigned long long* acc;  // [O_pad / 2] split-k accumulator words, zero between launches
    int O, n_sets;
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

// the two values of a packed fp16x2 / bf16x2 word as floats
template <int DT> __device__ __forceinline__ float2 unpack2(uint32_t w);
template <> __device__ __forceinline__ float2 unpack2<LUTVQ_F16>(uint32_t w)  { return __half22float2(*reinterpret_cast<const __half2*>(&w)); }
template <> __device__ __forceinline__ float2 unpack2<LUTVQ_BF16>(uint32_t w) { return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xFFFF0000u)); }
// sum of squares of the 8 values of one 16-byte group / codebook row, fixed order (every CTA must get the same bits)
template <int DT> __device__ __forceinline__ float sumsq8(uint4 v) {
    float s = 0.f;
    float2 f;
    f = unpack2<DT>(v.x); s = fmaf(f.x, f.x, s); s = fmaf(f.y, f.y, s);
    f = unpack2<DT>(v.y); s = fmaf(f.x, f.x, s); s = fmaf(f.y, f.y, s);
    f = unpack2<DT>(v.z); s = fmaf(f.x, f.x, s); s = fmaf(f.y, f.y, s);
    f = unpack2<DT>(v.w); s = fmaf(f.x, f.x, s); s = fmaf(f.y, f.y, s);
    return s;
}

__device__ __forceinline__ void sts_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" :: "r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ uint4 ldcg_u128(const void* p) {            // L2-coherent: x may be the previous kernel's output
    uint4 v;
    asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ unsigned long long atom_add_u64(unsigned long long* p, unsigned long long v) {
    unsigned long long old;
    asm volatile("atom.relaxed.gpu.global.add.u64 %0, [%1], %2;" : "=l"(old) : "l"(p), "l"(v) : "memory");
    return old;
}
// asynchronous HBM -> L2 prefetch of a contiguous range (no registers, no shared memory)
__device__ __forceinline__ void l2_prefetch_bulk(const void* p, uint32_t bytes, uint64_t pol) {
    asm volatile("cp.async.bulk.prefetch.L2.global.L2::cache_hint [%0], %1, %2;" :: "l"(p), "r"(bytes), "l"(pol) : "memory");
}
__device__ __forceinline__ void l2_prefetch_line(const void* p) {
    asm volatile("prefetch.global.L2 [%0];" :: "l"(p));
}
// L2 residency: the code stream is read exactly once per GEMV (evict_first), while the few KB of codebook / scales / bias
// of every layer are re-read every token (evict_last): without the hints the 11 MB stream of each launch pushes them out
// and every launch pays HBM latency for its codebook in the prologue and for its scales in the tail.
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
// ---- mbarrier + 1-D bulk async copy (TMA engine, no tensor map): global -> this CTA's shared memory ----
__device__ __forceinline__ void mbar_init(uint32_t mbar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(mbar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t mbar, uint64_t pol) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(mbar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 :: "r"(dst), "l"(src), "r"(bytes), "r"(mbar), "l"(pol) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t mbar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}\n" :: "r"(mbar), "r"(parity) : "memory");
}

// ---- fixed-point LUT entries (see the file header) ----
constexpr float    kFxTarget   = 32704.f;                            // |L * S| <= 32704 < 32767: margin for the fp32 roundings of bound and S
constexpr float    kFxMagicB   = 12582912.f + 32768.f;               // 1.5 * 2^23 + bias: float bits' low 16 = round(v) + 32768
constexpr int      kFxBias     = 32768;
constexpr unsigned long long kFieldMask = (1ull << 29) - 1ull;

// store one finished output element locally and into every peer's gather buffer (sharded decode)
template <int DT>
__device__ __forceinline__ void store_y(const GemvParams& p, const GemvProblem& PB, int row, typename Dtype<DT>::T v) {
    using T = typename Dtype<DT>::T;
    static_cast<T*>(PB.y)[row] = v;
    for (int r = 0; r < p.n_push; ++r) static_cast<T*>(p.push_y[r])[row] = v;       // NVLink stores (peer memory)
}

// Fallback when the grid is larger than the tabulated ranges (never on a 148-SM B200): the linear plan, one problem.
__device__ __forceinline__ void worker_range(const GemvParams& p, int w, int W, int& begin, int& end) {
    const long long T = (long long)p.n_k * p.prob[0].n_sets;
    begin = (int)((T * w) / W);
    end   = (int)((T * (w + 1)) / W);
}

// One lookup = one full-rate 32-bit shared load of the word {codebook 0 | codebook 1 << 16} of (group, idx) + one IDP.2A that adds
// the wanted 16-bit half to the accumulator (multipliers {1, 0} / {0, 1}).  (16-bit shared loads, the obvious alternative, ran
// at about half the wavefront rate of 32-bit ones: the gather saturated at 0.55 wavefronts per clock and SM.)
__device__ __forceinline__ uint32_t dp2a_lo(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t lds_u32(const uint8_t* p) { return *reinterpret_cast<const uint32_t*>(p); }

// 32 lookups of one tile: the lane's two 16-byte code images (halves p = 0, 1) against the slice LUT at `lut`
__device__ __forceinline__ uint32_t gather_tile(const uint8_t* lut, const uint4& w0, const uint4& w1, const uint32_t (&lm)[8]) {
    const uint32_t w[2][4] = {{w0.x, w0.y, w0.z, w0.w}, {w1.x, w1.y, w1.z, w1.w}};
    uint32_t acc0 = 0, acc1 = 0;
    #pragma unroll
    for (int p = 0; p < 2; ++p)
        #pragma unroll
        for (int j = 0; j < 4; ++j)
            #pragma unroll
            for (int b = 0; b < 4; b += 2) {
                const int u = 2 * j + (b >> 1);                      // image bytes k = 4j + b, 4j + b + 1: (u, h = 0) and (u, h = 1)
                const uint32_t ad0 = __byte_perm(w[p][j], lm[u], 0x6504u | (uint32_t)(b << 4));
                const uint32_t ad1 = __byte_perm(w[p][j], lm[u], 0x6504u | (uint32_t)((b + 1) << 4));
                acc0 = dp2a_lo(lds_u32(lut + ad0 + (128 * p)), 0x0001u, acc0);      // codebook 0: low half
                acc1 = dp2a_lo(lds_u32(lut + ad1 + (128 * p)), 0x0100u, acc1);      // codebook 1: high half
            }
    return acc0 + acc1;
}

// 16 lookups of one half tile (half p of the lane's image) against the slice LUT at `lut`
template <int P>
__device__ __forceinline__ uint32_t gather_half(const uint8_t* lut, const uint4& wv, const uint32_t (&lm)[8]) {
    const uint32_t w[4] = {wv.x, wv.y, wv.z, wv.w};
    uint32_t acc0 = 0, acc1 = 0;
    #pragma unroll
    for (int j = 0; j < 4; ++j)
        #pragma unroll
        for (int b = 0; b < 4; b += 2) {
            const int u = 2 * j + (b >> 1);
            const uint32_t ad0 = __byte_perm(w[j], lm[u], 0x6504u | (uint32_t)(b << 4));
            const uint32_t ad1 = __byte_perm(w[j], lm[u], 0x6504u | (uint32_t)((b + 1) << 4));
            acc0 = dp2a_lo(lds_u32(lut + ad0 + (128 * P)), 0x0001u, acc0);
            acc1 = dp2a_lo(lds_u32(lut + ad1 + (128 * P)), 0x0100u, acc1);
        }
    return acc0 + acc1;
}

// Code streaming modes.
//   kStreamRegs  128-bit no-allocate loads into a rolling register ring of kRingPieces half tiles (512 B per warp and piece):
//                no shared memory besides the LUT, <= 80 registers => THREE CTAs (of consecutive launches) per SM.
//   kStreamBulk  one 4 KiB bulk async copy (cp.async.bulk + mbarrier) per set into a per-warp landing slot, moved to registers
//                on arrival and refilled at once: 96 KiB of shared memory and 128 registers => two CTAs per SM.
constexpr int kStreamRegs = 0, kStreamBulk = 1;
constexpr int kRingPieces = 4;
// dynamic shared memory: [LUT 64 KiB]( + [NW landing slots of one set = 4 KiB each][NW mbarriers x 8 B])
template <int NW, int MODE> constexpr size_t gemv_smem_bytes() { return (size_t)kSliceBytes + (MODE == kStreamBulk ? (size_t)NW * kSetBytes + (size_t)NW * 8 : 0); }

template <int DT, int NW, int MODE>
__global__ void __launch_bounds__(NW * 32, MODE == kStreamRegs ? 3 : 2)
lut_gemv_2x8_kernel(const __grid_constant__ GemvParams p) {
    static_assert(16 % NW == 0, "NW divides the 16 idx blocks");
    static_assert(NW * 32 >= kSliceGroups, "one thread per group of the slice loads x");
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t lut_base = smem_u32(smem_raw);
    const GemvProblem& PB = p.prob[p.n_prob > 1 ? (int)p.tab_prob[blockIdx.x] : 0];       // indexed constant-bank loads

    TL_STAMP(0);
#if LUTVQ_DBG
    if (p.flags & 0x1000u) pdl_launch_dependents();
    if (p.tl != nullptr && threadIdx.x == 0) {
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        p.tl[(size_t)blockIdx.x * 16 + 15] = smid + 1;
    }
#endif
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int begin, end;
    if (p.has_table) { begin = p.tab_begin[blockIdx.x]; end = p.tab_end[blockIdx.x]; }          // two constant-bank loads
    else worker_range(p, blockIdx.x, gridDim.x, begin, end);

    // kStreamBulk: this warp's landing slot (one 4 KiB set, filled by ONE bulk copy: small copies are bound by the per-SM issue
    // rate of the copy engine, measured 2.4-3.3 TB/s chip-wide with 1 KiB copies) + its mbarrier.  A set is moved from the slot
    // into registers as soon as it has landed and the slot is refilled at once.
    const uint32_t slot_base = lut_base + kSliceBytes + (uint32_t)warp * kSetBytes;
    const uint32_t mbar = lut_base + kSliceBytes + (uint32_t)NW * kSetBytes + (uint32_t)warp * 8;

    // The workers of different k-slices walk the same rows; each starts at a slice-dependent rotation of its own
    // range so that the n_k arrivals of one row set are spread over the whole gather instead of colliding at one instant.
    auto rot_of = [&](int ks_, int n) { return (int)(((unsigned)ks_ * (unsigned)n) / (unsigned)p.n_k); };
    // first segment of this worker's range: slice ks, sets [lo, hi)
    int pos = begin;
    int ks = p.has_table ? (int)p.tab_ks[blockIdx.x] : (begin < end) ? (int)((unsigned)begin / (unsigned)PB.n_sets) : 0;
    int lo = begin - ks * PB.n_sets;
    int hi = min(PB.n_sets, lo + (end - begin));
    int seg_rot = rot_of(ks, hi - lo);
    // set (index inside the problem) that local index li (rotated) of the current segment stands for
    auto set_of = [&](int li) -> int {
        int v = li + seg_rot;
        const int n = hi - lo;
        if (v >= n) v -= n;
        return lo + v;
    };
    auto set_ptr = [&](int li) -> const uint8_t* { return PB.codes + ((size_t)ks * PB.n_sets + (size_t)set_of(li)) * kSetBytes; };
    // one set into this warp's landing slot (lane 0; the slot is free)
    const uint64_t pol_stream = l2_policy_evict_first(), pol_keep = l2_policy_evict_last();
    auto fill_slot = [&](const uint8_t* sp) { bulk_load(slot_base, sp, kSetBytes, mbar, pol_stream); };

    // ---------------- prologue that does not depend on the previous kernel in the stream ----------------
    // Codes and codebooks are weights: the first code pieces (registers / landing slot), the L2 prefetch of the worker's
    // share and the codebook MMA fragments are all requested BEFORE the programmatic-dependency wait.
    // kStreamRegs: the rolling register ring: piece i of this warp's stream = 16 bytes per lane at set_ptr(set) + (i % 8) * 512
    uint4 ring[MODE == kStreamRegs ? kRingPieces : 1];
    if (MODE == kStreamRegs) {
        const uint8_t* sp = set_ptr(warp < hi - lo ? warp : 0) + lane * 16;
        #pragma unroll
        for (int i = 0; i < kRingPieces; ++i) ring[i] = (begin < end) ? ldg_stream_u128_pol(sp + i * 512, pol_stream) : make_uint4(0u, 0u, 0u, 0u);
    }
    if (lane == 0) {
        if (MODE == kStreamBulk) {
            mbar_init(mbar, 1);
            fence_mbar_init();
            if (warp < hi - lo) fill_slot(set_ptr(warp));
        }
        // this worker's share is ONE contiguous range of the packed image: warp w prefetches items w, w + NW, ...
        const uint8_t* base = PB.codes + (size_t)begin * kSetBytes;
        if (p.flags & kFlagL2Prefetch LUTVQ_DBG_AND(!(p.flags & 0x800u)))
            for (int c = warp; c < end - begin; c += NW) l2_prefetch_bulk(base + (size_t)c * kSetBytes, kSetBytes, pol_stream);
    }
    TL_STAMP(9);
    if (p.n_k > 1 && begin < end) {
        // scale / bias of this worker's first-segment rows: 64 B per set each, pulled into L2 for the finishing rows
        const int nl = (hi - lo + 1) / 2;                             // 128-byte lines
        if ((int)threadIdx.x < nl) {
            const size_t off = (size_t)lo * 64 + (size_t)threadIdx.x * 128;
            if (off < (size_t)PB.O * 2) {
                if (PB.scales) l2_prefetch_line(static_cast<const uint8_t*>(PB.scales) + off);
                if (PB.bias)   l2_prefetch_line(static_cast<const uint8_t*>(PB.bias) + off);
            }
        }
    }
    const int r = lane >> 2, q = lane & 3;
    constexpr int MB_PER_WARP = 16 / NW;                                   // 16 idx blocks of 16 per codebook
    constexpr int NT = NW * 32;
    // Every CTA of a launch reads the same 8 KiB codebook and the same x at the same moment: those few hot lines are served by
    // a few L2 slices, and the launch's whole "skeleton" was bound by that traffic (measured: 3 us per launch with lookups,
    // streaming and atomics switched off).  So each line is requested ONCE per CTA: the codebook only as MMA fragments (its row
    // norms come out of the same registers), x only as the CTA's own 1 KiB slice through shared memory.
    uint32_t afrag[MB_PER_WARP][2][2];
    {
        const uint32_t* cb32 = static_cast<const uint32_t*>(PB.codebook);  // [2][256][4] words (2 values each)
        // A fragments for this warp's idx blocks: rows 16m + r and 16m + r + 8, k = 2q, 2q+1
        #pragma unroll
        for (int mi = 0; mi < MB_PER_WARP; ++mi) {
            const int m = warp * MB_PER_WARP + mi;
            #pragma unroll
            for (int n = 0; n < 2; ++n) {
                afrag[mi][n][0] = ldg_u32_pol(cb32 + (n * 256 + 16 * m + r) * 4 + q, pol_keep);
                afrag[mi][n][1] = ldg_u32_pol(cb32 + (n * 256 + 16 * m + r + 8) * 4 + q, pol_keep);
            }
        }
    }
    __shared__ uint32_t red_c[NW], red_x[NW];
    __shared__ __align__(16) uint4 xs[kSliceGroups];                      // the slice's 64 activation groups (silu(gate) * x applied)
    {
        // max over this warp's codebook rows of |row|_2^2, from the fragments: a row's 8 values sit in the 4 lanes of a quad.
        // (bits of a non-negative float order like the float; NaN highest)
        uint32_t mc = 0;
        #pragma unroll
        for (int mi = 0; mi < MB_PER_WARP; ++mi)
            #pragma unroll
            for (int n = 0; n < 2; ++n)
                #pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const float2 f = unpack2<DT>(afrag[mi][n][h]);
                    float s2 = fmaf(f.y, f.y, f.x * f.x);
                    s2 += __shfl_xor_sync(0xffffffffu, s2, 1);
                    s2 += __shfl_xor_sync(0xffffffffu, s2, 2);
                    mc = max(mc, __float_as_uint(s2));
                }
        mc = __reduce_max_sync(0xffffffffu, mc);
        if (lane == 0) red_c[warp] = mc;
    }
    TL_STAMP(1);
    // The next launch may become resident now that OUR weight requests are in flight (it still waits for us below).
#if LUTVQ_DBG
    if (!(p.flags & 0x1000u))                                      // DEBUG 0x1000: the trigger was issued at kernel entry instead
#endif
    pdl_launch_dependents();
    // X_READY: x is not the previous kernel's output and the workspace is exclusively ours (include/lutvq.h), so
    // nothing but the final y stores has to be ordered behind the previous kernel.
    const bool xready = (p.flags & LUTVQ_FLAG_X_READY) != 0;
    bool waited = !xready;
    if (!xready) pdl_wait();    // x and the workspace belong to the previous kernel until here
    TL_STAMP(2);

    const bool gated = (DT == LUTVQ_F16) && p.x_gate != nullptr;
    auto load_group = [&](int g) -> uint4 {                        // 16 bytes = one group of 8 activations, SwiGLU applied on the way in
        uint4 v = ldcg_u128(static_cast<const uint4*>(p.x) + g);
        if (gated) {
            const uint4 gt = ldcg_u128(static_cast<const uint4*>(p.x_gate) + g);
            v.x = silu_mul_halves(gt.x, v.x); v.y = silu_mul_halves(gt.y, v.y);
            v.z = silu_mul_halves(gt.z, v.z); v.w = silu_mul_halves(gt.w, v.w);
        }
        return v;
    };
    // the CTA's slice of x for a later segment: one coalesced 16-byte load per group (8 lines for the whole CTA)
    auto load_x_slice = [&](int ks_) {
        if (threadIdx.x < kSliceGroups) {
            const int g = ks_ * kSliceGroups + (int)threadIdx.x;
            xs[threadIdx.x] = (g < p.G) ? load_group(g) : make_uint4(0u, 0u, 0u, 0u);
        }
    };
    // ---------------- the activation-dependent part of the scale: max_g |x_g|_2^2 over ALL groups ----------------
    // ONE coalesced pass per CTA (every line of x requested once), XU loads in flight per thread; the groups of the CTA's
    // own slice are dropped into shared memory on the way (the B fragments come from there, not from L2 again).
    {
        uint32_t m = 0;
        constexpr int XU = 4;
        if (threadIdx.x < kSliceGroups && ks * kSliceGroups + (int)threadIdx.x >= p.G) xs[threadIdx.x] = make_uint4(0u, 0u, 0u, 0u);
        for (int g0 = threadIdx.x; g0 < p.G; g0 += XU * NT) {
            uint4 v[XU];
            #pragma unroll
            for (int u = 0; u < XU; ++u) {
                const int g = g0 + u * NT;
                v[u] = (g < p.G) ? load_group(g) : make_uint4(0u, 0u, 0u, 0u);
            }
            #pragma unroll
            for (int u = 0; u < XU; ++u) {
                const int g = g0 + u * NT;
                const unsigned loc = (unsigned)(g - ks * kSliceGroups);
                if (g < p.G && loc < (unsigned)kSliceGroups) xs[loc] = v[u];
                m = max(m, __float_as_uint(sumsq8<DT>(v[u])));
            }
        }
        m = __reduce_max_sync(0xffffffffu, m);
        if (lane == 0) red_x[warp] = m;
    }

    uint32_t lm[8];
    #pragma unroll
    for (int u = 0; u < 8; ++u) lm[u] = stream_lm(lane, u);
    const bool b3 = (lane >> 3) & 1, b4 = (lane >> 4) & 1;
    const bool odd = r & 1;
    const bool round16 = (p.flags & LUTVQ_FLAG_LUT_FP16) != 0;

    // ---- split-k completion: see the file header ----
    using D = Dtype<DT>;
    using T = typename D::T;
    const int bias_total = p.n_k * kSliceTables * kFxBias;               // every table of every slice contributes the bias once
    unsigned long long pend_old = 0, pend_add = 0;
    int pend_row = 0;
    bool pend_valid = false;
    const unsigned last_count = (unsigned)p.n_k - 1u;
    float fx_scale = 0.f, fx_inv = 0.f;
    bool bad = false;
    auto finish_one = [&](unsigned field, int row) {
        if (row < PB.O) {
            float v = (float)((int)field - bias_total) * fx_inv;
            if (bad) v = __uint_as_float(0x7FC00000u);
            T rr = D::from_f(v);                                                    // Code2x8MatVec writes round(acc)   (cuda_kernel.cu:221-231)
            if (PB.scales) { const uint16_t b = ldg_u16_pol(static_cast<const T*>(PB.scales) + row, pol_keep); rr = D::mul(rr, *reinterpret_cast<const T*>(&b)); }   // flat_output *= scales (cuda_kernel.cpp:101)
            if (PB.bias)   { const uint16_t b = ldg_u16_pol(static_cast<const T*>(PB.bias) + row, pol_keep);   rr = D::add(rr, *reinterpret_cast<const T*>(&b)); }   // flat_output += bias   (cuda_kernel.cpp:103)
            store_y<DT>(p, PB, row, rr);
        }
    };
    auto finish_pair = [&](unsigned long long old, unsigned long long add, int row) {   // row = even row of the pair
        if ((unsigned)(old >> 58) == last_count) {                 // this lane delivered the last partial of the pair
            if (!waited) { pdl_wait(); waited = true; }            // y stores are ordered behind the previous kernel
            const unsigned long long total = old + add;
            if (p.n_k > 1) asm volatile("st.global.cg.b64 [%0], %1;" :: "l"(PB.acc + (row >> 1)), "l"(0ull) : "memory");   // hand the word back
            finish_one((unsigned)(total & kFieldMask), row);
            finish_one((unsigned)((total >> 29) & kFieldMask), row + 1);
        }
    };

    bool first_seg = true;
    uint32_t phase = 0;                                            // kStreamBulk: parity of this warp's slot barrier
    while (pos < end) {
        const int n_seg = hi - lo;
        // ---------------- K1: build this worker's 64 KiB LUT for slice ks ----------------
        if (!first_seg) {
            __syncthreads();                                       // everyone is done reading the previous slice's LUT and xs
            if (MODE == kStreamBulk) { if (lane == 0 && warp < n_seg) fill_slot(set_ptr(warp)); }
            else if (warp < n_seg) {
                const uint8_t* sp = set_ptr(warp) + lane * 16;
                #pragma unroll
                for (int i = 0; i < kRingPieces; ++i) ring[i] = ldg_stream_u128_pol(sp + i * 512, pol_stream);
            }
            load_x_slice(ks);
        }
        __syncthreads();                                           // xs and red_c are complete
        if (first_seg) TL_STAMP(10);
        if (first_seg) {
            // ONE fixed-point scale per GEMV: S = 32704 / (max_k |C_k|_2 * max_g |x_g|_2) >= |L| by Cauchy-Schwarz, computed
            // identically by every CTA from (x, codebook) alone
            uint32_t mc = 0, mx = 0;
            #pragma unroll
            for (int i = 0; i < NW; ++i) { mc = max(mc, red_c[i]); mx = max(mx, red_x[i]); }
            const float bound = sqrtf(__uint_as_float(mc) * __uint_as_float(mx));
            bad = !(bound < __uint_as_float(0x7F800000u));          // inf or NaN activations / codebooks: the outputs are NaN
            const bool tiny = !(bound > 1e-30f);
            fx_scale = (bad || tiny) ? 0.f : __fdiv_rn(kFxTarget, bound);
            fx_inv   = (bad || tiny) ? 0.f : __fdiv_rn(bound, kFxTarget);
        }
        // B fragments of the slice: column (lane>>2) of N-tile j is local group build_group_of_col(j, r); k = 2q, 2q+1
        uint32_t bfrag[8];
        #pragma unroll
        for (int j = 0; j < 8; ++j) bfrag[j] = reinterpret_cast<const uint32_t*>(xs)[build_group_of_col(j, r) * 4 + q];
        if (first_seg) TL_STAMP(11);
        #pragma unroll
        for (int mi = 0; mi < MB_PER_WARP; ++mi) {
            const int m = warp * MB_PER_WARP + mi;
            #pragma unroll
            for (int hb = 0; hb < 2; ++hb) {                       // N-tiles 4hb .. 4hb+3 = tile pairs 2hb, 2hb+1
                float c[2][4][4];
                #pragma unroll
                for (int n = 0; n < 2; ++n)
                    #pragma unroll
                    for (int jj = 0; jj < 4; ++jj) {
                        c[n][jj][0] = c[n][jj][1] = c[n][jj][2] = c[n][jj][3] = 0.f;
                        MmaK8<DT>::run(c[n][jj], afrag[mi][n][0], afrag[mi][n][1], bfrag[4 * hb + jj]);
                    }
                // words[ii][hh][e*2+f]: tile pair ii of this half, idx row r + 8hh, groups +0..3 = (tile 2ii+e, column 2q+f)
                #pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                    uint32_t wd[2][4];
                    #pragma unroll
                    for (int ii = 0; ii < 2; ++ii)
                        #pragma unroll
                        for (int e = 0; e < 2; ++e)
                            #pragma unroll
                            for (int f = 0; f < 2; ++f) {
                                float v0 = c[0][2 * ii + e][2 * hh + f], v1 = c[1][2 * ii + e][2 * hh + f];
                                if (round16) { v0 = __half2float(__float2half_rn(v0)); v1 = __half2float(__float2half_rn(v1)); }
                                const uint32_t u0 = __float_as_uint(fmaf(v0, fx_scale, kFxMagicB));   // low 16 bits = round(v * S) + 32768
                                const uint32_t u1 = __float_as_uint(fmaf(v1, fx_scale, kFxMagicB));
                                wd[ii][2 * e + f] = __byte_perm(u0, u1, 0x5410);                       // {codebook 0 | codebook 1 << 16}
                            }
                    const int idx = 16 * m + r + 8 * hh;
                    // first store: odd lanes take pair 1, even lanes pair 0 (8 distinct 16-byte columns per quarter-warp phase)
                    const uint32_t first  = lut_base + build_store_offset(idx, 2 * hb + (odd ? 1 : 0), q);
                    const uint32_t second = lut_base + build_store_offset(idx, 2 * hb + (odd ? 0 : 1), q);
                    sts_v4(first,  odd ? wd[1][0] : wd[0][0], odd ? wd[1][1] : wd[0][1], odd ? wd[1][2] : wd[0][2], odd ? wd[1][3] : wd[0][3]);
                    sts_v4(second, odd ? wd[0][0] : wd[1][0], odd ? wd[0][1] : wd[1][1], odd ? wd[0][2] : wd[1][2], odd ? wd[0][3] : wd[1][3]);
                }
            }
        }
        if (first_seg) TL_STAMP(12);
        __syncthreads();
        if (first_seg) TL_STAMP(3);
        first_seg = false;

        // ---------------- K2: gather-accumulate over this warp's sets of the segment ----------------
        for (int li = warp; li < n_seg; li += NW) {
            const bool has_next = li + NW < n_seg;
            uint32_t acc[4];
            if (MODE == kStreamBulk) {
                uint4 w[4][2];
#if LUTVQ_DBG
                if (!(p.flags & 0x200u))                           // DEBUG 0x200: no code streaming at all (stale slot contents)
#endif
                mbar_wait(mbar, phase);
                phase ^= 1u;
                #pragma unroll
                for (int t = 0; t < 4; ++t) {
                    w[t][0] = lds_v4(slot_base + t * kTileBytes + lane * 16);
                    w[t][1] = lds_v4(slot_base + t * kTileBytes + 512 + lane * 16);
                }
                #pragma unroll
                for (int t = 0; t < 4; ++t) {
#if LUTVQ_DBG
                    if (p.flags & 0x100u) acc[t] = w[t][0].x ^ w[t][0].y ^ w[t][0].z ^ w[t][0].w ^ w[t][1].x ^ w[t][1].y ^ w[t][1].z ^ w[t][1].w;   // DEBUG: no lookups
                    else
#endif
                    acc[t] = gather_tile(smem_raw, w[t][0], w[t][1], lm);
                    if (t == 0) {
                        // Refill the slot with this warp's next set.  Every lane's 8 slot reads were issued before tile 0's
                        // lookups consumed the first of them (in-order return), ~100 cycles ago; the copy lands >= an L2 round
                        // trip from now.
                        __syncwarp();
                        if (lane == 0 && has_next LUTVQ_DBG_AND(!(p.flags & 0x200u))) fill_slot(set_ptr(li + NW));
                    }
                }
            } else {
                const uint8_t* cur = set_ptr(li) + lane * 16;
                const uint8_t* nxt = has_next ? set_ptr(li + NW) + lane * 16 : cur;
                #pragma unroll
                for (int pc = 0; pc < 8; ++pc) {
                    const uint4 wv = ring[pc % kRingPieces];
                    // the piece kRingPieces ahead in this warp's stream takes the freed ring slot (the last set re-reads
                    // its own pieces: harmless, never consumed)
                    const int ahead = pc + kRingPieces;
#if LUTVQ_DBG
                    if (!(p.flags & 0x200u))
#endif
                    ring[pc % kRingPieces] = ldg_stream_u128_pol(ahead < 8 ? cur + ahead * 512 : nxt + (ahead - 8) * 512, pol_stream);
                    uint32_t a;
#if LUTVQ_DBG
                    if (p.flags & 0x100u) a = wv.x ^ wv.y ^ wv.z ^ wv.w;
                    else
#endif
                    a = (pc & 1) ? gather_half<1>(smem_raw, wv, lm) : gather_half<0>(smem_raw, wv, lm);
                    if (pc & 1) acc[pc >> 1] += a; else acc[pc >> 1] = a;
                }
            }
            // transposed reduce over the 4 chunk lanes (lane bits 3,4): lane ends with tile (2*b4+b3), row lane&7
            uint32_t k0 = b4 ? acc[2] : acc[0], k1 = b4 ? acc[3] : acc[1];
            uint32_t g0 = b4 ? acc[0] : acc[2], g1 = b4 ? acc[1] : acc[3];
            k0 += __shfl_xor_sync(0xffffffffu, g0, 16);
            k1 += __shfl_xor_sync(0xffffffffu, g1, 16);
            uint32_t keep = b3 ? k1 : k0, give = b3 ? k0 : k1;
            keep += __shfl_xor_sync(0xffffffffu, give, 8);            // row (set row0 + lane): sum of its 128 entries, < 2^23
            const uint32_t nb = __shfl_xor_sync(0xffffffffu, keep, 1);
            const int row = set_of(li) * kSetRows + lane;
            // finish the atomic issued one set ago (its L2 round trip has long completed), then issue this set's
            if (pend_valid) finish_pair(pend_old, pend_add, pend_row);
            pend_valid = false;
            if (!(lane & 1)) {
                const unsigned long long add = (unsigned long long)keep + ((unsigned long long)nb << 29) + (1ull << 58);
                pend_add = add; pend_row = row; pend_valid = true;
                if (p.n_k == 1) pend_old = 0;
#if LUTVQ_DBG
                else if (p.flags & 0x400u) pend_old = 0;              // DEBUG: no atomics
#endif
                else pend_old = atom_add_u64(PB.acc + (row >> 1), add);
            }
        }
        pos += n_seg;
        if (pos < end) {                                           // next segment (linear plan only)
            ks = pos / PB.n_sets;
            lo = pos - ks * PB.n_sets;
            hi = min(PB.n_sets, lo + (end - pos));
            seg_rot = rot_of(ks, hi - lo);
        }
    }
    TL_STAMP(4);
    if (pend_valid) finish_pair(pend_old, pend_add, pend_row);
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
            if (threadIdx.x == 0) *p.cta_arrivals = 0u;                               // self-cleaning, like the counters
            __threadfence_system();
            const unsigned epoch = *p.push_sent + 1u;
            if ((int)threadIdx.x < p.n_push)
                asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p.push_flag[threadIdx.x]), "r"(epoch) : "memory");
            __syncthreads();
            if (threadIdx.x == 0) *p.push_sent = epoch;
        }
    }
}

// packed image from reference-layout codes: [k-slice][set][tile][half][lane] x 16 B in lane-walk order (lut_layout.h)
__global__ void prepack_2x8_kernel(const uint8_t* __restrict__ codes, uint4* __restrict__ packed, int O, int G,
                                   int n_k, int n_sets_total, uint4* __restrict__ zero, size_t zero16) {
    // staging use (gemv_stage_codes): the split-k words in front of the staged image may hold a previous call's staged
    // codes (another shape, same cached workspace): hand them over all-zero, as the GEMV kernel expects them
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < zero16; i += (size_t)gridDim.x * blockDim.x)
        zero[i] = make_uint4(0u, 0u, 0u, 0u);
    const size_t total = (size_t)n_k * n_sets_total * (kSetBytes / 16);
    const size_t row_bytes = (size_t)G * 2;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int lane = (int)(i & 31), ph = (int)((i >> 5) & 1), t = (int)((i >> 6) & 3);
        const size_t ss = i >> 8;
        const int set = (int)(ss % n_sets_total), ks = (int)(ss / n_sets_total);
        const int row = set * kSetRows + t * kTileRows + stream_row_in_tile(lane);
        const size_t off = (size_t)ks * kSliceTables + 16 * stream_chunk(lane, ph);
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (row < O && off < row_bytes) v = *reinterpret_cast<const uint4*>(codes + (size_t)row * row_bytes + off);
        const uint32_t in[4] = {v.x, v.y, v.z, v.w};
        uint32_t out[4];
        pack_lane_image(in, lane, out);                            // image byte k = 2u + h holds chunk byte 2(u ^ r) + h
        packed[i] = make_uint4(out[0], out[1], out[2], out[3]);
    }
}

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------
struct GemvTuning { int nw = 0, nslot = 0; };   // 0 = choose per launch mode (launch_tuned)
static GemvTuning g_tuning;
static int g_dep_workers_per_sm = 2;   // workers (CTAs) per SM of a DEPENDENT launch, see gemv_2x8_launch_group
void gemv_set_dependent_workers(int m) { g_dep_workers_per_sm = m < 1 ? 1 : (m > 2 ? 2 : m); }
static unsigned long long* g_timeline = nullptr;   // debug: [64 launches][256 CTAs][16] stamps
static unsigned g_timeline_seq = 0;
void gemv_set_timeline(void* buf) { g_timeline = static_cast<unsigned long long*>(buf); g_timeline_seq = 0; }

struct GemvPlan { int n_k, n_sets, W, aligned, q, r, max_range; };

// One worker (CTA) per SM.  Cost model in units of one 32-row set of gather work (128 lookups per lane); a LUT build
// (256 MMAs + 64 KiB of stores per CTA) costs about 6 of them.
static GemvPlan plan_gemv(int O, int G, int sm_count) {
    GemvPlan pl;
    pl.n_k = ceil_div(G * 2, kSliceTables);
    pl.n_sets = ceil_div(O, kSetRows);
    const long long T = (long long)pl.n_k * pl.n_sets;
    pl.W = (int)(T < sm_count ? T : sm_count);
    if (pl.W > 1024) pl.W = 1024;
    const int kBuildCost = 6;
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

static size_t align256(size_t v) { return (v + 255) / 256 * 256; }

// header + one split-k word per PAIR of (padded) rows.  (Kept at 8 bytes per row: callers size rings from this.)
size_t gemv_packed_workspace_bytes(int O, int G) {
    (void)G;
    const size_t O_pad = (size_t)ceil_div(O, kSetRows) * kSetRows;
    return kWsHeaderBytes + O_pad * sizeof(unsigned long long);
}

size_t gemv_packed_bytes(int O, int G) {
    const int n_k = ceil_div(G * 2, kSliceTables);
    return (size_t)n_k * ceil_div(O, kSetRows) * kSetBytes;
}

// the reference-layout entry points stage a packed image behind the split-k words
size_t gemv_workspace_bytes(int O, int G) { return align256(gemv_packed_workspace_bytes(O, G)) + gemv_packed_bytes(O, G); }

int gemv_prepack_launch(const void* codes, void* packed, int O, int I, cudaStream_t stream, void* zero, size_t zero_bytes) {
    const int G = I / 8;
    const int n_k = ceil_div(G * 2, kSliceTables), n_sets_total = ceil_div(O, kSetRows);
    const size_t total = (size_t)n_k * n_sets_total * (kSetBytes / 16);
    int grid = (int)((total + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    prepack_2x8_kernel<<<grid, 256, 0, stream>>>(static_cast<const uint8_t*>(codes), static_cast<uint4*>(packed), O, G, n_k, n_sets_total,
                                                 static_cast<uint4*>(zero), zero_bytes / 16);
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

template <int DT, int NW, int MODE>
static int launch_variant(const GemvParams& p, int grid, cudaStream_t stream, bool pdl) {
    auto kern = lut_gemv_2x8_kernel<DT, NW, MODE>;
    constexpr size_t smem = gemv_smem_bytes<NW, MODE>();
    static thread_local int attr_device = -1;
    int dev = 0;
    LUTVQ_CUDA_OK(cudaGetDevice(&dev));
    if (attr_device != dev) {
        LUTVQ_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
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

template <int DT>
static int launch_tuned(const GemvParams& p, int grid, cudaStream_t stream, bool pdl) {
    GemvTuning t = g_tuning;
    if (t.nw == 0) { t.nw = 8; t.nslot = kStreamRegs; }
    if (t.nw == 8 && t.nslot == kStreamRegs) return launch_variant<DT, 8, kStreamRegs>(p, grid, stream, pdl);   // 3 CTAs / SM: 64 KiB LUT, register ring
    if (t.nw == 8 && t.nslot == kStreamBulk) return launch_variant<DT, 8, kStreamBulk>(p, grid, stream, pdl);   // 2 CTAs / SM: + 32 KiB of landing slots
    return fail(LUTVQ_EINVAL, "no kernel variant for tuning nw=%d mode=%d", t.nw, t.nslot);
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
// its share of the SMs.  x_gate != null (fp16): the activations are silu(x_gate) * x, evaluated in the kernel's x loads.
// Validation is done by the caller (api.cu / shard.cu).
int gemv_2x8_launch_group(int n, const GemvItem* items, const void* x, const void* x_gate, int I, int dtype, unsigned flags,
                          void* ws, size_t ws_bytes, cudaStream_t stream, const GemvPush* push) {
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    LUTVQ_REQUIRE(n >= 1 && n <= kMaxProblems, "a group holds 1..%d linears, got %d", kMaxProblems, n);
    LUTVQ_REQUIRE(aligned16(x) && (x_gate == nullptr || aligned16(x_gate)), "x must be 16-byte aligned (the reference reads it as int4 too, cuda_kernel.cu:147)");
    // An independent (X_READY) ring fills the SMs by overlapping consecutive launches, one worker per SM each.  A dependent
    // launch runs alone: it may be planned over g_dep_workers_per_sm workers per SM.
    const int workers = di.sm_count * ((flags & LUTVQ_FLAG_X_READY) ? 1 : g_dep_workers_per_sm);
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
    LUTVQ_REQUIRE(p.x_gate == nullptr || (dtype == LUTVQ_F16 && !(flags & LUTVQ_FLAG_X_READY)), "the fused SwiGLU input is fp16 only and never X_READY");
    p.has_table = 0;
    int W_total = 0, left = workers;
    uint8_t* accp = static_cast<uint8_t*>(ws) + kWsHeaderBytes;
    for (int i = 0; i < kMaxProblems; ++i) p.prob[i] = GemvProblem{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, 0};
    for (int i = 0; i < n; ++i) {
        // SM share of problem i: proportional to its row sets, at least one, the last problem takes the remainder
        const int n_sets_i = ceil_div(items[i].O, kSetRows);
        int share = (i == n - 1) ? left : (int)((long long)workers * n_sets_i / sets_total);
        if (share < 1) share = 1;
        if (share > left - (n - 1 - i)) share = left - (n - 1 - i);
        left -= share;
        GemvPlan pl = plan_gemv(items[i].O, G, share);
        if (pl.n_k > kMaxSlices)
            return fail(LUTVQ_EUNSUPPORTED, "prob_k=%d needs %d k-slices; the split-k arrival counter holds %d (prob_k <= %d)", I, pl.n_k, kMaxSlices, kMaxSlices * kSliceGroups * 8);
        p.n_k = pl.n_k;
        GemvProblem& pb = p.prob[i];
        pb.codes = static_cast<const uint8_t*>(items[i].codes); pb.codebook = items[i].codebook; pb.y = items[i].y;
        pb.scales = items[i].scales; pb.bias = items[i].bias;
        pb.acc = reinterpret_cast<unsigned long long*>(accp);
        pb.O = items[i].O; pb.n_sets = pl.n_sets;
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
    // A dependent launch sits behind griddepcontrol.wait while its predecessor runs: pushing its share HBM -> L2 meanwhile is
    // free.  An independent (X_READY) launch streams at once, where the extra prefetch requests only cost bandwidth
    // (tools/stream_probe: 5.9 vs 7.3 TB/s).
    if (!(p.flags & LUTVQ_FLAG_X_READY)) p.flags |= kFlagL2Prefetch;
    p.tl = (g_timeline && W_total <= 256) ? g_timeline + (size_t)(g_timeline_seq++ % 64) * 256 * 16 : nullptr;
    if (dtype == LUTVQ_F16)  return launch_tuned<LUTVQ_F16>(p, W_total, stream, pdl);
    if (dtype == LUTVQ_BF16) return launch_tuned<LUTVQ_BF16>(p, W_total, stream, pdl);
    return fail(LUTVQ_EUNSUPPORTED, "dtype %d not supported (fp16=0, bf16=1)", dtype);
}

// One GEMV on a packed code image: y = epilogue(W(codes) . x).
int gemv_2x8_launch(const void* packed, const void* x, void* y, const void* codebook, const void* scales,
                    const void* bias, int O, int I, int dtype, unsigned flags, void* ws, size_t ws_bytes,
                    cudaStream_t stream, const GemvPush* push) {
    const GemvItem it{packed, codebook, y, scales, bias, O};
    return gemv_2x8_launch_group(1, &it, x, nullptr, I, dtype, flags, ws, ws_bytes, stream, push);
}

// Reference-layout codes [O][G][2] -> a packed image in the staging area behind the split-k words of `ws` (one extra pass
// over the codes; QuantizedLinear and every hot caller prepack once at load time instead).  The caller must launch the
// first GEMV that reads the staged image WITHOUT programmatic overlap (LUTVQ_FLAG_NO_PDL), or it could read the staging
// area before the prepack kernel has finished.
int gemv_stage_codes(const void* codes, int O, int I, void* ws, size_t ws_bytes, cudaStream_t stream, const void** staged, size_t* small_bytes) {
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;                  // no device: fail loudly, there is no CPU fallback
    const int G = I / 8;
    const size_t small = align256(gemv_packed_workspace_bytes(O, G));
    const size_t need = small + gemv_packed_bytes(O, G);
    LUTVQ_REQUIRE(ws != nullptr && ws_bytes >= need, "workspace too small: the reference-layout entry points need %zu bytes (lutvq_workspace_bytes), got %zu", need, ws_bytes);
    LUTVQ_REQUIRE(aligned16(ws), "workspace must be 16-byte aligned");
    void* st = static_cast<uint8_t*>(ws) + small;
    if (int rc = gemv_prepack_launch(codes, st, O, I, stream, ws, small)) return rc;     // small is a multiple of 256
    *staged = st; *small_bytes = small;
    return LUTVQ_OK;
}

void gemv_set_tuning(int nw, int nslot, int unused) {
    (void)unused;
    g_tuning.nw = nw > 0 ? nw : 0;
    g_tuning.nslot = nslot > 0 ? nslot : 0;
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
