// gemm_tc.cu -- prefill path: y[b][o] = epilogue( sum_k x[b][k] * W[o][k] ) with W dequantized ON THE FLY into TENSOR
// MEMORY and contracted on the 5th-generation tensor cores (tcgen05.mma with A from TMEM, fp32 accumulators in TMEM).
//
// Replaces code2x8_matmat_dequant (reference benchmark/AQLM/inference_lib/src/aqlm/inference_kernels/
// cuda_kernel.cpp:450-484): Code2x8Dequant writes the whole fp16 W to HBM (2*O*I bytes), cuBLAS F.linear reads it
// back, three elementwise kernels apply scale / bias.  Here W never exists in HBM: the only weight traffic is the
// code stream (O*I/4 bytes), and the epilogue (round to dtype, * scale, + bias; cuda_kernel.cpp:95-111) is fused.
//
// One CTA = one tile of up to 128 output channels x the whole (padded) batch, over a k-range of the split:
//   D[128 x N] (TMEM, fp32)  +=  A[128 x 64] (TMEM, dequantized weights, lane = row)  *  B[N x 64]^T (smem, activations)
//   warps 0-15 producers: each thread owns one output row and one quarter of every 64-wide k block (2 groups = 4 code bytes);
//              per group two conflict-free 128-bit reads of the 8x replicated codebooks in shared memory, one HADD2 x4 (the
//              SAME single RN add as Code2x8Dequant, cuda_kernel.cu:283-289); the 8 resulting 32-bit words are 8 consecutive
//              columns of the thread's own tensor-memory lane: one tcgen05.st per k block.  Its code words come from the
//              chunk ring (warp 18): wait for the chunk's barrier, one 32-bit shared load per k block, release the slot.
//              Afterwards the same warps run the epilogue (tcgen05.ld 32x32b: thread = output channel, registers = batch
//              columns).
//   warp 16    one elected thread issues tcgen05.mma (M=128, N=batch padded to 16, K=16; A operand = tensor-memory address,
//              B operand = SWIZZLE_128B shared-memory descriptor) four times per k block and tcgen05.commit's the stage
//              back; it also owns the TMEM allocation (all 512 columns: accumulator first, the A ring behind it).
//   warp 17    one thread brings in the activations: the tile of a k block (batch x 64, SWIZZLE_128B K-major) is ONE 2-D TMA
//              tensor copy (cp.async.bulk.tensor) completing on the stage's full barrier, issued as soon as the stage is
//              empty, i.e. up to the whole ring ahead of the dequantization; rows past the batch arrive as zeros.
//   warp 18    one thread brings in the CODES: per chunk of 4 (8) k blocks one TMA tensor copy of {64 (128) bytes x the CTA's
//              rows} of the [O][I/4] byte matrix (64- / 128-byte swizzle) into a ring of kCodeRing shared buffers, up to the whole
//              ring ahead and starting before anything else in the kernel.  No register staging, no block-wide barrier.
//   mbarrier ring of `stages` stages of 1 / 2 / 4 k blocks (full: one arrival per producer warp + the copy thread's
//   expect_tx arrival and the copy's bytes; empty: tcgen05.commit).  Programmatic dependent launch: everything before
//   griddepcontrol.wait reads weights only.
//
// Split-k (few row tiles, e.g. 4096 outputs = 32 tiles on 148 SMs): partial accumulators go to the workspace in
// fp32, a per-tile arrival counter elects the last CTA, which adds the partials IN FIXED ORDER (deterministic) and
// runs the epilogue; it resets the counter, so the workspace is all-zero again between launches.
//
// Bound: the tensor pipe at 256 columns (62 % of the active cycles, 80 of 128 MMA rows live); stage hand-overs
// and latency below.  Algorithmic flops 2*batch*I*O; algorithmic HBM bytes as the GEMV + 2*batch*(I+O).
#include "common.cuh"

#include <cuda.h>     // CUtensorMap and its enums only: the encoder is looked up through the runtime, no libcuda link

namespace lutvq {

constexpr int kBM = 128;                 // output channels per CTA (UMMA M)
constexpr int kBK = 64;                  // k per stage = 8 AQLM groups = 16 code bytes per row
constexpr int kProducerThreads = 512;      // 16 warps: thread = (output row, quarter of the k block); four warps per scheduler hide the LDS / STTM latency
constexpr int kGemmThreads = 608;          // + the MMA warp + the activation-copy warp + the code-copy warp
constexpr int kCodeRing = 4;               // code chunks in flight per CTA (shared-memory ring filled by TMA)
constexpr int kCbSmemBytes = 2 * 256 * 8 * 16;     // [cb][idx][copy] x 16 B (same replication as dequant.cu)
constexpr int kATmemCols = kBK / 2;                // a k block of A in tensor memory: 128 lanes (rows) x 32 columns (2 fp16 each)
constexpr int kTmemCols = 512;                     // the whole tensor memory of the SM: accumulator columns, then the A ring

struct GemmParams {
    const uint8_t* codes;   // [O][G][2]
    const void* x;          // [batch][I]
    const uint4* codebook;  // [2][256] x 16 B
    const void* scales;     // [O] or null
    const void* bias;       // [O] or null
    void* y;                // [batch][O]
    unsigned int* counters; // [n_tiles] split-k arrival counters (zero between launches)
    float* partial;         // [ksplit][n_pad][O_pad] fp32 partial accumulators (ksplit > 1)
    int O, O_pad, I, batch, n_pad, nkb, ksplit, stages, a_col0;    // a_col0 = first tensor-memory column of the A ring (after the n_pad accumulator columns)
    int rpc;                // output rows this CTA owns (<= 128, multiple of 8): rows of the M=128 tile past it are never dequantized or stored
    int chunk;              // k blocks per staged code chunk: 8 (128 B per row) or 4
    unsigned long long* stamps;   // experiments only (lutvq_debug_set_gemm_stamps): [CTA][16]: %globaltimer stamps of thread 0 (0..6), cycle totals (8..12), or null
    int dbg;                // experiments only (lutvq_debug_set_gemm_mode): 1 = no MMAs, 2 = no dequantization, 4 = no activation copies
};

// ---- PTX wrappers (forms as in CUTLASS cute/arch/{copy_sm100,mma_sm100_umma,tmem_allocator_sm100}.hpp) ----
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}"
        :: "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar) : "memory");
}
// D (tensor memory) (+)= A (TENSOR MEMORY: lane = row, 8 columns = 16 K-major 16-bit elements) * B (shared memory descriptor):
// the four K=16 steps of one k block, executed by a CONVERGED warp behind ONE election (one lane issues); A advances by 8 columns,
// B's descriptor low word by 2 (32 bytes) per step
__device__ __forceinline__ void tc_mma4_f16_ts_elect(uint32_t tmem_d, uint32_t tmem_a, uint32_t b_lo, uint32_t b_hi, uint32_t idesc, uint32_t accumulate_first) {
    asm volatile(
        "{\n\t.reg .pred p, e, t;\n\t.reg .b64 d0, d1, d2, d3;\n\t.reg .b32 a1, a2, a3, l1, l2, l3;\n\t"
        "elect.sync _|e, 0xffffffff;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "setp.eq.b32 t, 0, 0;\n\t"
        "add.u32 a1, %1, 8;\n\tadd.u32 a2, %1, 16;\n\tadd.u32 a3, %1, 24;\n\t"
        "add.u32 l1, %2, 2;\n\tadd.u32 l2, %2, 4;\n\tadd.u32 l3, %2, 6;\n\t"
        "mov.b64 d0, {%2, %3};\n\tmov.b64 d1, {l1, %3};\n\tmov.b64 d2, {l2, %3};\n\tmov.b64 d3, {l3, %3};\n\t"
        "@e tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], d0, %4, p;\n\t"
        "@e tcgen05.mma.cta_group::1.kind::f16 [%0], [a1], d1, %4, t;\n\t"
        "@e tcgen05.mma.cta_group::1.kind::f16 [%0], [a2], d2, %4, t;\n\t"
        "@e tcgen05.mma.cta_group::1.kind::f16 [%0], [a3], d3, %4, t;\n\t}"
        :: "r"(tmem_d), "r"(tmem_a), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate_first) : "memory");
}
__device__ __forceinline__ void tc_commit_elect(uint32_t bar) {
    asm volatile(
        "{\n\t.reg .pred e;\n\t"
        "elect.sync _|e, 0xffffffff;\n\t"
        "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}"
        :: "r"(bar) : "memory");
}
// 8 consecutive 32-bit columns of this thread's lane (warp w reaches lanes 32*(w%4) .. +31); warp-collective
__device__ __forceinline__ void tc_st8(uint32_t taddr, const uint32_t (&v)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 :: "r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
}
__device__ __forceinline__ void tc_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}
// one box of the 2-D tensor map (coordinates: c0 = element along k, c1 = batch row) -> shared memory; its bytes complete on `bar`
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 :: "r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}

__device__ __forceinline__ uint4 lds128(uint32_t a) {       // read-only data (the codebook copies): free to be scheduled
    uint4 v;
    asm("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ uint4 lds128_sync(uint32_t a) {  // data another thread wrote before a barrier
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t lds32_sync(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts128(uint32_t a, const uint4 v) {
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" :: "r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// K-major SWIZZLE_128B descriptor: rows of 128 bytes, 8-row atoms 1024 bytes apart (SBO), LBO unused (1), layout_type 2;
// the atom base must be 1024-byte aligned; a K=16 step inside the 128-byte span advances the start address by 32 bytes.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | (1ull << 16) | ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// cute::UMMA::InstrDescriptor for kind::f16: D = F32, A/B = F16 or BF16, both K-major, dense
__device__ __forceinline__ uint32_t umma_idesc(int bf16, int M, int N) {
    return (1u << 4) | ((uint32_t)bf16 << 7) | ((uint32_t)bf16 << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// Timeline instrumentation (tools/prefill_timeline.py): compiled in only with -DLUTVQ_GEMM_TIMELINE (make EXTRA=-DLUTVQ_GEMM_TIMELINE);
// otherwise all of it folds away.
#ifdef LUTVQ_GEMM_TIMELINE
constexpr bool kTimeline = true;
#else
constexpr bool kTimeline = false;
#endif
__device__ __forceinline__ long long tl_clock() { return kTimeline ? clock64() : 0ll; }
__device__ __forceinline__ void gemm_stamp(const GemmParams& p, int i) {
    if (kTimeline && p.stamps != nullptr && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.stamps[((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 16 + i] = t;
    }
}
// cycle totals of one thread (experiments): slot i of the CTA's row
__device__ __forceinline__ void gemm_total(const GemmParams& p, int i, long long cycles) {
    if (kTimeline && p.stamps != nullptr) p.stamps[((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 16 + i] = (unsigned long long)cycles;
}

template <int DT, int P, int BPS>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_dequant_tc_kernel(const GemmParams p, const __grid_constant__ CUtensorMap xmap, const __grid_constant__ CUtensorMap cmap) {
    using D = Dtype<DT>;
    using T = typename D::T;
    using T2 = typename D::T2;
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    uint4* sh_cb = reinterpret_cast<uint4*>(smem);
    const int S = p.stages;
    const uint32_t sC = sbase + kCbSmemBytes;                               // ring of kCodeRing code chunks of 128 rows x P x 16 B
    constexpr uint32_t cbuf = (uint32_t)kBM * P * 16u;
    // a pipeline stage = BPS k blocks (every mbarrier round trip costs ~150 cycles per warp: small batches take 2 per stage)
    const uint32_t bblock = (uint32_t)p.n_pad * 128u, bstage = (uint32_t)BPS * bblock;
    const uint32_t sB = sC + (uint32_t)kCodeRing * cbuf;                   // activation stages (1024-byte aligned: SWIZZLE_128B)
    const uint32_t bars = sB + (uint32_t)S * bstage;
    auto full_bar  = [&](int i) { return bars + 8u * (uint32_t)i; };
    auto empty_bar = [&](int i) { return bars + 8u * (uint32_t)(S + i); };
    const uint32_t accum_bar = bars + 16u * (uint32_t)S;
    const uint32_t tmem_slot = accum_bar + 8u;
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem + (tmem_slot - sbase));
    volatile uint32_t* last_flag_ptr = tmem_slot_ptr + 1;
    auto code_full  = [&](int i) { return tmem_slot + 8u + 8u * (uint32_t)i; };
    auto code_empty = [&](int i) { return tmem_slot + 8u + 8u * (uint32_t)(kCodeRing + i); };

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int tile = blockIdx.x, z = blockIdx.y;
    const int row0 = tile * p.rpc;
    const int kb0 = (int)(((long long)z * p.nkb) / p.ksplit), kb1 = (int)(((long long)(z + 1) * p.nkb) / p.ksplit);

    // The code stream arrives per CHUNK of P k blocks = P*16 bytes of each of the CTA's rows: ONE TMA tensor copy per chunk
    // (box {P*16 bytes, rpc rows} of the [O][I/4] byte matrix, 64- / 128-byte swizzle = the XOR the per-row reads undo) into a
    // ring of kCodeRing shared buffers, issued by a warp of its own up to the whole ring ahead.  No register staging, no
    // block-wide barrier: a producer warp waits for the chunk's barrier, reads its rows' words and moves on at its own pace.
    // The first chunks are requested before anything else (weights: nothing to wait for).
    constexpr int lgP = P == 8 ? 3 : 2;
    const uint32_t chunk_bytes = (uint32_t)p.rpc * P * 16u;
    int cw_c = 0, cw_b = 0;                                              // code warp: next chunk / its ring slot
    uint32_t cw_par = 0u;
    auto issue_code = [&]() {                                            // one thread
        mbar_arrive_expect_tx(code_full(cw_b), chunk_bytes);
        tma_load_2d(sC + (uint32_t)cw_b * cbuf, &cmap, code_full(cw_b), (kb0 + cw_c * P) * 16, row0);
        ++cw_c;
        if (++cw_b == kCodeRing) cw_b = 0;
    };
    const int n_chunks = (kb1 - kb0 + P - 1) / P;
    if (tid == kProducerThreads + 64) {
        for (int i = 0; i < kCodeRing; ++i) { mbar_init(code_full(i), 1); mbar_init(code_empty(i), kProducerThreads / 32); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        while (cw_c < n_chunks && cw_c < kCodeRing) issue_code();
    }

    gemm_stamp(p, 0);
    // ---------------- one-time setup ----------------
    if (tid == kProducerThreads) {
        for (int i = 0; i < S; ++i) { mbar_init(full_bar(i), kProducerThreads / 32 + 1); mbar_init(empty_bar(i), 1); }
        mbar_init(accum_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kProducerThreads / 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(tmem_slot), "r"((uint32_t)kTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    } else if (warp < kProducerThreads / 32) {
        #pragma unroll 8
        for (int i = tid; i < 2 * 256 * 8; i += kProducerThreads) sh_cb[i] = __ldg(p.codebook + (i >> 3));
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;
    // Up to here only weights were read (code chunk 0, codebook).  The successor may be released right away -- it cannot
    // co-reside (shared memory, tensor memory), it takes the SM when this CTA exits; activations, outputs and the split-k
    // workspace are touched only once the predecessor has completed.
    gemm_stamp(p, 1);
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    gemm_stamp(p, 2);

    if (warp < kProducerThreads / 32) {
        const int row = tid & (kBM - 1), hk = tid >> 7;                  // this thread's row of the tile and quarter (2 groups = 4 code bytes) of every k block
        // =============================== producers: dequantize A, copy B ===============================
        const uint32_t cbA = sbase + (uint32_t)(lane & 7) * 16u;         // this lane's copy of codebook 0; entry i at + i*128
        const uint32_t cbB = cbA + 256u * 8u * 16u;
        auto swz = [](int r, int c) { return P == 8 ? (c ^ (r & 7)) : (c ^ ((r >> 1) & 3)); };
        uint32_t cur[P];
        int rb = 0;                                                      // ring slot of the chunk being read
        uint32_t rpar = 0u;
        auto read_chunk = [&]() {
            mbar_wait(code_full(rb), rpar);
            const uint32_t buf = sC + (uint32_t)rb * cbuf;
            #pragma unroll
            for (int j = 0; j < P; ++j) cur[j] = lds32_sync(buf + (uint32_t)((row << lgP) + swz(row, j)) * 16u + (uint32_t)hk * 4u);
            __syncwarp();
            if (lane == 0) mbar_arrive(code_empty(rb));                  // this warp's words are in registers: one arrival per warp
            if (++rb == kCodeRing) { rb = 0; rpar ^= 1u; }
        };
        // A goes to TENSOR MEMORY, not shared memory: lane = output row of the tile, a k block = 32 columns (two 16-bit
        // elements each, K-major), ring of S * BPS k blocks after the accumulator's columns.  A thread's 2 groups are 8
        // consecutive columns of its own lane: one tcgen05.st 32x32b.x8 per k block (a warp reaches lanes 32*(warp%4)..+31,
        // which is how rows are dealt to warps already).  The MMA reads it as its A operand ([tmem] form), so the
        // dequantized tile costs no shared-memory store and no shared-memory read by the tensor core.
        // Whole warps past the CTA's rows skip the work; dead rows inside a live warp run on zero codes (never stored).
        const bool warp_live = (warp & 3) * 32 < p.rpc;
        // Rows past the CTA's share inside a live warp: in the 1-k-block-per-stage build (> 128 columns), where the tensor core's reads of the
        // activation tile (32 KiB per k block) compete with the codebook reads for shared memory, their lookups are predicated off (28.5 -> 27.5 us
        // at 256 rows); in the other builds the branch costs more than the wavefronts (+0.7 us at 16 rows even when no lane takes it), so every
        // lane of a live warp looks up (zero codes).
        const uint32_t a_lane = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)p.a_col0 + (uint32_t)hk * 8u;
        long long t_empty = 0, t_stage = 0, t_st = 0;                    // experiments: cycles in the empty wait / code staging / STTM wait
        int s = 0;                                                       // stage of this k block
        uint32_t parE = 1u;                                              // parity to wait for on empty[s] (a fresh barrier passes parity 1)
        auto do_stage = [&](int kb, const uint32_t* cw) {                   // k blocks kb .. kb+BPS-1, code words cw[0 .. BPS-1]
            const long long tw0 = tl_clock();
            mbar_wait(empty_bar(s), parE);                               // the MMAs that read this stage's last contents have retired
            t_empty += tl_clock() - tw0;
            tc_fence_after();
            #pragma unroll
            for (int sb = 0; sb < BPS; ++sb) {
                if (kb + sb >= kb1) break;
                const uint32_t w = cw[sb];
                if (!(p.dbg & 2) && warp_live) {                         // 2 groups: 4 reads in flight, then adds and ONE tensor-memory store
                    uint4 a[2], b[2];
                    uint32_t v[8];
                    #pragma unroll
                    for (int g = 0; g < 2; ++g) a[g] = b[g] = make_uint4(0u, 0u, 0u, 0u);
                    if (BPS != 1 || row < p.rpc) {                                // > 128 columns: dead rows of a live warp (80 rows per CTA: rows 80..95) cost no wavefronts
                        #pragma unroll
                        for (int g = 0; g < 2; ++g) {                    // bytes 2g / 2g+1 of the quarter = codebook-0 / codebook-1 index
                            a[g] = lds128(cbA + __byte_perm(w, 0u, 0x4440u + 2u * g) * 128u);
                            b[g] = lds128(cbB + __byte_perm(w, 0u, 0x4441u + 2u * g) * 128u);
                        }
                    }
                    #pragma unroll
                    for (int g = 0; g < 2; ++g) {
                        const T2* a2 = reinterpret_cast<const T2*>(&a[g]);
                        const T2* b2 = reinterpret_cast<const T2*>(&b[g]);
                        #pragma unroll
                        for (int jj = 0; jj < 4; ++jj) {                 // column 4g + jj of the quarter = elements k = 8g + 2jj, + 1
                            const T2 r = D::add2(a2[jj], b2[jj]);
                            v[4 * g + jj] = *reinterpret_cast<const uint32_t*>(&r);
                        }
                    }
                    tc_st8(a_lane + (uint32_t)(s * BPS + sb) * kATmemCols, v);
                }
            }
            const long long ts0 = tl_clock();
            tc_wait_st();                                                // this warp's rows of the stage are in tensor memory
            t_st += tl_clock() - ts0;
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(full_bar(s));                     // one arrival per warp
            if (++s == S) { s = 0; parE ^= 1u; }
        };
        for (int kbc = kb0, c = 0; kbc < kb1; kbc += P, ++c) {
            const long long tg0 = tl_clock();
            read_chunk();
            t_stage += tl_clock() - tg0;
            #pragma unroll
            for (int j = 0; j < P; j += BPS)
                if (kbc + j < kb1) do_stage(kbc + j, &cur[j]);
            if (c == 0) gemm_stamp(p, 3);
        }

        // =============================== epilogue ===============================
        if (tid == 0) { gemm_total(p, 8, t_empty); gemm_total(p, 9, t_stage); gemm_total(p, 10, t_st); }
        gemm_stamp(p, 4);
        mbar_wait(accum_bar, 0u);
        tc_fence_after();
        gemm_stamp(p, 5);
        const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);   // a warp reaches TMEM lanes 32*(warp%4)..+31
        const int o = row0 + row;
        const bool live = row < p.rpc && o < p.O;                        // the two warps of a row split the batch columns
        T sc = D::from_f(1.f), bi = D::from_f(0.f);
        if (live && p.scales) sc = static_cast<const T*>(p.scales)[o];
        if (live && p.bias)   bi = static_cast<const T*>(p.bias)[o];
        auto finish = [&](float acc, int b) {
            T rr = D::from_f(acc);                                       // round(acc)              (Code2x8MatVec / F.linear output dtype)
            if (p.scales) rr = D::mul(rr, sc);                           // flat_output *= scales   (cuda_kernel.cpp:101)
            if (p.bias)   rr = D::add(rr, bi);                           // flat_output += bias     (cuda_kernel.cpp:103)
            static_cast<T*>(p.y)[(size_t)b * p.O + o] = rr;
        };
        if (p.ksplit == 1) {
            for (int c0 = hk * 16; c0 < p.n_pad; c0 += 64) {
                uint32_t v[16];
                tc_ld16(taddr + (uint32_t)c0, v);
                #pragma unroll
                for (int j = 0; j < 16; ++j)
                    if (live && c0 + j < p.batch) finish(__uint_as_float(v[j]), c0 + j);
            }
        } else {
            float* mine = p.partial + ((size_t)z * p.n_pad) * p.O_pad + o;
            for (int c0 = hk * 16; c0 < p.n_pad; c0 += 64) {
                uint32_t v[16];
                tc_ld16(taddr + (uint32_t)c0, v);
                #pragma unroll
                for (int j = 0; j < 16; ++j)
                    if (live && c0 + j < p.batch) __stcg(mine + (size_t)(c0 + j) * p.O_pad, __uint_as_float(v[j]));
            }
            __threadfence();
            asm volatile("bar.sync 1, 512;" ::: "memory");
            if (tid == 0) *last_flag_ptr = (atomicAdd(p.counters + tile, 1u) == (unsigned)p.ksplit - 1u) ? 1u : 0u;
            asm volatile("bar.sync 1, 512;" ::: "memory");
            if (*last_flag_ptr) {                                        // last CTA of this tile: fixed-order sum of the partials
                __threadfence();
                // The reduction is a read of ksplit x batch x rows fp32 partials by ONE CTA: bound by the bytes it keeps in flight.  Any
                // thread may take any element here, so a thread takes 4 consecutive rows (one 16-byte load per partial) of 4 batch
                // columns at a time: 512 threads x 4 x 16 B = 32 KiB in flight.  (One column of one row per round trip -- the first
                // version, the output stores keeping the compiler from hoisting the next loads -- cost 48 us for 256 rows of a
                // 4096 x 4096 layer; 8 scalar columns in flight 29.7.)  Per element the partials are still added in split order.
                constexpr int UB = 4;
                const int og = tid & 31, bq = tid >> 5;                  // rows row0 + 4 og .. + 3; batch columns bq, bq + 16, ...
                const int o4 = row0 + og * 4;
                if (og * 4 < p.rpc && o4 < p.O) {
                    T sc4[4], bi4[4];
                    #pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        sc4[e] = (p.scales && o4 + e < p.O) ? static_cast<const T*>(p.scales)[o4 + e] : D::from_f(1.f);
                        bi4[e] = (p.bias && o4 + e < p.O) ? static_cast<const T*>(p.bias)[o4 + e] : D::from_f(0.f);
                    }
                    for (int b0 = bq; b0 < p.batch; b0 += 16 * UB) {
                        float4 acc[UB];
                        #pragma unroll
                        for (int u = 0; u < UB; ++u) acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                        for (int zz = 0; zz < p.ksplit; ++zz) {
                            float4 v[UB];
                            #pragma unroll
                            for (int u = 0; u < UB; ++u) {
                                const int bb = b0 + 16 * u;
                                v[u] = bb < p.batch ? __ldcg(reinterpret_cast<const float4*>(p.partial + ((size_t)zz * p.n_pad + bb) * p.O_pad + o4))
                                                    : make_float4(0.f, 0.f, 0.f, 0.f);
                            }
                            #pragma unroll
                            for (int u = 0; u < UB; ++u) { acc[u].x += v[u].x; acc[u].y += v[u].y; acc[u].z += v[u].z; acc[u].w += v[u].w; }
                        }
                        #pragma unroll
                        for (int u = 0; u < UB; ++u) {
                            const int bb = b0 + 16 * u;
                            if (bb >= p.batch) continue;
                            const float f[4] = {acc[u].x, acc[u].y, acc[u].z, acc[u].w};
                            #pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                if (o4 + e >= p.O) continue;
                                T rr = D::from_f(f[e]);                  // the reference epilogue, as in finish()
                                if (p.scales) rr = D::mul(rr, sc4[e]);
                                if (p.bias)   rr = D::add(rr, bi4[e]);
                                static_cast<T*>(p.y)[(size_t)bb * p.O + o4 + e] = rr;
                            }
                        }
                    }
                }
                if (tid == 0) p.counters[tile] = 0u;                     // hand the counter back
            }
        }
    } else if (warp == kProducerThreads / 32 + 1) {
        // =============================== activation copies (warp 9, one thread) ===============================
        // Activation tile: SWIZZLE_128B K-major -- batch row n = 128 contiguous bytes (64 k), its 16-byte pieces XOR-permuted
        // by n & 7: exactly what the copy engine writes for a {64, n_pad} box of the [batch][I] tensor map with
        // CU_TENSOR_MAP_SWIZZLE_128B into a 1024-byte aligned block.  Per stage: wait until it is empty, post the byte
        // count on its full barrier, one tensor copy per k block; rows n >= batch lie outside the tensor and arrive as zeros.
        if (lane == 0) {
            int s = 0;
            uint32_t par = 1u;
            for (int kb = kb0; kb < kb1; kb += BPS) {
                mbar_wait(empty_bar(s), par);
                const int nblk = kb1 - kb < BPS ? kb1 - kb : BPS;
                if (p.dbg & 4) {
                    mbar_arrive(full_bar(s));
                } else {
                    mbar_arrive_expect_tx(full_bar(s), (uint32_t)nblk * bblock);
                    #pragma unroll
                    for (int b = 0; b < BPS; ++b)
                        if (b < nblk) tma_load_2d(sB + (uint32_t)s * bstage + (uint32_t)b * bblock, &xmap, full_bar(s), (kb + b) * kBK, 0);
                }
                if (++s == S) { s = 0; par ^= 1u; }
            }
        }
        __syncwarp();
    } else if (warp == kProducerThreads / 32 + 2) {
        // =============================== code copies (one thread): the chunks past the first ring ===============================
        if (lane == 0) {
            while (cw_c < n_chunks) {
                mbar_wait(code_empty(cw_b), cw_par);                     // every producer warp has read the chunk that was here
                const bool wrap = cw_b == kCodeRing - 1;
                issue_code();
                if (wrap) cw_par ^= 1u;
            }
        }
        __syncwarp();
    } else {
        // =============================== MMA issuer ===============================
        // The whole warp runs this loop CONVERGED and one elected lane issues (elect.sync inside the asm): under `if (lane == 0)`
        // the compiler wraps every tcgen05 instruction in an ELECT / vote / branch loop, and with the descriptors rebuilt per MMA
        // (shift, mask, or: a chain of dependent uniform-datapath ops) one MMA took ~96 cycles to ISSUE -- 256 of them were the
        // whole loop time at <= 64 columns.  The descriptors' low words now advance by adds: B by 2 per k step (32 bytes), by
        // bblock/16 per k block; A by 8 columns per k step; the 4 k steps of a k block share one election (14.1 -> 13.9 us at 16 rows).
        const uint32_t idesc = umma_idesc(DT == LUTVQ_BF16 ? 1 : 0, kBM, p.n_pad);
        const uint64_t bdesc0 = umma_desc_sw128(sB);
        const uint32_t bdesc_hi = (uint32_t)(bdesc0 >> 32), bstage16 = bstage >> 4, bblock16 = bblock >> 4;
        const bool do_mma = !(p.dbg & 1);
        uint32_t b_lo = (uint32_t)bdesc0;                                // low word of the stage's first descriptor
        uint32_t a_t = tmem_base + (uint32_t)p.a_col0;                   // the stage's first A column
        int s = 0;
        uint32_t par = 0u;
        long long t_full = 0;
        const long long tl0 = tl_clock();
        for (int kb = kb0; kb < kb1; kb += BPS) {
            const long long tf0 = tl_clock();
            mbar_wait(full_bar(s), par);
            t_full += tl_clock() - tf0;
            tc_fence_after();
            #pragma unroll
            for (int sb = 0; sb < BPS; ++sb) {
                if (kb + sb >= kb1) break;
                if (do_mma) tc_mma4_f16_ts_elect(tmem_base, a_t + (uint32_t)(sb * kATmemCols), b_lo + (uint32_t)sb * bblock16, bdesc_hi, idesc, kb + sb > kb0 ? 1u : 0u);
            }
            tc_commit_elect(empty_bar(s));                               // smem stage free once these MMAs have read it
            b_lo += bstage16;
            a_t += (uint32_t)(BPS * kATmemCols);
            if (++s == S) { s = 0; par ^= 1u; b_lo = (uint32_t)bdesc0; a_t = tmem_base + (uint32_t)p.a_col0; }
        }
        if (lane == 0) { gemm_total(p, 11, t_full); gemm_total(p, 12, tl_clock() - tl0); }
        tc_commit_elect(accum_bar);                                      // accumulator complete
    }

    gemm_stamp(p, 6);
    // ---------------- teardown ----------------
    tc_fence_before();
    __syncthreads();
    if (warp == kProducerThreads / 32)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "r"((uint32_t)kTmemCols) : "memory");
}

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------

static int g_gemm_dbg = 0;
void gemm_set_debug_mode(int m) { g_gemm_dbg = m; }
static unsigned long long* g_gemm_stamps = nullptr;
void gemm_set_stamps(void* dev_ptr) { g_gemm_stamps = static_cast<unsigned long long*>(dev_ptr); }   // no effect without LUTVQ_GEMM_TIMELINE

struct GemmPlan { int n_tiles, rpc, O_pad, nkb, ksplit; };

static GemmPlan plan_gemm(int O, int I, int sm_count) {
    GemmPlan pl;
    // The producers' shared-memory traffic (per ROW) bounds the kernel, the M=128 MMA is nearly free: when 128-row tiles
    // would leave SMs idle (11008 rows = 86 tiles on 148 SMs) every CTA takes fewer rows of its tile instead.
    pl.rpc = kBM;
    const int tiles128 = ceil_div(O, kBM);
    if (tiles128 * 2 > sm_count && !(g_gemm_dbg & 8)) {
        const int waves = ceil_div(tiles128, sm_count);
        pl.rpc = ceil_div(ceil_div(O, waves * sm_count), 8) * 8;
        if (pl.rpc > kBM) pl.rpc = kBM;
    }
    pl.n_tiles = ceil_div(O, pl.rpc);
    pl.O_pad = pl.n_tiles * pl.rpc;
    pl.nkb = I / kBK;
    int ks = sm_count / pl.n_tiles;                       // fill the SMs when there are few row tiles
    if (ks > pl.nkb / 4) ks = pl.nkb / 4;                 // at least 4 k blocks per CTA
    if (ks > 16) ks = 16;
    if (ks < 1) ks = 1;
    pl.ksplit = ks;
    return pl;
}

constexpr int kGemmMaxBatch = 256;                        // UMMA N <= 256; larger batches are chunked by the launcher

static size_t gemv_acc_bytes(int O) { return ((size_t)ceil_div(O, 32) * 32 * 8 + 255) / 256 * 256; }

// workspace = [256 B header][GEMV accumulator words, also the GEMM tile counters][GEMM split-k partials]
size_t gemm_workspace_bytes_for(int O, int I, int max_batch) {
    int sm = 148;
    DeviceInfo di;
    if (get_device_info(&di) == 0) sm = di.sm_count;
    const GemmPlan pl = plan_gemm(O, I, sm);
    if (max_batch <= 0) return 0;
    int nb = max_batch < kGemmMaxBatch ? max_batch : kGemmMaxBatch;
    nb = ceil_div(nb, 16) * 16;
    const size_t partial = pl.ksplit > 1 ? (size_t)pl.ksplit * nb * pl.O_pad * sizeof(float) : 0;
    return 256 + gemv_acc_bytes(O) + partial;
}
size_t gemm_workspace_bytes(int O, int I) { return gemm_workspace_bytes_for(O, I, kGemmMaxBatch); }

// cuTensorMapEncodeTiled, looked up once through the runtime (the library does not link libcuda)
typedef CUresult (*TensorMapEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                           const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                           CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static TensorMapEncodeTiledFn tensor_map_encoder() {
    static TensorMapEncodeTiledFn fn = [] {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) f = nullptr;
        return reinterpret_cast<TensorMapEncodeTiledFn>(f);
    }();
    return fn;
}

// tensor map of the activations x[batch][I] (16-bit elements) with a {64 k, n_pad rows} box, 128-byte swizzle
static int make_x_map(CUtensorMap* map, const void* x, int batch, int I, int n_pad, int dtype) {
    TensorMapEncodeTiledFn enc = tensor_map_encoder();
    if (!enc) return fail(LUTVQ_ECUDA, "cuTensorMapEncodeTiled not available from this driver");
    LUTVQ_REQUIRE(aligned16(x), "x must be 16-byte aligned");
    const cuuint64_t dims[2] = {(cuuint64_t)I, (cuuint64_t)batch};
    const cuuint64_t strides[1] = {(cuuint64_t)I * 2};
    const cuuint32_t box[2] = {(cuuint32_t)kBK, (cuuint32_t)n_pad};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(map, dtype == LUTVQ_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(x),
                           dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(LUTVQ_ECUDA, "cuTensorMapEncodeTiled failed (CUresult %d) for x[%d][%d], box {%d, %d}", (int)r, batch, I, kBK, n_pad);
    return LUTVQ_OK;
}

// tensor map of the codes as a byte matrix [O][I/4] with a {chunk*16 bytes, rpc rows} box; the swizzle (64 bytes for chunks of 4
// k blocks, 128 for 8) is the XOR the producers' per-row reads undo; rows past O arrive as zeros
static int make_code_map(CUtensorMap* map, const void* codes, int O, int I, int chunk, int rpc) {
    TensorMapEncodeTiledFn enc = tensor_map_encoder();
    if (!enc) return fail(LUTVQ_ECUDA, "cuTensorMapEncodeTiled not available from this driver");
    LUTVQ_REQUIRE(aligned16(codes), "codes must be 16-byte aligned");
    const cuuint64_t dims[2] = {(cuuint64_t)I / 4, (cuuint64_t)O};
    const cuuint64_t strides[1] = {(cuuint64_t)I / 4};
    const cuuint32_t box[2] = {(cuuint32_t)chunk * 16u, (cuuint32_t)rpc};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(codes), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           chunk == 8 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(LUTVQ_ECUDA, "cuTensorMapEncodeTiled failed (CUresult %d) for codes[%d][%d], box {%d, %d}", (int)r, O, I / 4, chunk * 16, rpc);
    return LUTVQ_OK;
}

template <int DT, int P, int BPS>
static int launch_gemm(const GemmParams& p, const CUtensorMap& xmap, const CUtensorMap& cmap, dim3 grid, size_t smem, cudaStream_t stream) {
    auto kern = gemm_dequant_tc_kernel<DT, P, BPS>;
    static thread_local int attr_device = -1;
    int dev = 0;
    LUTVQ_CUDA_OK(cudaGetDevice(&dev));
    if (attr_device != dev) {
        LUTVQ_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_device = dev;
    }
    // programmatic dependent launch: the CTAs of this launch take their SMs as the previous kernel's CTAs leave them (no
    // grid-wide drain + launch latency in between); everything before griddepcontrol.wait touches weights only
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = (p.dbg & 2048) ? 0 : 1;
    LUTVQ_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, p, xmap, cmap));
    return LUTVQ_OK;
}

int gemm_dequant_tc_launch(const void* x, const void* codes, const void* codebooks, const void* scales,
                           const void* bias, void* y, int batch, int O, int I, int dtype, void* ws, size_t ws_bytes,
                           cudaStream_t stream) {
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    if (dtype != LUTVQ_F16 && dtype != LUTVQ_BF16) return fail(LUTVQ_EUNSUPPORTED, "dtype %d not supported (fp16=0, bf16=1)", dtype);
    const GemmPlan pl = plan_gemm(O, I, di.sm_count);
    for (int b0 = 0; b0 < batch; b0 += kGemmMaxBatch) {
        const int nb = batch - b0 < kGemmMaxBatch ? batch - b0 : kGemmMaxBatch;
        GemmParams p;
        p.codes = static_cast<const uint8_t*>(codes);
        p.x = static_cast<const uint8_t*>(x) + (size_t)b0 * I * 2;
        p.codebook = static_cast<const uint4*>(codebooks);
        p.scales = scales; p.bias = bias;
        p.y = static_cast<uint8_t*>(y) + (size_t)b0 * O * 2;
        p.O = O; p.O_pad = pl.O_pad; p.I = I; p.batch = nb; p.rpc = pl.rpc;
        p.n_pad = ceil_div(nb, 16) * 16;
        p.nkb = pl.nkb; p.ksplit = pl.ksplit;
        p.dbg = g_gemm_dbg;
        p.stamps = g_gemm_stamps;
        p.a_col0 = ceil_div(p.n_pad, 32) * 32;                          // accumulator: columns [0, n_pad); A ring behind it
        int bps = p.n_pad <= 64 ? 4 : p.n_pad <= 128 ? 2 : 1;             // k blocks per pipeline stage (a stage hand-over costs ~0.25 us)
        if (g_gemm_dbg & 128) bps = 1;
        if (g_gemm_dbg & 256) bps = 2;
        if ((g_gemm_dbg & 512) && bps == 1) bps = 2;
        const size_t stage_bytes = (size_t)bps * (size_t)p.n_pad * 128;       // activations only: A lives in tensor memory
        p.chunk = (p.n_pad <= 128 && bps == 1) ? 8 : 4;
        const size_t code_bytes = (size_t)kCodeRing * kBM * p.chunk * 16;
        const size_t budget = (size_t)227 * 1024 - kCbSmemBytes - code_bytes - 1024;
        int S = (int)(budget / stage_bytes);
        if (S > 8) S = 8;
        const int s_tmem = (kTmemCols - p.a_col0) / (bps * kATmemCols);   // k blocks of A the tensor memory holds
        if (S > s_tmem) S = s_tmem;
        if (S < 2) return fail(LUTVQ_EUNSUPPORTED, "not enough shared memory for 2 pipeline stages");
        p.stages = S;
        size_t smem = kCbSmemBytes + code_bytes + (size_t)S * stage_bytes + 16 * (size_t)S + 32 + 16 * kCodeRing;
        if (smem < (size_t)116 * 1024) smem = (size_t)116 * 1024;         // one CTA per SM: each allocates the whole tensor memory
        p.counters = nullptr; p.partial = nullptr;
        if (pl.ksplit > 1) {
            const size_t need = gemm_workspace_bytes_for(O, I, nb);
            LUTVQ_REQUIRE(ws != nullptr && ws_bytes >= need, "workspace too small: need %zu bytes, got %zu", need, ws_bytes);
            LUTVQ_REQUIRE(aligned16(ws), "workspace must be 16-byte aligned");
            p.counters = reinterpret_cast<unsigned int*>(static_cast<uint8_t*>(ws) + 256);
            p.partial = reinterpret_cast<float*>(static_cast<uint8_t*>(ws) + 256 + gemv_acc_bytes(O));
        }
        const dim3 grid(pl.n_tiles, pl.ksplit);
        CUtensorMap xmap, cmap;
        if (int rc = make_x_map(&xmap, p.x, nb, I, p.n_pad, dtype)) return rc;
        if (int rc = make_code_map(&cmap, codes, O, I, p.chunk, p.rpc)) return rc;
        int rc;
        if (bps == 4)          rc = dtype == LUTVQ_F16 ? launch_gemm<LUTVQ_F16, 4, 4>(p, xmap, cmap, grid, smem, stream) : launch_gemm<LUTVQ_BF16, 4, 4>(p, xmap, cmap, grid, smem, stream);
        else if (bps == 2)     rc = dtype == LUTVQ_F16 ? launch_gemm<LUTVQ_F16, 4, 2>(p, xmap, cmap, grid, smem, stream) : launch_gemm<LUTVQ_BF16, 4, 2>(p, xmap, cmap, grid, smem, stream);
        else if (p.chunk == 8) rc = dtype == LUTVQ_F16 ? launch_gemm<LUTVQ_F16, 8, 1>(p, xmap, cmap, grid, smem, stream) : launch_gemm<LUTVQ_BF16, 8, 1>(p, xmap, cmap, grid, smem, stream);
        else                   rc = dtype == LUTVQ_F16 ? launch_gemm<LUTVQ_F16, 4, 1>(p, xmap, cmap, grid, smem, stream) : launch_gemm<LUTVQ_BF16, 4, 1>(p, xmap, cmap, grid, smem, stream);
        if (rc) return rc;
    }
    return LUTVQ_OK;
}

}  // namespace lutvq
