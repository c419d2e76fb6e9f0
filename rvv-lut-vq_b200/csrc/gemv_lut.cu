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
// Grid = (row chunks, k-slices), at most one CTA per SM so that the whole grid is co-resident.  The split-k
// partial sums travel through L2 as 8-byte {value, epoch} words (flag-in-data, no fence / no atomic on the
// critical path); the 32-row sets of a row chunk are dealt round-robin to the (k-slice CTA, warp) pairs of
// that chunk, which poll the n_k partials of a set, add them IN FIXED k ORDER (deterministic) and apply the
// reference epilogue order (round to dtype, * scale, + bias; cuda_kernel.cpp:95-111).
//
// Two code layouts feed the same gather loop:
//   reference  int8 [O][G][2] exactly as QuantizedLinear.codes: a warp's 128-bit load touches 8 rows x 64 B;
//   packed     lutvq_code2x8_prepack(): [k-slice][32-row set][8-row tile][lane] x 16 B, the lane's register
//              image, so a warp's load is 512 contiguous bytes and a CTA's whole share is one contiguous range.
//
// Algorithmic HBM bytes per launch: O*(I/8)*2 codes + 8192 codebooks + 2I x + 2O scales + 2O y.
#include "common.cuh"
#include "lut_layout.h"

namespace lutvq {

struct WsHeader {               // first 256 bytes of the workspace
    unsigned int epoch;         // bumped once per launch by the last CTA to finish; partial tags are epoch + 1
    unsigned int done;          // CTAs finished in the current launch
};
constexpr size_t kWsHeaderBytes = 256;

struct GemvParams {
    const uint8_t* codes;     // reference layout [O][G][2]  or packed layout (see above)
    const void* x;            // [I]
    const void* codebook;     // [2][256][8]
    void* y;                  // [O]
    const void* scales;       // [O] or null
    const void* bias;         // [O] or null
    WsHeader* hdr;
    uint2* partial;           // [n_k][O_pad] {value bits, tag}
    int O, O_pad, G, n_k, sets_per_chunk, n_sets_total;
    unsigned flags;
};

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

__device__ __forceinline__ float lds_f32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts_v4(uint32_t addr, float a, float b, float c, float d) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" :: "r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
// flag-in-data partial sums: one 8-byte word is written / read as a unit
__device__ __forceinline__ void st_partial(uint2* p, float v, uint32_t tag) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" :: "l"(p), "r"(__float_as_uint(v)), "r"(tag) : "memory");
}
__device__ __forceinline__ uint2 ld_partial(const uint2* p) {
    uint2 r;
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p) : "memory");
    return r;
}
// asynchronous HBM -> L2 prefetch of a contiguous range (no registers, no shared memory)
__device__ __forceinline__ void l2_prefetch_bulk(const void* p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(p), "r"(bytes) : "memory");
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

// One warp tile set = 32 rows x 64 bytes: 4 tiles of 8 rows; per lane 4 x 16 bytes.
struct CodeRegs { uint32_t w[4][4]; };

template <int DT, bool PACKED, int NW, int PF>
__global__ void __launch_bounds__(NW * 32, 2)
lut_gemv_2x8_kernel(const GemvParams p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t lut_base = smem_u32(smem_raw);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int chunk = blockIdx.x;
    const int my = blockIdx.y, ny = gridDim.y;            // this CTA owns k-slices my, my+ny, ... (ny == n_k unless n_k > #SMs)
    const int set0 = chunk * p.sets_per_chunk;            // first 32-row set of this row chunk
    const int n_sets = min(p.sets_per_chunk, p.n_sets_total - set0);
    const int row0 = set0 * kSetRows;
    const size_t row_bytes = (size_t)p.G * 2;
    const int last_row = p.O - 1;

    // ---------------- streaming-side lane constants ----------------
    // Reference layout: no predicated loads.  Rows past the end are clamped to the last row (result discarded)
    // and lanes whose 16-byte chunk lies past the end of the row (last, partial k-slice) re-read chunk 0 --
    // their tables were built from x = 0, so whatever index they fetch contributes exactly 0.
    const int e = stream_first_half(lane);
    auto load_set = [&](CodeRegs& r, int ks, int s) {
        if (PACKED) {
            const uint4* base = reinterpret_cast<const uint4*>(p.codes) +
                                (((size_t)ks * p.n_sets_total + (set0 + s)) * 4) * 32 + lane;
            #pragma unroll
            for (int t = 0; t < 4; ++t) {
                const uint4 v = ldg_stream_u128(base + t * 32);
                r.w[t][0] = v.x; r.w[t][1] = v.y; r.w[t][2] = v.z; r.w[t][3] = v.w;
            }
        } else {
            const int slice_byte0 = ks * kSliceTables;
            const bool chunk_valid = (size_t)(slice_byte0 + 16 * stream_chunk(lane)) < row_bytes;
            const uint8_t* lane_base = p.codes + slice_byte0 + (chunk_valid ? 16 * stream_chunk(lane) : 0);
            const int rbase = row0 + s * kSetRows + stream_row_in_tile(lane);
            #pragma unroll
            for (int t = 0; t < 4; ++t) {
                const uint32_t row = (uint32_t)min(rbase + t * kTileRows, last_row);
                const uint4 v = ldg_stream_u128(lane_base + (uint64_t)row * (uint32_t)row_bytes);
                // register word c holds memory word c ^ 2e (lut_layout.h)
                r.w[t][0] = e ? v.z : v.x; r.w[t][1] = e ? v.w : v.y;
                r.w[t][2] = e ? v.x : v.z; r.w[t][3] = e ? v.y : v.w;
            }
        }
    };

    // ---------------- prologue that does not depend on the previous kernel in the stream ----------------
    // Codes and codebooks are weights.  (1) the first PF tile sets are requested into registers, (2) the rest
    // of this CTA's share is pushed HBM -> L2 by bulk prefetches, (3) the codebook MMA fragments are loaded --
    // all BEFORE the programmatic-dependency wait, i.e. while the previous kernel is still running.
    CodeRegs buf[PF];
    #pragma unroll
    for (int i = 0; i < PF; ++i)
        if (warp + i * NW < n_sets) load_set(buf[i], my, warp + i * NW);
    if (PACKED) {
        for (int ks = my; ks < p.n_k; ks += ny) {
            const uint8_t* base = p.codes + ((size_t)ks * p.n_sets_total + set0) * 2048;
            for (int s = threadIdx.x; s < n_sets; s += NW * 32) l2_prefetch_bulk(base + (size_t)s * 2048, 2048);
        }
    } else {
        const int rows_here = min(n_sets * kSetRows, p.O - row0);
        for (int ks = my; ks < p.n_k; ks += ny) {
            const size_t off = (size_t)ks * kSliceTables;
            const uint32_t seg = (uint32_t)min((size_t)kSliceTables, row_bytes - off);
            for (int r = threadIdx.x; r < rows_here; r += NW * 32)
                l2_prefetch_bulk(p.codes + (size_t)(row0 + r) * row_bytes + off, seg);
        }
    }
    const int r = lane >> 2, q = lane & 3;
    constexpr int MB_PER_WARP = 16 / NW;                                   // 16 idx blocks of 16 per codebook
    uint32_t afrag[MB_PER_WARP][2][2];
    {
        const uint32_t* cb32 = static_cast<const uint32_t*>(p.codebook);  // [2][256][4] words (2 values each)
        // A fragments for this warp's idx blocks: rows 16m + r and 16m + r + 8, k = 2q, 2q+1
        #pragma unroll
        for (int mi = 0; mi < MB_PER_WARP; ++mi) {
            const int m = warp * MB_PER_WARP + mi;
            #pragma unroll
            for (int n = 0; n < 2; ++n) {
                afrag[mi][n][0] = __ldg(cb32 + (n * 256 + 16 * m + r) * 4 + q);
                afrag[mi][n][1] = __ldg(cb32 + (n * 256 + 16 * m + r + 8) * 4 + q);
            }
        }
    }
    pdl_launch_dependents();    // the next kernel may start ITS weight prefetch now (it still waits for us below)
    pdl_wait();                 // x and the workspace belong to the previous kernel until here
    const uint32_t tag = ldcg_u32(&p.hdr->epoch) + 1u;

    // PRMT operands (lut_layout.h): byte0 <- lm, byte1 <- code byte, bytes 2,3 <- 0.  The slice base and the
    // (c&1)*128 half-select are compile-time constants for ptxas and ride in the LDS immediate.
    uint32_t sel[4], lm[2][4];
    #pragma unroll
    for (int d = 0; d < 4; ++d) {
        sel[d] = stream_sel(lane, d);
        lm[0][d] = stream_lm(lane, 0, d);
        lm[1][d] = stream_lm(lane, 1, d);
    }
    const bool b3 = (lane >> 3) & 1, b4 = (lane >> 4) & 1;
    const bool odd = r & 1;
    const bool round16 = (p.flags & LUTVQ_FLAG_LUT_FP16) != 0;

    for (int ks = my; ks < p.n_k; ks += ny) {
        // ---------------- K1: build this CTA's 64 KiB LUT slice ----------------
        {
            // B fragments: column r of N-tile j is local group build_group_of_col(j, r); k = 2q, 2q+1
            const uint32_t* x32 = static_cast<const uint32_t*>(p.x);              // 4 words per group
            uint32_t bfrag[4];
            #pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int g = ks * 32 + build_group_of_col(j, r);
                bfrag[j] = (g < p.G) ? ldcg_u32(x32 + g * 4 + q) : 0u;   // L2-coherent: x is the previous kernel's output
            }
            #pragma unroll
            for (int mi = 0; mi < MB_PER_WARP; ++mi) {
                const int m = warp * MB_PER_WARP + mi;
                float c[2][4][4];
                #pragma unroll
                for (int n = 0; n < 2; ++n)
                    #pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        c[n][j][0] = c[n][j][1] = c[n][j][2] = c[n][j][3] = 0.f;
                        MmaK8<DT>::run(c[n][j], afrag[mi][n][0], afrag[mi][n][1], bfrag[j]);
                    }
                if (round16) {
                    #pragma unroll
                    for (int n = 0; n < 2; ++n)
                        #pragma unroll
                        for (int j = 0; j < 4; ++j)
                            #pragma unroll
                            for (int i = 0; i < 4; ++i) c[n][j][i] = __half2float(__float2half_rn(c[n][j][i]));
                }
                // c[n][j][2h+u]: idx = 16m + r + 8h, table bytes beta = 2u+n of (chunk q, word w_j)
                #pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int idx = 16 * m + r + 8 * h;
                    #pragma unroll
                    for (int jp = 0; jp < 2; ++jp) {
                        const int j0 = 2 * jp, j1 = 2 * jp + 1;
                        // first store: odd lanes take tile j1, even lanes tile j0 (8 distinct 16-byte columns per phase)
                        float f0 = odd ? c[0][j1][2 * h] : c[0][j0][2 * h];
                        float f1 = odd ? c[1][j1][2 * h] : c[1][j0][2 * h];
                        float f2 = odd ? c[0][j1][2 * h + 1] : c[0][j0][2 * h + 1];
                        float f3 = odd ? c[1][j1][2 * h + 1] : c[1][j0][2 * h + 1];
                        float s0 = odd ? c[0][j0][2 * h] : c[0][j1][2 * h];
                        float s1 = odd ? c[1][j0][2 * h] : c[1][j1][2 * h];
                        float s2 = odd ? c[0][j0][2 * h + 1] : c[0][j1][2 * h + 1];
                        float s3 = odd ? c[1][j0][2 * h + 1] : c[1][j1][2 * h + 1];
                        const uint32_t first  = lut_base + build_store_offset(idx, odd ? j1 : j0, q);
                        const uint32_t second = lut_base + build_store_offset(idx, odd ? j0 : j1, q);
                        sts_v4(first, f0, f1, f2, f3);
                        sts_v4(second, s0, s1, s2, s3);
                    }
                }
            }
        }
        __syncthreads();

        // ---------------- K2: stream codes, gather-accumulate ----------------
        const int ks_next = ks + ny;
        for (int it = 0; it * NW + warp < n_sets; it += PF) {
            #pragma unroll
            for (int st = 0; st < PF; ++st) {
                const int s = (it + st) * NW + warp;
                if (s < n_sets) {                                   // warp-uniform
                    float acc[4][2];
                    #pragma unroll
                    for (int c = 0; c < 4; ++c)
                        #pragma unroll
                        for (int d = 0; d < 4; ++d)
                            #pragma unroll
                            for (int t = 0; t < 4; ++t) {
                                const uint32_t ad = __byte_perm(buf[st].w[t][c], lm[c >> 1][d], sel[d]);
                                const float v = lds_f32(lut_base + ad + stream_imm(c));
                                if (c == 0 && d < 2) acc[t][d] = v;
                                else                 acc[t][d & 1] += v;
                            }
                    // refill this stage with the set PF iterations ahead before the (short) reduction
                    if (s + PF * NW < n_sets) load_set(buf[st], ks, s + PF * NW);
                    float a0 = acc[0][0] + acc[0][1], a1 = acc[1][0] + acc[1][1];
                    float a2 = acc[2][0] + acc[2][1], a3 = acc[3][0] + acc[3][1];
                    // transposed reduce over the 4 chunk lanes (lane bits 3,4): lane ends with tile (2*b4+b3), row lane&7
                    float k0 = b4 ? a2 : a0, k1 = b4 ? a3 : a1;
                    float g0 = b4 ? a0 : a2, g1 = b4 ? a1 : a3;
                    k0 += __shfl_xor_sync(0xffffffffu, g0, 16);
                    k1 += __shfl_xor_sync(0xffffffffu, g1, 16);
                    float keep = b3 ? k1 : k0, give = b3 ? k0 : k1;
                    keep += __shfl_xor_sync(0xffffffffu, give, 8);
                    const int row = row0 + s * kSetRows + lane;
                    if (p.n_k == 1) { if (row < p.O) epilogue_store<DT>(p.y, p.scales, p.bias, row, keep); }
                    else            st_partial(p.partial + (size_t)ks * p.O_pad + row, keep, tag);
                }
            }
        }
        if (ks_next < p.n_k) {                                     // only when n_k exceeds the SM count
            #pragma unroll
            for (int i = 0; i < PF; ++i)
                if (warp + i * NW < n_sets) load_set(buf[i], ks_next, warp + i * NW);
            __syncthreads();                                       // everyone is done reading this slice's LUT
        }
    }

    if (p.n_k > 1) {
        // ---------------- deterministic split-k reduction, dealt over the (k-slice CTA, warp) pairs of the chunk ----------------
        for (int j = my + ny * warp; j < n_sets; j += ny * NW) {
            const int row = row0 + j * kSetRows + lane;
            const uint2* src = p.partial + row;
            float acc = 0.f;
            for (int k0 = 0; k0 < p.n_k; k0 += 8) {
                uint2 v[8];
                bool ok;
                do {
                    ok = true;
                    #pragma unroll
                    for (int i = 0; i < 8; ++i)
                        if (k0 + i < p.n_k) {
                            v[i] = ld_partial(src + (size_t)(k0 + i) * p.O_pad);
                            ok = ok && (v[i].y == tag);
                        }
                } while (!__all_sync(0xffffffffu, ok));
                #pragma unroll
                for (int i = 0; i < 8; ++i)
                    if (k0 + i < p.n_k) acc += __uint_as_float(v[i].x);
            }
            if (row < p.O) epilogue_store<DT>(p.y, p.scales, p.bias, row, acc);
        }
    }
    // ---------------- last CTA out bumps the epoch (every CTA has read it by then) ----------------
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned total = gridDim.x * gridDim.y;
        const unsigned prev = atomicAdd(&p.hdr->done, 1u);
        if (prev == total - 1) {
            p.hdr->done = 0u;
            p.hdr->epoch = tag;
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
struct GemvTuning { int nw = 8, pf = 2, ctas_per_sm = 1; };
static GemvTuning g_tuning;

struct GemvPlan { int n_k, ny, n_chunks, sets_per_chunk, n_sets_total; };

static GemvPlan plan_gemv(int O, int G, int sm_count) {
    GemvPlan pl;
    pl.n_k = ceil_div(G * 2, kSliceTables);
    pl.ny = pl.n_k < sm_count ? pl.n_k : sm_count;      // whole grid co-resident: the split-k reducers poll their peers
    pl.n_sets_total = ceil_div(O, kSetRows);
    int n_r = sm_count / pl.ny;
    if (n_r < 1) n_r = 1;
    pl.sets_per_chunk = ceil_div(pl.n_sets_total, n_r);
    pl.n_chunks = ceil_div(pl.n_sets_total, pl.sets_per_chunk);
    return pl;
}

size_t gemv_workspace_bytes(int O, int G) {
    const int n_k = ceil_div(G * 2, kSliceTables);
    const size_t O_pad = (size_t)ceil_div(O, kSetRows) * kSetRows;
    return kWsHeaderBytes + (size_t)n_k * O_pad * sizeof(uint2);
}

size_t gemv_packed_bytes(int O, int G) {
    const int n_k = ceil_div(G * 2, kSliceTables);
    return (size_t)n_k * ceil_div(O, kSetRows) * 2048;
}

int gemv_prepack_launch(const void* codes, void* packed, int O, int I, cudaStream_t stream) {
    const int G = I / 8;
    const int n_k = ceil_div(G * 2, kSliceTables), n_sets_total = ceil_div(O, kSetRows);
    const size_t total = (size_t)n_k * n_sets_total * 128;
    int grid = (int)((total + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    prepack_2x8_kernel<<<grid, 256, 0, stream>>>(static_cast<const uint8_t*>(codes), static_cast<uint4*>(packed), O, G, n_k, n_sets_total);
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

template <int DT, bool PACKED, int NW, int PF>
static int launch_variant(const GemvParams& p, dim3 grid, cudaStream_t stream, bool pdl) {
    auto kern = lut_gemv_2x8_kernel<DT, PACKED, NW, PF>;
    static thread_local int attr_device = -1;
    int dev = 0;
    LUTVQ_CUDA_OK(cudaGetDevice(&dev));
    if (attr_device != dev) {
        LUTVQ_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSliceBytes));
        attr_device = dev;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(NW * 32);
    cfg.dynamicSmemBytes = kSliceBytes;
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
static int launch_tuned(const GemvParams& p, dim3 grid, cudaStream_t stream, bool pdl) {
    const GemvTuning t = g_tuning;
    if (t.nw == 8 && t.pf == 2)  return launch_variant<DT, PACKED, 8, 2>(p, grid, stream, pdl);
    if (t.nw == 8 && t.pf == 3)  return launch_variant<DT, PACKED, 8, 3>(p, grid, stream, pdl);
    if (t.nw == 8 && t.pf == 1)  return launch_variant<DT, PACKED, 8, 1>(p, grid, stream, pdl);
    if (t.nw == 16 && t.pf == 1) return launch_variant<DT, PACKED, 16, 1>(p, grid, stream, pdl);
    if (t.nw == 16 && t.pf == 2) return launch_variant<DT, PACKED, 16, 2>(p, grid, stream, pdl);
    if (t.nw == 4 && t.pf == 2)  return launch_variant<DT, PACKED, 4, 2>(p, grid, stream, pdl);
    if (t.nw == 4 && t.pf == 4)  return launch_variant<DT, PACKED, 4, 4>(p, grid, stream, pdl);
    return fail(LUTVQ_EINVAL, "no kernel variant for tuning nw=%d pf=%d", t.nw, t.pf);
}

// One GEMV: y = epilogue(W(codes) . x).  Validation is done by the caller (api.cu).
int gemv_2x8_launch(const void* codes, int packed, const void* x, void* y, const void* codebook, const void* scales,
                    const void* bias, int O, int I, int dtype, unsigned flags, void* ws, size_t ws_bytes,
                    cudaStream_t stream) {
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    const int G = I / 8;
    const GemvPlan pl = plan_gemv(O, G, di.sm_count);
    const size_t need = gemv_workspace_bytes(O, G);
    LUTVQ_REQUIRE(ws != nullptr && ws_bytes >= need, "workspace too small: need %zu bytes, got %zu", need, ws_bytes);
    LUTVQ_REQUIRE(aligned16(ws), "workspace must be 16-byte aligned");
    GemvParams p;
    p.codes = static_cast<const uint8_t*>(codes);
    p.x = x; p.codebook = codebook; p.y = y; p.scales = scales; p.bias = bias;
    p.hdr = static_cast<WsHeader*>(ws);
    p.partial = reinterpret_cast<uint2*>(static_cast<uint8_t*>(ws) + kWsHeaderBytes);
    p.O = O; p.O_pad = pl.n_sets_total * kSetRows; p.G = G; p.n_k = pl.n_k;
    p.sets_per_chunk = pl.sets_per_chunk; p.n_sets_total = pl.n_sets_total; p.flags = flags;
    const dim3 grid(pl.n_chunks, pl.ny);
    const bool pdl = !(flags & LUTVQ_FLAG_NO_PDL);
    if (dtype == LUTVQ_F16)  return packed ? launch_tuned<LUTVQ_F16, true>(p, grid, stream, pdl)  : launch_tuned<LUTVQ_F16, false>(p, grid, stream, pdl);
    if (dtype == LUTVQ_BF16) return packed ? launch_tuned<LUTVQ_BF16, true>(p, grid, stream, pdl) : launch_tuned<LUTVQ_BF16, false>(p, grid, stream, pdl);
    return fail(LUTVQ_EUNSUPPORTED, "dtype %d not supported (fp16=0, bf16=1)", dtype);
}

void gemv_set_tuning(int nw, int pf, int ctas_per_sm) {
    if (nw > 0) g_tuning.nw = nw;
    if (pf > 0) g_tuning.pf = pf;
    if (ctas_per_sm > 0) g_tuning.ctas_per_sm = ctas_per_sm;
}
void gemv_get_plan(int O, int I, int* n_k, int* n_chunks, int* rows_per_chunk) {
    DeviceInfo di;
    int sm = 148;
    if (get_device_info(&di) == 0) sm = di.sm_count;
    GemvPlan pl = plan_gemv(O, I / 8, sm);
    if (n_k) *n_k = pl.n_k;
    if (n_chunks) *n_chunks = pl.n_chunks;
    if (rows_per_chunk) *rows_per_chunk = pl.sets_per_chunk * kSetRows;
}

}  // namespace lutvq
