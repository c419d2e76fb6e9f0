// gemv_lut.cu -- fused LUT-build + code-indexed gather-accumulate GEMV for AQLM 2x8 codes (sm_100a).
//
// Replaces Code2x8MatVec (reference benchmark/AQLM/inference_lib/src/aqlm/inference_kernels/
// cuda_kernel.cu:144-233) with the LUT formulation of the golden model (vq_dataflow_sim/vq.py:280-298):
//
//   K1  L[g][n][k] = sum_j C[n][k][j] * x[8g+j]      built per CTA for ITS 32-group k-slice with
//                                                     mma.sync.m16n8k8 (fp32 accumulate) into 64 KiB of smem
//   K2  y[o] = sum_g sum_n L[g][n][code(o,g,n)]       codes streamed from HBM straight into registers
//                                                     (2 x 64-bit no-allocate loads per lane per 8-row tile),
//                                                     one PRMT + one LDS + one FADD per code byte,
//                                                     conflict-free for any index pattern (lut_layout.h)
//
// Grid = (row chunks, k-slices).  Every CTA writes fp32 partial sums for its rows; the LAST k-slice CTA
// to arrive for a row chunk (atomic ticket) adds the partials in fixed k order -- deterministic -- and
// applies the reference epilogue order (round to dtype, * scale, + bias; cuda_kernel.cpp:95-111).
//
// Algorithmic HBM bytes per launch: O*(I/8)*2 codes + 8192 codebooks + 2I x + 2O scales + 2O y.
#include "common.cuh"
#include "lut_layout.h"

namespace lutvq {

struct GemvParams {
    const uint8_t* codes;     // [O][G][2]
    const void* x;            // [I]
    const void* codebook;     // [2][256][8]
    void* y;                  // [O]
    const void* scales;       // [O] or null
    const void* bias;         // [O] or null
    float* partial;           // [n_k][O]
    unsigned int* counters;   // [n_chunks]
    int O, G, n_k, rows_per_chunk;
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

template <int DT>
__device__ __forceinline__ void epilogue_store(void* y, const void* scales, const void* bias, int row, float acc) {
    using D = Dtype<DT>;
    using T = typename D::T;
    T r = D::from_f(acc);                                                  // Code2x8MatVec writes round(acc)   (cuda_kernel.cu:221-231)
    if (scales) r = D::mul(r, static_cast<const T*>(scales)[row]);         // flat_output *= scales             (cuda_kernel.cpp:101)
    if (bias)   r = D::add(r, static_cast<const T*>(bias)[row]);           // flat_output += bias               (cuda_kernel.cpp:103)
    static_cast<T*>(y)[row] = r;
}

// One warp tile set = 32 rows x 64 bytes: 4 tiles of 8 rows; per lane 4 x (2 x 8 bytes).
struct CodeRegs { uint32_t w[4][4]; };

template <int DT, int NW, int PF, int MINB>
__global__ void __launch_bounds__(NW * 32, MINB)
lut_gemv_2x8_kernel(const GemvParams p) {
    using D = Dtype<DT>;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ unsigned int s_is_last;
    const uint32_t lut_base = smem_u32(smem_raw);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int chunk = blockIdx.x, ks = blockIdx.y;
    const int row0 = chunk * p.rows_per_chunk;
    const int row_end = min(row0 + p.rows_per_chunk, p.O);
    const int n_sets = (row_end - row0 + kSetRows - 1) / kSetRows;
    const size_t row_bytes = (size_t)p.G * 2;
    const int slice_byte0 = ks * kSliceTables;

    // ---------------- streaming-side lane constants ----------------
    // No predicated loads: rows past the end are clamped to the last row (result discarded) and lanes whose
    // 16-byte chunk lies past the end of the row (last, partial k-slice) re-read chunk 0 -- their tables were
    // built from x = 0, so whatever index they fetch contributes exactly 0.
    const int e = stream_first_half(lane);
    const bool chunk_valid = (size_t)(slice_byte0 + 16 * stream_chunk(lane)) < row_bytes;
    const uint8_t* lane_base = p.codes + slice_byte0 + (chunk_valid ? 16 * stream_chunk(lane) : 0) + 8 * e;
    const int second_half_off = 8 - 16 * e;
    const uint32_t row_bytes32 = (uint32_t)row_bytes;
    const int last_row = p.O - 1;

    auto load_set = [&](CodeRegs& r, int s) {
        const int rbase = row0 + s * kSetRows + stream_row_in_tile(lane);
        #pragma unroll
        for (int t = 0; t < 4; ++t) {
            const uint32_t row = (uint32_t)min(rbase + t * kTileRows, last_row);
            const uint8_t* ptr = lane_base + (uint64_t)row * row_bytes32;
            const uint2 lo = ldg_stream_u64(ptr);
            const uint2 hi = ldg_stream_u64(ptr + second_half_off);
            r.w[t][0] = lo.x; r.w[t][1] = lo.y; r.w[t][2] = hi.x; r.w[t][3] = hi.y;
        }
    };

    // Codes and codebooks are weights: they do not depend on the previous kernel in the stream, so the
    // first PF tile sets are requested BEFORE the programmatic-dependency wait (x may still be in flight).
    CodeRegs buf[PF];
    #pragma unroll
    for (int i = 0; i < PF; ++i)
        if (warp + i * NW < n_sets) load_set(buf[i], warp + i * NW);

    // ---------------- K1: build this CTA's 64 KiB LUT slice ----------------
    {
        const int r = lane >> 2, q = lane & 3;
        const uint32_t* cb32 = static_cast<const uint32_t*>(p.codebook);      // [2][256][4] words (2 values each)
        // A fragments for this warp's idx blocks: rows 16m + r and 16m + r + 8, k = 2q, 2q+1
        constexpr int MB_PER_WARP = 16 / NW;                                   // 16 idx blocks of 16 per codebook
        uint32_t afrag[MB_PER_WARP][2][2];
        #pragma unroll
        for (int mi = 0; mi < MB_PER_WARP; ++mi) {
            const int m = warp * MB_PER_WARP + mi;
            #pragma unroll
            for (int n = 0; n < 2; ++n) {
                afrag[mi][n][0] = __ldg(cb32 + (n * 256 + 16 * m + r) * 4 + q);
                afrag[mi][n][1] = __ldg(cb32 + (n * 256 + 16 * m + r + 8) * 4 + q);
            }
        }
        pdl_wait();                 // x (and the shared workspace) belong to the previous kernel until here
        pdl_launch_dependents();
        // B fragments: column r of N-tile j is local group build_group_of_col(j, r); k = 2q, 2q+1
        const uint32_t* x32 = static_cast<const uint32_t*>(p.x);              // 4 words per group
        uint32_t bfrag[4];
        #pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int g = ks * 32 + build_group_of_col(j, r);
            bfrag[j] = (g < p.G) ? ldcg_u32(x32 + g * 4 + q) : 0u;   // L2-coherent: x is the previous kernel's output
        }
        const bool odd = r & 1;
        const bool round16 = (p.flags & LUTVQ_FLAG_LUT_FP16) != 0;
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
                if (s + PF * NW < n_sets) load_set(buf[st], s + PF * NW);
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
                if (row < row_end) {
                    if (p.n_k == 1) epilogue_store<DT>(p.y, p.scales, p.bias, row, keep);
                    else            p.partial[(size_t)ks * p.O + row] = keep;
                }
            }
        }
    }
    if (p.n_k == 1) return;

    // ---------------- deterministic split-k reduction by the last-arriving CTA of this row chunk ----------------
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned prev = atomicAdd(&p.counters[chunk], 1u);
        s_is_last = (prev == (unsigned)(p.n_k - 1));
    }
    __syncthreads();
    if (s_is_last) {
        __threadfence();
        for (int row = row0 + threadIdx.x; row < row_end; row += NW * 32) {
            float acc = 0.f;
            for (int k = 0; k < p.n_k; ++k) acc += ldcg_f32(p.partial + (size_t)k * p.O + row);
            epilogue_store<DT>(p.y, p.scales, p.bias, row, acc);
        }
        if (threadIdx.x == 0) p.counters[chunk] = 0u;          // re-arm for the next call on this stream
    }
}

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------
struct GemvTuning { int nw = 8, pf = 2, ctas_per_sm = 1; };
static GemvTuning g_tuning;

struct GemvPlan { int n_k, n_chunks, rows_per_chunk; };

static GemvPlan plan_gemv(int O, int G, int sm_count, int ctas_per_sm) {
    GemvPlan pl;
    pl.n_k = ceil_div(G * 2, kSliceTables);
    int n_r = (sm_count * ctas_per_sm) / pl.n_k;
    if (n_r < 1) n_r = 1;
    int rows = ceil_div(ceil_div(O, n_r), kSetRows) * kSetRows;
    pl.rows_per_chunk = rows;
    pl.n_chunks = ceil_div(O, rows);
    return pl;
}

size_t gemv_workspace_bytes(int O, int G) {
    const int n_k = ceil_div(G * 2, kSliceTables);
    // partial sums + one ticket per possible row chunk (chunks are >= 32 rows)
    return (size_t)n_k * O * sizeof(float) + (size_t)(ceil_div(O, kSetRows) + 1) * sizeof(unsigned int) + 256;
}

template <int DT, int NW, int PF, int MINB>
static int launch_variant(const GemvParams& p, dim3 grid, cudaStream_t stream, bool pdl) {
    auto kern = lut_gemv_2x8_kernel<DT, NW, PF, MINB>;
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

template <int DT>
static int launch_dtype(const GemvParams& p, dim3 grid, cudaStream_t stream, bool pdl) {
    const GemvTuning t = g_tuning;
    if (t.nw == 8 && t.pf == 2)  return launch_variant<DT, 8, 2, 2>(p, grid, stream, pdl);
    if (t.nw == 8 && t.pf == 3)  return launch_variant<DT, 8, 3, 2>(p, grid, stream, pdl);
    if (t.nw == 8 && t.pf == 1)  return launch_variant<DT, 8, 1, 2>(p, grid, stream, pdl);
    if (t.nw == 16 && t.pf == 1) return launch_variant<DT, 16, 1, 1>(p, grid, stream, pdl);
    if (t.nw == 16 && t.pf == 2) return launch_variant<DT, 16, 2, 1>(p, grid, stream, pdl);
    if (t.nw == 4 && t.pf == 2)  return launch_variant<DT, 4, 2, 3>(p, grid, stream, pdl);
    if (t.nw == 4 && t.pf == 4)  return launch_variant<DT, 4, 4, 3>(p, grid, stream, pdl);
    return fail(LUTVQ_EINVAL, "no kernel variant for tuning nw=%d pf=%d", t.nw, t.pf);
}

// One GEMV: y = epilogue(W(codes) . x).  Validation is done by the caller (api.cu).
int gemv_2x8_launch(const void* codes, const void* x, void* y, const void* codebook, const void* scales,
                    const void* bias, int O, int I, int dtype, unsigned flags, void* ws, size_t ws_bytes,
                    cudaStream_t stream) {
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    const int G = I / 8;
    const GemvPlan pl = plan_gemv(O, G, di.sm_count, g_tuning.ctas_per_sm);
    const size_t need = gemv_workspace_bytes(O, G);
    if (pl.n_k > 1) {
        LUTVQ_REQUIRE(ws != nullptr && ws_bytes >= need, "workspace too small: need %zu bytes, got %zu", need, ws_bytes);
        LUTVQ_REQUIRE(aligned16(ws), "workspace must be 16-byte aligned");
    }
    GemvParams p;
    p.codes = static_cast<const uint8_t*>(codes);
    p.x = x; p.codebook = codebook; p.y = y; p.scales = scales; p.bias = bias;
    p.partial = static_cast<float*>(ws);
    p.counters = reinterpret_cast<unsigned int*>(static_cast<uint8_t*>(ws) + (size_t)pl.n_k * O * sizeof(float));
    p.O = O; p.G = G; p.n_k = pl.n_k; p.rows_per_chunk = pl.rows_per_chunk; p.flags = flags;
    const dim3 grid(pl.n_chunks, pl.n_k);
    const bool pdl = !(flags & LUTVQ_FLAG_NO_PDL);
    if (dtype == LUTVQ_F16)  return launch_dtype<LUTVQ_F16>(p, grid, stream, pdl);
    if (dtype == LUTVQ_BF16) return launch_dtype<LUTVQ_BF16>(p, grid, stream, pdl);
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
    GemvPlan pl = plan_gemv(O, I / 8, sm, g_tuning.ctas_per_sm);
    if (n_k) *n_k = pl.n_k;
    if (n_chunks) *n_chunks = pl.n_chunks;
    if (rows_per_chunk) *rows_per_chunk = pl.rows_per_chunk;
}

}  // namespace lutvq
