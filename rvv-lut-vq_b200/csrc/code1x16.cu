// code1x16.cu -- AQLM 1x16 format (one codebook of 65536 vectors of 8 or 16 values, 16-bit codes): SURVEY 8a row a12.
//
// Replaces Code1x16MatVec / Code1x16Dequant + launchers (reference inference_kernels/cuda_kernel.cu:7-142, 470-553) and
// the host wrappers code1x16_matvec / _matmat / _dequant (cuda_kernel.cpp:113-209).  This format is NOT LUT-able (a LUT
// would need G x 65536 entries per activation vector): it stays a dequantize-and-FMA GEMV whose bound is
// the L2 gather of one 16- or 32-byte codebook row per 2-byte code (the 1-2 MiB codebook lives in the 126 MB L2), i.e.
// 8-16x the HBM code stream in L2->SM traffic.
//
//   GEMV  y[o] = sum_g  C[code(o,g)][0..GS) . x[g*GS .. (g+1)*GS)            fp32 accumulate (the reference: fp16 HFMA2 per group)
//         one warp per output row, a lane owns 8 consecutive codes per trip (one 128-bit code load) and issues their 8
//         (16) codebook-row loads back to back before the first FMA; x sits in shared memory, padded by 16 B per 128 B so
//         that the 128-bit reads of a quarter warp hit 32 distinct banks; up to 4 batch rows share every fetched codebook
//         row (code1x16_matmat loops one launch per row in the reference, cuda_kernel.cpp:160-170).
//         Epilogue in the reference's rounding order: round(acc) -> * scale -> + bias (cuda_kernel.cpp:95-111).
//   Dequant  W[o][g*GS + j] = C[code(o,g)][j] (* scale[o]: the reference's separate `weight *= scales`, one RN multiply): bit-exact.
#include "common.cuh"

namespace lutvq {

constexpr int k1x16Warps = 16;               // 16 rows per CTA trip: half as many copies of x into shared memory as with 8
constexpr int k1x16MaxBatch = 4;

// codebook rows: 16 random bytes out of 1-2 MiB per code.  They are served by L2; letting them allocate in L1 (ld.global.nc)
// streams whole sectors through a cache that cannot hold the table -- measured 15 / 28 / 39 us per GEMV on the three 7B shapes where
// the reference's kernel, which bypasses L1 (cuda_kernel.cu:48-50), takes 12.5 / 26.6 / 27.6.
__device__ __forceinline__ uint4 ldg_nc_u128(const void* p) {
    uint4 r;
    asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
template <int DT> __device__ __forceinline__ float2 cvt2(uint32_t w);
template <> __device__ __forceinline__ float2 cvt2<LUTVQ_F16>(uint32_t w)  { return __half22float2(*reinterpret_cast<const __half2*>(&w)); }
template <> __device__ __forceinline__ float2 cvt2<LUTVQ_BF16>(uint32_t w) { return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xFFFF0000u)); }

template <int DT> __device__ __forceinline__ float dot8(const uint4& c, const uint4& x, float acc) {
    float2 a, b;
    a = cvt2<DT>(c.x); b = cvt2<DT>(x.x); acc = fmaf(a.x, b.x, acc); acc = fmaf(a.y, b.y, acc);
    a = cvt2<DT>(c.y); b = cvt2<DT>(x.y); acc = fmaf(a.x, b.x, acc); acc = fmaf(a.y, b.y, acc);
    a = cvt2<DT>(c.z); b = cvt2<DT>(x.z); acc = fmaf(a.x, b.x, acc); acc = fmaf(a.y, b.y, acc);
    a = cvt2<DT>(c.w); b = cvt2<DT>(x.w); acc = fmaf(a.x, b.x, acc); acc = fmaf(a.y, b.y, acc);
    return acc;
}

// shared-memory offset (bytes) of the 16-byte piece `v` of x: one 16-byte pad per lane's worth of pieces (8 groups of V
// pieces), so the lane stride is 8V + 1 pieces = an odd multiple of 4 banks and a quarter warp covers all 32 banks
template <int V> __device__ __forceinline__ uint32_t xpad(uint32_t v) { return (v + (v / (8u * V))) * 16u; }

static int g_1x16_mode = 0;      // experiments (fp16, group 8, one row): 1 = the reference's HFMA2 arithmetic, 2 = gather only (no math)
void code1x16_set_mode(int m) { g_1x16_mode = m; }
static int g_1x16_split = 0, g_1x16_per_sm = 0;      // experiments: override the launch plan (0 = automatic)
void code1x16_set_plan(int split, int per_sm) { g_1x16_split = split; g_1x16_per_sm = per_sm; }

// `split` (1, 2 or 4) warps share one output row (contiguous K ranges, partial sums through shared memory): layers with few rows per
// SM would otherwise leave the SM with too few gathers in flight (the kernel is bound by the latency of the codebook gather).
template <int DT, int GS, int NB, int MODE = 0>
__global__ void __launch_bounds__(k1x16Warps * 32, (NB == 1 && GS == 8) ? 3 : 1)
code1x16_gemv_kernel(const uint16_t* __restrict__ codes, const void* __restrict__ x, void* __restrict__ y,
                     const uint4* __restrict__ codebook, const void* __restrict__ scales, const void* __restrict__ bias,
                     int O, int G, int I, int nb, int raw, int split) {
    using D = Dtype<DT>;
    using T = typename D::T;
    extern __shared__ __align__(16) uint8_t sx[];
    constexpr int V = GS / 8;                                      // 16-byte pieces per group
    const int n_v = I / 8;                                         // pieces per batch row
    const uint32_t row_stride = xpad<V>((uint32_t)n_v) + 16u;
    pdl_launch_dependents();
    pdl_wait();                                                    // x may be the previous kernel's output
    for (int b = 0; b < NB; ++b) {
        const uint4* xb = reinterpret_cast<const uint4*>(static_cast<const T*>(x) + (size_t)(b < nb ? b : nb - 1) * I);
        for (int v = threadIdx.x; v < n_v; v += blockDim.x) *reinterpret_cast<uint4*>(sx + b * row_stride + xpad<V>((uint32_t)v)) = xb[v];
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // balanced contiguous row ranges: CTA c owns rows [c * O / grid, (c + 1) * O / grid), its warps take them in turn
    const int row0 = (int)((long long)blockIdx.x * O / gridDim.x), row1 = (int)((long long)(blockIdx.x + 1) * O / gridDim.x);
    __shared__ float red[k1x16Warps][NB];
    const int teams = k1x16Warps / split, team = warp / split, part = warp - team * split;
    const int chunks = G / 8, cpp = ((chunks + split - 1) / split + 31) / 32 * 32;       // 8-code chunks per part: whole warp iterations
    const int g_begin = part * cpp * 8, g_end = min(G, (part + 1) * cpp * 8);
    const int n_trips = (row1 - row0 + teams - 1) / teams;                             // uniform over the CTA (barriers below)
    for (int trip = 0; trip < n_trips; ++trip) {
        const int row = row0 + trip * teams + team;
        const bool valid = row < row1;
        const uint16_t* crow = codes + (size_t)(valid ? row : row0) * G;
        float acc[NB];
        #pragma unroll
        for (int b = 0; b < NB; ++b) acc[b] = 0.f;
        for (int g0 = g_begin + lane * 8; valid && g0 < g_end; g0 += 256) {
            // 8 codes = one 128-bit load when the row allows it (G % 8 == 0 is required by the host)
            const uint4 cw = ldg_stream_u128(crow + g0);
            const uint32_t w[4] = {cw.x, cw.y, cw.z, cw.w};
            uint4 cb[8][V];
            #pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint32_t code = (w[i >> 1] >> (16 * (i & 1))) & 0xFFFFu;
                #pragma unroll
                for (int v = 0; v < V; ++v) cb[i][v] = ldg_nc_u128(codebook + (size_t)code * V + v);
            }
            #pragma unroll
            for (int i = 0; i < 8; ++i)
                #pragma unroll
                for (int v = 0; v < V; ++v) {
                    const uint32_t off = xpad<V>((uint32_t)((g0 + i) * V + v));
                    if (MODE == 2) { acc[0] += __uint_as_float((cb[i][v].x ^ cb[i][v].y ^ cb[i][v].z ^ cb[i][v].w) & 0x3f800000u); continue; }
                    if (MODE == 1) {
                        const uint4 xv = *reinterpret_cast<const uint4*>(sx + off);
                        const __half2* a = reinterpret_cast<const __half2*>(&cb[i][v]);
                        const __half2* bb = reinterpret_cast<const __half2*>(&xv);
                        __half2 r2 = __hmul2(a[0], bb[0]);
                        r2 = __hfma2(a[1], bb[1], r2); r2 = __hfma2(a[2], bb[2], r2); r2 = __hfma2(a[3], bb[3], r2);
                        acc[0] += __half2float(r2.x) + __half2float(r2.y);
                        continue;
                    }
                    #pragma unroll
                    for (int b = 0; b < NB; ++b) acc[b] = dot8<DT>(cb[i][v], *reinterpret_cast<const uint4*>(sx + b * row_stride + off), acc[b]);
                }
        }
        #pragma unroll
        for (int b = 0; b < NB; ++b) {
            #pragma unroll
            for (int s_ = 16; s_ > 0; s_ >>= 1) acc[b] += __shfl_xor_sync(0xffffffffu, acc[b], s_);
        }
        if (split > 1) {                                                                   // uniform over the grid
            if (lane == 0) {
                #pragma unroll
                for (int b = 0; b < NB; ++b) red[warp][b] = acc[b];
            }
            __syncthreads();
            if (part == 0 && lane == 0) {
                #pragma unroll
                for (int b = 0; b < NB; ++b)
                    for (int q = 1; q < split; ++q) acc[b] += red[warp + q][b];               // fixed order: deterministic
            }
            __syncthreads();
        }
        if (valid && part == 0 && lane == 0) {
            #pragma unroll
            for (int b = 0; b < NB; ++b) {
                if (b >= nb) break;
                T r = D::from_f(acc[b]);                                                   // Code1x16MatVec writes round(acc) (cuda_kernel.cu:86-92)
                if (!raw) {
                    if (scales) r = D::mul(r, static_cast<const T*>(scales)[row]);         // flat_output *= scales (cuda_kernel.cpp:101)
                    if (bias)   r = D::add(r, static_cast<const T*>(bias)[row]);           // flat_output += bias   (cuda_kernel.cpp:103)
                }
                static_cast<T*>(y)[(size_t)b * O + row] = r;
            }
        }
    }
}

// one thread per code: 2-byte code -> GS values (x scale), 16-byte stores; consecutive threads = consecutive groups
template <int DT, int GS>
__global__ void code1x16_dequant_kernel(const uint16_t* __restrict__ codes, uint4* __restrict__ W, const uint4* __restrict__ codebook,
                                        const void* __restrict__ scales, int O, int G) {
    using D = Dtype<DT>;
    using T = typename D::T;
    using T2 = typename D::T2;
    constexpr int V = GS / 8;
    const size_t total = (size_t)O * G;
    for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const uint32_t code = codes[t];
        const int row = (int)(t / G);
        #pragma unroll
        for (int v = 0; v < V; ++v) {
            uint4 c = ldg_nc_u128(codebook + (size_t)code * V + v);
            if (scales) {
                const T2 s2 = D::dup(static_cast<const T*>(scales)[row]);
                T2* h = reinterpret_cast<T2*>(&c);
                #pragma unroll
                for (int j = 0; j < 4; ++j) h[j] = D::mul2(h[j], s2);
            }
            __stcs(W + t * V + v, c);
        }
    }
}

template <int DT, int GS, int NB>
static int launch_gemv(const void* codes, const void* x, void* y, const void* codebook, const void* scales, const void* bias,
                       int O, int I, int nb, int raw, cudaStream_t stream, const DeviceInfo& di) {
    auto kern = code1x16_gemv_kernel<DT, GS, NB>;
    if (DT == LUTVQ_F16 && GS == 8 && NB == 1 && g_1x16_mode == 1) kern = code1x16_gemv_kernel<DT, GS, NB, (DT == LUTVQ_F16 && GS == 8 && NB == 1) ? 1 : 0>;
    if (DT == LUTVQ_F16 && GS == 8 && NB == 1 && g_1x16_mode == 2) kern = code1x16_gemv_kernel<DT, GS, NB, (DT == LUTVQ_F16 && GS == 8 && NB == 1) ? 2 : 0>;
    const int n_v = I / 8;
    const size_t smem = (size_t)NB * ((size_t)(n_v + (n_v >> 3)) * 16 + 16);
    if (smem > 48 * 1024) {
        LUTVQ_REQUIRE(smem <= di.smem_optin, "prob_k=%d needs %zu bytes of shared memory for x (limit %zu)", I, smem, di.smem_optin);
        LUTVQ_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    // Launch plan, measured on B200 (tools/sweep_1x16.py, us per GEMV at 4096->4096 / 4096->11008 / 11008->4096, fp16, group 8):
    // exactly as many CTAs per SM as are resident (3 x 16 warps at <= 42 registers) = one full wave with balanced row ranges:
    // 11.2 / 24.3 / 25.8; 2 per SM: 14.6 / 35.1 / 35.1; 6 per SM: 14.1 / 25.5 / 31.6; rows split over 2 or 4 warps: never better.
    // (The reference's Code1x16MatVec on the same GPU: 12.5 / 26.5 / 27.6.)
    int split = 1;
    if (g_1x16_split == 1 || g_1x16_split == 2 || g_1x16_split == 4) split = g_1x16_split;
    while (split > 1 && (I / GS) / split < 256) split >>= 1;
    int per_sm = NB == 1 ? 3 : 2;
    if (g_1x16_per_sm > 0) per_sm = g_1x16_per_sm;
    int ctas = di.sm_count * per_sm;
    if (ctas > O) ctas = O;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(ctas); cfg.blockDim = dim3(k1x16Warps * 32); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    LUTVQ_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, static_cast<const uint16_t*>(codes), x, y, static_cast<const uint4*>(codebook), scales, bias,
                                     O, I / GS, I, nb, raw, split));
    return LUTVQ_OK;
}

template <int DT, int GS>
static int gemv_batch(const void* codes, const void* x, void* y, const void* codebook, const void* scales, const void* bias,
                      int batch, int O, int I, int raw, cudaStream_t stream, const DeviceInfo& di) {
    const size_t es = 2;
    for (int b0 = 0; b0 < batch; b0 += k1x16MaxBatch) {
        const int nb = batch - b0 < k1x16MaxBatch ? batch - b0 : k1x16MaxBatch;
        const void* xb = static_cast<const uint8_t*>(x) + (size_t)b0 * I * es;
        void* yb = static_cast<uint8_t*>(y) + (size_t)b0 * O * es;
        int rc;
        if (nb == 1) rc = launch_gemv<DT, GS, 1>(codes, xb, yb, codebook, scales, bias, O, I, nb, raw, stream, di);
        else if (nb == 2) rc = launch_gemv<DT, GS, 2>(codes, xb, yb, codebook, scales, bias, O, I, nb, raw, stream, di);
        else rc = launch_gemv<DT, GS, 4>(codes, xb, yb, codebook, scales, bias, O, I, nb, raw, stream, di);
        if (rc) return rc;
    }
    return LUTVQ_OK;
}

static int check_1x16(const void* codes, const void* codebook, int O, int I, int gs, int dtype) {
    // the reference throws c10::NotImplementedError for other group sizes (cuda_kernel.cpp:136-145) and dtypes (:9-25)
    if (gs != 8 && gs != 16) return fail(LUTVQ_EUNSUPPORTED, "AQLM CUDA kernels only support codebooks with 8 or 16 features. Got %d.", gs);
    if (dtype != LUTVQ_F16 && dtype != LUTVQ_BF16) return fail(LUTVQ_EUNSUPPORTED, "dtype %d not supported (fp16=0, bf16=1)", dtype);
    LUTVQ_REQUIRE(codes && codebook, "null pointer");
    LUTVQ_REQUIRE(O > 0 && I > 0 && I % (8 * gs) == 0, "prob_k=%d must be a positive multiple of %d (8 codes of %d values per 128-bit load; the reference reads int4 too, cuda_kernel.cu:41)", I, 8 * gs, gs);
    LUTVQ_REQUIRE(aligned16(codes) && aligned16(codebook), "codes and codebook must be 16-byte aligned");
    return LUTVQ_OK;
}

}  // namespace lutvq

using namespace lutvq;

extern "C" {

void lutvq_debug_set_1x16_mode(int mode) { code1x16_set_mode(mode); }
void lutvq_debug_set_1x16_plan(int split, int per_sm) { code1x16_set_plan(split, per_sm); }

int lutvq_code1x16_matmat(const void* x, const void* codes, const void* codebooks, const void* scales, const void* bias, void* y,
                          int batch, int prob_m, int prob_k, int group_size, int dtype, void* stream) {
    if (int rc = check_1x16(codes, codebooks, prob_m, prob_k, group_size, dtype)) return rc;
    LUTVQ_REQUIRE(x && y && batch >= 0, "null pointer or negative batch");
    LUTVQ_REQUIRE(aligned16(x), "x must be 16-byte aligned");
    if (batch == 0) return LUTVQ_OK;
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == LUTVQ_F16)
        return group_size == 8 ? gemv_batch<LUTVQ_F16, 8>(codes, x, y, codebooks, scales, bias, batch, prob_m, prob_k, 0, st, di)
                               : gemv_batch<LUTVQ_F16, 16>(codes, x, y, codebooks, scales, bias, batch, prob_m, prob_k, 0, st, di);
    return group_size == 8 ? gemv_batch<LUTVQ_BF16, 8>(codes, x, y, codebooks, scales, bias, batch, prob_m, prob_k, 0, st, di)
                           : gemv_batch<LUTVQ_BF16, 16>(codes, x, y, codebooks, scales, bias, batch, prob_m, prob_k, 0, st, di);
}

int lutvq_code1x16_matvec(const void* codes, const void* x, void* y, const void* codebook, int prob_m, int prob_k, int group_size,
                          int dtype, void* stream) {
    if (int rc = check_1x16(codes, codebook, prob_m, prob_k, group_size, dtype)) return rc;
    LUTVQ_REQUIRE(x && y && aligned16(x), "x / y null or x not 16-byte aligned");
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == LUTVQ_F16)
        return group_size == 8 ? gemv_batch<LUTVQ_F16, 8>(codes, x, y, codebook, nullptr, nullptr, 1, prob_m, prob_k, 1, st, di)
                               : gemv_batch<LUTVQ_F16, 16>(codes, x, y, codebook, nullptr, nullptr, 1, prob_m, prob_k, 1, st, di);
    return group_size == 8 ? gemv_batch<LUTVQ_BF16, 8>(codes, x, y, codebook, nullptr, nullptr, 1, prob_m, prob_k, 1, st, di)
                           : gemv_batch<LUTVQ_BF16, 16>(codes, x, y, codebook, nullptr, nullptr, 1, prob_m, prob_k, 1, st, di);
}

int lutvq_code1x16_dequant(const void* codes, void* W, const void* codebook, const void* scales, int prob_m, int prob_k, int group_size,
                           int dtype, void* stream) {
    if (int rc = check_1x16(codes, codebook, prob_m, prob_k, group_size, dtype)) return rc;
    LUTVQ_REQUIRE(W && aligned16(W), "W null or not 16-byte aligned");
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int G = prob_k / group_size;
    const size_t total = (size_t)prob_m * G;
    size_t blocks = (total + 255) / 256;
    if (blocks > (size_t)di.sm_count * 16) blocks = (size_t)di.sm_count * 16;
    const uint16_t* c = static_cast<const uint16_t*>(codes);
    uint4* w = static_cast<uint4*>(W);
    const uint4* cb = static_cast<const uint4*>(codebook);
    if (dtype == LUTVQ_F16) {
        if (group_size == 8) code1x16_dequant_kernel<LUTVQ_F16, 8><<<(unsigned)blocks, 256, 0, st>>>(c, w, cb, scales, prob_m, G);
        else code1x16_dequant_kernel<LUTVQ_F16, 16><<<(unsigned)blocks, 256, 0, st>>>(c, w, cb, scales, prob_m, G);
    } else {
        if (group_size == 8) code1x16_dequant_kernel<LUTVQ_BF16, 8><<<(unsigned)blocks, 256, 0, st>>>(c, w, cb, scales, prob_m, G);
        else code1x16_dequant_kernel<LUTVQ_BF16, 16><<<(unsigned)blocks, 256, 0, st>>>(c, w, cb, scales, prob_m, G);
    }
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

}  // extern "C"
