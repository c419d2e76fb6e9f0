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

constexpr int k1x16Warps = 8;
constexpr int k1x16MaxBatch = 4;

__device__ __forceinline__ uint4 ldg_nc_u128(const void* p) {      // read-only path; L1 may keep hot codebook rows
    uint4 r;
    asm volatile("ld.global.nc.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
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

template <int DT, int GS, int NB>
__global__ void __launch_bounds__(k1x16Warps * 32)
code1x16_gemv_kernel(const uint16_t* __restrict__ codes, const void* __restrict__ x, void* __restrict__ y,
                     const uint4* __restrict__ codebook, const void* __restrict__ scales, const void* __restrict__ bias,
                     int O, int G, int I, int nb, int raw) {
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
    const int n_warps = gridDim.x * k1x16Warps;
    for (int row = blockIdx.x * k1x16Warps + warp; row < O; row += n_warps) {
        const uint16_t* crow = codes + (size_t)row * G;
        float acc[NB];
        #pragma unroll
        for (int b = 0; b < NB; ++b) acc[b] = 0.f;
        for (int g0 = lane * 8; g0 < G; g0 += 256) {
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
                    #pragma unroll
                    for (int b = 0; b < NB; ++b) acc[b] = dot8<DT>(cb[i][v], *reinterpret_cast<const uint4*>(sx + b * row_stride + off), acc[b]);
                }
        }
        #pragma unroll
        for (int b = 0; b < NB; ++b) {
            float a = acc[b];
            #pragma unroll
            for (int s = 16; s > 0; s >>= 1) a += __shfl_xor_sync(0xffffffffu, a, s);
            if (lane == 0 && b < nb) {
                T r = D::from_f(a);                                                        // Code1x16MatVec writes round(acc) (cuda_kernel.cu:86-92)
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
    const int n_v = I / 8;
    const size_t smem = (size_t)NB * ((size_t)(n_v + (n_v >> 3)) * 16 + 16);
    if (smem > 48 * 1024) {
        LUTVQ_REQUIRE(smem <= di.smem_optin, "prob_k=%d needs %zu bytes of shared memory for x (limit %zu)", I, smem, di.smem_optin);
        LUTVQ_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    int ctas = ceil_div(O, k1x16Warps);
    const int cap = di.sm_count * (NB == 1 ? 6 : 3);
    if (ctas > cap) ctas = cap;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(ctas); cfg.blockDim = dim3(k1x16Warps * 32); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    LUTVQ_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, static_cast<const uint16_t*>(codes), x, y, static_cast<const uint4*>(codebook), scales, bias,
                                     O, I / GS, I, nb, raw));
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
