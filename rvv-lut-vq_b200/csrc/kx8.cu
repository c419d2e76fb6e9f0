// kx8.cu -- the Kx8 generalisation of the AQLM path (num_codebooks N in 1..8, 256 entries, groups of 8; fp16 / bf16):
// reference CodeKx8MatVec / CodeKx8Dequant + codeKx8_matmat / code1x8_* host wrappers
// (benchmark/AQLM/inference_lib/src/aqlm/inference_kernels/cuda_kernel.cu:296-468, cuda_kernel.cpp:521-684).
//
// GEMV: N in {1, 2, 4, 8} with whole 16-byte code rows runs on the FUSED LUT-GEMV of gemv_lut.cu (the kernel is templated on
// the codebooks per group: a k-slice is 64 code bytes whatever N is, only the tensor-core LUT build knows the format), fp16
// and bf16, scale / bias in the kernel's epilogue in the reference's rounding order.  Other N (3, 5, 6, 7 -- the reference's
// host code accepts them through its generic template) and ragged rows keep the unfused fp32 K1/K2 pair of vq_golden.cu
// (fp16): LUT build into the workspace, one-warp-per-row gather, epilogue kernel.
// Dequantize: sequential adds in codebook order in the storage type, exactly the `c = __hadd2(c, a)` loop of CodeKx8Dequant
// (cuda_kernel.cu:437-453); for N = 2 identical to code2x8_dequant.
#include "common.cuh"

namespace lutvq {

int lut_build_launch(const void* x, const void* codebook, void* lut, int M, int N, int K, int d, int cb_shared,
                     int lut_mode, cudaStream_t stream);
int lut_gemv_launch(const void* codeword, const void* lut, void* out, int d_out, int M, int N, int K,
                    int lut_is_fp32, int acc_mode, cudaStream_t stream);
size_t gemv_workspace_bytes(int O, int G);
int gemv_2x8_launch_rows(const void* codes, int packed, const void* x, void* y, const void* codebook, const void* scales, const void* bias,
                         int rows, int O, int I, int dtype, unsigned flags, void* ws, size_t ws_bytes, cudaStream_t stream, int ncb);
int gemv_max_rows_per_launch();
size_t gemv_rows_workspace_bytes(int rows, int O);
int gemv_2x8_launch_ref(const void* codes, const void* x, void* y, const void* codebook, const void* scales,
                        const void* bias, int O, int I, int dtype, unsigned flags, void* ws, size_t ws_bytes, cudaStream_t stream);
int gemv_kx8_launch_ref(const void* codes, const void* x, void* y, const void* codebook, const void* scales,
                        const void* bias, int O, int I, int ncb, int dtype, unsigned flags, void* ws, size_t ws_bytes,
                        cudaStream_t stream);

template <typename H, typename H2>
__global__ void __launch_bounds__(256) kx8_dequant_kernel(const uint8_t* __restrict__ codes, H* __restrict__ W,
                                                          const uint4* __restrict__ codebook, const H* __restrict__ scales,
                                                          int O, int G, int N) {
    const size_t total = (size_t)O * G;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t o = i / G;
        const uint8_t* c = codes + i * N;
        uint4 acc = codebook[c[0]];
        H2* a2 = reinterpret_cast<H2*>(&acc);
        for (int n = 1; n < N; ++n) {
            const uint4 v = codebook[(size_t)n * 256 + c[n]];
            const H2* v2 = reinterpret_cast<const H2*>(&v);
            #pragma unroll
            for (int j = 0; j < 4; ++j) a2[j] = __hadd2(a2[j], v2[j]);
        }
        if (scales) {
            const H2 s = H2{scales[o], scales[o]};
            #pragma unroll
            for (int j = 0; j < 4; ++j) a2[j] = __hmul2(a2[j], s);
        }
        reinterpret_cast<uint4*>(W)[i] = acc;
    }
}

__global__ void kx8_epilogue_kernel(const float* __restrict__ acc, const __half* __restrict__ scales, const __half* __restrict__ bias,
                                    __half* __restrict__ y, int O) {
    const int o = blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= O) return;
    __half r = __float2half_rn(acc[o]);
    if (scales) r = __hmul(r, scales[o]);
    if (bias) r = __hadd(r, bias[o]);
    y[o] = r;
}

static size_t kx8_lut_bytes(int I, int N) { return ((size_t)(I / 8) * N * 256 * sizeof(float) + 255) / 256 * 256; }
// the fused kernel takes N in {1, 2, 4, 8} when a code row is a whole number of 16-byte chunks
static bool kx8_fused(int prob_k, int N) { return (N == 1 || N == 2 || N == 4 || N == 8) && ((prob_k / 8) * N) % 16 == 0; }

}  // namespace lutvq

using namespace lutvq;

extern "C" {

size_t lutvq_codeKx8_workspace_bytes(int prob_m, int prob_k, int num_codebooks) {
    if (prob_m <= 0 || prob_k <= 0 || num_codebooks <= 0) return 0;
    if (kx8_fused(prob_k, num_codebooks)) return gemv_workspace_bytes(prob_m, prob_k / 8);
    return kx8_lut_bytes(prob_k, num_codebooks) + (size_t)prob_m * sizeof(float);
}

int lutvq_codeKx8_matmat(const void* x, const void* codes, const void* codebooks, const void* scales, const void* bias, void* y,
                         int batch, int prob_m, int prob_k, int num_codebooks, int dtype, void* ws, size_t ws_bytes, void* stream) {
    LUTVQ_REQUIRE(codes && codebooks, "null codes/codebooks pointer");
    LUTVQ_REQUIRE(prob_m > 0 && prob_k > 0 && prob_k % 8 == 0, "bad shape prob_m=%d prob_k=%d", prob_m, prob_k);
    LUTVQ_REQUIRE(num_codebooks >= 1 && num_codebooks <= 8, "num_codebooks=%d must be in 1..8", num_codebooks);
    LUTVQ_REQUIRE(batch >= 0, "bad batch %d", batch);
    if (dtype != LUTVQ_F16 && dtype != LUTVQ_BF16)
        return fail(LUTVQ_EUNSUPPORTED, "AQLM kernels only support float16 (0) and bfloat16 (1); got dtype %d", dtype);
    const bool fused = kx8_fused(prob_k, num_codebooks) && aligned16(codes) && aligned16(codebooks);
    if (!fused && dtype != LUTVQ_F16)
        return fail(LUTVQ_EUNSUPPORTED, "bf16 Kx8 needs num_codebooks in {1, 2, 4, 8} and 16-byte code rows (got %d codebooks, prob_k=%d)", num_codebooks, prob_k);
    if (batch == 0) return LUTVQ_OK;
    LUTVQ_REQUIRE(x && y, "null x/y pointer");
    const size_t need = lutvq_codeKx8_workspace_bytes(prob_m, prob_k, num_codebooks);
    LUTVQ_REQUIRE(ws != nullptr && ws_bytes >= need && aligned16(ws), "workspace too small or misaligned: need %zu bytes, got %zu", need, ws_bytes);
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (fused) {
        // up to 8 rows per launch when the workspace holds their split-k words (as lutvq_code2x8_matmat does, lutvq.h) ...
        const int per = batch < gemv_max_rows_per_launch() ? batch : gemv_max_rows_per_launch();
        if (batch > 1 && ws_bytes >= gemv_rows_workspace_bytes(per, prob_m)) {
            for (int b0 = 0; b0 < batch; b0 += per) {
                const int nb = batch - b0 < per ? batch - b0 : per;
                if (int rc = gemv_2x8_launch_rows(codes, 0, static_cast<const uint8_t*>(x) + (size_t)b0 * prob_k * 2,
                                                  static_cast<uint8_t*>(y) + (size_t)b0 * prob_m * 2, codebooks, scales, bias, nb, prob_m, prob_k,
                                                  dtype, 0u, ws, ws_bytes, st, num_codebooks))
                    return rc;
            }
            return LUTVQ_OK;
        }
        // ... else one fused launch per batch row, like cuda_kernel.cpp:540-552 loops CodeKx8MatVec; scale / bias ride in the epilogue
        for (int b = 0; b < batch; ++b) {
            const void* xb = static_cast<const uint8_t*>(x) + (size_t)b * prob_k * 2;
            void* yb = static_cast<uint8_t*>(y) + (size_t)b * prob_m * 2;
            const int rc = num_codebooks == 2
                ? gemv_2x8_launch_ref(codes, xb, yb, codebooks, scales, bias, prob_m, prob_k, dtype, 0u, ws, ws_bytes, st)
                : gemv_kx8_launch_ref(codes, xb, yb, codebooks, scales, bias, prob_m, prob_k, num_codebooks, dtype, 0u, ws, ws_bytes, st);
            if (rc) return rc;
        }
        return LUTVQ_OK;
    }
    float* lut = static_cast<float*>(ws);
    float* acc = reinterpret_cast<float*>(static_cast<uint8_t*>(ws) + kx8_lut_bytes(prob_k, num_codebooks));
    const int G = prob_k / 8;
    for (int b = 0; b < batch; ++b) {
        const __half* xb = static_cast<const __half*>(x) + (size_t)b * prob_k;
        if (int rc = lut_build_launch(xb, codebooks, lut, G, num_codebooks, 256, 8, /*cb_shared=*/1, LUTVQ_LUT_FP32, st)) return rc;
        if (int rc = lut_gemv_launch(codes, lut, acc, prob_m, G, num_codebooks, 256, /*lut_is_fp32=*/1, LUTVQ_ACC_F32, st)) return rc;
        kx8_epilogue_kernel<<<(prob_m + 255) / 256, 256, 0, st>>>(acc, static_cast<const __half*>(scales), static_cast<const __half*>(bias),
                                                                  static_cast<__half*>(y) + (size_t)b * prob_m, prob_m);
        LUTVQ_CUDA_OK(cudaGetLastError());
    }
    return LUTVQ_OK;
}

int lutvq_codeKx8_dequant(const void* codes, void* W, const void* codebooks, const void* scales, int prob_m, int prob_k,
                          int num_codebooks, int dtype, void* stream) {
    LUTVQ_REQUIRE(codes && W && codebooks, "null pointer");
    LUTVQ_REQUIRE(prob_m > 0 && prob_k > 0 && prob_k % 8 == 0, "bad shape prob_m=%d prob_k=%d", prob_m, prob_k);
    LUTVQ_REQUIRE(num_codebooks >= 1 && num_codebooks <= 8, "num_codebooks=%d must be in 1..8", num_codebooks);
    LUTVQ_REQUIRE(aligned16(W) && aligned16(codebooks), "W and codebooks must be 16-byte aligned");
    if (dtype != LUTVQ_F16 && dtype != LUTVQ_BF16)
        return fail(LUTVQ_EUNSUPPORTED, "AQLM kernels only support float16 (0) and bfloat16 (1); got dtype %d", dtype);
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    const size_t total = (size_t)prob_m * (prob_k / 8);
    int grid = (int)((total + 255) / 256);
    if (grid > di.sm_count * 16) grid = di.sm_count * 16;
    if (dtype == LUTVQ_F16)
        kx8_dequant_kernel<__half, __half2><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
            static_cast<const uint8_t*>(codes), static_cast<__half*>(W), static_cast<const uint4*>(codebooks), static_cast<const __half*>(scales),
            prob_m, prob_k / 8, num_codebooks);
    else
        kx8_dequant_kernel<__nv_bfloat16, __nv_bfloat162><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
            static_cast<const uint8_t*>(codes), static_cast<__nv_bfloat16*>(W), static_cast<const uint4*>(codebooks),
            static_cast<const __nv_bfloat16*>(scales), prob_m, prob_k / 8, num_codebooks);
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

}  // extern "C"
