// api.cu -- the extern "C" surface of liblutvq.so (include/lutvq.h): argument validation, error strings,
// device-info cache, dispatch to the kernel launchers.  No torch types, no allocation, no host sync.
#include "common.cuh"

#include <mutex>
#include <string.h>

namespace lutvq {

// launchers (gemv_lut.cu, dequant.cu, vq_golden.cu, gemm_tc.cu)
int gemv_2x8_launch(const void* packed, const void* x, void* y, const void* codebook, const void* scales,
                    const void* bias, int O, int I, int dtype, unsigned flags, void* ws, size_t ws_bytes,
                    cudaStream_t stream, const GemvPush* push = nullptr);
int gemv_2x8_launch_rows(const void* codes, int packed, const void* x, void* y, const void* codebook, const void* scales, const void* bias,
                         int rows, int O, int I, int dtype, unsigned flags, void* ws, size_t ws_bytes, cudaStream_t stream, int ncb = 2);
int gemv_max_rows_per_launch();
size_t gemv_rows_workspace_bytes(int rows, int O);
int gemv_2x8_launch_ref(const void* codes, const void* x, void* y, const void* codebook, const void* scales,
                        const void* bias, int O, int I, int dtype, unsigned flags, void* ws, size_t ws_bytes, cudaStream_t stream);
size_t gemv_workspace_bytes(int O, int G);
size_t gemv_packed_workspace_bytes(int O, int G);
size_t gemv_group_workspace_bytes(int n, const int* O);
void gemv_set_dependent_workers(int m);
int gemv_2x8_launch_group(int n, const GemvItem* items, const void* x, const void* x_gate, int I, int dtype, unsigned flags,
                          void* ws, size_t ws_bytes, cudaStream_t stream, const GemvPush* push, const GemvLL* ll = nullptr);
size_t gemv_packed_bytes(int O, int G);
int gemv_prepack_launch(const void* codes, void* packed, int O, int I, cudaStream_t stream, void* zero, size_t zero_bytes);
void gemv_set_tuning(int nw, int pf, int pa);
void gemv_set_timeline(void* buf);
void gemm_set_debug_mode(int m);
void gemm_set_stamps(void* dev_ptr);
void gemv_get_plan(int O, int I, int* n_k, int* n_chunks, int* rows_per_chunk);
int code2x8_dequant_launch(const void* codes, void* W, const void* codebook, const void* scales, int O, int I,
                           int dtype, cudaStream_t stream);
int lut_build_launch(const void* x, const void* codebook, void* lut, int M, int N, int K, int d, int cb_shared,
                     int lut_mode, cudaStream_t stream);
int lut_gemv_launch(const void* codeword, const void* lut, void* out, int d_out, int M, int N, int K,
                    int lut_is_fp32, int acc_mode, cudaStream_t stream);
int vq_dequantize_launch(const void* codebook, const void* codeword, void* W, int d_out, int M, int N, int K, int d,
                         int cb_shared, cudaStream_t stream);
int gemm_dequant_tc_launch(const void* x, const void* codes, const void* codebooks, const void* scales,
                           const void* bias, void* y, int batch, int O, int I, int dtype, void* ws, size_t ws_bytes,
                           cudaStream_t stream);
size_t gemm_workspace_bytes(int O, int I);
size_t gemm_workspace_bytes_for(int O, int I, int max_batch);

char* err_buf() {
    static thread_local char buf[512] = "";
    return buf;
}
int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err_buf(), 512, fmt, ap);
    va_end(ap);
    return code;
}

int get_device_info(DeviceInfo* out) {
    static std::mutex mu;
    static DeviceInfo cache[64];
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess)
        return fail(LUTVQ_ECUDA, "no usable CUDA device (%s); liblutvq has no CPU fallback", cudaGetErrorString(e));
    if (dev < 0 || dev >= 64) return fail(LUTVQ_ECUDA, "device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(mu);
    DeviceInfo& c = cache[dev];
    if (c.device != dev) {
        int v = 0;
        LUTVQ_CUDA_OK(cudaDeviceGetAttribute(&c.sm_count, cudaDevAttrMultiProcessorCount, dev));
        LUTVQ_CUDA_OK(cudaDeviceGetAttribute(&c.cc_major, cudaDevAttrComputeCapabilityMajor, dev));
        LUTVQ_CUDA_OK(cudaDeviceGetAttribute(&c.cc_minor, cudaDevAttrComputeCapabilityMinor, dev));
        LUTVQ_CUDA_OK(cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        c.smem_optin = (size_t)v;
        if (c.cc_major != 10)
            return fail(LUTVQ_EUNSUPPORTED, "liblutvq is built for sm_100a only; device %d is sm_%d%d", dev, c.cc_major, c.cc_minor);
        c.device = dev;
    }
    *out = c;
    return LUTVQ_OK;
}

static int check_2x8_common(const void* codes, const void* codebook, int prob_m, int prob_k, int dtype) {
    LUTVQ_REQUIRE(codes && codebook, "null codes/codebook pointer");
    LUTVQ_REQUIRE(prob_m > 0 && prob_k > 0, "bad shape prob_m=%d prob_k=%d", prob_m, prob_k);
    LUTVQ_REQUIRE(prob_k % 64 == 0, "prob_k=%d must be a multiple of 64 (16-byte code rows, as cuda_kernel.cu:154 assumes)", prob_k);
    LUTVQ_REQUIRE(aligned16(codes) && aligned16(codebook), "codes and codebook must be 16-byte aligned");
    if (dtype != LUTVQ_F16 && dtype != LUTVQ_BF16)
        return fail(LUTVQ_EUNSUPPORTED, "AQLM kernels only support float16 (0) and bfloat16 (1); got dtype %d", dtype);
    return LUTVQ_OK;
}

}  // namespace lutvq

using namespace lutvq;

extern "C" {

const char* lutvq_last_error(void) { return err_buf(); }
int lutvq_version(void) { return 10000 * 0 + 100 * 1 + 0; }

int lutvq_device_info(int* sm_count, int* cc_major, int* cc_minor, size_t* smem_optin_bytes) {
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    if (sm_count) *sm_count = di.sm_count;
    if (cc_major) *cc_major = di.cc_major;
    if (cc_minor) *cc_minor = di.cc_minor;
    if (smem_optin_bytes) *smem_optin_bytes = di.smem_optin;
    return LUTVQ_OK;
}

size_t lutvq_workspace_bytes(int prob_m, int prob_k) {
    if (prob_m <= 0 || prob_k <= 0) return 0;
    size_t a = gemv_workspace_bytes(prob_m, prob_k / 8);
    size_t b = gemm_workspace_bytes(prob_m, prob_k);
    return a > b ? a : b;
}

size_t lutvq_packed_workspace_bytes(int prob_m, int prob_k) {
    if (prob_m <= 0 || prob_k <= 0) return 0;
    return gemv_packed_workspace_bytes(prob_m, prob_k / 8);
}

size_t lutvq_workspace_bytes_ex(int prob_m, int prob_k, int max_batch) {
    if (prob_m <= 0 || prob_k <= 0) return 0;
    size_t a = gemv_workspace_bytes(prob_m, prob_k / 8);
    size_t b = max_batch > 6 ? gemm_workspace_bytes_for(prob_m, prob_k, max_batch) : 0;   // batch <= 6 never takes the GEMM path
    return a > b ? a : b;
}

int lutvq_code2x8_matvec(const void* codes, const void* x, void* y, const void* codebook, int prob_m, int prob_k,
                         int dtype, void* ws, size_t ws_bytes, void* stream) {
    if (int rc = check_2x8_common(codes, codebook, prob_m, prob_k, dtype)) return rc;
    LUTVQ_REQUIRE(x && y, "null x/y pointer");
    return gemv_2x8_launch_ref(codes, x, y, codebook, nullptr, nullptr, prob_m, prob_k, dtype, 0u, ws, ws_bytes,
                               static_cast<cudaStream_t>(stream));
}

static int matmat_rows(const void* x, const void* codes, int packed, const void* codebooks, const void* scales,
                       const void* bias, void* y, int batch, int prob_m, int prob_k, int dtype, unsigned flags,
                       void* ws, size_t ws_bytes, void* stream) {
    if (int rc = check_2x8_common(codes, codebooks, prob_m, prob_k, dtype)) return rc;
    LUTVQ_REQUIRE(batch >= 0, "bad batch %d", batch);
    if (batch == 0) return LUTVQ_OK;                                  // empty input: nothing to launch (pointers may be null)
    LUTVQ_REQUIRE(x && y, "null x/y pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t region = ws_bytes;
    // cuda_kernel.cpp:404-414 loops the matvec over the flattened batch rows; so do we (one fused launch per row,
    // scale/bias folded into the kernel's epilogue instead of three extra elementwise kernels).
    // Small batches go out as ONE launch per gemv_max_rows_per_launch() rows when the workspace holds their split-k words
    // (gemv_rows_workspace_bytes; a ring of min(batch, LUTVQ_XREADY_WS_RING) single-row regions is enough): the rows share the
    // kernel's fixed cost and the code stream (gemv_lut.cu, gemv_2x8_launch_rows).
    {
        const int per = batch < gemv_max_rows_per_launch() ? batch : gemv_max_rows_per_launch();
        if (batch > 1 && ws_bytes >= gemv_rows_workspace_bytes(per, prob_m)) {
            for (int b0 = 0; b0 < batch; b0 += per) {
                const int nb = batch - b0 < per ? batch - b0 : per;
                unsigned f = flags;
                if (b0 > 0) f &= ~LUTVQ_FLAG_X_READY;                 // the chunks of one call share the workspace: they run in turn
                if (int rc = gemv_2x8_launch_rows(codes, packed, static_cast<const uint8_t*>(x) + (size_t)b0 * prob_k * 2,
                                                  static_cast<uint8_t*>(y) + (size_t)b0 * prob_m * 2, codebooks, scales, bias, nb, prob_m, prob_k,
                                                  dtype, f, ws, ws_bytes, st))
                    return rc;
            }
            return LUTVQ_OK;
        }
    }
    // Otherwise one launch per row.  X_READY launches overlap under programmatic dependent launch, so consecutive rows must not
    // share split-k words: with a single-region workspace the promise is dropped for the rows after the first.
    for (int b = 0; b < batch; ++b) {
        const void* xb = static_cast<const uint8_t*>(x) + (size_t)b * prob_k * 2;
        void* yb = static_cast<uint8_t*>(y) + (size_t)b * prob_m * 2;
        unsigned f = flags;
        if (b > 0) f &= ~LUTVQ_FLAG_X_READY;
        if (int rc = packed ? gemv_2x8_launch(codes, xb, yb, codebooks, scales, bias, prob_m, prob_k, dtype, f, ws, region, st)
                            : gemv_2x8_launch_ref(codes, xb, yb, codebooks, scales, bias, prob_m, prob_k, dtype, f, ws, region, st))
            return rc;
    }
    return LUTVQ_OK;
}

int lutvq_code2x8_matmat_ex(const void* x, const void* codes, const void* codebooks, const void* scales,
                            const void* bias, void* y, int batch, int prob_m, int prob_k, int dtype, unsigned flags,
                            void* ws, size_t ws_bytes, void* stream) {
    return matmat_rows(x, codes, 0, codebooks, scales, bias, y, batch, prob_m, prob_k, dtype, flags, ws, ws_bytes, stream);
}

size_t lutvq_code2x8_packed_bytes(int prob_m, int prob_k) {
    if (prob_m <= 0 || prob_k <= 0) return 0;
    return gemv_packed_bytes(prob_m, prob_k / 8);
}

int lutvq_code2x8_prepack(const void* codes, void* packed, int prob_m, int prob_k, void* stream) {
    LUTVQ_REQUIRE(codes && packed, "null codes/packed pointer");
    LUTVQ_REQUIRE(prob_m > 0 && prob_k > 0 && prob_k % 64 == 0, "bad shape prob_m=%d prob_k=%d (prob_k %% 64 == 0)", prob_m, prob_k);
    LUTVQ_REQUIRE(aligned16(codes) && aligned16(packed), "codes and packed must be 16-byte aligned");
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    return gemv_prepack_launch(codes, packed, prob_m, prob_k, static_cast<cudaStream_t>(stream), nullptr, 0);
}

int lutvq_code2x8_matmat_packed(const void* x, const void* packed, const void* codebooks, const void* scales,
                                const void* bias, void* y, int batch, int prob_m, int prob_k, int dtype,
                                unsigned flags, void* ws, size_t ws_bytes, void* stream) {
    return matmat_rows(x, packed, 1, codebooks, scales, bias, y, batch, prob_m, prob_k, dtype, flags, ws, ws_bytes, stream);
}

size_t lutvq_group_workspace_bytes(int n, const int* prob_m, int prob_k) {
    if (n < 1 || n > 3 || !prob_m || prob_k <= 0) return 0;
    return gemv_group_workspace_bytes(n, prob_m);
}

int lutvq_code2x8_matmat_group_packed(const void* x, int n, const void* const* packed, const void* const* codebooks,
                                      const void* const* scales, const void* const* bias, void* const* y, const int* prob_m,
                                      int prob_k, int dtype, unsigned flags, void* ws, size_t ws_bytes, void* stream) {
    LUTVQ_REQUIRE(n >= 1 && n <= 3, "a group holds 1..3 linears, got %d", n);
    LUTVQ_REQUIRE(x && packed && codebooks && y && prob_m, "null pointer");
    GemvItem items[3];
    for (int i = 0; i < n; ++i) {
        if (int rc = check_2x8_common(packed[i], codebooks[i], prob_m[i], prob_k, dtype)) return rc;
        LUTVQ_REQUIRE(y[i] != nullptr, "null y pointer of linear %d", i);
        items[i] = GemvItem{packed[i], codebooks[i], y[i], scales ? scales[i] : nullptr, bias ? bias[i] : nullptr, prob_m[i]};
    }
    return gemv_2x8_launch_group(n, items, x, nullptr, prob_k, dtype, flags, ws, ws_bytes, static_cast<cudaStream_t>(stream), nullptr);
}

int lutvq_code2x8_matmat_group_packed_swiglu(const void* gate, const void* up, int n, const void* const* packed,
                                             const void* const* codebooks, const void* const* scales, const void* const* bias,
                                             void* const* y, const int* prob_m, int prob_k, int dtype, unsigned flags, void* ws,
                                             size_t ws_bytes, void* stream) {
    LUTVQ_REQUIRE(gate && up, "null gate/up pointer");
    LUTVQ_REQUIRE(dtype == LUTVQ_F16, "the fused SwiGLU input is fp16 only (dtype %d)", dtype);
    LUTVQ_REQUIRE(n >= 1 && n <= 3, "a group holds 1..3 linears, got %d", n);
    LUTVQ_REQUIRE(packed && codebooks && y && prob_m, "null pointer");
    GemvItem items[3];
    for (int i = 0; i < n; ++i) {
        if (int rc = check_2x8_common(packed[i], codebooks[i], prob_m[i], prob_k, dtype)) return rc;
        LUTVQ_REQUIRE(y[i] != nullptr, "null y pointer of linear %d", i);
        items[i] = GemvItem{packed[i], codebooks[i], y[i], scales ? scales[i] : nullptr, bias ? bias[i] : nullptr, prob_m[i]};
    }
    return gemv_2x8_launch_group(n, items, up, gate, prob_k, dtype, flags, ws, ws_bytes, static_cast<cudaStream_t>(stream), nullptr);
}

int lutvq_code2x8_matmat(const void* x, const void* codes, const void* codebooks, const void* scales, const void* bias,
                         void* y, int batch, int prob_m, int prob_k, int dtype, void* ws, size_t ws_bytes, void* stream) {
    return lutvq_code2x8_matmat_ex(x, codes, codebooks, scales, bias, y, batch, prob_m, prob_k, dtype, 0u, ws, ws_bytes, stream);
}

int lutvq_code2x8_dequant(const void* codes, void* W, const void* codebook, const void* scales, int prob_m, int prob_k,
                          int dtype, void* stream) {
    if (int rc = check_2x8_common(codes, codebook, prob_m, prob_k, dtype)) return rc;
    LUTVQ_REQUIRE(W && aligned16(W), "W must be non-null and 16-byte aligned");
    return code2x8_dequant_launch(codes, W, codebook, scales, prob_m, prob_k, dtype, static_cast<cudaStream_t>(stream));
}

int lutvq_code2x8_matmat_dequant(const void* x, const void* codes, const void* codebooks, const void* scales,
                                 const void* bias, void* y, int batch, int prob_m, int prob_k, int dtype, void* ws,
                                 size_t ws_bytes, void* stream) {
    if (int rc = check_2x8_common(codes, codebooks, prob_m, prob_k, dtype)) return rc;
    LUTVQ_REQUIRE(batch >= 0, "bad batch %d", batch);
    if (batch == 0) return LUTVQ_OK;
    LUTVQ_REQUIRE(x && y, "null x/y pointer");
    LUTVQ_REQUIRE(aligned16(x) && aligned16(y), "x and y must be 16-byte aligned");
    return gemm_dequant_tc_launch(x, codes, codebooks, scales, bias, y, batch, prob_m, prob_k, dtype, ws, ws_bytes,
                                  static_cast<cudaStream_t>(stream));
}

int lutvq_lut_build(const void* x, const void* codebook, void* lut, int M, int N, int K, int d, int cb_shared,
                    int lut_mode, void* stream) {
    LUTVQ_REQUIRE(x && codebook && lut, "null pointer");
    LUTVQ_REQUIRE(M > 0 && N > 0 && K > 0 && d > 0 && d <= 32 && K <= 65536, "bad shape M=%d N=%d K=%d d=%d", M, N, K, d);
    LUTVQ_REQUIRE(lut_mode == LUTVQ_LUT_VQPY_FP16 || lut_mode == LUTVQ_LUT_FP32, "bad lut_mode %d", lut_mode);
    return lut_build_launch(x, codebook, lut, M, N, K, d, cb_shared, lut_mode, static_cast<cudaStream_t>(stream));
}

int lutvq_lut_gemv(const void* codeword, const void* lut, void* out, int d_out, int M, int N, int K, int lut_is_fp32,
                   int acc_mode, void* stream) {
    LUTVQ_REQUIRE(codeword && lut && out, "null pointer");
    LUTVQ_REQUIRE(d_out > 0 && M > 0 && N > 0 && K > 0 && K <= 256, "bad shape d_out=%d M=%d N=%d K=%d (uint8 codes: K<=256)", d_out, M, N, K);
    LUTVQ_REQUIRE(acc_mode == LUTVQ_ACC_F64 || acc_mode == LUTVQ_ACC_F32, "bad acc_mode %d", acc_mode);
    return lut_gemv_launch(codeword, lut, out, d_out, M, N, K, lut_is_fp32, acc_mode, static_cast<cudaStream_t>(stream));
}

int lutvq_vq_dequantize(const void* codebook, const void* codeword, void* W, int d_out, int M, int N, int K, int d,
                        int cb_shared, void* stream) {
    LUTVQ_REQUIRE(codebook && codeword && W, "null pointer");
    LUTVQ_REQUIRE(d_out > 0 && M > 0 && N > 0 && K > 0 && K <= 256 && d > 0, "bad shape");
    return vq_dequantize_launch(codebook, codeword, W, d_out, M, N, K, d, cb_shared, static_cast<cudaStream_t>(stream));
}

// ---- not part of the public header: tuning hooks used by bench.py's sweep and the tests ----
void lutvq_debug_set_gemv_tuning(int nw, int pf, int pa) { gemv_set_tuning(nw, pf, pa); }
void lutvq_debug_set_timeline(void* buf) { gemv_set_timeline(buf); }
void lutvq_debug_set_dependent_workers(int per_sm) { gemv_set_dependent_workers(per_sm); }
void lutvq_debug_set_gemm_mode(int m) { gemm_set_debug_mode(m); }
void lutvq_debug_set_gemm_stamps(void* dev_ptr) { gemm_set_stamps(dev_ptr); }
void lutvq_debug_get_gemv_plan(int prob_m, int prob_k, int* n_k, int* n_chunks, int* rows_per_chunk) {
    gemv_get_plan(prob_m, prob_k, n_k, n_chunks, rows_per_chunk);
}

}  // extern "C"
