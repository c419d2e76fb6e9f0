// vq_golden.cu -- the golden model's three functions as separate CUDA kernels, in vq.py's own layouts
// (per-subvector codebooks (M,N,K,d), codeword uint8 (d_out,M,N)) and with vq.py's own rounding, so the
// GPU path can be checked BIT-FOR-BIT against the reference's golden code:
//
//   lut_build_kernel     VQ.compute_lut        vq_dataflow_sim/vq.py:280-285
//                        (fp32 mode: the LUT line of numba_gemv_lut_, aqlm numba_kernel.py:39-40)
//   lut_gemv_kernel      VQ.compute_lut_gemm   vq_dataflow_sim/vq.py:293-298   (float64 accumulate: exact)
//                        (fp32 mode: numba_kernel.py:43-46, sequential j-then-c order per output)
//   vq_dequantize_kernel VQ.dequantize         vq_dataflow_sim/vq.py:269-278   (float64 sum, one fp16 rounding)
//
// These are the unfused K1 / K2 pair of the north star (usable on their own and under ncu); the fused
// production GEMV lives in gemv_lut.cu.  Both kernels are HBM/L2-bound gathers: one warp per output row,
// coalesced code reads, fp64 warp-shuffle reduction.
#include "common.cuh"

namespace lutvq {

// numpy's pairwise fp16 sum for n <= 128 (what ndarray.sum(axis=-1) runs per LUT entry): fp32 lanes r[0..7],
// fixed tree ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)), tail added sequentially.
__device__ __forceinline__ float np_pairwise_sum(const float* v, int n) {
    if (n < 8) {
        float res = 0.f;
        for (int i = 0; i < n; ++i) res = __fadd_rn(res, v[i]);
        return res;
    }
    float r[8];
    #pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = v[i];
    int i = 8;
    for (; i < n - (n % 8); i += 8)
        #pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = __fadd_rn(r[j], v[i + j]);
    float res = __fadd_rn(__fadd_rn(__fadd_rn(r[0], r[1]), __fadd_rn(r[2], r[3])),
                          __fadd_rn(__fadd_rn(r[4], r[5]), __fadd_rn(r[6], r[7])));
    for (; i < n; ++i) res = __fadd_rn(res, v[i]);
    return res;
}

// one thread per LUT entry (m, n, k)
template <int MODE>
__global__ void __launch_bounds__(256)
lut_build_kernel(const __half* __restrict__ x, const __half* __restrict__ codebook, void* __restrict__ lut,
                 int M, int N, int K, int d, int cb_shared) {
    const size_t total = (size_t)M * N * K;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int k = (int)(e % K);
        const int n = (int)((e / K) % N);
        const int m = (int)(e / ((size_t)K * N));
        const __half* c = codebook + (((size_t)(cb_shared ? 0 : m) * N + n) * K + k) * d;
        const __half* xs = x + (size_t)m * d;
        if (MODE == LUTVQ_LUT_VQPY_FP16) {
            float prod[32];
            for (int j = 0; j < d; ++j) prod[j] = __half2float(__hmul(c[j], xs[j]));     // fp16 product, rounded (numpy fp16 multiply)
            const float s = __fadd_rn(0.f, np_pairwise_sum(prod, d));
            static_cast<__half*>(lut)[e] = __float2half_rn(s);
        } else {
            float s = 0.f;
            for (int j = 0; j < d; ++j) s = __fadd_rn(s, __fmul_rn(__half2float(xs[j]), __half2float(c[j])));
            static_cast<float*>(lut)[e] = s;
        }
    }
}

// float64 accumulate: one warp per output row, lanes stride over the M*N tables (exact => order-free)
template <bool LUT_F32>
__global__ void __launch_bounds__(256)
lut_gemv_f64_kernel(const uint8_t* __restrict__ codeword, const void* __restrict__ lut, double* __restrict__ out,
                    int d_out, int M, int N, int K) {
    const int lane = threadIdx.x & 31;
    const int warps_per_block = blockDim.x >> 5;
    const int T = M * N;
    for (int row = blockIdx.x * warps_per_block + (threadIdx.x >> 5); row < d_out; row += gridDim.x * warps_per_block) {
        const uint8_t* cw = codeword + (size_t)row * T;
        double acc = 0.0;
        for (int t = lane; t < T; t += 32) {
            const size_t idx = (size_t)t * K + cw[t];
            acc += LUT_F32 ? (double)static_cast<const float*>(lut)[idx]
                           : (double)__half2float(static_cast<const __half*>(lut)[idx]);
        }
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) out[row] = acc;
    }
}

// float32 accumulate in the reference's order (groups ascending, codebooks inner): one thread per row
template <bool LUT_F32>
__global__ void __launch_bounds__(128)
lut_gemv_f32_kernel(const uint8_t* __restrict__ codeword, const void* __restrict__ lut, float* __restrict__ out,
                    int d_out, int M, int N, int K) {
    const int T = M * N;
    for (int row = blockIdx.x * blockDim.x + threadIdx.x; row < d_out; row += gridDim.x * blockDim.x) {
        const uint8_t* cw = codeword + (size_t)row * T;
        float acc = 0.f;
        for (int t = 0; t < T; ++t) {
            const size_t idx = (size_t)t * K + cw[t];
            acc = __fadd_rn(acc, LUT_F32 ? static_cast<const float*>(lut)[idx]
                                         : __half2float(static_cast<const __half*>(lut)[idx]));
        }
        out[row] = acc;
    }
}

// one thread per (row j, subvector m): d outputs, float64 sum over the N codebooks, one fp16 rounding
__global__ void __launch_bounds__(256)
vq_dequantize_kernel(const __half* __restrict__ codebook, const uint8_t* __restrict__ codeword, __half* __restrict__ W,
                     int d_out, int M, int N, int K, int d, int cb_shared) {
    const size_t total = (size_t)d_out * M;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int m = (int)(e % M);
        const uint8_t* cw = codeword + e * N;
        for (int i = 0; i < d; ++i) {
            double s = 0.0;
            for (int n = 0; n < N; ++n)
                s += (double)__half2float(codebook[(((size_t)(cb_shared ? 0 : m) * N + n) * K + cw[n]) * d + i]);
            W[e * d + i] = __double2half(s);
        }
    }
}

static int grid_for(size_t work_items, int block, int sm_count) {
    size_t g = (work_items + block - 1) / block;
    const size_t cap = (size_t)sm_count * 16;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

int lut_build_launch(const void* x, const void* codebook, void* lut, int M, int N, int K, int d, int cb_shared,
                     int lut_mode, cudaStream_t stream) {
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    const int grid = grid_for((size_t)M * N * K, 256, di.sm_count);
    if (lut_mode == LUTVQ_LUT_VQPY_FP16)
        lut_build_kernel<LUTVQ_LUT_VQPY_FP16><<<grid, 256, 0, stream>>>(
            static_cast<const __half*>(x), static_cast<const __half*>(codebook), lut, M, N, K, d, cb_shared);
    else
        lut_build_kernel<LUTVQ_LUT_FP32><<<grid, 256, 0, stream>>>(
            static_cast<const __half*>(x), static_cast<const __half*>(codebook), lut, M, N, K, d, cb_shared);
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

int lut_gemv_launch(const void* codeword, const void* lut, void* out, int d_out, int M, int N, int K,
                    int lut_is_fp32, int acc_mode, cudaStream_t stream) {
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    const uint8_t* cw = static_cast<const uint8_t*>(codeword);
    if (acc_mode == LUTVQ_ACC_F64) {
        const int grid = grid_for((size_t)d_out * 32, 256, di.sm_count);
        if (lut_is_fp32) lut_gemv_f64_kernel<true><<<grid, 256, 0, stream>>>(cw, lut, static_cast<double*>(out), d_out, M, N, K);
        else             lut_gemv_f64_kernel<false><<<grid, 256, 0, stream>>>(cw, lut, static_cast<double*>(out), d_out, M, N, K);
    } else {
        const int grid = grid_for((size_t)d_out, 128, di.sm_count);
        if (lut_is_fp32) lut_gemv_f32_kernel<true><<<grid, 128, 0, stream>>>(cw, lut, static_cast<float*>(out), d_out, M, N, K);
        else             lut_gemv_f32_kernel<false><<<grid, 128, 0, stream>>>(cw, lut, static_cast<float*>(out), d_out, M, N, K);
    }
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

int vq_dequantize_launch(const void* codebook, const void* codeword, void* W, int d_out, int M, int N, int K, int d,
                         int cb_shared, cudaStream_t stream) {
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    const int grid = grid_for((size_t)d_out * M, 256, di.sm_count);
    vq_dequantize_kernel<<<grid, 256, 0, stream>>>(static_cast<const __half*>(codebook),
                                                   static_cast<const uint8_t*>(codeword),
                                                   static_cast<__half*>(W), d_out, M, N, K, d, cb_shared);
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

}  // namespace lutvq
