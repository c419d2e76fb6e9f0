// simd_lut.cu -- SURVEY 8a row a13: the integer-LUT "lookup and add" microkernel of the reference's
// benchmark/kernel/simd_lut/bench_avx_lut.cpp (lookup_and_add_simd, :26-138), _SCALAR_ semantics (:98-107):
//     for run < n_runs:  for j < n_lut_per_run:  for k < CB:   output[k] += lut[ idx[j*CB + k] ]      (int16 wraparound)
// The reference walks this on ONE core with the LUT in a vector register (vpshufb); the sums are additions modulo 2^16,
// so they can be split over any number of workers and added in any order without changing a bit.  Here: every CTA takes
// a share of the runs (each run really walks all the indices again -- no algebraic shortcut, the benchmark's unit is
// the lookup), a thread owns 4 adjacent byte lanes k (one 32-bit word of indices per row) and a strided set of rows,
// the table sits sign-extended in shared memory (one LDS.32 per lookup, the bound of this kernel: 32 lookups per clock
// and SM), partial sums are 32-bit wraparound (exact modulo 2^16), combined through shared then global atomics; a
// second tiny kernel folds them into the int16 output and hands the workspace back zeroed.
#include "common.cuh"

namespace lutvq {

constexpr int kSimdLutThreads = 256;

__global__ void __launch_bounds__(kSimdLutThreads) simd_lut_kernel(const uint32_t* __restrict__ idx32, const int8_t* __restrict__ lut,
                                                                  unsigned* __restrict__ acc, int wpr, int n_entries,
                                                                  unsigned long long n_runs, unsigned long long n_rows) {
    __shared__ int s_lut[256];
    __shared__ unsigned s_acc[256];
    for (int i = threadIdx.x; i < 256; i += kSimdLutThreads) { s_lut[i] = i < n_entries ? (int)lut[i] : 0; s_acc[i] = 0u; }
    __syncthreads();
    const int w = threadIdx.x % wpr, r0 = threadIdx.x / wpr, rstep = kSimdLutThreads / wpr;
    unsigned a0 = 0u, a1 = 0u, a2 = 0u, a3 = 0u;
    if (r0 < rstep)                                                  // threads past the last whole row group idle
        for (unsigned long long run = blockIdx.x; run < n_runs; run += gridDim.x)
        {
            // 4 rows per trip: the index loads of a trip are issued together, then their 16 table reads
            const uint32_t* pidx = idx32 + w;
            const unsigned n = (unsigned)n_rows, step = (unsigned)rstep;
            unsigned j = (unsigned)r0;
            for (; j + 3u * step < n; j += 4u * step) {
                uint32_t v[4];
                #pragma unroll
                for (int u = 0; u < 4; ++u) v[u] = __ldg(pidx + (size_t)(j + u * step) * wpr);
                #pragma unroll
                for (int u = 0; u < 4; ++u) {
                    a0 += (unsigned)s_lut[v[u] & 0xFFu];
                    a1 += (unsigned)s_lut[(v[u] >> 8) & 0xFFu];
                    a2 += (unsigned)s_lut[(v[u] >> 16) & 0xFFu];
                    a3 += (unsigned)s_lut[v[u] >> 24];
                }
            }
            for (; j < n; j += step) {
                const uint32_t v = __ldg(pidx + (size_t)j * wpr);
                a0 += (unsigned)s_lut[v & 0xFFu];
                a1 += (unsigned)s_lut[(v >> 8) & 0xFFu];
                a2 += (unsigned)s_lut[(v >> 16) & 0xFFu];
                a3 += (unsigned)s_lut[v >> 24];
            }
        }
    atomicAdd(&s_acc[4 * w + 0], a0); atomicAdd(&s_acc[4 * w + 1], a1);
    atomicAdd(&s_acc[4 * w + 2], a2); atomicAdd(&s_acc[4 * w + 3], a3);
    __syncthreads();
    if (threadIdx.x < 4 * wpr && s_acc[threadIdx.x] != 0u) atomicAdd(acc + threadIdx.x, s_acc[threadIdx.x]);
}

__global__ void simd_lut_finish_kernel(unsigned* acc, int16_t* output, int cb) {
    const int k = threadIdx.x;
    if (k < cb) {
        output[k] = (int16_t)(uint16_t)((unsigned)(uint16_t)output[k] + acc[k]);   // modulo 2^16, like the int16 += of the reference
        acc[k] = 0u;
    }
}

}  // namespace lutvq

using namespace lutvq;

extern "C" {

size_t lutvq_simd_lut_workspace_bytes(int cb_size) { return cb_size > 0 ? 256 * sizeof(unsigned) : 0; }

int lutvq_simd_lut_lookup_add(const void* idx, const void* lut, void* output, int cb_size, int n_entries,
                              unsigned long long n_runs, unsigned long long n_lut_per_run, void* ws, size_t ws_bytes, void* stream) {
    LUTVQ_REQUIRE(idx && lut && output, "null pointer");
    LUTVQ_REQUIRE(cb_size > 0 && cb_size <= 256 && cb_size % 4 == 0 && kSimdLutThreads % (cb_size / 4) == 0,
                  "cb_size=%d must be a multiple of 4 that divides 1024 (16 / 32 / 64 in the reference)", cb_size);
    LUTVQ_REQUIRE(n_entries > 0 && n_entries <= 256, "n_entries=%d must be in 1..256", n_entries);
    LUTVQ_REQUIRE(n_lut_per_run < (1ull << 31), "n_lut_per_run must be below 2^31");
    LUTVQ_REQUIRE((reinterpret_cast<uintptr_t>(idx) & 3u) == 0, "idx must be 4-byte aligned");
    LUTVQ_REQUIRE(ws != nullptr && ws_bytes >= lutvq_simd_lut_workspace_bytes(cb_size) && (reinterpret_cast<uintptr_t>(ws) & 3u) == 0,
                  "workspace too small or misaligned (zero-initialised, %zu bytes)", lutvq_simd_lut_workspace_bytes(cb_size));
    if (n_runs == 0 || n_lut_per_run == 0) return LUTVQ_OK;
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    unsigned long long grid = n_runs < (unsigned long long)di.sm_count * 8 ? n_runs : (unsigned long long)di.sm_count * 8;
    simd_lut_kernel<<<(unsigned)grid, kSimdLutThreads, 0, st>>>(static_cast<const uint32_t*>(idx), static_cast<const int8_t*>(lut),
                                                               static_cast<unsigned*>(ws), cb_size / 4, n_entries, n_runs, n_lut_per_run);
    LUTVQ_CUDA_OK(cudaGetLastError());
    simd_lut_finish_kernel<<<1, 256, 0, st>>>(static_cast<unsigned*>(ws), static_cast<int16_t*>(output), cb_size);
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

}  // extern "C"
