// dequant.cu -- AQLM 2x8 codebook dequantize: W[o][8g..8g+7] = cb0[c0(o,g)] + cb1[c1(o,g)]  (then * scale[o]).
//
// Replaces Code2x8Dequant + the torch `weight *= scales` that follows it (reference benchmark/AQLM/
// inference_lib/src/aqlm/inference_kernels/cuda_kernel.cu:235-294, cuda_kernel.cpp:423-448).  One RN add
// in the storage dtype per element (vq.py:276-278 for N=2: float64 sum then one rounding == one IEEE add),
// and the scale is a second, separate RN multiply exactly like the reference -- both fused into the
// store here, so the result is bit-identical while W is written once instead of three times.
//
// HBM-bound on the 2*O*I-byte write stream.  Persistent CTAs (a multiple of the SM count), one warp per
// output row at a time, one thread per 8-element group: a 2-byte code load, two 128-bit shared-memory
// reads, one 128-bit coalesced global store.  The two 256x8 codebooks are replicated 8x in shared memory,
// copy (lane & 7) private to each lane of a quarter-warp, so the 128-bit reads never bank-conflict
// whatever the indices are.
// Algorithmic bytes per launch: O*(I/8)*2 (codes) + 8192 (codebooks) + 2*O (scales) + 2*O*I (W).
#include "common.cuh"

namespace lutvq {

constexpr int kDqThreads = 512;                  // 16 warps x 2 CTAs / SM: enough loads in flight for the dependent code -> smem -> store chain
constexpr int kDqSmemBytes = 2 * 256 * 8 * 16;   // 64 KiB: [cb][idx][copy] x 16 bytes

template <int DT>
__global__ void __launch_bounds__(kDqThreads, 2)
code2x8_dequant_kernel(const uint16_t* __restrict__ codes, const uint4* __restrict__ codebook,
                       const void* __restrict__ scales, uint4* __restrict__ W, int O, int G) {
    using D = Dtype<DT>;
    using T = typename D::T;
    using T2 = typename D::T2;
    extern __shared__ __align__(16) uint4 sh_cb[];          // [2*256][8 copies]
    for (int i = threadIdx.x; i < 2 * 256 * 8; i += kDqThreads) sh_cb[i] = __ldg(codebook + (i >> 3));
    __syncthreads();

    const int lane = threadIdx.x & 31, copy = lane & 7;
    const int warps = kDqThreads / 32;
    const uint4* cb0 = sh_cb + copy;
    const uint4* cb1 = sh_cb + 256 * 8 + copy;
    for (int row = blockIdx.x * warps + (threadIdx.x >> 5); row < O; row += gridDim.x * warps) {
        const uint16_t* crow = codes + (size_t)row * G;
        uint4* wrow = W + (size_t)row * G;
        T2 s2;
        const bool has_scale = scales != nullptr;
        if (has_scale) s2 = D::dup(static_cast<const T*>(scales)[row]);
        // 4 groups per lane and trip: the 4 code loads are issued together (a lone 2-byte load per trip left the kernel
        // latency-bound at 2.2 TB/s of writes with 16 warps per SM)
        for (int g0 = lane; g0 < G; g0 += 128) {
            uint32_t cc[4];
            #pragma unroll
            for (int u = 0; u < 4; ++u) cc[u] = (g0 + 32 * u < G) ? crow[g0 + 32 * u] : 0u;   // byte 0 = codebook-0 index, byte 1 = codebook-1 index
            #pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int g = g0 + 32 * u;
                if (g >= G) break;
                const uint4 a = cb0[(cc[u] & 0xFFu) * 8];
                const uint4 b = cb1[(cc[u] >> 8) * 8];
                uint4 r;
                const T2* a2 = reinterpret_cast<const T2*>(&a);
                const T2* b2 = reinterpret_cast<const T2*>(&b);
                T2* r2 = reinterpret_cast<T2*>(&r);
                #pragma unroll
                for (int j = 0; j < 4; ++j) {
                    T2 v = D::add2(a2[j], b2[j]);
                    if (has_scale) v = D::mul2(v, s2);
                    r2[j] = v;
                }
                __stcs(wrow + g, r);                         // streaming store: W is consumed by another kernel
            }
        }
    }
}

int code2x8_dequant_launch(const void* codes, void* W, const void* codebook, const void* scales, int O, int I,
                           int dtype, cudaStream_t stream) {
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    const int G = I / 8;
    const int warps = kDqThreads / 32;
    int grid = di.sm_count * 2;
    if (grid > ceil_div(O, warps)) grid = ceil_div(O, warps);
    static thread_local int attr_device = -1;
    int dev = 0;
    LUTVQ_CUDA_OK(cudaGetDevice(&dev));
    if (attr_device != dev) {
        LUTVQ_CUDA_OK(cudaFuncSetAttribute(code2x8_dequant_kernel<LUTVQ_F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, kDqSmemBytes));
        LUTVQ_CUDA_OK(cudaFuncSetAttribute(code2x8_dequant_kernel<LUTVQ_BF16>, cudaFuncAttributeMaxDynamicSharedMemorySize, kDqSmemBytes));
        attr_device = dev;
    }
    if (dtype == LUTVQ_F16)
        code2x8_dequant_kernel<LUTVQ_F16><<<grid, kDqThreads, kDqSmemBytes, stream>>>(
            static_cast<const uint16_t*>(codes), static_cast<const uint4*>(codebook), scales, static_cast<uint4*>(W), O, G);
    else if (dtype == LUTVQ_BF16)
        code2x8_dequant_kernel<LUTVQ_BF16><<<grid, kDqThreads, kDqSmemBytes, stream>>>(
            static_cast<const uint16_t*>(codes), static_cast<const uint4*>(codebook), scales, static_cast<uint4*>(W), O, G);
    else
        return fail(LUTVQ_EUNSUPPORTED, "dtype %d not supported (fp16=0, bf16=1)", dtype);
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

}  // namespace lutvq
