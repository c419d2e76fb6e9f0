// noalloc.cu -- lutvq_code2x8_matmat_noalloc: the fused LUT-GEMV of lutvq_code2x8_matmat for callers that cannot provide (or
// allocate) a workspace: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
// No counterpart in the reference.
//
// The split-k accumulator words the GEMV kernel needs (256-byte header + one 64-bit word per output row, rounded up to 32 rows)
// live in a file-scope `static __device__` array of LUTVQ_NOALLOC_MAX_ROWS rows (257 KiB; every GPU that loads the library holds
// it until exit).  In host code the array's NAME is only a host-side stand-in, not a device address: the address the kernel gets
// is looked up once per device with cudaGetSymbolAddress while that device is current.  The entry point calls no cudaMalloc /
// cudaFree / allocating library routine.  Every call shares the one array, which keeps whatever the last call left: the region
// a call uses is cleared (stream-ordered memset) before its launches.  Calls that overlap on the GPU (two streams, or two
// processes' worth of threads on one device) race on it -- serialise them, or use lutvq_code2x8_matmat with private workspaces.
#include "common.cuh"

#include <mutex>

namespace lutvq {

size_t gemv_workspace_bytes(int O, int G);
int gemv_2x8_launch_ref(const void* codes, const void* x, void* y, const void* codebook, const void* scales,
                        const void* bias, int O, int I, int dtype, unsigned flags, void* ws, size_t ws_bytes, cudaStream_t stream);

constexpr size_t kNoAllocScratchBytes = 256 + (size_t)LUTVQ_NOALLOC_MAX_ROWS * sizeof(unsigned long long);
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

constexpr int kNoAllocMaxDevices = 64;
static std::mutex g_noalloc_mutex;
static void* g_noalloc_ptr[kNoAllocMaxDevices] = {};

// device address of the scratch array on the CURRENT device (looked up once per device, cached)
static int noalloc_scratch(void** out) {
    int dev = -1;
    LUTVQ_CUDA_OK(cudaGetDevice(&dev));
    LUTVQ_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, "device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_ptr[dev] == nullptr) {
        void* p = nullptr;
        LUTVQ_CUDA_OK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        LUTVQ_REQUIRE(p != nullptr && (reinterpret_cast<uintptr_t>(p) & 255u) == 0, "scratch symbol resolved to a bad address");
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return LUTVQ_OK;
}

}  // namespace lutvq

using namespace lutvq;

extern "C" {

size_t lutvq_noalloc_scratch_bytes(void) { return kNoAllocScratchBytes; }

int lutvq_code2x8_matmat_noalloc(const void* x, const void* codes, const void* codebooks, const void* scales, const void* bias,
                                 void* y, int batch, int prob_m, int prob_k, int dtype, void* stream) {
    LUTVQ_REQUIRE(codes && codebooks, "null codes/codebook pointer");
    LUTVQ_REQUIRE(prob_m > 0 && prob_k > 0, "bad shape prob_m=%d prob_k=%d", prob_m, prob_k);
    LUTVQ_REQUIRE(prob_m <= LUTVQ_NOALLOC_MAX_ROWS,
                  "prob_m=%d exceeds the built-in scratch of the no-allocation entry point (LUTVQ_NOALLOC_MAX_ROWS = %d rows); "
                  "use lutvq_code2x8_matmat with a caller-provided workspace", prob_m, LUTVQ_NOALLOC_MAX_ROWS);
    LUTVQ_REQUIRE(prob_k % 64 == 0, "prob_k=%d must be a multiple of 64 (16-byte code rows, as cuda_kernel.cu:154 assumes)", prob_k);
    LUTVQ_REQUIRE(aligned16(codes) && aligned16(codebooks), "codes and codebook must be 16-byte aligned");
    if (dtype != LUTVQ_F16 && dtype != LUTVQ_BF16)
        return fail(LUTVQ_EUNSUPPORTED, "AQLM kernels only support float16 (0) and bfloat16 (1); got dtype %d", dtype);
    LUTVQ_REQUIRE(batch >= 0, "bad batch %d", batch);
    if (batch == 0) return LUTVQ_OK;
    LUTVQ_REQUIRE(x && y, "null x/y pointer");
    const size_t need = gemv_workspace_bytes(prob_m, prob_k / 8);
    LUTVQ_REQUIRE(need <= kNoAllocScratchBytes, "internal: workspace of %zu bytes exceeds the scratch array (%zu)", need, kNoAllocScratchBytes);
    void* ws = nullptr;
    if (int rc = noalloc_scratch(&ws)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // the array keeps the last call's contents (another shape, or a call that was cut short): the words must start at zero
    LUTVQ_CUDA_OK(cudaMemsetAsync(ws, 0, need, st));
    // one fused launch per batch row, like lutvq_code2x8_matmat (cuda_kernel.cpp:404-414); rows are dependent launches on the
    // stream, and each hands the words back zeroed, so they can share the scratch
    for (int b = 0; b < batch; ++b) {
        const void* xb = static_cast<const uint8_t*>(x) + (size_t)b * prob_k * 2;
        void* yb = static_cast<uint8_t*>(y) + (size_t)b * prob_m * 2;
        if (int rc = gemv_2x8_launch_ref(codes, xb, yb, codebooks, scales, bias, prob_m, prob_k, dtype, 0u, ws, need, st)) return rc;
    }
    return LUTVQ_OK;
}

}  // extern "C"
