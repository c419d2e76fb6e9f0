// common.cuh -- shared helpers for liblutvq.so (error reporting, dtype traits, small PTX wrappers)
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/lutvq.h"

namespace lutvq {

// ---- thread-local error string (lutvq_last_error) ----
char* err_buf();
int   fail(int code, const char* fmt, ...);

#define LUTVQ_CUDA_OK(expr)                                                                     \
    do {                                                                                        \
        cudaError_t _e = (expr);                                                                \
        if (_e != cudaSuccess)                                                                  \
            return ::lutvq::fail(LUTVQ_ECUDA, "%s failed: %s (%s:%d)", #expr,                   \
                                 cudaGetErrorString(_e), __FILE__, __LINE__);                   \
    } while (0)

#define LUTVQ_REQUIRE(cond, ...)                                                                \
    do {                                                                                        \
        if (!(cond)) return ::lutvq::fail(LUTVQ_EINVAL, __VA_ARGS__);                           \
    } while (0)

struct DeviceInfo {
    int device = -1, sm_count = 0, cc_major = 0, cc_minor = 0;
    size_t smem_optin = 0;
};
// Cached per device; queries the CURRENT device (not device 0 as cuda_kernel.cu:486,497 does).
int get_device_info(DeviceInfo* out);

// Peer targets of a sharded GEMV (shard.cu -> gemv_lut.cu): where each finished output element is ALSO stored, and
// which flags the launch's last CTA publishes.  Pointers are device addresses valid on the launching GPU (local or
// peer memory mapped through CUDA IPC / peer access).
struct GemvPush {
    int n = 0;
    void* y[8] = {};             // peer r's gather buffer, already offset to this rank's first row
    unsigned* flag[8] = {};      // peer r's epoch flag for (slot, this rank)
    unsigned* sent = nullptr;    // local pushes-completed counter of the slot
};

// ---- "LL" gather words of the sharded decode (shard.cu) ----
// One 32-bit word per output row: {value: fp16/bf16 bits | epoch16 << 16}, written with ONE store into every rank's gather
// slot (the own rank's too).  A word is valid once its upper half equals the expected epoch, so a consumer needs neither a
// flag nor a fence nor a separate exchange launch: it polls the very words it is about to use (local memory, written remotely
// over NVLink) and a 32-bit store is atomic.  epoch16 = (tick * 1024 + seq) mod 65535 + 1 is never 0 (= "never written").
struct GemvLL {
    int nranks = 0;                      // 0: no LL output
    uint8_t* peer_base[8] = {};          // rank r's LL area as mapped here (peer_base[rank] = the local one)
    size_t out_off[3] = {};              // per linear of a group: byte offset of (slot, this rank's first row) inside the LL area
    unsigned out_seq[3] = {};            // per linear: sequence number of this exchange inside the step
    const unsigned* tick = nullptr;      // device counter of steps (lutvq_shard_ll_tick), read after the dependency wait
    const uint32_t* x_ll = nullptr;      // input activations as LL words of an earlier exchange (local slot), or null: plain x
    unsigned x_seq = 0;
    const uint32_t* gate_ll = nullptr;   // SwiGLU gate as LL words, or null
    unsigned gate_seq = 0;
    unsigned* error = nullptr;           // local error word: set when a bounded poll gives up
};
__device__ __forceinline__ unsigned ll_epoch16(unsigned tick, unsigned seq) { return (tick * 1024u + seq) % 65535u + 1u; }
__device__ __forceinline__ uint4 ll_ld_v4(const uint32_t* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void ll_st(uint8_t* base, size_t off, uint32_t word) {
    asm volatile("st.volatile.global.u32 [%0], %1;" :: "l"(base + off), "r"(word) : "memory");
}
// 8 consecutive LL words (one group of 8 activations) -> the 8 values packed as 16 bytes; polls until all 8 carry `ep`.
// Bounded: ~2^21 polls (seconds), then the error word is set and whatever is there is returned instead of hanging the GPU.
__device__ __forceinline__ uint4 ll_load8(const uint32_t* w, unsigned ep, unsigned* error) {
    uint4 a, b;
    for (unsigned spin = 0;; ++spin) {
        a = ll_ld_v4(w);
        b = ll_ld_v4(w + 4);
        const unsigned bad = ((a.x >> 16) ^ ep) | ((a.y >> 16) ^ ep) | ((a.z >> 16) ^ ep) | ((a.w >> 16) ^ ep) |
                             ((b.x >> 16) ^ ep) | ((b.y >> 16) ^ ep) | ((b.z >> 16) ^ ep) | ((b.w >> 16) ^ ep);
        if (bad == 0u) break;
        if (spin > (1u << 21)) { if (error) *error = 0xFFFFFFFDu; break; }
        if (spin > 64u) __nanosleep(100);
    }
    uint4 v;
    v.x = (a.x & 0xFFFFu) | (a.y << 16); v.y = (a.z & 0xFFFFu) | (a.w << 16);
    v.z = (b.x & 0xFFFFu) | (b.y << 16); v.w = (b.z & 0xFFFFu) | (b.w << 16);
    return v;
}
// one LL word -> its 16 value bits
__device__ __forceinline__ unsigned short ll_load1(const uint32_t* w, unsigned ep, unsigned* error) {
    uint32_t v;
    for (unsigned spin = 0;; ++spin) {
        asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(w) : "memory");
        if ((v >> 16) == ep) break;
        if (spin > (1u << 21)) { if (error) *error = 0xFFFFFFFDu; break; }
        if (spin > 64u) __nanosleep(100);
    }
    return (unsigned short)(v & 0xFFFFu);
}

// one linear of a grouped GEMV launch (gemv_lut.cu gemv_2x8_launch_group)
// one linear of a grouped GEMV launch.  x_off: byte offset of THIS item's activation row from the launch's x pointer (0 for the
// linears of a layer that share one input; b * prob_k * 2 for the rows of a multi-row call, which share the codes instead)
struct GemvItem { const void* codes; const void* codebook; void* y; const void* scales; const void* bias; int O; size_t x_off = 0; };

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---- dtype traits: storage type <-> float, one RN op in the storage type ----
template <int DT> struct Dtype;
template <> struct Dtype<LUTVQ_F16> {
    using T = __half;
    using T2 = __half2;
    static __device__ __forceinline__ float to_f(T v) { return __half2float(v); }
    static __device__ __forceinline__ T from_f(float v) { return __float2half_rn(v); }
    static __device__ __forceinline__ T mul(T a, T b) { return __hmul(a, b); }
    static __device__ __forceinline__ T add(T a, T b) { return __hadd(a, b); }
    static __device__ __forceinline__ T2 add2(T2 a, T2 b) { return __hadd2(a, b); }
    static __device__ __forceinline__ T2 mul2(T2 a, T2 b) { return __hmul2(a, b); }
    static __device__ __forceinline__ T2 dup(T a) { return __half2half2(a); }
};
template <> struct Dtype<LUTVQ_BF16> {
    using T = __nv_bfloat16;
    using T2 = __nv_bfloat162;
    static __device__ __forceinline__ float to_f(T v) { return __bfloat162float(v); }
    static __device__ __forceinline__ T from_f(float v) { return __float2bfloat16_rn(v); }
    static __device__ __forceinline__ T mul(T a, T b) { return __hmul(a, b); }
    static __device__ __forceinline__ T add(T a, T b) { return __hadd(a, b); }
    static __device__ __forceinline__ T2 add2(T2 a, T2 b) { return __hadd2(a, b); }
    static __device__ __forceinline__ T2 mul2(T2 a, T2 b) { return __hmul2(a, b); }
    static __device__ __forceinline__ T2 dup(T a) { return __bfloat162bfloat162(a); }
};

// ---- PTX wrappers ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
// streaming 64-bit global load that does not allocate in L1 (codes are read exactly once)
__device__ __forceinline__ uint2 ldg_stream_u64(const void* p) {
    uint2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
    return r;
}
__device__ __forceinline__ uint4 ldg_stream_u128(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
// L2-coherent load (bypasses L1): for data written by other CTAs of the same grid
__device__ __forceinline__ float ldcg_f32(const float* p) {
    float v;
    asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ uint32_t ldcg_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
// programmatic dependent launch (no-ops when the launch did not opt in)
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

}  // namespace lutvq
