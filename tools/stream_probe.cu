// stream_probe.cu -- what can ONE short kernel per 11.27 MB cold buffer stream on this GPU?  (experiment, not product)
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/stream_probe tools/stream_probe.cu
// A ring of NBUF distinct buffers (> 2x L2) is read by a graph of NBUF launches (programmatic dependent launch, like the
// GEMV ring); methods: plain 128-bit loads with U in flight per thread; L2 bulk prefetch of the whole share followed by the
// loads; bulk async copies into a per-warp shared-memory ring.  Prints achieved GB/s per method / CTAs per SM.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint4 ldg_stream(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// mode 0: loads only; mode 1: bulk L2 prefetch of the CTA's share first; U loads in flight per thread
template <int U>
__global__ void __launch_bounds__(256) k_ldg(const uint8_t* __restrict__ buf, size_t bytes, uint32_t* out, int mode, int pdl_early) {
    const size_t per = (bytes / gridDim.x) & ~(size_t)4095;
    const uint8_t* base = buf + per * blockIdx.x;
    if (mode == 1 && (threadIdx.x & 31) == 0) {
        const int warp = threadIdx.x >> 5;
        for (size_t off = (size_t)warp * 4096; off < per; off += 8 * 4096)
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(base + off), "r"(4096u) : "memory");
    }
    if (pdl_early) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    uint32_t acc = 0;
    const size_t n16 = per / 16;
    for (size_t i = threadIdx.x; i < n16; i += 256 * U) {
        uint4 v[U];
        #pragma unroll
        for (int u = 0; u < U; ++u) v[u] = (i + (size_t)u * 256 < n16) ? ldg_stream(base + (i + (size_t)u * 256) * 16) : make_uint4(0, 0, 0, 0);
        #pragma unroll
        for (int u = 0; u < U; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (acc == 0x12345678u) out[blockIdx.x] = acc;
}

// per-warp ring of S slots of 4 KiB filled by cp.async.bulk
template <int S>
__global__ void __launch_bounds__(256) k_bulk(const uint8_t* __restrict__ buf, size_t bytes, uint32_t* out, int pdl_early) {
    extern __shared__ __align__(1024) uint8_t sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t per = (bytes / gridDim.x) & ~(size_t)4095;
    const uint8_t* base = buf + per * blockIdx.x;
    const int n = (int)(per / 4096);                               // 4 KiB items of this CTA; warp w takes w, w+8, ...
    const uint32_t slot0 = smem_u32(sm) + warp * S * 4096, mb0 = smem_u32(sm) + 8 * S * 4096 + warp * S * 8;
    auto issue = [&](int item, int s) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(mb0 + s * 8), "r"(4096u) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     :: "r"(slot0 + s * 4096), "l"(base + (size_t)item * 4096), "r"(4096u), "r"(mb0 + s * 8) : "memory");
    };
    if (lane == 0) {
        for (int s = 0; s < S; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(mb0 + s * 8), "r"(1u) : "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int s = 0; s < S; ++s) if (warp + s * 8 < n) issue(warp + s * 8, s);
    }
    __syncwarp();
    if (pdl_early) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    uint32_t acc = 0;
    int k = 0;
    for (int item = warp; item < n; item += 8, ++k) {
        const int s = k % S;
        const uint32_t parity = (k / S) & 1;
        asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}\n" :: "r"(mb0 + s * 8), "r"(parity) : "memory");
        #pragma unroll
        for (int j = 0; j < 8; ++j) {
            uint4 v;
            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(slot0 + s * 4096 + j * 512 + lane * 16) : "memory");
            acc ^= v.x ^ v.y ^ v.z ^ v.w;
        }
        __syncwarp();
        if (lane == 0 && item + S * 8 < n) issue(item + S * 8, s);
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (acc == 0x12345678u) out[blockIdx.x] = acc;
}

int main() {
    const size_t bytes = 11272192;                                  // codes of one 4096 -> 11008 GEMV
    const int NBUF = 64;
    std::vector<uint8_t*> bufs(NBUF);
    for (auto& b : bufs) { CK(cudaMalloc(&b, bytes)); CK(cudaMemset(b, 1, bytes)); }
    uint32_t* out;
    CK(cudaMalloc(&out, 4096 * 4));
    cudaStream_t st;
    CK(cudaStreamCreate(&st));
    auto run = [&](const char* name, auto launch) {
        cudaGraph_t g; cudaGraphExec_t ex;
        CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeGlobal));
        for (int b = 0; b < NBUF; ++b) launch(bufs[b]);
        CK(cudaStreamEndCapture(st, &g));
        CK(cudaGraphInstantiate(&ex, g, 0));
        for (int i = 0; i < 3; ++i) CK(cudaGraphLaunch(ex, st));
        CK(cudaStreamSynchronize(st));
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        CK(cudaEventRecord(e0, st));
        const int reps = 20;
        for (int i = 0; i < reps; ++i) CK(cudaGraphLaunch(ex, st));
        CK(cudaEventRecord(e1, st));
        CK(cudaStreamSynchronize(st));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        const double us = ms * 1e3 / (reps * NBUF);
        printf("{\"method\": \"%s\", \"us_per_buffer\": %.3f, \"gbps\": %.1f}\n", name, us, bytes / us / 1e3);
        fflush(stdout);
        CK(cudaGraphExecDestroy(ex)); CK(cudaGraphDestroy(g));
    };
    auto cfg_launch = [&](auto kern, int grid, size_t smem, bool pdl, auto... args) {
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(grid); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = smem; cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
        CK(cudaLaunchKernelEx(&cfg, kern, args...));
    };
    char name[128];
    for (int pdl = 0; pdl < 2; ++pdl)
        for (int cps = 1; cps <= 4; cps *= 2) {
            for (int mode = 0; mode < 2; ++mode) {
                snprintf(name, sizeof name, "ldg U=4 mode=%d ctas/sm=%d pdl=%d", mode, cps, pdl);
                run(name, [&](uint8_t* b) { cfg_launch(k_ldg<4>, 148 * cps, 0, pdl, (const uint8_t*)b, bytes, out, mode, pdl); });
                snprintf(name, sizeof name, "ldg U=8 mode=%d ctas/sm=%d pdl=%d", mode, cps, pdl);
                run(name, [&](uint8_t* b) { cfg_launch(k_ldg<8>, 148 * cps, 0, pdl, (const uint8_t*)b, bytes, out, mode, pdl); });
            }
            if (cps <= 2) {
                CK(cudaFuncSetAttribute(k_bulk<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 2 * 4096 + 256));
                CK(cudaFuncSetAttribute(k_bulk<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 4 * 4096 + 256));
                snprintf(name, sizeof name, "bulk 4KiB S=2 ctas/sm=%d pdl=%d", cps, pdl);
                run(name, [&](uint8_t* b) { cfg_launch(k_bulk<2>, 148 * cps, (size_t)8 * 2 * 4096 + 256, pdl, (const uint8_t*)b, bytes, out, pdl); });
                snprintf(name, sizeof name, "bulk 4KiB S=4 ctas/sm=%d pdl=%d", cps, pdl);
                run(name, [&](uint8_t* b) { cfg_launch(k_bulk<4>, 148 * cps, (size_t)8 * 4 * 4096 + 256, pdl, (const uint8_t*)b, bytes, out, pdl); });
            }
        }
    return 0;
}
