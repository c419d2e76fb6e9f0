// gemv_bench.cu -- stand-alone microbenchmark of liblutvq's AQLM 2x8 LUT-GEMV through the C-ABI (no Python).
//
//   gemv_bench [--shapes 4096x11008,4096x4096,11008x4096] [--packed 0|1|2] [--tuning nw,pf] [--nopdl] [--reps R]
//              [--check] [--ncu]
//
// Per shape: a ring of distinct layers whose codes total > 2x the 126 MB L2 (so every launch streams its codes
// from HBM), captured as ONE CUDA graph of back-to-back GEMVs, replayed R times between CUDA events.
// --check compares layer 0's output with a naive fp32 dequant-dot kernel (max |err| / max |ref|).
// --ncu   runs only a handful of eager launches (for use under ncu).
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <algorithm>
#include <cmath>
#include "../include/lutvq.h"

extern "C" void lutvq_debug_set_gemv_tuning(int nw, int pf, int ctas_per_sm);
extern "C" void lutvq_debug_set_timeline(void* buf);
extern "C" void lutvq_debug_set_dependent_workers(int m);

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(2); } } while (0)
#define LK(x) do { int rc_ = (x); if (rc_ != 0) { fprintf(stderr, "lutvq error %d: %s at %s:%d\n", rc_, lutvq_last_error(), __FILE__, __LINE__); exit(3); } } while (0)

__device__ __forceinline__ uint32_t hash32(uint32_t v) {
    v ^= v >> 16; v *= 0x7feb352dU; v ^= v >> 15; v *= 0x846ca68bU; v ^= v >> 16;
    return v;
}
__global__ void fill_codes(uint32_t* p, size_t n_words, uint32_t seed) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n_words; i += (size_t)gridDim.x * blockDim.x)
        p[i] = hash32((uint32_t)i * 2654435761U + seed);
}
__global__ void fill_half(__half* p, size_t n, uint32_t seed, float scale) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        // sum of 4 uniforms ~ roughly normal, unit-ish variance
        uint32_t h = hash32((uint32_t)i + seed);
        float u = ((h & 0xFF) + ((h >> 8) & 0xFF) + ((h >> 16) & 0xFF) + (h >> 24)) / 255.f - 2.f;
        p[i] = __float2half_rn(u * 1.7f * scale);
    }
}
// naive reference: y[o] = sum_g sum_j (cb0[c0][j] + cb1[c1][j]) * x[8g+j] in fp32 (weights as one fp16 add, like the reference)
__global__ void ref_gemv(const uint8_t* codes, const __half* x, const __half* cb, float* y, int O, int G) {
    const int o = blockIdx.x;
    float acc = 0.f;
    for (int g = threadIdx.x; g < G; g += blockDim.x) {
        const int c0 = codes[((size_t)o * G + g) * 2], c1 = codes[((size_t)o * G + g) * 2 + 1];
        for (int j = 0; j < 8; ++j) {
            const float w = __half2float(cb[c0 * 8 + j]) + __half2float(cb[(256 + c1) * 8 + j]);
            acc += w * __half2float(x[g * 8 + j]);
        }
    }
    __shared__ float red[256];
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int s = blockDim.x / 2; s > 0; s >>= 1) { if (threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s]; __syncthreads(); }
    if (threadIdx.x == 0) y[o] = red[0];
}

struct Layer { void* codes; void* packed; __half* cb; __half* scales; __half* x; __half* y; void* ws; };

int main(int argc, char** argv) {
    std::string shapes = "4096x4096,4096x11008,11008x4096";
    int packed = 2, reps = 40, nw = 0, pf = 0, pa = 0;
    bool nopdl = false, check = false, ncu = false, timeline = false, xready = false;
    bool noscales = false;
    int wsring = 8;   // workspaces rotate through a ring of this many regions of ONE allocation (0: one allocation per layer)
    unsigned dbg = 0;
    double peak = 6536.7;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto next = [&]() { return std::string(i + 1 < argc ? argv[++i] : ""); };
        if (a == "--shapes") shapes = next();
        else if (a == "--packed") packed = atoi(next().c_str());
        else if (a == "--reps") reps = atoi(next().c_str());
        else if (a == "--tuning") { std::string t = next(); sscanf(t.c_str(), "%d,%d,%d", &nw, &pf, &pa); }
        else if (a == "--nopdl") nopdl = true;
        else if (a == "--xready") xready = true;
        else if (a == "--noscales") noscales = true;
        else if (a == "--wsring") wsring = atoi(next().c_str());
        else if (a == "--dbg") dbg = (unsigned)strtoul(next().c_str(), nullptr, 0);
        else if (a == "--check") check = true;
        else if (a == "--ncu") ncu = true;
        else if (a == "--timeline") timeline = true;
        else if (a == "--peak") peak = atof(next().c_str());
        else if (a == "--depworkers") lutvq_debug_set_dependent_workers(atoi(next().c_str()));
    }
    if (nw) lutvq_debug_set_gemv_tuning(nw, pf, pa);
    int sms = 0;
    LK(lutvq_device_info(&sms, nullptr, nullptr, nullptr));
    cudaStream_t st;
    CK(cudaStreamCreate(&st));
    const unsigned flags = (nopdl ? LUTVQ_FLAG_NO_PDL : 0u) | (xready ? LUTVQ_FLAG_X_READY : 0u) | dbg;

    size_t pos = 0;
    while (pos < shapes.size()) {
        size_t comma = shapes.find(',', pos);
        std::string sh = shapes.substr(pos, comma == std::string::npos ? std::string::npos : comma - pos);
        pos = comma == std::string::npos ? shapes.size() : comma + 1;
        int I = 0, O = 0;
        if (sscanf(sh.c_str(), "%dx%d", &I, &O) != 2) { fprintf(stderr, "bad shape %s\n", sh.c_str()); return 1; }
        const int G = I / 8;
        const size_t code_bytes = (size_t)O * G * 2;
        int nbuf = (int)std::max<size_t>(8, (size_t)(280e6 / code_bytes) + 1);
        if (ncu) nbuf = 4;
        if (wsring > 0) nbuf = (nbuf + wsring - 1) / wsring * wsring;
        const size_t pk_bytes = lutvq_code2x8_packed_bytes(O, I);
        const size_t ws_bytes = lutvq_workspace_bytes(O, I);
        std::vector<Layer> L(nbuf);
        // small operands live in ONE slab (a handful of pages, TLB-warm like a real model's activations);
        // only the weights (codes) get an allocation per layer
        auto al = [](size_t v) { return (v + 255) / 256 * 256; };
        const size_t small_per = al(2 * 256 * 8 * 2) + al((size_t)O * 2) * 2 + al((size_t)I * 2);
        const int n_ws = wsring > 0 ? wsring : nbuf;
        uint8_t* slab;
        CK(cudaMalloc(&slab, small_per * nbuf + al(ws_bytes) * n_ws));
        CK(cudaMemset(slab, 0, small_per * nbuf + al(ws_bytes) * n_ws));
        for (int b = 0; b < nbuf; ++b) {
            CK(cudaMalloc(&L[b].codes, code_bytes));
            CK(cudaMalloc(&L[b].packed, pk_bytes));
            uint8_t* q = slab + small_per * b;
            L[b].cb = (__half*)q;      q += al(2 * 256 * 8 * 2);
            L[b].scales = (__half*)q;  q += al((size_t)O * 2);
            L[b].y = (__half*)q;       q += al((size_t)O * 2);
            L[b].x = (__half*)q;
            L[b].ws = slab + small_per * nbuf + al(ws_bytes) * (b % n_ws);
            fill_codes<<<592, 256, 0, st>>>((uint32_t*)L[b].codes, code_bytes / 4, 1000u * b + 7u);
            fill_half<<<16, 256, 0, st>>>(L[b].cb, 2 * 256 * 8, 31u * b + 1u, 1.f);
            fill_half<<<16, 256, 0, st>>>(L[b].x, I, 77u * b + 5u, 1.f);
            fill_half<<<16, 256, 0, st>>>(L[b].scales, O, 13u * b + 3u, 0.f);   // zeros, then set to 1 below
            LK(lutvq_code2x8_prepack(L[b].codes, L[b].packed, O, I, st));
        }
        CK(cudaStreamSynchronize(st));
        {   // scales = 1 like benchmark_rq_gemm_gpu.py:97
            std::vector<__half> ones(O, __float2half(1.f));
            for (int b = 0; b < nbuf; ++b) CK(cudaMemcpy(L[b].scales, ones.data(), (size_t)O * 2, cudaMemcpyHostToDevice));
        }
        const double alg_bytes = (double)code_bytes + 8192 + 2.0 * I + 4.0 * O;

        for (int mode = 0; mode < 2; ++mode) {
            if (packed != 2 && packed != mode) continue;
            auto launch = [&](int b) {
                if (mode) LK(lutvq_code2x8_matmat_packed(L[b].x, L[b].packed, L[b].cb, noscales ? nullptr : L[b].scales, nullptr, L[b].y, 1, O, I, LUTVQ_F16, flags, L[b].ws, ws_bytes, st));
                else      LK(lutvq_code2x8_matmat_ex(L[b].x, L[b].codes, L[b].cb, L[b].scales, nullptr, L[b].y, 1, O, I, LUTVQ_F16, flags, L[b].ws, ws_bytes, st));
            };
            if (check) {
                float* yref;
                CK(cudaMalloc(&yref, (size_t)O * 4));
                ref_gemv<<<O, 256, 0, st>>>((const uint8_t*)L[0].codes, L[0].x, L[0].cb, yref, O, G);
                CK(cudaMemsetAsync(L[0].y, 0xFF, (size_t)O * 2, st));
                launch(0);
                for (int b = 1; b < std::min(nbuf, 9); ++b) launch(b);   // overlapping successors (X_READY: distinct workspaces)
                if (!xready) launch(0);                                  // again on the same workspace: exercises the epoch hand-over
                CK(cudaStreamSynchronize(st));
                std::vector<float> r(O);
                std::vector<__half> y(O);
                CK(cudaMemcpy(r.data(), yref, (size_t)O * 4, cudaMemcpyDeviceToHost));
                CK(cudaMemcpy(y.data(), L[0].y, (size_t)O * 2, cudaMemcpyDeviceToHost));
                double maxref = 0, maxerr = 0;
                int bad = -1;
                for (int o = 0; o < O; ++o) {
                    maxref = std::max(maxref, (double)fabsf(r[o]));
                    double e = fabs((double)__half2float(y[o]) - r[o]);
                    if (!(e <= maxerr)) { maxerr = e; bad = o; }
                }
                printf("check %s mode=%s: max|err|/max|ref| = %.3e (row %d: got %f want %f)\n", sh.c_str(), mode ? "packed" : "reference-layout",
                       maxerr / maxref, bad, bad >= 0 ? __half2float(y[bad]) : 0.f, bad >= 0 ? r[bad] : 0.f);
                CK(cudaFree(yref));
            }
            if (timeline) {
                // per-CTA %globaltimer stamps of 8 back-to-back launches inside one graph (debug hook of the library)
                const int NL = 8;
                unsigned long long* tl;
                const size_t tl_bytes = (size_t)64 * 256 * 16 * 8;
                CK(cudaMalloc(&tl, tl_bytes));
                CK(cudaMemset(tl, 0, tl_bytes));
                for (int b = 0; b < nbuf; ++b) launch(b);        // warm
                lutvq_debug_set_timeline(tl);
                cudaGraph_t g; cudaGraphExec_t ex;
                CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeGlobal));
                for (int b = 0; b < NL; ++b) launch(b % nbuf);
                CK(cudaStreamEndCapture(st, &g));
                CK(cudaGraphInstantiate(&ex, g, 0));
                lutvq_debug_set_timeline(nullptr);
                for (int i = 0; i < 3; ++i) CK(cudaGraphLaunch(ex, st));
                CK(cudaStreamSynchronize(st));
                std::vector<unsigned long long> h(tl_bytes / 8);
                CK(cudaMemcpy(h.data(), tl, tl_bytes, cudaMemcpyDeviceToHost));
                unsigned long long t00 = ~0ull;
                for (int c = 0; c < 256; ++c) if (h[c * 16]) t00 = std::min(t00, h[c * 16]);
                printf("timeline %s mode=%s (ns relative to first CTA entry of launch 0; per stamp min/median/max over CTAs)\n", sh.c_str(), mode ? "packed" : "reference");
                const char* names[15] = {"entry", "prologue-issued", "pdl-wait-done", "lut-built", "stream-done", "reduce-done", "exit", "", "range-done", "loads-issued", "xnorm-done", "scale-done", "mma-done", "poll-issue", "poll1-done"};
                const int order[14] = {0, 8, 9, 1, 2, 10, 11, 12, 3, 4, 13, 14, 5, 6};
                for (int l = 0; l < NL; ++l) {
                    for (int oi = 0; oi < 14; ++oi) {
                        const int s = order[oi];
                        std::vector<long long> v;
                        for (int c = 0; c < 256; ++c) { unsigned long long t = h[((size_t)l * 256 + c) * 16 + s]; if (t) v.push_back((long long)(t - t00)); }
                        if (v.empty()) continue;
                        std::sort(v.begin(), v.end());
                        printf("  L%d %-16s n=%3zu min %7lld med %7lld max %7lld\n", l, names[s], v.size(), v.front(), v[v.size() / 2], v.back());
                    }
                    {
                        std::vector<long long> v;
                        for (int c = 0; c < 256; ++c) { unsigned long long t = h[((size_t)l * 256 + c) * 16 + 7]; if (t) v.push_back((long long)t); }
                        std::sort(v.begin(), v.end());
                        if (!v.empty()) printf("  L%d poll rounds      n=%3zu min %7lld med %7lld max %7lld\n", l, v.size(), v.front(), v[v.size() / 2], v.back());
                    }
                }
                {   // per-CTA detail of launch 5, relative to its own pdl-wait release
                    const int l = 5;
                    unsigned long long w0 = ~0ull;
                    for (int c = 0; c < 256; ++c) { unsigned long long t = h[((size_t)l * 256 + c) * 16 + 2]; if (t) w0 = std::min(w0, t); }
                    printf("  per-CTA L5 (cta smid | entry prologue wait lut stream reduce exit, ns rel. to first wait release)\n");
                    for (int c = 0; c < 256; ++c) {
                        const unsigned long long* e = &h[((size_t)l * 256 + c) * 16];
                        if (!e[0]) continue;
                        if (c >= 12) break;
                        printf("   cta %3d sm %3llu |", c, e[15] - 1);
                        for (int s = 0; s < 7; ++s) printf(" %7lld", (long long)(e[s] - w0));
                        printf("\n");
                    }
                }
                CK(cudaGraphExecDestroy(ex)); CK(cudaGraphDestroy(g)); CK(cudaFree(tl));
                continue;
            }
            if (ncu) {
                for (int b = 0; b < nbuf; ++b) launch(b);
                CK(cudaStreamSynchronize(st));
                continue;
            }
            cudaGraph_t graph;
            cudaGraphExec_t exec;
            CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeGlobal));
            for (int b = 0; b < nbuf; ++b) launch(b);
            CK(cudaStreamEndCapture(st, &graph));
            CK(cudaGraphInstantiate(&exec, graph, 0));
            for (int i = 0; i < 5; ++i) CK(cudaGraphLaunch(exec, st));
            CK(cudaStreamSynchronize(st));
            cudaEvent_t e0, e1;
            CK(cudaEventCreate(&e0));
            CK(cudaEventCreate(&e1));
            CK(cudaEventRecord(e0, st));
            for (int i = 0; i < reps; ++i) CK(cudaGraphLaunch(exec, st));
            CK(cudaEventRecord(e1, st));
            CK(cudaStreamSynchronize(st));
            float ms = 0;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            const double us = ms * 1e3 / ((double)reps * nbuf);
            const double gbps = alg_bytes / (us * 1e-6) / 1e9;
            printf("{\"shape\": \"%s\", \"layout\": \"%s\", \"pdl\": %d, \"xready\": %d, \"dbg\": %u, \"tuning\": \"%d,%d,%d\", \"us\": %.3f, \"gbps\": %.1f, \"frac\": %.4f, \"ring\": %d, \"sms\": %d}\n",
                   sh.c_str(), mode ? "packed" : "reference", nopdl ? 0 : 1, xready ? 1 : 0, dbg, nw, pf, pa, us, gbps, gbps / peak, nbuf, sms);
            CK(cudaGraphExecDestroy(exec));
            CK(cudaGraphDestroy(graph));
        }
        for (int b = 0; b < nbuf; ++b) {
            cudaFree(L[b].codes); cudaFree(L[b].packed);
        }
        cudaFree(slab);
    }
    return 0;
}
