// decode_glue.cu -- the small non-GEMV steps of a Llama-shaped decode step (batch 1), so that the whole token is a
// chain of OUR launches inside one CUDA graph: residual add + RMSNorm, RoPE + KV append + single-query attention,
// SwiGLU.  The reference gets these from HF transformers (benchmark/llm/generate_benchmark.py); they are not LUT-VQ
// work and are outside the graded hot path, but the decode tok/s metric (BASELINE.json configs[3], [4]) runs through them.
// fp16 only (the decode benchmark's dtype).  All kernels honour programmatic dependent launch like the GEMV.
#include "common.cuh"

namespace lutvq {

// Optional LL inputs (common.cuh "LL gather words"): operand k of a glue kernel comes from the local gather slot w[k] of an earlier
// sharded GEMV (exchange sequence number seq[k]) instead of a plain buffer; the kernel polls exactly the words it consumes.
struct LLIn {
    const uint32_t* w[3] = {nullptr, nullptr, nullptr};
    unsigned seq[3] = {0, 0, 0};
    const unsigned* tick = nullptr;
    unsigned* error = nullptr;
    // optional LL OUTPUT: the result also goes out as {value | epoch} words into a LOCAL gather slot, so that the GEMV that
    // consumes it can poll for it instead of waiting for this kernel to complete (DESIGN 4.5 "dataflow launches")
    uint32_t* out = nullptr;
    unsigned out_seq = 0;
};
__device__ __forceinline__ void ll_store8(uint32_t* dst, const uint4& halves, unsigned ep) {
    const uint32_t e = ep << 16;
    const uint4 lo = make_uint4((halves.x & 0xffffu) | e, (halves.x >> 16) | e, (halves.y & 0xffffu) | e, (halves.y >> 16) | e);
    const uint4 hi = make_uint4((halves.z & 0xffffu) | e, (halves.z >> 16) | e, (halves.w & 0xffffu) | e, (halves.w >> 16) | e);
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" :: "l"(dst), "r"(lo.x), "r"(lo.y), "r"(lo.z), "r"(lo.w) : "memory");
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" :: "l"(dst + 4), "r"(hi.x), "r"(hi.y), "r"(hi.z), "r"(hi.w) : "memory");
}
__device__ __forceinline__ __half ll_half(const LLIn& ll, int k, unsigned tick, int i) {
    const unsigned short b = ll_load1(ll.w[k] + i, ll_epoch16(tick, ll.seq[k]), ll.error);
    return *reinterpret_cast<const __half*>(&b);
}

__device__ __forceinline__ float warp_sum(float v) {
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// resid (+= delta, if given);  y = resid * rsqrt(mean(resid^2) + eps) * w.   One CTA, n <= 1024 * 16.
__global__ void __launch_bounds__(1024) add_rmsnorm_kernel(__half* resid, const __half* delta, const __half* w, __half* y, int n, float eps, LLIn ll) {
    pdl_launch_dependents();
    pdl_wait();
    const bool has_delta = delta != nullptr || ll.w[0] != nullptr;
    const unsigned tick = ll.w[0] ? *ll.tick : 0u;
    __shared__ float red[32];
    float v[16];
    float ss = 0.f;
    int cnt = 0;
    for (int i = threadIdx.x; i < n; i += 1024, ++cnt) {
        float a = __half2float(resid[i]);
        if (has_delta) { a = __half2float(__float2half_rn(a + __half2float(ll.w[0] ? ll_half(ll, 0, tick, i) : delta[i]))); resid[i] = __float2half_rn(a); }
        v[cnt] = a;
        ss += a * a;
    }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
    #pragma unroll
    for (int i = 0; i < 32; ++i) tot += red[i];
    const float r = rsqrtf(tot / (float)n + eps);
    cnt = 0;
    const unsigned ep_out = ll.out ? ll_epoch16(*ll.tick, ll.out_seq) : 0u;
    for (int i = threadIdx.x; i < n; i += 1024, ++cnt) {
        const __half o = __float2half_rn(v[cnt] * r * __half2float(w[i]));
        y[i] = o;
        if (ll.out) asm volatile("st.volatile.global.u32 [%0], %1;" :: "l"(ll.out + i), "r"((uint32_t)__half_as_ushort(o) | (ep_out << 16)) : "memory");
    }
}

// The same for n % 8 == 0, n <= 8192 (every Llama hidden size): one 16-byte load per operand and thread, all of them
// in flight at once (the norm weights before griddepcontrol.wait: no kernel of the chain writes them), values kept
// in registers -- the generic kernel above pays three dependent L2 round trips and a local-memory array.
__global__ void __launch_bounds__(1024) add_rmsnorm_vec8_kernel(__half* resid, const __half* delta, const __half* w, __half* y, int n, float eps, LLIn ll) {
    pdl_launch_dependents();
    __shared__ float red[32];
    const int i0 = threadIdx.x * 8;
    const bool act = i0 < n;
    uint4 wv = make_uint4(0u, 0u, 0u, 0u), rv = wv, dv = wv;
    if (act) wv = __ldg(reinterpret_cast<const uint4*>(w + i0));
    pdl_wait();
    const bool has_delta = delta != nullptr || ll.w[0] != nullptr;
    if (act) {
        rv = __ldcg(reinterpret_cast<const uint4*>(resid + i0));
        if (ll.w[0]) dv = ll_load8(ll.w[0] + i0, ll_epoch16(*ll.tick, ll.seq[0]), ll.error);
        else if (delta) dv = __ldcg(reinterpret_cast<const uint4*>(delta + i0));
    }
    const __half* rh = reinterpret_cast<const __half*>(&rv);
    const __half* dh = reinterpret_cast<const __half*>(&dv);
    const __half* wh = reinterpret_cast<const __half*>(&wv);
    float a[8], ss = 0.f;
    #pragma unroll
    for (int j = 0; j < 8; ++j) {
        a[j] = __half2float(rh[j]);
        if (has_delta) a[j] = __half2float(__float2half_rn(a[j] + __half2float(dh[j])));
        ss += a[j] * a[j];
    }
    if (act && has_delta) {
        uint4 o;
        __half* oh = reinterpret_cast<__half*>(&o);
        #pragma unroll
        for (int j = 0; j < 8; ++j) oh[j] = __float2half_rn(a[j]);
        *reinterpret_cast<uint4*>(resid + i0) = o;
    }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
    const int nw = blockDim.x >> 5;
    for (int i = 0; i < nw; ++i) tot += red[i];
    const float r = rsqrtf(tot / (float)n + eps);
    if (act) {
        uint4 o;
        __half* oh = reinterpret_cast<__half*>(&o);
        #pragma unroll
        for (int j = 0; j < 8; ++j) oh[j] = __float2half_rn(a[j] * r * __half2float(wh[j]));
        *reinterpret_cast<uint4*>(y + i0) = o;
        if (ll.out) ll_store8(ll.out + i0, o, ll_epoch16(*ll.tick, ll.out_seq));
    }
}

__global__ void silu_mul_kernel(const __half* g, const __half* u, __half* y, int n) {
    pdl_launch_dependents();
    pdl_wait();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        const float a = __half2float(g[i]);
        y[i] = __float2half_rn(a / (1.f + __expf(-a)) * __half2float(u[i]));
    }
}

// One CTA (8 warps) per query head.  RoPE (HF rotate_half convention) on q and on the new k, the new k/v appended to the
// cache at *pos by the first head of each KV group, then softmax(q.K^T / sqrt(d)) V over [0, *pos].  Positions are dealt
// to the warps round-robin in both passes (lane = 4 head dims, 8-byte coalesced cache reads, 4 positions in flight per
// warp), the 8 partial outputs are added through shared memory: the first version walked the positions on one thread
// per head dim, one dependent L2 round trip after the other (+1.4 us per layer for every 32 positions).
constexpr int kHeadDim = 128;
constexpr int kAttnWarps = 8;
__global__ void __launch_bounds__(kAttnWarps * 32) attn_decode_kernel(const __half* q, const __half* k_new, const __half* v_new,
                                                                     __half* kcache, __half* vcache, const int* pos_ptr, __half* out,
                                                                     int n_heads, int n_kv, int max_seq, float theta, LLIn ll) {
    pdl_launch_dependents();
    extern __shared__ float sc[];                 // [max_seq] scores, then [kAttnWarps][kHeadDim] partial outputs
    float* part = sc + max_seq;
    __shared__ float qs[kHeadDim], ks[kHeadDim], vs[kHeadDim], red[kAttnWarps];
    const int h = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int kvh = h / (n_heads / n_kv);
    // *pos and the cache rows below it were written by earlier TOKENS (complete before this token's first launch), not
    // by a kernel this launch may overlap with: the first kPre positions of every warp are fetched before
    // griddepcontrol.wait, i.e. while the q/k/v GEMV is still running.
    const int pos = *pos_ptr;
    const uint2* kbase = reinterpret_cast<const uint2*>(kcache + (size_t)kvh * max_seq * kHeadDim) + lane;   // + t * 32
    const uint2* vbase = reinterpret_cast<const uint2*>(vcache + (size_t)kvh * max_seq * kHeadDim) + lane;
    constexpr int kPre = 16;                      // positions per warp held in registers (covers pos <= 128)
    uint2 kpre[kPre], vpre[kPre];
    #pragma unroll
    for (int u = 0; u < kPre; ++u) {
        const int t = warp + u * kAttnWarps;
        kpre[u] = t < pos ? __ldcg(kbase + (size_t)t * 32) : make_uint2(0u, 0u);
        vpre[u] = t < pos ? __ldcg(vbase + (size_t)t * 32) : make_uint2(0u, 0u);
    }
    pdl_wait();
    if (tid < kHeadDim) {
        const int d = tid, half_d = kHeadDim / 2, i = d % half_d;
        const float freq = __powf(theta, -2.f * (float)i / (float)kHeadDim);
        float sn, cs;
        __sincosf((float)pos * freq, &sn, &cs);
        const int dr = d < half_d ? d + half_d : d - half_d;          // rotate_half partner
        const unsigned tick = ll.w[0] ? *ll.tick : 0u;
        auto ld = [&](int k, const __half* plain, int i) { return ll.w[k] ? ll_half(ll, k, tick, i) : plain[i]; };
        const float a = __half2float(ld(0, q, h * kHeadDim + d)), b = __half2float(ld(0, q, h * kHeadDim + dr));
        qs[d] = (d < half_d) ? a * cs - b * sn : a * cs + b * sn;
        const float c = __half2float(ld(1, k_new, kvh * kHeadDim + d)), e = __half2float(ld(1, k_new, kvh * kHeadDim + dr));
        const __half kh = __float2half_rn((d < half_d) ? c * cs - e * sn : c * cs + e * sn);
        const __half vh = ld(2, v_new, kvh * kHeadDim + d);
        ks[d] = __half2float(kh);                 // the current position comes from shared memory (rounded like the cache copy)
        vs[d] = __half2float(vh);
        if (h % (n_heads / n_kv) == 0) {
            kcache[((size_t)kvh * max_seq + pos) * kHeadDim + d] = kh;
            vcache[((size_t)kvh * max_seq + pos) * kHeadDim + d] = vh;
        }
    }
    __syncthreads();
    const float scale = rsqrtf((float)kHeadDim);
    const float q0 = qs[4 * lane], q1 = qs[4 * lane + 1], q2 = qs[4 * lane + 2], q3 = qs[4 * lane + 3];
    auto dot4 = [&](const uint2 k) {
        const float2 k0 = __half22float2(*reinterpret_cast<const __half2*>(&k.x)), k1 = __half22float2(*reinterpret_cast<const __half2*>(&k.y));
        return q0 * k0.x + q1 * k0.y + q2 * k1.x + q3 * k1.y;
    };
    float mx = -1e30f;
    #pragma unroll
    for (int u = 0; u < kPre; ++u) {
        const int t = warp + u * kAttnWarps;
        if (t < pos) {                            // warp-uniform
            const float s_ = warp_sum(dot4(kpre[u])) * scale;
            if (lane == 0) sc[t] = s_;
            mx = fmaxf(mx, s_);
        }
    }
    for (int t0 = warp + kPre * kAttnWarps; t0 < pos; t0 += 4 * kAttnWarps) {
        uint2 kv[4];
        #pragma unroll
        for (int u = 0; u < 4; ++u) { const int t = t0 + u * kAttnWarps; kv[u] = t < pos ? __ldcg(kbase + (size_t)t * 32) : make_uint2(0u, 0u); }
        #pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int t = t0 + u * kAttnWarps;
            const float s_ = warp_sum(dot4(kv[u])) * scale;
            if (t < pos) { if (lane == 0) sc[t] = s_; mx = fmaxf(mx, s_); }
        }
    }
    if (warp == pos % kAttnWarps) {
        const float s_ = warp_sum(q0 * ks[4 * lane] + q1 * ks[4 * lane + 1] + q2 * ks[4 * lane + 2] + q3 * ks[4 * lane + 3]) * scale;
        if (lane == 0) sc[pos] = s_;
        mx = fmaxf(mx, s_);
    }
    if (lane == 0) red[warp] = mx;
    __syncthreads();
    mx = red[0];
    #pragma unroll
    for (int w = 1; w < kAttnWarps; ++w) mx = fmaxf(mx, red[w]);
    __syncthreads();
    float sum = 0.f;
    for (int t = tid; t <= pos; t += kAttnWarps * 32) { const float e = __expf(sc[t] - mx); sc[t] = e; sum += e; }
    sum = warp_sum(sum);
    if (lane == 0) red[warp] = sum;
    __syncthreads();
    float tot = 0.f;
    #pragma unroll
    for (int w = 0; w < kAttnWarps; ++w) tot += red[w];
    const float inv = 1.f / tot;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    auto fma4 = [&](float p_, const uint2 v) {
        const float2 v0 = __half22float2(*reinterpret_cast<const __half2*>(&v.x)), v1 = __half22float2(*reinterpret_cast<const __half2*>(&v.y));
        acc[0] += p_ * v0.x; acc[1] += p_ * v0.y; acc[2] += p_ * v1.x; acc[3] += p_ * v1.y;
    };
    #pragma unroll
    for (int u = 0; u < kPre; ++u) {
        const int t = warp + u * kAttnWarps;
        if (t < pos) fma4(sc[t], vpre[u]);
    }
    for (int t0 = warp + kPre * kAttnWarps; t0 < pos; t0 += 4 * kAttnWarps) {
        uint2 vv[4];
        #pragma unroll
        for (int u = 0; u < 4; ++u) { const int t = t0 + u * kAttnWarps; vv[u] = t < pos ? __ldcg(vbase + (size_t)t * 32) : make_uint2(0u, 0u); }
        #pragma unroll
        for (int u = 0; u < 4; ++u) { const int t = t0 + u * kAttnWarps; if (t < pos) fma4(sc[t], vv[u]); }
    }
    if (warp == pos % kAttnWarps) {
        const float p_ = sc[pos];
        #pragma unroll
        for (int j = 0; j < 4; ++j) acc[j] += p_ * vs[4 * lane + j];
    }
    #pragma unroll
    for (int j = 0; j < 4; ++j) part[warp * kHeadDim + 4 * lane + j] = acc[j];
    __syncthreads();
    if (tid < kHeadDim) {
        float o = 0.f;
        #pragma unroll
        for (int w = 0; w < kAttnWarps; ++w) o += part[w * kHeadDim + tid];
        const __half oh = __float2half_rn(o * inv);
        out[h * kHeadDim + tid] = oh;
        if (ll.out) asm volatile("st.volatile.global.u32 [%0], %1;" :: "l"(ll.out + h * kHeadDim + tid),
                                 "r"((uint32_t)__half_as_ushort(oh) | (ll_epoch16(*ll.tick, ll.out_seq) << 16)) : "memory");
    }
}

template <typename... Args>
static int launch_pdl(void (*kern)(Args...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    LUTVQ_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, args...));
    return LUTVQ_OK;
}

// shard.cu: the local LL slot of a shard handle as a glue-kernel input
int shard_ll_input(lutvq_shard_t* h, int slot, const uint32_t** words, const unsigned** tick, unsigned** error, size_t* capacity_rows);

}  // namespace lutvq

using namespace lutvq;

extern "C" {

static int add_rmsnorm_impl(void* resid, const void* delta, const void* w, void* y, int n, float eps, void* stream, const LLIn& ll) {
    LUTVQ_REQUIRE(resid && w && y && n > 0 && n <= 16384, "bad arguments (n=%d, n <= 16384)", n);
    const bool vec8 = n % 8 == 0 && n <= 8192 && aligned16(resid) && aligned16(w) && aligned16(y) && (!delta || aligned16(delta));
    if (vec8)
        return launch_pdl(add_rmsnorm_vec8_kernel, dim3(1), dim3((n / 8 + 31) / 32 * 32), 0, static_cast<cudaStream_t>(stream), static_cast<__half*>(resid),
                          static_cast<const __half*>(delta), static_cast<const __half*>(w), static_cast<__half*>(y), n, eps, ll);
    return launch_pdl(add_rmsnorm_kernel, dim3(1), dim3(1024), 0, static_cast<cudaStream_t>(stream), static_cast<__half*>(resid),
                      static_cast<const __half*>(delta), static_cast<const __half*>(w), static_cast<__half*>(y), n, eps, ll);
}

int lutvq_decode_add_rmsnorm(void* resid, const void* delta, const void* w, void* y, int n, float eps, void* stream) {
    return add_rmsnorm_impl(resid, delta, w, y, n, eps, stream, LLIn{});
}

int lutvq_decode_add_rmsnorm_ll(lutvq_shard_t* h, void* resid, int delta_slot, unsigned delta_seq, const void* w, void* y, int n, float eps,
                                void* stream) {
    LLIn ll;
    size_t cap = 0;
    if (int rc = shard_ll_input(h, delta_slot, &ll.w[0], &ll.tick, &ll.error, &cap)) return rc;
    LUTVQ_REQUIRE((size_t)n <= cap, "n=%d exceeds the gather slot (%zu rows)", n, cap);
    ll.seq[0] = delta_seq;
    return add_rmsnorm_impl(resid, nullptr, w, y, n, eps, stream, ll);
}

// The same with the normalised vector ALSO written as LL words into the local slot y_slot (sequence number y_seq); delta_slot < 0:
// no delta (first layer of a token).
int lutvq_decode_add_rmsnorm_ll_out(lutvq_shard_t* h, void* resid, int delta_slot, unsigned delta_seq, const void* w, void* y, int y_slot,
                                    unsigned y_seq, int n, float eps, void* stream) {
    LLIn ll;
    size_t cap = 0;
    if (delta_slot >= 0) {
        if (int rc = shard_ll_input(h, delta_slot, &ll.w[0], &ll.tick, &ll.error, &cap)) return rc;
        LUTVQ_REQUIRE((size_t)n <= cap, "n=%d exceeds the gather slot (%zu rows)", n, cap);
        ll.seq[0] = delta_seq;
    }
    const uint32_t* ow = nullptr;
    if (int rc = shard_ll_input(h, y_slot, &ow, &ll.tick, &ll.error, &cap)) return rc;
    LUTVQ_REQUIRE((size_t)n <= cap, "n=%d exceeds the gather slot (%zu rows)", n, cap);
    ll.out = const_cast<uint32_t*>(ow);
    ll.out_seq = y_seq;
    return add_rmsnorm_impl(resid, nullptr, w, y, n, eps, stream, ll);
}

int lutvq_decode_silu_mul(const void* gate, const void* up, void* y, int n, void* stream) {
    LUTVQ_REQUIRE(gate && up && y && n > 0, "bad arguments");
    return launch_pdl(silu_mul_kernel, dim3((n + 255) / 256), dim3(256), 0, static_cast<cudaStream_t>(stream),
                      static_cast<const __half*>(gate), static_cast<const __half*>(up), static_cast<__half*>(y), n);
}

static int attention_impl(const void* q, const void* k_new, const void* v_new, void* kcache, void* vcache, const int* pos,
                          void* out, int n_heads, int n_kv_heads, int head_dim, int max_seq, float rope_theta, void* stream, const LLIn& ll) {
    LUTVQ_REQUIRE(kcache && vcache && pos && out, "null pointer");
    LUTVQ_REQUIRE(head_dim == kHeadDim, "head_dim must be %d", kHeadDim);
    LUTVQ_REQUIRE(n_heads > 0 && n_kv_heads > 0 && n_heads % n_kv_heads == 0 && max_seq > 0 && max_seq <= 8192, "bad head counts / max_seq");
    return launch_pdl(attn_decode_kernel, dim3(n_heads), dim3(kAttnWarps * 32), ((size_t)max_seq + kAttnWarps * kHeadDim) * sizeof(float), static_cast<cudaStream_t>(stream),
                      static_cast<const __half*>(q), static_cast<const __half*>(k_new), static_cast<const __half*>(v_new),
                      static_cast<__half*>(kcache), static_cast<__half*>(vcache), pos, static_cast<__half*>(out), n_heads, n_kv_heads,
                      max_seq, rope_theta, ll);
}

int lutvq_decode_attention(const void* q, const void* k_new, const void* v_new, void* kcache, void* vcache, const int* pos,
                           void* out, int n_heads, int n_kv_heads, int head_dim, int max_seq, float rope_theta, void* stream) {
    LUTVQ_REQUIRE(q && k_new && v_new, "null pointer");
    return attention_impl(q, k_new, v_new, kcache, vcache, pos, out, n_heads, n_kv_heads, head_dim, max_seq, rope_theta, stream, LLIn{});
}

int lutvq_decode_attention_ll(lutvq_shard_t* h, int q_slot, int k_slot, int v_slot, unsigned q_seq, unsigned k_seq, unsigned v_seq,
                              void* kcache, void* vcache, const int* pos, void* out, int n_heads, int n_kv_heads, int head_dim, int max_seq,
                              float rope_theta, void* stream) {
    LLIn ll;
    size_t cap = 0;
    const int slots[3] = {q_slot, k_slot, v_slot};
    const unsigned seqs[3] = {q_seq, k_seq, v_seq};
    for (int k = 0; k < 3; ++k) {
        if (int rc = shard_ll_input(h, slots[k], &ll.w[k], &ll.tick, &ll.error, &cap)) return rc;
        ll.seq[k] = seqs[k];
    }
    LUTVQ_REQUIRE((size_t)n_heads * head_dim <= cap, "q (%d values) exceeds the gather slot (%zu rows)", n_heads * head_dim, cap);
    return attention_impl(nullptr, nullptr, nullptr, kcache, vcache, pos, out, n_heads, n_kv_heads, head_dim, max_seq, rope_theta, stream, ll);
}

// The same with the attention output ALSO written as LL words into the local slot out_slot (sequence number out_seq).
int lutvq_decode_attention_ll_out(lutvq_shard_t* h, int q_slot, int k_slot, int v_slot, unsigned q_seq, unsigned k_seq, unsigned v_seq,
                                  void* kcache, void* vcache, const int* pos, void* out, int out_slot, unsigned out_seq, int n_heads,
                                  int n_kv_heads, int head_dim, int max_seq, float rope_theta, void* stream) {
    LLIn ll;
    size_t cap = 0;
    const int slots[3] = {q_slot, k_slot, v_slot};
    const unsigned seqs[3] = {q_seq, k_seq, v_seq};
    for (int k = 0; k < 3; ++k) {
        if (int rc = shard_ll_input(h, slots[k], &ll.w[k], &ll.tick, &ll.error, &cap)) return rc;
        ll.seq[k] = seqs[k];
    }
    const uint32_t* ow = nullptr;
    if (int rc = shard_ll_input(h, out_slot, &ow, &ll.tick, &ll.error, &cap)) return rc;
    LUTVQ_REQUIRE((size_t)n_heads * head_dim <= cap, "q (%d values) exceeds the gather slot (%zu rows)", n_heads * head_dim, cap);
    ll.out = const_cast<uint32_t*>(ow);
    ll.out_seq = out_seq;
    return attention_impl(nullptr, nullptr, nullptr, kcache, vcache, pos, out, n_heads, n_kv_heads, head_dim, max_seq, rope_theta, stream, ll);
}

}  // extern "C"
