// tmac.cu -- SURVEY 8a row a14 / 8f rank 1: the T-MAC bit-serial integer-LUT entry points of the reference's
// benchmark/kernel/bs_lut_gemm/{tbl.cc, lut_ctor.cc} on the GPU, x86 semantics (float_type = float, tbl.cc:10-15):
//
//   tbl_g4_int8_int32_update_impl  (tbl.cc:536-630)   c[2i + l] += sum_k lut[k][a[iK + 16k + l] & 15]   (int16 wraparound k-sum)
//                                                     c[2i + 16 + l] += sum_k lut[k][a[iK + 16k + l] >> 4]
//   tbl_g4_int8_float_update_impl  (tbl.cc:323-532)   the same sums, scaled per act_k block by lut_scales / lut_biases and per
//                                                     output by `scales`, accumulated in fp32 in the reference's own order
//   tbl_*_reset                    (tbl.cc:697-715)
//   partial_max_g4_int8_k8 / partial_max_reset        (lut_ctor.cc:232-266)
//   lut_ctor_g4_int8_impl          (lut_ctor.cc:38-221)  16 signed sums +-b0+-b1+-b2+-b3 per group of 4 activations, int8
//   preprocessor_int8 / qgemm_lut_int8                (T-MAC deploy/tuned/kernels.cc:1002-1037, 827-960: the call structure
//                                                     the generated kernels wrap around the functions above)
//
// Integer results are exact; the fp32 results reproduce the reference BIT FOR BIT, so every floating-point operation is
// written with an explicit rounding intrinsic (__fadd_rn / __fmul_rn / __fdiv_rn: never contracted) and fmaf appears
// exactly where the reference writes _mm256_fmadd_ps.  A 16-entry table is ONE 128-bit register quad per thread: the
// lookup is a byte extract (PRMT) from the quad selected by the nibble, no shared-memory traffic at all.
//
// Work split: a thread owns one byte lane (block of 32 outputs, l) = two outputs (low nibble, high nibble); the k loop
// stays sequential inside the thread because the reference's int16 wraparound and fp32 order are order-dependent.
#include "common.cuh"

namespace lutvq {

// one lookup in the 16-entry int8 table held as 4 words: entry n = byte (n & 3) of word (n >> 2)
__device__ __forceinline__ int lut16(const uint4& t, unsigned n) {
    const uint32_t lo = (n & 4u) ? t.y : t.x, hi = (n & 4u) ? t.w : t.z;
    const uint32_t w = (n & 8u) ? hi : lo;
    return (int)(int8_t)__byte_perm(w, 0u, n & 3u);
}
__device__ __forceinline__ uint4 ld_table(const int8_t* lut, int k) { return __ldg(reinterpret_cast<const uint4*>(lut) + k); }

// int16-wraparound sums of K lookups for one byte lane (SignedWideningAdder, tbl.cc:216-235: _mm256_add_epi16 wraps)
__device__ __forceinline__ void lane_sums_i16(const int8_t* lut, const uint8_t* a, int k0, int nk, int& bot, int& top) {
    int b = 0, t = 0;
    for (int k = 0; k < nk; ++k) {
        const uint4 tab = ld_table(lut, k0 + k);
        const unsigned v = a[(size_t)(k0 + k) * 16];
        b += lut16(tab, v & 15u);
        t += lut16(tab, v >> 4);
    }
    bot = (int)(int16_t)b;                       // a wrapped int16 sum equals the exact sum reduced modulo 2^16
    top = (int)(int16_t)t;
}

// tbl.cc:586-627
__global__ void tbl_int32_update_kernel(int m, int32_t* __restrict__ c, const int8_t* __restrict__ lut,
                                        const uint8_t* __restrict__ a, int K) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;             // byte lane: block t / 16, l = t % 16
    if (t >= m / 2) return;
    const int i = (t >> 4) << 4, l = t & 15;
    int bot, top;
    lane_sums_i16(lut, a + (size_t)i * K + l, 0, K, bot, top);
    c[i * 2 + l] += bot;
    c[i * 2 + 16 + l] += top;
}

__device__ __forceinline__ int get_bias_scale(int bits) { return bits == 4 ? 15 : bits == 3 ? 7 : bits == 2 ? 3 : bits == 1 ? 1 : 0; }   // tbl.cc:300-317
__device__ __forceinline__ int mylog2(int k) { int r = -1; while (k) { ++r; k >>= 1; } return r; }                                         // tbl.cc:239-250

// SignedHalvingAdder (tbl.cc:169-215): a binary tree of _mm256_avg_epu8 = UNSIGNED rounding averages of the raw bytes
__device__ __forceinline__ void lane_sums_halving(const int8_t* lut, const uint8_t* a, int k0, int act_k, int& bot, int& top) {
    uint8_t bb[64], tt[64];
    for (int k = 0; k < act_k; ++k) {
        const uint4 tab = ld_table(lut, k0 + k);
        const unsigned v = a[(size_t)(k0 + k) * 16];
        bb[k] = (uint8_t)lut16(tab, v & 15u);
        tt[k] = (uint8_t)lut16(tab, v >> 4);
    }
    for (int n = act_k; n > 1; n >>= 1)
        for (int j = 0; j < n / 2; ++j) {
            bb[j] = (uint8_t)(((unsigned)bb[2 * j] + (unsigned)bb[2 * j + 1] + 1u) >> 1);
            tt[j] = (uint8_t)(((unsigned)tt[2 * j] + (unsigned)tt[2 * j + 1] + 1u) >> 1);
        }
    bot = (int)(int8_t)bb[0];
    top = (int)(int8_t)tt[0];
}

// One byte lane of tbl_g4_int8_float_update_impl (tbl.cc:433-520): returns the two finished outputs.
// Lane (block i/16, l): low nibble -> sub-block s = l / 8, high nibble -> s = 2 + l / 8, element e = l % 8; ib = i/4 + s.
struct TblFloatCfg { int K, bits, act_k, fast_aggregation, zero_point; };
__device__ __forceinline__ void tbl_float_lane(const TblFloatCfg& g, int i, int l, float c_bot, float c_top, const int8_t* lut,
                                               const uint8_t* a_lane, const float* scales, const float* lut_scales,
                                               const float* lut_biases, float& out_bot, float& out_top) {
    const int e = l & 7, ib_bot = i / 4 + (l >> 3), ib_top = ib_bot + 2;
    float vb = 0.f, vt = 0.f;
    float partial_sum = -0.0f;
    for (int kk = 0; kk < g.K; kk += g.act_k) {
        int bot, top;
        if (g.fast_aggregation) lane_sums_halving(lut, a_lane, kk, g.act_k, bot, top);
        else lane_sums_i16(lut, a_lane, kk, g.act_k, bot, top);
        float lut_s = __ldg(lut_scales + kk / g.act_k);
        float lut_b = __ldg(lut_biases + kk / g.act_k);
        partial_sum = __fadd_rn(partial_sum, lut_b);
        if (g.fast_aggregation) {
            lut_s = __fmul_rn(lut_s, (float)g.act_k);
            lut_b = __fsub_rn(lut_b, __fmul_rn(lut_s, (float)(mylog2(g.act_k) / 4 * get_bias_scale(g.bits))));
        }
        const float fb = (float)bot, ft = (float)top;
        const float xb = (ib_bot % g.bits) ? __fmul_rn(fb, lut_s) : fmaf(fb, lut_s, lut_b);
        const float xt = (ib_top % g.bits) ? __fmul_rn(ft, lut_s) : fmaf(ft, lut_s, lut_b);
        vb = (kk == 0) ? xb : __fadd_rn(vb, xb);
        vt = (kk == 0) ? xt : __fadd_rn(vt, xt);
    }
    if (g.zero_point) {
        partial_sum = __fmul_rn(partial_sum, 2.f);
        const float* sb = scales + (ib_bot / g.bits) * 16;
        const float* st = scales + (ib_top / g.bits) * 16;
        const float tb = fmaf(vb, __ldg(sb + e), c_bot), tt = fmaf(vt, __ldg(st + e), c_top);
        out_bot = (ib_bot % g.bits) ? tb : fmaf(__ldg(sb + 8 + e), partial_sum, tb);
        out_top = (ib_top % g.bits) ? tt : fmaf(__ldg(st + 8 + e), partial_sum, tt);
    } else {
        out_bot = fmaf(vb, __ldg(scales + (ib_bot / g.bits) * 8 + e), c_bot);
        out_top = fmaf(vt, __ldg(scales + (ib_top / g.bits) * 8 + e), c_top);
    }
}

__global__ void tbl_float_update_kernel(int m, float* __restrict__ c, const int8_t* __restrict__ lut, const uint8_t* __restrict__ a,
                                        const float* __restrict__ scales, const float* __restrict__ lut_scales,
                                        const float* __restrict__ lut_biases, TblFloatCfg g) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= m / 2) return;
    const int i = (t >> 4) << 4, l = t & 15;
    float ob, ot;
    tbl_float_lane(g, i, l, c[i * 2 + l], c[i * 2 + 16 + l], lut, a + (size_t)i * g.K + l, scales, lut_scales, lut_biases, ob, ot);
    c[i * 2 + l] = ob;
    c[i * 2 + 16 + l] = ot;
}

// ---- lut_ctor.cc ----
// abs-sum maximum of 8 groups of 4 activations / 127 (lut_ctor.cc:241-259); std::max(*lut_scales, scales) = (a < b) ? b : a
__device__ __forceinline__ float partial_max32(const float* b) {
    float mx = -INFINITY;
    for (int j = 0; j < 8; ++j) {
        const float s = __fadd_rn(__fadd_rn(fabsf(b[4 * j]), fabsf(b[4 * j + 1])), __fadd_rn(fabsf(b[4 * j + 2]), fabsf(b[4 * j + 3])));
        mx = fmaxf(mx, s);                      // _mm_max_ps on non-NaN data
    }
    return __fdiv_rn(mx, 127.f);
}
__global__ void partial_max_kernel(float* lut_scales, const float* __restrict__ b) {
    const float s = partial_max32(b);
    if (*lut_scales < s) *lut_scales = s;
}

// One CTA of 128 threads per independent (scale, bias) segment of act_k activations; thread = (group j, entry g).
// preprocess != 0: the segment's scale is first recomputed as partial_max_reset + partial_max (act_k must be 32).
__global__ void __launch_bounds__(128) lut_ctor_kernel(int act_k, int8_t* __restrict__ qlut, const float* __restrict__ b,
                                                       float* __restrict__ lut_scales, float* __restrict__ lut_biases, int preprocess) {
    const int seg = blockIdx.x, j = threadIdx.x >> 4, g = threadIdx.x & 15;
    b += (size_t)seg * act_k;
    qlut += (size_t)seg * act_k * 4;
    __shared__ float e0[8];
    __shared__ float s_scale;
    if (threadIdx.x == 0) {
        float s = lut_scales[seg];
        if (preprocess) { const float p = partial_max32(b); s = (0.0f < p) ? p : 0.0f; }
        s_scale = s;
    }
    __syncthreads();
    const float scales = s_scale;
    const float t_scales = scales ? __fdiv_rn(1.0f, scales) : 0.0f;
    float biases = 0.0f;
    for (int k = 0; k < act_k / 32; ++k) {
        const float* bj = b + k * 32 + 4 * j;
        // odd g: ((b0 +- b1) +- b2) +- b3 with + where bit 1 / 2 / 3 of g is set; even g: -(entry 15 - g)   (lut_ctor.cc:133-176)
        const int go = (g & 1) ? g : 15 - g;
        float v = bj[0];
        v = (go & 2) ? __fadd_rn(v, bj[1]) : __fsub_rn(v, bj[1]);
        v = (go & 4) ? __fadd_rn(v, bj[2]) : __fsub_rn(v, bj[2]);
        v = (go & 8) ? __fadd_rn(v, bj[3]) : __fsub_rn(v, bj[3]);
        if (!(g & 1)) v = -v;
        if (g == 0) e0[j] = v;
        float r = rintf(__fmul_rn(v, t_scales));            // _mm256_round_ps(NEAREST) -> cvtps_epi32 -> packs (saturating)
        r = fminf(fmaxf(r, -128.f), 127.f);
        qlut[k * 128 + j * 16 + g] = (int8_t)(int)r;
        __syncthreads();
        // _mm256_addv_ps order (lut_ctor.cc:26-32): ((v0+v4) + (v2+v6)) + ((v1+v5) + (v3+v7))
        if (threadIdx.x == 0)
            biases = __fadd_rn(biases, __fadd_rn(__fadd_rn(__fadd_rn(e0[0], e0[4]), __fadd_rn(e0[2], e0[6])),
                                                 __fadd_rn(__fadd_rn(e0[1], e0[5]), __fadd_rn(e0[3], e0[7]))));
        __syncthreads();
    }
    if (threadIdx.x == 0) { lut_scales[seg] = scales; lut_biases[seg] = biases; }
}

// qgemm_lut_t1_int8_m{bm}_k{K}_n1_b{bits} (kernels.cc:827-960): per byte lane (two bit-plane rows) and k-block of 32 activations
//     x = sum of 8 table lookups (int16) * lut_scale (+ lut_bias on the first bit plane);   c = fma(x, scale, c)
// then C = ((c0 * 0.5 + c1) + c2 * 2) + c3 * 4 over the bit planes.  The fp32 chain over the k-blocks is order-dependent, the
// lookups are not: a CTA owns 16 byte lanes (48 for 3 bits: the bit planes of an output then lie inside the CTA) and
//   phase 1  all its threads (lane x 32 k-slots) compute x and fetch the scale for a chunk of k-blocks into shared memory,
//   phase 2  one thread per (lane, nibble) runs the chunk's fma chain in the reference's order (loads hoisted, 4 clocks a step),
// so the grid is M * bits / 32 CTAs instead of M * bits / bm and a lane's lookups are no longer one serial chain of K / 32 global
// round trips (first version: 128 us per 4096 x 4096 4-bit GEMV whatever M was).  Results are bit-identical to that version.
constexpr int kTmacThreads = 512;
__global__ void __launch_bounds__(kTmacThreads) tmac_qgemm_kernel(int K, int bits, int bm, int nblk, int chunk, const uint8_t* __restrict__ A,
                                                                  const int8_t* __restrict__ qlut, const float* __restrict__ scales,
                                                                  const float* __restrict__ lut_scales, const float* __restrict__ lut_biases,
                                                                  float* __restrict__ C) {
    extern __shared__ float sm_f[];
    const int LC = 16 * nblk, W = 2 * LC;                            // byte lanes and cbits of this CTA
    float* sv = sm_f;                                                // [chunk][W]  x values
    float* ss = sm_f + (size_t)chunk * W;                            // [chunk][W]  scales
    __shared__ float cb[96];
    const int ctas_per_tile = bm / 2 / LC, tile = blockIdx.x / ctas_per_tile, sub = blockIdx.x - tile * ctas_per_tile;
    const int kt_n = K / 32, KP = kTmacThreads / LC;
    const int ll = threadIdx.x % LC, q = threadIdx.x / LC;           // phase 1: local lane, k-slot
    const int lane_t = sub * LC + ll, i = (lane_t >> 4) << 4, l = lane_t & 15, e = l & 7;
    const int ib_bot = i / 4 + (l >> 3), ib_top = ib_bot + 2;
    const bool plain_b = (ib_bot % bits) != 0, plain_t = (ib_top % bits) != 0;
    float c = 0.f;                                                   // tbl_float_reset; phase-2 threads only
    for (int kc = 0; kc < kt_n; kc += chunk) {
        const int kend = min(kt_n, kc + chunk);
        if (q < KP)
            for (int kt = kc + q; kt < kend; kt += KP) {
                const size_t blk = (size_t)tile * kt_n + kt;
                const uint8_t* a = A + blk * (size_t)(bm / 2 * 8) + (size_t)i * 8 + l;
                const int8_t* lut = qlut + (size_t)kt * 128;
                int bot = 0, top = 0;
                #pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const uint4 tab = ld_table(lut, k);
                    const unsigned v = __ldg(a + k * 16);
                    bot += lut16(tab, v & 15u);
                    top += lut16(tab, v >> 4);
                }
                const float fb = (float)(int)(int16_t)bot, ft = (float)(int)(int16_t)top;
                const float lut_s = __ldg(lut_scales + kt), lut_b = __ldg(lut_biases + kt);
                const float* sc = scales + blk * (size_t)(bm / bits);
                float* pv = sv + (size_t)(kt - kc) * W + 2 * ll;
                float* ps = ss + (size_t)(kt - kc) * W + 2 * ll;
                pv[0] = plain_b ? __fmul_rn(fb, lut_s) : fmaf(fb, lut_s, lut_b);
                pv[1] = plain_t ? __fmul_rn(ft, lut_s) : fmaf(ft, lut_s, lut_b);
                ps[0] = __ldg(sc + (ib_bot / bits) * 8 + e);
                ps[1] = __ldg(sc + (ib_top / bits) * 8 + e);
            }
        __syncthreads();
        if (threadIdx.x < W) {
            const float* pv = sv + threadIdx.x;
            const float* ps = ss + threadIdx.x;
            const int n = kend - kc;
            int k = 0;
            for (; k + 8 <= n; k += 8) {
                float v[8], sc8[8];
                #pragma unroll
                for (int u = 0; u < 8; ++u) { v[u] = pv[(size_t)(k + u) * W]; sc8[u] = ps[(size_t)(k + u) * W]; }
                #pragma unroll
                for (int u = 0; u < 8; ++u) c = fmaf(v[u], sc8[u], c);
            }
            for (; k < n; ++k) c = fmaf(pv[(size_t)k * W], ps[(size_t)k * W], c);
        }
        __syncthreads();
    }
    if (threadIdx.x < W) {
        const int lane_local = threadIdx.x >> 1, which = threadIdx.x & 1;
        cb[(lane_local >> 4) * 32 + which * 16 + (lane_local & 15)] = c;   // cbits[i * 2 + l] (low nibble), cbits[i * 2 + 16 + l] (high)
    }
    __syncthreads();
    const float alphas[4] = {0.5f, 1.f, 2.f, 4.f};
    const int n_out_cta = W / bits, n_out_tile = bm / bits;
    for (int o = threadIdx.x; o < n_out_cta; o += blockDim.x) {
        const int j = o >> 3, eo = o & 7;
        float acc = __fmul_rn(cb[8 * bits * j + eo], alphas[0]);
        for (int bp = 1; bp < bits; ++bp) acc = __fadd_rn(acc, __fmul_rn(cb[8 * bits * j + 8 * bp + eo], alphas[bp]));
        C[(size_t)tile * n_out_tile + (size_t)sub * n_out_cta + o] = acc;
    }
}

// Second version of the same kernel: phase 1 does FOUR lookups per PRMT triple (pq4.cu's trick).  A thread owns one (k-block, block of
// 16 byte lanes): it reads the block's 128 bytes [table k][lane] as 8 x 16 bytes and the 8 tables as 8 x 16 bytes; 16 bits of a row =
// two byte lanes = 4 codes that share table k, so  prmt(prmt(T0,T1,s&0x7777), prmt(T2,T3,s&0x7777), 0x3210|bit3<<2)  returns
// (low, high nibble of lane 2j, low, high of lane 2j+1).  Tables are biased by +128 (xor 0x80) so the 8 results per lane add up as
// unsigned bytes packed in one register (R += r; H += odd bytes; even sums = R - (H << 8)); 8 * 255 < 2^16, the bias 8 * 128 comes
// off at the end.  ~3 instructions per lookup instead of ~30.  Phase 2 (the order-dependent fma chain) is unchanged.
constexpr int kTmacV2Threads = 256;
__global__ void __launch_bounds__(kTmacV2Threads) tmac_qgemm_v2_kernel(int K, int bits, int bm, int nblk, int chunk, const uint8_t* __restrict__ A,
                                                                       const int8_t* __restrict__ qlut, const float* __restrict__ scales,
                                                                       const float* __restrict__ lut_scales, const float* __restrict__ lut_biases,
                                                                       float* __restrict__ C) {
    extern __shared__ __align__(16) float sm_f[];
    const int LC = 16 * nblk, W = 2 * LC, SW = (LC / 4 / bits) * 8;  // byte lanes, cbits and scale floats (per k-block) of this CTA
    float* sv = sm_f;                                                // [chunk][W]   x values
    float* ssc = sm_f + (size_t)chunk * W;                           // [chunk][SW]  scales
    __shared__ float cb[96];
    const int ctas_per_tile = bm / 2 / LC, tile = blockIdx.x / ctas_per_tile, sub = blockIdx.x - tile * ctas_per_tile;
    const int kt_n = K / 32;
    const int ib0 = sub * LC / 4, s0 = (ib0 / bits) * 8;            // first scale float of the CTA inside a k-block's scale row
    // phase-2 identity of this thread: (byte lane, nibble) -> its scale column
    const int lane2 = threadIdx.x >> 1, which = threadIdx.x & 1;
    const int i2 = ((sub * LC + lane2) >> 4) << 4, l2 = lane2 & 15;
    const int ib2 = i2 / 4 + (l2 >> 3) + 2 * which, soff2 = (ib2 / bits) * 8 + (l2 & 7) - s0;
    float c = 0.f;                                                   // tbl_float_reset
    for (int kc = 0; kc < kt_n; kc += chunk) {
        // an item = (k-block, block of 16 byte lanes, half of the block): 8 lanes x 8 tables per thread -- twice the threads of a
        // whole-block item (the grid is only M * bits / 32 CTAs: more warps per SM, not fewer instructions, is what helps here)
        const int kend = min(kt_n, kc + chunk), n_items = (kend - kc) * nblk * 2;
        for (int it = threadIdx.x; it < n_items; it += kTmacV2Threads) {
            const int hb = it & 1, itb = it >> 1;
            const int kt = kc + itb / nblk, b = itb - (itb / nblk) * nblk;
            const int i = sub * LC + b * 16;                         // first byte lane of the block inside the tile (multiple of 16)
            const size_t blk = (size_t)tile * kt_n + kt;
            const uint2* a = reinterpret_cast<const uint2*>(A + blk * (size_t)(bm / 2 * 8) + (size_t)i * 8 + hb * 8);   // + k * 16 bytes
            const uint4* tabs = reinterpret_cast<const uint4*>(qlut + (size_t)kt * 128);
            uint32_t R[4], H[4];
            #pragma unroll
            for (int j = 0; j < 4; ++j) R[j] = H[j] = 0u;
            #pragma unroll
            for (int k = 0; k < 8; ++k) {
                uint4 T = __ldg(tabs + k);
                T.x ^= 0x80808080u; T.y ^= 0x80808080u; T.z ^= 0x80808080u; T.w ^= 0x80808080u;
                const uint2 av = __ldg(a + 2 * k);
                const uint32_t w[2] = {av.x, av.y};
                #pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint32_t sj = (j & 1) ? (w[j >> 1] >> 16) : w[j >> 1];
                    const uint32_t m = sj & 0x7777u;
                    const uint32_t r = __byte_perm(__byte_perm(T.x, T.y, m), __byte_perm(T.z, T.w, m), ((sj >> 1) & 0x4444u) | 0x3210u);
                    R[j] += r;
                    H[j] += __byte_perm(r, 0u, 0x4341);
                }
            }
            const float lut_s = __ldg(lut_scales + kt), lut_b = __ldg(lut_biases + kt);
            float4* out = reinterpret_cast<float4*>(sv + (size_t)(kt - kc) * W + b * 32 + hb * 16);
            const int ib_bot = i / 4 + hb, ib_top = ib_bot + 2;      // the 8 lanes of a half share l >> 3 = hb
            const bool plain_b = (ib_bot % bits) != 0, plain_t = (ib_top % bits) != 0;
            #pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t E = R[j] - (H[j] << 8);               // even bytes: low nibbles of two lanes;  H: their high nibbles
                const float f0 = (float)((int)(E & 0xffffu) - 1024), f1 = (float)((int)(H[j] & 0xffffu) - 1024);
                const float f2 = (float)((int)(E >> 16) - 1024), f3 = (float)((int)(H[j] >> 16) - 1024);
                float4 o;
                o.x = plain_b ? __fmul_rn(f0, lut_s) : fmaf(f0, lut_s, lut_b);
                o.y = plain_t ? __fmul_rn(f1, lut_s) : fmaf(f1, lut_s, lut_b);
                o.z = plain_b ? __fmul_rn(f2, lut_s) : fmaf(f2, lut_s, lut_b);
                o.w = plain_t ? __fmul_rn(f3, lut_s) : fmaf(f3, lut_s, lut_b);
                out[j] = o;
            }
            if (b == 0 && hb == 0) {                                 // the k-block's scales of this CTA: SW contiguous floats
                const float4* sc = reinterpret_cast<const float4*>(scales + blk * (size_t)(bm / bits) + s0);
                float4* dst = reinterpret_cast<float4*>(ssc + (size_t)(kt - kc) * SW);
                for (int v = 0; v < SW / 4; ++v) dst[v] = __ldg(sc + v);
            }
        }
        __syncthreads();
        if (threadIdx.x < W) {
            const float* pv = sv + threadIdx.x;
            const float* ps = ssc + soff2;
            const int n = kend - kc;
            int k = 0;
            for (; k + 8 <= n; k += 8) {
                float v[8], sc8[8];
                #pragma unroll
                for (int u = 0; u < 8; ++u) { v[u] = pv[(size_t)(k + u) * W]; sc8[u] = ps[(size_t)(k + u) * SW]; }
                #pragma unroll
                for (int u = 0; u < 8; ++u) c = fmaf(v[u], sc8[u], c);
            }
            for (; k < n; ++k) c = fmaf(pv[(size_t)k * W], ps[(size_t)k * SW], c);
        }
        __syncthreads();
    }
    if (threadIdx.x < W) cb[(lane2 >> 4) * 32 + which * 16 + (lane2 & 15)] = c;   // cbits[i * 2 + l] (low nibble), cbits[i * 2 + 16 + l] (high)
    __syncthreads();
    const float alphas[4] = {0.5f, 1.f, 2.f, 4.f};
    const int n_out_cta = W / bits, n_out_tile = bm / bits;
    for (int o = threadIdx.x; o < n_out_cta; o += blockDim.x) {
        const int j = o >> 3, eo = o & 7;
        float acc = __fmul_rn(cb[8 * bits * j + eo], alphas[0]);
        for (int bp = 1; bp < bits; ++bp) acc = __fadd_rn(acc, __fmul_rn(cb[8 * bits * j + 8 * bp + eo], alphas[bp]));
        C[(size_t)tile * n_out_tile + (size_t)sub * n_out_cta + o] = acc;
    }
}

static int need_device() {
    DeviceInfo di;
    return get_device_info(&di);                     // no device: LUTVQ_ECUDA, there is no CPU fallback
}

}  // namespace lutvq

using namespace lutvq;

extern "C" {

int lutvq_tbl_g4_int8_int32_update(int m, void* c, const void* lut, const void* a, int K, void* stream) {
    LUTVQ_REQUIRE(c && lut && a, "null pointer");
    LUTVQ_REQUIRE(m > 0 && m % 32 == 0 && K > 0, "m=%d must be a positive multiple of 32 and K=%d positive (tbl.cc:586 walks blocks of 32 outputs)", m, K);
    LUTVQ_REQUIRE(aligned16(lut), "lut must be 16-byte aligned (one table = one 128-bit load)");
    if (int rc = need_device()) return rc;
    const int lanes = m / 2;
    tbl_int32_update_kernel<<<ceil_div(lanes, 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(
        m, static_cast<int32_t*>(c), static_cast<const int8_t*>(lut), static_cast<const uint8_t*>(a), K);
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

int lutvq_tbl_g4_int8_float_update(int m, void* c, const void* lut, const void* a, const void* scales, const void* lut_scales,
                                   const void* lut_biases, int K, int bits, int act_k, int fast_aggregation, int zero_point,
                                   void* stream) {
    LUTVQ_REQUIRE(c && lut && a && scales && lut_scales && lut_biases, "null pointer");
    LUTVQ_REQUIRE(m > 0 && m % 32 == 0, "m=%d must be a positive multiple of 32", m);
    LUTVQ_REQUIRE(bits >= 1 && bits <= 4, "bits=%d must be 1..4 (tbl.cc:300-317)", bits);
    LUTVQ_REQUIRE(act_k > 0 && K > 0 && K % act_k == 0, "K=%d must be a positive multiple of act_k=%d", K, act_k);
    LUTVQ_REQUIRE(!fast_aggregation || (act_k <= 64 && (act_k & (act_k - 1)) == 0), "fast aggregation needs a power-of-two act_k <= 64, got %d", act_k);
    LUTVQ_REQUIRE(aligned16(lut), "lut must be 16-byte aligned (one table = one 128-bit load)");
    if (int rc = need_device()) return rc;
    const TblFloatCfg g{K, bits, act_k, fast_aggregation ? 1 : 0, zero_point ? 1 : 0};
    tbl_float_update_kernel<<<ceil_div(m / 2, 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(
        m, static_cast<float*>(c), static_cast<const int8_t*>(lut), static_cast<const uint8_t*>(a), static_cast<const float*>(scales),
        static_cast<const float*>(lut_scales), static_cast<const float*>(lut_biases), g);
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

static int reset_bytes(void* c, long long bytes, void* stream) {
    LUTVQ_REQUIRE(c != nullptr && bytes >= 0, "null pointer or negative size");
    if (int rc = need_device()) return rc;
    LUTVQ_CUDA_OK(cudaMemsetAsync(c, 0, (size_t)bytes, static_cast<cudaStream_t>(stream)));
    return LUTVQ_OK;
}
int lutvq_tbl_int8_reset(int m, void* c, void* stream)  { return reset_bytes(c, (long long)m, stream); }
int lutvq_tbl_int16_reset(int m, void* c, void* stream) { return reset_bytes(c, (long long)m * 2, stream); }
int lutvq_tbl_int32_reset(int m, void* c, void* stream) { return reset_bytes(c, (long long)m * 4, stream); }
int lutvq_tbl_float_reset(int m, void* c, void* stream) { return reset_bytes(c, (long long)m * 4, stream); }
int lutvq_partial_max_reset(void* lut_scales, void* stream) { return reset_bytes(lut_scales, 4, stream); }

int lutvq_partial_max_g4_int8_k8(void* lut_scales, const void* b, void* stream) {
    LUTVQ_REQUIRE(lut_scales && b, "null pointer");
    if (int rc = need_device()) return rc;
    partial_max_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<float*>(lut_scales), static_cast<const float*>(b));
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

int lutvq_lut_ctor_g4_int8(int act_k, void* qlut, const void* b, void* lut_scales, void* lut_biases, void* stream) {
    LUTVQ_REQUIRE(qlut && b && lut_scales && lut_biases, "null pointer");
    LUTVQ_REQUIRE(act_k > 0 && act_k % 32 == 0, "act_k=%d must be a positive multiple of 32 (lut_ctor.cc:127)", act_k);
    if (int rc = need_device()) return rc;
    lut_ctor_kernel<<<1, 128, 0, static_cast<cudaStream_t>(stream)>>>(act_k, static_cast<int8_t*>(qlut), static_cast<const float*>(b),
                                                                     static_cast<float*>(lut_scales), static_cast<float*>(lut_biases), 0);
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

int lutvq_tmac_preprocessor_int8(int K, const void* B, void* lut_scales, void* lut_biases, void* qlut, void* stream) {
    LUTVQ_REQUIRE(B && lut_scales && lut_biases && qlut, "null pointer");
    LUTVQ_REQUIRE(K > 0 && K % 32 == 0, "K=%d must be a positive multiple of 32", K);
    if (int rc = need_device()) return rc;
    lut_ctor_kernel<<<K / 32, 128, 0, static_cast<cudaStream_t>(stream)>>>(32, static_cast<int8_t*>(qlut), static_cast<const float*>(B),
                                                                          static_cast<float*>(lut_scales), static_cast<float*>(lut_biases), 1);
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

int lutvq_tmac_qgemm_lut_int8(int M, int K, int bits, int bm, const void* A, const void* qlut, const void* scales,
                              const void* lut_scales, const void* lut_biases, void* C, void* stream) {
    LUTVQ_REQUIRE(A && qlut && scales && lut_scales && lut_biases && C, "null pointer");
    LUTVQ_REQUIRE(bits >= 1 && bits <= 4, "bits=%d must be 1..4", bits);
    LUTVQ_REQUIRE(K > 0 && K % 32 == 0, "K=%d must be a positive multiple of 32", K);
    // the reference returns -1 when no tuned kernel matches (deploy/tuned/kernels.h): here the shape rules are explicit
    if (!(bm > 0 && bm % 32 == 0 && bm % (8 * bits) == 0 && bm <= 8192 && M > 0 && ((long long)M * bits) % bm == 0))
        return fail(LUTVQ_EUNSUPPORTED, "no kernel for M=%d bits=%d bm=%d: bm must be a multiple of 32 and of 8*bits, <= 8192, and divide M*bits", M, bits, bm);
    LUTVQ_REQUIRE(aligned16(qlut), "qlut must be 16-byte aligned");
    if (int rc = need_device()) return rc;
    const int tiles = (int)((long long)M * bits / bm);
    const int nblk = bits == 3 ? 3 : 1, LC = 16 * nblk;             // bm % (8 * bits) == 0 and bm % 32 == 0 => bm / 2 % LC == 0
    const int chunk = nblk == 1 ? 128 : 40;                           // k-blocks per shared-memory chunk (<= 32 KiB)
    if (aligned16(A) && aligned16(scales)) {                         // 128-bit loads of the weight bytes and the scales
        const int SW = (LC / 4 / bits) * 8;
        const size_t smem2 = (size_t)chunk * (2 * LC + SW) * sizeof(float);
        tmac_qgemm_v2_kernel<<<tiles * (bm / 2 / LC), kTmacV2Threads, smem2, static_cast<cudaStream_t>(stream)>>>(
            K, bits, bm, nblk, chunk, static_cast<const uint8_t*>(A), static_cast<const int8_t*>(qlut), static_cast<const float*>(scales),
            static_cast<const float*>(lut_scales), static_cast<const float*>(lut_biases), static_cast<float*>(C));
        LUTVQ_CUDA_OK(cudaGetLastError());
        return LUTVQ_OK;
    }
    const size_t smem = (size_t)chunk * 2 * LC * 2 * sizeof(float);
    tmac_qgemm_kernel<<<tiles * (bm / 2 / LC), kTmacThreads, smem, static_cast<cudaStream_t>(stream)>>>(
        K, bits, bm, nblk, chunk, static_cast<const uint8_t*>(A), static_cast<const int8_t*>(qlut), static_cast<const float*>(scales),
        static_cast<const float*>(lut_scales), static_cast<const float*>(lut_biases), static_cast<float*>(C));
    LUTVQ_CUDA_OK(cudaGetLastError());
    return LUTVQ_OK;
}

}  // extern "C"
