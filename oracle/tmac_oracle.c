/* tmac_oracle.c -- TEST INFRASTRUCTURE ONLY (see lutvq_oracle.c header).
 *
 * Scalar C restatement of the reference's T-MAC bit-serial integer-LUT kernels, x86 semantics (float_type = float,
 * benchmark/kernel/bs_lut_gemm/tbl.cc:10-15).  Each function cites the lines of the reference's AVX2 path it follows.
 * Parity: PINNED -- oracle/gen_golden_tmac.py runs the reference's own code (oracle/_ref/libtmac_ref.so, compiled from
 * /root/reference in place) and freezes inputs + outputs under tests/golden/tmac_*.npz, including the deterministic
 * known answer of test_lut_ctor.cc (b[i] = i); tests/test_oracle_golden.py replays them bit for bit.
 * Built with -ffp-contract=off: fused multiply-adds appear only where the reference writes _mm256_fmadd_ps.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

/* tbl.cc:586-627  tbl_g4_int8_int32_update_impl<K, Bits>, AVX2:
 * block i (32 outputs): c[2i + l] += sum_k lut[k][a[iK + 16k + l] & 15], c[2i + 16 + l] += sum_k lut[k][a[..] >> 4];
 * the k-sum runs in int16 (SignedWideningAdder, tbl.cc:216-235: _mm256_add_epi16 wraps), then widens to int32. */
void orc_tbl_g4_int8_int32_update(int m, int32_t* c, const int8_t* lut, const uint8_t* a, int K) {
    for (int i = 0; i < m / 2; i += 16)
        for (int l = 0; l < 16; ++l) {
            int16_t bot = 0, top = 0;
            for (int k = 0; k < K; ++k) {
                const uint8_t v = a[i * K + k * 16 + l];
                bot = (int16_t)(bot + lut[k * 16 + (v & 15)]);
                top = (int16_t)(top + lut[k * 16 + (v >> 4)]);
            }
            c[i * 2 + l] += bot;
            c[i * 2 + 16 + l] += top;
        }
}

/* tbl.cc:169-215  SignedHalvingAdder<N>, AVX2: a binary tree of _mm256_avg_epu8 -- UNSIGNED rounding average of the raw
 * bytes ((x + y + 1) >> 1 on uint8), the result reinterpreted as int8 by _mm256_cvtepi8_epi16 (tbl.cc:165-166). */
static uint8_t halving_tree(const uint8_t* v, int n) {
    if (n == 1) return v[0];
    const uint8_t lo = halving_tree(v, n / 2), hi = halving_tree(v + n / 2, n / 2);
    return (uint8_t)(((unsigned)lo + (unsigned)hi + 1u) >> 1);
}
static int get_bias_scale(int bits) { return bits == 4 ? 15 : bits == 3 ? 7 : bits == 2 ? 3 : bits == 1 ? 1 : 0; }   /* tbl.cc:300-317 */
static int mylog2(int k) { return k == 0 ? -1 : 1 + mylog2(k / 2); }                                                  /* tbl.cc:239-250 */

/* tbl.cc:433-520  tbl_g4_int8_float_update_impl<has_scale, K, Bits, ActK, FastAggregation, ZeroPoint, OneScale>, AVX2
 * (has_scale and OneScale are not consulted on this path).  Sub-block s of block i covers 8 outputs; ib = i/4 + s. */
void orc_tbl_g4_int8_float_update(int m, float* c, const int8_t* lut, const uint8_t* a, const float* scales,
                                  const float* lut_scales, const float* lut_biases, int K, int bits, int act_k,
                                  int fast_aggregation, int zero_point) {
    for (int i = 0; i < m / 2; i += 16) {
        float vec_c[4][8];
        float partial_sum = -0.0f;
        for (int kk = 0; kk < K; kk += act_k) {
            float v[4][8];                                        /* bot 0..7, bot 8..15, top 0..7, top 8..15 */
            for (int l = 0; l < 16; ++l) {
                int bot, top;
                if (!fast_aggregation) {
                    int16_t b16 = 0, t16 = 0;
                    for (int k = 0; k < act_k; ++k) {
                        const uint8_t x = a[i * K + (kk + k) * 16 + l];
                        b16 = (int16_t)(b16 + lut[(kk + k) * 16 + (x & 15)]);
                        t16 = (int16_t)(t16 + lut[(kk + k) * 16 + (x >> 4)]);
                    }
                    bot = b16; top = t16;
                } else {
                    uint8_t bb[64], tt[64];
                    for (int k = 0; k < act_k; ++k) {
                        const uint8_t x = a[i * K + (kk + k) * 16 + l];
                        bb[k] = (uint8_t)lut[(kk + k) * 16 + (x & 15)];
                        tt[k] = (uint8_t)lut[(kk + k) * 16 + (x >> 4)];
                    }
                    bot = (int8_t)halving_tree(bb, act_k); top = (int8_t)halving_tree(tt, act_k);
                }
                v[l / 8][l % 8] = (float)bot;
                v[2 + l / 8][l % 8] = (float)top;
            }
            float lut_s = lut_scales[kk / act_k];
            float lut_b = lut_biases[kk / act_k];
            partial_sum += lut_b;
            if (fast_aggregation) {
                lut_s = lut_s * act_k;
                lut_b -= lut_s * (float)(mylog2(act_k) / 4 * get_bias_scale(bits));
            }
            for (int s = 0; s < 4; ++s) {
                const int ib = i / 4 + s;
                for (int e = 0; e < 8; ++e) {
                    const float val = (ib % bits) ? v[s][e] * lut_s : fmaf(v[s][e], lut_s, lut_b);
                    vec_c[s][e] = (kk == 0) ? val : vec_c[s][e] + val;
                }
            }
        }
        if (zero_point) {
            partial_sum *= 2;
            for (int s = 0; s < 4; ++s) {
                const int ib = i / 4 + s;
                const float* vs = scales + (ib / bits) * 16;
                for (int e = 0; e < 8; ++e) {
                    const float t = fmaf(vec_c[s][e], vs[e], c[i * 2 + 8 * s + e]);
                    c[i * 2 + 8 * s + e] = (ib % bits) ? t : fmaf(vs[8 + e], partial_sum, t);
                }
            }
        } else {
            for (int s = 0; s < 4; ++s) {
                const int ib = i / 4 + s;
                const float* vs = scales + (ib / bits) * 8;
                for (int e = 0; e < 8; ++e) c[i * 2 + 8 * s + e] = fmaf(vec_c[s][e], vs[e], c[i * 2 + 8 * s + e]);
            }
        }
    }
}

/* lut_ctor.cc:232-266  partial_max_g4_int8_k8 (AVX2 :241-259) over 32 activations + partial_max_reset */
void orc_partial_max_g4_int8_k8(float* lut_scales, const float* b) {
    float mx = -INFINITY;
    for (int j = 0; j < 8; ++j) {
        const float s = (fabsf(b[4 * j]) + fabsf(b[4 * j + 1])) + (fabsf(b[4 * j + 2]) + fabsf(b[4 * j + 3]));
        if (s > mx) mx = s;
    }
    const float scales = mx / 127;
    if (*lut_scales < scales) *lut_scales = scales;             /* std::max(*lut_scales, scales) */
}

static int8_t sat_round_i8(float v) {                            /* _mm256_round_ps(NEAREST) -> cvtps_epi32 -> packs_epi32 -> packs_epi16 */
    float r = rintf(v);
    if (r > 127.f) r = 127.f;
    if (r < -128.f) r = -128.f;
    return (int8_t)r;
}

/* lut_ctor.cc:120-216  lut_ctor_g4_int8_impl, AVX2: per 32 activations (8 groups of 4), 16 signed sums per group:
 * odd g: ((b0 +- b1) +- b2) +- b3 with + where bit 1/2/3 of g is set; even g: -(entry 15 - g).
 * qlut[k*128 + j*16 + g] = sat_int8(round(entry * (1 / scales)));  biases += horizontal sum of entry 0 over the 8 groups
 * in _mm256_addv_ps order (lut_ctor.cc:26-32): ((v0+v4) + (v2+v6)) + ((v1+v5) + (v3+v7)). */
void orc_lut_ctor_g4_int8(int act_k, int8_t* qlut, const float* b, float* lut_scales, float* lut_biases) {
    float biases = 0.0f;
    const float scales = *lut_scales;
    const float t_scales = scales ? 1.0f / scales : 0.0f;
    for (int k = 0; k < act_k / 32; ++k) {
        float e0[8];
        for (int j = 0; j < 8; ++j) {
            const float* bj = b + k * 32 + 4 * j;
            float ent[16];
            for (int g = 1; g < 16; g += 2) {
                float v = bj[0];
                v = (g & 2) ? v + bj[1] : v - bj[1];
                v = (g & 4) ? v + bj[2] : v - bj[2];
                v = (g & 8) ? v + bj[3] : v - bj[3];
                ent[g] = v;
            }
            for (int g = 0; g < 16; g += 2) ent[g] = -ent[15 - g];
            e0[j] = ent[0];
            for (int g = 0; g < 16; ++g) qlut[k * 128 + j * 16 + g] = sat_round_i8(ent[g] * t_scales);
        }
        biases += ((e0[0] + e0[4]) + (e0[2] + e0[6])) + ((e0[1] + e0[5]) + (e0[3] + e0[7]));
    }
    *lut_scales = scales;
    *lut_biases = biases;
}

/* T-MAC deploy/tuned/kernels.cc:1002-1037  preprocessor_t1_int8_m*_k{K}_n1_b*: per 32 activations
 * partial_max_reset -> partial_max_g4_int8_k8 -> lut_ctor_g4_int8_k0_b*(32, ...) */
void orc_tmac_preprocessor_int8(int K, const float* B, float* lut_scales, float* lut_biases, int8_t* qlut) {
    for (int kk = 0; kk < K / 32; ++kk) {
        lut_scales[kk] = 0.0f;
        orc_partial_max_g4_int8_k8(lut_scales + kk, B + kk * 32);
    }
    for (int kk = 0; kk < K / 32; ++kk) orc_lut_ctor_g4_int8(32, qlut + kk * 128, B + kk * 32, lut_scales + kk, lut_biases + kk);
}

/* T-MAC deploy/tuned/kernels.cc:827-960  qgemm_lut_t1_int8_m{bm}_k{K}_n1_b{bits} with float_type = float, generalised over
 * M: per tile of bm bit-plane rows: tbl_float_reset; for each k tile of (kfactor = 8 tables = 32 activations):
 * tbl_g4_int8_float_update_strue_k8_b{bits}_ak8_fafalse(bm, CBits, LUT + kt*128, A + (tile, kt), Scales + (tile, kt), ...);
 * then C[8j + e] = ((c0 * 0.5 + c1) + c2 * 2) + c3 * 4 over the bit planes (utils.py get_bits_alphas).
 * A: [M*bits/bm][K/32][bm/2 * 8] bytes; Scales: [M*bits/bm][K/32][bm/bits] floats (one scale per output and k tile). */
void orc_tmac_qgemm_lut_int8(int M, int K, int bits, int bm, const uint8_t* A, const int8_t* qlut, const float* scales,
                             const float* lut_scales, const float* lut_biases, float* C) {
    const int tiles = M * bits / bm, kt_n = K / 32;
    const float alphas[4] = {0.5f, 1.f, 2.f, 4.f};
    float cbits[1024];
    for (int t = 0; t < tiles; ++t) {
        memset(cbits, 0, sizeof(float) * bm);
        for (int kt = 0; kt < kt_n; ++kt)
            orc_tbl_g4_int8_float_update(bm, cbits, qlut + kt * 128, A + ((size_t)t * kt_n + kt) * (bm / 2 * 8),
                                         scales + ((size_t)t * kt_n + kt) * (bm / bits), lut_scales + kt, lut_biases + kt, 8, bits, 8, 0, 0);
        for (int j = 0; j < bm / bits / 8; ++j)
            for (int e = 0; e < 8; ++e) {
                float acc = cbits[8 * bits * j + e] * alphas[0];
                for (int bp = 1; bp < bits; ++bp) acc = acc + cbits[8 * bits * j + 8 * bp + e] * alphas[bp];
                C[t * (bm / bits) + 8 * j + e] = acc;
            }
    }
}
