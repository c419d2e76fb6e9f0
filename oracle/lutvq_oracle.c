/*
 * lutvq_oracle.c -- CPU restatement of the reference's LUT-VQ hot path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it.  The product path
 * (rvv-lut-vq_b200/, liblutvq.so) never links or calls anything in oracle/.
 *
 * Parity status: PINNED.  Every function below is checked bit-for-bit (or within the
 * stated tolerance) against the reference's own Python code run in the build container
 * (oracle/gen_golden.py imports /root/reference in place and writes tests/golden/*.npz;
 * tests/test_oracle_golden.py replays them without the reference).
 *
 * Reference anchors (paths relative to /root/reference):
 *   vq_dataflow_sim/vq.py:280-285    VQ.compute_lut        -> orc_lut_build_vqpy
 *   vq_dataflow_sim/vq.py:287-302    VQ.compute_lut_gemm   -> orc_lut_gemv_f64
 *   vq_dataflow_sim/vq.py:269-278    VQ.dequantize         -> orc_dequantize
 *   vq_dataflow_sim/vq.py:304-307    VQ.compute_fp_gemm    -> orc_dense_gemv_f64
 *   benchmark/AQLM/inference_lib/src/aqlm/inference_kernels/numba_kernel.py:38-48
 *                                    numba_gemv_lut_       -> orc_aqlm_gemv_f32
 *   benchmark/AQLM/inference_lib/src/aqlm/utils.py:24-31   pack/unpack_int_data -> orc_unpack_codes
 *   benchmark/AQLM/inference_lib/src/aqlm/inference_kernels/cuda_kernel.cpp:95-111
 *                                    scale_bias_unflatten_output -> orc_scale_bias_epilogue
 *   benchmark/kernel/simd_lut/bench_avx_lut.cpp:98-107     scalar lookup_and_add -> orc_simd_lut_scalar
 *
 * Build: make -C oracle   (gcc -O3 -fopenmp -shared)
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <math.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef _Float16 f16;

static inline float h2f(uint16_t h) { f16 v; memcpy(&v, &h, 2); return (float)v; }
static inline uint16_t f2h(float f) { f16 v = (f16)f; uint16_t h; memcpy(&h, &v, 2); return h; }
static inline uint16_t d2h(double d) { f16 v = (f16)d; uint16_t h; memcpy(&h, &v, 2); return h; }

static inline float bf2f(uint16_t b) { uint32_t u = (uint32_t)b << 16; float f; memcpy(&f, &u, 4); return f; }
static inline uint16_t f2bf(float f) {
    uint32_t u; memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);   /* quiet NaN */
    u += 0x7fffu + ((u >> 16) & 1u);                                           /* RN-even */
    return (uint16_t)(u >> 16);
}

/* dtype: 0 = fp16, 1 = bf16 (the two dtypes cuda_kernel.cpp:9-25 accepts) */
static inline float ld16(uint16_t v, int dtype) { return dtype ? bf2f(v) : h2f(v); }
static inline uint16_t st16(float v, int dtype) { return dtype ? f2bf(v) : f2h(v); }

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void orc_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* numpy's HALF_pairwise_sum for n <= 128 (numpy/_core/src/umath/loops_utils.h.src), which is
 * what `lut.sum(axis=-1)` in vq.py:285 runs per LUT entry: fp16 inputs widened to fp32,
 * 8 running lanes, then the fixed tree ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)). */
static float np_half_pairwise_sum(const uint16_t* a, int n) {
    if (n < 8) {
        float res = 0.f;
        for (int i = 0; i < n; ++i) res += h2f(a[i]);
        return res;
    }
    float r[8];
    for (int i = 0; i < 8; ++i) r[i] = h2f(a[i]);
    int i;
    for (i = 8; i < n - (n % 8); i += 8)
        for (int j = 0; j < 8; ++j) r[j] += h2f(a[i + j]);
    float res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
    for (; i < n; ++i) res += h2f(a[i]);
    return res;
}

/* vq.py:280-285.  x fp16 [M*d]; codebook fp16 [M][N][K][d] (cb_shared=0) or [N][K][d]
 * broadcast over M (cb_shared=1: the AQLM case, SURVEY 8 "AQLM subset of vq.py");
 * lut fp16 [M][N][K].  Products are rounded to fp16 one by one (numpy fp16 multiply),
 * then summed with np_half_pairwise_sum and rounded once to fp16 (0 + sum, as HALF_add reduce). */
void orc_lut_build_vqpy(const uint16_t* x, const uint16_t* codebook, int cb_shared,
                        int M, int N, int K, int d, uint16_t* lut) {
    #pragma omp parallel for schedule(static)
    for (int m = 0; m < M; ++m) {
        uint16_t prod[128];
        for (int n = 0; n < N; ++n)
            for (int k = 0; k < K; ++k) {
                const uint16_t* c = codebook + ((size_t)(cb_shared ? 0 : m) * N + n) * K * d + (size_t)k * d;
                for (int j = 0; j < d; ++j) {
                    f16 a, b; memcpy(&a, &c[j], 2); memcpy(&b, &x[m * d + j], 2);
                    f16 p = (f16)((float)a * (float)b);        /* fp16*fp16 -> exact in fp32 -> RN to fp16 */
                    memcpy(&prod[j], &p, 2);
                }
                float s = 0.f + np_half_pairwise_sum(prod, d);
                lut[((size_t)m * N + n) * K + k] = f2h(s);
            }
    }
}

/* fp32 LUT (numba_kernel.py:39-40): lut[g][n][k] = sum_j x[g*d+j]*C[n][k][j], fp32, j ascending.
 * (numba calls BLAS there; summation order is unspecified, parity is by tolerance.) */
void orc_lut_build_f32(const float* x, const float* codebook, int G, int N, int K, int d, float* lut) {
    #pragma omp parallel for schedule(static)
    for (int g = 0; g < G; ++g)
        for (int n = 0; n < N; ++n)
            for (int k = 0; k < K; ++k) {
                const float* c = codebook + ((size_t)n * K + k) * d;
                float s = 0.f;
                for (int j = 0; j < d; ++j) s += x[g * d + j] * c[j];
                lut[((size_t)g * N + n) * K + k] = s;
            }
}

/* vq.py:293-298.  out[j] = sum_m sum_n lut[m][n][codeword[j][m][n]] accumulated in float64
 * (exact for fp16 addends at these sizes). codeword uint8 [d_out][M][N]. */
void orc_lut_gemv_f64(const uint16_t* lut, const uint8_t* codeword, int d_out, int M, int N, int K,
                      double* out) {
    #pragma omp parallel for schedule(static)
    for (int j = 0; j < d_out; ++j) {
        double acc = 0.0;
        const uint8_t* cw = codeword + (size_t)j * M * N;
        for (int m = 0; m < M; ++m)
            for (int n = 0; n < N; ++n)
                acc += (double)h2f(lut[((size_t)m * N + n) * K + cw[m * N + n]]);
        out[j] = acc;
    }
}

/* vq.py:269-278 (+ cuda_kernel.cu:283-289 semantics for N=2: one RN add in the storage dtype).
 * W[j][m*d + i] = round_dtype( sum_n codebook[m][n][code[j][m][n]][i] ), float64 sum.
 * If scales != NULL: W = round_dtype(W * scales[j])   (cuda_kernel.cpp:445, a separate rounding). */
void orc_dequantize(const uint16_t* codebook, int cb_shared, const uint8_t* codeword,
                    const uint16_t* scales, int d_out, int M, int N, int K, int d, int dtype,
                    uint16_t* W) {
    #pragma omp parallel for schedule(static)
    for (int j = 0; j < d_out; ++j)
        for (int m = 0; m < M; ++m)
            for (int i = 0; i < d; ++i) {
                double s = 0.0;
                for (int n = 0; n < N; ++n) {
                    int k = codeword[((size_t)j * M + m) * N + n];
                    s += (double)ld16(codebook[(((size_t)(cb_shared ? 0 : m) * N + n) * K + k) * d + i], dtype);
                }
                uint16_t w = dtype ? f2bf((float)s) : d2h(s);
                if (scales) w = st16(ld16(w, dtype) * ld16(scales[j], dtype), dtype);
                W[(size_t)j * M * d + m * d + i] = w;
            }
}

/* vq.py:304-307 in float64: out[j] = sum_i x[i] * W[j][i]  (x, W in storage dtype) */
void orc_dense_gemv_f64(const uint16_t* x, const uint16_t* W, int d_out, int D, int dtype, double* out) {
    #pragma omp parallel for schedule(static)
    for (int j = 0; j < d_out; ++j) {
        double acc = 0.0;
        for (int i = 0; i < D; ++i) acc += (double)ld16(x[i], dtype) * (double)ld16(W[(size_t)j * D + i], dtype);
        out[j] = acc;
    }
}

/* Exact (float64) evaluation of the AQLM GEMV from the raw operands, the y64 that SURVEY 8(c)
 * measures tolerance against: y64[o] = sum_g sum_n sum_j C[n][idx(o,g,n)][j] * x[g*d+j].
 * codes are int8 holding unsigned indices (utils.py:24-31), layout [O][G][N]. */
void orc_aqlm_gemv_f64(const uint16_t* x, const int8_t* codes, const uint16_t* codebooks,
                       int O, int G, int N, int d, int dtype, double* out) {
    #pragma omp parallel for schedule(static)
    for (int o = 0; o < O; ++o) {
        double acc = 0.0;
        for (int g = 0; g < G; ++g)
            for (int n = 0; n < N; ++n) {
                int k = (uint8_t)codes[((size_t)o * G + g) * N + n];
                const uint16_t* c = codebooks + ((size_t)n * 256 + k) * d;
                for (int j = 0; j < d; ++j) acc += (double)ld16(c[j], dtype) * (double)ld16(x[g * d + j], dtype);
            }
        out[o] = acc;
    }
}

/* numba_kernel.py:38-48 restated: fp32 LUT, then for every output i: sum over groups j and
 * codebooks c of lut[j][c][codes_alt[j][i][c]], fp32, scaled last.  codes_alt is the CPU
 * layout [G][O][N] (inference.py:78-83 permutes to it).  threads=1 follows the reference's
 * only race-free setting; this port parallelises over OUTPUT rows instead,
 * which is race-free at any thread count and keeps the per-row j-ascending order. */
void orc_aqlm_gemv_f32(const float* x, const uint8_t* codes_alt, const float* codebooks, const float* scales,
                       const float* bias, int O, int G, int N, int d, float* lut_scratch, float* out) {
    orc_lut_build_f32(x, codebooks, G, N, 256, d, lut_scratch);
    #pragma omp parallel for schedule(static)
    for (int i = 0; i < O; ++i) {
        float acc = 0.f;
        for (int j = 0; j < G; ++j)
            for (int c = 0; c < N; ++c)
                acc += lut_scratch[((size_t)j * N + c) * 256 + codes_alt[((size_t)j * O + i) * N + c]];
        acc *= scales[i];
        if (bias) acc += bias[i];
        out[i] = acc;
    }
}

/* Same math on the GPU-side layout codes[O][G][N] (what QuantizedLinear holds on CUDA,
 * inference.py:47-54) -- the CPU baseline bench.py times: fp16 operands widened once, fp32 LUT,
 * fp32 gather-accumulate, then the reference's epilogue rounding (fp16(acc) -> *scale -> +bias). */
void orc_aqlm_gemv_rowmajor(const uint16_t* x, const int8_t* codes, const uint16_t* codebooks,
                            const uint16_t* scales, const uint16_t* bias, int O, int G, int N, int d,
                            int dtype, float* lut_scratch, uint16_t* y) {
    #pragma omp parallel for schedule(static)
    for (int g = 0; g < G; ++g)
        for (int n = 0; n < N; ++n)
            for (int k = 0; k < 256; ++k) {
                const uint16_t* c = codebooks + ((size_t)n * 256 + k) * d;
                float s = 0.f;
                for (int j = 0; j < d; ++j) s += ld16(x[g * d + j], dtype) * ld16(c[j], dtype);
                lut_scratch[((size_t)g * N + n) * 256 + k] = s;
            }
    #pragma omp parallel for schedule(static)
    for (int o = 0; o < O; ++o) {
        const uint8_t* cw = (const uint8_t*)codes + (size_t)o * G * N;
        float acc = 0.f;
        for (int t = 0; t < G * N; ++t) acc += lut_scratch[(size_t)t * 256 + cw[t]];
        uint16_t r = st16(acc, dtype);
        if (scales) r = st16(ld16(r, dtype) * ld16(scales[o], dtype), dtype);
        if (bias) r = st16(ld16(r, dtype) + ld16(bias[o], dtype), dtype);
        y[o] = r;
    }
}

/* cuda_kernel.cpp:95-111: flat_output *= scales; += bias, each an RN op in the storage dtype,
 * applied to an already-rounded output. */
void orc_scale_bias_epilogue(uint16_t* y, const uint16_t* scales, const uint16_t* bias, int batch, int O,
                             int dtype) {
    for (int b = 0; b < batch; ++b)
        for (int o = 0; o < O; ++o) {
            uint16_t r = y[(size_t)b * O + o];
            if (scales) r = st16(ld16(r, dtype) * ld16(scales[o], dtype), dtype);
            if (bias) r = st16(ld16(r, dtype) + ld16(bias[o], dtype), dtype);
            y[(size_t)b * O + o] = r;
        }
}

/* utils.py:29-31 unpack_int_data for nbits=8: idx = int8 value mod 256 */
void orc_unpack_codes(const int8_t* codes, size_t n, uint8_t* idx) {
    for (size_t i = 0; i < n; ++i) idx[i] = (uint8_t)codes[i];
}

/* bench_avx_lut.cpp:98-107 (_SCALAR_ build): output[k] += lut[idx[j*CB+k]], int16 wraparound,
 * repeated n_runs times. */
void orc_simd_lut_scalar(const uint8_t* idx, const int8_t* lut, int16_t* output, int cb_size,
                         uint64_t n_runs, uint64_t n_lut_per_run) {
    for (uint64_t i = 0; i < n_runs; ++i)
        for (uint64_t j = 0; j < n_lut_per_run; ++j)
            for (int k = 0; k < cb_size; ++k)
                output[k] = (int16_t)(output[k] + lut[idx[j * cb_size + k]]);
}

/* Batched dense reference for the prefill path (cuda_kernel.cpp:450-484 = dequant + F.linear):
 * y64[b][o] = sum_i x[b][i] * W[o][i], float64, W already dequantized in the storage dtype. */
void orc_dense_gemm_f64(const uint16_t* x, const uint16_t* W, int B, int O, int I, int dtype, double* out) {
    #pragma omp parallel for schedule(static) collapse(2)
    for (int b = 0; b < B; ++b)
        for (int o = 0; o < O; ++o) {
            double acc = 0.0;
            for (int i = 0; i < I; ++i)
                acc += (double)ld16(x[(size_t)b * I + i], dtype) * (double)ld16(W[(size_t)o * I + i], dtype);
            out[(size_t)b * O + o] = acc;
        }
}
