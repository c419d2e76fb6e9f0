// ref_wrap_simd_lut.cc -- TEST INFRASTRUCTURE ONLY.  The reference's lookup_and_add_simd (benchmark/kernel/simd_lut/
// bench_avx_lut.cpp:26-138), _SCALAR_ build, compiled from where it lies into oracle/_ref/libsimdlut_ref.so.  The file has
// its own main(); it is renamed away.  The scalar build fixes CB_SIZE = 16 (bench_avx_lut.cpp:21-23).
#define _SCALAR_
#define main ref_bench_main
#include "bench_avx_lut.cpp"
#undef main

extern "C" void ref_lookup_and_add_scalar(uint8_t* idx, int8_t* lut, int16_t* output, uint64_t n_runs, uint64_t n_lut_per_run) {
    lookup_and_add_simd(idx, lut, output, n_runs, n_lut_per_run);
}
extern "C" int ref_simd_lut_cb_size() { return CB_SIZE; }
