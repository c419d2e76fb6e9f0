// ref_wrap_tmac.cc -- TEST INFRASTRUCTURE ONLY.  Compiles the reference's OWN T-MAC integer-LUT kernels
// (benchmark/kernel/bs_lut_gemm/tbl.cc, lut_ctor.cc: included where they lie under /root/reference, never copied)
// into oracle/_ref/libtmac_ref.so and instantiates the name-mangling macros (tbl.cc:673-691, lut_ctor.cc:223-226)
// for the configurations the parity suite pins.  Used only to validate oracle/tmac_oracle.c and to generate the
// committed fixtures (oracle/gen_golden_tmac.py); nothing under rvv-lut-vq_b200/ links or loads it.
//   build: make -C oracle ref     (g++ -O3 -mavx2 -mfma, AVX2 path => float_type = float)
#include <stdio.h>
#include <stdint.h>
#include "tbl.cc"
#include "lut_ctor.cc"

extern "C" {
// (has_scale, K, Bits, ActK, FastAggregation, ZeroPoint, OneScale)
tbl_g4_int8_int32_update(true, 16, 2, 16, false, false, false)
tbl_g4_int8_int32_update(true, 8, 4, 8, false, false, false)
tbl_g4_int8_float_update(true, 16, 2, 16, false, false, false)   // test_tbl.cc's instance
tbl_g4_int8_float_update(true, 8, 4, 8, false, false, false)     // the one deploy/tuned/kernels.cc:868 calls
tbl_g4_int8_float_update(true, 16, 4, 16, false, true, false)    // ZeroPoint
tbl_g4_int8_float_update(true, 16, 2, 16, true, false, false)    // FastAggregation (halving adder)
lut_ctor(0, 4)
lut_ctor(0, 2)
}
