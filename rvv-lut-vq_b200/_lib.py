"""Loader for liblutvq.so (the C-ABI in include/lutvq.h).  There is NO fallback: if the library is missing
or a call fails the error is raised -- nothing here can compute on the CPU."""
from __future__ import annotations

import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(CSRC, "liblutvq.so")

LUTVQ_OK, LUTVQ_EINVAL, LUTVQ_EUNSUPPORTED, LUTVQ_ECUDA = 0, -1, -2, -3
LUTVQ_F16, LUTVQ_BF16 = 0, 1
FLAG_LUT_FP16, FLAG_NO_PDL, FLAG_X_READY, FLAG_PUSH_EPILOGUE = 1, 2, 4, 8
LUT_VQPY_FP16, LUT_FP32 = 0, 1
ACC_F64, ACC_F32 = 0, 1

P, I, U, SZ = ctypes.c_void_p, ctypes.c_int, ctypes.c_uint, ctypes.c_size_t
U64 = ctypes.c_ulonglong

# name -> (restype, argtypes); must list every symbol include/lutvq.h declares (tests/test_host_logic.py checks)
SIGNATURES = {
    "lutvq_last_error": (ctypes.c_char_p, []),
    "lutvq_version": (I, []),
    "lutvq_device_info": (I, [P, P, P, P]),
    "lutvq_workspace_bytes": (SZ, [I, I]),
    "lutvq_workspace_bytes_ex": (SZ, [I, I, I]),
    "lutvq_packed_workspace_bytes": (SZ, [I, I]),
    "lutvq_code2x8_matvec": (I, [P, P, P, P, I, I, I, P, SZ, P]),
    "lutvq_code2x8_matmat": (I, [P, P, P, P, P, P, I, I, I, I, P, SZ, P]),
    "lutvq_code2x8_matmat_ex": (I, [P, P, P, P, P, P, I, I, I, I, U, P, SZ, P]),
    "lutvq_code2x8_packed_bytes": (SZ, [I, I]),
    "lutvq_code2x8_prepack": (I, [P, P, I, I, P]),
    "lutvq_code2x8_matmat_packed": (I, [P, P, P, P, P, P, I, I, I, I, U, P, SZ, P]),
    "lutvq_group_workspace_bytes": (SZ, [I, P, I]),
    "lutvq_code2x8_matmat_group_packed": (I, [P, I, P, P, P, P, P, P, I, I, U, P, SZ, P]),
    "lutvq_code2x8_matmat_group_packed_swiglu": (I, [P, P, I, P, P, P, P, P, P, I, I, U, P, SZ, P]),
    "lutvq_code2x8_dequant": (I, [P, P, P, P, I, I, I, P]),
    "lutvq_code2x8_matmat_dequant": (I, [P, P, P, P, P, P, I, I, I, I, P, SZ, P]),
    "lutvq_simd_lut_workspace_bytes": (SZ, [I]),
    "lutvq_simd_lut_lookup_add": (I, [P, P, P, I, I, U64, U64, P, SZ, P]),
    "lutvq_code1x16_matvec": (I, [P, P, P, P, I, I, I, I, P]),
    "lutvq_code1x16_matmat": (I, [P, P, P, P, P, P, I, I, I, I, I, P]),
    "lutvq_code1x16_dequant": (I, [P, P, P, P, I, I, I, I, P]),
    "lutvq_tbl_g4_int8_int32_update": (I, [I, P, P, P, I, P]),
    "lutvq_tbl_g4_int8_float_update": (I, [I, P, P, P, P, P, P, I, I, I, I, I, P]),
    "lutvq_tbl_int8_reset": (I, [I, P, P]),
    "lutvq_tbl_int16_reset": (I, [I, P, P]),
    "lutvq_tbl_int32_reset": (I, [I, P, P]),
    "lutvq_tbl_float_reset": (I, [I, P, P]),
    "lutvq_partial_max_g4_int8_k8": (I, [P, P, P]),
    "lutvq_partial_max_reset": (I, [P, P]),
    "lutvq_lut_ctor_g4_int8": (I, [I, P, P, P, P, P]),
    "lutvq_noalloc_scratch_bytes": (SZ, []),
    "lutvq_code2x8_matmat_noalloc": (I, [P, P, P, P, P, P, I, I, I, I, P]),
    "lutvq_pq4_pack_codes": (I, [P, SZ, SZ, SZ, SZ, SZ, P, P]),
    "lutvq_pq4_pack_codes_range": (I, [P, SZ, SZ, SZ, SZ, SZ, P, P]),
    "lutvq_pq4_pack_LUT": (I, [I, I, P, P, P]),
    "lutvq_pq4_qbs_to_nq": (I, [I]),
    "lutvq_pq4_pack_LUT_qbs": (I, [I, I, P, P, P, P]),
    "lutvq_pq4_accumulate_loop": (I, [I, SZ, I, I, P, P, P, SZ, I, P]),
    "lutvq_pq4_accumulate_loop_qbs": (I, [I, SZ, I, P, P, P, SZ, I, P]),
    "lutvq_tmac_preprocessor_int8": (I, [I, P, P, P, P, P]),
    "lutvq_tmac_qgemm_lut_int8": (I, [I, I, I, I, P, P, P, P, P, P, P]),
    "lutvq_codeKx8_workspace_bytes": (SZ, [I, I, I]),
    "lutvq_codeKx8_matmat": (I, [P, P, P, P, P, P, I, I, I, I, I, P, SZ, P]),
    "lutvq_codeKx8_dequant": (I, [P, P, P, P, I, I, I, I, P]),
    "lutvq_lut_build": (I, [P, P, P, I, I, I, I, I, I, P]),
    "lutvq_lut_gemv": (I, [P, P, P, I, I, I, I, I, I, P]),
    "lutvq_vq_dequantize": (I, [P, P, P, I, I, I, I, I, I, P]),
    "lutvq_shard_create": (I, [P, I, I, I, SZ]),
    "lutvq_shard_export": (I, [P, P]),
    "lutvq_shard_attach": (I, [P, P]),
    "lutvq_shard_attach_ptrs": (I, [P, P]),
    "lutvq_shard_base": (P, [P]),
    "lutvq_shard_slot": (P, [P, I]),
    "lutvq_shard_slot_bytes": (SZ, [P]),
    "lutvq_sharded_matmat_packed": (I, [P, I, SZ, P, P, P, P, P, I, I, I, U, P, SZ, P]),
    "lutvq_sharded_matmat_group_packed": (I, [P, I, P, P, P, P, P, P, P, P, I, I, U, P, SZ, P]),
    "lutvq_sharded_matmat_group_packed_swiglu": (I, [P, I, P, P, P, P, P, P, P, P, P, I, I, U, P, SZ, P]),
    "lutvq_shard_wait": (I, [P, I, P]),
    "lutvq_shard_wait_mask": (I, [P, U64, P]),
    "lutvq_shard_push_mask": (I, [P, U64, P]),
    "lutvq_shard_error": (I, [P, P]),
    "lutvq_shard_destroy": (I, [P]),
    "lutvq_shard_ll_tick": (I, [P, P]),
    "lutvq_sharded_matmat_group_ll": (I, [P, I, P, P, P, P, I, U, I, U, P, P, P, P, P, I, I, U, P, SZ, P]),
    "lutvq_shard_ll_unpack": (I, [P, I, U, P, I, P]),
    "lutvq_decode_add_rmsnorm_ll": (I, [P, P, I, U, P, P, I, ctypes.c_float, P]),
    "lutvq_decode_attention_ll": (I, [P, I, I, I, U, U, U, P, P, P, P, I, I, I, I, ctypes.c_float, P]),
    "lutvq_decode_add_rmsnorm_ll_out": (I, [P, P, I, U, P, P, I, U, I, ctypes.c_float, P]),
    "lutvq_decode_attention_ll_out": (I, [P, I, I, I, U, U, U, P, P, P, P, I, U, I, I, I, I, ctypes.c_float, P]),
    "lutvq_decode_add_rmsnorm": (I, [P, P, P, P, I, ctypes.c_float, P]),
    "lutvq_decode_silu_mul": (I, [P, P, P, I, P]),
    "lutvq_decode_attention": (I, [P, P, P, P, P, P, P, I, I, I, I, ctypes.c_float, P]),
}
DEBUG_SIGNATURES = {
    "lutvq_debug_set_gemv_tuning": (None, [I, I, I]),
    "lutvq_debug_set_dependent_workers": (None, [I]),
    "lutvq_debug_get_gemv_plan": (None, [I, I, P, P, P]),
    "lutvq_debug_set_gemm_mode": (None, [I]),
    "lutvq_debug_set_gemm_stamps": (None, [P]),
    "lutvq_debug_set_pq4_mode": (None, [I]),
    "lutvq_debug_set_1x16_mode": (None, [I]),
    "lutvq_debug_set_1x16_plan": (None, [I, I]),
}


class LutvqError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"liblutvq error {code}: {msg}")
        self.code = code


def build(force: bool = False) -> str:
    """Compile liblutvq.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", CSRC, "-j8"] + (["-B"] if force else [])
    subprocess.check_call(cmd, stdout=subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)")
        _lib = ctypes.CDLL(LIB_PATH)
        for table in (SIGNATURES, DEBUG_SIGNATURES):
            for name, (res, args) in table.items():
                fn = getattr(_lib, name)
                fn.restype, fn.argtypes = res, args
        if os.environ.get("LUTVQ_DEP_WORKERS"):          # experiment switch: workers per SM of a dependent GEMV launch
            _lib.lutvq_debug_set_dependent_workers(int(os.environ["LUTVQ_DEP_WORKERS"]))
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise LutvqError(rc, lib().lutvq_last_error().decode())
