#!/usr/bin/env python
"""Generate tests/golden/*.npz by running THE REFERENCE'S OWN CODE in the build container.

Run once here (needs /root/reference; it does not exist on the GPU box):
    python oracle/gen_golden.py
The fixtures are small and committed; tests/test_oracle_golden.py replays them against the oracle
and tests/test_gpu_parity.py against the CUDA path.  Nothing at test/bench time reads /root/reference.

Reference functions executed (paths relative to /root/reference):
  vq_dataflow_sim/vq.py                  VQ.__init__/compute_lut/compute_lut_gemm/dequantize/compute_fp_gemm
  benchmark/AQLM/inference_lib/src/aqlm  utils.pack_int_data/unpack_int_data/_dequantize_weight,
                                         inference_kernels.numba_kernel.numba_gemm_lut
  benchmark/kernel/bs_lut_gemm           lut_ctor.cc / tbl.cc through oracle/_ref/libtmac_ref.so (if built)
"""
import contextlib
import io
import os
import sys

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def import_vq():
    sys.path.insert(0, os.path.join(REF, "vq_dataflow_sim"))
    with contextlib.redirect_stdout(io.StringIO()):   # module-level demo prints (vq.py:311-329)
        import vq  # noqa
    return vq


def gen_vqpy(vq):
    """Per-subvector-codebook configs through the golden model itself (slow Python loops: keep small)."""
    cases = {
        # name: (D, M, N, K, d_out, batch, seed)
        "vq_d128_m16_n2_k256": (128, 16, 2, 256, 40, 2, 42),
        "vq_d64_m8_n1_k256": (64, 8, 1, 256, 24, 1, 7),
        "vq_d64_m8_n4_k256": (64, 8, 4, 256, 16, 1, 11),
        "vq_d96_m6_n2_k256_d16": (96, 6, 2, 256, 16, 1, 3),     # subvector dim 16
    }
    for name, (D, M, N, K, d_out, B, seed) in cases.items():
        with contextlib.redirect_stdout(io.StringIO()):
            m = vq.VQ((D, M, N, K), d_out=d_out, seed=seed)
        x = np.random.randn(B, D).astype(np.float16)          # continues the seeded stream, like vq.py:322
        out = m.compute_lut_gemm(x)                            # float64 (B, d_out); leaves m.lut = LUT of last row
        lut_last = m.lut.copy()
        fp = m.compute_fp_gemm(x.astype(np.float32))
        np.savez_compressed(
            os.path.join(OUT, name + ".npz"),
            cfg=np.array([D, M, N, K, d_out, B, seed]),
            codebook=m.codebook, codeword=m.codeword, x=x,
            lut_last=lut_last, out_lut_gemm=out, weight_mat=m.weight_mat, out_fp_gemm=fp,
        )
        print(f"[vq.py] {name}: lut {lut_last.shape} out {out.shape} |lut-fp|max = {np.abs(out - fp).max():.4f}")


def gen_aqlm():
    sys.path.insert(0, os.path.join(REF, "benchmark/AQLM/inference_lib/src"))
    import numba
    import torch
    from aqlm.inference_kernels.numba_kernel import numba_gemm_lut
    from aqlm.utils import _dequantize_weight, pack_int_data, unpack_int_data

    numba.set_num_threads(1)   # the reference kernel races at >1 thread
    torch.set_num_threads(1)
    cases = {
        # name: (in_features, out_features, num_codebooks, batch, seed, with_bias)
        "aqlm_2x8_i256_o96": (256, 96, 2, 3, 0, True),
        "aqlm_2x8_i1024_o64": (1024, 64, 2, 1, 1, False),
        "aqlm_1x8_i128_o32": (128, 32, 1, 2, 2, True),
        "aqlm_4x8_i128_o48": (128, 48, 4, 1, 3, False),
    }
    for name, (I, O, N, B, seed, with_bias) in cases.items():
        torch.manual_seed(seed)
        x = torch.randn((B, I), dtype=torch.float32)
        idx = torch.randint(256, (O, I // 8, N))
        codes = pack_int_data(idx.clone(), 8)                                   # int8 [O, G, N]
        assert torch.equal(unpack_int_data(codes, 8), idx)
        codebooks = torch.randn((N, 256, 1, 8), dtype=torch.float32)
        scales = torch.randn((O, 1, 1, 1), dtype=torch.float32)
        bias = torch.randn(O, dtype=torch.float32) if with_bias else None
        # fp16-representable operands so the same fixture drives the fp16 CUDA path exactly
        x, codebooks, scales = (t.half().float() for t in (x, codebooks, scales))
        if bias is not None:
            bias = bias.half().float()
        w32 = _dequantize_weight(unpack_int_data(codes, 8), codebooks, scales).contiguous()
        w16_noscale = _dequantize_weight(unpack_int_data(codes, 8), codebooks.half(), None).contiguous()   # fp16 embedding_bag sum
        w16_scaled = _dequantize_weight(unpack_int_data(codes, 8), codebooks.half(), scales.half()).contiguous()
        codes_alt = torch.permute(codes, (1, 0, 2)).contiguous()                # CPU layout (inference.py:78-83)
        y = numba_gemm_lut(x, codes_alt, codebooks, scales, bias)
        y_dense = torch.nn.functional.linear(x, w32, bias)
        np.savez_compressed(
            os.path.join(OUT, name + ".npz"),
            cfg=np.array([I, O, N, B, seed, int(with_bias)]),
            x=x.numpy(), codes=codes.numpy(), idx=idx.numpy().astype(np.uint8), codebooks=codebooks.numpy(),
            scales=scales.numpy(), bias=(bias.numpy() if bias is not None else np.zeros(0, np.float32)),
            weight_f32=w32.numpy(), weight_f16_noscale=w16_noscale.numpy(), weight_f16_scaled=w16_scaled.numpy(),
            y_numba=y.numpy(), y_dense=y_dense.numpy(),
        )
        rel = (y - y_dense).abs().mean() / y_dense.abs().mean()
        print(f"[aqlm] {name}: numba-vs-dense rel {rel:.2e}")


def codebook_1x16(seed: int, gs: int) -> np.ndarray:
    """The 65536-entry codebook of a 1x16 fixture, regenerated from its seed (PCG64 stream: stable across numpy versions;
    the fixture stores a checksum) instead of committing 1-2 MiB of random halves per case."""
    return np.random.default_rng(seed).standard_normal((1, 65536, 1, gs)).astype(np.float16)


def gen_aqlm_1x16():
    """AQLM 1x16 (one 65536-entry codebook, int16 codes; inference.py:39-61) through the reference's own
    pack_int_data / unpack_int_data / _dequantize_weight (aqlm/utils.py:11-70)."""
    sys.path.insert(0, os.path.join(REF, "benchmark/AQLM/inference_lib/src"))
    import torch
    from aqlm.utils import _dequantize_weight, pack_int_data, unpack_int_data
    torch.set_num_threads(1)
    cases = {"aqlm_1x16_g8_i256_o48": (256, 48, 8, 3, 21, True), "aqlm_1x16_g16_i512_o40": (512, 40, 16, 2, 22, False)}
    for name, (I, O, gs, B, seed, with_bias) in cases.items():
        torch.manual_seed(seed)
        x = torch.randn((B, I)).half()
        idx = torch.randint(65536, (O, I // gs, 1))
        idx[0, 0, 0], idx[1, 0, 0] = 65535, 32768                              # exercise the sign wrap of pack_int_data
        codes = pack_int_data(idx.clone(), 16)                                 # int16 [O, G, 1]
        assert codes.dtype == torch.int16 and torch.equal(unpack_int_data(codes, 16), idx)
        cb = torch.from_numpy(codebook_1x16(seed, gs))
        scales = torch.randn((O, 1, 1, 1)).half()
        bias = torch.randn(O).half() if with_bias else None
        w16_noscale = _dequantize_weight(unpack_int_data(codes, 16), cb, None).contiguous()
        w16_scaled = _dequantize_weight(unpack_int_data(codes, 16), cb, scales).contiguous()
        w32 = _dequantize_weight(unpack_int_data(codes, 16), cb.float(), scales.float()).contiguous()
        y_dense = torch.nn.functional.linear(x.float(), w32, None if bias is None else bias.float())
        np.savez_compressed(
            os.path.join(OUT, name + ".npz"),
            cfg=np.array([I, O, gs, B, seed, int(with_bias)]),
            x=x.numpy(), codes=codes.numpy(), idx=idx.numpy().astype(np.uint16), codebook_checksum=np.array([int(cb.numpy().view(np.uint16).astype(np.uint64).sum())]),
            scales=scales.numpy(), bias=(bias.numpy() if bias is not None else np.zeros(0, np.float16)),
            weight_f16_noscale=w16_noscale.numpy(), weight_f16_scaled=w16_scaled.numpy(), y_dense=y_dense.numpy(),
        )
        print(f"[aqlm 1x16] {name}: W {tuple(w16_scaled.shape)}")


def main():
    os.makedirs(OUT, exist_ok=True)
    vq = import_vq()
    gen_vqpy(vq)
    gen_aqlm()
    gen_aqlm_1x16()
    try:
        from gen_golden_tmac import gen_tmac
        gen_tmac(OUT)
    except ImportError:
        pass
    sz = sum(os.path.getsize(os.path.join(OUT, f)) for f in os.listdir(OUT))
    print(f"golden dir: {len(os.listdir(OUT))} files, {sz/1024:.0f} KiB")


if __name__ == "__main__":
    main()
