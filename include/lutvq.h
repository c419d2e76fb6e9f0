/*
 * lutvq.h -- C-ABI of liblutvq.so: the B200-native (sm_100a) LUT vector-quantization hot path.
 *
 * Drop-in boundary for the reference's kernel entry points (paths relative to the reference tree
 * wh-xu/RVV-LUT-VQ; "aqlm/" = benchmark/AQLM/inference_lib/src/aqlm/inference_kernels/):
 *
 *   lutvq_code2x8_matvec          <- code2x8_matvec_cuda<bf16>(A,B,C,codebook,prob_m,prob_k)   aqlm/cuda_kernel.cu:555-563
 *                                    + code2x8_matvec(A,B,C,codebook,use_bfloat16)             aqlm/cuda_kernel.cpp:356-385
 *   lutvq_code2x8_matmat          <- code2x8_matmat(input,codes,codebooks,scales,bias)         aqlm/cuda_kernel.cpp:387-421
 *                                    + scale_bias_unflatten_output                             aqlm/cuda_kernel.cpp:95-111
 *   lutvq_code2x8_dequant         <- code2x8_dequant_cuda(A,C,codebook,prob_m,prob_k,bf16)     aqlm/cuda_kernel.cu:622-629
 *                                    + code2x8_dequant(codes,codebooks,scales)                 aqlm/cuda_kernel.cpp:423-448
 *   lutvq_code2x8_matmat_dequant  <- code2x8_matmat_dequant(input,codes,codebooks,scales,bias) aqlm/cuda_kernel.cpp:450-484
 *   lutvq_lut_build               <- VQ.compute_lut                                            vq_dataflow_sim/vq.py:280-285
 *                                    (and the LUT line of numba_gemv_lut_, aqlm/numba_kernel.py:39-40)
 *   lutvq_lut_gemv                <- VQ.compute_lut_gemm inner loops                           vq_dataflow_sim/vq.py:293-298
 *                                    (and numba_gemv_lut_ gather loop, aqlm/numba_kernel.py:43-46)
 *   lutvq_vq_dequantize           <- VQ.dequantize (per-subvector codebooks)                   vq_dataflow_sim/vq.py:269-278
 *
 * Conventions (differences from the reference launchers are deliberate and listed in INTEGRATION.md):
 *   - every entry point is extern "C", takes plain device pointers + sizes, an explicit CUDA stream
 *     (cudaStream_t passed as void*), never allocates, never synchronises the host, and runs on the
 *     CURRENT device of the calling thread (the reference queries device 0, cuda_kernel.cu:486,497);
 *   - returns 0 on success, a negative LUTVQ_E* code otherwise; lutvq_last_error() gives a thread-local
 *     message.  There is no CPU fallback: without a CUDA device every compute call fails with LUTVQ_ECUDA;
 *   - scratch memory is passed in (ws, ws_bytes); size it with lutvq_workspace_bytes().  A workspace must
 *     be zero-filled once before first use (cudaMemset); every call leaves its first lutvq_packed_workspace_bytes()
 *     bytes (the split-k words) all-zero again, so it may be reused by any later call ON THE SAME STREAM (any shape).
 *     (The GEMV needs exactly those bytes for packed and reference-layout codes alike; the rest of lutvq_workspace_bytes()
 *     is the split-k scratch of the batch > 6 tensor-core path);
 *   - dtype is the storage/arithmetic type of x, codebooks, scales, bias, y (fp16 or bf16, the two types
 *     cuda_kernel.cpp:9-25 accepts).  Codes are int8 holding unsigned 8-bit indices
 *     (aqlm/../utils.py:24-31), layout [prob_m][prob_k/8][num_codebooks] exactly as QuantizedLinear.codes.
 */
#ifndef LUTVQ_H_
#define LUTVQ_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LUTVQ_OK            0
#define LUTVQ_EINVAL       (-1)   /* bad pointer / shape / alignment / workspace too small */
#define LUTVQ_EUNSUPPORTED (-2)   /* valid request this build does not implement (e.g. dtype, codebook count) */
#define LUTVQ_ECUDA        (-3)   /* CUDA runtime error (message carries cudaGetErrorString) */

typedef enum { LUTVQ_F16 = 0, LUTVQ_BF16 = 1 } lutvq_dtype_t;

/* flags for the *_ex entry points */
#define LUTVQ_FLAG_LUT_FP16   1u  /* round LUT entries to fp16 (vq.py:285 semantics) before accumulation */
#define LUTVQ_FLAG_NO_PDL     2u  /* do not use programmatic dependent launch for this call */
#define LUTVQ_FLAG_X_READY    4u  /* caller's promise: (1) x was complete BEFORE the run of back-to-back launches this call belongs
                                     to began -- i.e. it was not written by any kernel still in flight when this one may start.
                                     Launches overlap under programmatic dependent launch, so "not the kernel right before"
                                     is NOT enough: with producer -> A -> B(X_READY), B can start while the producer of x is
                                     still running.  Valid: a microbenchmark loop over resident inputs, weights-only
                                     prefetch runs, inputs produced before a stream/event synchronisation.
                                     (2) ws is not used by any of the LUTVQ_XREADY_WS_RING - 1 calls before or after this one
                                     in the stream (rotate through a ring of workspaces).
                                     The LUT build, the gather and the split-k reduction then overlap the previous kernels;
                                     the y stores and the completion of the call stay stream-ordered. */
#define LUTVQ_FLAG_PUSH_EPILOGUE 8u /* lutvq_sharded_matmat_packed only: push the rows to the peers from the GEMV's own epilogue
                                     (fused variant) instead of from the exchange kernel behind lutvq_shard_wait */
#define LUTVQ_XREADY_WS_RING  8

/* lutvq_lut_build lut_mode */
#define LUTVQ_LUT_VQPY_FP16   0   /* bit-exact vq.py: fp16 products, numpy pairwise fp32 sum, fp16 entries */
#define LUTVQ_LUT_FP32        1   /* fp32 products and sum (numba semantics), fp32 entries */
/* lutvq_lut_gemv acc_mode */
#define LUTVQ_ACC_F64         0   /* float64 accumulate, float64 out (vq.py:297; exact for fp16 entries) */
#define LUTVQ_ACC_F32         1   /* float32 accumulate, float32 out (numba_kernel.py:46) */

const char* lutvq_last_error(void);
int         lutvq_version(void);                     /* 10000*major + 100*minor + patch */
/* fills what the launch heuristics see; any pointer may be NULL */
int         lutvq_device_info(int* sm_count, int* cc_major, int* cc_minor, size_t* smem_optin_bytes);

/* Scratch needed by matvec / matmat / matmat_dequant for an [prob_m x prob_k] layer (any batch). */
size_t lutvq_workspace_bytes(int prob_m, int prob_k);
/* Scratch needed by the *_packed / group / sharded entry points: the split-k words only (8 bytes per output row + 256). */
size_t lutvq_packed_workspace_bytes(int prob_m, int prob_k);
/* Same, when the caller knows its largest batch: batch <= 6 only ever needs the (small) GEMV scratch. */
size_t lutvq_workspace_bytes_ex(int prob_m, int prob_k, int max_batch);

/* y[prob_m] = W(codes, codebook) . x[prob_k]            (no scale, like Code2x8MatVec)
 * codes    int8  [prob_m][prob_k/8][2]   (16-byte aligned rows: prob_k % 64 == 0, as the reference requires)
 * x        dtype [prob_k]                 (16-byte aligned: Code2x8MatVec reads it as int4 too, cuda_kernel.cu:147)
 * y        dtype [prob_m]
 * codebook dtype [2][256][8]
 * Arithmetic: LUT entries L = C . x_group from tensor-core MMAs (fp32 accumulate), stored as fixed point with one power-of-two
 * scale per k-slice of 256 activations (|L * 2^e| <= 2^22: about 22 significant bits, far finer than the fp16 LUT of vq.py:285);
 * the gather adds integers and the k-slices of a row meet in a 64-bit fixed-point word (exact, order-free => bit-reproducible);
 * one rounding to dtype.  Tolerance (tests): max|y - y64| <= 1e-3 * max|y64| against the float64 evaluation of the same
 * operands.  prob_k <= 65280 (255 k-slices).  A non-finite activation or codebook entry saturates the rows it touches to +-inf
 * (the reference's FMA chain yields NaN there). */
int lutvq_code2x8_matvec(const void* codes, const void* x, void* y, const void* codebook,
                         int prob_m, int prob_k, int dtype, void* ws, size_t ws_bytes, void* stream);

/* y[b][o] = round(round(acc) * scales[o]) + bias[o]   for b < batch  (reference epilogue order, fused)
 * x dtype [batch][prob_k]; scales dtype [prob_m] (NULL = no scaling); bias dtype [prob_m] or NULL; y dtype [batch][prob_m].
 * cuda_kernel.cpp:404-414 launches its kernel once per row.  Here up to 8 rows go out as ONE launch (each row takes a share of the
 * SMs; the rows share the kernel's fixed cost and read the layer's codes at the same pace, so HBM sees them once) when ws holds
 * min(batch, LUTVQ_XREADY_WS_RING) regions of ceil256(lutvq_packed_workspace_bytes(prob_m, prob_k)) bytes (zero-initialised, handed
 * back zero); with a single region it is one launch per row.  Results are bit-identical either way.  (_matmat_packed: the same.) */
int lutvq_code2x8_matmat(const void* x, const void* codes, const void* codebooks, const void* scales,
                         const void* bias, void* y, int batch, int prob_m, int prob_k, int dtype,
                         void* ws, size_t ws_bytes, void* stream);
int lutvq_code2x8_matmat_ex(const void* x, const void* codes, const void* codebooks, const void* scales,
                            const void* bias, void* y, int batch, int prob_m, int prob_k, int dtype,
                            unsigned flags, void* ws, size_t ws_bytes, void* stream);

/* Optional load-time re-layout of the codes for the GEMV (the reference does the same kind of one-off permutation
 * for its CPU kernel in QuantizedLinear.prepare_matmul_op, aqlm/../inference.py:78-83).  The packed image holds the
 * same code bytes, ordered [k-slice of 64 code bytes][32-row set][8-row tile][lane] x 16 B = each lane's register image
 * (rvv-lut-vq_b200/csrc/lut_layout.h), so that a warp's load is 512 contiguous bytes and a worker's share ONE contiguous range;
 * rows / k-slices are zero-padded to whole sets (lutvq_code2x8_packed_bytes >= prob_m*prob_k/4).
 * lutvq_code2x8_matmat_packed == lutvq_code2x8_matmat_ex on the packed image (bit-identical results; the reference-layout
 * entry points read [prob_m][prob_k/8][2] directly: a warp's 128-bit load then touches 8 rows x 64 B). */
size_t lutvq_code2x8_packed_bytes(int prob_m, int prob_k);
int lutvq_code2x8_prepack(const void* codes, void* packed, int prob_m, int prob_k, void* stream);
int lutvq_code2x8_matmat_packed(const void* x, const void* packed, const void* codebooks, const void* scales,
                                const void* bias, void* y, int batch, int prob_m, int prob_k, int dtype,
                                unsigned flags, void* ws, size_t ws_bytes, void* stream);

/* Grouped launch: n <= 3 linears that read the SAME x (q/k/v, gate/up of a decoder layer; each with its own packed codes,
 * codebooks, scales, bias, output) in ONE kernel launch, batch 1.  Results are bit-identical to n separate
 * lutvq_code2x8_matmat_packed calls; a dependent decode chain saves n-1 launch latencies.  ws: at least
 * lutvq_group_workspace_bytes(n, prob_m, prob_k), zero-filled once, left all-zero. */
size_t lutvq_group_workspace_bytes(int n, const int* prob_m, int prob_k);
int lutvq_code2x8_matmat_group_packed(const void* x, int n, const void* const* packed, const void* const* codebooks,
                                      const void* const* scales, const void* const* bias, void* const* y, const int* prob_m,
                                      int prob_k, int dtype, unsigned flags, void* ws, size_t ws_bytes, void* stream);

/* Grouped launch whose activations are computed on the fly: x[i] = fp16( silu(gate[i]) * up[i] ), the MLP activation of a
 * Llama-shaped decoder (down_proj input), evaluated inside the kernel's x load with the arithmetic of
 * lutvq_decode_silu_mul -- results are bit-identical to lutvq_decode_silu_mul followed by
 * lutvq_code2x8_matmat_group_packed, one launch and one [prob_k] round trip fewer in a dependent decode chain.
 * fp16 only; LUTVQ_FLAG_X_READY is rejected.  Not a reference entry point (the reference gets SwiGLU from HF
 * transformers' LlamaMLP); used by the decode loop (BASELINE configs[3], [4]). */
int lutvq_code2x8_matmat_group_packed_swiglu(const void* gate, const void* up, int n, const void* const* packed,
                                             const void* const* codebooks, const void* const* scales,
                                             const void* const* bias, void* const* y, const int* prob_m, int prob_k,
                                             int dtype, unsigned flags, void* ws, size_t ws_bytes, void* stream);

/* W[o][8g+j] = codebook[0][c0(o,g)][j] + codebook[1][c1(o,g)][j]   (one RN add in dtype; bit-exact class)
 * then, if scales != NULL, W[o][:] = round(W[o][:] * scales[o])     (cuda_kernel.cpp:445)
 * W dtype [prob_m][prob_k] row-major. */
int lutvq_code2x8_dequant(const void* codes, void* W, const void* codebook, const void* scales,
                          int prob_m, int prob_k, int dtype, void* stream);

/* Kx8 generalisation (num_codebooks N in 1..8 codebooks of 256 entries, groups of 8; fp16; batch >= 0):
 *   lutvq_codeKx8_matmat   <- codeKx8_matmat / code1x8_matmat + CodeKx8MatVec   aqlm/cuda_kernel.cpp:521-684, cuda_kernel.cu:296-390
 *   lutvq_codeKx8_dequant  <- CodeKx8Dequant (sequential fp16 adds in codebook order) aqlm/cuda_kernel.cu:392-468
 * codes int8 [prob_m][prob_k/8][N]; codebooks fp16 / bf16 [N][256][8].  N = 1, 2, 4, 8 run on the tuned fused LUT-GEMV (kernel
 * templated on the codebooks per group); N = 3, 5, 6, 7 take the unfused pair (fp32 LUT in the workspace + gather + reference
 * epilogue).  ws >= lutvq_codeKx8_workspace_bytes. */
size_t lutvq_codeKx8_workspace_bytes(int prob_m, int prob_k, int num_codebooks);
int lutvq_codeKx8_matmat(const void* x, const void* codes, const void* codebooks, const void* scales, const void* bias, void* y,
                         int batch, int prob_m, int prob_k, int num_codebooks, int dtype, void* ws, size_t ws_bytes, void* stream);
int lutvq_codeKx8_dequant(const void* codes, void* W, const void* codebooks, const void* scales, int prob_m, int prob_k,
                          int num_codebooks, int dtype, void* stream);

/* Prefill path: y = x . W^T with W dequantized on the fly into TENSOR MEMORY (A operand of tcgen05.mma; the activations arrive
 * by TMA tensor copies) and contracted on the tcgen05 tensor cores (fp32 accumulation in TMEM), then the matmat epilogue.  Same
 * arguments as matmat; x must be 16-byte aligned.  A launch takes up to 256 rows; larger batches are chunked (the host-side op
 * switches to lutvq_code2x8_dequant + a library GEMM above 256 rows).  Launched with programmatic stream
 * serialization: reads codes / codebooks before the previous kernel on the stream has completed (weights must not be written by it). */
int lutvq_code2x8_matmat_dequant(const void* x, const void* codes, const void* codebooks, const void* scales,
                                 const void* bias, void* y, int batch, int prob_m, int prob_k, int dtype,
                                 void* ws, size_t ws_bytes, void* stream);

/* Golden-model pair (vq.py layouts; unfused; used by the parity tests and for ncu of the two halves):
 * lutvq_lut_build:  x dtype-fp16 [M*d]; codebook fp16 [M][N][K][d] or, if cb_shared, [N][K][d];
 *                   lut [M][N][K] fp16 (LUTVQ_LUT_VQPY_FP16) or fp32 (LUTVQ_LUT_FP32).   d <= 32, K <= 65536.
 * lutvq_lut_gemv:   codeword uint8 [d_out][M][N]; lut as above (lut_is_fp32 selects);
 *                   out float64[d_out] (LUTVQ_ACC_F64) or float32[d_out] (LUTVQ_ACC_F32). */
int lutvq_lut_build(const void* x, const void* codebook, void* lut, int M, int N, int K, int d,
                    int cb_shared, int lut_mode, void* stream);
int lutvq_lut_gemv(const void* codeword, const void* lut, void* out, int d_out, int M, int N, int K,
                   int lut_is_fp32, int acc_mode, void* stream);
/* W fp16 [d_out][M*d] = fp16( sum_n codebook[m][n][code[j][m][n]][:] ) with the sum in float64 (vq.py:276-278) */
int lutvq_vq_dequantize(const void* codebook, const void* codeword, void* W, int d_out, int M, int N, int K,
                        int d, int cb_shared, void* stream);


/* ---- output-channel-sharded decode over the GPUs of one NVSwitch domain (one process per GPU) --------------------
 * Replaces nothing in the reference (its kernels are single-GPU; cuda_kernel.cu:486,497 even queries device 0): this
 * is the multi-GPU row of the hot path (BASELINE.json north_star: "weight codes sharded by output channel across 2, 4
 * and 8 GPUs ... all-gather of output slices over NVLink").  Rank r of P owns rows [r*O/P, (r+1)*O/P) of a layer.
 * lutvq_sharded_matmat_packed runs the fused LUT-GEMV on the local rows and, in the SAME kernel, stores every finished
 * row into the gather slot of every peer over NVLink and publishes the slot's epoch; lutvq_shard_wait (stream-ordered,
 * graph-capturable, bounded spin) returns once all P slices of the slot have landed in the LOCAL slot, which then holds
 * the whole y.  The NCCL all-gather variant of the same step lives in the host plumbing (rvv-lut-vq_b200/sharded.py).
 *
 *   create (cudaMalloc of {flags | n_slots x slot_bytes}) -> export (64-byte CUDA IPC handle) -> exchange the handles
 *   with any host transport -> attach (maps the peers) -> [sharded_matmat_packed -> wait -> read lutvq_shard_slot()]* -> destroy
 * A slot may be pushed again once every rank has consumed its previous epoch; rotate through >= 2 slots. */
typedef struct lutvq_shard lutvq_shard_t;
#define LUTVQ_SHARD_HANDLE_BYTES 64
int    lutvq_shard_create(lutvq_shard_t** out, int nranks, int rank, int n_slots, size_t slot_bytes);
int    lutvq_shard_export(lutvq_shard_t* s, void* handle64);
int    lutvq_shard_attach(lutvq_shard_t* s, const void* handles /* nranks x 64 bytes, indexed by rank */);
/* same-process alternative to export/attach: bases[r] = lutvq_shard_base() of rank r's object (peer access enabled) */
int    lutvq_shard_attach_ptrs(lutvq_shard_t* s, void* const* bases);
void*  lutvq_shard_base(lutvq_shard_t* s);
void*  lutvq_shard_slot(lutvq_shard_t* s, int slot);              /* local device pointer of a gather slot */
size_t lutvq_shard_slot_bytes(lutvq_shard_t* s);
/* y_slot[row_offset + i] (on EVERY rank) = epilogue(W_local(packed) . x)[i], i < rows_local; batch 1.
 * packed = lutvq_code2x8_prepack image of the LOCAL rows; other arguments as lutvq_code2x8_matmat_packed. */
int    lutvq_sharded_matmat_packed(lutvq_shard_t* s, int slot, size_t row_offset, const void* x, const void* packed,
                                   const void* codebooks, const void* scales, const void* bias, int rows_local,
                                   int prob_k, int dtype, unsigned flags, void* ws, size_t ws_bytes, void* stream);
/* grouped form (see lutvq_code2x8_matmat_group_packed): linear i fills rows [row_offsets[i], +rows_local[i]) of slots[i] */
int    lutvq_sharded_matmat_group_packed(lutvq_shard_t* s, int n, const int* slots, const size_t* row_offsets, const void* x,
                                         const void* const* packed, const void* const* codebooks, const void* const* scales,
                                         const void* const* bias, const int* rows_local, int prob_k, int dtype, unsigned flags,
                                         void* ws, size_t ws_bytes, void* stream);
/* the same with the SwiGLU fused into the activation load: x[i] = silu(gate[i]) * up[i] (see
 * lutvq_code2x8_matmat_group_packed_swiglu) */
int    lutvq_sharded_matmat_group_packed_swiglu(lutvq_shard_t* s, int n, const int* slots, const size_t* row_offsets,
                                                const void* gate, const void* up, const void* const* packed,
                                                const void* const* codebooks, const void* const* scales, const void* const* bias,
                                                const int* rows_local, int prob_k, int dtype, unsigned flags, void* ws,
                                                size_t ws_bytes, void* stream);
int    lutvq_shard_wait(lutvq_shard_t* s, int slot, void* stream);
int    lutvq_shard_wait_mask(lutvq_shard_t* s, unsigned long long slot_mask, void* stream);   /* all slots of the mask, one launch */
/* software pipelining: forward + publish the slots now (non-blocking), lutvq_shard_wait[_mask] on them later only waits */
int    lutvq_shard_push_mask(lutvq_shard_t* s, unsigned long long slot_mask, void* stream);
int    lutvq_shard_error(lutvq_shard_t* s, unsigned* out);        /* nonzero: a wait timed out (epoch in *out); syncs the device */
/* ---- "LL" exchange: the all-gather of a sharded decode without exchange launches, flags or fences ----
 * Every finished output row of a sharded GEMV is written as ONE 32-bit word {value | epoch16 << 16} into the gather slot of
 * EVERY rank (own rank included; peer memory over NVLink).  A word is valid once its upper half carries the expected epoch
 * = f(step counter, sequence number of the exchange inside the step), so the consumer -- the next GEMV's activation load, the
 * residual-add + RMSNorm, the attention kernel -- polls exactly the words it is about to use, in its own prologue.  Slot
 * reuse needs no acknowledgement in a decode chain (a rank can only be one exchange ahead of the slowest peer; rotate through
 * >= 2 slots).  Everything is device-side state: replays inside CUDA graphs.
 *   lutvq_shard_ll_tick            ++step counter; call once per step (token) on every rank, before the step's first exchange
 *   lutvq_sharded_matmat_group_ll  n <= 3 linears sharing an input, outputs as LL words into out_slots[i] at row_offsets[i]
 *                                  with sequence numbers out_seq[i] (< 1024, unique within a step); input: plain x (x_slot < 0)
 *                                  or the LL words of the local slot x_slot / x_seq; gate_slot >= 0: silu(gate) * x, both LL (fp16)
 *   lutvq_decode_add_rmsnorm_ll / lutvq_decode_attention_ll   the glue kernels with their sharded operands taken from LL slots
 *   lutvq_decode_add_rmsnorm_ll_out / lutvq_decode_attention_ll_out   the same, with the result ALSO written as LL words into the
 *                                  LOCAL slot y_slot / out_slot (delta_slot < 0: no delta).  This makes "dataflow launches" possible:
 *                                  lutvq_sharded_matmat_group_ll with LUTVQ_FLAG_X_READY and x_slot >= 0 does not wait for the
 *                                  launch before it to complete before it reads its activations -- it streams its codes, then
 *                                  polls the LL words (they carry their own validity) -- and only orders its stores behind it.
 *                                  Its workspace must not be shared with the LUTVQ_XREADY_WS_RING - 1 launches around it.
 *   lutvq_shard_ll_unpack          LL slot -> plain vector (for consumers that are not ours, and for tests)
 * Polls are bounded (seconds): on a timeout the shard's error word (lutvq_shard_error) is set and the kernel returns. */
int    lutvq_shard_ll_tick(lutvq_shard_t* s, void* stream);
int    lutvq_sharded_matmat_group_ll(lutvq_shard_t* s, int n, const int* out_slots, const size_t* row_offsets, const unsigned* out_seq,
                                     const void* x, int x_slot, unsigned x_seq, int gate_slot, unsigned gate_seq,
                                     const void* const* packed, const void* const* codebooks, const void* const* scales,
                                     const void* const* bias, const int* rows_local, int prob_k, int dtype, unsigned flags,
                                     void* ws, size_t ws_bytes, void* stream);
int    lutvq_shard_ll_unpack(lutvq_shard_t* s, int slot, unsigned seq, void* y, int n, void* stream);
int    lutvq_decode_add_rmsnorm_ll(lutvq_shard_t* s, void* resid, int delta_slot, unsigned delta_seq, const void* w, void* y, int n,
                                   float eps, void* stream);
int    lutvq_decode_attention_ll(lutvq_shard_t* s, int q_slot, int k_slot, int v_slot, unsigned q_seq, unsigned k_seq, unsigned v_seq,
                                 void* kcache, void* vcache, const int* pos, void* out, int n_heads, int n_kv_heads, int head_dim,
                                 int max_seq, float rope_theta, void* stream);
int    lutvq_decode_add_rmsnorm_ll_out(lutvq_shard_t* s, void* resid, int delta_slot, unsigned delta_seq, const void* w, void* y,
                                       int y_slot, unsigned y_seq, int n, float eps, void* stream);
int    lutvq_decode_attention_ll_out(lutvq_shard_t* s, int q_slot, int k_slot, int v_slot, unsigned q_seq, unsigned k_seq, unsigned v_seq,
                                     void* kcache, void* vcache, const int* pos, void* out, int out_slot, unsigned out_seq, int n_heads,
                                     int n_kv_heads, int head_dim, int max_seq, float rope_theta, void* stream);
int    lutvq_shard_destroy(lutvq_shard_t* s);

/* ---- integer LUT: lookup-and-add ----
 * Replaces lookup_and_add_simd of benchmark/kernel/simd_lut/bench_avx_lut.cpp:26-138 with the _SCALAR_ build's semantics
 * (:98-107):  for run < n_runs, j < n_lut_per_run, k < cb_size:  output[k] += lut[idx[j*cb_size + k]]  with int16
 * wraparound (additions modulo 2^16: split over all SMs, any order, same bits).  idx: uint8 [n_lut_per_run][cb_size]
 * (4-byte aligned, values < n_entries), lut: int8 [n_entries <= 256], output: int16 [cb_size] (accumulated in place),
 * cb_size 16 / 32 / 64 (any multiple of 4 dividing 1024).  ws: lutvq_simd_lut_workspace_bytes() bytes, zero on first
 * use, handed back zeroed.  The reference's AVX builds use SATURATING adds and differ from its own scalar build; the
 * scalar build is the oracle (oracle/lutvq_oracle.c orc_simd_lut_scalar). */
size_t lutvq_simd_lut_workspace_bytes(int cb_size);
int    lutvq_simd_lut_lookup_add(const void* idx, const void* lut, void* output, int cb_size, int n_entries,
                                 unsigned long long n_runs, unsigned long long n_lut_per_run, void* ws, size_t ws_bytes,
                                 void* stream);

/* ---- AQLM 1x16 (one codebook of 65536 vectors, 16-bit codes) ----
 * Replaces code1x16_matvec_cuda<bf16, group_size> / code1x16_dequant_cuda<group_size> (cuda_kernel.cu:470-553 launching
 * Code1x16MatVec :7-95 / Code1x16Dequant :98-142) and the host wrappers code1x16_matvec / code1x16_matmat /
 * code1x16_dequant (cuda_kernel.cpp:113-209).  codes int16 [prob_m][prob_k / group_size] (unsigned index in two's
 * complement, utils.py:24-31), codebook [65536][group_size] fp16/bf16, group_size 8 or 16 (anything else:
 * LUTVQ_EUNSUPPORTED, the reference's NotImplementedError cuda_kernel.cpp:136-145), prob_k a multiple of 8 * group_size.
 *   matvec : y[prob_m] = round(W x)                          (the raw launcher, no scale)
 *   matmat : y[batch][prob_m] = round(W x) * scales + bias   (the reference's rounding order, cuda_kernel.cpp:95-111;
 *            up to 4 batch rows share each fetched codebook row instead of one launch per row)
 *   dequant: W[prob_m][prob_k] = codebook[code] (* scales[row] when scales != NULL: `weight *= scales`, one RN multiply); bit-exact.
 * Not LUT-able (G x 65536 entries per activation vector): a dequantize-and-FMA GEMV, fp32 accumulation, bound by the
 * L2 gather of one codebook row per code.  No workspace. */
int lutvq_code1x16_matvec(const void* codes, const void* x, void* y, const void* codebook, int prob_m, int prob_k, int group_size,
                          int dtype, void* stream);
int lutvq_code1x16_matmat(const void* x, const void* codes, const void* codebooks, const void* scales, const void* bias, void* y,
                          int batch, int prob_m, int prob_k, int group_size, int dtype, void* stream);
int lutvq_code1x16_dequant(const void* codes, void* W, const void* codebook, const void* scales, int prob_m, int prob_k, int group_size,
                           int dtype, void* stream);

/* ---- T-MAC bit-serial integer LUT ----
 * The entry points of benchmark/kernel/bs_lut_gemm/tbl.cc and lut_ctor.cc with the reference's argument lists (x86
 * semantics: float_type = float, tbl.cc:10-15) plus a stream; the template parameters the reference bakes into the
 * symbol name (tbl.cc:673-691 `tbl_g4_int8_float_update_s{s}_k{K}_b{bits}_ak{act_k}_fa{fa}_z{z}_os{os}`, lut_ctor.cc:223-226
 * `lut_ctor_g4_int8_k{fak}_b{bits}`) are trailing arguments.  All pointers are device pointers.  Integer results are exact,
 * fp32 results reproduce the reference's AVX2 build bit for bit (oracle/tmac_oracle.c, pinned by oracle/_ref/libtmac_ref.so).
 *   tbl_g4_int8_int32_update (tbl.cc:536-630): block i of 32 outputs: c[2i+l] += sum_k lut[k][a[iK+16k+l] & 15],
 *       c[2i+16+l] += sum_k lut[k][a[iK+16k+l] >> 4], k-sum in wrapping int16.  c int32[m], lut int8[K][16] (16-byte
 *       aligned), a uint8[m/2 * K]; m % 32 == 0.
 *   tbl_g4_int8_float_update (tbl.cc:323-532): same sums per act_k block, * lut_scales[kk] (+ lut_biases[kk] on bit plane 0),
 *       then c = fma(sum, scales[...], c); fast_aggregation = the halving adder (tbl.cc:169-215), zero_point = tbl.cc:493-507.
 *       (has_scale and one_scale are not consulted on the x86 path and have no argument.)
 *   tbl_{int8,int16,int32,float}_reset (tbl.cc:697-715), partial_max_g4_int8_k8 / partial_max_reset (lut_ctor.cc:232-266):
 *       *lut_scales = max(*lut_scales, max_j(|b[4j]|+|b[4j+1]|+|b[4j+2]|+|b[4j+3]|) / 127) over 32 activations.
 *   lut_ctor_g4_int8 (lut_ctor.cc:38-221): qlut int8[act_k/4][16] = sat(round(entry / *lut_scales)), *lut_biases = sum of
 *       the all-minus entries; *lut_scales is read (and written back unchanged).
 *   tmac_preprocessor_int8 / tmac_qgemm_lut_int8: the call structure T-MAC's generated kernels wrap around the above
 *       (deploy/tuned/kernels.cc:1002-1037 / 827-960; TMACGeMMWrapper calls them): per 32 activations reset + partial_max +
 *       lut_ctor(32); per tile of bm bit-plane rows tbl_float_reset, K/32 x tbl_g4_int8_float_update(k8, ak8), then
 *       C = ((c0 * 0.5 + c1) + c2 * 2) + c3 * 4 over the bit planes.  A uint8[M*bits/bm][K/32][bm/2 * 8],
 *       scales float[M*bits/bm][K/32][bm/bits], C float[M].  Shapes without a kernel return LUTVQ_EUNSUPPORTED (the
 *       reference's generated dispatcher returns -1). */
int lutvq_tbl_g4_int8_int32_update(int m, void* c, const void* lut, const void* a, int K, void* stream);
int lutvq_tbl_g4_int8_float_update(int m, void* c, const void* lut, const void* a, const void* scales, const void* lut_scales,
                                   const void* lut_biases, int K, int bits, int act_k, int fast_aggregation, int zero_point,
                                   void* stream);
int lutvq_tbl_int8_reset(int m, void* c, void* stream);
int lutvq_tbl_int16_reset(int m, void* c, void* stream);
int lutvq_tbl_int32_reset(int m, void* c, void* stream);
int lutvq_tbl_float_reset(int m, void* c, void* stream);
int lutvq_partial_max_g4_int8_k8(void* lut_scales, const void* b, void* stream);
int lutvq_partial_max_reset(void* lut_scales, void* stream);
int lutvq_lut_ctor_g4_int8(int act_k, void* qlut, const void* b, void* lut_scales, void* lut_biases, void* stream);
int lutvq_tmac_preprocessor_int8(int K, const void* B, void* lut_scales, void* lut_biases, void* qlut, void* stream);
int lutvq_tmac_qgemm_lut_int8(int M, int K, int bits, int bm, const void* A, const void* qlut, const void* scales,
                              const void* lut_scales, const void* lut_biases, void* C, void* stream);

/* ---- no-allocation variant of lutvq_code2x8_matmat (csrc/noalloc.cu) ----
 * No counterpart in the reference: added for callers that cannot call cudaMalloc or carry
 * a workspace (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 * Same arguments and results as lutvq_code2x8_matmat (reference-layout codes int8 [prob_m][prob_k/8][2], bit-identical outputs)
 * minus the workspace: the split-k words live in a file-scope static __device__ array inside the library, sized for
 * prob_m <= LUTVQ_NOALLOC_MAX_ROWS (lutvq_noalloc_scratch_bytes() = 256 + 8 * LUTVQ_NOALLOC_MAX_ROWS bytes = 257 KiB, held by every
 * GPU that loads the library until exit).  Larger prob_m is rejected with LUTVQ_EINVAL and a message naming the limit.
 * The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine): one stream-ordered memset of the region
 * it uses + one kernel launch per batch row.  ALL calls on a device share the one array: calls that overlap on the GPU (two
 * streams) race -- serialise them on one stream, or use lutvq_code2x8_matmat with private workspaces. */
#define LUTVQ_NOALLOC_MAX_ROWS 32768
size_t lutvq_noalloc_scratch_bytes(void);
int lutvq_code2x8_matmat_noalloc(const void* x, const void* codes, const void* codebooks, const void* scales, const void* bias,
                                 void* y, int batch, int prob_m, int prob_k, int dtype, void* stream);

/* ---- PQ4 fast-scan distance accumulation (csrc/pq4.cu) ----
 * Replaces, with the same packed layouts (images packed by either side are interchangeable) and bit-identical uint16 results:
 *   pq4_pack_codes            benchmark/kernel/faiss-simd/pq4_fast_scan.cpp:45-91    codes [ntotal][(M+1)/2] -> blocks [nb * nsq / 2]
 *   pq4_pack_codes_range      pq4_fast_scan.cpp:93-137   ORs vectors [i0, i1) into blocks (zero-initialised by the caller)
 *   pq4_pack_LUT              pq4_fast_scan.cpp:253-264  src [nq][nsq][16] -> dest [(sq/2 * nq + q) * 32 + 16 * (sq & 1)]
 *   pq4_pack_LUT_qbs(_q_map)  pq4_fast_scan.cpp:266-322  one pq4_pack_LUT per base-16 digit of qbs; q_map: DEVICE int array or NULL
 *   pq4_accumulate_loop       pq4_fast_scan_search_1.cpp:30-222 with a store-to-memory result handler:
 *                             accu[q * ld + j] = sum_sq LUT[q][sq][code(j, sq)] modulo 2^16; scale >= 0 = NormTableScaler(scale):
 *                             the last two sub-quantizers' values are multiplied by scale (uint16 wrap-around); scale < 0: none
 *   pq4_accumulate_loop_qbs   pq4_fast_scan.h:170-188 (declared there, defined in a file the reference does not carry): bbs = 32,
 *                             query blocks of 1..4 (the digits of qbs), LUT packed by pq4_pack_LUT_qbs, rows in digit order
 * All pointers are device memory.  Errors mirror the reference's FAISS_THROW_IF_NOT checks (32-byte aligned codes / LUT,
 * bbs % 32, nb % bbs, nsq % 2); the reference instantiates (nq, bbs/32) in {1x1..1x5, 2x1, 2x2, 3x1, 4x1} only, this
 * library takes every bbs with nq = 1..4.  pq4_get/set_packed_element and CodePackerPQ4 (single-element host accessors)
 * are mirrored in the Python host layer (pq4.py).  pq4_preferred_qbs is declared by the reference but defined nowhere in
 * it: not provided. */
int lutvq_pq4_pack_codes(const void* codes, size_t ntotal, size_t M, size_t nb, size_t bbs, size_t nsq, void* blocks, void* stream);
int lutvq_pq4_pack_codes_range(const void* codes, size_t M, size_t i0, size_t i1, size_t bbs, size_t nsq, void* blocks, void* stream);
int lutvq_pq4_pack_LUT(int nq, int nsq, const void* src, void* dest, void* stream);
int lutvq_pq4_qbs_to_nq(int qbs);
int lutvq_pq4_pack_LUT_qbs(int qbs, int nsq, const void* src, const int* q_map, void* dest, void* stream);
int lutvq_pq4_accumulate_loop(int nq, size_t nb, int bbs, int nsq, const void* codes, const void* LUT, void* accu, size_t ld,
                              int scale, void* stream);
int lutvq_pq4_accumulate_loop_qbs(int qbs, size_t nb, int nsq, const void* codes, const void* LUT, void* accu, size_t ld, int scale,
                                  void* stream);

/* ---- decode glue (fp16, batch 1): the non-GEMV steps of a Llama-shaped decode step, so a whole token is one chain of
 * launches of this library inside a CUDA graph.  Not reference entry points (the reference takes these from HF
 * transformers, benchmark/llm/generate_benchmark.py) and not LUT-VQ work; they exist for the decode tok/s metric.
 *   add_rmsnorm : resid += delta (delta may be NULL); y = resid * rsqrt(mean(resid^2) + eps) * w        (n <= 16384)
 *   silu_mul    : y = silu(gate) * up
 *   attention   : RoPE(q, k_new) at position *pos (device int), append k/v to the caches [n_kv][max_seq][128],
 *                 out = softmax(q K^T / sqrt(128)) V over positions 0..*pos; head_dim must be 128 */
int lutvq_decode_add_rmsnorm(void* resid, const void* delta, const void* w, void* y, int n, float eps, void* stream);
int lutvq_decode_silu_mul(const void* gate, const void* up, void* y, int n, void* stream);
int lutvq_decode_attention(const void* q, const void* k_new, const void* v_new, void* kcache, void* vcache, const int* pos,
                           void* out, int n_heads, int n_kv_heads, int head_dim, int max_seq, float rope_theta, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* LUTVQ_H_ */
