// shard.cu -- output-channel-sharded decode across the GPUs of one NVSwitch domain.
//
// Rank r of P holds codes[r*O/P : (r+1)*O/P] and computes that slice of y with the SAME fused LUT-GEMV kernel
// (prob_m = O/P).  The exchange step of the path -- every rank needs the whole y as the next layer's input -- is fused
// into the kernel's epilogue instead of a separate collective: the lane that finishes a row stores it into every peer's
// gather buffer over NVLink (plain peer stores, 2 bytes x P per row), and the last CTA of the launch publishes the
// slot's epoch to every peer with one st.release.sys each.  A consumer waits (lutvq_shard_wait: one warp spinning on
// ITS OWN memory with ld.acquire.sys) until all P epochs of the slot have arrived.  No NCCL call, no extra copy,
// no host involvement; everything is graph-capturable because the epochs live in device memory.
//
// One process per GPU: each rank cudaMalloc's one symmetric buffer {flags | gather slots}, exports it as a CUDA IPC
// handle, the handles are exchanged by the host plumbing (torch.distributed in rvv-lut-vq_b200/sharded.py) and every
// rank maps its peers' buffers (lutvq_shard_attach).  lutvq_shard_attach_ptrs takes raw device pointers instead, for a
// single process that drives several GPUs with peer access enabled (and for the one-GPU protocol test).
//
// Slot reuse: a slot may be pushed again once every rank has consumed its previous epoch.  In a decode chain that is
// implied with >= 2 slots (a rank can only be one gather ahead of the slowest peer); the python
// layer rotates through all slots.
#include "common.cuh"

#include <new>
#include <string.h>

namespace lutvq {

int gemv_2x8_launch(const void* packed, const void* x, void* y, const void* codebook, const void* scales,
                    const void* bias, int O, int I, int dtype, unsigned flags, void* ws, size_t ws_bytes,
                    cudaStream_t stream, const GemvPush* push);

int gemv_2x8_launch_group(int n, const GemvItem* items, const void* x, const void* x_gate, int I, int dtype, unsigned flags,
                          void* ws, size_t ws_bytes, cudaStream_t stream, const GemvPush* push, const GemvLL* ll = nullptr);

constexpr int    kShardMaxRanks  = 8;
constexpr int    kShardMaxSlots  = 64;      // = bits of a slot mask; [slot][rank] epochs fill exactly the first kLocalOff bytes
constexpr size_t kShardFlagBytes = 4096;       // [slot][rank] uint32 epochs written by peers, then local counters
constexpr unsigned long long kWaitTimeoutNs = 4000000000ull;   // a wait gives up (and records an error) after 4 s
constexpr size_t kLocalOff       = 2048;       // local-only area inside the flag page: sent[slot], consumed[slot], error

struct Shard {
    int nranks = 0, rank = 0, n_slots = 0, device = -1;
    size_t slot_bytes = 0, total_bytes = 0;
    size_t ll_off = 0, ll_slot_bytes = 0;      // LL gather words (common.cuh): n_slots slots of one 32-bit word per output row, behind the plain slots
    uint8_t* local = nullptr;                   // this rank's symmetric buffer
    uint8_t* peer[kShardMaxRanks] = {};         // peer[r] = rank r's buffer as mapped here (peer[rank] == local)
    bool ipc_opened[kShardMaxRanks] = {};
    bool attached = false;
    // what the last lutvq_sharded_matmat_packed on each slot produced (host state; baked into the exchange launch)
    struct Pending { unsigned row_off = 0, rows = 0; int fused = 0, pushed = 0; } pending[kShardMaxSlots];
};

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// One warp.  Lane r < nranks spins on flag[slot][r] (local memory, written remotely) until it reaches the slot's next
// expected epoch.  Bounded: after `timeout_ns` the kernel records an error word and returns instead of hanging the GPU.
__global__ void shard_wait_kernel(const unsigned* flags, unsigned* consumed, unsigned* error, int nranks,
                                  unsigned long long timeout_ns) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // the next launch may prefetch its weights while we spin
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const unsigned want = *consumed + 1u;
    const int lane = threadIdx.x;
    bool ok = true;
    if (lane < nranks) {
        unsigned long long t0, t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        while ((int)(ld_acquire_sys(flags + lane) - want) < 0) {
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t - t0 > timeout_ns) { ok = false; break; }
            __nanosleep(64);
        }
    }
    const unsigned all_ok = __all_sync(0xffffffffu, ok);
    __threadfence_system();
    if (lane == 0) {
        *consumed = want;
        if (!all_ok) *error = want;
    }
}

// Several slots at once: thread t = 8 * slot + rank (blockDim = 8 * kShardMaxSlots); slots outside `mask` are skipped.
__global__ void shard_wait_mask_kernel(const unsigned* flags, unsigned* consumed, unsigned* error, int nranks, unsigned long long mask,
                                       unsigned long long timeout_ns) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int slot = threadIdx.x >> 3, r = threadIdx.x & 7;
    const bool active = ((mask >> slot) & 1ull) != 0 && r < nranks;
    bool ok = true;
    unsigned want = 0;
    if (active) {
        want = consumed[slot] + 1u;
        unsigned long long t0, t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        while ((int)(ld_acquire_sys(flags + slot * kShardMaxRanks + r) - want) < 0) {
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t - t0 > timeout_ns) { ok = false; break; }
            __nanosleep(64);
        }
    }
    const int all_ok = __syncthreads_and(ok ? 1 : 0);
    __threadfence_system();
    if (active && r == 0) consumed[slot] = want;
    if (threadIdx.x == 0 && !all_ok) *error = 0xFFFFFFFFu;
}

// ---- exchange kernel: push + wait of up to kShardMaxSlots slots in ONE launch (the default exchange) -------------------
// grid = nranks CTAs.  CTA c != rank forwards this rank's finished slice of every listed slot into peer c's gather slot
// (coalesced 128-bit NVLink stores from L2-resident data), fences once, and publishes the epochs to peer c.  CTA
// c == rank publishes the local epochs and then spins (bounded) until every rank's epoch of every listed slot has arrived.
// The GEMV launches themselves stay untouched: no fences, no arrival counting on their critical path.  Measured on
// 2 x B200 (4096x11008 ring): fused epilogue push (LUTVQ_FLAG_PUSH_EPILOGUE) 12.8 us per GEMV -- the per-CTA
// system-scope fence dominates, it is the same 13.5 us on ONE GPU with no peer at all -- vs this kernel: see DESIGN.md 4.5.
struct ExchangeList { int n; int slot[kShardMaxSlots]; unsigned row_off[kShardMaxSlots]; unsigned rows[kShardMaxSlots]; };
struct PeerBases { uint8_t* base[kShardMaxRanks]; };

// grid = (nranks, listed slots).  CTA (c, e): c != rank forwards slice e to peer c and publishes its epoch there;
// c == rank publishes the local epoch and (do_wait) spins, bounded, until all ranks' epochs of slot e have arrived.
// Epochs live in device memory (graph replays need no host state): sent[slot] counts pushes, consumed[slot] waits.
// Launch-queue note: three resident GEMV CTAs hold 61K of an SM's 64K registers, so an exchange CTA only becomes resident
// when a GEMV CTA exits, and everything queued behind it in the SAME stream waits with it.  Measured (4096x11008 X_READY
// ring, 2 x B200): exchange launches in the GEMV stream 8.2 us per GEMV; on a side stream (parallel graph branch) 5.5 us.
// Throughput callers therefore issue the exchange from a side stream; a decode chain is dependent anyway.
constexpr int kExThreads = 256;
__global__ void __launch_bounds__(kExThreads) shard_exchange_kernel(uint8_t* local, PeerBases peers, int nranks, int rank, size_t slot_bytes,
                                                            ExchangeList ex, int do_push, int do_wait, unsigned long long timeout_ns) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");                 // the GEMVs that filled the local slices are complete
    unsigned* lflags = reinterpret_cast<unsigned*>(local);
    unsigned* lstate = reinterpret_cast<unsigned*>(local + kLocalOff);
    unsigned* sent = lstate;
    unsigned* consumed = lstate + kShardMaxSlots;
    unsigned* error = lstate + 2 * kShardMaxSlots;
    unsigned* arrive = lstate + 2 * kShardMaxSlots + 1;                // CTAs of this launch that have finished pushing
    const int c = blockIdx.x, e = blockIdx.y, slot = ex.slot[e];
    const unsigned epoch_push = sent[slot] + 1u;                        // read by every CTA before any CTA can arrive below
    if (do_push) {
        if (c != rank) {
            uint8_t* pb = peers.base[c];
            const size_t off = kShardFlagBytes + slot_bytes * slot + (size_t)ex.row_off[e] * 2;
            const uint4* src = reinterpret_cast<const uint4*>(local + off);
            uint4* dst = reinterpret_cast<uint4*>(pb + off);
            const int n16 = (int)(ex.rows[e] * 2 / 16);
            #pragma unroll 8
            for (int i = threadIdx.x; i < n16; i += kExThreads) dst[i] = __ldcg(src + i);      // L2-coherent reads, coalesced NVLink stores
            const unsigned short* s2 = reinterpret_cast<const unsigned short*>(local + off);
            unsigned short* d2 = reinterpret_cast<unsigned short*>(pb + off);
            for (int i = n16 * 8 + threadIdx.x; i < (int)ex.rows[e]; i += kExThreads) d2[i] = __ldcg(s2 + i);
            __threadfence_system();
            __syncthreads();
            if (threadIdx.x == 0) {
                unsigned* f = reinterpret_cast<unsigned*>(pb) + (size_t)slot * kShardMaxRanks + rank;
                asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(f), "r"(epoch_push) : "memory");
            }
        } else if (threadIdx.x == 0) {
            asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(lflags + (size_t)slot * kShardMaxRanks + rank), "r"(epoch_push) : "memory");
        }
    }
    if (do_wait && c == rank) {
        const unsigned want = consumed[slot] + 1u;
        bool ok = true;
        if ((int)threadIdx.x < nranks) {
            const unsigned* f = lflags + (size_t)slot * kShardMaxRanks + threadIdx.x;
            unsigned long long t0, t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            while ((int)(ld_acquire_sys(f) - want) < 0) {
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                if (t - t0 > timeout_ns) { ok = false; break; }
                __nanosleep(32);
            }
        }
        const int all_ok = __syncthreads_and(ok ? 1 : 0);
        __threadfence_system();
        if (threadIdx.x == 0) {
            consumed[slot] = want;
            if (!all_ok) *error = 0xFFFFFFFEu;
        }
    }
    if (do_push && threadIdx.x == 0) {
        // the last CTA of the launch advances the push epochs (every CTA has read its epoch by the time it arrives here)
        const unsigned total = gridDim.x * gridDim.y;
        __threadfence();
        if (atomicAdd(arrive, 1u) == total - 1) {
            for (int k = 0; k < ex.n; ++k) sent[ex.slot[k]] += 1u;
            __threadfence();
            *arrive = 0u;
        }
    }
}

// local state words behind sent[] / consumed[] / error / arrive: the LL step counter
constexpr int kLLTickWord = 2 * kShardMaxSlots + 2;

__global__ void shard_ll_tick_kernel(unsigned* tick) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    *tick += 1u;
}

// LL words of a slot -> a plain fp16 / bf16 vector (tests, and consumers that are not ours, e.g. a library GEMM)
__global__ void shard_ll_unpack_kernel(const uint32_t* words, const unsigned* tick, unsigned seq, unsigned* error, unsigned short* y, int n) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const unsigned ep = ll_epoch16(*tick, seq);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) y[i] = ll_load1(words + i, ep, error);
}

static int launch_exchange(Shard* s, const ExchangeList& ex, cudaStream_t stream, int do_push = 1, int do_wait = 1) {
    PeerBases pb;
    for (int r = 0; r < kShardMaxRanks; ++r) pb.base[r] = r < s->nranks ? s->peer[r] : nullptr;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(s->nranks, ex.n); cfg.blockDim = dim3(kExThreads); cfg.dynamicSmemBytes = 0; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    LUTVQ_CUDA_OK(cudaLaunchKernelEx(&cfg, shard_exchange_kernel, s->local, pb, s->nranks, s->rank, s->slot_bytes, ex, do_push, do_wait, kWaitTimeoutNs));
    return LUTVQ_OK;
}

}  // namespace lutvq

using namespace lutvq;

extern "C" {

int lutvq_shard_create(lutvq_shard_t** out, int nranks, int rank, int n_slots, size_t slot_bytes) {
    LUTVQ_REQUIRE(out != nullptr, "null out pointer");
    LUTVQ_REQUIRE(nranks >= 1 && nranks <= kShardMaxRanks && rank >= 0 && rank < nranks, "bad nranks=%d rank=%d (<= %d ranks)", nranks, rank, kShardMaxRanks);
    LUTVQ_REQUIRE(n_slots >= 1 && n_slots <= kShardMaxSlots, "n_slots=%d must be in 1..%d", n_slots, kShardMaxSlots);
    LUTVQ_REQUIRE(slot_bytes > 0, "slot_bytes must be positive");
    DeviceInfo di;
    if (int rc = get_device_info(&di)) return rc;
    Shard* s = new (std::nothrow) Shard();
    LUTVQ_REQUIRE(s != nullptr, "out of host memory");
    s->nranks = nranks; s->rank = rank; s->n_slots = n_slots; s->device = di.device;
    s->slot_bytes = (slot_bytes + 255) / 256 * 256;
    s->ll_off = kShardFlagBytes + s->slot_bytes * n_slots;
    s->ll_slot_bytes = 2 * s->slot_bytes;
    s->total_bytes = s->ll_off + s->ll_slot_bytes * n_slots;
    cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&s->local), s->total_bytes);
    if (e != cudaSuccess) { delete s; return fail(LUTVQ_ECUDA, "cudaMalloc(%zu) failed: %s", s->total_bytes, cudaGetErrorString(e)); }
    e = cudaMemset(s->local, 0, s->total_bytes);
    if (e != cudaSuccess) { cudaFree(s->local); delete s; return fail(LUTVQ_ECUDA, "cudaMemset failed: %s", cudaGetErrorString(e)); }
    s->peer[rank] = s->local;
    s->attached = (nranks == 1);
    *out = reinterpret_cast<lutvq_shard_t*>(s);
    return LUTVQ_OK;
}

int lutvq_shard_export(lutvq_shard_t* h, void* handle64) {
    Shard* s = reinterpret_cast<Shard*>(h);
    LUTVQ_REQUIRE(s && handle64, "null pointer");
    static_assert(sizeof(cudaIpcMemHandle_t) == LUTVQ_SHARD_HANDLE_BYTES, "IPC handle size");
    cudaIpcMemHandle_t ih;
    LUTVQ_CUDA_OK(cudaIpcGetMemHandle(&ih, s->local));
    memcpy(handle64, &ih, sizeof(ih));
    return LUTVQ_OK;
}

int lutvq_shard_attach(lutvq_shard_t* h, const void* handles) {
    Shard* s = reinterpret_cast<Shard*>(h);
    LUTVQ_REQUIRE(s && handles, "null pointer");
    for (int r = 0; r < s->nranks; ++r) {
        if (r == s->rank || s->peer[r] != nullptr) continue;
        cudaIpcMemHandle_t ih;
        memcpy(&ih, static_cast<const uint8_t*>(handles) + (size_t)r * LUTVQ_SHARD_HANDLE_BYTES, sizeof(ih));
        void* p = nullptr;
        LUTVQ_CUDA_OK(cudaIpcOpenMemHandle(&p, ih, cudaIpcMemLazyEnablePeerAccess));
        s->peer[r] = static_cast<uint8_t*>(p);
        s->ipc_opened[r] = true;
    }
    s->attached = true;
    return LUTVQ_OK;
}

int lutvq_shard_attach_ptrs(lutvq_shard_t* h, void* const* bases) {
    Shard* s = reinterpret_cast<Shard*>(h);
    LUTVQ_REQUIRE(s && bases, "null pointer");
    for (int r = 0; r < s->nranks; ++r) {
        if (r == s->rank) continue;
        LUTVQ_REQUIRE(bases[r] != nullptr, "null base pointer for rank %d", r);
        s->peer[r] = static_cast<uint8_t*>(bases[r]);
    }
    s->attached = true;
    return LUTVQ_OK;
}

void* lutvq_shard_base(lutvq_shard_t* h) {
    Shard* s = reinterpret_cast<Shard*>(h);
    return s ? s->local : nullptr;
}

void* lutvq_shard_slot(lutvq_shard_t* h, int slot) {
    Shard* s = reinterpret_cast<Shard*>(h);
    if (!s || slot < 0 || slot >= s->n_slots) return nullptr;
    return s->local + kShardFlagBytes + s->slot_bytes * slot;
}

size_t lutvq_shard_slot_bytes(lutvq_shard_t* h) {
    Shard* s = reinterpret_cast<Shard*>(h);
    return s ? s->slot_bytes : 0;
}

int lutvq_sharded_matmat_packed(lutvq_shard_t* h, int slot, size_t row_offset, const void* x, const void* packed,
                                const void* codebooks, const void* scales, const void* bias, int rows_local,
                                int prob_k, int dtype, unsigned flags, void* ws, size_t ws_bytes, void* stream) {
    Shard* s = reinterpret_cast<Shard*>(h);
    LUTVQ_REQUIRE(s != nullptr, "null shard handle");
    LUTVQ_REQUIRE(s->attached, "lutvq_shard_attach has not been called");
    LUTVQ_REQUIRE(slot >= 0 && slot < s->n_slots, "slot %d out of range (n_slots=%d)", slot, s->n_slots);
    LUTVQ_REQUIRE(rows_local > 0 && prob_k > 0 && prob_k % 64 == 0, "bad shape rows_local=%d prob_k=%d", rows_local, prob_k);
    LUTVQ_REQUIRE((row_offset + (size_t)rows_local) * 2 <= s->slot_bytes, "rows [%zu, %zu) do not fit a %zu-byte slot", row_offset, row_offset + rows_local, s->slot_bytes);
    LUTVQ_REQUIRE(row_offset % 8 == 0, "row_offset %zu must be a multiple of 8 (the exchange moves 16-byte pieces)", row_offset);
    LUTVQ_REQUIRE(x && packed && codebooks, "null pointer");
    if (dtype != LUTVQ_F16 && dtype != LUTVQ_BF16) return fail(LUTVQ_EUNSUPPORTED, "dtype %d not supported (fp16=0, bf16=1)", dtype);
    const size_t slot_off = kShardFlagBytes + s->slot_bytes * slot + row_offset * 2;
    void* y_local = s->local + slot_off;
    s->pending[slot].row_off = (unsigned)row_offset;
    s->pending[slot].rows = (unsigned)rows_local;
    s->pending[slot].fused = (flags & LUTVQ_FLAG_PUSH_EPILOGUE) ? 1 : 0;
    s->pending[slot].pushed = 0;
    if (!(flags & LUTVQ_FLAG_PUSH_EPILOGUE))
        // default: the GEMV fills the local slice; lutvq_shard_wait[_mask] forwards it to the peers and waits (one kernel)
        return gemv_2x8_launch(packed, x, y_local, codebooks, scales, bias, rows_local, prob_k, dtype, flags, ws, ws_bytes,
                               static_cast<cudaStream_t>(stream), nullptr);
    GemvPush push;
    push.n = 0;
    for (int r = 0; r < s->nranks; ++r) {
        if (r == s->rank) continue;                                  // the local copy is the kernel's ordinary y store
        push.y[push.n] = s->peer[r] + slot_off;
        push.flag[push.n] = reinterpret_cast<unsigned*>(s->peer[r]) + (size_t)slot * kShardMaxRanks + s->rank;
        ++push.n;
    }
    // the local flag goes through the same mechanism (own buffer is "peer" rank): keeps one code path for P == 1
    push.y[push.n] = s->local + slot_off;                            // harmless duplicate of the y store
    push.flag[push.n] = reinterpret_cast<unsigned*>(s->local) + (size_t)slot * kShardMaxRanks + s->rank;
    ++push.n;
    push.sent = reinterpret_cast<unsigned*>(s->local + kLocalOff) + slot;
    return gemv_2x8_launch(packed, x, y_local, codebooks, scales, bias, rows_local, prob_k, dtype, flags, ws, ws_bytes,
                           static_cast<cudaStream_t>(stream), &push);
}

static int sharded_group_impl(lutvq_shard_t* h, int n, const int* slots, const size_t* row_offsets, const void* x, const void* gate,
                              const void* const* packed, const void* const* codebooks, const void* const* scales,
                              const void* const* bias, const int* rows_local, int prob_k, int dtype, unsigned flags,
                              void* ws, size_t ws_bytes, void* stream) {
    Shard* s = reinterpret_cast<Shard*>(h);
    LUTVQ_REQUIRE(s != nullptr && s->attached, "shard not attached");
    LUTVQ_REQUIRE(n >= 1 && n <= 3, "a group holds 1..3 linears, got %d", n);
    LUTVQ_REQUIRE(slots && row_offsets && x && packed && codebooks && rows_local, "null pointer");
    LUTVQ_REQUIRE(prob_k > 0 && prob_k % 64 == 0, "bad prob_k=%d", prob_k);
    LUTVQ_REQUIRE(!(flags & LUTVQ_FLAG_PUSH_EPILOGUE), "the fused push epilogue takes one linear per launch");
    if (dtype != LUTVQ_F16 && dtype != LUTVQ_BF16) return fail(LUTVQ_EUNSUPPORTED, "dtype %d not supported (fp16=0, bf16=1)", dtype);
    GemvItem items[3];
    for (int i = 0; i < n; ++i) {
        const int slot = slots[i];
        LUTVQ_REQUIRE(slot >= 0 && slot < s->n_slots, "slot %d out of range (n_slots=%d)", slot, s->n_slots);
        for (int j = 0; j < i; ++j) LUTVQ_REQUIRE(slots[j] != slot, "the linears of a group need distinct slots");
        LUTVQ_REQUIRE(rows_local[i] > 0 && (row_offsets[i] + (size_t)rows_local[i]) * 2 <= s->slot_bytes, "rows of linear %d do not fit a slot", i);
        LUTVQ_REQUIRE(row_offsets[i] % 8 == 0, "row offset of linear %d must be a multiple of 8", i);
        LUTVQ_REQUIRE(packed[i] && codebooks[i], "null pointer in linear %d", i);
        void* y_local = s->local + kShardFlagBytes + s->slot_bytes * slot + row_offsets[i] * 2;
        items[i] = GemvItem{packed[i], codebooks[i], y_local, scales ? scales[i] : nullptr, bias ? bias[i] : nullptr, rows_local[i]};
        s->pending[slot].row_off = (unsigned)row_offsets[i];
        s->pending[slot].rows = (unsigned)rows_local[i];
        s->pending[slot].fused = 0;
        s->pending[slot].pushed = 0;
    }
    return gemv_2x8_launch_group(n, items, x, gate, prob_k, dtype, flags, ws, ws_bytes, static_cast<cudaStream_t>(stream), nullptr);
}

int lutvq_sharded_matmat_group_packed(lutvq_shard_t* h, int n, const int* slots, const size_t* row_offsets, const void* x,
                                      const void* const* packed, const void* const* codebooks, const void* const* scales,
                                      const void* const* bias, const int* rows_local, int prob_k, int dtype, unsigned flags,
                                      void* ws, size_t ws_bytes, void* stream) {
    return sharded_group_impl(h, n, slots, row_offsets, x, nullptr, packed, codebooks, scales, bias, rows_local, prob_k, dtype, flags, ws, ws_bytes, stream);
}

int lutvq_sharded_matmat_group_packed_swiglu(lutvq_shard_t* h, int n, const int* slots, const size_t* row_offsets, const void* gate,
                                             const void* up, const void* const* packed, const void* const* codebooks,
                                             const void* const* scales, const void* const* bias, const int* rows_local, int prob_k,
                                             int dtype, unsigned flags, void* ws, size_t ws_bytes, void* stream) {
    LUTVQ_REQUIRE(gate && up, "null gate/up pointer");
    LUTVQ_REQUIRE(dtype == LUTVQ_F16, "the fused SwiGLU input is fp16 only (dtype %d)", dtype);
    return sharded_group_impl(h, n, slots, row_offsets, up, gate, packed, codebooks, scales, bias, rows_local, prob_k, dtype, flags, ws, ws_bytes, stream);
}

}  // extern "C"

namespace lutvq {
int shard_ll_input(lutvq_shard_t* h, int slot, const uint32_t** words, const unsigned** tick, unsigned** error, size_t* capacity_rows) {
    Shard* s = reinterpret_cast<Shard*>(h);
    LUTVQ_REQUIRE(s != nullptr && s->attached, "shard not attached");
    LUTVQ_REQUIRE(slot >= 0 && slot < s->n_slots, "slot %d out of range (n_slots=%d)", slot, s->n_slots);
    unsigned* lstate = reinterpret_cast<unsigned*>(s->local + kLocalOff);
    *words = reinterpret_cast<const uint32_t*>(s->local + s->ll_off + s->ll_slot_bytes * slot);
    *tick = lstate + kLLTickWord;
    *error = lstate + 2 * kShardMaxSlots;
    *capacity_rows = s->ll_slot_bytes / 4;
    return LUTVQ_OK;
}
}  // namespace lutvq

extern "C" {

int lutvq_shard_ll_tick(lutvq_shard_t* h, void* stream) {
    Shard* s = reinterpret_cast<Shard*>(h);
    LUTVQ_REQUIRE(s != nullptr && s->attached, "shard not attached");
    unsigned* lstate = reinterpret_cast<unsigned*>(s->local + kLocalOff);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(1); cfg.blockDim = dim3(1); cfg.dynamicSmemBytes = 0; cfg.stream = static_cast<cudaStream_t>(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    LUTVQ_CUDA_OK(cudaLaunchKernelEx(&cfg, shard_ll_tick_kernel, lstate + kLLTickWord));
    return LUTVQ_OK;
}

int lutvq_shard_ll_unpack(lutvq_shard_t* h, int slot, unsigned seq, void* y, int n, void* stream) {
    const uint32_t* words; const unsigned* tick; unsigned* error; size_t cap;
    if (int rc = shard_ll_input(h, slot, &words, &tick, &error, &cap)) return rc;
    LUTVQ_REQUIRE(y != nullptr && n > 0 && (size_t)n <= cap, "bad y / n=%d (slot holds %zu rows)", n, cap);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((n + 255) / 256 < 64 ? (n + 255) / 256 : 64); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = 0; cfg.stream = static_cast<cudaStream_t>(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    LUTVQ_CUDA_OK(cudaLaunchKernelEx(&cfg, shard_ll_unpack_kernel, words, tick, seq, error, static_cast<unsigned short*>(y), n));
    return LUTVQ_OK;
}

// n <= 3 linears that share an input in ONE launch; every finished output row goes out as an LL word into ALL ranks' gather
// slots (out_slots[i], at this rank's row offset) -- no exchange launch, no flag, no fence.  The input is either a plain
// device vector x (x_slot < 0) or the LL words of an earlier exchange in the local slot x_slot (sequence number x_seq); with
// gate_slot >= 0 the activations are silu(gate) * x, both from LL slots (fp16).
int lutvq_sharded_matmat_group_ll(lutvq_shard_t* h, int n, const int* out_slots, const size_t* row_offsets, const unsigned* out_seq,
                                  const void* x, int x_slot, unsigned x_seq, int gate_slot, unsigned gate_seq,
                                  const void* const* packed, const void* const* codebooks, const void* const* scales,
                                  const void* const* bias, const int* rows_local, int prob_k, int dtype, unsigned flags,
                                  void* ws, size_t ws_bytes, void* stream) {
    Shard* s = reinterpret_cast<Shard*>(h);
    LUTVQ_REQUIRE(s != nullptr && s->attached, "shard not attached");
    LUTVQ_REQUIRE(n >= 1 && n <= 3, "a group holds 1..3 linears, got %d", n);
    LUTVQ_REQUIRE(out_slots && row_offsets && out_seq && packed && codebooks && rows_local, "null pointer");
    LUTVQ_REQUIRE(prob_k > 0 && prob_k % 64 == 0, "bad prob_k=%d", prob_k);
    LUTVQ_REQUIRE(!(flags & LUTVQ_FLAG_PUSH_EPILOGUE), "LL launches do not take the flag-based push");
    // X_READY on an LL launch = "dataflow launch": the activations are LL words, which carry their own validity, so the kernel
    // does not wait for the launch before it to complete before reading them (it polls); it still orders its stores behind it.
    LUTVQ_REQUIRE(!(flags & LUTVQ_FLAG_X_READY) || x_slot >= 0, "LUTVQ_FLAG_X_READY on an LL launch needs the activations as LL words (x_slot >= 0)");
    LUTVQ_REQUIRE((x != nullptr) != (x_slot >= 0), "exactly one of x / x_slot");
    LUTVQ_REQUIRE(gate_slot < 0 || x_slot >= 0, "an LL gate needs an LL x");
    if (dtype != LUTVQ_F16 && dtype != LUTVQ_BF16) return fail(LUTVQ_EUNSUPPORTED, "dtype %d not supported (fp16=0, bf16=1)", dtype);
    unsigned* lstate = reinterpret_cast<unsigned*>(s->local + kLocalOff);
    GemvLL ll;
    ll.nranks = s->nranks;
    for (int r = 0; r < s->nranks; ++r) ll.peer_base[r] = s->peer[r] + s->ll_off;
    ll.tick = lstate + kLLTickWord;
    ll.error = lstate + 2 * kShardMaxSlots;
    GemvItem items[3];
    for (int i = 0; i < n; ++i) {
        const int slot = out_slots[i];
        LUTVQ_REQUIRE(slot >= 0 && slot < s->n_slots, "slot %d out of range (n_slots=%d)", slot, s->n_slots);
        LUTVQ_REQUIRE(slot != x_slot && slot != gate_slot, "a launch must not write the slot it reads");
        for (int j = 0; j < i; ++j) LUTVQ_REQUIRE(out_slots[j] != slot, "the linears of a group need distinct slots");
        LUTVQ_REQUIRE(rows_local[i] > 0 && (row_offsets[i] + (size_t)rows_local[i]) * 4 <= s->ll_slot_bytes, "rows of linear %d do not fit a slot", i);
        LUTVQ_REQUIRE(out_seq[i] < 1024u, "sequence numbers are < 1024");
        LUTVQ_REQUIRE(packed[i] && codebooks[i], "null pointer in linear %d", i);
        ll.out_off[i] = s->ll_slot_bytes * slot + row_offsets[i] * 4;
        ll.out_seq[i] = out_seq[i];
        // y is unused in LL mode (store_y writes the LL words only); keep a valid pointer for the launcher's checks
        items[i] = GemvItem{packed[i], codebooks[i], s->local + s->ll_off, scales ? scales[i] : nullptr, bias ? bias[i] : nullptr, rows_local[i]};
    }
    if (x_slot >= 0) {
        LUTVQ_REQUIRE(x_slot < s->n_slots && (size_t)prob_k * 4 <= s->ll_slot_bytes, "x slot %d out of range or too small for prob_k=%d", x_slot, prob_k);
        LUTVQ_REQUIRE(x_seq < 1024u && gate_seq < 1024u, "sequence numbers are < 1024");
        ll.x_ll = reinterpret_cast<const uint32_t*>(s->local + s->ll_off + s->ll_slot_bytes * x_slot);
        ll.x_seq = x_seq;
        if (gate_slot >= 0) {
            LUTVQ_REQUIRE(gate_slot < s->n_slots && dtype == LUTVQ_F16, "gate slot out of range, or not fp16");
            ll.gate_ll = reinterpret_cast<const uint32_t*>(s->local + s->ll_off + s->ll_slot_bytes * gate_slot);
            ll.gate_seq = gate_seq;
        }
    }
    return gemv_2x8_launch_group(n, items, x, nullptr, prob_k, dtype, flags, ws, ws_bytes, static_cast<cudaStream_t>(stream), nullptr, &ll);
}

int lutvq_shard_wait(lutvq_shard_t* h, int slot, void* stream) {
    Shard* s = reinterpret_cast<Shard*>(h);
    LUTVQ_REQUIRE(s != nullptr && s->attached, "shard not attached");
    LUTVQ_REQUIRE(slot >= 0 && slot < s->n_slots, "slot %d out of range", slot);
    if (!s->pending[slot].fused) {
        ExchangeList ex{};
        ex.n = 1; ex.slot[0] = slot; ex.row_off[0] = s->pending[slot].row_off; ex.rows[0] = s->pending[slot].rows;
        const int pushed = s->pending[slot].pushed;
        s->pending[slot].pushed = 0;
        return launch_exchange(s, ex, static_cast<cudaStream_t>(stream), pushed ? 0 : 1, 1);
    }
    const unsigned* flags = reinterpret_cast<const unsigned*>(s->local) + (size_t)slot * kShardMaxRanks;
    unsigned* local = reinterpret_cast<unsigned*>(s->local + kLocalOff);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(1); cfg.blockDim = dim3(32); cfg.dynamicSmemBytes = 0; cfg.stream = static_cast<cudaStream_t>(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    LUTVQ_CUDA_OK(cudaLaunchKernelEx(&cfg, shard_wait_kernel, flags, local + kShardMaxSlots + slot, local + 2 * kShardMaxSlots,
                                     s->nranks, kWaitTimeoutNs));
    return LUTVQ_OK;
}

int lutvq_shard_wait_mask(lutvq_shard_t* h, unsigned long long slot_mask, void* stream) {
    Shard* s = reinterpret_cast<Shard*>(h);
    LUTVQ_REQUIRE(s != nullptr && s->attached, "shard not attached");
    LUTVQ_REQUIRE(slot_mask != 0 && (s->n_slots == 64 || (slot_mask >> s->n_slots) == 0), "slot mask 0x%llx names slots outside 0..%d", slot_mask, s->n_slots - 1);
    {
        ExchangeList ex{};
        int fused = 0, pushed = 0;
        for (int sl = 0; sl < s->n_slots; ++sl)
            if ((slot_mask >> sl) & 1ull) {
                fused += s->pending[sl].fused;
                pushed += s->pending[sl].pushed;
                s->pending[sl].pushed = 0;
                ex.slot[ex.n] = sl; ex.row_off[ex.n] = s->pending[sl].row_off; ex.rows[ex.n] = s->pending[sl].rows;
                ++ex.n;
            }
        LUTVQ_REQUIRE(fused == 0 || fused == ex.n, "slot mask 0x%llx mixes fused-epilogue and exchange-kernel slots", slot_mask);
        LUTVQ_REQUIRE(pushed == 0 || pushed == ex.n, "slot mask 0x%llx mixes already-pushed and unpushed slots", slot_mask);
        if (fused == 0) return launch_exchange(s, ex, static_cast<cudaStream_t>(stream), pushed ? 0 : 1, 1);
    }
    const unsigned* flags = reinterpret_cast<const unsigned*>(s->local);
    unsigned* local = reinterpret_cast<unsigned*>(s->local + kLocalOff);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(1); cfg.blockDim = dim3(8 * kShardMaxSlots); cfg.dynamicSmemBytes = 0; cfg.stream = static_cast<cudaStream_t>(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    LUTVQ_CUDA_OK(cudaLaunchKernelEx(&cfg, shard_wait_mask_kernel, flags, local + kShardMaxSlots, local + 2 * kShardMaxSlots, s->nranks,
                                     slot_mask, kWaitTimeoutNs));
    return LUTVQ_OK;
}

int lutvq_shard_push_mask(lutvq_shard_t* h, unsigned long long slot_mask, void* stream) {
    Shard* s = reinterpret_cast<Shard*>(h);
    LUTVQ_REQUIRE(s != nullptr && s->attached, "shard not attached");
    LUTVQ_REQUIRE(slot_mask != 0 && (s->n_slots == 64 || (slot_mask >> s->n_slots) == 0), "slot mask 0x%llx names slots outside 0..%d", slot_mask, s->n_slots - 1);
    ExchangeList ex{};
    for (int sl = 0; sl < s->n_slots; ++sl)
        if ((slot_mask >> sl) & 1ull) {
            LUTVQ_REQUIRE(!s->pending[sl].fused && !s->pending[sl].pushed, "slot %d was already pushed", sl);
            ex.slot[ex.n] = sl; ex.row_off[ex.n] = s->pending[sl].row_off; ex.rows[ex.n] = s->pending[sl].rows;
            s->pending[sl].pushed = 1;
            ++ex.n;
        }
    return launch_exchange(s, ex, static_cast<cudaStream_t>(stream), 1, 0);
}

int lutvq_shard_error(lutvq_shard_t* h, unsigned* out) {
    Shard* s = reinterpret_cast<Shard*>(h);
    LUTVQ_REQUIRE(s && out, "null pointer");
    LUTVQ_CUDA_OK(cudaMemcpy(out, s->local + kLocalOff + 2 * kShardMaxSlots * sizeof(unsigned), sizeof(unsigned), cudaMemcpyDeviceToHost));
    return LUTVQ_OK;
}

int lutvq_shard_destroy(lutvq_shard_t* h) {
    Shard* s = reinterpret_cast<Shard*>(h);
    if (!s) return LUTVQ_OK;
    for (int r = 0; r < s->nranks; ++r)
        if (s->ipc_opened[r] && s->peer[r]) cudaIpcCloseMemHandle(s->peer[r]);
    if (s->local) cudaFree(s->local);
    delete s;
    return LUTVQ_OK;
}

}  // extern "C"
