// launch_rate.cu -- how fast can the GPU front end retire a linear chain of tiny kernels inside a CUDA graph?
// (the floor under any per-call GEMV: period per launch with and without programmatic dependent launch)
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(2); } } while (0)

template <int MODE>   // 0: empty; 1: trigger at entry + wait at exit; 2: trigger, spin ~1 us, wait
__global__ void k(int* sink, int spin_ns) {
    extern __shared__ int sm[];
    if (MODE >= 1) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (MODE == 2) {
        unsigned long long t0, t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        do { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); } while (t - t0 < (unsigned long long)spin_ns);
    }
    if (MODE >= 1) asm volatile("griddepcontrol.wait;" ::: "memory");
    if (sink && threadIdx.x == 9999) sink[0] = sm[0];
}

template <int MODE>
static void run(const char* name, int grid, int block, size_t smem, bool pdl, int spin_ns, cudaStream_t st) {
    CK(cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int N = 64, reps = 50;
    cudaGraph_t g; cudaGraphExec_t ex;
    CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeGlobal));
    for (int i = 0; i < N; ++i) {
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
        CK(cudaLaunchKernelEx(&cfg, k<MODE>, (int*)nullptr, spin_ns));
    }
    CK(cudaStreamEndCapture(st, &g));
    CK(cudaGraphInstantiate(&ex, g, 0));
    for (int i = 0; i < 5; ++i) CK(cudaGraphLaunch(ex, st));
    CK(cudaStreamSynchronize(st));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0, st));
    for (int i = 0; i < reps; ++i) CK(cudaGraphLaunch(ex, st));
    CK(cudaEventRecord(e1, st));
    CK(cudaStreamSynchronize(st));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    printf("%-44s grid %4d block %4d smem %6zu pdl %d spin %5d ns: %.3f us per launch\n", name, grid, block, smem, pdl ? 1 : 0, spin_ns, ms * 1e3 / (reps * N));
    CK(cudaGraphExecDestroy(ex)); CK(cudaGraphDestroy(g));
}

int main() {
    cudaStream_t st; CK(cudaStreamCreate(&st));
    run<0>("empty, no pdl", 148, 256, 0, false, 0, st);
    run<0>("empty, pdl attr", 148, 256, 0, true, 0, st);
    run<1>("trigger+wait, pdl", 148, 256, 0, true, 0, st);
    run<1>("trigger+wait, pdl, 64K smem", 148, 256, 65536, true, 0, st);
    run<1>("trigger+wait, pdl, 64K smem, 1 CTA", 1, 256, 65536, true, 0, st);
    for (int ns : {500, 1000, 2000, 4000}) {
        run<2>("trigger, spin, wait, pdl, 64K", 148, 256, 65536, true, ns, st);
        run<2>("trigger, spin, wait, NO pdl, 64K", 148, 256, 65536, false, ns, st);
    }
    run<2>("trigger, spin, wait, pdl, 64K, grid 296", 296, 256, 65536, true, 2000, st);
    run<2>("trigger, spin, wait, pdl, 32K, grid 148", 148, 256, 32768, true, 4000, st);
    run<2>("trigger, spin, wait, pdl, 32K, grid 148, 128thr", 148, 128, 32768, true, 4000, st);
    return 0;
}
