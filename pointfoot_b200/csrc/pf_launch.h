// Programmatic dependent launch (PDL) for the chain of small kernels one policy step is made of (4 x torques,
// post_physics, finalize, act): a kernel launched with the attribute may become resident while its predecessor in
// the stream is still running; it must execute pf_grid_dep_wait() before its first access to global memory, which
// returns once the predecessor has completed and its writes are visible.  Every kernel fires
// pf_grid_dep_launch() at its very start, so the successor's CTAs are already sitting at their wait (launch latency,
// prologue, shared-memory setup paid) when the predecessor's last CTA retires.  Without the attribute both
// instructions are no-ops.  Stream capture turns the pairs into programmatic graph edges.  PF_NO_PDL=1 disables it.
#pragma once
#include <cuda_runtime.h>
#include <stdlib.h>
#include <utility>

#ifdef __CUDACC__
__device__ __forceinline__ void pf_grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pf_grid_dep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif

static inline bool pf_pdl_enabled() {
    static const bool on = [] { const char* e = getenv("PF_NO_PDL"); return !(e && atoi(e) != 0); }();
    return on;
}

template <typename... KArgs, typename... Args>
static inline cudaError_t pf_launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                        Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pf_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}
