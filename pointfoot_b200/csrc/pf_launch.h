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
// Optional device timeline (tools/step_timeline.py), compiled in with -DPF_TIMELINE (PF_BUILD_TIMELINE=1 python
// __graft_entry__.py --force): when a log buffer is installed (pf_debug_timeline), thread 0 of CTA 0 of every kernel
// appends (tag << 48 | globaltimer ns) records -- log[0] is the record count.  Each translation unit has its own copy
// of the pointer.  Without the define every mark is an empty function.
static __device__ unsigned long long* g_pf_timeline = nullptr;
__device__ __forceinline__ void pf_timeline_raw(unsigned tag) {      // the calling thread appends a record
#ifndef PF_TIMELINE
    (void)tag; return;          // compiled out of the product build: even the NULL test of the log pointer (a global load
                                // per mark on the critical thread) was measured at ~4 us per policy step
#endif
    unsigned long long* p = g_pf_timeline;
    if (p) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        const unsigned long long i = atomicAdd(p, 1ull);
        if (i < 4094) p[1 + i] = ((unsigned long long)tag << 48) | (t & 0xFFFFFFFFFFFFull);
    }
}
__device__ __forceinline__ void pf_timeline_raw_if(bool on, unsigned tag) { if (on) pf_timeline_raw(tag); }
__device__ __forceinline__ void pf_timeline(unsigned tag) {
    if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) pf_timeline_raw(tag);
}
static inline cudaError_t pf_timeline_install(unsigned long long* device_log) {
#ifdef PF_TIMELINE
    return cudaMemcpyToSymbol(g_pf_timeline, &device_log, sizeof(device_log));
#else
    return device_log ? cudaErrorNotSupported : cudaSuccess;
#endif
}
__device__ __forceinline__ void pf_grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pf_grid_dep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif

// Launch counter behind bench.py's `gpu_launches`: every kernel launch of the library (pf_launch_pdl and the plain <<<>>> sites,
// through PF_LAUNCHED()) bumps it on the host; pf_debug_launch_count() reads it.  Launches recorded during a stream
// capture count once, at capture; a graph replay is counted by the caller as that many launches per replay.
extern unsigned long long g_pf_launch_count;
#define PF_LAUNCHED() ((void)++g_pf_launch_count)

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
    PF_LAUNCHED();
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}
