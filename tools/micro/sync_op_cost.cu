// Issue-side cost (SM cycles per operation, one converged warp, the rest of the SM idle) of the synchronisation operations the
// MMA-issuing warp of pf_mlp_tc.cu executes per weight tile / per group.  Numbers feed DESIGN.md section 4.3.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o sync_op_cost tools/micro/sync_op_cost.cu && ./sync_op_cost
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ uint32_t try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok;
}
__device__ __forceinline__ uint32_t test_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok;
}
__device__ __forceinline__ void commit(uint32_t bar) { asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory"); }

#define REPS 64
__global__ void __launch_bounds__(32, 1) bench(long long* out) {
    __shared__ uint64_t bars[4];
    __shared__ uint32_t tmem_slot;
    const int lane = threadIdx.x;
    if (lane == 0) {
        for (int i = 0; i < 4; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[i])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(32u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    __syncwarp();
    // phase 0 of bars[0] completes now: every wait below is on a completed phase
    if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&bars[0])) : "memory");
    __syncwarp();
    const uint32_t b0 = smem_u32(&bars[0]), b1 = smem_u32(&bars[1]);
    uint32_t acc = 0;
    long long t[8];
    t[0] = clock64();
    for (int i = 0; i < REPS; ++i) acc += try_wait(b0, 0);
    t[1] = clock64();
    for (int i = 0; i < REPS; ++i) acc += test_wait(b0, 0);
    t[2] = clock64();
    for (int i = 0; i < REPS; ++i) acc += elect_one() ? 1u : 0u;
    t[3] = clock64();
    for (int i = 0; i < REPS; ++i) asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    t[4] = clock64();
    for (int i = 0; i < REPS; ++i) { if (elect_one()) commit(b1); }      // count 1: every commit completes a phase of bars[1]
    t[5] = clock64();
    for (int i = 0; i < REPS; ++i) { if (elect_one()) { commit(b1); commit(smem_u32(&bars[2])); } }
    t[6] = clock64();
    for (int i = 0; i < REPS; ++i) __syncwarp();
    t[7] = clock64();
    if (lane == 0) { for (int i = 0; i < 7; ++i) out[i] = t[i + 1] - t[i]; out[7] = acc; }
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_slot), "r"(32u) : "memory");
}

int main() {
    long long* d; long long h[8];
    cudaMalloc(&d, 64);
    bench<<<1, 32>>>(d); bench<<<1, 32>>>(d);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
    cudaMemcpy(h, d, 64, cudaMemcpyDeviceToHost);
    const char* names[7] = {"mbarrier.try_wait (completed phase)", "mbarrier.test_wait (completed phase)", "elect.sync", "tcgen05.fence::after_thread_sync",
                            "elect + tcgen05.commit", "elect + 2 x tcgen05.commit", "__syncwarp"};
    for (int i = 0; i < 7; ++i) printf("%-40s %7.1f cycles each (loop of %d, includes ~6 cycles of loop overhead)\n", names[i], (double)h[i] / REPS, REPS);
    return 0;
}
