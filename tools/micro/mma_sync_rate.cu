// Microbenchmark: sustained mma.sync.m16n8k16 bf16 rate per SM on sm_100a (decides whether a register-resident warp-level
// MLP can carry the small flat policy nets).  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_sync_rate mma_sync_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <cstdint>

__global__ void __launch_bounds__(512) k(float* out, int iters, uint32_t seed) {
    uint32_t a0 = seed + threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
    float c[8][4];
    for (int j = 0; j < 8; ++j) for (int i = 0; i < 4; ++i) c[j][i] = 0.f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(c[j][0]), "+f"(c[j][1]), "+f"(c[j][2]), "+f"(c[j][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    float s = 0.f;
    for (int j = 0; j < 8; ++j) for (int i = 0; i < 4; ++i) s += c[j][i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
    float* out; cudaMalloc(&out, 148 * 4 * 512 * sizeof(float));
    for (int warps = 4; warps <= 16; warps *= 2) {
        const int iters = 4096;
        k<<<148, warps * 32>>>(out, 16, 1u);
        cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
        cudaEventRecord(a); k<<<148, warps * 32>>>(out, iters, 1u); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        const double mmas = 148.0 * warps * iters * 8, flops = mmas * 16 * 8 * 16 * 2;
        printf("warps/SM %2d: %.3f ms, %.1f TFLOP/s, %.2f MMA/clk/SM (at 1.965 GHz)\n", warps, ms, flops / ms / 1e9,
               mmas / 148 / (ms * 1e-3 * 1.965e9));
    }
    return 0;
}
