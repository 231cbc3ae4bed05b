// Tensor-core path of ActorCritic (rsl_rl modules/actor_critic.py): the whole MLP -- every Linear + bias + ELU of
// one network -- fused into ONE sm_100a kernel built on tcgen05.mma with accumulators in TMEM.
//
//   * a CTA owns a tile of 128 rows (envs / samples): UMMA M = 128, cta_group::1, bf16 operands, fp32 accumulate;
//   * activations never leave the SM: each layer's output tile goes TMEM -> registers (tcgen05.ld) -> bias + ELU
//     -> bf16 -> shared memory, already in the K-major core-matrix layout the next layer's A descriptor reads;
//   * weights are pre-packed (pf_tc_pack_weights, after every optimizer step) into that same layout, tile by tile
//     in consumption order, so a weight tile is ONE contiguous bulk async copy (cp.async.bulk, the TMA engine,
//     completion by mbarrier transaction bytes) into a small ring that runs ahead of the MMAs;
//   * the layer output is produced in chunks of <= 128 columns, double-buffered in TMEM (2 x 128 columns): the
//     MMAs of chunk c+1 are issued before the epilogue of chunk c, so the tensor pipe works under the epilogue.
//
// Shared-memory operand layout (no swizzle, "interleave"): 8 rows x 16 bytes core matrices, stored contiguously
// (128 B); for an operand with R rows: element (r, k) of bf16 at  (k/8) * (R*16) + r*16 + (k%8)*2  bytes.
// UMMA descriptor: leading byte offset = R*16 (next 8 K-columns), stride byte offset = 128 (next 8 rows).
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdio.h>
#include <stdlib.h>
#include "pf_rl_internal.h"

void pf_set_error(const char* msg);

#define TC_ROWS 128          // rows per CTA (UMMA M)
#define TC_NCHUNK 128        // output columns per accumulator chunk (UMMA N <= 128)
#define TC_KTILE 32          // K extent of one weight tile
#define TC_TILE_BYTES (TC_NCHUNK * TC_KTILE * 2)
#define TC_MAX_STAGES 4
#define TC_EPI_WARPS 8       // staging + epilogue warps: 4 TMEM lane quarters x 2 column halves
#define TC_THREADS (TC_EPI_WARPS * 32 + 32)   // + one warp that owns TMEM, issues the bulk copies and the MMAs

struct TcPlan {
    int layers;
    int dims[PF_MAX_MLP_LAYERS + 1];     // true widths
    int pdims[PF_MAX_MLP_LAYERS + 1];    // padded to a multiple of 16
    int stages;
    int bufA_bytes, bufB_bytes;          // activations of odd / even layer index
    int smem_bytes;
    long long w_off[PF_MAX_MLP_LAYERS];      // byte offsets of each layer's packed weight tiles inside the blob
    long long blob_bytes;
};

struct TcState { TcPlan plan[2]; };

static inline int pad16(int x) { return (x + 15) & ~15; }

// ------------------------------------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// bounded wait: a protocol bug must end in a trap, never in a hung GPU
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok = 0;
    for (uint32_t it = 0; it < 50000000u; ++it) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (ok) return;
    }
    printf("pf_mlp_tc: mbarrier wait timed out (block %d thread %d bar %u parity %u)\n", blockIdx.x, threadIdx.x, bar, parity);
    __trap();
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, no-swizzle shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
// instruction descriptor: D fp32, A/B bf16, both K-major, M = 128 (cute::UMMA::InstrDescriptor bit layout)
__device__ __forceinline__ uint32_t make_idesc(int n) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(TC_ROWS >> 4) << 24);
}

// ------------------------------------------------------------------------------------------------ kernel
struct TcArgs {
    TcPlan plan;
    const unsigned char* blob;   // packed bf16 weight + bias tiles
    const float* x; long long ldx; int rows;
    float* out;                  // [rows, dims[layers]] fp32
    float* copy_to;              // optional [rows, dims[0]] copy of the input rows (rollout storage)
};

// Tiles of one (layer l, column chunk c), contiguous in the blob: first the BIAS tile (nt x 16: column 0 = bf16(b),
// column 1 = bf16(b - bf16(b)), multiplied by a constant "ones" A operand, so the bias add is one extra K=16 MMA
// per chunk instead of epilogue work), then the K tiles of TC_KTILE columns.
__device__ __forceinline__ long long chunk_offset(const TcPlan& P, int l, int c) {
    return P.w_off[l] + (long long)c * TC_NCHUNK * (P.pdims[l] + 16) * 2;
}

__device__ __forceinline__ float elu_fast(float x) { return x > 0.0f ? x : __expf(x) - 1.0f; }

__global__ void __launch_bounds__(TC_THREADS, 1) pf_mlp_tc_kernel(const TcArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    const TcPlan& P = a.plan;
    unsigned char* bufA = smem;
    unsigned char* bufB = smem + P.bufA_bytes;
    unsigned char* ones = bufB + P.bufB_bytes;                 // [128 x 16] bf16: columns 0,1 = 1
    unsigned char* ring = ones + TC_ROWS * 16 * 2;
    uint64_t* bars = reinterpret_cast<uint64_t*>(ring + P.stages * TC_TILE_BYTES);
    const int S = P.stages;
    const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + TC_MAX_STAGES), accum0 = smem_u32(bars + 2 * TC_MAX_STAGES);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * TC_MAX_STAGES + 2);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const bool leader = tid == TC_EPI_WARPS * 32;              // lane 0 of the last warp
    const int row0 = blockIdx.x * TC_ROWS;
    const int nrows = min(TC_ROWS, a.rows - row0);
    const int L = P.layers;

    if (leader) {
        for (int s = 0; s < S; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 1); }
        mbar_init(accum0, 1); mbar_init(accum0 + 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == TC_EPI_WARPS) {
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(256u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    } else {
        // ---- stage the input rows, coalesced: fp32 global -> bf16 core-matrix layout (layer 0 reads bufB) ----
        const int K0 = P.dims[0], half = P.pdims[0] >> 1;                    // pairs of columns per padded row
        const uint32_t magic = (uint32_t)((0x100000000ull + (uint32_t)half - 1) / (uint32_t)half);
        for (int i = tid; i < TC_ROWS * half; i += TC_EPI_WARPS * 32) {
            const int r = (int)__umulhi((uint32_t)i, magic), k = (i - r * half) * 2;
            float v0 = 0.0f, v1 = 0.0f;
            if (r < nrows) {
                const float* xr = a.x + (long long)(row0 + r) * a.ldx;
                float* cr = a.copy_to ? a.copy_to + (long long)(row0 + r) * K0 : nullptr;
                if (k < K0) { v0 = __ldg(xr + k); if (cr) cr[k] = v0; }
                if (k + 1 < K0) { v1 = __ldg(xr + k + 1); if (cr) cr[k + 1] = v1; }
            }
            __nv_bfloat162 pr = __floats2bfloat162_rn(v0, v1);
            *reinterpret_cast<uint32_t*>(bufB + (size_t)(k >> 3) * (TC_ROWS * 16) + r * 16 + (k & 7) * 2) = *reinterpret_cast<uint32_t*>(&pr);
        }
        for (int i = tid; i < TC_ROWS * 2; i += TC_EPI_WARPS * 32)           // ones operand: 2 column blocks x 128 rows x 16 B
            *reinterpret_cast<uint4*>(ones + i * 16) = i < TC_ROWS ? make_uint4(0x3F803F80u, 0u, 0u, 0u) : make_uint4(0u, 0u, 0u, 0u);
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // total number of tiles, and the loader's cursor over (layer, chunk, tile-in-chunk)
    int total_tiles = 0;
    for (int l = 0; l < L; ++l) total_tiles += ((P.pdims[l + 1] + TC_NCHUNK - 1) / TC_NCHUNK) * ((P.pdims[l] + TC_KTILE - 1) / TC_KTILE + 1);
    int ld_idx = 0, ld_l = 0, ld_c = 0, ld_t = 0;
    auto issue_load = [&]() {         // leader only: tile ld_idx -> slot ld_idx % S
        const int Kp = P.pdims[ld_l], Np = P.pdims[ld_l + 1];
        const int nt = min(TC_NCHUNK, Np - ld_c * TC_NCHUNK);
        const int kt = ld_t == 0 ? 16 : min(TC_KTILE, Kp - (ld_t - 1) * TC_KTILE);
        const long long off = chunk_offset(P, ld_l, ld_c) + (ld_t == 0 ? 0 : (long long)nt * 32 + (long long)(ld_t - 1) * nt * TC_KTILE * 2);
        const uint32_t bytes = (uint32_t)(nt * kt * 2);
        const int slot = ld_idx % S;
        mbar_expect_tx(full0 + 8 * slot, bytes);
        bulk_g2s(smem_u32(ring + slot * TC_TILE_BYTES), a.blob + off, bytes, full0 + 8 * slot);
        ++ld_idx;
        if (++ld_t == (Kp + TC_KTILE - 1) / TC_KTILE + 1) {
            ld_t = 0;
            if (++ld_c == (Np + TC_NCHUNK - 1) / TC_NCHUNK) { ld_c = 0; ++ld_l; }
        }
    };
    if (leader)
        for (int s = 0; s < S && ld_idx < total_tiles; ++s) issue_load();

    int use_idx = 0, chunk_seq = 0;
    for (int l = 0; l < L; ++l) {
        unsigned char* in = (l & 1) ? bufA : bufB;
        unsigned char* outb = (l & 1) ? bufB : bufA;
        const int Kp = P.pdims[l], Np = P.pdims[l + 1];
        const int nch = (Np + TC_NCHUNK - 1) / TC_NCHUNK, ntl = (Kp + TC_KTILE - 1) / TC_KTILE + 1;
        const bool last = l == L - 1;
        for (int c = 0; c <= nch; ++c) {
            // ---- MMAs of chunk c (async), issued by one thread ----
            if (c < nch) {
                const int nt = min(TC_NCHUNK, Np - c * TC_NCHUNK);
                const int buf = chunk_seq & 1;
                if (leader) {
                    tc_fence_after();
                    const uint32_t idesc = make_idesc(nt);
                    // leading byte offset = next 8 K-columns, stride byte offset = next 8 rows (verified on B200)
                    const uint32_t a_lbo = (uint32_t)(TC_ROWS * 16), a_sbo = 128u, b_lbo = (uint32_t)(nt * 16), b_sbo = 128u;
                    for (int t = 0; t < ntl; ++t) {
                        const int kt = t == 0 ? 16 : min(TC_KTILE, Kp - (t - 1) * TC_KTILE);
                        const int slot = use_idx % S;
                        mbar_wait(full0 + 8 * slot, (uint32_t)((use_idx / S) & 1));
                        tc_fence_after();
                        const uint32_t bs = smem_u32(ring + slot * TC_TILE_BYTES);
                        const uint32_t as = t == 0 ? smem_u32(ones) : smem_u32(in) + (uint32_t)((t - 1) * (TC_KTILE / 8) * (TC_ROWS * 16));
                        for (int kk = 0; kk < kt / 16; ++kk) {
                            const uint64_t ad = make_desc(as + (uint32_t)(kk * 2 * (TC_ROWS * 16)), a_lbo, a_sbo);
                            const uint64_t bd = make_desc(bs + (uint32_t)(kk * 2 * (nt * 16)), b_lbo, b_sbo);
                            umma_bf16(tmem_base + (uint32_t)(buf * TC_NCHUNK), ad, bd, idesc, (t > 0 || kk > 0) ? 1u : 0u);
                        }
                        umma_commit(empty0 + 8 * slot);            // slot is free once these MMAs have read it
                        // refill the slot of the PREVIOUS tile (its MMAs are done or nearly so): keeps the issue
                        // of this tile's MMAs off the wait
                        if (use_idx >= 1 && ld_idx < total_tiles && ld_idx == use_idx - 1 + S) {
                            const int ps = (use_idx - 1) % S;
                            mbar_wait(empty0 + 8 * ps, (uint32_t)(((use_idx - 1) / S) & 1));
                            issue_load();
                        }
                        ++use_idx;
                    }
                    umma_commit(accum0 + 8 * buf);
                }
                ++chunk_seq;
            }
            // ---- epilogue of chunk c-1: 8 warps = 4 TMEM lane quarters x 2 column halves; the issuing warp only
            //      joins the barrier ----
            if (c > 0 && warp == TC_EPI_WARPS) { tc_fence_before(); __syncthreads(); }
            if (c > 0 && warp < TC_EPI_WARPS) {
                const int pc = c - 1;
                const int seq = chunk_seq - (c < nch ? 2 : 1);
                const int buf = seq & 1;
                const int nt = min(TC_NCHUNK, Np - pc * TC_NCHUNK);
                const int q = warp & 3, h = warp >> 2, row = q * 32 + lane;
                mbar_wait(accum0 + 8 * buf, (uint32_t)((seq >> 1) & 1));
                tc_fence_after();
                const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * TC_NCHUNK);
                for (int g = h * 32; g < nt; g += 64) {
                    uint32_t v[32];
                    const int w = min(32, nt - g);
                    if (w == 32) tmem_ld32(trow + g, v); else tmem_ld16(trow + g, v);
                    tmem_ld_wait();
                    const int n0 = pc * TC_NCHUNK + g;
                    if (last) {
                        const int no = P.dims[L];
                        if (row < nrows)
                            for (int j = 0; j < w; ++j)
                                if (n0 + j < no) a.out[(long long)(row0 + row) * no + n0 + j] = __uint_as_float(v[j]);
                    } else {
                        for (int j = 0; j < w; j += 8) {
                            uint32_t pk[4];
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                __nv_bfloat162 p = __floats2bfloat162_rn(elu_fast(__uint_as_float(v[j + 2 * i])),
                                                                         elu_fast(__uint_as_float(v[j + 2 * i + 1])));
                                pk[i] = *reinterpret_cast<uint32_t*>(&p);
                            }
                            *reinterpret_cast<uint4*>(outb + (size_t)((n0 + j) >> 3) * (TC_ROWS * 16) + row * 16) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                        }
                    }
                }
                fence_async_smem();
                tc_fence_before();
                __syncthreads();
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == TC_EPI_WARPS) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(256u) : "memory");
    }
}

// ------------------------------------------------------------------------------------------------ packing
// fp32 nn.Linear weight [out, in] + bias -> bf16 tiles in consumption order (see chunk_offset)
__global__ void pf_tc_pack_kernel(const TcPlan P, int l, const float* __restrict__ W, const float* __restrict__ b,
                                  unsigned char* blob) {
    const int Kp = P.pdims[l], Np = P.pdims[l + 1], K = P.dims[l], N = P.dims[l + 1];
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)Np * Kp) return;
    const int n = (int)(i / Kp), k = (int)(i % Kp);
    const int c = n / TC_NCHUNK, nl = n % TC_NCHUNK, nt = min(TC_NCHUNK, Np - c * TC_NCHUNK);
    const int kt_i = k / TC_KTILE, kk = k % TC_KTILE;
    const long long base = P.w_off[l] + (long long)c * TC_NCHUNK * (Kp + 16) * 2;
    const long long off = base + (long long)nt * 32 + (long long)kt_i * nt * TC_KTILE * 2 +
                          ((long long)(kk >> 3) * (nt >> 3) + (nl >> 3)) * 128 + (nl & 7) * 16 + (kk & 7) * 2;
    const float v = (n < N && k < K) ? W[(long long)n * K + k] : 0.0f;
    *reinterpret_cast<__nv_bfloat16*>(blob + off) = __float2bfloat16_rn(v);
    if (k == 0) {           // bias tile: columns 0 / 1 = high / low bf16 parts of the bias, the rest stays zero
        const float bv = n < N ? b[n] : 0.0f;
        const __nv_bfloat16 hi = __float2bfloat16_rn(bv);
        const __nv_bfloat16 lo = __float2bfloat16_rn(bv - __bfloat162float(hi));
        __nv_bfloat16* d = reinterpret_cast<__nv_bfloat16*>(blob + base + (long long)(nl >> 3) * 128 + (nl & 7) * 16);
        d[0] = hi; d[1] = lo;
    }
}

static int tc_fail(int code, const char* msg) { pf_set_error(msg); return code; }

static int build_plan(TcPlan& P, const PfNet& n) {
    P.layers = n.layers;
    for (int i = 0; i <= n.layers; ++i) { P.dims[i] = n.dims[i]; P.pdims[i] = pad16(n.dims[i]); }
    int wa = 16, wb = 16;
    for (int i = 0; i < n.layers; ++i) {
        if (i & 1) wa = P.pdims[i] > wa ? P.pdims[i] : wa;
        else wb = P.pdims[i] > wb ? P.pdims[i] : wb;
    }
    P.bufA_bytes = TC_ROWS * wa * 2;
    P.bufB_bytes = TC_ROWS * wb * 2;
    const int limit = 227 * 1024 - 256 - TC_ROWS * 16 * 2;
    int S = (limit - P.bufA_bytes - P.bufB_bytes) / TC_TILE_BYTES;
    if (S < 2) return 1;
    P.stages = S > TC_MAX_STAGES ? TC_MAX_STAGES : S;
    P.smem_bytes = P.bufA_bytes + P.bufB_bytes + TC_ROWS * 16 * 2 + P.stages * TC_TILE_BYTES + 256;
    long long off = 0;
    for (int l = 0; l < n.layers; ++l) { P.w_off[l] = off; off += (long long)(P.pdims[l] + 16) * P.pdims[l + 1] * 2; }
    P.blob_bytes = off;
    return 0;
}

int pf_tc_create(pf_rl_ctx* c) {
    TcState* t = new TcState();
    int max_smem = 0;
    for (int k = 0; k < 2; ++k) {
        if (build_plan(t->plan[k], c->net[k])) {
            delete t;
            return tc_fail(6, "pf_rl_create: MLP too wide for the fused tensor-core kernel (activations of one 128-row tile must fit in shared memory)");
        }
        if (cudaMalloc(&c->net[k].tc_weights, (size_t)t->plan[k].blob_bytes) != cudaSuccess) { delete t; return tc_fail(2, "pf_rl_create: cudaMalloc of the packed weights failed"); }
        cudaMemset(c->net[k].tc_weights, 0, (size_t)t->plan[k].blob_bytes);      // padding columns of the bias tiles stay zero
        c->net[k].tc_weight_bytes = (size_t)t->plan[k].blob_bytes;
        max_smem = t->plan[k].smem_bytes > max_smem ? t->plan[k].smem_bytes : max_smem;
    }
    if (cudaFuncSetAttribute(pf_mlp_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem) != cudaSuccess) {
        delete t;
        return tc_fail(2, "pf_rl_create: cannot reserve shared memory for the tensor-core kernel");
    }
    c->tc = t;
    return 0;
}

void pf_tc_destroy(pf_rl_ctx* c) {
    for (int k = 0; k < 2; ++k) cudaFree(c->net[k].tc_weights);
    delete static_cast<TcState*>(c->tc);
    c->tc = nullptr;
}

int pf_tc_pack_weights(pf_rl_ctx* c, const float* params, cudaStream_t st) {
    TcState* t = static_cast<TcState*>(c->tc);
    for (int k = 0; k < 2; ++k) {
        const TcPlan& P = t->plan[k];
        for (int l = 0; l < P.layers; ++l) {
            const long long n = (long long)P.pdims[l] * P.pdims[l + 1];
            pf_tc_pack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(P, l, params + c->net[k].w_off[l], params + c->net[k].b_off[l],
                                                                          static_cast<unsigned char*>(c->net[k].tc_weights));
        }
    }
    return cudaGetLastError() == cudaSuccess ? 0 : tc_fail(2, "pf_tc_pack_weights: launch failed");
}

int pf_tc_mlp_forward(pf_rl_ctx* c, int net, const float* params, const float* x, long long ldx, int rows, float* out,
                      float* copy_to, cudaStream_t st) {
    (void)params;
    TcState* t = static_cast<TcState*>(c->tc);
    TcArgs a;
    a.plan = t->plan[net];
    a.blob = static_cast<const unsigned char*>(c->net[net].tc_weights);
    a.x = x; a.ldx = ldx; a.rows = rows; a.out = out; a.copy_to = copy_to;
    pf_mlp_tc_kernel<<<(rows + TC_ROWS - 1) / TC_ROWS, TC_THREADS, a.plan.smem_bytes, st>>>(a);
    return cudaGetLastError() == cudaSuccess ? 0 : tc_fail(2, "pf_tc_mlp_forward: launch failed");
}
