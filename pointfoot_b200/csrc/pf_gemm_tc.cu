// Tensor-core GEMM of the PPO update (rsl_rl algorithms/ppo.py update(): the forward over a minibatch, loss.backward()'s
// dgrad and wgrad), PF_RL_BF16_TC precision: bf16 operands, fp32 accumulation in TMEM, tcgen05.mma.
//
//   C[M,N] = epilogue( sum_k A(m,k) * B(k,n) )      same contract as the fp32 kernel's GemmArgs (pf_rl.cu)
//
// Two steps per GEMM:
//   1. pf_tc_pack_operand_kernel turns each operand (any element strides, optional row gather through the minibatch
//      permutation) into bf16 tiles in the UMMA "interleave" layout, tile by tile in consumption order:
//      [row tile][k block of 8][row in tile][8 k]  -- so that
//   2. pf_tc_gemm_kernel moves one pipeline stage (A: 128 rows x 64 k = 16 KB, B: 128 or 256 rows x 64 k) with ONE bulk
//      async copy per operand into a 4-deep shared-memory ring and issues four K = 16 MMAs (N = 128 / 256) per stage.
//      Warp roles: 4 epilogue warps (TMEM lane quarters), 1 MMA warp, 1 loader warp; split-K over blockIdx.z for
//      the weight gradients, partials reduced in fixed order by pf_reduce_partials_kernel (deterministic).
// The epilogue applies bias / ELU / ELU'(saved activation) exactly like the fp32 kernel and writes fp32 row-major.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdlib.h>
#include <string.h>
#include "pf_rl_internal.h"
#include "pf_tc_ptx.cuh"
#include "pf_launch.h"

void pf_set_error(const char* msg);

#define GT_ROWS 128
#define GT_KSTEP 64
#define GT_STAGES 4                          // deepest ring; short-K launches use fewer (two CTAs per SM then fit)
#define GT_EPI_WARPS 4
#define GT_THREADS ((GT_EPI_WARPS + 2) * 32)

struct TcGemmState {
    __nv_bfloat16* ap; __nv_bfloat16* bp;      // packed operands
    size_t ap_bytes, bp_bytes;
};

// ------------------------------------------------------------------------------------------------ packing
// dst 16-byte slot index = ((row_tile * kblocks + kb) * TH + row_in_tile): thread idx writes slot idx.
// value(r, k) = src[(gr ? gr[r] : r) * s_r + (gk ? gk[k] : k) * s_k], zero outside [R) x [k_lo, k_hi).
__global__ void __launch_bounds__(256) pf_tc_pack_operand_kernel(const float* __restrict__ src, long long s_r, long long s_k,
                                                                  const int* __restrict__ gr, const int* __restrict__ gk,
                                                                  int R, int K, int TH, int kblocks, long long slots, uint4* dst) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= slots) return;
    const int rl = (int)(idx % TH);
    const long long t = idx / TH;
    const int kb = (int)(t % kblocks);
    const int r = (int)(t / kblocks) * TH + rl;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = 0.0f;
    if (r < R) {
        const float* row = src + (long long)(gr ? gr[r] : r) * s_r;
        const int k0 = kb * 8;
        if (!gk && s_k == 1 && k0 + 8 <= K && ((reinterpret_cast<uintptr_t>(row + k0) & 15) == 0)) {
            const float4 a = __ldg(reinterpret_cast<const float4*>(row + k0)), b = __ldg(reinterpret_cast<const float4*>(row + k0 + 4));
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (k0 + j < K) v[j] = __ldg(row + (long long)(gk ? gk[k0 + j] : k0 + j) * s_k);
        }
    }
    uint32_t pk[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        __nv_bfloat162 p = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        pk[i] = *reinterpret_cast<uint32_t*>(&p);
    }
    dst[idx] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
}

// ------------------------------------------------------------------------------------------------ GEMM
struct TcGemmArgs {
    const unsigned char* ap; const unsigned char* bp;
    int kblocks;                  // 8-column blocks per packed row tile (K padded to a multiple of GT_KSTEP)
    int stages_per_split;         // K steps of GT_KSTEP per blockIdx.z slice
    int nt;                       // B rows per tile: 128 or 256 (UMMA N)
    int stages;                   // ring depth of this launch (<= GT_STAGES)
    float* C; long long ldc; long long split_stride;
    int M, N;
    const float* bias; int elu;
    const float* elu_out; long long ld_elu;
};

// ELU with the exponential on the MUFU unit (ex2.approx), branch-free; the operands were rounded to bf16 anyway
__device__ __forceinline__ float gt_elu(float x) {
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(fminf(x, 0.0f) * 1.4426950408889634f));
    return fmaxf(x, e - 1.0f);          // x > 0: e = 1 -> max(x, 0) = x;  x <= 0: exp(x) - 1 >= x
}

__global__ void __launch_bounds__(GT_THREADS, 2) pf_tc_gemm_kernel(const __grid_constant__ TcGemmArgs g) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int NT = g.nt;
    const uint32_t a_bytes = GT_ROWS * GT_KSTEP * 2, b_bytes = (uint32_t)NT * GT_KSTEP * 2, stage_bytes = a_bytes + b_bytes;
    const int S = g.stages;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S * stage_bytes);
    const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + GT_STAGES), acc_bar = smem_u32(bars + 2 * GT_STAGES);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * GT_STAGES + 1);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int mt = blockIdx.y, ntile = blockIdx.x, z = blockIdx.z;
    const int total_stages = g.kblocks / (GT_KSTEP / 8);
    const int s_beg = z * g.stages_per_split;
    const int s_end = min(total_stages, s_beg + g.stages_per_split);
    const int nst = s_end - s_beg;

    if (warp == GT_EPI_WARPS + 1 && lane == 0) {
        for (int s = 0; s < S; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 1); }
        mbar_init(acc_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == GT_EPI_WARPS) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"((uint32_t)NT) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == GT_EPI_WARPS + 1) {
        // ---- loader: one bulk copy per operand per stage ----
        const unsigned char* a_src = g.ap + ((size_t)mt * g.kblocks + (size_t)s_beg * (GT_KSTEP / 8)) * (GT_ROWS * 16);
        const unsigned char* b_src = g.bp + ((size_t)ntile * g.kblocks + (size_t)s_beg * (GT_KSTEP / 8)) * ((size_t)NT * 16);
        for (int s = 0; s < nst; ++s) {
            const int slot = s % S;
            if (s >= S) mbar_wait(empty0 + 8 * slot, (uint32_t)((s / S) - 1) & 1u);
            if (elect_one()) {
                mbar_expect_tx(full0 + 8 * slot, stage_bytes);
                bulk_g2s(smem_u32(smem + slot * stage_bytes), a_src + (size_t)s * a_bytes, a_bytes, full0 + 8 * slot);
                bulk_g2s(smem_u32(smem + slot * stage_bytes + a_bytes), b_src + (size_t)s * b_bytes, b_bytes, full0 + 8 * slot);
            }
        }
    } else if (warp == GT_EPI_WARPS) {
        // ---- MMA issue: four K = 16 steps per stage ----
        const uint32_t idesc = make_idesc(NT);
        const uint64_t b_lbo = (uint64_t)((uint32_t)(NT * 16) >> 4) << 16;
        for (int s = 0; s < nst; ++s) {
            const int slot = s % S;
            mbar_wait(full0 + 8 * slot, (uint32_t)(s / S) & 1u);
            if (elect_one()) {
                const uint64_t ad = make_desc(smem_u32(smem + slot * stage_bytes), GT_ROWS * 16, 128u);
                const uint64_t bd = make_desc(smem_u32(smem + slot * stage_bytes + a_bytes), 0u, 128u) + b_lbo;
                if (s == 0) umma_bf16(tmem_base, ad, bd, idesc, 0u); else umma_bf16_acc(tmem_base, ad, bd, idesc);
                if (NT == 256) issue_mmas<256, 512, 3>(tmem_base, ad + 256, bd + 512, idesc);
                else issue_mmas<256, 256, 3>(tmem_base, ad + 256, bd + 256, idesc);
                umma_commit(empty0 + 8 * slot);
                if (s == nst - 1) umma_commit(acc_bar);
            }
            __syncwarp();
        }
    } else {
        // ---- epilogue: 4 warps, warp q owns TMEM lanes [32q, 32q+32) = rows of the tile ----
        // TMEM hands a thread one ROW (32 consecutive columns); written as is, every store instruction of a warp
        // would touch 32 different rows.  Each 32 x 32 piece therefore goes through a padded shared-memory square
        // (the pipeline ring is idle by now) and leaves with lanes along the columns: 128-byte coalesced stores,
        // and the bias / ELU' operands are read coalesced in the same pass.
        const int row0 = mt * GT_ROWS + warp * 32;
        if (nst > 0) {
            mbar_wait(acc_bar, 0u);
            tc_fence_after();
        }
        float* sq = reinterpret_cast<float*>(smem) + warp * (32 * 33);
        float* C = g.C + (long long)z * g.split_stride;
        const int n_base = ntile * NT;
        for (int c0 = 0; c0 < NT; c0 += 32) {
            if (n_base + c0 >= g.N) break;                       // warp-uniform
            uint32_t v[32];
            if (nst > 0) {
                tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = 0u;
            }
#pragma unroll
            for (int j = 0; j < 32; ++j) sq[lane * 33 + j] = __uint_as_float(v[j]);
            __syncwarp();
            const int n = n_base + c0 + lane;
            const bool nv = n < g.N;
            const float bias = (g.bias && nv) ? g.bias[n] : 0.0f;
            const int rmax = min(32, g.M - row0);
            for (int rb = 0; rb < 32; rb += 8) {                 // 8 rows per pass: the ELU' operand loads are all in flight
                float y[8];
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    y[u] = (g.elu_out && nv && rb + u < rmax) ? __ldg(g.elu_out + (long long)(row0 + rb + u) * g.ld_elu + n) : 1.0f;
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    if (nv && rb + u < rmax) {
                        float x = sq[(rb + u) * 33 + lane] + bias;
                        if (g.elu) x = gt_elu(x);
                        x *= y[u] > 0.0f ? 1.0f : y[u] + 1.0f;
                        C[(long long)(row0 + rb + u) * g.ldc + n] = x;
                    }
                }
            }
            __syncwarp();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == GT_EPI_WARPS) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)NT) : "memory");
    }
}

// ------------------------------------------------------------------------------------------------ host
static int gt_fail(int code, const char* msg) { pf_set_error(msg); return code; }

int pf_tc_gemm_create(pf_rl_ctx* c) {
    TcGemmState* t = new TcGemmState();
    // operands of every GEMM of a minibatch step: rows up to rows_cap (padded to 256), K up to the widest layer or,
    // for the weight gradients, rows up to the widest layer and K up to rows_cap
    int maxw = 1;
    for (int k = 0; k < 2; ++k)
        for (int l = 0; l <= c->net[k].layers; ++l) maxw = c->net[k].dims[l] > maxw ? c->net[k].dims[l] : maxw;
    const size_t rows_p = ((size_t)c->rows_cap + 255) / 256 * 256 + 256, w_p = ((size_t)maxw + 255) / 256 * 256;
    t->ap_bytes = t->bp_bytes = rows_p * w_p * 2;
    if (cudaMalloc(&t->ap, t->ap_bytes) != cudaSuccess || cudaMalloc(&t->bp, t->bp_bytes) != cudaSuccess) {
        cudaFree(t->ap); delete t;
        return gt_fail(2, "pf_rl_create: cudaMalloc of the packed GEMM operands failed");
    }
    const int smem = GT_STAGES * (GT_ROWS * GT_KSTEP * 2 + 256 * GT_KSTEP * 2) + 256;
    if (cudaFuncSetAttribute(pf_tc_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
        cudaFree(t->ap); cudaFree(t->bp); delete t;
        return gt_fail(2, "pf_rl_create: cannot reserve shared memory for the tensor-core GEMM");
    }
    c->tc_gemm = t;
    return 0;
}

void pf_tc_gemm_destroy(pf_rl_ctx* c) {
    TcGemmState* t = static_cast<TcGemmState*>(c->tc_gemm);
    if (!t) return;
    cudaFree(t->ap); cudaFree(t->bp);
    delete t;
    c->tc_gemm = nullptr;
}

// worth the two packing passes once there is real work; thin extents (6 actions, 1 value) are padded to a tile
int pf_tc_gemm_supported(const pf_rl_ctx* c, int M, int N, int K) {
    return c->tc_gemm && (long long)M * N * K >= (1ll << 20) && (M >= 64 || N >= 64);
}

int pf_tc_gemm(pf_rl_ctx* c, const GemmArgs& d, int splits, cudaStream_t st) {
    TcGemmState* t = static_cast<TcGemmState*>(c->tc_gemm);
    const int NT = d.N > 128 ? 256 : 128;
    // K range of one split, in pipeline stages; the fp32 path's k_per_split is a multiple of 16, here of GT_KSTEP
    const int total_stages = (d.K + GT_KSTEP - 1) / GT_KSTEP;
    int sps = total_stages;
    if (splits > 1) {
        if (d.k_per_split % GT_KSTEP) return gt_fail(1, "pf_tc_gemm: split size must be a multiple of 64");
        sps = d.k_per_split / GT_KSTEP;
    }
    const int kblocks = total_stages * (GT_KSTEP / 8);
    const int mtiles = (d.M + GT_ROWS - 1) / GT_ROWS, ntiles = (d.N + NT - 1) / NT;
    const long long a_slots = (long long)mtiles * kblocks * GT_ROWS, b_slots = (long long)ntiles * kblocks * NT;
    if ((size_t)a_slots * 16 > t->ap_bytes || (size_t)b_slots * 16 > t->bp_bytes) return gt_fail(1, "pf_tc_gemm: operand larger than the packing workspace");
    PF_LAUNCHED(), pf_tc_pack_operand_kernel<<<(unsigned)((a_slots + 255) / 256), 256, 0, st>>>(d.A, d.sam, d.sak, d.a_rows, nullptr, d.M, d.K, GT_ROWS,
                                                                                 kblocks, a_slots, reinterpret_cast<uint4*>(t->ap));
    // B(k, n) = B[krow(k) * sbk + n * sbn]: the packed operand is B^T, rows = n, columns = k
    PF_LAUNCHED(), pf_tc_pack_operand_kernel<<<(unsigned)((b_slots + 255) / 256), 256, 0, st>>>(d.B, d.sbn, d.sbk, nullptr, d.b_rows, d.N, d.K, NT,
                                                                                 kblocks, b_slots, reinterpret_cast<uint4*>(t->bp));
    TcGemmArgs g;
    memset(&g, 0, sizeof(g));
    g.ap = reinterpret_cast<const unsigned char*>(t->ap); g.bp = reinterpret_cast<const unsigned char*>(t->bp);
    g.kblocks = kblocks; g.stages_per_split = sps; g.nt = NT;
    g.C = d.C; g.ldc = d.ldc; g.split_stride = d.split_stride; g.M = d.M; g.N = d.N;
    g.bias = d.bias; g.elu = d.elu; g.elu_out = d.elu_out; g.ld_elu = d.ld_elu;
    const int per_split = sps < total_stages ? sps : total_stages;
    g.stages = per_split < 2 ? 2 : per_split > GT_STAGES ? GT_STAGES : per_split;
    if (NT == 256 && g.stages == 3) g.stages = 2;         // 2 x 48 KB stages: two CTAs per SM; 3 would only fit one
    const int smem = g.stages * (GT_ROWS * GT_KSTEP * 2 + NT * GT_KSTEP * 2) + 256;
    PF_LAUNCHED(), pf_tc_gemm_kernel<<<dim3(ntiles, mtiles, splits), GT_THREADS, smem, st>>>(g);
    return cudaGetLastError() == cudaSuccess ? 0 : gt_fail(2, "pf_tc_gemm: launch failed");
}
