// Tensor-core path of PPO.update (rsl_rl algorithms/ppo.py update(): forward over a minibatch, loss.backward()'s dgrad and
// wgrad), PF_RL_BF16_TC precision: bf16 operands, fp32 accumulation in TMEM, tcgen05.mma.
//
// ONE layout for every matrix the update touches -- activations X_l, output gradients dY_l, weights W_l -- so that no
// operand is ever re-packed or transposed between the forward, dgrad and wgrad GEMMs:
//
//     "row-tile interleave", bf16:  element (r, c) of a [rows, cols] matrix with tile height TH lives at byte
//         ((r / TH) * (cols_p / 8) + c / 8) * TH * 16  +  (r % TH) * 16  +  (c % 8) * 2
//
// i.e. tiles of TH rows; inside a tile, blocks of 8 columns ("col blocks") of TH x 16 bytes each.  A col block is a stack
// of UMMA core matrices (8 rows x 16 bytes, contiguous), which the tensor core can read in BOTH orientations:
//   * K-major  (contraction along the columns):  forward A = X, forward B = W, dgrad A = dY
//   * MN-major (contraction along the rows):     dgrad B = W (dX = dY W contracts over W's rows), wgrad A = dY and
//                                                wgrad B = X (dW = dY^T X contracts over the batch rows)
// the orientation is two bits of the instruction descriptor (a_major / b_major) and which of the descriptor's two strides
// (LBO / SBO) walks the col blocks.  Every pipeline stage is one or a few cp.async.bulk copies straight from the tensors.
//
// Bias without an epilogue operand: every activation tensor carries two columns of ones behind its real columns and every
// weight tensor the bias split in (hi, lo) bf16 in the same two columns, so the forward GEMM adds the bias (to ~2^-17) and
// the wgrad GEMM produces the bias gradient as one more column of dW.
//
// Epilogues write the next GEMM's operand directly: forward = ELU -> bf16 -> packed Y_l; dgrad = acc * ELU'(Y_{l-1}) ->
// bf16 -> packed dY_{l-1}; wgrad = fp32 split-K partials, reduced in a fixed order (deterministic) by one kernel for all
// layers.  Actor and critic go through the same launches (a launch carries up to four problems); a minibatch step of
// the [512,256,128] nets is 15 launches, all chained by programmatic dependent launch.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdlib.h>
#include <string.h>
#include "pf_rl_internal.h"
#include "pf_tc_ptx.cuh"
#include "pf_launch.h"

void pf_set_error(const char* msg);

#define TT_TH 128                      // rows per tile of every activation / gradient tensor (= UMMA M)
#define TT_KS 128                      // K elements per pipeline stage = the tile height: a stage of an MN-major operand (wgrad's
                                       // dY^T and X: batch rows are K) is then ONE contiguous bulk copy of whole col blocks; with 64 it was
                                       // up to 48 copies of 1 KB per stage and 0.7 us per 48 KB (profiles/r2p_train_trace_*.txt)
#define TT_A_BYTES 32768               // staged A tile: 128 x 128 bf16 in either orientation
#define TT_B_BYTES 65536               // staged B tile: up to 256 x 128 bf16
#define TT_STAGE_BYTES (TT_A_BYTES + TT_B_BYTES)
#define TT_STAGES 2
#define TT_MAXPROB 4
#define TT_EPI_WARPS 16                // 4 TMEM lane quarters x 4 column groups: the epilogue (ELU + pack + store), not the MMAs, bounds a tile
#define TT_EPI_H (TT_EPI_WARPS / 4)
#define TT_MMA_WARP TT_EPI_WARPS
#define TT_LOAD_WARP (TT_EPI_WARPS + 1)
#define TT_THREADS ((TT_EPI_WARPS + 2) * 32)
#define TT_SMEM (TT_STAGES * TT_STAGE_BYTES + 256)
enum { TT_FWD = 0, TT_DGRAD = 1, TT_WGRAD = 2 };

static inline int pad16i(int x) { return (x + 15) & ~15; }

// one GEMM of a launch.  Its output is cut into tiles of 128 rows x up to 256 columns (x split-K slices for wgrad); the
// launch's tiles are numbered problem after problem and dealt round-robin to the persistent CTAs.
struct TtProb {
    const unsigned char* a; const unsigned char* b;
    int a_cbs, b_cbs, b_th;            // col blocks per tile of the A / B source tensors; tile height of the B source
    int mode;
    int m_tiles, n_tiles;              // output tiles: m_tiles x n_tiles (x splits)
    int n_cbs_total;                   // dgrad / wgrad: N extent in col blocks (a tile takes up to 32 of them)
    int k_stages, k_last;              // K stages of TT_KS elements; elements in the last one (multiple of 16)
    int splits, stages_per_split;      // wgrad: split-K over the batch
    int tile0, tiles;
    unsigned char* out_pk; int out_cbs; int n_valid; int ones; int elu;     // packed bf16 output (fwd hidden, dgrad)
    const unsigned char* y_pk; int y_cbs;                                   // dgrad: saved activation for ELU'
    float* out_f32; int ld_f32; int rows_valid;                             // fwd, last layer: plain fp32 [rows, n_valid]
    float* part;                       // wgrad: partials [split][m_tile][n_tile][128][256] fp32
};
struct TtArgs { TtProb p[TT_MAXPROB]; int nprob; int total_tiles; };

__device__ __forceinline__ uint32_t tt_idesc(int n, int a_mn, int b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
           ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}
__device__ __forceinline__ float tt_elu(float x) {
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(fminf(x, 0.0f) * 1.4426950408889634f));
    return fmaxf(x, e - 1.0f);
}
__device__ __forceinline__ uint32_t tt_pack2(float a, float b) {
    __nv_bfloat162 p = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&p);
}

// what a CTA needs to know about one tile
struct TtTile { int pi, mt, jn, z, s_beg, nst, w; };
__device__ __forceinline__ TtTile tt_tile(const TtArgs& g, int t) {
    TtTile T;
    T.pi = 0;
#pragma unroll
    for (int i = 1; i < TT_MAXPROB; ++i) if (i < g.nprob && t >= g.p[i].tile0) T.pi = i;
    const TtProb& P = g.p[T.pi];
    int idx = t - P.tile0;
    T.mt = idx % P.m_tiles; idx /= P.m_tiles;
    T.jn = idx % P.n_tiles;
    T.z = idx / P.n_tiles;
    T.s_beg = T.z * P.stages_per_split;
    T.nst = min(P.k_stages, T.s_beg + P.stages_per_split) - T.s_beg;
    if (P.mode == TT_FWD) T.w = P.b_th;                                       // a tile of the weight tensor's rows
    else { const int cb = P.n_cbs_total - T.jn * 32; T.w = (cb < 32 ? cb : 32) * 8; }
    return T;
}

// Persistent: one CTA per SM walks its tiles.  Loader warp -> 4-stage shared-memory ring -> MMA warp -> two 256-column TMEM
// accumulators -> 8 epilogue warps, so that the epilogue of tile i runs under the loads and MMAs of tile i + 1.
__global__ void __launch_bounds__(TT_THREADS, 1) pf_train_gemm_kernel(const __grid_constant__ TtArgs g) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    pf_grid_dep_launch();
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + TT_STAGES * TT_STAGE_BYTES);
    const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + TT_STAGES);
    const uint32_t accf0 = smem_u32(bars + 2 * TT_STAGES), acce0 = smem_u32(bars + 2 * TT_STAGES + 2);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * TT_STAGES + 4);

    if (warp == TT_LOAD_WARP && lane == 0) {
        for (int s = 0; s < TT_STAGES; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(accf0 + 8 * b, 1); mbar_init(acce0 + 8 * b, TT_EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == TT_MMA_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pf_timeline_raw_if(blockIdx.x == 0 && tid == 0, 0x340);
    pf_grid_dep_wait();              // everything above ran under the predecessor's tail; its outputs are visible from here on
    pf_timeline_raw_if(blockIdx.x == 0 && tid == 0, 0x341);

    if (warp == TT_LOAD_WARP) {
        // ================================================================ loader: the whole warp issues a stage's copies
        uint32_t it = 0;                                             // stages issued so far by this CTA (ring position)
        for (int t = (int)blockIdx.x; t < g.total_tiles; t += (int)gridDim.x) {
            const TtTile T = tt_tile(g, t);
            const TtProb& P = g.p[T.pi];
            const int mode = P.mode;
            pf_timeline_raw_if(blockIdx.x == 0 && lane == 0, 0x300 + P.mode);          // loader: tile begins
            for (int s = 0; s < T.nst; ++s, ++it) {
                const uint32_t slot = it % TT_STAGES, gs = T.s_beg + s;
                const int kthis = (int)gs == P.k_stages - 1 ? P.k_last : TT_KS;
                if (it >= TT_STAGES) mbar_wait(empty0 + 8 * slot, ((it / TT_STAGES) - 1u) & 1u);
                const uint32_t fb = full0 + 8 * slot;
                const uint32_t a_dst = smem_u32(smem + slot * TT_STAGE_BYTES), b_dst = a_dst + TT_A_BYTES;
                const int a_ncb = mode == TT_WGRAD ? min(16, P.a_cbs - T.mt * 16) : kthis >> 3;
                const uint32_t a_bytes = (uint32_t)a_ncb * TT_TH * 16u;      // whole col blocks of the 128-row tile, either orientation
                const uint32_t b_bytes = mode == TT_FWD ? (uint32_t)(kthis >> 3) * (uint32_t)P.b_th * 16u
                                       : mode == TT_WGRAD ? (uint32_t)(T.w >> 3) * TT_TH * 16u : (uint32_t)(T.w >> 3) * (uint32_t)kthis * 16u;
                if (lane == 0) mbar_expect_tx(fb, a_bytes + b_bytes);
                __syncwarp();
                // ---- A ----
                if (mode != TT_WGRAD) {           // K-major rows of X / dY: the stage's col blocks are contiguous
                    if (lane == 0) bulk_g2s(a_dst, P.a + ((size_t)T.mt * P.a_cbs + (size_t)gs * (TT_KS / 8)) * (TT_TH * 16), a_bytes, fb);
                } else {                          // MN-major: row tile gs (a stage IS a row tile), the M tile's col blocks: contiguous
                    if (lane == 0) bulk_g2s(a_dst, P.a + ((size_t)gs * P.a_cbs + (size_t)T.mt * 16) * (TT_TH * 16), a_bytes, fb);
                }
                // ---- B ----
                if (mode == TT_FWD) {             // K-major rows of W: tile jn, the stage's col blocks
                    if (lane == 0) bulk_g2s(b_dst, P.b + ((size_t)T.jn * P.b_cbs + (size_t)gs * (TT_KS / 8)) * ((size_t)P.b_th * 16), b_bytes, fb);
                } else if (mode == TT_WGRAD) {    // MN-major batch rows of X: row tile gs, the N tile's col blocks: contiguous
                    if (lane == 0) bulk_g2s(b_dst, P.b + ((size_t)gs * P.b_cbs + (size_t)T.jn * 32) * (TT_TH * 16), b_bytes, fb);
                } else {
                    // dgrad, MN-major: rows [gs*128, +kthis) of W (tile height b_th), col blocks of tile jn; a col block is TT_KS rows apart in the stage
                    const int th = P.b_th;
                    const int row = gs * TT_KS, rt = row / th, r0 = row - rt * th;
                    for (int cb = lane; cb < (T.w >> 3); cb += 32)
                        bulk_g2s(b_dst + cb * (TT_KS * 16), P.b + ((size_t)rt * P.b_cbs + (size_t)T.jn * 32 + cb) * ((size_t)th * 16) + r0 * 16,
                                 (uint32_t)kthis * 16u, fb);
                }
            }
        }
    } else if (warp == TT_MMA_WARP) {
        // ================================================================ MMA issue
        uint32_t it = 0, j = 0;
        for (int t = (int)blockIdx.x; t < g.total_tiles; t += (int)gridDim.x, ++j) {
            const TtTile T = tt_tile(g, t);
            const TtProb& P = g.p[T.pi];
            const bool a_mn = P.mode == TT_WGRAD, b_mn = P.mode != TT_FWD;
            const uint32_t idesc = tt_idesc(T.w, a_mn, b_mn);
            // descriptor strides (bytes): K-major = (LBO: next 8 K columns, SBO: next 8 rows); MN-major = (LBO: next 8 K rows, SBO: next 8 MN)
            const uint32_t a_lbo = a_mn ? 128u : TT_TH * 16u, a_sbo = a_mn ? TT_KS * 16u : 128u;
            const uint32_t b_lbo = b_mn ? 128u : (uint32_t)P.b_th * 16u, b_sbo = b_mn ? TT_KS * 16u : 128u;
            const uint64_t a_step = a_mn ? 16u : (2u * TT_TH * 16u) >> 4;            // K = 16 further, in 16-byte units
            const uint64_t b_step = b_mn ? 16u : (2u * (uint32_t)P.b_th * 16u) >> 4;
            const uint32_t acc = j & 1u, td = tmem_base + acc * 256u;
            pf_timeline_raw_if(blockIdx.x == 0 && lane == 0, 0x310);                     // MMA: tile begins (waits for its accumulator)
            if (j >= 2) mbar_wait(acce0 + 8 * acc, ((j >> 1) - 1u) & 1u);            // the epilogue has drained this accumulator
            tc_fence_after();
            pf_timeline_raw_if(blockIdx.x == 0 && lane == 0, 0x311);                     // MMA: accumulator free
            for (int s = 0; s < T.nst; ++s, ++it) {
                const uint32_t slot = it % TT_STAGES;
                const int nk16 = (T.s_beg + s == P.k_stages - 1 ? P.k_last : TT_KS) >> 4;
                mbar_wait(full0 + 8 * slot, (it / TT_STAGES) & 1u);
                tc_fence_after();
                pf_timeline_raw_if(blockIdx.x == 0 && lane == 0 && (s == 0 || s == T.nst - 1), s == 0 ? 0x312 : 0x313);   // first / last stage landed
                if (elect_one()) {
                    const uint32_t base = smem_u32(smem + slot * TT_STAGE_BYTES);
                    const uint64_t ad = make_desc(base, a_lbo, a_sbo), bd = make_desc(base + TT_A_BYTES, b_lbo, b_sbo);
                    if (s == 0) umma_bf16(td, ad, bd, idesc, 0u); else umma_bf16_acc(td, ad, bd, idesc);
                    for (int k = 1; k < nk16; ++k) umma_bf16_acc(td, ad + k * a_step, bd + k * b_step, idesc);
                    umma_commit(empty0 + 8 * slot);
                    if (s == T.nst - 1) umma_commit(accf0 + 8 * acc);
                }
                __syncwarp();
            }
        }
    } else {
        // ================================================================ epilogue: warp (q, h) owns TMEM lanes [32q, 32q + 32) = tile rows,
        // column half h of the accumulator
        const int q = warp & 3, h = warp >> 2, row = q * 32 + lane;
        uint32_t j = 0;
        for (int t = (int)blockIdx.x; t < g.total_tiles; t += (int)gridDim.x, ++j) {
            const TtTile T = tt_tile(g, t);
            const TtProb& P = g.p[T.pi];
            const int mode = P.mode, mt = T.mt, w = T.w;
            // the problem's fields used per element, ONCE per tile into registers: `P` is a run-time index into the grid constant, and
            // the compiler re-issued a constant load per use -- ~100 dependent LDCs per group of 32 columns made the epilogue
            // 10-13 us per tile against 2-3 us of loads + MMAs (tools/train_trace.py, profiles/r2p_train_trace_rough4096_before.txt)
            const int n_valid = P.n_valid, out_cbs = P.out_cbs, y_cbs = P.y_cbs, n_tiles = P.n_tiles, m_tiles = P.m_tiles;
            const int rows_valid = P.rows_valid, ld_f32 = P.ld_f32, b_th = P.b_th;
            const bool do_elu = P.elu != 0, do_ones = P.ones != 0;
            unsigned char* const out_pk = P.out_pk;
            const unsigned char* const y_pk = P.y_pk;
            float* const out_f32 = P.out_f32;
            float* const part = P.part;
            const uint32_t acc = j & 1u;
            pf_timeline_raw_if(blockIdx.x == 0 && tid == 0, 0x320);                      // epilogue: waits for the accumulator
            mbar_wait(accf0 + 8 * acc, (j >> 1) & 1u);
            tc_fence_after();
            pf_timeline_raw_if(blockIdx.x == 0 && tid == 0, 0x321);                      // epilogue: accumulator complete
            const uint32_t trow = tmem_base + acc * 256u + ((uint32_t)(q * 32) << 16);
            const int n0 = T.jn * (mode == TT_FWD ? b_th : 256);
            // forward into a packed tensor: the last N tile also writes the columns behind the real ones (ones, zero padding)
            const int cols_end = (mode == TT_FWD && out_pk && T.jn == n_tiles - 1) ? out_cbs * 8 - n0 : w;
            // this warp's column range: groups of 32 columns, alternating between the two halves
            const int ngroups = (cols_end + 31) >> 5;
            bool released = false;
            for (int gi = h; gi < ngroups || !released; gi += TT_EPI_H) {
                const int c0 = gi * 32;
                uint32_t v[32];
                const bool real = gi < ngroups && c0 < w;
                if (real) {
                    if (w - c0 >= 32) tmem_ld32(trow + c0, v);
                    else { tmem_ld16(trow + c0, v);
#pragma unroll
                        for (int jj = 16; jj < 32; ++jj) v[jj] = 0u; }
                    tmem_ld_wait();
                } else {
#pragma unroll
                    for (int jj = 0; jj < 32; ++jj) v[jj] = 0u;
                }
                // the accumulator is free once this warp's last real group sits in registers
                if (!released && (gi + TT_EPI_H) * 32 >= w) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(acce0 + 8 * acc);
                    released = true;
                }
                if (gi >= ngroups) break;
                if (mode == TT_WGRAD) {
                    float4* dst = reinterpret_cast<float4*>(part + ((((size_t)T.z * m_tiles + mt) * n_tiles + T.jn) * 128 + row) * 256 + c0);
                    const int nq = min(32, w - c0) >> 2;
#pragma unroll
                    for (int qq = 0; qq < 8; ++qq)
                        if (qq < nq) dst[qq] = make_float4(__uint_as_float(v[4 * qq]), __uint_as_float(v[4 * qq + 1]), __uint_as_float(v[4 * qq + 2]), __uint_as_float(v[4 * qq + 3]));
                } else if (out_f32) {       // forward, last layer: plain fp32 rows (mu / value)
                    const int grow = mt * TT_TH + row;
                    if (grow < rows_valid) {
#pragma unroll
                        for (int jj = 0; jj < 32; ++jj)
                            if (n0 + c0 + jj < n_valid) out_f32[(size_t)grow * ld_f32 + n0 + c0 + jj] = __uint_as_float(v[jj]);
                    }
                } else {
                    const int ncb = min(4, (cols_end - c0 + 7) >> 3);
                    uint4 yv[4];                  // dgrad: the group's four saved-activation units, all in flight before the first use
                    if (mode == TT_DGRAD) {
#pragma unroll
                        for (int qq = 0; qq < 4; ++qq) {
                            const int cb = (n0 + c0) / 8 + qq;
                            yv[qq] = make_uint4(0u, 0u, 0u, 0u);
                            if (qq < ncb && cb < out_cbs)
                                yv[qq] = __ldg(reinterpret_cast<const uint4*>(y_pk + ((size_t)mt * y_cbs + cb) * (TT_TH * 16) + row * 16));
                        }
                    }
#pragma unroll
                    for (int qq = 0; qq < 4; ++qq) {
                        if (qq >= ncb) continue;
                        const int cb = (n0 + c0) / 8 + qq;
                        if (cb >= out_cbs) continue;
                        float x[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) x[i] = __uint_as_float(v[8 * qq + i]);
                        if (mode == TT_FWD) {
                            if (cb * 8 + 8 <= n_valid) {          // (warp-uniform) a block of real columns: no per-element tests
                                if (do_elu) {
#pragma unroll
                                    for (int i = 0; i < 8; ++i) x[i] = tt_elu(x[i]);
                                }
                            } else {
#pragma unroll
                                for (int i = 0; i < 8; ++i) {
                                    const int n = cb * 8 + i;
                                    x[i] = n < n_valid ? (do_elu ? tt_elu(x[i]) : x[i]) : ((do_ones && n < n_valid + 2) ? 1.0f : 0.0f);
                                }
                            }
                        } else {              // dgrad: times ELU'(y) recovered from the saved activation: y > 0 ? 1 : y + 1
                            const uint4 yy = yv[qq];
                            const uint32_t yw[4] = {yy.x, yy.y, yy.z, yy.w};
                            const bool full = cb * 8 + 8 <= n_valid;          // (warp-uniform)
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const float y = __uint_as_float((i & 1) ? (yw[i >> 1] & 0xffff0000u) : (yw[i >> 1] << 16));
                                const float d = x[i] * (y > 0.0f ? 1.0f : y + 1.0f);
                                x[i] = (full || cb * 8 + i < n_valid) ? d : 0.0f;
                            }
                        }
                        *reinterpret_cast<uint4*>(out_pk + ((size_t)mt * out_cbs + cb) * (TT_TH * 16) + row * 16) =
                            make_uint4(tt_pack2(x[0], x[1]), tt_pack2(x[2], x[3]), tt_pack2(x[4], x[5]), tt_pack2(x[6], x[7]));
                    }
                }
            }
        }
    }
    pf_timeline_raw_if(blockIdx.x == 0 && (tid == 0 || tid == TT_MMA_WARP * 32 || tid == TT_LOAD_WARP * 32), 0x330 + (tid == 0 ? 0 : tid == TT_MMA_WARP * 32 ? 1 : 2));
    tc_fence_before();
    __syncthreads();
    if (warp == TT_MMA_WARP) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

extern "C" int pf_tt_debug_timeline(unsigned long long* device_log) {       // diagnostic builds only (tools/train_trace.py)
    return pf_timeline_install(device_log) == cudaSuccess ? 0 : 1;
}

// ------------------------------------------------------------------------------------------------ small kernels
// packed offset of the 16-byte unit (row r, col block cb) of a tensor with tile height th and cbs col blocks per tile
__host__ __device__ inline size_t tt_unit(int r, int cb, int th, int cbs) {
    return ((size_t)(r / th) * cbs + cb) * ((size_t)th * 16) + (size_t)(r % th) * 16;
}

// fp32 nn.Linear weight [out, in] + bias -> packed bf16 [rows_p, in_p]: columns [0, in) = W, column in / in + 1 = the bias
// split in (hi, lo) bf16 (met by the two ones-columns of the activations), everything else zero.  All layers of both
// nets in one launch.
struct TtLayerTab {
    int n;
    struct E { long long w_off, b_off; int out, in, rows_p, cbs, th; unsigned char* pk; long long unit0; } e[2 * PF_MAX_MLP_LAYERS];
    long long units;
};
__global__ void __launch_bounds__(256) pf_tt_pack_weights_kernel(const __grid_constant__ TtLayerTab t, const float* __restrict__ params) {
    pf_grid_dep_launch();
    pf_grid_dep_wait();
    const long long u = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= t.units) return;
    int li = 0;
    for (int i = 1; i < t.n; ++i) if (u >= t.e[i].unit0) li = i;
    const TtLayerTab::E& L = t.e[li];
    const long long lu = u - L.unit0;
    const int r = (int)(lu / L.cbs), cb = (int)(lu % L.cbs);
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int c = cb * 8 + i;
        float x = 0.0f;
        if (r < L.out) {
            if (c < L.in) x = params[L.w_off + (long long)r * L.in + c];
            else if (c == L.in || c == L.in + 1) {
                const float b = params[L.b_off + r];
                const float hi = __bfloat162float(__float2bfloat16_rn(b));
                x = c == L.in ? hi : b - hi;
            }
        }
        v[i] = x;
    }
    *reinterpret_cast<uint4*>(L.pk + tt_unit(r, cb, L.th, L.cbs)) =
        make_uint4(tt_pack2(v[0], v[1]), tt_pack2(v[2], v[3]), tt_pack2(v[4], v[5]), tt_pack2(v[6], v[7]));
}

// mini_batch_generator's gather (rollout_storage.py): rows perm[b] of the flattened storage -> packed bf16 X0 of the
// minibatch, with the two ones-columns behind the observation; rows past the minibatch are zero.  Both nets in one launch.
struct TtX0Args { const float* src[2]; int width[2]; unsigned char* dst[2]; int cbs[2]; int rows_p, rows; const int* perm; };
__global__ void __launch_bounds__(256) pf_tt_pack_x0_kernel(const TtX0Args a) {
    pf_grid_dep_launch();
    pf_grid_dep_wait();
    const int k = blockIdx.y;
    const long long u = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int cbs = a.cbs[k];
    if (u >= (long long)a.rows_p * cbs) return;
    const int r = (int)(u / cbs), cb = (int)(u % cbs), W = a.width[k];
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = 0.0f;
    if (r < a.rows) {
        const float* row = a.src[k] + (long long)a.perm[r] * W;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int c = cb * 8 + i;
            v[i] = c < W ? __ldg(row + c) : (c < W + 2 ? 1.0f : 0.0f);
        }
    }
    *reinterpret_cast<uint4*>(a.dst[k] + tt_unit(r, cb, TT_TH, cbs)) =
        make_uint4(tt_pack2(v[0], v[1]), tt_pack2(v[2], v[3]), tt_pack2(v[4], v[5]), tt_pack2(v[6], v[7]));
}

// loss gradients d loss / d mu [rows, A] and d loss / d value [rows] (fp32, written by pf_ppo_loss_kernel) -> the packed bf16
// dY of the two last layers; rows past the minibatch are zero
struct TtDoutArgs { const float* src[2]; int width[2]; unsigned char* dst[2]; int cbs[2]; int rows_p, rows; };
__global__ void __launch_bounds__(256) pf_tt_pack_dout_kernel(const TtDoutArgs a) {
    pf_grid_dep_launch();
    pf_grid_dep_wait();
    const int k = blockIdx.y;
    const long long u = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int cbs = a.cbs[k];
    if (u >= (long long)a.rows_p * cbs) return;
    const int r = (int)(u / cbs), cb = (int)(u % cbs), W = a.width[k];
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int c = cb * 8 + i;
        v[i] = (r < a.rows && c < W) ? a.src[k][(long long)r * W + c] : 0.0f;
    }
    *reinterpret_cast<uint4*>(a.dst[k] + tt_unit(r, cb, TT_TH, cbs)) =
        make_uint4(tt_pack2(v[0], v[1]), tt_pack2(v[2], v[3]), tt_pack2(v[4], v[5]), tt_pack2(v[6], v[7]));
}

// split-K partials of every layer's wgrad -> the flat gradient buffer (weight [out, in] and bias [out]: column `in` of dW),
// summed over the splits in a fixed order: run-to-run deterministic.
struct TtReduceTab {
    int n;
    struct E { long long w_off, b_off, elem0; int out, in, splits, m_tiles, n_tiles; const float* part; } e[2 * PF_MAX_MLP_LAYERS];
    long long elems;
};
__global__ void __launch_bounds__(256) pf_tt_reduce_kernel(const __grid_constant__ TtReduceTab t, float* __restrict__ grads) {
    pf_grid_dep_launch();
    pf_grid_dep_wait();
    const long long u = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= t.elems) return;
    int li = 0;
    for (int i = 1; i < t.n; ++i) if (u >= t.e[i].elem0) li = i;
    const TtReduceTab::E& L = t.e[li];
    const long long lu = u - L.elem0;
    const int o = (int)(lu / (L.in + 1)), i = (int)(lu % (L.in + 1));          // i == in: the bias column
    const size_t tile = ((size_t)(o >> 7) * L.n_tiles + (i >> 8)) * 32768 + (size_t)(o & 127) * 256 + (i & 255);
    const size_t stride = (size_t)L.m_tiles * L.n_tiles * 32768;
    float s = 0.0f;
    for (int z = 0; z < L.splits; ++z) s += L.part[(size_t)z * stride + tile];
    if (i < L.in) grads[L.w_off + (long long)o * L.in + i] = s;
    else grads[L.b_off + o] = s;
}

// ------------------------------------------------------------------------------------------------ host
struct TtTensor { unsigned char* p; int rows_p, cbs, th; size_t bytes; };
struct TtNet {
    int L;
    int in[PF_MAX_MLP_LAYERS], out[PF_MAX_MLP_LAYERS];            // true widths of layer l
    TtTensor W[PF_MAX_MLP_LAYERS];        // [pad(out), pad16(in + 2)]: weights + bias columns
    TtTensor X[PF_MAX_MLP_LAYERS];        // input of layer l: X[0] = gathered observations, X[l] = ELU output of layer l - 1 (+ ones)
    TtTensor dY[PF_MAX_MLP_LAYERS];       // gradient w.r.t. the pre-activation output of layer l: [rows_p, pad16(out)]
    float* part[PF_MAX_MLP_LAYERS]; int splits[PF_MAX_MLP_LAYERS], sps[PF_MAX_MLP_LAYERS], mt_w[PF_MAX_MLP_LAYERS], nt_w[PF_MAX_MLP_LAYERS];
};
struct TtState { TtNet net[2]; int rows_p, row_tiles, rows; int num_sms; };

static int tt_fail(int code, const char* msg) { pf_set_error(msg); return code; }

static int tt_alloc(TtTensor& t, int rows, int cols_p, int th) {
    t.th = th; t.cbs = cols_p / 8; t.rows_p = (rows + th - 1) / th * th;
    t.bytes = (size_t)t.rows_p * cols_p * 2;
    if (cudaMalloc(&t.p, t.bytes) != cudaSuccess) return 1;
    return cudaMemset(t.p, 0, t.bytes) != cudaSuccess;
}

void pf_tt_destroy(pf_rl_ctx* c) {
    TtState* t = static_cast<TtState*>(c->tc_gemm);
    if (!t) return;
    for (int k = 0; k < 2; ++k)
        for (int l = 0; l < PF_MAX_MLP_LAYERS; ++l) {
            cudaFree(t->net[k].W[l].p); cudaFree(t->net[k].X[l].p); cudaFree(t->net[k].dY[l].p); cudaFree(t->net[k].part[l]);
        }
    delete t;
    c->tc_gemm = nullptr;
}

int pf_tt_create(pf_rl_ctx* c) {
    TtState* t = new TtState();
    memset(t, 0, sizeof(*t));
    c->tc_gemm = t;                          // (pf_rl_destroy -> pf_tt_destroy frees whatever was allocated if we fail below)
    t->rows = c->mb_size;
    t->row_tiles = (c->mb_size + TT_TH - 1) / TT_TH;
    if (t->row_tiles < 1) t->row_tiles = 1;          // (a storage smaller than num_mini_batches: empty minibatches, as upstream)
    t->rows_p = t->row_tiles * TT_TH;
    t->num_sms = c->num_sms;
    const int k_stages_batch = t->row_tiles;             // wgrad: K = the batch, one stage per 128-row tile
    // tiles of all wgrad problems of one layer index (both nets): decides the split-K factor of that launch
    int tiles_at[PF_MAX_MLP_LAYERS] = {0};
    for (int k = 0; k < 2; ++k) {
        TtNet& n = t->net[k];
        n.L = c->net[k].layers;
        for (int l = 0; l < n.L; ++l) {
            n.in[l] = c->net[k].dims[l]; n.out[l] = c->net[k].dims[l + 1];
            const int in_p = pad16i(n.in[l] + 2), out_p = pad16i(n.out[l]);
            // weights: tile height = the N of one forward MMA (<= 128: the whole layer; wider: 256-row tiles)
            if (tt_alloc(n.W[l], out_p, in_p, out_p <= 128 ? out_p : 256) || tt_alloc(n.X[l], t->rows_p, in_p, TT_TH) ||
                tt_alloc(n.dY[l], t->rows_p, out_p, TT_TH))
                return tt_fail(2, "pf_rl_create: cudaMalloc of the packed training tensors failed");
            n.mt_w[l] = (out_p + 127) / 128; n.nt_w[l] = (in_p + 255) / 256;
            tiles_at[l] += n.mt_w[l] * n.nt_w[l];
        }
    }
    for (int k = 0; k < 2; ++k) {
        TtNet& n = t->net[k];
        for (int l = 0; l < n.L; ++l) {
            int splits = (t->num_sms + tiles_at[l] - 1) / tiles_at[l];
            if (splits > k_stages_batch / 2) splits = k_stages_batch / 2;       // at least two stages per split
            if (splits < 1) splits = 1;
            const int sps = (k_stages_batch + splits - 1) / splits;
            n.sps[l] = sps; n.splits[l] = (k_stages_batch + sps - 1) / sps;
            if (cudaMalloc(&n.part[l], sizeof(float) * (size_t)n.splits[l] * n.mt_w[l] * n.nt_w[l] * 32768) != cudaSuccess)
                return tt_fail(2, "pf_rl_create: cudaMalloc of the wgrad partials failed");
        }
    }
    if (cudaFuncSetAttribute(pf_train_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TT_SMEM) != cudaSuccess)
        return tt_fail(2, "pf_rl_create: cannot reserve shared memory for the tensor-core training GEMM");
    return 0;
}

// packed bf16 copies of the weights after every change of the fp32 parameters (init, load, Adam step)
int pf_tt_pack_weights(pf_rl_ctx* c, const float* params, cudaStream_t st) {
    TtState* t = static_cast<TtState*>(c->tc_gemm);
    TtLayerTab tab;
    memset(&tab, 0, sizeof(tab));
    long long units = 0;
    for (int k = 0; k < 2; ++k)
        for (int l = 0; l < t->net[k].L; ++l) {
            TtLayerTab::E& e = tab.e[tab.n++];
            const TtTensor& W = t->net[k].W[l];
            e.w_off = c->net[k].w_off[l]; e.b_off = c->net[k].b_off[l];
            e.out = t->net[k].out[l]; e.in = t->net[k].in[l]; e.rows_p = W.rows_p; e.cbs = W.cbs; e.th = W.th; e.pk = W.p;
            e.unit0 = units; units += (long long)W.rows_p * W.cbs;
        }
    tab.units = units;
    return pf_launch_pdl(pf_tt_pack_weights_kernel, dim3((unsigned)((units + 255) / 256)), dim3(256), 0, st, tab, params) == cudaSuccess
               ? 0 : tt_fail(2, "pf_tt_pack_weights: launch failed");
}

static int tt_launch(TtState* t, TtArgs& a, cudaStream_t st) {
    int tiles = 0;
    for (int i = 0; i < a.nprob; ++i) { a.p[i].tile0 = tiles; tiles += a.p[i].tiles; }
    if (tiles == 0) return 0;
    a.total_tiles = tiles;
    const int grid = tiles < t->num_sms ? tiles : t->num_sms;         // persistent: one CTA per SM
    return pf_launch_pdl(pf_train_gemm_kernel, dim3(grid), dim3(TT_THREADS), (size_t)TT_SMEM, st, a) == cudaSuccess
               ? 0 : tt_fail(2, "pf_train_gemm_kernel: launch failed");
}

static void tt_k(TtProb& p, int k_elems) { p.k_stages = (k_elems + TT_KS - 1) / TT_KS; p.k_last = k_elems - (p.k_stages - 1) * TT_KS; p.splits = 1; p.stages_per_split = p.k_stages; }

// forward of layer l of net n: Y = ELU(X W^T + b) into X[l + 1], or the plain fp32 output of the last layer
static TtProb tt_fwd(TtState* t, TtNet& n, int l, float* out_f32) {
    TtProb p; memset(&p, 0, sizeof(p));
    const TtTensor &X = n.X[l], &W = n.W[l];
    p.mode = TT_FWD; p.a = X.p; p.a_cbs = X.cbs; p.b = W.p; p.b_cbs = W.cbs; p.b_th = W.th;
    p.m_tiles = t->row_tiles; p.n_tiles = W.rows_p / W.th;
    tt_k(p, X.cbs * 8);
    p.tiles = p.m_tiles * p.n_tiles;
    p.n_valid = n.out[l];
    if (l < n.L - 1) { p.out_pk = n.X[l + 1].p; p.out_cbs = n.X[l + 1].cbs; p.ones = 1; p.elu = 1; }
    else { p.out_f32 = out_f32; p.ld_f32 = n.out[l]; p.rows_valid = t->rows; }
    return p;
}
// dgrad through layer l (l >= 1): dY[l - 1] = (dY[l] W_l) * ELU'(X[l])
static TtProb tt_dgrad(TtState* t, TtNet& n, int l) {
    TtProb p; memset(&p, 0, sizeof(p));
    const TtTensor &G = n.dY[l], &W = n.W[l], &D = n.dY[l - 1];
    p.mode = TT_DGRAD; p.a = G.p; p.a_cbs = G.cbs; p.b = W.p; p.b_cbs = W.cbs; p.b_th = W.th;
    p.m_tiles = t->row_tiles;
    p.n_cbs_total = D.cbs;                       // the real inputs of layer l (the bias columns of W are not propagated)
    p.n_tiles = (D.cbs + 31) / 32;
    tt_k(p, G.cbs * 8);
    p.tiles = p.m_tiles * p.n_tiles;
    p.out_pk = D.p; p.out_cbs = D.cbs; p.n_valid = n.in[l];
    p.y_pk = n.X[l].p; p.y_cbs = n.X[l].cbs;
    return p;
}
// wgrad of layer l: dW = dY[l]^T X[l] (the ones columns of X make column `in` of dW the bias gradient)
static TtProb tt_wgrad(TtState* t, TtNet& n, int l) {
    TtProb p; memset(&p, 0, sizeof(p));
    const TtTensor &G = n.dY[l], &X = n.X[l];
    p.mode = TT_WGRAD; p.a = G.p; p.a_cbs = G.cbs; p.b = X.p; p.b_cbs = X.cbs; p.b_th = TT_TH;
    p.m_tiles = n.mt_w[l]; p.n_tiles = n.nt_w[l]; p.n_cbs_total = X.cbs;
    tt_k(p, t->rows_p);
    p.splits = n.splits[l]; p.stages_per_split = n.sps[l];
    p.tiles = p.m_tiles * p.n_tiles * p.splits;
    p.part = n.part[l];
    return p;
}

// One minibatch of PPO.update up to the optimizer step (ppo.py update(), the body of the generator loop): gather, forward of
// both nets, loss + its gradients, backward.  `idx` = this minibatch's slice of the permutation.
int pf_tt_minibatch(pf_rl_ctx* c, const pf_rl_storage* storage, const pf_rl_learner* L, const int* idx, cudaStream_t st,
                    int (*loss)(pf_rl_ctx*, const pf_rl_storage*, const pf_rl_learner*, const int*, cudaStream_t)) {
    TtState* t = static_cast<TtState*>(c->tc_gemm);
    const pf_rl_params& P = c->P;
    // ---- gather + pack the observations of the minibatch ----
    {
        TtX0Args a;
        a.src[0] = storage->observations; a.width[0] = P.num_obs;
        a.src[1] = storage->privileged_observations ? storage->privileged_observations : storage->observations; a.width[1] = P.num_critic_obs;
        for (int k = 0; k < 2; ++k) { a.dst[k] = t->net[k].X[0].p; a.cbs[k] = t->net[k].X[0].cbs; }
        a.rows_p = t->rows_p; a.rows = t->rows; a.perm = idx;
        const int cbs = a.cbs[0] > a.cbs[1] ? a.cbs[0] : a.cbs[1];
        if (pf_launch_pdl(pf_tt_pack_x0_kernel, dim3((unsigned)(((long long)t->rows_p * cbs + 255) / 256), 2), dim3(256), 0, st, a) != cudaSuccess)
            return tt_fail(2, "pf_tt_pack_x0_kernel: launch failed");
    }
    // ---- forward, layer by layer, both nets per launch ----
    const int Lmax = t->net[0].L > t->net[1].L ? t->net[0].L : t->net[1].L;
    for (int l = 0; l < Lmax; ++l) {
        TtArgs a; memset(&a, 0, sizeof(a));
        for (int k = 1; k >= 0; --k)              // critic first: the wider input
            if (l < t->net[k].L) a.p[a.nprob++] = tt_fwd(t, t->net[k], l, k == 0 ? c->mu : c->val);
        if (int rc = tt_launch(t, a, st)) return rc;
    }
    // ---- loss + d loss / d (mu, value, std) (fp32, pf_rl.cu) -> packed bf16 gradients of the two output layers ----
    if (int rc = loss(c, storage, L, idx, st)) return rc;
    {
        TtDoutArgs a;
        a.src[0] = c->dmu; a.width[0] = P.num_actions; a.src[1] = c->dval; a.width[1] = 1;
        for (int k = 0; k < 2; ++k) { const TtTensor& D = t->net[k].dY[t->net[k].L - 1]; a.dst[k] = D.p; a.cbs[k] = D.cbs; }
        a.rows_p = t->rows_p; a.rows = t->rows;
        const int cbs = a.cbs[0] > a.cbs[1] ? a.cbs[0] : a.cbs[1];
        if (pf_launch_pdl(pf_tt_pack_dout_kernel, dim3((unsigned)(((long long)t->rows_p * cbs + 255) / 256), 2), dim3(256), 0, st, a) != cudaSuccess)
            return tt_fail(2, "pf_tt_pack_dout_kernel: launch failed");
    }
    // ---- backward: per layer index one launch with the wgrads (long split-K CTAs first) and the dgrads of both nets ----
    for (int l = Lmax - 1; l >= 0; --l) {
        TtArgs a; memset(&a, 0, sizeof(a));
        for (int k = 1; k >= 0; --k) if (l < t->net[k].L) a.p[a.nprob++] = tt_wgrad(t, t->net[k], l);
        for (int k = 1; k >= 0; --k) if (l >= 1 && l < t->net[k].L) a.p[a.nprob++] = tt_dgrad(t, t->net[k], l);
        if (int rc = tt_launch(t, a, st)) return rc;
    }
    // ---- split-K partials -> flat gradient buffer ----
    TtReduceTab tab; memset(&tab, 0, sizeof(tab));
    long long elems = 0;
    for (int k = 0; k < 2; ++k)
        for (int l = 0; l < t->net[k].L; ++l) {
            TtReduceTab::E& e = tab.e[tab.n++];
            TtNet& n = t->net[k];
            e.w_off = c->net[k].w_off[l]; e.b_off = c->net[k].b_off[l]; e.out = n.out[l]; e.in = n.in[l];
            e.splits = n.splits[l]; e.m_tiles = n.mt_w[l]; e.n_tiles = n.nt_w[l]; e.part = n.part[l];
            e.elem0 = elems; elems += (long long)n.out[l] * (n.in[l] + 1);
        }
    tab.elems = elems;
    return pf_launch_pdl(pf_tt_reduce_kernel, dim3((unsigned)((elems + 255) / 256)), dim3(256), 0, st, tab, L->grads) == cudaSuccess
               ? 0 : tt_fail(2, "pf_tt_reduce_kernel: launch failed");
}
