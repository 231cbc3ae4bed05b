// Register-resident path of ActorCritic for SMALL networks (the pointfoot_flat policy: 27 -> 128 -> 64 -> 32 -> 6 / 1):
// every Linear + bias + ELU of both nets, the Normal sample / log-prob of PPO.act and the rollout-storage writes in one
// launch, one WARP per (net, tile of 16 rows), activations never leaving the register file.
//
// Why not the tcgen05 kernel (pf_mlp_tc.cu) here: these nets are 55 kFLOP per row against ~470 B of HBM traffic -- memory
// bound -- and an item of the tcgen05 kernel is a serial chain of four MMA -> TMEM -> epilogue -> shared-memory hand-offs
// (~10k cycles) of which TMEM can hold two or three at a time per SM: 161 us for 262,144 rows against a 20 us HBM bound.
// A warp-level m16n8k16 MMA keeps its accumulator fragment in registers, and that fragment IS the A fragment of the next
// layer after ELU + fp16 packing (rows g / g+8, column pairs 2t / 2t+1 of n-tiles 2s, 2s+1 -> k-step s), so a layer
// boundary costs no memory traffic and no barrier; 16 independent warps per SM hide each other's latencies.  Measured
// mma.sync 16-bit rate on this part: 0.46 MMA/clk/SM = 550 TFLOP/s (tools/micro/mma_sync_rate.cu) -- 28 us for the 3.7 M
// MMAs of 262,144 rows, next to the HBM time.  Larger nets (K or N > 128, i.e. the rough policy) stay on tcgen05.
//
// Weights: fp16, pre-packed by pf_wm_pack_kernel in B-fragment order -- for layer l, k-step s, n-tile j the 32 lanes'
// {b0, b1} = W[8j + g][16s + 2t + {0,1}], W[8j + g][16s + 2t + 8 + {0,1}] (nn.Linear weight [out, in] IS the col-major B) as
// one uint2 per lane -- copied verbatim into shared memory once per CTA (57 KB for both flat nets); biases stay fp32: the
// accumulators start from them.  Inputs: the tile's rows are one contiguous block of global memory, read coalesced, copied to
// the rollout-storage row, and staged as they are in a per-warp buffer the A fragments of layer 0 are built from.
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "pf_rl_internal.h"
#include "pf_rng.cuh"
#include "pf_launch.h"

void pf_set_error(const char* msg);

#define WM_ROWS 16                   // rows per warp task (MMA M)
#define WM_MAX_W 128                 // widest layer (K or N): 8 k-steps / 16 n-tiles of register fragments
#define WM_WARPS 8                   // warps per CTA (two CTAs per SM)
#define WM_MAX_SMEM (100 * 1024)

struct WmNet {
    int layers;
    int k_true0;                               // true input width
    int kp[PF_MAX_MLP_LAYERS];                 // padded K of layer l (multiple of 16)
    int np[PF_MAX_MLP_LAYERS];                 // padded N of layer l (multiple of 8; hidden layers: multiple of 16)
    int n_out;                                 // true output width (<= 8)
    uint32_t w_off[PF_MAX_MLP_LAYERS];         // byte offsets inside the net's blob
    uint32_t b_off[PF_MAX_MLP_LAYERS];
    uint32_t blob_bytes;
};
struct WmState { WmNet net[2]; unsigned char* blob[2]; };

enum { WM_OUT_PLAIN = 0, WM_OUT_SAMPLE = 1, WM_OUT_VALUE = 2 };
struct WmNetArgs {
    WmNet net;
    const unsigned char* blob;
    uint32_t smem_off;                         // where the blob sits in the CTA's shared memory
    const float* x; long long ldx;
    float* out; float* copy_to;
    int mode;
    int flat3;                                 // the net is [<= 32, 128, 64, 32, <= 8]: layers called directly (set by wm_launch)
};
struct WmArgs {
    WmNetArgs n[2];
    int nnets, rows, tiles, stage_stride;      // stage_stride: floats of one warp's staging area
    int weights_early;                         // the packed weights are NOT the product of the kernel right in front: stage them before the dependency wait
    uint32_t stage_off;                        // per-warp staging area offset in shared memory
    uint32_t seed, policy_step; int env_id_offset, num_actions;
    const float* std;
    pf_rl_storage S; int has_storage; long long storage_row0;
    float* actions_out; float* values_out; float* logp_out;
};

// Operands are fp16 (inputs are clipped observations, |x| <= 100, and pre-activations of a trained policy are O(10): inside
// fp16's range with three more mantissa bits than bf16; conversions saturate instead of overflowing), accumulation is fp32.
__device__ __forceinline__ uint32_t wm_pack(float lo, float hi) {            // {lo, hi} -> f16x2 (lo in the low half)
    uint32_t r;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
// ELU on a packed pair, branch-free: max(z, exp2(min(z, 0) * log2 e) - 1).  Packed conversion / min / mul / sub / max, two
// MUFU.EX2.F16: 8 issue slots for two elements, and the result IS the packed A-fragment register of the next layer.
__device__ __forceinline__ uint32_t wm_elu2(float z0, float z1) {
    const uint32_t z = wm_pack(z0, z1), zero = 0u, log2e = 0x3DC53DC5u /* 1.4424 | 1.4424 */, one = 0x3C003C00u;
    uint32_t m, e, r;
    asm("min.f16x2 %0, %1, %2;" : "=r"(m) : "r"(z), "r"(zero));
    asm("mul.rn.f16x2 %0, %1, %2;" : "=r"(m) : "r"(m), "r"(log2e));
    asm("ex2.approx.f16x2 %0, %1;" : "=r"(e) : "r"(m));
    asm("sub.rn.f16x2 %0, %1, %2;" : "=r"(e) : "r"(e), "r"(one));
    asm("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(z), "r"(e));
    return r;
}
__device__ __forceinline__ void wm_mma(float (&c)[4], const uint32_t (&a)[4], uint2 b) {
    asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y));
}

// One layer with COMPILE-TIME shape: KS k-steps, NT n-tiles -> straight-line code (immediate shared-memory offsets, no
// predicates, no register shuffling).  Accumulators start from the bias; unless it is the output layer the ELU'd, packed
// accumulators overwrite af[] as the next layer's A fragments.  (A first version with run-time bounds on a fully unrolled
// 8 x 16 nest issued 5,200 instructions per task, 2,800 with uniform branches: profiles/r2m_wm_flat262k_v*.txt.)
template <int KS, int NT>
__device__ __forceinline__ void wm_layer(uint32_t (&af)[WM_MAX_W / 16][4], const uint2* __restrict__ wl, const float* __restrict__ bl,
                                         int t, bool last, float* mrow0, float* mrow8) {
    float acc[NT][4];
#pragma unroll
    for (int j = 0; j < NT; ++j) {
        const float2 b = *reinterpret_cast<const float2*>(bl + 8 * j + 2 * t);
        acc[j][0] = acc[j][2] = b.x; acc[j][1] = acc[j][3] = b.y;
    }
#pragma unroll
    for (int s = 0; s < KS; ++s)
#pragma unroll
        for (int j = 0; j < NT; ++j) wm_mma(acc[j], af[s], wl[(s * NT + j) * 32]);
    if (NT >= 2 && !last) {
#pragma unroll
        for (int s = 0; s < NT / 2; ++s) {
            af[s][0] = wm_elu2(acc[2 * s][0], acc[2 * s][1]);
            af[s][1] = wm_elu2(acc[2 * s][2], acc[2 * s][3]);
            af[s][2] = wm_elu2(acc[2 * s + 1][0], acc[2 * s + 1][1]);
            af[s][3] = wm_elu2(acc[2 * s + 1][2], acc[2 * s + 1][3]);
        }
    } else {                                          // output layer: rows g, g + 8; columns 2t, 2t + 1 of n-tile 0
        *reinterpret_cast<float2*>(mrow0) = make_float2(acc[0][0], acc[0][1]);
        *reinterpret_cast<float2*>(mrow8) = make_float2(acc[0][2], acc[0][3]);
    }
}

// Normal sample + log-prob + rollout-storage row of one env over actions [j0, j1); returns the partial log-prob.  Same
// statement as pf_mlp_tc.cu::tc_sample_row with the hardware transcendentals (lg2 / cos / rsqrt units, relative error
// ~1e-6: three orders below this path's tolerance) -- the precise libm versions were 14 % of the kernel's instructions.
__device__ __forceinline__ float wm_sample_part(const WmArgs& a, long long er, const float* mean, int j0, int j1) {
    const int A = a.num_actions;
    const uint32_t ek = pf_rng_env_key(pf_rng_key(a.seed, a.policy_step, PF_RNG_POLICY), (uint32_t)(a.env_id_offset + (int)er));
    const long long srow = a.storage_row0 + er;
    float logp = 0.0f;
#pragma unroll 1
    for (int j = j0; j < j1; ++j) {
        const float m = mean[j], sd = a.std[j];
        const float u1 = ((float)(pf_rng_u32(ek, 2u * (uint32_t)j) >> 8) + 1.0f) * (1.0f / 16777216.0f);
        const float u2 = (float)(pf_rng_u32(ek, 2u * (uint32_t)j + 1u) >> 8) * (1.0f / 16777216.0f);
        const float nrm = sqrtf(-2.0f * __logf(u1)) * __cosf(6.283185307179586f * u2);
        const float act = m + sd * nrm;
        const float d = act - m;
        logp += -__fdividef(d * d, 2.0f * (sd * sd)) - __logf(sd) - 0.9189385332046727f;
        if (a.actions_out) a.actions_out[er * A + j] = act;
        if (a.has_storage) {
            a.S.actions[srow * A + j] = act;
            a.S.mu[srow * A + j] = m;
            a.S.sigma[srow * A + j] = sd;
        }
    }
    return logp;
}

// Input rows of one task -> the warp's fp32 staging tile [16][K0] (dense) + the rollout-storage copy.  The tile is one
// contiguous block of global memory: 128-bit coalesced reads, 4 in flight per lane, stored to shared memory and to the
// storage row as they are; rows beyond the batch are zeros.
__device__ __forceinline__ void wm_stage_rows(const float* __restrict__ x, long long ldx, float* __restrict__ copy_to, int K0,
                                              int row0, int nrows, float* stage, int lane) {
    const int nel = nrows * K0, nall = WM_ROWS * K0;
    const float* __restrict__ xt = x + (long long)row0 * ldx;
    float* __restrict__ ct = copy_to ? copy_to + (long long)row0 * K0 : nullptr;
    const bool vec = ldx == K0 && ((reinterpret_cast<uintptr_t>(xt) | (ct ? reinterpret_cast<uintptr_t>(ct) : 0)) & 15) == 0;
    if (vec) {
        for (int base = 0; base < nall; base += 512) {
            float4 q[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = base + (u * 32 + lane) * 4;
                q[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (i + 3 < nel) q[u] = __ldg(reinterpret_cast<const float4*>(xt + i));
                else if (i < nel) { q[u].x = xt[i]; if (i + 1 < nel) q[u].y = xt[i + 1]; if (i + 2 < nel) q[u].z = xt[i + 2]; }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = base + (u * 32 + lane) * 4;
                if (i < nall) {
                    *reinterpret_cast<float4*>(stage + i) = q[u];
                    if (ct) {
                        if (i + 3 < nel) *reinterpret_cast<float4*>(ct + i) = q[u];
                        else { if (i < nel) ct[i] = q[u].x; if (i + 1 < nel) ct[i + 1] = q[u].y; if (i + 2 < nel) ct[i + 2] = q[u].z; }
                    }
                }
            }
        }
    } else {
        const uint32_t rcp = (4194304u + (uint32_t)K0 - 1u) / (uint32_t)K0;       // i / K0 by multiply-shift (exact for i < 2048)
        for (int base = 0; base < nall; base += 256) {
            float v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int i = base + u * 32 + lane;
                v[u] = 0.0f;
                if (i < nel) { const int r = (int)(((uint32_t)i * rcp) >> 22); v[u] = __ldg(xt + (long long)r * ldx + (i - r * K0)); }
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int i = base + u * 32 + lane;
                if (i < nall) { stage[i] = v[u]; if (ct && i < nel) ct[i] = v[u]; }
            }
        }
    }
}

#define WM_CASE(KS, NT) case (KS) * 32 + (NT): wm_layer<KS, NT>(af, wl, bl, t, last, mrow0, mrow8); break;
#define WM_CASES(KS) WM_CASE(KS, 1) WM_CASE(KS, 2) WM_CASE(KS, 4) WM_CASE(KS, 8) WM_CASE(KS, 16)

__global__ void __launch_bounds__(WM_WARPS * 32, 2) pf_mlp_wm_kernel(const __grid_constant__ WmArgs a) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    pf_grid_dep_launch();
    if (!a.weights_early) pf_grid_dep_wait();            // the weights may have been packed by the kernel in front of this one
    for (int i = 0; i < a.nnets; ++i) {
        const uint4* src = reinterpret_cast<const uint4*>(a.n[i].blob);
        uint4* dst = reinterpret_cast<uint4*>(smem + a.n[i].smem_off);
        const int n16 = (int)(a.n[i].net.blob_bytes >> 4);
        for (int k0 = 0; k0 < n16; k0 += 4 * WM_WARPS * 32) {          // 4 independent 128-bit loads in flight per thread
            uint4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) { const int k = k0 + u * WM_WARPS * 32 + tid; if (k < n16) v[u] = __ldg(src + k); }
#pragma unroll
            for (int u = 0; u < 4; ++u) { const int k = k0 + u * WM_WARPS * 32 + tid; if (k < n16) dst[k] = v[u]; }
        }
    }
    if (a.weights_early) pf_grid_dep_wait();             // ... or not (host-side bookkeeping): then they were staged under the predecessor's tail
    __syncthreads();
    float* stage = reinterpret_cast<float*>(smem + a.stage_off) + (size_t)warp * a.stage_stride;      // [16][K0] fp32 + [16][8] outputs
    const int ntasks = a.nnets * a.tiles;
    // task = warp * gridDim.x + blockIdx.x: a small batch spreads over every SM before any CTA gets a second warp's worth
    for (int task = warp * gridDim.x + blockIdx.x; task < ntasks; task += gridDim.x * WM_WARPS) {
        const int ni = task >= a.tiles ? 1 : 0, tile = task - ni * a.tiles;          // (at most two nets per launch)
        const WmNetArgs& N = a.n[ni];
        const WmNet& P = N.net;
        const int row0 = tile * WM_ROWS, nrows = min(WM_ROWS, a.rows - row0), K0 = P.k_true0, K0p = P.kp[0];
        float* mtile = stage + WM_ROWS * K0;
        wm_stage_rows(N.x, N.ldx, N.copy_to, K0, row0, nrows, stage, lane);
        __syncwarp();
        // ---- A fragments of layer 0: fp32 pairs from the staged tile -> f16x2 (columns >= K0 are zeros) ----
        uint32_t af[WM_MAX_W / 16][4];
#pragma unroll
        for (int s = 0; s < WM_MAX_W / 16; ++s) {
            if (s * 16 >= K0p) break;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int c = 16 * s + 8 * h + 2 * t;
                const float* r0 = stage + g * K0 + c;
                const float* r8 = r0 + 8 * K0;
                const float x00 = c < K0 ? r0[0] : 0.0f, x01 = c + 1 < K0 ? r0[1] : 0.0f;
                const float x80 = c < K0 ? r8[0] : 0.0f, x81 = c + 1 < K0 ? r8[1] : 0.0f;
                af[s][2 * h] = wm_pack(x00, x01);
                af[s][2 * h + 1] = wm_pack(x80, x81);
            }
        }
        const unsigned char* blob = smem + N.smem_off;
        float* mrow0 = mtile + g * 8 + 2 * t;
        float* mrow8 = mrow0 + 64;
        if (N.flat3) {
            // the pointfoot_flat policy [<= 32, 128, 64, 32, <= 8]: the four layers called directly, no per-layer dispatch
            auto W = [&](int l) { return reinterpret_cast<const uint2*>(blob + P.w_off[l]) + lane; };
            auto Bv = [&](int l) { return reinterpret_cast<const float*>(blob + P.b_off[l]); };
            wm_layer<2, 16>(af, W(0), Bv(0), t, false, mrow0, mrow8);
            wm_layer<8, 8>(af, W(1), Bv(1), t, false, mrow0, mrow8);
            wm_layer<4, 4>(af, W(2), Bv(2), t, false, mrow0, mrow8);
            wm_layer<2, 1>(af, W(3), Bv(3), t, true, mrow0, mrow8);
        } else {
#pragma unroll 1
            for (int l = 0; l < P.layers; ++l) {
                const uint2* wl = reinterpret_cast<const uint2*>(blob + P.w_off[l]) + lane;
                const float* bl = reinterpret_cast<const float*>(blob + P.b_off[l]);
                const bool last = l + 1 == P.layers;
                switch ((P.kp[l] >> 4) * 32 + (P.np[l] >> 3)) {          // warp-uniform: (k-steps, n-tiles) of this layer
                    WM_CASES(1) WM_CASES(2) WM_CASES(4) WM_CASES(8)
                    default: break;                                       // (pf_wm_create admits power-of-two widths only)
                }
            }
        }
        __syncwarp();
        // ---- outputs ----
        if (N.mode == WM_OUT_PLAIN) {
            const int no = P.n_out;
            for (int i = lane; i < nrows * no; i += 32) { const int r = i / no, c = i - r * no; N.out[(long long)(row0 + r) * no + c] = mtile[r * 8 + c]; }
        } else if (N.mode == WM_OUT_VALUE) {
            if (lane < nrows) {
                const long long er = (long long)row0 + lane;
                const float val = mtile[lane * 8];
                if (a.values_out) a.values_out[er] = val;
                if (a.has_storage) a.S.values[a.storage_row0 + er] = val;
            }
        } else {                                          // two lanes per row, half of the actions each
            const int r = lane & 15, hA = (a.num_actions + 1) >> 1, j0 = (lane >> 4) * hA, j1 = min(a.num_actions, j0 + hA);
            const long long er = (long long)row0 + r;
            float lp = 0.0f;
            if (r < nrows) lp = wm_sample_part(a, er, mtile + r * 8, j0, j1);
            lp += __shfl_xor_sync(0xffffffffu, lp, 16);
            if (lane < nrows) {
                if (a.logp_out) a.logp_out[er] = lp;
                if (a.has_storage) a.S.actions_log_prob[a.storage_row0 + er] = lp;
            }
        }
        __syncwarp();
    }
}

// fp32 nn.Linear weight [out, in] + bias -> the blob: B fragments in consumption order (k-step major, n-tile minor), then
// the fp32 bias (zero padded).  One thread per packed 32-bit word.
struct WmPackArgs { WmNet net[2]; unsigned char* blob[2]; long long w_off[2][PF_MAX_MLP_LAYERS], b_off[2][PF_MAX_MLP_LAYERS]; int dims[2][PF_MAX_MLP_LAYERS + 1]; };

__global__ void __launch_bounds__(256) pf_wm_pack_kernel(const WmPackArgs a, const float* __restrict__ params) {
    pf_grid_dep_launch();
    pf_grid_dep_wait();
    const int k = blockIdx.y;
    const WmNet& P = a.net[k];
    for (int l = 0; l < P.layers; ++l) {
        const int ks = P.kp[l] >> 4, nt = P.np[l] >> 3, K = a.dims[k][l], Nn = a.dims[k][l + 1];
        const float* W = params + a.w_off[k][l];
        uint32_t* dst = reinterpret_cast<uint32_t*>(a.blob[k] + P.w_off[l]);
        const int words = ks * nt * 64;
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < words; i += gridDim.x * blockDim.x) {
            const int half = i & 1, lane = (i >> 1) & 31, sj = i >> 6, s = sj / nt, j = sj - s * nt;
            const int n = 8 * j + (lane >> 2), kk = 16 * s + 2 * (lane & 3) + 8 * half;
            const float lo = (n < Nn && kk < K) ? W[(long long)n * K + kk] : 0.0f;
            const float hi = (n < Nn && kk + 1 < K) ? W[(long long)n * K + kk + 1] : 0.0f;
            dst[i] = wm_pack(lo, hi);
        }
        float* bd = reinterpret_cast<float*>(a.blob[k] + P.b_off[l]);
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P.np[l]; i += gridDim.x * blockDim.x)
            bd[i] = i < Nn ? params[a.b_off[k][l] + i] : 0.0f;
    }
}

// ================================================================================================ host
static int wm_fail(int code, const char* msg) { pf_set_error(msg); return code; }
static inline int rup(int x, int m) { return (x + m - 1) / m * m; }

static bool wm_plan(const PfNet& n, WmNet& P) {
    memset(&P, 0, sizeof(P));
    if (n.layers < 1 || n.layers > PF_MAX_MLP_LAYERS) return false;
    P.layers = n.layers; P.k_true0 = n.dims[0]; P.n_out = n.dims[n.layers];
    if (n.dims[0] > WM_MAX_W || P.n_out > 8) return false;
    uint32_t off = 0;
    for (int l = 0; l < n.layers; ++l) {
        const bool last = l == n.layers - 1;
        const int h = n.dims[l + 1];
        if (!last && !(h == 16 || h == 32 || h == 64 || h == 128)) return false;                  // hidden widths: the instantiated layer shapes
        P.kp[l] = l == 0 ? (n.dims[0] <= 16 ? 16 : n.dims[0] <= 32 ? 32 : n.dims[0] <= 64 ? 64 : 128) : n.dims[l];
        P.np[l] = last ? 8 : n.dims[l + 1];
        P.w_off[l] = off; off += (uint32_t)(P.kp[l] / 16) * (P.np[l] / 8) * 256;
        P.b_off[l] = off; off += (uint32_t)rup(P.np[l] * 4, 16);
    }
    P.blob_bytes = rup(off, 16);
    return true;
}

static int wm_stage_stride(int k0_max) { return rup(WM_ROWS * k0_max + WM_ROWS * 8, 4); }       // floats per warp: input tile + output tile
static size_t wm_smem(const WmNet* nets, int n, int k0_max) {
    size_t b = 0;
    for (int i = 0; i < n; ++i) b += nets[i].blob_bytes;
    return b + (size_t)WM_WARPS * wm_stage_stride(k0_max) * 4;
}

int pf_wm_create(pf_rl_ctx* c) {
    c->wm = nullptr;
    { const char* e = getenv("PF_RL_WM"); if (e && atoi(e) == 0) return 0; }
    WmState* t = new WmState();
    memset(t, 0, sizeof(*t));
    if (!wm_plan(c->net[0], t->net[0]) || !wm_plan(c->net[1], t->net[1]) || c->net[1].dims[c->net[1].layers] != 1 ||
        wm_smem(t->net, 2, t->net[0].k_true0 > t->net[1].k_true0 ? t->net[0].k_true0 : t->net[1].k_true0) > WM_MAX_SMEM) {
        delete t;
        return 0;                                  // not a small-net policy: the tcgen05 path takes it
    }
    for (int k = 0; k < 2; ++k)
        if (cudaMalloc(&t->blob[k], t->net[k].blob_bytes) != cudaSuccess) {
            cudaFree(t->blob[0]); delete t;
            return wm_fail(2, "pf_wm_create: cudaMalloc failed");
        }
    if (cudaFuncSetAttribute(pf_mlp_wm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WM_MAX_SMEM) != cudaSuccess) {
        cudaFree(t->blob[0]); cudaFree(t->blob[1]); delete t;
        return wm_fail(2, "pf_wm_create: cudaFuncSetAttribute failed");
    }
    c->wm = t;
    return 0;
}

void pf_wm_destroy(pf_rl_ctx* c) {
    WmState* t = static_cast<WmState*>(c->wm);
    if (!t) return;
    cudaFree(t->blob[0]); cudaFree(t->blob[1]);
    delete t;
    c->wm = nullptr;
}

int pf_wm_pack_weights(pf_rl_ctx* c, const float* params, cudaStream_t st) {
    WmState* t = static_cast<WmState*>(c->wm);
    WmPackArgs a;
    memset(&a, 0, sizeof(a));
    for (int k = 0; k < 2; ++k) {
        a.net[k] = t->net[k]; a.blob[k] = t->blob[k];
        for (int l = 0; l < PF_MAX_MLP_LAYERS; ++l) { a.w_off[k][l] = c->net[k].w_off[l]; a.b_off[k][l] = c->net[k].b_off[l]; }
        for (int l = 0; l <= PF_MAX_MLP_LAYERS; ++l) a.dims[k][l] = c->net[k].dims[l];
    }
    return pf_launch_pdl(pf_wm_pack_kernel, dim3(8, 2), dim3(256), 0, st, a, params) == cudaSuccess
               ? 0 : wm_fail(2, "pf_wm_pack_weights: launch failed");
}

static int wm_launch(pf_rl_ctx* c, WmArgs& a, cudaStream_t st) {
    if (a.rows <= 0) return 0;
    a.tiles = (a.rows + WM_ROWS - 1) / WM_ROWS;
    int k0 = 0;
    uint32_t off = 0;
    WmNet nets[2];
    for (int i = 0; i < a.nnets; ++i) {
        a.n[i].smem_off = off; off += a.n[i].net.blob_bytes;
        { const WmNet& w = a.n[i].net;
          a.n[i].flat3 = w.layers == 4 && w.kp[0] == 32 && w.np[0] == 128 && w.kp[1] == 128 && w.np[1] == 64 && w.kp[2] == 64 && w.np[2] == 32 &&
                         w.kp[3] == 32 && w.np[3] == 8 && getenv("PF_WM_NO_FLAT3") == nullptr; }
        if (a.n[i].net.k_true0 > k0) k0 = a.n[i].net.k_true0;
        nets[i] = a.n[i].net;
    }
    a.stage_off = off;
    a.stage_stride = wm_stage_stride(k0);
    const size_t smem = wm_smem(nets, a.nnets, k0);
    const int ntasks = a.nnets * a.tiles;
    // CTAs: full ones (8 tasks, one per warp) rather than one task per CTA -- every CTA stages both nets' weights (57 KB), and
    // at 4,096 rows 296 CTAs doing that for 512 tasks cost more than the SMs they fill (PF_WM_SPREAD=1: the old spreading)
    static const int spread = [] { const char* e = getenv("PF_WM_SPREAD"); return e ? atoi(e) : 0; }();
    int grid = spread ? ntasks : (ntasks + WM_WARPS - 1) / WM_WARPS;
    if (grid > 2 * c->num_sms) grid = 2 * c->num_sms;
    return pf_launch_pdl(pf_mlp_wm_kernel, dim3(grid), dim3(WM_WARPS * 32), smem, st, a) == cudaSuccess
               ? 0 : wm_fail(2, "pf_mlp_wm_kernel: launch failed");
}

int pf_wm_mlp_forward(pf_rl_ctx* c, int net, const float* x, long long ldx, int rows, float* out, float* copy_to, cudaStream_t st) {
    WmState* t = static_cast<WmState*>(c->wm);
    WmArgs a;
    memset(&a, 0, sizeof(a));
    a.n[0].net = t->net[net]; a.n[0].blob = t->blob[net];
    a.n[0].x = x; a.n[0].ldx = ldx; a.n[0].out = out; a.n[0].copy_to = copy_to; a.n[0].mode = WM_OUT_PLAIN;
    a.nnets = 1; a.rows = rows;
    a.weights_early = pf_rl_weights_early(c, st);
    return wm_launch(c, a, st);
}

// PPO.act in one launch: critic tasks first, then the actor tasks with the sample / log-prob tail
int pf_wm_act(pf_rl_ctx* c, const float* params, const float* obs, const float* critic_obs, const pf_rl_storage* storage,
              int t_row, int64_t policy_step, float* actions_out, float* values_out, float* logp_out, cudaStream_t st) {
    WmState* t = static_cast<WmState*>(c->wm);
    const pf_rl_params& P = c->P;
    WmArgs a;
    memset(&a, 0, sizeof(a));
    const long long row0 = (long long)t_row * P.num_envs;
    for (int i = 0; i < 2; ++i) {
        const int k = 1 - i;
        a.n[i].net = t->net[k]; a.n[i].blob = t->blob[k];
        a.n[i].x = k == 0 ? obs : critic_obs;
        a.n[i].ldx = c->net[k].dims[0];
        a.n[i].mode = k == 0 ? WM_OUT_SAMPLE : WM_OUT_VALUE;
        if (storage) {
            if (k == 0) a.n[i].copy_to = storage->observations + row0 * P.num_obs;
            else if (storage->privileged_observations) a.n[i].copy_to = storage->privileged_observations + row0 * P.num_critic_obs;
        }
    }
    a.nnets = 2; a.rows = P.num_envs;
    a.weights_early = pf_rl_weights_early(c, st);
    a.seed = (uint32_t)P.seed; a.policy_step = (uint32_t)policy_step; a.env_id_offset = (int)P.env_id_offset; a.num_actions = P.num_actions;
    a.std = params + c->std_off;
    if (storage) { a.S = *storage; a.has_storage = 1; }
    a.storage_row0 = row0;
    a.actions_out = actions_out; a.values_out = values_out; a.logp_out = logp_out;
    return wm_launch(c, a, st);
}
