// Tensor-core path of ActorCritic (rsl_rl modules/actor_critic.py): every Linear + bias + ELU of BOTH networks, the
// Normal sample / log-prob of PPO.act and the rollout-storage writes fused into ONE persistent sm_100a kernel built
// on tcgen05.mma with accumulators in TMEM.
//
//   * a work item is (net, tile of 128 rows): UMMA M = 128, cta_group::1, fp16 operands, fp32 accumulate; one CTA
//     per SM walks the items round-robin, so weights of the next item stream in under the tail of the current one;
//   * activations never leave the SM and are never held whole: the output of a layer is produced in chunks of
//     <= 128 columns (TMEM -> registers -> ELU -> fp16 -> shared memory, in the K-major core-matrix layout an A
//     descriptor reads) and every chunk is consumed at once as a K-slice of the NEXT layer, whose accumulator
//     (<= 256 columns) sits in the other half of TMEM.  Two 32 KB chunk buffers + the staged input rows are all the
//     activation storage there is, which leaves ~110 KB for the weight ring;
//   * weights are pre-packed (pf_tc_pack_weights, after every optimizer step) in exactly the order the MMAs consume
//     them, <= 8 KB tiles, so a tile is ONE contiguous bulk async copy (cp.async.bulk / TMA engine, completion by
//     mbarrier transaction bytes) issued by a dedicated loader thread that runs up to 14 tiles ahead;
//   * the bias add is one extra K = 16 MMA per accumulator against a constant "ones" operand (bias split in
//     hi + lo fp16), so the epilogue is only ELU + pack;
//   * warp roles: warps 0-7 run the epilogues (4 TMEM lane quarters x 2 column halves), warps 8-15 stage the input
//     rows of the NEXT item (fp32 -> fp16 operand layout + the rollout-storage copy) as soon as the current item's
//     first layer has consumed its rows, warp 16 owns TMEM and issues the MMAs, warp 17 issues the bulk copies.
//     They only meet at mbarriers; the schedule (which group waits for / signals which barrier, with which phase)
//     is a small static PROGRAM built on the host.
//
// Shared-memory operand layout (no swizzle, "interleave"): 8 rows x 16 bytes core matrices, stored contiguously
// (128 B); for an operand with R rows: element (r, k) of fp16 at  (k/8) * (R*16) + r*16 + (k%8)*2  bytes.
// UMMA descriptor: leading byte offset = R*16 (next 8 K-columns), stride byte offset = 128 (next 8 rows).
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "pf_rl_internal.h"
#include "pf_rng.cuh"

void pf_set_error(const char* msg);

#define TC_ROWS 128                          // rows per work item (UMMA M)
#define TC_CHUNK 128                         // columns of one activation chunk
#define TC_ACHUNK_BYTES (TC_ROWS * TC_CHUNK * 2)
#define TC_X0_LBO (TC_ROWS * 16 + 16)        // staged input: byte distance between 8-column blocks, padded (see tc_stage_item)
#define TC_SLOT_BYTES 32768                  // weight ring slot = largest tile (N x K x 2 B: 256x64, 128x128 ...), one CTA per SM
#define TC_SLOT_BYTES_COMPACT 16384          // ... and in the compact variant (two CTAs per SM)
#define TC_MAX_STAGES 16
#define TC_X0_KSTEP ((2 * TC_X0_LBO) / 16)    // K = 16 further along the staged input, in descriptor units
// Warp roles are template parameters of the kernel: kEpi epilogue warps (8 = 4 TMEM lane quarters x 2 column halves; the ELU is
// MUFU-bound, more do not help; 4 in the compact variant), kStg staging warps (input rows of the NEXT item, under the tail of
// the current one), then the MMA warp and the loader warp.
// Warp ids: epilogue warps 0 .. kEpi-1 (a warp reads the TMEM lane quarter `id % 4`, so they cover every scheduler); the MMA warp is
// kEpi and the loader kEpi + 4 -- both on scheduler 0 --, the staging warps are the following ids that are NOT on scheduler 0, so the
// issuing warp shares its scheduler with two epilogue warps and the loader only.
constexpr int tc_last_stg_offset(int kStg) { int d = 0, n = 0; while (n < kStg) { ++d; if (d % 4) ++n; } return d; }
#define TC_THREADS_OF(kEpi, kStg) (((kEpi) + tc_last_stg_offset(kStg) + 1) * 32)
#define TC_MAX_GROUPS 24
#define TC_MAX_EPIS 16
#define TC_TMEM_COLS 512
#define TC_TRACE_CAP 512
#define TC_TRACE_ROLES 4

// program barriers (besides the ring's full / empty pairs).  A program names the ones it uses; what arrives on a barrier
// (one tcgen05.commit, every epilogue warp, every staging warp) is the program's bar_kind[].
enum { BAR_ACC0 = 0, BAR_ACC1 = 1, BAR_ACC2 = 2,      // accumulator complete: TMEM cols [0,128) or [0,256) / [128,256) / [256,512)
       BAR_AFREE0 = 3, BAR_AFREE1 = 4,               // MMAs that read activation chunk buffer b have completed
       BAR_E0 = 5, BAR_E1 = 6,                       // epilogue wrote chunk buffer b (and drained its accumulator)
       BAR_XFULL = 7,                                // input rows of the item are staged
       BAR_XFREE = 8,                                // the MMAs that read the staged rows have completed
       // two-phase programs only (build_program_pipe)
       BAR_ACC3 = 9,                                 // last layer's accumulator complete
       BAR_AFREE2 = 10, BAR_E2 = 11,                 // chunk buffer 2 (the second phase's)
       BAR_ODONE = 12,                               // the final epilogue has read its accumulator
       BAR_XFULL1 = 13, BAR_XFREE1 = 14,             // second staged-row buffer (items with odd index); = BAR_XFULL / XFREE + TC_XBAR_STEP
       TC_NBARS = 15 };
#define TC_XBAR_STEP 6
enum { TC_BK_COMMIT = 0, TC_BK_EPI = 1, TC_BK_STG = 2 };

// One accumulation step: D[128 x nt] (+)= A[128 x kc] * W^T, A = staged input rows or an activation chunk.
struct TcGroup {
    uint32_t blob_off;                 // first tile of the group inside the net's packed blob
    uint16_t nt, kc, kt;               // UMMA N, K extent of the group, K per weight tile
    uint16_t tmem_col;                 // accumulator column
    uint16_t n_base, k_base;           // origin inside the layer's weight matrix (packing)
    int8_t layer;
    int8_t a_src;                      // 0 staged input, 1 + b activation chunk buffer b
    int8_t bias;                       // starts the accumulation with the bias tile
    int8_t wait_bar, wait_k;           // barrier (+ completion index inside the item) to wait for before the first MMA
    int8_t commit_bar, commit2_bar;    // barriers the completion of the group's MMAs arrives on (-1: none)
    int8_t phase;                      // two-phase programs: 0 = first half of the item, 1 = second half, run one iteration later
    int8_t prev_bar, prev_k;           // also wait for completion prev_k of this barrier by the PREVIOUS item of the same phase (-1: none)
    uint16_t kt0;                      // != 0: the bias tile and the first kt0 K-columns travel as ONE weight tile (they are contiguous in the blob)
};
// One epilogue step: accumulator columns [tmem_col, tmem_col + nt) -> activation chunk buffer or the final output.
struct TcEpi {
    uint16_t tmem_col, nt, n0;         // n0: first output column (final layer)
    int8_t acc_bar, acc_k;             // accumulator-complete barrier + completion index
    int8_t dst;                        // -1 final output, else chunk buffer
    int8_t free_bar, free_k;           // MMAs still reading the chunk buffer (-1: none)
    int8_t done_bar;                   // arrives here when written (final output: when the accumulator has been read), -1: none
    int8_t phase;
    int8_t pad;
};
struct TcProgram {
    int odd_col;                         // TMEM column of the accumulators of odd layers: 256, or 128 in the compact variant
    int layers;
    int dims[PF_MAX_MLP_LAYERS + 1];     // true widths
    int pdims[PF_MAX_MLP_LAYERS + 1];    // padded to a multiple of 16
    int ngroups, nepis;
    uint32_t parity_mask;                // bit b: barrier b completes an odd number of times per item
    int phases;                          // 1, or 2: steps of phase 1 belong to the item of the PREVIOUS iteration (software pipeline)
    int nbuf;                            // activation chunk buffers
    int x0_nbuf;                         // staged-row buffers: 1, or 2 in two-phase programs (item m uses buffer m & 1 and barrier set m & 1)
    int8_t bar_kind[TC_NBARS];           // TC_BK_*
    int x0_bytes;
    long long blob_bytes;
    TcGroup g[TC_MAX_GROUPS];
    TcEpi e[TC_MAX_EPIS];
};

// [actor only, critic only, both].  The COMPACT variant (progc, stages_c ...) exists when both nets are small enough for
// TWO CTAs per SM: every accumulator within 256 TMEM columns, <= 113 KB of shared memory, 4 + 4 worker warps.
struct TcState {
    TcProgram prog[2]; int stages[3]; int smem_bytes[3]; int x0_bytes[3];
    int compact; TcProgram progc[2]; int stages_c[3]; int smem_bytes_c[3];
    int slot_bytes;
    int x0_nbuf_p;
    int pipe; TcProgram progp[2]; int stages_p[3]; int smem_bytes_p[3]; int slot_bytes_p;      // two-phase programs (large batches)
    unsigned long long* trace; int rehearse;
};

static inline int pad16(int x) { return (x + 15) & ~15; }

#include "pf_tc_ptx.cuh"
#include "pf_launch.h"

// ------------------------------------------------------------------------------------------------ kernel
enum { TC_OUT_PLAIN = 0, TC_OUT_SAMPLE = 1, TC_OUT_VALUE = 2 };

struct TcNetArgs {
    const unsigned char* blob;   // packed fp16 weight + bias tiles
    const float* x; long long ldx;
    float* out;                  // TC_OUT_PLAIN: [rows, dims[layers]] fp32
    float* copy_to;              // optional [rows, dims[0]] copy of the input rows (rollout storage)
    int mode;
    uint32_t magic;              // ceil(2^32 / dims[0]): element index -> row by multiply-high (filled by tc_launch)
};
struct TcArgs {
    TcProgram prog[2];
    TcNetArgs net[2];
    int nnets, rows, tiles, stages, x0_bytes, slot_bytes, nbuf, x0_nbuf;
    int tmem_cols;               // 512, or 256 in the compact variant (two CTAs per SM)
    // PPO.act tail (TC_OUT_SAMPLE / TC_OUT_VALUE)
    uint32_t seed, policy_step; int env_id_offset, num_actions;
    const float* std;
    pf_rl_storage S; int has_storage; long long storage_row0;
    float* actions_out; float* values_out; float* logp_out;
    // Hard-wired MMA role for the two-phase program of [*, 512, 256, 128, *] nets (tc_mma_fast3): per net, layer 0's
    // K = 16 steps in its first (bias + K) tile and in its second tile, the output layer's UMMA N, and whether its bias
    // travels with its one K tile.  fast3 = 0: the program interpreter.
    int fast3, f3_first[2], f3_rest[2], f3_nto[2], f3_merged[2];
    int weights_early;           // the loader may stream weight tiles before the dependency wait (pf_rl_weights_early)
    int rehearse;                // run one DRY item (no input loads, no outputs) before the dependency wait: see pf_rl_tc_rehearse
    int debug;                   // experiments only (PF_TC_DEBUG): bit 0 no storage copy, bit 1 no global loads, bit 3 no weight copies, bit 4 no MMAs,
                                 // bit 5 no barrier peek, bit 6 epilogue steps do no work, bit 7 epilogue converts without the ELU
    unsigned long long* trace;   // optional timeline of CTA 0: TC_TRACE_ROLES x TC_TRACE_CAP (tag << 48 | clock)
};

struct TcTrace {
    unsigned long long* p; int n;
    unsigned long long* tl;            // step timeline log (pf_rl_debug_timeline), read ONCE: events cost nothing when it is off
    __device__ __forceinline__ void init(unsigned long long* base, int role, bool on) {
        p = on && base ? base + role * TC_TRACE_CAP : nullptr; n = 0;
#ifdef PF_TIMELINE
        tl = on ? g_pf_timeline : nullptr;
#else
        tl = nullptr;
#endif
    }
    __device__ __forceinline__ void ev(unsigned tag) {
        if (p && n < TC_TRACE_CAP) p[n++] = ((unsigned long long)tag << 48) | ((unsigned long long)clock64() & 0xFFFFFFFFFFFFull);
        if (tl) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            const unsigned long long i = atomicAdd(tl, 1ull);
            if (i < 4094) tl[1 + i] = ((unsigned long long)(0x1000u + tag) << 48) | (t & 0xFFFFFFFFFFFFull);
        }
    }
};

// branch-free ELU: the exponential is evaluated unconditionally (one MUFU) and selected, so that 32 of them schedule
// as straight-line code (a data-dependent branch per element serialises the epilogue)
__device__ __forceinline__ float elu_fast(float x) {
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(fminf(x, 0.0f) * 1.4426950408889634f));
    return fmaxf(x, e - 1.0f);          // x > 0: e = 1 -> max(x, 0) = x;  x <= 0: exp(x) - 1 >= x
}

// ---- fp16 operands.  The tensor cores take f16 and bf16 at the same rate; f16 keeps 10 mantissa bits (bf16: 7) and lets the
// whole ELU run on PACKED pairs (below).  Its range is what it costs: conversions saturate at +-65504 instead of
// overflowing (observations are clipped to +-100 long before), and magnitudes under 6e-8 flush to zero.
__device__ __forceinline__ uint32_t tc_pack_h2(float lo, float hi) {           // {lo, hi} -> f16x2 (lo in the low half)
    uint32_t r;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
__device__ __forceinline__ unsigned short tc_h(float v) { return (unsigned short)(tc_pack_h2(v, 0.0f) & 0xFFFFu); }
// ELU of two accumulators, packed: convert first, then min / mul / 2x MUFU.EX2.F16 / sub / max on the pair -- 7 issue slots
// for two elements (the fp32 form: 11), and the result is the operand word.  |error| <= 2^-11 (the -1 after an f16 exp).
__device__ __forceinline__ uint32_t tc_elu_h2(uint32_t a0, uint32_t a1) {
    const uint32_t z = tc_pack_h2(__uint_as_float(a0), __uint_as_float(a1)), zero = 0u, log2e = 0x3DC53DC5u, one = 0x3C003C00u;
    uint32_t m, e, r;
    asm("min.f16x2 %0, %1, %2;" : "=r"(m) : "r"(z), "r"(zero));
    asm("mul.rn.f16x2 %0, %1, %2;" : "=r"(m) : "r"(m), "r"(log2e));
    asm("ex2.approx.f16x2 %0, %1;" : "=r"(e) : "r"(m));
    asm("sub.rn.f16x2 %0, %1, %2;" : "=r"(e) : "r"(e), "r"(one));
    asm("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(z), "r"(e));
    return r;
}
// instruction descriptor: D fp32, A/B f16 (format 0), both K-major, M = 128
__device__ __forceinline__ uint32_t make_idesc_f16(int n) {
    return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

// ---- input rows of item w: fp32 global -> fp16 core-matrix layout (+ the rollout-storage copy); all epilogue warps.
// The tile is one contiguous block of global memory, read (and copied to the storage row) with fully coalesced
// 128-bit accesses, 4 loads in flight per thread.  The four fp16 of a float4 land in one core-matrix row as a 64-bit
// shared store; consecutive lanes hold consecutive 8-column blocks, which is why the staged operand uses a PADDED
// leading byte offset (TC_X0_LBO = 2048 + 16): with the natural 2048 every lane of the warp would hit the same
// bank.  Out of line so that the loads in flight do not add to the epilogue's register pressure.
template <int TC_STG_THREADS>
__device__ __noinline__ void tc_stage_item(const TcArgs& a, int w, unsigned char* x0, uint32_t xfull_bar, int tid, TcTrace& tr, bool dry) {
    const int lane = tid & 31;
    const int k = w >= a.tiles ? 1 : 0, tile = w - k * a.tiles;          // (at most two nets per launch)
    const TcProgram& P = a.prog[k];
    const TcNetArgs& N = a.net[k];
    const int row0 = tile * TC_ROWS;
    const int nrows = min(TC_ROWS, a.rows - row0);
    const int K0 = P.dims[0], Kp = P.pdims[0];
    const uint32_t magic = N.magic;
    const float* xb = N.x + (long long)row0 * N.ldx;
    float* cb = (N.copy_to && !(a.debug & 1) && !dry) ? N.copy_to + (long long)row0 * K0 : nullptr;
    const bool noload = (a.debug & 2) || dry;       // dry item: same code, zeros instead of the rows
    const int total = nrows * K0;
    auto put = [&](int e, float v) {            // element e of the contiguous tile -> (row, column)
        const int r = (int)__umulhi((uint32_t)e, magic), kx = e - r * K0;
        *reinterpret_cast<unsigned short*>(x0 + (size_t)(kx >> 3) * TC_X0_LBO + r * 16 + (kx & 7) * 2) = tc_h(v);
    };
    if (N.ldx == K0 && ((reinterpret_cast<uintptr_t>(xb) | reinterpret_cast<uintptr_t>(cb)) & 15) == 0) {
        const int n4 = total >> 2;
        const float4* x4 = reinterpret_cast<const float4*>(xb);
        float4* c4 = reinterpret_cast<float4*>(cb);
        const bool row4 = (K0 & 3) == 0;        // a float4 never straddles rows or 8-column blocks
        for (int base = tid; base < n4; base += 4 * TC_STG_THREADS) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int idx = base + u * TC_STG_THREADS;
                if (idx < n4) v[u] = noload ? make_float4(0.f, 0.f, 0.f, 0.f) : __ldg(x4 + idx);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int idx = base + u * TC_STG_THREADS;
                if (idx < n4) {
                    if (c4) c4[idx] = v[u];
                    if (row4) {
                        const int e = idx * 4, r = (int)__umulhi((uint32_t)e, magic), kx = e - r * K0;
                        *reinterpret_cast<uint2*>(x0 + (size_t)(kx >> 3) * TC_X0_LBO + r * 16 + (kx & 7) * 2) =
                            make_uint2(tc_pack_h2(v[u].x, v[u].y), tc_pack_h2(v[u].z, v[u].w));
                    } else {
                        put(idx * 4, v[u].x); put(idx * 4 + 1, v[u].y); put(idx * 4 + 2, v[u].z); put(idx * 4 + 3, v[u].w);
                    }
                }
            }
        }
        for (int e = n4 * 4 + tid; e < total; e += TC_STG_THREADS) {
            const float v = noload ? 0.0f : __ldg(xb + e);
            if (cb) cb[e] = v;
            put(e, v);
        }
    } else {
        for (int e = tid; e < total; e += TC_STG_THREADS) {
            const int r = (int)__umulhi((uint32_t)e, magic), kx = e - r * K0;
            const float v = noload ? 0.0f : __ldg(xb + (long long)r * N.ldx + kx);
            if (cb) cb[e] = v;
            put(e, v);
        }
    }
    tr.ev(0x503);
    // zero padding: columns [K0, Kp) of the valid rows, and whole rows past the end of the batch
    // (no division: thread (row, column) pairs walk rows by TC_STG_THREADS / 16 and the <= 15 pad columns by 16)
    {
        const int padw = Kp - K0, pc = tid & 15;
        if (pc < padw) {
            const int kx = K0 + pc;
            unsigned char* col = x0 + (size_t)(kx >> 3) * TC_X0_LBO + (kx & 7) * 2;
            for (int r = tid >> 4; r < nrows; r += TC_STG_THREADS / 16) *reinterpret_cast<unsigned short*>(col + r * 16) = 0;
        }
        const int nblk = Kp >> 3;
        for (int r = nrows + (tid >> 5); r < TC_ROWS; r += TC_STG_THREADS / 32)
            for (int kb = lane; kb < nblk; kb += 32) *reinterpret_cast<uint4*>(x0 + (size_t)kb * TC_X0_LBO + r * 16) = make_uint4(0u, 0u, 0u, 0u);
    }
    tr.ev(0x504);
    fence_async_smem();
    __syncwarp();
    if (lane == 0) mbar_arrive(xfull_bar);
    tr.ev(0x505);
    // Pull the rows of the item after this one into L2 now.  All CTAs run in step, so without this every CTA would
    // hit HBM in the same instant one item later and sit out the burst inside its critical path; the prefetch spreads
    // that traffic under a whole item of MMAs and turns the next staging pass into L2 hits.
    const int w2 = w + (int)gridDim.x;
    if (!dry && w2 < a.nnets * a.tiles) {
        const int k2 = w2 >= a.tiles ? 1 : 0, tile2 = w2 - k2 * a.tiles;
        const TcNetArgs& N2 = a.net[k2];
        const int rows2 = min(TC_ROWS, a.rows - tile2 * TC_ROWS);
        const char* base = reinterpret_cast<const char*>(N2.x + (long long)tile2 * TC_ROWS * N2.ldx);
        const long long bytes = ((long long)(rows2 - 1) * N2.ldx + a.prog[k2].dims[0]) * 4;
        for (long long o = (long long)tid * 128; o < bytes; o += TC_STG_THREADS * 128)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(base + o));
    }
}

// PPO.act tail of one row: Normal(mean, std).sample() under the RNG contract + log_prob(actions).sum(-1)
// (actor_critic.py act / get_actions_log_prob) written straight into the rollout storage (rollout_storage.py
// add_transitions).  Out of line and rolled: logf / cosf / sqrtf are long, sixteen inlined copies of them made
// the kernel several times larger than the instruction cache.
__device__ __noinline__ void tc_sample_row(const TcArgs& a, long long er, const float* mean) {
    const int A = a.num_actions;
    const uint32_t ek = pf_rng_env_key(pf_rng_key(a.seed, a.policy_step, PF_RNG_POLICY), (uint32_t)(a.env_id_offset + (int)er));
    const long long srow = a.storage_row0 + er;
    float logp = 0.0f;
#pragma unroll 1
    for (int j = 0; j < A; ++j) {
        const float m = mean[j], sd = a.std[j];
        const float act = m + sd * pf_rng_normal(ek, (uint32_t)j);
        const float d = act - m;
        logp += -(d * d) / (2.0f * (sd * sd)) - logf(sd) - 0.9189385332046727f;
        if (a.actions_out) a.actions_out[er * A + j] = act;
        if (a.has_storage) {
            a.S.actions[srow * A + j] = act;
            a.S.mu[srow * A + j] = m;
            a.S.sigma[srow * A + j] = sd;
        }
    }
    if (a.logp_out) a.logp_out[er] = logp;
    if (a.has_storage) a.S.actions_log_prob[srow] = logp;
}

// A run of n accumulating N = 128 MMAs over the staged rows, the counts the usual input widths give unrolled (27 / 148 / 214
// inputs: 2 | 7 + 3 | 7 + 7 K = 16 steps): three instructions per MMA instead of a loop's six
__device__ __forceinline__ void tc_issue_x0_run(int n, uint32_t tmem_d, uint64_t ad, uint64_t bd, uint32_t idesc) {
    if (n == 7) issue_mmas<TC_X0_KSTEP, 256, 7>(tmem_d, ad, bd, idesc);
    else if (n == 3) issue_mmas<TC_X0_KSTEP, 256, 3>(tmem_d, ad, bd, idesc);
    else if (n == 2) issue_mmas<TC_X0_KSTEP, 256, 2>(tmem_d, ad, bd, idesc);
    else for (int k = 0; k < n; ++k) { umma_bf16_acc(tmem_d, ad, bd, idesc); ad += TC_X0_KSTEP; bd += 256; }
}

// The MMA role of the two-phase program for the nets every task of the reference uses ([in, 512, 256, 128, out]), written
// out: the eleven groups of an iteration in the order build_program_pipe gives them, every descriptor stride, tile count,
// barrier and parity a constant of the code -- no step table, no dispatch.  The program interpreter spends ~1.3k cycles
// per group on itself with NOTHING to issue (profiles/r2q_mlp_tc_rough_investigation.txt), 11 groups per 128 rows; this
// role is the only one whose time a two-phase iteration cannot hide.  The loader, the epilogue and the staging warps keep
// walking the program, so the two must agree: pf_tc launch code sets fast3 only after comparing the program's groups
// (order, UMMA N, K extents, tile sizes, barriers) with what is written here.
//   B = the item of this iteration's phase 0 (number mB = it), A = the item of its phase 1 (mA = it - 1).
//   Completions per item: ACC0 4, E0 4, AFREE0 4, ACC1 1, E1 3, AFREE1 1, ACC2 1, ACC3 1, ODONE 1; XFULL / XFREE per buffer.
__device__ __noinline__ void tc_mma_fast3(const TcArgs& a, int n_local, int n_iter, uint32_t x0_addr, uint32_t ach_addr, uint32_t ones_addr,
                                          uint32_t ring_addr, uint32_t full0, uint32_t empty0, uint32_t pb0, uint32_t tmem_base) {
    const uint32_t S = (uint32_t)a.stages;
    uint32_t slot = 0, rpar = 0, full_bar = full0, empty_bar = empty0;
    const uint64_t ring_desc0 = make_desc(ring_addr, 0u, 128u), slot_step = (uint64_t)((uint32_t)a.slot_bytes >> 4);
    uint64_t ring_desc = ring_desc0;
    const uint64_t ones_desc = make_desc(ones_addr, TC_ROWS * 16, 128u);
    const uint64_t ach0_desc = make_desc(ach_addr, TC_ROWS * 16, 128u), ach1_desc = make_desc(ach_addr + TC_ACHUNK_BYTES, TC_ROWS * 16, 128u);
    const uint64_t x0_desc0 = make_desc(x0_addr, TC_X0_LBO, 128u), x0_desc1 = make_desc(x0_addr + (uint32_t)a.x0_bytes, TC_X0_LBO, 128u);
    const uint32_t idesc128 = make_idesc_f16(128), idesc256 = make_idesc_f16(256);
    const uint64_t lbo128 = (uint64_t)(128u * 16u >> 4) << 16, lbo256 = (uint64_t)(256u * 16u >> 4) << 16;
    const uint32_t tS = tmem_base, tT = tmem_base + 128, tA1 = tmem_base + 256;
    auto bar = [&](int b) { return pb0 + 8u * (uint32_t)b; };
    auto next_slot = [&]() {
        full_bar += 8; empty_bar += 8; ring_desc += slot_step;
        if (++slot == S) { slot = 0; rpar ^= 1u; full_bar = full0; empty_bar = empty0; ring_desc = ring_desc0; }
    };
    const int gridn = (int)gridDim.x, cta = (int)blockIdx.x, tiles = a.tiles;
    TcTrace tr; tr.init(a.trace, 1, blockIdx.x == 0 && (threadIdx.x & 31) == 0);
    for (int it = 0; it < n_iter; ++it) {
        if (it == 0) pf_grid_dep_wait();
        const bool hasB = it < n_local, hasA = it >= 1;
        const int kB = (hasB && cta + it * gridn >= tiles) ? 1 : 0, kA = (hasA && cta + (it - 1) * gridn >= tiles) ? 1 : 0;
        const uint32_t mB = (uint32_t)it, mA = (uint32_t)(it - 1), xb = a.x0_nbuf == 2 ? (mB & 1u) : 0u, xuse = a.x0_nbuf == 2 ? (mB >> 1) : mB;
        const uint64_t xdesc = xb ? x0_desc1 : x0_desc0;
        const int nfirst = kB ? a.f3_first[1] : a.f3_first[0], nrest = kB ? a.f3_rest[1] : a.f3_rest[0];
        // ---- layer 0, chunk c of item B: S (+)= x0 * W0[c]^T, bias + first K columns in one tile, the rest in a second
        auto l0 = [&](int c, int gi) {
            if (!hasB) return;
            tr.ev(0x200 + gi);
            if (c == 0) mbar_wait(bar(BAR_XFULL) + 8u * xb * TC_XBAR_STEP, xuse & 1u);
            else mbar_wait(bar(BAR_E0), (uint32_t)(c - 1) & 1u);                    // S drained (and chunk c - 1 in buffer 0)
            tc_fence_after();
            tr.ev(0x300 + gi);
            mbar_wait(full_bar, rpar);
            if (elect_one()) {
                uint64_t bd = ring_desc + lbo128, ad = xdesc;
                umma_bf16(tS, ones_desc, bd, idesc128, 0u);
                bd += 256;
                tc_issue_x0_run(nfirst, tS, ad, bd, idesc128);
                umma_commit(empty_bar);
            }
            next_slot();
            if (nrest) {
                mbar_wait(full_bar, rpar);
                if (elect_one()) {
                    uint64_t bd = ring_desc + lbo128, ad = xdesc + (uint64_t)(nfirst * TC_X0_KSTEP);
                    tc_issue_x0_run(nrest, tS, ad, bd, idesc128);
                    umma_commit(empty_bar);
                }
                next_slot();
            }
            if (elect_one()) {
                umma_commit(bar(BAR_ACC0));
                if (c == 3) umma_commit(bar(BAR_XFREE) + 8u * xb * TC_XBAR_STEP);
            }
            __syncwarp();
            tr.ev(0x400 + gi);
        };
        // ---- layer 1, K-slice c of item B (chunk c from buffer 0): A1 (+)= chunk * W1[:, 128c : 128c + 128]^T, two tiles of 64
        auto l1 = [&](int c, int gi) {
            if (!hasB) return;
            tr.ev(0x200 + gi);
            mbar_wait(bar(BAR_E0), (uint32_t)c & 1u);
            tc_fence_after();
            tr.ev(0x300 + gi);
            if (c == 0) {
                mbar_wait(full_bar, rpar);
                if (elect_one()) { umma_bf16(tA1, ones_desc, ring_desc + lbo256, idesc256, 0u); umma_commit(empty_bar); }
                next_slot();
            }
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                mbar_wait(full_bar, rpar);
                if (elect_one()) {
                    issue_mmas<256, 512, 4>(tA1, ach0_desc + (uint64_t)(t * 4 * 256), ring_desc + lbo256, idesc256);
                    umma_commit(empty_bar);
                }
                next_slot();
            }
            if (elect_one()) {
                umma_commit(bar(BAR_AFREE0));
                if (c == 3) umma_commit(bar(BAR_ACC1));
            }
            __syncwarp();
            tr.ev(0x400 + gi);
        };
        // ---- layer 2, K-slice j of item A (layer-1 chunk j from buffer 1): T (+)= chunk * W2[:, 128j : 128j + 128]^T
        auto l2 = [&](int j, int gi) {
            if (!hasA) return;
            tr.ev(0x200 + gi);
            mbar_wait(bar(BAR_E1), (mA + (uint32_t)j) & 1u);                        // completion 3 mA + j
            if (j == 0 && mA >= 1u) mbar_wait(bar(BAR_ODONE), (mA - 1u) & 1u);        // T: the previous item's output accumulator has been read
            tc_fence_after();
            tr.ev(0x300 + gi);
            if (j == 0) {
                mbar_wait(full_bar, rpar);
                if (elect_one()) { umma_bf16(tT, ones_desc, ring_desc + lbo128, idesc128, 0u); umma_commit(empty_bar); }
                next_slot();
            }
            mbar_wait(full_bar, rpar);
            if (elect_one()) {
                issue_mmas<256, 256, 8>(tT, ach1_desc, ring_desc + lbo128, idesc128);
                umma_commit(empty_bar);
            }
            next_slot();
            if (elect_one()) umma_commit(bar(j == 0 ? BAR_AFREE1 : BAR_ACC2));
            __syncwarp();
            tr.ev(0x400 + gi);
        };
        // ---- output layer of item A (layer-2 activations from buffer 1): T = h3 * W3^T + b3, K = 128
        auto l3 = [&](int gi) {
            if (!hasA) return;
            tr.ev(0x200 + gi);
            mbar_wait(bar(BAR_E1), mA & 1u);                                         // completion 3 mA + 2
            tc_fence_after();
            tr.ev(0x300 + gi);
            const int nt = kA ? a.f3_nto[1] : a.f3_nto[0], merged = kA ? a.f3_merged[1] : a.f3_merged[0];
            const uint32_t idesc = make_idesc_f16(nt);
            const uint64_t lbo = (uint64_t)((uint32_t)(nt * 16) >> 4) << 16, bstep = (uint64_t)(nt * 2);
            mbar_wait(full_bar, rpar);
            if (!merged) {
                if (elect_one()) { umma_bf16(tT, ones_desc, ring_desc + lbo, idesc, 0u); umma_commit(empty_bar); }
                next_slot();
                mbar_wait(full_bar, rpar);
            }
            if (elect_one()) {
                uint64_t bd = ring_desc + lbo, ad = ach1_desc;
                if (merged) { umma_bf16(tT, ones_desc, bd, idesc, 0u); bd += bstep; }
#pragma unroll
                for (int k = 0; k < 8; ++k) { umma_bf16_acc(tT, ad, bd, idesc); ad += 256; bd += bstep; }
                umma_commit(empty_bar);
            }
            next_slot();
            if (elect_one()) umma_commit(bar(BAR_ACC3));
            __syncwarp();
            tr.ev(0x400 + gi);
        };
        l0(0, 0); l2(0, 1); l0(1, 2); l2(1, 3); l1(0, 4); l0(2, 5); l1(1, 6); l3(7); l0(3, 8); l1(2, 9); l1(3, 10);
    }
}

// The MMA role of the SINGLE-phase program of the same nets, written out the same way (single-item launches, inputs too wide
// for two staged-row buffers): groups in build_program's order, two layer-0 accumulator slots at TMEM [0, 128) / [128, 256),
// layer 1 at [256, 512), layer 2 at [0, 128), the output layer at [256, ...), chunk buffers alternating.
//   Completions per item: ACC0 3, ACC1 2, ACC2 2, AFREE0 4, AFREE1 3, E0 4, E1 3, XFULL 1, XFREE 1.
__device__ __noinline__ void tc_mma_fast1(const TcArgs& a, int n_local, int reh, uint32_t x0_addr, uint32_t ach_addr, uint32_t ones_addr,
                                          uint32_t ring_addr, uint32_t full0, uint32_t empty0, uint32_t pb0, uint32_t tmem_base) {
    const uint32_t S = (uint32_t)a.stages;
    uint32_t slot = 0, rpar = 0, full_bar = full0, empty_bar = empty0;
    const uint64_t ring_desc0 = make_desc(ring_addr, 0u, 128u), slot_step = (uint64_t)((uint32_t)a.slot_bytes >> 4);
    uint64_t ring_desc = ring_desc0;
    const uint64_t ones_desc = make_desc(ones_addr, TC_ROWS * 16, 128u);
    const uint64_t ach0_desc = make_desc(ach_addr, TC_ROWS * 16, 128u), ach1_desc = make_desc(ach_addr + TC_ACHUNK_BYTES, TC_ROWS * 16, 128u);
    const uint64_t xdesc = make_desc(x0_addr, TC_X0_LBO, 128u);
    const uint32_t idesc128 = make_idesc_f16(128), idesc256 = make_idesc_f16(256);
    const uint64_t lbo128 = (uint64_t)(128u * 16u >> 4) << 16, lbo256 = (uint64_t)(256u * 16u >> 4) << 16;
    auto bar = [&](int b) { return pb0 + 8u * (uint32_t)b; };
    auto next_slot = [&]() {
        full_bar += 8; empty_bar += 8; ring_desc += slot_step;
        if (++slot == S) { slot = 0; rpar ^= 1u; full_bar = full0; empty_bar = empty0; ring_desc = ring_desc0; }
    };
    const int gridn = (int)gridDim.x, cta = (int)blockIdx.x, tiles = a.tiles;
    TcTrace tr; tr.init(a.trace, 1, blockIdx.x == 0 && (threadIdx.x & 31) == 0);
    for (int j = -reh; j < n_local; ++j) {
        if (j == 0) pf_grid_dep_wait();
        const uint32_t me = (uint32_t)(j + reh);                                     // items walked before this one (the dry one included)
        const int k = (cta + (j < 0 ? 0 : j) * gridn >= tiles) ? 1 : 0;
        const int nfirst = k ? a.f3_first[1] : a.f3_first[0], nrest = k ? a.f3_rest[1] : a.f3_rest[0];
        auto l0 = [&](int c, int gi) {
            tr.ev(0x200 + gi);
            if (c == 0) mbar_wait(bar(BAR_XFULL), me & 1u);
            else if (c == 2) mbar_wait(bar(BAR_E0), 0u);                              // slot 0 drained: E0 completion 4 me + 0
            else if (c == 3) mbar_wait(bar(BAR_E1), me & 1u);                         // slot 1 drained: E1 completion 3 me + 0
            tc_fence_after();
            tr.ev(0x300 + gi);
            const uint32_t tS = tmem_base + (uint32_t)(c & 1) * 128u;
            mbar_wait(full_bar, rpar);
            if (elect_one()) {
                uint64_t bd = ring_desc + lbo128, ad = xdesc;
                umma_bf16(tS, ones_desc, bd, idesc128, 0u);
                bd += 256;
                tc_issue_x0_run(nfirst, tS, ad, bd, idesc128);
                umma_commit(empty_bar);
            }
            next_slot();
            if (nrest) {
                mbar_wait(full_bar, rpar);
                if (elect_one()) {
                    uint64_t bd = ring_desc + lbo128, ad = xdesc + (uint64_t)(nfirst * TC_X0_KSTEP);
                    tc_issue_x0_run(nrest, tS, ad, bd, idesc128);
                    umma_commit(empty_bar);
                }
                next_slot();
            }
            if (elect_one()) {
                umma_commit(bar(BAR_ACC0 + (c & 1)));
                if (c == 3) umma_commit(bar(BAR_XFREE));
            }
            __syncwarp();
            tr.ev(0x400 + gi);
        };
        auto l1 = [&](int c, int gi) {
            tr.ev(0x200 + gi);
            // chunk c sits in buffer c & 1: E0 completions 4 me + {0, 1}, E1 completions 3 me + {0, 1}
            if (c & 1) mbar_wait(bar(BAR_E1), (me + (uint32_t)(c >> 1)) & 1u);
            else mbar_wait(bar(BAR_E0), (uint32_t)(c >> 1) & 1u);
            tc_fence_after();
            tr.ev(0x300 + gi);
            const uint32_t tA1 = tmem_base + 256;
            const uint64_t adesc = (c & 1) ? ach1_desc : ach0_desc;
            if (c == 0) {
                mbar_wait(full_bar, rpar);
                if (elect_one()) { umma_bf16(tA1, ones_desc, ring_desc + lbo256, idesc256, 0u); umma_commit(empty_bar); }
                next_slot();
            }
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                mbar_wait(full_bar, rpar);
                if (elect_one()) {
                    issue_mmas<256, 512, 4>(tA1, adesc + (uint64_t)(t * 4 * 256), ring_desc + lbo256, idesc256);
                    umma_commit(empty_bar);
                }
                next_slot();
            }
            if (elect_one()) {
                umma_commit(bar(BAR_AFREE0 + (c & 1)));
                if (c == 3) umma_commit(bar(BAR_ACC2));
            }
            __syncwarp();
            tr.ev(0x400 + gi);
        };
        auto l2 = [&](int jj, int gi) {
            tr.ev(0x200 + gi);
            if (jj) mbar_wait(bar(BAR_E1), (me + 2u) & 1u);                           // layer-1 chunk 1 in buffer 1: E1 completion 3 me + 2
            else mbar_wait(bar(BAR_E0), 0u);                                          // layer-1 chunk 0 in buffer 0: E0 completion 4 me + 2
            tc_fence_after();
            tr.ev(0x300 + gi);
            const uint32_t tT = tmem_base;
            if (jj == 0) {
                mbar_wait(full_bar, rpar);
                if (elect_one()) { umma_bf16(tT, ones_desc, ring_desc + lbo128, idesc128, 0u); umma_commit(empty_bar); }
                next_slot();
            }
            mbar_wait(full_bar, rpar);
            if (elect_one()) {
                issue_mmas<256, 256, 8>(tT, jj ? ach1_desc : ach0_desc, ring_desc + lbo128, idesc128);
                umma_commit(empty_bar);
            }
            next_slot();
            if (elect_one()) {
                umma_commit(bar(BAR_AFREE0 + jj));
                if (jj == 1) umma_commit(bar(BAR_ACC0));
            }
            __syncwarp();
            tr.ev(0x400 + gi);
        };
        auto l3 = [&](int gi) {
            tr.ev(0x200 + gi);
            mbar_wait(bar(BAR_E0), 1u);                                               // layer-2 activations in buffer 0: E0 completion 4 me + 3
            tc_fence_after();
            tr.ev(0x300 + gi);
            const uint32_t tO = tmem_base + 256;
            const int nt = k ? a.f3_nto[1] : a.f3_nto[0], merged = k ? a.f3_merged[1] : a.f3_merged[0];
            const uint32_t idesc = make_idesc_f16(nt);
            const uint64_t lbo = (uint64_t)((uint32_t)(nt * 16) >> 4) << 16, bstep = (uint64_t)(nt * 2);
            mbar_wait(full_bar, rpar);
            if (!merged) {
                if (elect_one()) { umma_bf16(tO, ones_desc, ring_desc + lbo, idesc, 0u); umma_commit(empty_bar); }
                next_slot();
                mbar_wait(full_bar, rpar);
            }
            if (elect_one()) {
                uint64_t bd = ring_desc + lbo, ad = ach0_desc;
                if (merged) { umma_bf16(tO, ones_desc, bd, idesc, 0u); bd += bstep; }
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) { umma_bf16_acc(tO, ad, bd, idesc); ad += 256; bd += bstep; }
                umma_commit(empty_bar);
            }
            next_slot();
            if (elect_one()) { umma_commit(bar(BAR_AFREE0)); umma_commit(bar(BAR_ACC2)); }
            __syncwarp();
            tr.ev(0x400 + gi);
        };
        l0(0, 0); l0(1, 1); l1(0, 2); l0(2, 3); l1(1, 4); l0(3, 5); l1(2, 6); l1(3, 7); l2(0, 8); l2(1, 9); l3(10);
    }
}

// The loader of the same program, written out the same way: the weight tiles of an iteration in the order tc_mma_fast3
// consumes them.  With two 32 KB ring slots (what two staged-row buffers leave) a tile's round trip -- the MMAs' commit, this
// thread's wake-up and bookkeeping, the bulk copy -- is on the MMA role's critical path, so its bookkeeping counts too.
__device__ __noinline__ void tc_load_fast3(const TcArgs& a, int n_local, int n_iter, uint32_t ring_addr, uint32_t full0, uint32_t empty0) {
    const uint32_t S = (uint32_t)a.stages, slot_bytes = (uint32_t)a.slot_bytes;
    uint32_t slot = 0, round = 0;
    const int gridn = (int)gridDim.x, cta = (int)blockIdx.x, tiles = a.tiles;
    auto tile = [&](const unsigned char* src, uint32_t bytes) {
        if (round) mbar_wait(empty0 + 8 * slot, (round - 1u) & 1u);
        if (elect_one()) {
            mbar_expect_tx(full0 + 8 * slot, bytes);
            bulk_g2s(ring_addr + slot * slot_bytes, src, bytes, full0 + 8 * slot);
        }
        if (++slot == S) { slot = 0; ++round; }
    };
    for (int it = 0; it < n_iter; ++it) {
        if (it == 0 && !a.weights_early) pf_grid_dep_wait();
        const bool hasB = it < n_local, hasA = it >= 1;
        const int kB = (hasB && cta + it * gridn >= tiles) ? 1 : 0, kA = (hasA && cta + (it - 1) * gridn >= tiles) ? 1 : 0;
        const unsigned char* blobB = a.net[kB].blob;
        const unsigned char* blobA = a.net[kA].blob;
        const TcProgram& PB = a.prog[kB];
        const TcProgram& PA = a.prog[kA];
        const uint32_t l0a = (uint32_t)(128 * (16 + 16 * (kB ? a.f3_first[1] : a.f3_first[0])) * 2), l0b = (uint32_t)(128 * 16 * (kB ? a.f3_rest[1] : a.f3_rest[0]) * 2);
        auto l0 = [&](int gi) { if (!hasB) return; const unsigned char* p = blobB + PB.g[gi].blob_off; tile(p, l0a); if (l0b) tile(p + l0a, l0b); };
        auto l1 = [&](int gi, bool bias) { if (!hasB) return; const unsigned char* p = blobB + PB.g[gi].blob_off; if (bias) { tile(p, 8192u); p += 8192; } tile(p, 32768u); tile(p + 32768, 32768u); };
        auto l2 = [&](int gi, bool bias) { if (!hasA) return; const unsigned char* p = blobA + PA.g[gi].blob_off; if (bias) { tile(p, 4096u); p += 4096; } tile(p, 32768u); };
        auto l3 = [&](int gi) {
            if (!hasA) return;
            const unsigned char* p = blobA + PA.g[gi].blob_off;
            const uint32_t nt = (uint32_t)(kA ? a.f3_nto[1] : a.f3_nto[0]);
            if (kA ? a.f3_merged[1] : a.f3_merged[0]) tile(p, nt * (16u + 128u) * 2u);
            else { tile(p, nt * 32u); tile(p + nt * 32u, nt * 256u); }
        };
        l0(0); l2(1, true); l0(2); l2(3, false); l1(4, true); l0(5); l1(6, false); l3(7); l0(8); l1(9, false); l1(10, false);
    }
}

template <int TC_EPI_WARPS, int TC_STG_WARPS, int kCtasPerSm>
__global__ void __launch_bounds__(TC_THREADS_OF(TC_EPI_WARPS, TC_STG_WARPS), kCtasPerSm) pf_mlp_tc_kernel(const __grid_constant__ TcArgs a) {
    constexpr int TC_EPI_THREADS = TC_EPI_WARPS * 32, TC_STG_THREADS = TC_STG_WARPS * 32;
    constexpr int TC_MMA_WARP = TC_EPI_WARPS, TC_LOAD_WARP = TC_EPI_WARPS + 4;
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* x0 = smem;                                  // staged input rows [128 x pdims[0]] fp16 (x 2 in two-phase programs)
    unsigned char* ach = smem + a.x0_nbuf * a.x0_bytes;        // activation chunk buffers [128 x 128] fp16
    unsigned char* ones = ach + a.nbuf * TC_ACHUNK_BYTES;      // [128 x 16] fp16: columns 0,1 = 1
    unsigned char* ring = ones + TC_ROWS * 32;
    uint64_t* bars = reinterpret_cast<uint64_t*>(ring + a.stages * a.slot_bytes);
    const int S = a.stages;
    const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + TC_MAX_STAGES), pb0 = smem_u32(bars + 2 * TC_MAX_STAGES);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * TC_MAX_STAGES + TC_NBARS);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int stg_d = warp - TC_EPI_WARPS;                                          // staging warps: offsets 1, 2, 3, 5, 6, 7, 9, ...
    const int stg_warp = (stg_d > 0 && (stg_d & 3)) ? stg_d - 1 - (stg_d >> 2) : 1 << 20, stg_tid = stg_warp * 32 + lane;
    const int items = a.nnets * a.tiles;
    pf_grid_dep_launch();       // the prologue below (barriers, TMEM, ones operand) runs under the predecessor's tail
    pf_timeline(0x50);

    if (warp == TC_LOAD_WARP && lane == 0) {
        for (int s = 0; s < S; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 1); }
        for (int b = 0; b < TC_NBARS; ++b) {                    // one arrival per warp, or one tcgen05.commit
            const int kind = a.prog[0].bar_kind[b];
            mbar_init(pb0 + 8 * b, kind == TC_BK_EPI ? TC_EPI_WARPS : kind == TC_BK_STG ? TC_STG_WARPS : 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == TC_MMA_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"((uint32_t)a.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    } else if (warp < TC_EPI_WARPS) {
        for (int i = tid; i < TC_ROWS * 2; i += TC_EPI_THREADS)             // ones operand: 2 column blocks x 128 rows x 16 B
            *reinterpret_cast<uint4*>(ones + i * 16) = i < TC_ROWS ? make_uint4(0x3C003C00u, 0u, 0u, 0u) : make_uint4(0u, 0u, 0u, 0u);
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pf_timeline(0x51);
    // Every role below walks ITERATIONS it = -reh, 0, 1, ..., and in each one the steps of the program in program order.
    // A step of phase ph belongs to the CTA's item number m = it - ph: in a two-phase program the second half of item m
    // (its last layers) is interleaved, step by step, with the first half of item m + 1, so that the epilogue warps and
    // the tensor pipe always have the other item's work while one item's chain waits for a hand-off.  A single-phase
    // program is the degenerate case (one item per iteration).
    // With a.rehearse the first pass (it = -1) is a DRY copy of the CTA's first item -- zeros instead of the input rows,
    // no outputs, every barrier used exactly as by a real item -- executed BEFORE the dependency wait: launched as a
    // programmatic dependent, the CTA would otherwise sit idle at that wait while its predecessor runs, and then execute
    // every role's code for the first time (instruction fetches all the way from DRAM after an L2 flush) on the critical
    // path.  Each role passes the wait at it = 0.  (Single-phase programs only.)
    const int reh = a.rehearse ? 1 : 0;
    const int n_local = (items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;      // items of this CTA (>= 1)
    const int n_iter = n_local + a.prog[0].phases - 1;
    // item of phase ph in iteration it: its work index, or -1 when that phase has no item in this iteration
    auto item_of = [&](int it, int ph) -> int {
        if (it < 0) return ph == 0 ? (int)blockIdx.x : -1;
        const int m = it - ph;
        return (m >= 0 && m < n_local) ? (int)blockIdx.x + m * (int)gridDim.x : -1;
    };
    // parity to wait for: completion k (inside the item) of barrier b, given the barrier completions of all earlier items
    auto par_of = [](uint32_t bits, int b, int k) -> uint32_t { return ((bits >> b) ^ (uint32_t)k) & 1u; };

    if (warp == TC_LOAD_WARP && a.fast3 == 1 && !(a.debug & 8)) {
        tc_load_fast3(a, n_local, n_iter, smem_u32(ring), full0, empty0);
    } else if (warp == TC_LOAD_WARP) {
        // ================================================================ loader: one thread, runs ahead of the MMAs
        {   // (the whole warp walks the loop, one elected lane issues: code the compiler KNOWS to be warp-uniform keeps
            // the copy's operands in uniform registers; a lone lane made it re-broadcast every operand per instruction)
            uint32_t slot = 0, round = 0;           // ring position: slot index and how many times the ring wrapped
            const uint32_t ring0 = smem_u32(ring), slot_bytes = (uint32_t)a.slot_bytes;
            TcTrace tr; tr.init(a.trace, 2, blockIdx.x == 0 && lane == 0);
            for (int it = -reh; it < n_iter; ++it) {
                if (it == 0 && !a.weights_early) pf_grid_dep_wait();      // (the dry item read the weights early: its results are discarded)
                const int w0 = item_of(it, 0), w1 = item_of(it, 1);
                const int k0n = w0 >= 0 ? (w0 >= a.tiles ? 1 : 0) : -1, k1n = w1 >= 0 ? (w1 >= a.tiles ? 1 : 0) : -1;
                const TcProgram& PR = a.prog[k0n >= 0 ? k0n : k1n];          // step list (the same for both nets when two-phase)
                const int ngroups = PR.ngroups;
                for (int gi = 0; gi < ngroups; ++gi) {
                    const int k = PR.g[gi].phase ? k1n : k0n;
                    if (k < 0) continue;
                    const TcGroup G = a.prog[k].g[gi];
                    const unsigned char* blob = a.net[k].blob;
                    tr.ev(0x100 + gi);
                    uint32_t off = G.blob_off;
                    const int nt = G.nt, kc = G.kc, kt = G.kt;
                    // tiles: [bias] [kt] [kt] ..., or with kt0: [bias + kt0] [kt] ...
                    for (int kk = (G.bias && !G.kt0) ? -16 : 0; kk < kc; kk += (kk < 0 ? 16 : (kk == 0 && G.kt0) ? (int)G.kt0 : kt)) {
                        const uint32_t bytes = (kk == 0 && G.kt0) ? (uint32_t)(nt * (16 + (int)G.kt0) * 2)
                                                                  : (uint32_t)(nt * (kk < 0 ? 16 : min(kt, kc - kk)) * 2);
                        if (round) mbar_wait(empty0 + 8 * slot, (round - 1u) & 1u);
                        if (elect_one()) {
                            if (a.debug & 8) mbar_arrive(full0 + 8 * slot);          // experiment: no weight traffic
                            else {
                                mbar_expect_tx(full0 + 8 * slot, bytes);
                                bulk_g2s(ring0 + slot * slot_bytes, blob + off, bytes, full0 + 8 * slot);
                            }
                        }
                        off += bytes;
                        if (++slot == (uint32_t)S) { slot = 0; ++round; }
                    }
                }
            }
        }
    } else if (warp == TC_MMA_WARP && a.fast3 == 2) {
        // ================================================================ MMA issue, hard-wired, single-phase program (tc_mma_fast1)
        tc_mma_fast1(a, n_local, reh, smem_u32(x0), smem_u32(ach), smem_u32(ones), smem_u32(ring), full0, empty0, pb0, tmem_base);
    } else if (warp == TC_MMA_WARP && a.fast3) {
        // ================================================================ MMA issue, hard-wired (see tc_mma_fast3 below)
        tc_mma_fast3(a, n_local, n_iter, smem_u32(x0), smem_u32(ach), smem_u32(ones), smem_u32(ring), full0, empty0, pb0, tmem_base);
    } else if (warp == TC_MMA_WARP) {
        // ================================================================ MMA issue: one thread
        // The issuing thread is the critical resource of the kernel: traced at 262,144 rows, every weight tile costs it a
        // barrier wait (>= 90 cycles even when complete) + ~90 of bookkeeping + the MMAs, against 64 cycles of tensor-pipe
        // time for an N = 128 MMA.  The whole warp walks the loop and one elected lane issues: code the compiler knows to
        // be warp-uniform keeps descriptors in uniform registers (three instructions per MMA from the compile-time
        // unrolled runs below); run by a lone lane, every MMA re-broadcast its seven operand registers first.
        {
            uint32_t slot = 0, round = 0, pbits = 0;
            uint32_t full_bar = full0, empty_bar = empty0;
            const uint64_t ring_desc0 = make_desc(smem_u32(ring), 0u, 128u);        // leading byte offset added per group
            const uint64_t slot_step = (uint64_t)((uint32_t)a.slot_bytes >> 4);
            uint64_t ring_desc = ring_desc0;
            const uint64_t ones_desc = make_desc(smem_u32(ones), TC_ROWS * 16, 128u);
            const uint32_t x0_addr = smem_u32(x0), ach_addr = smem_u32(ach);
            const bool no_mma = (a.debug & 16) != 0;                                 // experiment: no MMAs
            const bool peek = (a.debug & 32) == 0;                                   // experiment: bit 5 = no barrier peeking
            TcTrace tr; tr.init(a.trace, 1, blockIdx.x == 0 && lane == 0);
            auto next_slot = [&]() {
                full_bar += 8; empty_bar += 8; ring_desc += slot_step;
                if (++slot == (uint32_t)S) { slot = 0; ++round; full_bar = full0; empty_bar = empty0; ring_desc = ring_desc0; }
            };
            uint32_t pbits1 = 0, pprev1 = 0, pprev0 = 0;  // parities before phase 1's current item; before the previous item of each phase
            for (int it = -reh; it < n_iter; ++it) {
                if (it == 0) pf_grid_dep_wait();
                const int w0 = item_of(it, 0), w1 = item_of(it, 1);
                const int k0n = w0 >= 0 ? (w0 >= a.tiles ? 1 : 0) : -1, k1n = w1 >= 0 ? (w1 >= a.tiles ? 1 : 0) : -1;
                const TcProgram& PR = a.prog[k0n >= 0 ? k0n : k1n];
                const int ngroups = PR.ngroups;
                for (int gi = 0; gi < ngroups; ++gi) {
                    const int ph = PR.g[gi].phase;
                    const int k = ph ? k1n : k0n;
                    if (k < 0) continue;
                    const TcGroup G = a.prog[k].g[gi];
                    tr.ev(0x200 + gi);
                    // two staged-row buffers (two-phase programs): item m of phase 0 uses buffer / barrier set m & 1, use number m >> 1
                    const uint32_t xb = a.x0_nbuf == 2 ? (uint32_t)(it & 1) : 0u;
                    if (G.wait_bar == BAR_XFULL && a.x0_nbuf == 2) mbar_wait(pb0 + 8 * (BAR_XFULL + xb * TC_XBAR_STEP), (uint32_t)(it >> 1) & 1u);
                    else if (G.wait_bar >= 0) mbar_wait(pb0 + 8 * G.wait_bar, par_of(ph ? pbits1 : pbits, G.wait_bar, G.wait_k));
                    // (two-phase programs) a resource the previous item of this phase still held: its one completion of
                    // prev_bar -- there is no such item in the phase's first iteration
                    if (G.prev_bar >= 0 && it - ph >= 1) mbar_wait(pb0 + 8 * G.prev_bar, par_of(ph ? pprev1 : pprev0, G.prev_bar, G.prev_k));
                    tc_fence_after();
                    tr.ev(0x300 + gi);
                    const int nt = G.nt;
                    const bool xs = G.a_src == 0;
                    const uint32_t idesc = make_idesc_f16(nt);
                    const uint32_t tmem_d = tmem_base + G.tmem_col;
                    const uint32_t a_base = xs ? x0_addr + xb * (uint32_t)a.x0_bytes : ach_addr + (uint32_t)(G.a_src - 1) * TC_ACHUNK_BYTES;
                    const uint32_t a_lbo = xs ? TC_X0_LBO : TC_ROWS * 16;
                    uint64_t adesc = make_desc(a_base, a_lbo, 128u);
                    const uint64_t a_step = (uint64_t)((2 * a_lbo) >> 4);               // K = 16 further along the A operand
                    const uint64_t b_lbo = (uint64_t)((uint32_t)(nt * 16) >> 4) << 16;  // next 8 K-columns of a weight tile
                    const uint64_t b_step = (uint64_t)(nt * 2);                         // K = 16 further inside a tile, in 16 B units
                    if (G.bias && !G.kt0) {
                        tr.ev(0xA00 + gi);
                        mbar_wait(full_bar, round & 1u);
                        tc_fence_after();
                        tr.ev(0xB00 + gi);
                        if (elect_one()) {
                            umma_bf16(tmem_d, ones_desc, ring_desc + b_lbo, idesc, 0u);
                            umma_commit(empty_bar);
                        }
                        next_slot();
                    }
                    int left = G.kc >> 4;                       // K = 16 steps still to issue
                    const int per_tile = G.kt >> 4;
                    uint32_t have = 0;                          // the next tile's barrier was already seen complete
                    bool first = true;
                    while (left > 0) {
                        tr.ev(0xA00 + gi);
                        if (!have) mbar_wait(full_bar, round & 1u);      // bulk copy -> MMA: both async proxy, the mbarrier orders them
                        tr.ev(0xB00 + gi);
                        uint64_t bdesc0 = ring_desc + b_lbo;
                        const uint64_t adesc0 = adesc;
                        const uint32_t cur_empty = empty_bar;
                        int nk = left < per_tile ? left : per_tile;
                        const bool with_bias = first && G.kt0;         // this tile starts with the bias columns
                        if (with_bias) { nk = G.kt0 >> 4; }
                        first = false;
                        adesc += (uint64_t)nk * a_step;
                        next_slot();
                        left -= nk;
                        // peek at the NEXT tile before this tile's MMAs go out: a barrier test costs this thread ~100-150
                        // cycles even when the tile has long landed, and here they pass under the MMA issue
                        have = (left > 0 && peek) ? mbar_test(full_bar, round & 1u) : 0u;
                        if (elect_one()) {
                            if (with_bias) {
                                if (!no_mma) umma_bf16(tmem_d, ones_desc, bdesc0, idesc, 0u);
                                bdesc0 += (uint64_t)(nt * 2);          // past the bias columns (nt x 16 fp16 = nt * 2 sixteen-byte units)
                            }
                            // the common tile shapes with compile-time descriptor strides, anything else through the generic loop
                            if (no_mma) { }
                            else if (nt == 128 && xs) issue_tile<TC_X0_KSTEP, 256>(nk, tmem_d, adesc0, bdesc0, idesc);
                            else if (nt == 128) issue_tile<256, 256>(nk, tmem_d, adesc0, bdesc0, idesc);
                            else if (nt == 256 && !xs) issue_tile<256, 512>(nk, tmem_d, adesc0, bdesc0, idesc);
                            else {
                                uint64_t ad = adesc0, bd = bdesc0;
                                for (int kk = 0; kk < nk; ++kk) {
                                    umma_bf16_acc(tmem_d, ad, bd, idesc);
                                    ad += a_step;
                                    bd += b_step;
                                }
                            }
                            umma_commit(cur_empty);             // slot is free once these MMAs have read it
                        }
                    }
                    if (elect_one()) {
                        if (G.commit_bar >= 0) umma_commit(pb0 + 8 * G.commit_bar);
                        if (G.commit2_bar >= 0) umma_commit(pb0 + 8 * (G.commit2_bar + (G.commit2_bar == BAR_XFREE ? xb * TC_XBAR_STEP : 0u)));
                    }
                    __syncwarp();
                    tr.ev(0x400 + gi);
                }
                if (k0n >= 0) { pprev0 = pbits; pbits ^= a.prog[k0n].parity_mask; }
                if (k1n >= 0) { pprev1 = pbits1; pbits1 ^= a.prog[k1n].parity_mask; }
            }
        }
    } else if (warp > TC_EPI_WARPS && stg_warp < TC_STG_WARPS) {
        // ================================================================ staging: 8 warps, one item ahead
        // The staged rows of item i are dead once its last layer-0 MMAs have completed (BAR_XFREE); the rows of
        // item i + 1 are staged right then, under the MMAs and epilogues of the remaining layers.
        TcTrace tr; tr.init(a.trace, 3, blockIdx.x == 0 && stg_tid == 0);
        uint32_t cnt = 0;                            // items staged so far (the staging role has no phases: one item per pass)
        for (int j = -reh; j < n_local; ++j) {
            const bool dry = j < 0;
            const int w = (int)blockIdx.x + (dry ? 0 : j) * (int)gridDim.x;
            if (j == 0) { pf_grid_dep_wait(); pf_timeline_raw_if(stg_tid == 0 && blockIdx.x == 0, 0x52); }
            // buffer (and barrier set) of this item; it is free once the MMAs of the item that used it last have completed
            const uint32_t nb = (uint32_t)a.x0_nbuf, xb = nb == 2 ? (cnt & 1u) : 0u, use = nb == 2 ? (cnt >> 1) : cnt;
            if (use) mbar_wait(pb0 + 8 * (BAR_XFREE + xb * TC_XBAR_STEP), (use - 1u) & 1u);
            tr.ev(0x500);
            tc_stage_item<TC_STG_THREADS>(a, w, x0 + xb * a.x0_bytes, pb0 + 8 * (BAR_XFULL + xb * TC_XBAR_STEP), stg_tid, tr, dry);
            tr.ev(0x502);
            ++cnt;
        }
    } else if (warp < TC_EPI_WARPS) {
        // ================================================================ epilogues: 8 warps
        const int q = warp & 3, h = warp >> 2, row = q * 32 + lane;          // (h == 0 for every warp when TC_EPI_WARPS == 4)
        uint32_t pbits = 0;
        TcTrace tr; tr.init(a.trace, 0, blockIdx.x == 0 && tid == 0);
        const bool raw = (a.debug & 128) != 0;          // experiment, bit 7: convert only, no ELU
        // ELU + fp16 pack of 16 accumulator columns [g, g + 16) of this thread's row into a chunk buffer
        auto pack16 = [&](const uint32_t (&v)[16], unsigned char* outb, int g) {
#pragma unroll
            for (int j = 0; j < 16; j += 8) {
                uint32_t pk[4];
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    pk[i] = raw ? tc_pack_h2(__uint_as_float(v[j + 2 * i]), __uint_as_float(v[j + 2 * i + 1])) : tc_elu_h2(v[j + 2 * i], v[j + 2 * i + 1]);
                *reinterpret_cast<uint4*>(outb + (size_t)((g + j) >> 3) * (TC_ROWS * 16) + row * 16) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            }
        };
        uint32_t pbits1 = 0;
        for (int it = -reh; it < n_iter; ++it) {
            if (it == 0) pf_grid_dep_wait();
            const bool dry = it < 0;
            const int w0 = item_of(it, 0), w1 = item_of(it, 1);
            const int k0n = w0 >= 0 ? (w0 >= a.tiles ? 1 : 0) : -1, k1n = w1 >= 0 ? (w1 >= a.tiles ? 1 : 0) : -1;
            const TcProgram& PR = a.prog[k0n >= 0 ? k0n : k1n];
            const int nepis = PR.nepis;
            for (int ei = 0; ei < nepis; ++ei) {
                const int ph = PR.e[ei].phase;
                const int k = ph ? k1n : k0n;
                if (k < 0) continue;
                const TcProgram& P = a.prog[k];
                const TcNetArgs& N = a.net[k];
                const TcEpi& E = P.e[ei];
                const int tile = (ph ? w1 : w0) - k * a.tiles;
                const int row0 = tile * TC_ROWS;
                const int nrows = dry ? 0 : min(TC_ROWS, a.rows - row0);      // dry item: no row is written
                const int L = P.layers;
                const uint32_t bits = ph ? pbits1 : pbits;
                tr.ev(0x600 + ei);
                mbar_wait(pb0 + 8 * E.acc_bar, par_of(bits, E.acc_bar, E.acc_k));
                if (E.free_bar >= 0) mbar_wait(pb0 + 8 * E.free_bar, par_of(bits, E.free_bar, E.free_k));
                tc_fence_after();
                tr.ev(0x700 + ei);
                const int nt = E.nt, units = nt >> 4, u0 = TC_EPI_WARPS == 4 ? units : (units + 1) >> 1;   // 16-column units over the column halves
                const int c_lo = h == 0 ? 0 : u0 * 16, c_hi = h == 0 ? u0 * 16 : nt;
                const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16) + E.tmem_col;
                if (E.dst >= 0) {
                    unsigned char* outb = ach + E.dst * TC_ACHUNK_BYTES;
                    // two register sets of 16 columns: the TMEM load of the next 16 is in flight while the ELUs of the
                    // current 16 issue (tcgen05.ld is asynchronous up to its wait); c_lo, c_hi are multiples of 16
                    if (c_lo < c_hi && !(a.debug & 64)) {              // (experiment, bit 6: the step does no work at all)
                        uint32_t va[16], vb[16];
                        int g = c_lo;
                        tr.ev(0xC00 + ei);
                        tmem_ld16(trow + g, va);
                        tmem_ld_wait_on(va);
                        tr.ev(0xD00 + ei);
                        if (c_hi - c_lo == 64) {                 // the usual share (half of a 128-column chunk): four rounds, straight-line
                            tmem_ld16(trow + g + 16, vb);
                            pack16(va, outb, g);
                            tmem_ld_wait_on(vb);
                            tmem_ld16(trow + g + 32, va);
                            pack16(vb, outb, g + 16);
                            tmem_ld_wait_on(va);
                            tmem_ld16(trow + g + 48, vb);
                            pack16(va, outb, g + 32);
                            tmem_ld_wait_on(vb);
                            pack16(vb, outb, g + 48);
                        } else for (;;) {
                            const int g2 = g + 16;
                            if (g2 < c_hi) tmem_ld16(trow + g2, vb);
                            pack16(va, outb, g);
                            if (g2 >= c_hi) break;
                            tmem_ld_wait_on(vb);
                            g = g2 + 16;
                            if (g < c_hi) tmem_ld16(trow + g, va);
                            pack16(vb, outb, g2);
                            if (g >= c_hi) break;
                            tmem_ld_wait_on(va);
                        }
                    }
                    tr.ev(0x800 + ei);
                    fence_async_smem();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(pb0 + 8 * E.done_bar);
                } else {
                    for (int g = c_lo; g < c_hi; g += 32) {
                        uint32_t v[32];
                        const int wd = min(32, c_hi - g);
                        if (wd == 32) tmem_ld32(trow + g, v); else tmem_ld16(trow + g, v);
                        tmem_ld_wait();
                        if (row >= nrows) continue;
                        const int no = P.dims[L], n0 = E.n0 + g;
                        const long long er = (long long)row0 + row;
                        if (N.mode == TC_OUT_PLAIN) {
                            float o32[32];
#pragma unroll
                            for (int j = 0; j < 32; ++j) o32[j] = __uint_as_float(v[j]);
                            float* dst = N.out + er * no + n0;
                            const int cnt = min(wd, no - n0);
#pragma unroll 1
                            for (int j = 0; j < cnt; ++j) dst[j] = o32[j];       // rolled on purpose (code size)
                        } else if (N.mode == TC_OUT_VALUE) {
                            if (n0 == 0) {
                                const float val = __uint_as_float(v[0]);
                                if (a.values_out) a.values_out[er] = val;
                                if (a.has_storage) a.S.values[a.storage_row0 + er] = val;
                            }
                        } else if (n0 == 0) {
                            float m16[16];
#pragma unroll
                            for (int j = 0; j < 16; ++j) m16[j] = __uint_as_float(v[j]);
                            tc_sample_row(a, er, m16);
                        }
                    }
                    tr.ev(0x800 + ei);
                    if (E.done_bar >= 0) {              // two-phase programs: the accumulator has been read, its columns are free
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(pb0 + 8 * E.done_bar);
                    }
                }
                tr.ev(0x900 + ei);
            }
            tc_fence_before();
            if (k0n >= 0) pbits ^= a.prog[k0n].parity_mask;
            if (k1n >= 0) pbits1 ^= a.prog[k1n].parity_mask;
        }
    }
    tc_fence_before();
    __syncthreads();
    pf_timeline(0x53);
    if (warp == TC_MMA_WARP) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)a.tmem_cols) : "memory");
    }
}

// the two instantiations: 8 + 8 worker warps, one CTA per SM, all of TMEM; and the compact one (4 + 4 warps, two CTAs per SM)
static const auto pf_mlp_tc_kernel_std = pf_mlp_tc_kernel<8, 8, 1>;
static const auto pf_mlp_tc_kernel_compact = pf_mlp_tc_kernel<4, 4, 2>;

// ------------------------------------------------------------------------------------------------ packing
// fp32 nn.Linear weight [out, in] + bias -> fp16 tiles in consumption order.  Tiles of a group, contiguous: the BIAS
// tile if the group starts an accumulation (nt x 16: column 0 = fp16(b), column 1 = fp16(b - fp16(b)), multiplied by
// the "ones" A operand), then the K tiles of kt columns.  blockIdx.y = group (both nets in one launch).
struct TcPackArgs { TcProgram prog[2]; unsigned char* blob[2]; long long w_off[2][PF_MAX_MLP_LAYERS], b_off[2][PF_MAX_MLP_LAYERS]; };

__global__ void pf_tc_pack_kernel(const __grid_constant__ TcPackArgs a, const float* __restrict__ params) {
    int gi = blockIdx.y, k = 0;
    if (gi >= a.prog[0].ngroups) { gi -= a.prog[0].ngroups; k = 1; }
    const TcProgram& P = a.prog[k];
    const TcGroup& G = P.g[gi];
    const int l = G.layer, nt = G.nt, kc = G.kc;
    const int K = P.dims[l], N = P.dims[l + 1];
    const float* W = params + a.w_off[k][l];
    const float* b = params + a.b_off[k][l];
    unsigned char* blob = a.blob[k] + G.blob_off;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nt * kc; i += gridDim.x * blockDim.x) {
        const int nl = i / kc, kl = i - nl * kc;
        const int n = G.n_base + nl, kg = G.k_base + kl;
        // K-major core matrices over the WHOLE group: any K range is contiguous, so the tile size (kt) is only the
        // loader's chunking -- the standard and the compact programs read the same blob with different tiles
        const long long off = (G.bias ? nt * 32 : 0) + (long long)(kl >> 3) * (nt * 16) + nl * 16 + (kl & 7) * 2;
        const float v = (n < N && kg < K) ? W[(long long)n * K + kg] : 0.0f;
        *reinterpret_cast<unsigned short*>(blob + off) = tc_h(v);
        if (kl == 0 && G.bias) {   // bias tile: columns 0 / 1 = high / low fp16 parts of the bias, the rest stays zero
            const float bv = n < N ? b[n] : 0.0f;
            const unsigned short hi = tc_h(bv);
            const unsigned short lo = tc_h(bv - __half2float(__ushort_as_half(hi)));
            unsigned short* d = reinterpret_cast<unsigned short*>(blob + (long long)(nl >> 3) * 128 + (nl & 7) * 16);
            d[0] = hi; d[1] = lo;
        }
    }
}

static int tc_fail(int code, const char* msg) { pf_set_error(msg); return code; }

// K-columns that fit a slot beside the bias tile (0: the bias travels alone -- nothing would be saved)
static int pick_kt0(int nt, int kc, int kt, int slot_bytes, int bias) {
    if (!bias || getenv("PF_TC_NO_KT0")) return 0;          // (experiment: the bias always in its own tile)
    int k0 = ((slot_bytes - nt * 32) / 2 / nt) & ~15;
    if (k0 > 128) k0 = 128;
    if (k0 > kc) k0 = kc;
    if (k0 < 16) return 0;
    const int merged = 1 + (kc - k0 + kt - 1) / kt, separate = 1 + (kc + kt - 1) / kt;
    return merged < separate ? k0 : 0;
}
static int pick_kt(int nt, int kc, int slot_bytes) {
    int kt = (slot_bytes / 2 / nt) & ~15;
    if (kt > 128) kt = 128;
    return kt < kc ? kt : kc;
}

// Last pass over a built program: blob offsets in CANONICAL order (layer, output column, K origin) -- whatever order and
// tile size a program issues its groups in, all programs of a net read the same packed blob -- plus the per-item barrier
// completion parities and what arrives on each barrier.
static void finalize_program(TcProgram& P) {
    if (P.phases == 0) { P.phases = 1; P.nbuf = 2; }
    int order[TC_MAX_GROUPS];
    for (int i = 0; i < P.ngroups; ++i) order[i] = i;
    auto key = [&](int i) { const TcGroup& g = P.g[i]; return ((long long)g.layer << 40) | ((long long)g.n_base << 20) | g.k_base; };
    for (int i = 1; i < P.ngroups; ++i) { const int x = order[i]; int j = i - 1; while (j >= 0 && key(order[j]) > key(x)) { order[j + 1] = order[j]; --j; } order[j + 1] = x; }
    long long off = 0;
    for (int i = 0; i < P.ngroups; ++i) {
        TcGroup& g = P.g[order[i]];
        g.blob_off = (uint32_t)off;
        off += (long long)g.nt * ((g.bias ? 16 : 0) + g.kc) * 2;
    }
    P.blob_bytes = off;
    int cnt[TC_NBARS] = {0};
    for (int i = 0; i < P.ngroups; ++i) { if (P.g[i].commit_bar >= 0) cnt[P.g[i].commit_bar]++; if (P.g[i].commit2_bar >= 0) cnt[P.g[i].commit2_bar]++; }
    for (int i = 0; i < P.nepis; ++i) if (P.e[i].done_bar >= 0) cnt[P.e[i].done_bar]++;
    cnt[BAR_XFULL]++;
    P.parity_mask = 0;
    for (int b = 0; b < TC_NBARS; ++b) if (cnt[b] & 1) P.parity_mask |= 1u << b;
    for (int b = 0; b < TC_NBARS; ++b) P.bar_kind[b] = TC_BK_COMMIT;
    for (int i = 0; i < P.nepis; ++i) if (P.e[i].done_bar >= 0) P.bar_kind[P.e[i].done_bar] = TC_BK_EPI;
    P.bar_kind[BAR_XFULL] = TC_BK_STG;
    if (P.x0_nbuf == 0) P.x0_nbuf = 1;
    if (P.x0_nbuf == 2) P.bar_kind[BAR_XFULL1] = TC_BK_STG;
}

// Builds the static schedule of one net (see the header comment).  Returns non-zero if the shape does not fit.
static int build_program(TcProgram& P, const PfNet& n, int odd_col = 256, int slot_bytes = TC_SLOT_BYTES) {
    memset(&P, 0, sizeof(P));
    const int L = n.layers;
    P.layers = L;
    P.odd_col = odd_col;
    for (int i = 0; i <= L; ++i) { P.dims[i] = n.dims[i]; P.pdims[i] = pad16(n.dims[i]); }
    if (P.pdims[0] > 256) return 1;                                   // staged input rows: <= 64 KB
    for (int l = 1; l < L; ++l) if (P.pdims[l + 1] > 256) return 1;    // accumulators of layers >= 1 live whole in half of TMEM
    if (odd_col == 128) {          // compact: layer 0 in ONE chunk at columns [0, 128), every other accumulator <= 128 wide
        if (P.pdims[1] > TC_CHUNK) return 1;
        for (int l = 1; l < L; ++l) if (P.pdims[l + 1] > 128) return 1;
    }
    const int nch0 = (P.pdims[1] + TC_CHUNK - 1) / TC_CHUNK;
    if (nch0 > 8 || (L == 1 && nch0 > 2)) return 1;
    P.x0_bytes = (((P.pdims[0] >> 3) * TC_X0_LBO) + 127) & ~127;
    int compl_[TC_NBARS] = {0};
    int last_free_k[2] = {-1, -1};
    int ach_next = 0;
    auto add_group = [&](TcGroup g) -> int {
        if (P.ngroups == TC_MAX_GROUPS) return 1;
        g.kt = (uint16_t)pick_kt(g.nt, g.kc, slot_bytes);
        g.kt0 = (uint16_t)pick_kt0(g.nt, g.kc, g.kt, slot_bytes, g.bias);
        P.g[P.ngroups++] = g;
        return 0;
    };
    auto add_epi = [&](TcEpi e) -> int {
        if (P.nepis == TC_MAX_EPIS) return 1;
        P.e[P.nepis++] = e;
        return 0;
    };
    auto region_col = [odd_col](int l) { return (l & 1) ? odd_col : 0; };
    auto region_bar = [](int l) { return (l & 1) ? BAR_ACC2 : BAR_ACC0; };

    // ---- layer 0: chunk c accumulates in TMEM slot c & 1 from the staged input ----
    int acc_k0[8] = {0}, e_k0[8] = {0}, e_b0[8] = {0};
    auto layer0_group = [&](int c) -> int {
        TcGroup g = {};
        g.layer = 0; g.n_base = (uint16_t)(c * TC_CHUNK); g.k_base = 0;
        g.nt = (uint16_t)((P.pdims[1] - c * TC_CHUNK) < TC_CHUNK ? (P.pdims[1] - c * TC_CHUNK) : TC_CHUNK);
        g.kc = (uint16_t)P.pdims[0]; g.a_src = 0; g.bias = 1; g.tmem_col = (uint16_t)((c & 1) * TC_CHUNK);
        g.wait_bar = -1; g.wait_k = 0; g.prev_bar = -1;
        // The accumulator slot may still hold what the PREVIOUS item's final epilogue reads: with an odd number of layers
        // the last layer accumulates at column 0 (chunk 0's slot), and a single-layer net's chunks ARE the final outputs.
        // (Even layer counts -- the three-hidden-layer nets of the reference -- end in the other half of TMEM.)
        if (L == 1) { g.prev_bar = BAR_ODONE; g.prev_k = (int8_t)c; }
        else if ((L & 1) && c == 0) { g.prev_bar = BAR_ODONE; g.prev_k = (int8_t)((P.pdims[L] + TC_CHUNK - 1) / TC_CHUNK - 1); }
        if (c == 0) { g.wait_bar = BAR_XFULL; g.wait_k = 0; }
        else if (c >= 2 && L > 1) { g.wait_bar = (int8_t)(BAR_E0 + e_b0[c - 2]); g.wait_k = (int8_t)e_k0[c - 2]; }   // slot drained
        g.commit_bar = (int8_t)(BAR_ACC0 + (c & 1));
        g.commit2_bar = (int8_t)(c == nch0 - 1 ? BAR_XFREE : -1);    // last reader of the staged rows
        acc_k0[c] = compl_[BAR_ACC0 + (c & 1)]++;
        return add_group(g);
    };
    // consumer of one produced chunk: K-slice j of layer l reading chunk buffer b
    auto consumer_group = [&](int l, int j, int width, int b, int e_k, bool last) -> int {
        TcGroup g = {};
        g.layer = (int8_t)l; g.n_base = 0; g.k_base = (uint16_t)(j * TC_CHUNK);
        g.nt = (uint16_t)P.pdims[l + 1]; g.kc = (uint16_t)width; g.a_src = (int8_t)(1 + b); g.bias = j == 0;
        g.tmem_col = (uint16_t)region_col(l);
        g.wait_bar = (int8_t)(BAR_E0 + b); g.wait_k = (int8_t)e_k; g.prev_bar = -1;
        g.commit_bar = (int8_t)(BAR_AFREE0 + b);
        last_free_k[b] = compl_[BAR_AFREE0 + b]++;
        g.commit2_bar = -1;
        if (last) { g.commit2_bar = (int8_t)region_bar(l); }
        return add_group(g);
    };
    if (layer0_group(0)) return 1;
    if (nch0 > 1 && layer0_group(1)) return 1;
    int acc_k_prev = 0;                         // completion index of the accumulator the next epilogues read (layers >= 1)
    for (int c = 0; c < nch0; ++c) {
        TcEpi e = {};
        const int nt = (P.pdims[1] - c * TC_CHUNK) < TC_CHUNK ? (P.pdims[1] - c * TC_CHUNK) : TC_CHUNK;
        e.tmem_col = (uint16_t)((c & 1) * TC_CHUNK); e.nt = (uint16_t)nt; e.n0 = (uint16_t)(c * TC_CHUNK);
        e.acc_bar = (int8_t)(BAR_ACC0 + (c & 1)); e.acc_k = (int8_t)acc_k0[c];
        e.dst = -1; e.free_bar = -1; e.free_k = 0; e.done_bar = -1;
        if (L > 1) {
            const int b = ach_next; ach_next ^= 1;
            e.dst = (int8_t)b; e.done_bar = (int8_t)(BAR_E0 + b);
            if (last_free_k[b] >= 0) { e.free_bar = (int8_t)(BAR_AFREE0 + b); e.free_k = (int8_t)last_free_k[b]; }
            e_b0[c] = b; e_k0[c] = compl_[BAR_E0 + b]++;
            if (add_epi(e)) return 1;
            if (consumer_group(1, c, nt, b, e_k0[c], c == nch0 - 1)) return 1;
            if (c == nch0 - 1) acc_k_prev = compl_[region_bar(1)]++;
        } else { e.done_bar = BAR_ODONE; if (add_epi(e)) return 1; }      // final output: tells the next item its accumulator is read
        if (c + 2 < nch0 && layer0_group(c + 2)) return 1;
    }
    // ---- layers >= 1: whole accumulator, epilogue in chunks, each chunk feeds layer l + 1 at once ----
    for (int l = 1; l < L; ++l) {
        const int Np = P.pdims[l + 1];
        const int nch = (Np + TC_CHUNK - 1) / TC_CHUNK;
        const int acc_k = acc_k_prev;
        for (int j = 0; j < nch; ++j) {
            TcEpi e = {};
            const int nt = (Np - j * TC_CHUNK) < TC_CHUNK ? (Np - j * TC_CHUNK) : TC_CHUNK;
            e.tmem_col = (uint16_t)(region_col(l) + j * TC_CHUNK); e.nt = (uint16_t)nt; e.n0 = (uint16_t)(j * TC_CHUNK);
            e.acc_bar = (int8_t)region_bar(l); e.acc_k = (int8_t)acc_k;
            e.dst = -1; e.free_bar = -1; e.free_k = 0; e.done_bar = -1;
            if (l < L - 1) {
                const int b = ach_next; ach_next ^= 1;
                e.dst = (int8_t)b; e.done_bar = (int8_t)(BAR_E0 + b);
                if (last_free_k[b] >= 0) { e.free_bar = (int8_t)(BAR_AFREE0 + b); e.free_k = (int8_t)last_free_k[b]; }
                const int e_k = compl_[BAR_E0 + b]++;
                if (add_epi(e)) return 1;
                if (consumer_group(l + 1, j, nt, b, e_k, j == nch - 1)) return 1;
                if (j == nch - 1) acc_k_prev = compl_[region_bar(l + 1)]++;
            } else { e.done_bar = BAR_ODONE; if (add_epi(e)) return 1; }
        }
    }
    (void)compl_;
    finalize_program(P);
    return 0;
}

// The same networks as a TWO-PHASE program (see the kernel's iteration loop): three hidden layers, the rsl_rl default.
//   phase 0 of an item: layer 0 in chunks of 128 columns through ONE accumulator slot S = TMEM [0, 128) and ONE chunk
//                       buffer (0), every chunk fed at once to layer 1, whose accumulator is TMEM [256, 512);
//   phase 1:            layer 1's epilogue (1-2 chunks) -> layer 2 -> its epilogue -> output layer, accumulators in
//                       T = TMEM [128, 256); activation chunks through buffer 1.
// Phase 1 of item m runs interleaved with phase 0 of item m + 1: the epilogue warps alternate between the two items'
// steps (so a chunk buffer's reader has a whole foreign step to finish in before its next writer asks), and the MMA
// thread issues every group right after the epilogue step it depends on, in the merged order.  The staged input rows
// are DOUBLE-buffered (the shared memory a third chunk buffer would take): they are held until the item's last layer-0
// chunk, most of an iteration, and with one buffer the next item's staging pass sat on the critical loop.  The hazards
// that program order does not already cover are explicit waits (comments below); check_program() replays the schedule.
static int build_program_pipe(TcProgram& P, const PfNet& n, int slot_bytes, int x0_nbuf = 2) {
    memset(&P, 0, sizeof(P));
    const int L = n.layers;
    if (L != 4) return 1;
    P.layers = L; P.odd_col = 256; P.phases = 2; P.nbuf = 2; P.x0_nbuf = x0_nbuf;
    for (int i = 0; i <= L; ++i) { P.dims[i] = n.dims[i]; P.pdims[i] = pad16(n.dims[i]); }
    if (P.pdims[0] > 256 || P.pdims[2] > 256 || P.pdims[3] > 128 || P.pdims[4] > 128) return 1;
    const int nch0 = (P.pdims[1] + TC_CHUNK - 1) / TC_CHUNK, nch1 = (P.pdims[2] + TC_CHUNK - 1) / TC_CHUNK;
    if (nch0 < 2 || nch0 > 8) return 1;
    P.x0_bytes = (((P.pdims[0] >> 3) * TC_X0_LBO) + 127) & ~127;
    auto width = [&](int l, int c) { const int w = P.pdims[l] - c * TC_CHUNK; return w < TC_CHUNK ? w : TC_CHUNK; };   // chunk c of the layer whose OUTPUT is pdims[l]
    const int S = 0, T = 128, A1 = 256;

    // ---- the epilogue steps of the two phases, then merged A, B, A, B ...
    struct Ep { TcEpi e; int pos; };
    Ep ep0[8], ep1[4]; int n0 = 0, n1 = 0;
    for (int c = 0; c < nch0; ++c) {                       // phase 0: chunk c of layer 0: S -> buffer 0
        TcEpi e = {};
        e.phase = 0; e.tmem_col = S; e.nt = (uint16_t)width(1, c); e.n0 = (uint16_t)(c * TC_CHUNK);
        e.acc_bar = BAR_ACC0; e.acc_k = (int8_t)c;
        e.dst = 0; e.done_bar = BAR_E0;
        e.free_bar = -1;
        // buffer 0 is read by layer-1 slice c - 1 of this item (AFREE0); the last slice of the PREVIOUS item completed with
        // its layer-1 accumulator, which phase 1's first step (earlier in the merged order) has waited for
        if (c >= 1) { e.free_bar = BAR_AFREE0; e.free_k = (int8_t)(c - 1); }
        ep0[n0++].e = e;
    }
    for (int j = 0; j < nch1; ++j) {                       // phase 1: chunk j of layer 1: A1 -> buffer 1
        TcEpi e = {};
        e.phase = 1; e.tmem_col = (uint16_t)(A1 + j * TC_CHUNK); e.nt = (uint16_t)width(2, j); e.n0 = (uint16_t)(j * TC_CHUNK);
        e.acc_bar = BAR_ACC1; e.acc_k = 0;
        e.dst = 1; e.done_bar = BAR_E1;
        e.free_bar = -1;
        if (j >= 1) { e.free_bar = BAR_AFREE1; e.free_k = (int8_t)(j - 1); }     // layer-2 slice j - 1 has read the buffer
        ep1[n1++].e = e;                                   // (j = 0: the previous item's output MMAs, which its final epilogue waited for)
    }
    {                                                      // layer 2: T -> buffer 1 (its last reader, the last layer-2 slice, completed with ACC2)
        TcEpi e = {};
        e.phase = 1; e.tmem_col = T; e.nt = (uint16_t)P.pdims[3]; e.n0 = 0;
        e.acc_bar = BAR_ACC2; e.acc_k = 0; e.dst = 1; e.done_bar = BAR_E1; e.free_bar = -1;
        ep1[n1++].e = e;
    }
    {                                                      // output layer: T -> global memory; ODONE: T is free for the next item
        TcEpi e = {};
        e.phase = 1; e.tmem_col = T; e.nt = (uint16_t)P.pdims[4]; e.n0 = 0;
        e.acc_bar = BAR_ACC3; e.acc_k = 0; e.dst = -1; e.done_bar = BAR_ODONE; e.free_bar = -1;
        ep1[n1++].e = e;
    }
    int pos = 0, i0 = 0, i1 = 0;
    while (i0 < n0 || i1 < n1) {
        if (i1 < n1) { ep1[i1].pos = pos++; P.e[P.nepis++] = ep1[i1++].e; }
        if (i0 < n0) { ep0[i0].pos = pos++; P.e[P.nepis++] = ep0[i0++].e; }
    }

    // ---- MMA groups, each placed after the last epilogue step it depends on (stable in the order created below)
    struct Gr { TcGroup g; int pos; int seq; };
    Gr gr[TC_MAX_GROUPS]; int ng = 0;
    auto add = [&](TcGroup g, int at) { g.kt = (uint16_t)pick_kt(g.nt, g.kc, slot_bytes); g.kt0 = (uint16_t)pick_kt0(g.nt, g.kc, g.kt, slot_bytes, g.bias); gr[ng].g = g; gr[ng].pos = at; gr[ng].seq = ng; ++ng; };
    auto l0_group = [&](int c) {
        TcGroup g = {};
        g.phase = 0; g.layer = 0; g.n_base = (uint16_t)(c * TC_CHUNK); g.k_base = 0; g.nt = (uint16_t)width(1, c);
        g.kc = (uint16_t)P.pdims[0]; g.a_src = 0; g.bias = 1; g.tmem_col = S;
        if (c == 0) { g.wait_bar = BAR_XFULL; g.wait_k = 0; }     // S: drained by the previous item's last chunk, which its last layer-1 slice waited for
        else { g.wait_bar = BAR_E0; g.wait_k = (int8_t)(c - 1); }   // S drained by chunk c - 1
        g.commit_bar = BAR_ACC0; g.commit2_bar = (int8_t)(c == nch0 - 1 ? BAR_XFREE : -1); g.prev_bar = -1;
        return g;
    };
    const int pos_l1_drained = ep1[nch1 - 1].pos;          // the previous item's last layer-1 chunk has left A1
    add(l0_group(0), -1);
    for (int j = 0; j < nch1; ++j) {                       // layer 2, slice j (reads buffer 1)
        TcGroup g = {};
        g.phase = 1; g.layer = 2; g.n_base = 0; g.k_base = (uint16_t)(j * TC_CHUNK); g.nt = (uint16_t)P.pdims[3];
        g.kc = (uint16_t)width(2, j); g.a_src = 2; g.bias = j == 0; g.tmem_col = T;
        g.wait_bar = BAR_E1; g.wait_k = (int8_t)j;
        g.commit_bar = (int8_t)(j < nch1 - 1 ? BAR_AFREE1 : BAR_ACC2); g.commit2_bar = -1;
        // T still holds the previous item's output accumulator until its final epilogue has read it
        g.prev_bar = (int8_t)(j == 0 ? BAR_ODONE : -1); g.prev_k = 0;
        add(g, ep1[j].pos);
    }
    for (int c = 0; c < nch0; ++c) {
        if (c + 1 < nch0) add(l0_group(c + 1), ep0[c].pos);
        TcGroup g = {};                                    // layer 1, slice c (reads buffer 0)
        g.phase = 0; g.layer = 1; g.n_base = 0; g.k_base = (uint16_t)(c * TC_CHUNK); g.nt = (uint16_t)P.pdims[2];
        g.kc = (uint16_t)width(1, c); g.a_src = 1; g.bias = c == 0; g.tmem_col = A1;
        g.wait_bar = BAR_E0; g.wait_k = (int8_t)c;
        g.commit_bar = BAR_AFREE0; g.commit2_bar = (int8_t)(c == nch0 - 1 ? BAR_ACC1 : -1); g.prev_bar = -1;
        // slice 0 overwrites A1: it goes after the phase-1 group that waited for the previous item's last layer-1 chunk
        add(g, c == 0 && pos_l1_drained > ep0[0].pos ? pos_l1_drained : ep0[c].pos);
    }
    {                                                      // output layer (reads buffer 1, written by the layer-2 epilogue)
        TcGroup g = {};
        g.phase = 1; g.layer = 3; g.n_base = 0; g.k_base = 0; g.nt = (uint16_t)P.pdims[4];
        g.kc = (uint16_t)P.pdims[3]; g.a_src = 2; g.bias = 1; g.tmem_col = T;
        g.wait_bar = BAR_E1; g.wait_k = (int8_t)nch1;
        g.commit_bar = BAR_ACC3; g.commit2_bar = -1; g.prev_bar = -1;
        add(g, ep1[nch1].pos);
    }
    // stable sort by (pos, phase 1 first, creation order)
    for (int i = 1; i < ng; ++i) {
        Gr x = gr[i]; int j = i - 1;
        auto before = [](const Gr& u, const Gr& v) { return u.pos != v.pos ? u.pos < v.pos : (u.g.phase != v.g.phase ? u.g.phase > v.g.phase : u.seq < v.seq); };
        while (j >= 0 && before(x, gr[j])) { gr[j + 1] = gr[j]; --j; }
        gr[j + 1] = x;
    }
    for (int i = 0; i < ng; ++i) P.g[P.ngroups++] = gr[i].g;
    finalize_program(P);
    return 0;
}

// Dry run of the schedule on the host: staging, MMA issuer and epilogue as three in-order agents that only advance when
// the barrier completion they wait for has happened, over three items (four iterations of a two-phase program: prologue,
// two steady iterations, drain).  A schedule that cannot finish, or in which a barrier runs more than one completion ahead
// of a waiter (the parity wait could not tell), is rejected at create time instead of trapping on the device.
static int check_program(const TcProgram& P) {
    const int NI = 3, n_iter = NI + P.phases - 1;
    int per[TC_NBARS] = {0};
    for (int i = 0; i < P.ngroups; ++i) { if (P.g[i].commit_bar >= 0) per[P.g[i].commit_bar]++; if (P.g[i].commit2_bar >= 0) per[P.g[i].commit2_bar]++; }
    for (int i = 0; i < P.nepis; ++i) if (P.e[i].done_bar >= 0) per[P.e[i].done_bar]++;
    per[BAR_XFULL] = 1;
    for (int b = 0; b < TC_NBARS; ++b) if ((per[b] & 1) != (int)((P.parity_mask >> b) & 1u)) return 1;
    int done[TC_NBARS] = {0};
    int s_item = 0, g_it = 0, g_i = 0, e_it = 0, e_i = 0;
    bool bad = false;
    const int nb = P.x0_nbuf;
    auto reached = [&](int b, int k, int m) {        // completion k of item m on barrier b
        const int target = m * per[b] + k;
        if (b == BAR_XFULL && nb == 2) return done[b] > target;
        if (done[b] > target + 1) { bad = true; if (getenv("PF_TC_CHECK_DEBUG")) fprintf(stderr, "check_program: barrier %d at %d, waiter wants %d (item %d k %d)\n", b, done[b], target, m, k); }
        return done[b] > target;
    };
    bool progress = true;
    while (progress && !bad) {
        progress = false;
        // staging: item s goes to buffer s % nb, free once item s - nb's layer-0 MMAs have completed (one logical XFULL /
        // XFREE counter here; the device uses a barrier pair per buffer, so "two ahead" cannot alias there)
        while (s_item < NI && (s_item < nb || done[BAR_XFREE] > s_item - nb)) { done[BAR_XFULL]++; ++s_item; progress = true; }
        while (g_it < n_iter) {
            if (g_i == P.ngroups) { g_i = 0; ++g_it; progress = true; continue; }
            const TcGroup& G = P.g[g_i];
            const int m = g_it - G.phase;
            if (m >= 0 && m < NI) {
                if (G.wait_bar >= 0 && !reached(G.wait_bar, G.wait_k, m)) break;
                if (G.prev_bar >= 0 && m >= 1 && !reached(G.prev_bar, G.prev_k, m - 1)) break;
                if (G.commit_bar >= 0) done[G.commit_bar]++;
                if (G.commit2_bar >= 0) done[G.commit2_bar]++;
            }
            ++g_i; progress = true;
        }
        while (e_it < n_iter) {
            if (e_i == P.nepis) { e_i = 0; ++e_it; progress = true; continue; }
            const TcEpi& E = P.e[e_i];
            const int m = e_it - E.phase;
            if (m >= 0 && m < NI) {
                if (!reached(E.acc_bar, E.acc_k, m)) break;
                if (E.free_bar >= 0 && !reached(E.free_bar, E.free_k, m)) break;
                if (E.done_bar >= 0) done[E.done_bar]++;
            }
            ++e_i; progress = true;
        }
    }
    if (getenv("PF_TC_CHECK_DEBUG")) fprintf(stderr, "check_program: bad=%d staged=%d mma it=%d step=%d epi it=%d step=%d\n", (int)bad, s_item, g_it, g_i, e_it, e_i);
    if (bad || s_item < NI || g_it < n_iter || e_it < n_iter) return 1;
    for (int b = 0; b < TC_NBARS; ++b) if (done[b] != NI * per[b]) return 1;
    return 0;
}

// Diagnostic: device buffer of 3 * TC_TRACE_CAP uint64 that CTA 0 of every following launch fills with its role
// timelines (tools/tc_trace.py); NULL switches it off.
// Opt-in: the fused act / MLP kernel runs one dry item before its dependency wait (see the kernel).  Worth it when the
// kernel is launched right behind a kernel that still runs for a few microseconds (a CUDA-graphed env step + act
// chain); without such a predecessor it adds one item's latency, hence off by default.
extern "C" int pf_rl_tc_rehearse(pf_rl_ctx* c, int on) {
    if (c && c->wm) return 0;                    // the small-net path has no chain to rehearse
    if (!c || !c->tc) return 1;
    static_cast<TcState*>(c->tc)->rehearse = on ? 1 : 0;
    return 0;
}

extern "C" int pf_rl_debug_timeline(unsigned long long* device_log) {
    return pf_timeline_install(device_log) == cudaSuccess ? 0 : 1;
}

extern "C" int pf_rl_tc_trace(pf_rl_ctx* c, unsigned long long* device_buf) {
    if (!c || !c->tc) return 1;
    static_cast<TcState*>(c->tc)->trace = device_buf;
    return 0;
}

// Diagnostic (no device needed): builds + dry-runs the schedule for an MLP shape and prints it.  Returns 0 if the
// fused kernel supports the shape, 6 if not.
extern "C" int pf_rl_tc_plan(const int32_t* dims, int32_t layers, char* text, int32_t cap) {
    if (!dims || layers < 1 || layers > PF_MAX_MLP_LAYERS) return 1;
    PfNet n = {};
    n.layers = layers;
    for (int i = 0; i <= layers; ++i) n.dims[i] = dims[i];
    TcProgram P;
    if (getenv("PF_TC_PLAN_PIPE")) {                 // the two-phase program of the same net (diagnostic)
        if (build_program_pipe(P, n, 32768) || check_program(P)) return 6;
    } else if (build_program(P, n) || check_program(P)) return 6;
    int w = 0;
    auto put = [&](const char* fmt, auto... v) { if (text && w < cap) { int r = snprintf(text + w, (size_t)(cap - w), fmt, v...); if (r > 0) w += r; } };
    put("groups=%d epilogues=%d blob_bytes=%lld x0_bytes=%d parity_mask=0x%x\n", P.ngroups, P.nepis, P.blob_bytes, P.x0_bytes, P.parity_mask);
    for (int i = 0; i < P.ngroups; ++i) {
        const TcGroup& G = P.g[i];
        put("G%d ph=%d layer=%d a_src=%d nt=%d kc=%d kt=%d col=%d bias=%d kt0=%d wait=%d@%d prev=%d commit=%d,%d off=%u\n", i, G.phase, G.layer, G.a_src, G.nt, G.kc, G.kt,
            G.tmem_col, G.bias, G.kt0, G.wait_bar, G.wait_k, G.prev_bar, G.commit_bar, G.commit2_bar, G.blob_off);
    }
    for (int i = 0; i < P.nepis; ++i) {
        const TcEpi& E = P.e[i];
        put("E%d ph=%d col=%d nt=%d n0=%d acc=%d@%d dst=%d free=%d@%d done=%d\n", i, E.phase, E.tmem_col, E.nt, E.n0, E.acc_bar, E.acc_k, E.dst, E.free_bar, E.free_k, E.done_bar);
    }
    return 0;
}

// budget_kb: 227 (one CTA per SM), or 113 for the compact variant (two CTAs per SM: 2 x (113 + 1 KB reserved) = 228 KB)
static int smem_layout(int x0_bytes, int* stages, int* smem_bytes, int budget_kb = 227, int min_stages = 2, int slot_bytes = TC_SLOT_BYTES, int nbuf = 2, int x0_nbuf = 1) {
    const int fixed = x0_nbuf * x0_bytes + nbuf * TC_ACHUNK_BYTES + TC_ROWS * 32 + 512;
    int S = (budget_kb * 1024 - fixed) / slot_bytes;
    if (S < min_stages) return 1;
    *stages = S > TC_MAX_STAGES ? TC_MAX_STAGES : S;
    *smem_bytes = fixed + *stages * slot_bytes;
    return 0;
}

int pf_tc_create(pf_rl_ctx* c) {
    TcState* t = new TcState();
    { const char* e = getenv("PF_TC_SLOT_KB"); t->slot_bytes = e ? atoi(e) * 1024 : TC_SLOT_BYTES; }       // experiments
    for (int k = 0; k < 2; ++k) {
        if (build_program(t->prog[k], c->net[k], 256, t->slot_bytes) || check_program(t->prog[k])) {
            delete t;
            return tc_fail(6, "pf_rl_create: MLP shape not supported by the fused tensor-core kernel (input <= 256 wide, "
                              "first hidden layer <= 1024, later layers <= 256)");
        }
    }
    int max_smem = 0;
    for (int v = 0; v < 3; ++v) {
        const int xa = t->prog[0].x0_bytes, xc = t->prog[1].x0_bytes;
        t->x0_bytes[v] = v == 0 ? xa : v == 1 ? xc : (xa > xc ? xa : xc);
        if (smem_layout(t->x0_bytes[v], &t->stages[v], &t->smem_bytes[v], 227, 2, t->slot_bytes)) { delete t; return tc_fail(6, "pf_rl_create: shared memory layout does not fit"); }
        max_smem = t->smem_bytes[v] > max_smem ? t->smem_bytes[v] : max_smem;
    }
    // compact variant (two CTAs per SM): both nets' accumulators within 256 TMEM columns, shared memory within 113 KB.  Same
    // groups, tiles and blob offsets as the standard programs (only the accumulator columns differ): the packed weights are shared.
    t->compact = getenv("PF_TC_NO_COMPACT") == nullptr;
    for (int k = 0; k < 2 && t->compact; ++k)
        if (build_program(t->progc[k], c->net[k], 128, TC_SLOT_BYTES_COMPACT) || check_program(t->progc[k]) || t->progc[k].blob_bytes != t->prog[k].blob_bytes ||
            t->progc[k].ngroups != t->prog[k].ngroups)
            t->compact = 0;
    for (int v = 0; v < 3 && t->compact; ++v)
        if (smem_layout(t->x0_bytes[v], &t->stages_c[v], &t->smem_bytes_c[v], 113, 2, TC_SLOT_BYTES_COMPACT)) t->compact = 0;
    if (t->compact) {
        int mc = 0;
        for (int v = 0; v < 3; ++v) mc = t->smem_bytes_c[v] > mc ? t->smem_bytes_c[v] : mc;
        if (cudaFuncSetAttribute(pf_mlp_tc_kernel_compact, cudaFuncAttributeMaxDynamicSharedMemorySize, mc) != cudaSuccess) { cudaGetLastError(); t->compact = 0; }
    }
    // two-phase variant: both nets with the same step list (so that one item's second half interleaves with the other
    // net's first half at the actor -> critic boundary), same packed weights
    t->pipe = getenv("PF_TC_NO_PIPE") == nullptr;
    { const char* e = getenv("PF_TC_PIPE_SLOT_KB"); t->slot_bytes_p = e ? atoi(e) * 1024 : 32768; }
    // staged rows double-buffered if that fits beside two chunk buffers and two ring slots (inputs up to ~190 wide), else single
    t->x0_nbuf_p = 2;
    { int st_, sm_; const int xmax = t->x0_bytes[2]; if (smem_layout(xmax, &st_, &sm_, 227, 2, t->slot_bytes_p, 2, 2)) t->x0_nbuf_p = 1; }
    for (int k = 0; k < 2 && t->pipe; ++k)
        if (build_program_pipe(t->progp[k], c->net[k], t->slot_bytes_p, t->x0_nbuf_p) || check_program(t->progp[k]) || t->progp[k].blob_bytes != t->prog[k].blob_bytes)
            t->pipe = 0;
    if (t->pipe) {
        const TcProgram &A = t->progp[0], &B = t->progp[1];
        if (A.ngroups != B.ngroups || A.nepis != B.nepis) t->pipe = 0;
        for (int i = 0; i < A.ngroups && t->pipe; ++i) if (A.g[i].phase != B.g[i].phase || A.g[i].layer != B.g[i].layer) t->pipe = 0;
        for (int i = 0; i < A.nepis && t->pipe; ++i) if (A.e[i].phase != B.e[i].phase) t->pipe = 0;
    }
    for (int v = 0; v < 3 && t->pipe; ++v) {
        if (smem_layout(t->x0_bytes[v], &t->stages_p[v], &t->smem_bytes_p[v], 227, 2, t->slot_bytes_p, 2, t->x0_nbuf_p)) t->pipe = 0;
        else max_smem = t->smem_bytes_p[v] > max_smem ? t->smem_bytes_p[v] : max_smem;
    }
    for (int k = 0; k < 2; ++k) {
        if (cudaMalloc(&c->net[k].tc_weights, (size_t)t->prog[k].blob_bytes) != cudaSuccess) { delete t; return tc_fail(2, "pf_rl_create: cudaMalloc of the packed weights failed"); }
        cudaMemset(c->net[k].tc_weights, 0, (size_t)t->prog[k].blob_bytes);      // padding columns of the bias tiles stay zero
        c->net[k].tc_weight_bytes = (size_t)t->prog[k].blob_bytes;
    }
    if (cudaFuncSetAttribute(pf_mlp_tc_kernel_std, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem) != cudaSuccess) {
        delete t;
        return tc_fail(2, "pf_rl_create: cannot reserve shared memory for the tensor-core kernel");
    }
    c->tc = t;
    return 0;
}

void pf_tc_destroy(pf_rl_ctx* c) {
    for (int k = 0; k < 2; ++k) { cudaFree(c->net[k].tc_weights); c->net[k].tc_weights = nullptr; }
    delete static_cast<TcState*>(c->tc);
    c->tc = nullptr;
}

int pf_tc_pack_weights(pf_rl_ctx* c, const float* params, cudaStream_t st) {
    TcState* t = static_cast<TcState*>(c->tc);
    TcPackArgs a;
    for (int k = 0; k < 2; ++k) {
        a.prog[k] = t->prog[k];
        a.blob[k] = static_cast<unsigned char*>(c->net[k].tc_weights);
        for (int l = 0; l < PF_MAX_MLP_LAYERS; ++l) { a.w_off[k][l] = c->net[k].w_off[l]; a.b_off[k][l] = c->net[k].b_off[l]; }
    }
    PF_LAUNCHED(), pf_tc_pack_kernel<<<dim3(16, (unsigned)(t->prog[0].ngroups + t->prog[1].ngroups)), 256, 0, st>>>(a, params);
    return cudaGetLastError() == cudaSuccess ? 0 : tc_fail(2, "pf_tc_pack_weights: launch failed");
}

// Does this two-phase program have exactly the shape tc_mma_fast3 is written for?  Fills the per-net parameters it reads.
static int tc_fast3_params(const TcProgram& P, int& nfirst, int& nrest, int& nto, int& merged) {
    if (P.phases != 2 || P.layers != 4 || P.pdims[1] != 512 || P.pdims[2] != 256 || P.pdims[3] != 128 || P.ngroups != 11 || P.nepis != 8) return 0;
    // the order the role issues in: (layer, slice) per group
    static const int want_layer[11] = {0, 2, 0, 2, 1, 0, 1, 3, 0, 1, 1}, want_idx[11] = {0, 0, 1, 1, 0, 2, 1, 0, 3, 2, 3};
    for (int i = 0; i < 11; ++i) {
        const TcGroup& g = P.g[i];
        if (g.layer != want_layer[i]) return 0;
        const int idx = g.layer == 0 ? g.n_base / TC_CHUNK : g.k_base / TC_CHUNK;
        if (idx != want_idx[i]) return 0;
        if (g.layer == 0) {           // bias + first K columns in one tile, at most one more tile
            if (g.nt != 128 || !g.bias || g.kt0 == 0 || g.kc - g.kt0 > g.kt || g.a_src != 0 || g.tmem_col != 0) return 0;
            nfirst = g.kt0 >> 4; nrest = (g.kc - g.kt0) >> 4;
        } else if (g.layer == 1) {    // [bias] + two tiles of 64
            if (g.nt != 256 || g.kc != 128 || g.kt != 64 || g.kt0 != 0 || g.bias != (idx == 0) || g.a_src != 1 || g.tmem_col != 256) return 0;
        } else if (g.layer == 2) {    // [bias] + one tile of 128
            if (g.nt != 128 || g.kc != 128 || g.kt != 128 || g.kt0 != 0 || g.bias != (idx == 0) || g.a_src != 2 || g.tmem_col != 128) return 0;
        } else {                      // output layer: one K tile of 128, the bias with it or in front of it
            if (g.kc != 128 || g.kt != 128 || !g.bias || (g.kt0 != 0 && g.kt0 != 128) || g.a_src != 2 || g.tmem_col != 128 || g.nt > 128) return 0;
            nto = g.nt; merged = g.kt0 != 0;
        }
    }
    return 1;
}

// ... and the single-phase program tc_mma_fast1 is written for.
static int tc_fast1_params(const TcProgram& P, int& nfirst, int& nrest, int& nto, int& merged) {
    if (P.phases != 1 || P.layers != 4 || P.pdims[1] != 512 || P.pdims[2] != 256 || P.pdims[3] != 128 || P.ngroups != 11 || P.nepis != 8 || P.odd_col != 256) return 0;
    static const int want_layer[11] = {0, 0, 1, 0, 1, 0, 1, 1, 2, 2, 3}, want_idx[11] = {0, 1, 0, 2, 1, 3, 2, 3, 0, 1, 0};
    for (int i = 0; i < 11; ++i) {
        const TcGroup& g = P.g[i];
        if (g.layer != want_layer[i]) return 0;
        const int idx = g.layer == 0 ? g.n_base / TC_CHUNK : g.k_base / TC_CHUNK;
        if (idx != want_idx[i]) return 0;
        if (g.layer == 0) {
            if (g.nt != 128 || !g.bias || g.kt0 == 0 || g.kc - g.kt0 > g.kt || g.a_src != 0 || g.tmem_col != (idx & 1) * 128) return 0;
            nfirst = g.kt0 >> 4; nrest = (g.kc - g.kt0) >> 4;
        } else if (g.layer == 1) {
            if (g.nt != 256 || g.kc != 128 || g.kt != 64 || g.kt0 != 0 || g.bias != (idx == 0) || g.a_src != 1 + (idx & 1) || g.tmem_col != 256) return 0;
        } else if (g.layer == 2) {
            if (g.nt != 128 || g.kc != 128 || g.kt != 128 || g.kt0 != 0 || g.bias != (idx == 0) || g.a_src != 1 + idx || g.tmem_col != 0) return 0;
        } else {
            if (g.kc != 128 || g.kt != 128 || !g.bias || (g.kt0 != 0 && g.kt0 != 128) || g.a_src != 1 || g.tmem_col != 256 || g.nt > 128) return 0;
            nto = g.nt; merged = g.kt0 != 0;
        }
    }
    return 1;
}

static int tc_launch(pf_rl_ctx* c, TcArgs& a, int variant, cudaStream_t st) {
    TcState* t = static_cast<TcState*>(c->tc);
    a.weights_early = pf_rl_weights_early(c, st);
    for (int i = 0; i < a.nnets; ++i) { const uint32_t K0 = (uint32_t)a.prog[i].dims[0]; a.net[i].magic = (uint32_t)((0x100000000ull + K0 - 1) / K0); }
    a.tiles = (a.rows + TC_ROWS - 1) / TC_ROWS;
    a.stages = t->stages[variant];
    a.slot_bytes = t->slot_bytes;
    a.nbuf = 2;
    a.x0_nbuf = 1;
    a.x0_bytes = t->x0_bytes[variant];
    a.trace = t->trace;
    a.rehearse = 0;
    { const char* e = getenv("PF_TC_DEBUG"); a.debug = e ? atoi(e) : 0; }
    const int items = a.nnets * a.tiles;
    if (items == 0) return 0;
    // the dry item pays off only while a CTA has a single item (its whole latency is exposed) -- measured: -5 us at
    // 4,096 envs, +3 us at 16,384 (two items per CTA, a shorter wait to hide it in)
    a.rehearse = t->rehearse && items <= c->num_sms;
    // An item is a serial chain of MMA -> epilogue hand-offs (~10k cycles for the flat nets, the tensor pipe busy for a few
    // percent of them): once a CTA would walk several items, two CTAs per SM (the compact variant) overlap two chains.
    // With at most one item per SM the 8 + 8 warp variant has the shorter single-item latency.
    static const int force = [] { const char* e = getenv("PF_TC_COMPACT"); return e ? atoi(e) : -1; }();    // experiments: 0 never, 1 always
    const bool compact = t->compact && (force < 0 ? items > c->num_sms : force != 0);
    if (compact) {
        for (int i = 0; i < a.nnets; ++i)
            for (int k = 0; k < 2; ++k)
                if (a.net[i].blob == static_cast<const unsigned char*>(c->net[k].tc_weights)) a.prog[i] = t->progc[k];
        a.stages = t->stages_c[variant];
        a.slot_bytes = TC_SLOT_BYTES_COMPACT;
        a.tmem_cols = 256;
        const int grid = items < 2 * c->num_sms ? items : 2 * c->num_sms;
        return pf_launch_pdl(pf_mlp_tc_kernel_compact, dim3(grid), dim3(TC_THREADS_OF(4, 4)), (size_t)t->smem_bytes_c[variant], st, a) == cudaSuccess
                   ? 0 : tc_fail(2, "pf_mlp_tc_kernel (compact): launch failed");
    }
    a.tmem_cols = TC_TMEM_COLS;
    const int grid = items < c->num_sms ? items : c->num_sms;
    // The two-phase programs: second half of item m interleaved with the first half of item m + 1.
    // The two-phase program is OPT-IN (PF_TC_PIPE=1, read per launch).  On the program interpreter it was a wash; with the
    // hard-wired MMA / loader roles (fast3, the [*, 512, 256, 128, *] nets) it beat the interpreted single-phase program by
    // 14-17 % -- and then the single-phase program got its own hard-wired role (tc_mma_fast1) and is 6-16 % faster than
    // the two-phase one at every size (262,144 rows: 421 vs 448 us; 16,384: 41.5 vs 49.3): two accumulator slots and two
    // chunk buffers for layer 0 are worth more than overlapping the two halves of consecutive items.  DESIGN.md 4.3.
    const char* pipe_env = getenv("PF_TC_PIPE");
    bool f3ok = t->pipe && getenv("PF_TC_NO_FAST3") == nullptr && t->slot_bytes_p == 32768;
    for (int i = 0; i < a.nnets && f3ok; ++i)
        for (int k = 0; k < 2; ++k)
            if (a.net[i].blob == static_cast<const unsigned char*>(c->net[k].tc_weights))
                f3ok = tc_fast3_params(t->progp[k], a.f3_first[i], a.f3_rest[i], a.f3_nto[i], a.f3_merged[i]) != 0;
    if (t->pipe && pipe_env && atoi(pipe_env) != 0) {
        for (int i = 0; i < a.nnets; ++i)
            for (int k = 0; k < 2; ++k)
                if (a.net[i].blob == static_cast<const unsigned char*>(c->net[k].tc_weights)) a.prog[i] = t->progp[k];
        a.stages = t->stages_p[variant];
        a.slot_bytes = t->slot_bytes_p;
        a.nbuf = 2;
        a.x0_nbuf = t->x0_nbuf_p;
        a.rehearse = 0;
        // the hard-wired MMA role, if every net of this launch has exactly the program it was written against
        a.fast3 = f3ok ? 1 : 0;
        if (a.nnets == 1) { a.f3_first[1] = a.f3_first[0]; a.f3_rest[1] = a.f3_rest[0]; a.f3_nto[1] = a.f3_nto[0]; a.f3_merged[1] = a.f3_merged[0]; }
        return pf_launch_pdl(pf_mlp_tc_kernel_std, dim3(grid), dim3(TC_THREADS_OF(8, 8)), (size_t)t->smem_bytes_p[variant], st, a) == cudaSuccess
                   ? 0 : tc_fail(2, "pf_mlp_tc_kernel (two-phase): launch failed");
    }
    // single-phase program: the hard-wired MMA role where the nets are the ones it is written for (the loader walks the program)
    a.fast3 = (getenv("PF_TC_NO_FAST3") == nullptr && t->slot_bytes == 32768) ? 2 : 0;
    for (int i = 0; i < a.nnets && a.fast3; ++i)
        if (!tc_fast1_params(a.prog[i], a.f3_first[i], a.f3_rest[i], a.f3_nto[i], a.f3_merged[i])) a.fast3 = 0;
    if (a.nnets == 1) { a.f3_first[1] = a.f3_first[0]; a.f3_rest[1] = a.f3_rest[0]; a.f3_nto[1] = a.f3_nto[0]; a.f3_merged[1] = a.f3_merged[0]; }
    return pf_launch_pdl(pf_mlp_tc_kernel_std, dim3(grid), dim3(TC_THREADS_OF(8, 8)), (size_t)t->smem_bytes[variant], st, a) == cudaSuccess
               ? 0 : tc_fail(2, "pf_mlp_tc_kernel: launch failed");
}

int pf_tc_mlp_forward(pf_rl_ctx* c, int net, const float* params, const float* x, long long ldx, int rows, float* out,
                      float* copy_to, cudaStream_t st) {
    (void)params;
    TcState* t = static_cast<TcState*>(c->tc);
    TcArgs a;
    memset(&a, 0, sizeof(a));
    a.prog[0] = t->prog[net];
    a.net[0].blob = static_cast<const unsigned char*>(c->net[net].tc_weights);
    a.net[0].x = x; a.net[0].ldx = ldx; a.net[0].out = out; a.net[0].copy_to = copy_to; a.net[0].mode = TC_OUT_PLAIN;
    a.nnets = 1; a.rows = rows;
    return tc_launch(c, a, net, st);
}

int pf_tc_act_fusable(const pf_rl_ctx* c) {
    const TcState* t = static_cast<const TcState*>(c->tc);
    return t && t->prog[0].pdims[t->prog[0].layers] == 16 && t->prog[1].dims[t->prog[1].layers] == 1;
}

// PPO.act in one launch: critic items first (the longer ones), then the actor items with the sample / log-prob tail.
int pf_tc_act(pf_rl_ctx* c, const float* params, const float* obs, const float* critic_obs, const pf_rl_storage* storage,
              int t_row, int64_t policy_step, float* actions_out, float* values_out, float* logp_out, cudaStream_t st) {
    TcState* t = static_cast<TcState*>(c->tc);
    const pf_rl_params& P = c->P;
    TcArgs a;
    memset(&a, 0, sizeof(a));
    const long long row0 = (long long)t_row * P.num_envs;
    for (int i = 0; i < 2; ++i) {
        const int k = 1 - i;                       // item order: critic, actor
        a.prog[i] = t->prog[k];
        a.net[i].blob = static_cast<const unsigned char*>(c->net[k].tc_weights);
        a.net[i].x = k == 0 ? obs : critic_obs;
        a.net[i].ldx = c->net[k].dims[0];
        a.net[i].out = nullptr;
        a.net[i].mode = k == 0 ? TC_OUT_SAMPLE : TC_OUT_VALUE;
        a.net[i].copy_to = nullptr;
        if (storage) {
            if (k == 0) a.net[i].copy_to = storage->observations + row0 * P.num_obs;
            else if (storage->privileged_observations) a.net[i].copy_to = storage->privileged_observations + row0 * P.num_critic_obs;
        }
    }
    a.nnets = 2; a.rows = P.num_envs;
    a.seed = (uint32_t)P.seed; a.policy_step = (uint32_t)policy_step; a.env_id_offset = (int)P.env_id_offset; a.num_actions = P.num_actions;
    a.std = params + c->std_off;
    if (storage) { a.S = *storage; a.has_storage = 1; }
    a.storage_row0 = row0;
    a.actions_out = actions_out; a.values_out = values_out; a.logp_out = logp_out;
    return tc_launch(c, a, 2, st);
}
