// Per-environment logic of PointFoot.step(), written once as host/device inline functions.
// The CUDA kernels (pf_env.cu) call it with rows staged in shared memory; tests/host_emu compiles the same
// header with g++ and calls it in a serial loop, so the logic can be checked without a GPU.  That emulation
// is test infrastructure: it is not linked into libpointfoot_b200.so.
//
// Floating point: the reference runs torch CPU eager (PF:41,50 hard-code 'cpu'), i.e. one rounding per op and
// no contraction, EXCEPT where ATen itself fuses: norm over 2/3 elements is an fma chain and cross() is
// fma(a1,b2,-(a2*b1)) (measured).  This file is compiled with -fmad=false / -ffp-contract=off
// and spells those fmas out, so integer results that depend on fp32 thresholds (termination, contacts,
// terrain indices, curriculum levels) come out bit-identical.
#pragma once
#include <stdint.h>
#include <math.h>
#include "../../include/pointfoot_b200.h"
#include "pf_rng.cuh"

#ifdef __CUDACC__
#define PF_HD __host__ __device__ __forceinline__
#define PF_CE __host__ __device__ constexpr
#else
#define PF_HD static inline
#define PF_CE constexpr
#endif

// min / max as the select the reference's clip ops amount to for ordinary numbers.  On the device they are the
// single-instruction FMNMX (fminf / fmaxf): identical for every pair of non-NaN operands except the sign of a zero
// result when both operands are zeros of opposite sign, and NaN in the FIRST operand yields the second either way.
#ifdef __CUDA_ARCH__
PF_HD float pf_minf(float a, float b) { return fminf(a, b); }
PF_HD float pf_maxf(float a, float b) { return fmaxf(a, b); }
#else
PF_HD float pf_minf(float a, float b) { return a < b ? a : b; }
PF_HD float pf_maxf(float a, float b) { return a > b ? a : b; }
#endif
// Pairs of neighbouring values of one env ([N,2] feet arrays, the three (x, y) pairs of an [N,6] row) as ONE 64-bit
// (16-bit for bool pairs) access on the device: half the load / store instructions and L1 lookups of the per-env state.
// The kernels take this path only when every such tensor is 8-byte aligned (pf_post_physics_impl checks).
#ifdef __CUDA_ARCH__
struct alignas(8) PfF2 { float x, y; };
PF_HD void pf_ld2(const float* p, float& a, float& b) { const PfF2 v = *reinterpret_cast<const PfF2*>(p); a = v.x; b = v.y; }
PF_HD void pf_st2(float* p, float a, float b) { PfF2 v; v.x = a; v.y = b; *reinterpret_cast<PfF2*>(p) = v; }
PF_HD void pf_ldb2(const uint8_t* p, bool& a, bool& b) { const uint16_t v = *reinterpret_cast<const uint16_t*>(p); a = (v & 0xffu) != 0; b = (v >> 8) != 0; }
PF_HD void pf_stb2(uint8_t* p, bool a, bool b) { *reinterpret_cast<uint16_t*>(p) = (uint16_t)((a ? 1u : 0u) | ((b ? 1u : 0u) << 8)); }
#else
PF_HD void pf_ld2(const float* p, float& a, float& b) { a = p[0]; b = p[1]; }
PF_HD void pf_st2(float* p, float a, float b) { p[0] = a; p[1] = b; }
PF_HD void pf_ldb2(const uint8_t* p, bool& a, bool& b) { a = p[0] != 0; b = p[1] != 0; }
PF_HD void pf_stb2(uint8_t* p, bool a, bool b) { p[0] = a; p[1] = b; }
#endif

PF_HD float pf_clipf(float x, float lo, float hi) { return pf_minf(pf_maxf(x, lo), hi); }
PF_HD float pf_sq(float x) { return x * x; }

// ATen CPU norm(p=2) over the last dim: 2 and 3 elements = fma chain, 4 elements = plain sequential sum.
PF_HD float pf_norm2(float x, float y) { return sqrtf(fmaf(y, y, x * x)); }
PF_HD float pf_norm3(float x, float y, float z) { return sqrtf(fmaf(z, z, fmaf(y, y, x * x))); }

PF_HD float pf_sumsq3(float x, float y, float z) { return fmaf(z, z, fmaf(y, y, x * x)); }   // pf_norm3 before the root
// x / d for a fixed divisor, r = RN(1/d): quotient estimate + one exact-remainder correction (Markstein), the
// correctly rounded quotient without the division sequence
PF_HD float pf_div_by(float x, float d, float r) {
    const float q = x * r;
    return fmaf(fmaf(-q, d, x), r, q);
}

// Reward terms by id: row of episode_sums (-1 = term not enabled) and scale.  Built once on the host; handed to the
// step as a kernel parameter, so that every lookup below is a constant-bank operand at a compile-time offset.
struct PfRewTab {
    int32_t row[PF_REW_COUNT];
    int32_t off[PF_REW_COUNT];        // row * num_envs: element offset of the term's episode_sums row
    float scale[PF_REW_COUNT];
    // sqrtf(x) > c  <=>  x >= sq_gt(c) for the two contact thresholds (PF:241: 1.0, PF:1167: 0.1): the squared
    // norms are compared directly, no square root on the way to a bool
    float sq_gt_termination, sq_gt_collision;
    float rdt;                        // RN(1 / dt)
    int32_t dt_div_exact;             // pf_div_by(x, dt, rdt) == x / dt for every x (checked exhaustively on the host)
    float r_hscale;                   // RN(1 / horizontal_scale)
    float rows_m2, cols_m2;           // (float)(hf_rows - 2), (float)(hf_cols - 2): index clamps of PF:1104-1105
    int32_t index_div_kind;           // PF:1102 (x / horizontal_scale): 0 IEEE division, 1 x * r_hscale (CUDA eager),
                                      // 2 pf_div_by (bit-identical to the division, checked exhaustively on the host)
};

// Is pf_div_by(x, d, RN(1/d)) the correctly rounded quotient for EVERY x?  One binade of x (all 2^23 significands)
// decides it: scaling x by a power of two scales every intermediate exactly.  ~10 ms with hardware FMA; cached.
#if defined(__x86_64__) && defined(__GNUC__)
__attribute__((target("fma")))
#endif
static bool pf_div_by_scan(float d, float r) {
    for (uint32_t m = 0; m < (1u << 23); ++m) {
        union { uint32_t u; float f; } a;
        a.u = 0x3f800000u + m;
        const float q = a.f * r;
        const float q1 = __builtin_fmaf(__builtin_fmaf(-q, d, a.f), r, q);
        if (q1 != a.f / d) return false;
    }
    return true;
}
static inline bool pf_div_by_is_exact(float d) {
    static float seen_d[8]; static int seen_ok[8]; static int n_seen = 0;      // benign race: same answer either way
    if (!(d > 1e-30f && d < 1e30f)) return false;
    for (int i = 0; i < n_seen && i < 8; ++i) if (seen_d[i] == d) return seen_ok[i] != 0;
    const bool ok = pf_div_by_scan(d, 1.0f / d);
    if (n_seen < 8) { seen_d[n_seen] = d; seen_ok[n_seen] = ok; ++n_seen; }
    return ok;
}


// smallest float t with sqrtf(t) > c (sqrtf is correctly rounded and monotone on both sides of the bus)
static inline float pf_sq_gt(float c) {
    float t = (float)((double)c * (double)c);
    while (sqrtf(t) > c) t = nextafterf(t, 0.0f);
    while (!(sqrtf(t) > c)) t = nextafterf(t, INFINITY);
    return t;
}

// torch.cross, ATen CPU: out[i] = fma(a[i+1], b[i+2], -(a[i+2]*b[i+1]))
PF_HD void pf_cross(const float a[3], const float b[3], float o[3]) {
    o[0] = fmaf(a[1], b[2], -(a[2] * b[1]));
    o[1] = fmaf(a[2], b[0], -(a[0] * b[2]));
    o[2] = fmaf(a[0], b[1], -(a[1] * b[0]));
}

// isaacgym.torch_utils.quat_rotate_inverse / quat_rotate; q is xyzw.
PF_HD void pf_quat_rotate(const float q[4], const float v[3], float o[3], bool inverse) {
    const float qw = q[3];
    const float s = 2.0f * (qw * qw) - 1.0f;
    float c[3];
    pf_cross(q, v, c);
    const float d = (q[0] * v[0] + q[1] * v[1]) + q[2] * v[2];     // bmm(q_vec, v)
    for (int i = 0; i < 3; ++i) {
        const float a = v[i] * s;
        const float b = (c[i] * qw) * 2.0f;
        const float cc = (q[i] * d) * 2.0f;
        o[i] = inverse ? (a - b) + cc : (a + b) + cc;
    }
}

// isaacgym.torch_utils.quat_apply
PF_HD void pf_quat_apply(const float q[4], const float b[3], float o[3]) {
    float t[3], u[3];
    pf_cross(q, b, t);
    t[0] *= 2.0f; t[1] *= 2.0f; t[2] *= 2.0f;
    pf_cross(q, t, u);
    for (int i = 0; i < 3; ++i) o[i] = (b[i] + q[3] * t[i]) + u[i];
}

// (x / horizontal_scale) of PF:1102.  CPU eager: true divide.  CUDA eager: x * (1.0f / scale).
PF_HD float pf_index_div(float x, float scale, int mode) {
    return mode == PF_DIV_TRUE ? x / scale : x * (1.0f / scale);
}

// _get_terrain_heights_from_points, PF:1100-1112: min of 3 neighbouring int16 samples (NOT bilinear).
PF_HD int16_t pf_terrain_min3(const pf_params& P, const int16_t* __restrict__ hs, float x, float y) {
    x = x + P.border_size;
    y = y + P.border_size;
    x = pf_index_div(x, P.horizontal_scale, P.index_div_mode);
    y = pf_index_div(y, P.horizontal_scale, P.index_div_mode);
    long long px = (long long)x;       // .long(): truncation toward zero
    long long py = (long long)y;
    const long long mx = P.hf_rows - 2, my = P.hf_cols - 2;
    px = px < 0 ? 0 : (px > mx ? mx : px);
    py = py < 0 ? 0 : (py > my ? my : py);
    const int16_t* r = hs + px * (long long)P.hf_cols + py;
#ifdef __CUDA_ARCH__
    const int16_t h1 = __ldg(r), h2 = __ldg(r + P.hf_cols), h3 = __ldg(r + 1);   // read-only path: may move across stores
#else
    const int16_t h1 = r[0], h2 = r[P.hf_cols], h3 = r[1];
#endif
    const int16_t m = h1 < h2 ? h1 : h2;
    return m < h3 ? m : h3;
}

// The same lookup with the table's constants: division by the scale without the division sequence where that is
// bit-identical (PfRewTab::index_div_kind), truncation + clamp done on the float (trunc and clamp commute, the
// bounds are small integers, NaN -> 0 either way), 32-bit indexing.
PF_HD int16_t pf_terrain_min3_fast(const pf_params& P, const PfRewTab& T, const int16_t* __restrict__ hs, float x, float y) {
    x = x + P.border_size;
    y = y + P.border_size;
    if (T.index_div_kind == 2) { x = pf_div_by(x, P.horizontal_scale, T.r_hscale); y = pf_div_by(y, P.horizontal_scale, T.r_hscale); }
    else if (T.index_div_kind == 1) { x = x * T.r_hscale; y = y * T.r_hscale; }
    else { x = x / P.horizontal_scale; y = y / P.horizontal_scale; }
    const int px = (int)pf_minf(pf_maxf(x, 0.0f), T.rows_m2);
    const int py = (int)pf_minf(pf_maxf(y, 0.0f), T.cols_m2);
    const int16_t* r = hs + (px * P.hf_cols + py);
#ifdef __CUDA_ARCH__
    const int16_t h1 = __ldg(r), h2 = __ldg(r + P.hf_cols), h3 = __ldg(r + 1);
#else
    const int16_t h1 = r[0], h2 = r[P.hf_cols], h3 = r[1];
#endif
    const int16_t m = h1 < h2 ? h1 : h2;
    return m < h3 ? m : h3;
}

// Normalised yaw-only quaternion of quat_apply_yaw (math.py:38-42): returns (z, w) of normalize((0,0,qz,qw)).
PF_HD void pf_yaw_quat(const float q[4], float& zn, float& wn) {
    float n = sqrtf(q[2] * q[2] + q[3] * q[3]);     // norm over 4 elements: plain sum; the two zeros are exact
    n = pf_maxf(n, 1e-9f);
    zn = q[2] / n;
    wn = q[3] / n;
}

// One height-scan sample, PF:1070-1075: rotate grid point (gx,gy,0) by the yaw quaternion, add base xyz,
// look the terrain up.  Returns the raw int16 sample.
PF_HD int16_t pf_scan_point(const pf_params& P, const int16_t* __restrict__ hs, float zn, float wn,
                            float gx, float gy, float base_x, float base_y, const PfRewTab* T = nullptr) {
    // quat_apply((0,0,zn,wn), (gx,gy,0)) with the cross() formula above; zero products dropped (exact)
    const float t0 = -(zn * gy) * 2.0f;
    const float t1 = (zn * gx) * 2.0f;
    const float ox = (gx + wn * t0) + (-(zn * t1));
    const float oy = (gy + wn * t1) + (zn * t0);
    return T ? pf_terrain_min3_fast(P, *T, hs, ox + base_x, oy + base_y) : pf_terrain_min3(P, hs, ox + base_x, oy + base_y);
}

// wrap_to_pi, math.py:45-48 (torch remainder = fmod with the divisor's sign)
PF_HD float pf_wrap_to_pi(float a) {
    const float two_pi = 6.283185307179586f;
    float m = fmodf(a, two_pi);
    if (m != 0.0f && m < 0.0f) m += two_pi;
    m -= two_pi * (m > 3.141592653589793f ? 1.0f : 0.0f);
    return m;
}

// ------------------------------------------------------------------------------------------------
// Inputs of one env for one step.  Row pointers may point into shared memory (kernel) or straight into the
// tensors (host emulation).
struct PfEnvIn {
    const float* root;        // 13
    const float* dof;         // D pairs (pos, vel)
    const float* cf;          // B*3 contact forces
    const float* feet_pos[PF_NUM_FEET];   // xyz of rigid_body_states[feet_indices[f]]
    float height_mean;        // mean_p (root_z - measured_heights[p]) over the P scan points: fp32 terms summed in
                              // double, rounded once (ATen's cascade sum order depends on the host ISA; this is
                              // the closest order-free statement of torch.mean, PF:1144-1147)
    bool have_heights;        // measured_heights is a tensor (rough) rather than the int 0 (PF:775)
};

// Everything the step leaves behind for the observation stage.
struct PfEnvOut {
    float ang_vel[3], gravity[3];     // pre-reset (PF:210-211)
    float dof_pos[PF_MAX_DOF], dof_vel[PF_MAX_DOF];   // post-reset
    float actions[PF_MAX_DOF];        // this step's clipped actions (PF:177)
    float root_z;                     // post-reset
    float root_vel[6];                // post-reset root lin / ang velocity (PF:227)
    float commands[4];                // after the heading update
    bool reset;
};

// The reward names in dir() (= alphabetical) order: the order class_to_dict hands them to
// _prepare_reward_function (PF:796-820), hence the summation order of compute_reward (PF:307-311).
#define PF_REWARD_CANON(X)                                                                                   \
    X(PF_REW_ACTION_RATE) X(PF_REW_ANG_VEL_XY) X(PF_REW_BASE_HEIGHT) X(PF_REW_COLLISION) X(PF_REW_DOF_ACC)    \
    X(PF_REW_DOF_POS_LIMITS) X(PF_REW_DOF_VEL) X(PF_REW_DOF_VEL_LIMITS) X(PF_REW_FEET_AIR_TIME)              \
    X(PF_REW_FEET_CONTACT_FORCES) X(PF_REW_FEET_DISTANCE) X(PF_REW_LIN_VEL_Z) X(PF_REW_NO_FLY)               \
    X(PF_REW_ORIENTATION) X(PF_REW_STAND_STILL) X(PF_REW_STUMBLE) X(PF_REW_SURVIVAL) X(PF_REW_TERMINATION)   \
    X(PF_REW_TORQUE_LIMITS) X(PF_REW_TORQUES) X(PF_REW_TRACKING_ANG_VEL) X(PF_REW_TRACKING_LIN_VEL)          \
    X(PF_REW_UNBALANCE_FEET_AIR_TIME) X(PF_REW_UNBALANCE_FEET_HEIGHT)

// True when P.reward_ids is a subsequence of the canonical order (what the reference's cfg always yields): the
// step then runs the straight-line statement of the reward sum, every term's code specialised at compile time.
static inline bool pf_rewards_canonical(const pf_params& P) {
    static const int canon[] = {
#define PF_X(id) id,
        PF_REWARD_CANON(PF_X)
#undef PF_X
    };
    int pos = 0;
    for (int k = 0; k < P.num_reward_terms; ++k) {
        while (pos < PF_REW_COUNT && canon[pos] != P.reward_ids[k]) ++pos;
        if (pos == PF_REW_COUNT) return false;
        ++pos;
    }
    return true;
}

// Position of a term in the canonical order, and the set of enabled terms as a bit mask over those positions.
// The step is compiled three ways (template parameter kMask of pf_env_step_t):
//   PF_MASK_GENERIC   any num_dof, terms in any order: guarded loops, order-agnostic reward loop
//   PF_MASK_TABLE     6 DOF, canonical order, any subset: straight-line sum, every term behind a warp-uniform test
//   a literal mask    6 DOF, exactly that subset (PF_MASK_POINTFOOT = the 19 terms of both pointfoot cfgs,
//                     pointfoot_rough_config.py:131-154): disabled terms are not compiled at all, rows of
//                     episode_sums are compile-time constants
PF_CE int pf_canon_pos(int id) {
    int p = 0;
#define PF_X(ID) if (id == (ID)) return p; ++p;
    PF_REWARD_CANON(PF_X)
#undef PF_X
    return -1;
}
PF_CE uint32_t pf_canon_bit(int id) { return 1u << pf_canon_pos(id); }
PF_CE int pf_popc_const(uint32_t v) { int n = 0; for (; v; v &= v - 1) ++n; return n; }
#define PF_MASK_GENERIC 0u
#define PF_MASK_TABLE 0x80000000u
constexpr uint32_t PF_MASK_POINTFOOT =
    pf_canon_bit(PF_REW_ACTION_RATE) | pf_canon_bit(PF_REW_ANG_VEL_XY) | pf_canon_bit(PF_REW_BASE_HEIGHT) |
    pf_canon_bit(PF_REW_COLLISION) | pf_canon_bit(PF_REW_DOF_ACC) | pf_canon_bit(PF_REW_FEET_AIR_TIME) |
    pf_canon_bit(PF_REW_FEET_CONTACT_FORCES) | pf_canon_bit(PF_REW_FEET_DISTANCE) | pf_canon_bit(PF_REW_LIN_VEL_Z) |
    pf_canon_bit(PF_REW_NO_FLY) | pf_canon_bit(PF_REW_ORIENTATION) | pf_canon_bit(PF_REW_STAND_STILL) |
    pf_canon_bit(PF_REW_SURVIVAL) | pf_canon_bit(PF_REW_TORQUE_LIMITS) | pf_canon_bit(PF_REW_TORQUES) |
    pf_canon_bit(PF_REW_TRACKING_ANG_VEL) | pf_canon_bit(PF_REW_TRACKING_LIN_VEL) |
    pf_canon_bit(PF_REW_UNBALANCE_FEET_AIR_TIME) | pf_canon_bit(PF_REW_UNBALANCE_FEET_HEIGHT);

// is term ID enabled / element offset of its episode_sums row: compile-time for a literal mask
template <uint32_t kMask, int ID>
PF_HD bool pf_term_on(const PfRewTab& T) {
    constexpr bool on = ((kMask >> pf_canon_pos(ID)) & 1u) != 0;
    return (kMask & PF_MASK_TABLE) ? T.row[ID] >= 0 : on;
}
template <uint32_t kMask, int ID>
PF_HD int pf_term_row(const PfRewTab& T) {
    constexpr int row = pf_popc_const(kMask & (pf_canon_bit(ID) - 1u));
    return (kMask & PF_MASK_TABLE) ? T.row[ID] : row;
}
template <uint32_t kMask, int ID>
PF_HD size_t pf_term_off(const PfRewTab& T, int N) {
    constexpr int row = pf_popc_const(kMask & (pf_canon_bit(ID) - 1u));
    return (kMask & PF_MASK_TABLE) ? (size_t)T.off[ID] : (size_t)row * (size_t)N;
}

// mask of P's enabled terms over the canonical positions (meaningful when the table is canonical)
static inline uint32_t pf_reward_mask(const pf_params& P) {
    uint32_t m = 0;
    for (int k = 0; k < P.num_reward_terms; ++k) {
        const int pos = pf_canon_pos(P.reward_ids[k]);
        if (pos >= 0) m |= 1u << pos;
    }
    return m;
}

static inline bool pf_build_reward_table(const pf_params& P, PfRewTab& T) {
    for (int i = 0; i < PF_REW_COUNT; ++i) { T.row[i] = -1; T.off[i] = 0; T.scale[i] = 0.0f; }
    T.sq_gt_termination = pf_sq_gt(1.0f);
    T.sq_gt_collision = pf_sq_gt(0.1f);
    T.rdt = 1.0f / P.dt;
    T.dt_div_exact = pf_div_by_is_exact(P.dt) ? 1 : 0;
    T.r_hscale = 1.0f / P.horizontal_scale;
    T.rows_m2 = (float)(P.hf_rows - 2); T.cols_m2 = (float)(P.hf_cols - 2);
    T.index_div_kind = P.index_div_mode == PF_DIV_RECIPROCAL ? 1 : (pf_div_by_is_exact(P.horizontal_scale) ? 2 : 0);
    if ((long long)P.num_envs * (PF_MAX_REWARD_TERMS + 1) > 0x7fffffffLL) return false;
    for (int k = 0; k < P.num_reward_terms; ++k) {
        const int id = P.reward_ids[k];
        if (id < 0 || id >= PF_REW_COUNT || T.row[id] >= 0) return false;      // unknown or repeated term: generic loop
        T.row[id] = k; T.off[id] = k * P.num_envs; T.scale[id] = P.reward_scales[k];
    }
    return pf_rewards_canonical(P);
}

// What the _reward_* terms read.  Pointers to the caller's local arrays: every index below is a compile-time
// constant once the guarded loops are unrolled, so the arrays stay in registers.
struct PfRewIn {
    const float *lin_vel, *ang_vel, *gravity, *commands;
    const float *tq, *ldv, *la, *ac;      // torques, last_dof_vel, last_actions, actions   [D]
    const float *dof, *cf;                // staged rows: (pos, vel) pairs; contact forces
    const float *fat, *lfat, *lmfh;       // feet_air_time, last_feet_air_time, last_max_feet_height   [F]
    const bool* first_contact;
    const float *feet0, *feet1;
    float bh;                             // base height (PF:1144-1147)
    bool reset, time_out;
};

// One _reward_* term, unscaled (PF:1132-1253).
template <int kD>
PF_HD float pf_reward_term(const pf_params& P, const PfRewTab& T, const int id, const PfRewIn& c) {
    const int D = kD ? kD : P.num_dof;
    const float dt = P.dt;
    float r = 0.0f;
    switch (id) {
    case PF_REW_LIN_VEL_Z: r = pf_sq(c.lin_vel[2]); break;
    case PF_REW_ANG_VEL_XY: r = pf_sq(c.ang_vel[0]) + pf_sq(c.ang_vel[1]); break;
    case PF_REW_ORIENTATION: r = pf_sq(c.gravity[0]) + pf_sq(c.gravity[1]); break;
    case PF_REW_BASE_HEIGHT: r = pf_sq(c.bh - P.base_height_target); break;
    case PF_REW_TORQUES:
#pragma unroll
        for (int d = 0; d < PF_MAX_DOF; ++d) if (d < D) r += pf_sq(c.tq[d]);
        break;
    case PF_REW_DOF_VEL:
#pragma unroll
        for (int d = 0; d < PF_MAX_DOF; ++d) if (d < D) r += pf_sq(c.dof[2 * d + 1]);
        break;
    case PF_REW_DOF_ACC:
#pragma unroll
        for (int d = 0; d < PF_MAX_DOF; ++d)
            if (d < D) r += pf_sq((kD && T.dt_div_exact) ? pf_div_by(c.ldv[d] - c.dof[2 * d + 1], dt, T.rdt)
                                                         : (c.ldv[d] - c.dof[2 * d + 1]) / dt);
        break;
    case PF_REW_ACTION_RATE:
#pragma unroll
        for (int d = 0; d < PF_MAX_DOF; ++d) if (d < D) r += pf_sq(c.la[d] - c.ac[d]);
        break;
    case PF_REW_COLLISION:
#pragma unroll
        for (int i = 0; i < PF_MAX_BODIES; ++i)
            if (i < P.num_penalised) {
                const float* F = c.cf + P.penalised_indices[i] * 3;
                const bool hit = kD ? pf_sumsq3(F[0], F[1], F[2]) >= T.sq_gt_collision : pf_norm3(F[0], F[1], F[2]) > 0.1f;
                r += hit ? 1.0f : 0.0f;
            }
        break;
    case PF_REW_TERMINATION: r = (c.reset && !c.time_out) ? 1.0f : 0.0f; break;
    case PF_REW_DOF_POS_LIMITS:
#pragma unroll
        for (int d = 0; d < PF_MAX_DOF; ++d)
            if (d < D) {
                const float qd = c.dof[2 * d];
                r += -pf_minf(qd - P.dof_pos_limits[d][0], 0.0f) + pf_maxf(qd - P.dof_pos_limits[d][1], 0.0f);
            }
        break;
    case PF_REW_DOF_VEL_LIMITS:
#pragma unroll
        for (int d = 0; d < PF_MAX_DOF; ++d)
            if (d < D) r += pf_clipf(fabsf(c.dof[2 * d + 1]) - P.dof_vel_limits[d] * P.soft_dof_vel_limit, 0.0f, 1.0f);
        break;
    case PF_REW_TORQUE_LIMITS:
#pragma unroll
        for (int d = 0; d < PF_MAX_DOF; ++d)
            if (d < D) r += pf_maxf(fabsf(c.tq[d]) - P.torque_limits[d] * P.soft_torque_limit, 0.0f);
        break;
    case PF_REW_TRACKING_LIN_VEL: {
        const float err = pf_sq(c.commands[0] - c.lin_vel[0]) + pf_sq(c.commands[1] - c.lin_vel[1]);
        r = expf(-err / P.tracking_sigma);
    } break;
    case PF_REW_TRACKING_ANG_VEL:
        r = expf(-pf_sq(c.commands[2] - c.ang_vel[2]) / P.tracking_sigma);
        break;
    case PF_REW_FEET_AIR_TIME: {
        float lo = 0.0f, hi = 0.0f;
#pragma unroll
        for (int f = 0; f < PF_NUM_FEET; ++f) {
            const float m = c.first_contact[f] ? 1.0f : 0.0f;
            lo += pf_minf(c.fat[f] - P.min_feet_air_time, 0.0f) * m;
            hi += pf_minf(P.max_feet_air_time - c.fat[f], 0.0f) * m;
        }
        r = lo + hi;
    } break;
    case PF_REW_NO_FLY: {
        int n = 0;
#pragma unroll
        for (int f = 0; f < PF_NUM_FEET; ++f) n += c.cf[P.feet_indices[f] * 3 + 2] > 0.1f;
        r = n == 1 ? 1.0f : 0.0f;
    } break;
    case PF_REW_UNBALANCE_FEET_AIR_TIME: {     // torch.var over 2 values, unbiased
        const float m = (c.lfat[0] + c.lfat[1]) / 2.0f;
        r = pf_sq(c.lfat[0] - m) + pf_sq(c.lfat[1] - m);
    } break;
    case PF_REW_UNBALANCE_FEET_HEIGHT: {
        const float m = (c.lmfh[0] + c.lmfh[1]) / 2.0f;
        r = pf_sq(c.lmfh[0] - m) + pf_sq(c.lmfh[1] - m);
    } break;
    case PF_REW_STUMBLE: {
        bool any = false;
#pragma unroll
        for (int f = 0; f < PF_NUM_FEET; ++f) {
            const float* F = c.cf + P.feet_indices[f] * 3;
            any = any || (pf_norm2(F[0], F[1]) > 5.0f * fabsf(F[2]));
        }
        r = any ? 1.0f : 0.0f;
    } break;
    case PF_REW_STAND_STILL:                   // signed comparison, PF:1233-1234
        r = (fabsf(c.lin_vel[0]) * (c.commands[0] < 0.1f ? 1.0f : 0.0f)
             + fabsf(c.lin_vel[1]) * (c.commands[1] < 0.1f ? 1.0f : 0.0f))
            + fabsf(c.ang_vel[2]) * (c.commands[2] < 0.1f ? 1.0f : 0.0f);
        break;
    case PF_REW_FEET_CONTACT_FORCES:
#pragma unroll
        for (int f = 0; f < PF_NUM_FEET; ++f) {
            const float* F = c.cf + P.feet_indices[f] * 3;
            r += pf_maxf(pf_norm3(F[0], F[1], F[2]) - P.max_contact_force, 0.0f);
        }
        break;
    case PF_REW_FEET_DISTANCE: {
        const float d = pf_norm2(c.feet0[0] - c.feet1[0], c.feet0[1] - c.feet1[1]);
        r = pf_clipf(P.min_feet_distance - d, 0.0f, 1.0f);
    } break;
    case PF_REW_SURVIVAL: r = (c.reset ? 0.0f : 1.0f) * dt; break;
    default: break;
    }
    return r;
}

// Everything the step reads from env-owned global buffers, as one bundle of independent loads: issued together (and,
// in the kernel, BEFORE the wait for the staged PhysX rows, so the two round trips to memory overlap).  Left where
// the values are used, each load sits behind an earlier store the compiler must assume aliases it, and the step pays
// ~60 memory latencies one after another.  Loops run to the compile-time maxima under a guard so that every array
// lives in registers.
struct PfEnvPre {
    float es[PF_REW_COUNT];           // episode sums by term id (kMask != GENERIC)
    float es_term;
    float tq[PF_MAX_DOF], ldv[PF_MAX_DOF], la[PF_MAX_DOF], ac[PF_MAX_DOF];
    bool fc_old[PF_NUM_FEET], cfilt_old[PF_NUM_FEET], lc_old[PF_NUM_FEET];
    float fat_old[PF_NUM_FEET], lfat_old[PF_NUM_FEET], lmfh_old[PF_NUM_FEET], cmfh_old[PF_NUM_FEET];
    float cmd_old[4];
    long long ep_len_old;
};

template <uint32_t kMask>
PF_HD void pf_env_preload(const pf_params& P, const pf_env_buffers& B, const PfRewTab& T, int e, PfEnvPre& pre) {
    constexpr bool kFast = kMask != PF_MASK_GENERIC;
    const int D = kFast ? PF_MAX_DOF : P.num_dof;
    const int N = P.num_envs;
    if (kFast) {
        const float* es_in = B.episode_sums + e;
#define PF_X(ID) pre.es[ID] = pf_term_on<kMask, ID>(T) ? es_in[pf_term_off<kMask, ID>(T, N)] : 0.0f;
        PF_REWARD_CANON(PF_X)
#undef PF_X
    }
    pre.es_term = P.has_termination_term ? B.episode_sums[(size_t)P.num_reward_terms * N + e] : 0.0f;
    if (kFast && PF_MAX_DOF == 6 && PF_NUM_FEET == 2) {
#pragma unroll
        for (int d = 0; d < 6; d += 2) {
            pf_ld2(B.torques + e * 6 + d, pre.tq[d], pre.tq[d + 1]);
            pf_ld2(B.last_dof_vel + e * 6 + d, pre.ldv[d], pre.ldv[d + 1]);
            pf_ld2(B.last_actions + e * 6 + d, pre.la[d], pre.la[d + 1]);
            pf_ld2(B.actions + e * 6 + d, pre.ac[d], pre.ac[d + 1]);
        }
        pf_ldb2(B.first_contact + e * 2, pre.fc_old[0], pre.fc_old[1]);
        pf_ldb2(B.contact_filt + e * 2, pre.cfilt_old[0], pre.cfilt_old[1]);
        pf_ldb2(B.last_contacts + e * 2, pre.lc_old[0], pre.lc_old[1]);
        pf_ld2(B.feet_air_time + e * 2, pre.fat_old[0], pre.fat_old[1]);
        pf_ld2(B.last_feet_air_time + e * 2, pre.lfat_old[0], pre.lfat_old[1]);
        pf_ld2(B.last_max_feet_height + e * 2, pre.lmfh_old[0], pre.lmfh_old[1]);
        pre.cmfh_old[0] = pre.cmfh_old[1] = 0.0f;
        if (P.include_feet_height_rewards) pf_ld2(B.current_max_feet_height + e * 2, pre.cmfh_old[0], pre.cmfh_old[1]);
    } else {
#pragma unroll
        for (int d = 0; d < PF_MAX_DOF; ++d) {
            const bool ok = d < D;
            pre.tq[d] = ok ? B.torques[e * D + d] : 0.0f; pre.ldv[d] = ok ? B.last_dof_vel[e * D + d] : 0.0f;
            pre.la[d] = ok ? B.last_actions[e * D + d] : 0.0f; pre.ac[d] = ok ? B.actions[e * D + d] : 0.0f;
        }
#pragma unroll
        for (int f = 0; f < PF_NUM_FEET; ++f) {
            const int k = e * PF_NUM_FEET + f;
            pre.fc_old[f] = B.first_contact[k] != 0; pre.cfilt_old[f] = B.contact_filt[k] != 0; pre.lc_old[f] = B.last_contacts[k] != 0;
            pre.fat_old[f] = B.feet_air_time[k]; pre.lfat_old[f] = B.last_feet_air_time[k];
            pre.lmfh_old[f] = B.last_max_feet_height[k];
            pre.cmfh_old[f] = P.include_feet_height_rewards ? B.current_max_feet_height[k] : 0.0f;
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) pre.cmd_old[i] = i < P.num_commands ? B.commands[e * P.num_commands + i] : 0.0f;
    pre.ep_len_old = B.episode_length_buf[e];
}

// post_physics_step up to (and including) reset_idx: PF:204-222.  `e` = local env index, writes go to B.
// kMask: PF_MASK_GENERIC, PF_MASK_TABLE or a literal term mask (see pf_canon_pos).  Everything but GENERIC is the
// common shape -- P.num_dof == PF_MAX_DOF and P.reward_ids in canonical order (pf_build_reward_table returned true
// for T) -> straight-line reward sum, no guards.
template <uint32_t kMask>
PF_HD void pf_env_step_t(const pf_params& P, const pf_env_buffers& B, const PfRewTab& T, int e, const PfEnvIn& in,
                         const PfEnvPre& pre, uint32_t step,
                         float* sums_out /* [R] scaled reward terms of this step or NULL */, PfEnvOut& out) {
    constexpr bool kFast = kMask != PF_MASK_GENERIC;
    constexpr bool kCanon = kFast;
    constexpr int kD = kFast ? PF_MAX_DOF : 0;
    const int D = kFast ? PF_MAX_DOF : P.num_dof;
    const float dt = P.dt;
    const int N = P.num_envs;
    const int R_terms = P.num_reward_terms;
    const float* es = pre.es;
    const float es_term = pre.es_term;
    const float *tq_ = pre.tq, *ldv_ = pre.ldv, *la_ = pre.la;
#pragma unroll
    for (int d = 0; d < PF_MAX_DOF; ++d) out.actions[d] = pre.ac[d];
    const bool *fc_old_ = pre.fc_old, *cfilt_old_ = pre.cfilt_old, *lc_old_ = pre.lc_old;
    const float *fat_old_ = pre.fat_old, *lfat_old_ = pre.lfat_old, *lmfh_old_ = pre.lmfh_old, *cmfh_old_ = pre.cmfh_old;
    const float* cmd_old_ = pre.cmd_old;
    // ---- PF:204 ----
    const long long ep_len = pre.ep_len_old + 1;

    // ---- PF:208-211 base quantities ----
    const float* q = in.root + 3;
    float lin_vel[3];
    const float grav[3] = {0.0f, 0.0f, -1.0f};
    pf_quat_rotate(q, in.root + 7, lin_vel, true);
    pf_quat_rotate(q, in.root + 10, out.ang_vel, true);
    pf_quat_rotate(q, grav, out.gravity, true);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        B.base_lin_vel[e * 3 + i] = lin_vel[i];
        B.base_ang_vel[e * 3 + i] = out.ang_vel[i];
        B.projected_gravity[e * 3 + i] = out.gravity[i];
    }

    // ---- PF:1114-1129 feet state machine ----
    float fat[PF_NUM_FEET], lfat[PF_NUM_FEET], lmfh[PF_NUM_FEET];
    float fh_[PF_NUM_FEET], cm_[PF_NUM_FEET];
    bool first_contact[PF_NUM_FEET], cfilt_[PF_NUM_FEET], contact_[PF_NUM_FEET];
#pragma unroll
    for (int f = 0; f < PF_NUM_FEET; ++f) {
        const bool fc_old = fc_old_[f];
        const bool cfilt_old = cfilt_old_[f];
        float fat_f = fat_old_[f];
        lfat[f] = fc_old ? fat_f : lfat_old_[f];
        fat_f = cfilt_old ? 0.0f : fat_f;
        lmfh[f] = lmfh_old_[f];
        if (P.include_feet_height_rewards) {
            float cm = cmfh_old_[f];
            lmfh[f] = fc_old ? cm : lmfh[f];
            cm = cfilt_old ? 0.0f : cm;
            float below = 0.0f;                                   // PF:1089-1090 plane
            if (P.terrain_kind == PF_TERRAIN_HEIGHTFIELD)
                below = (float)(kFast ? pf_terrain_min3_fast(P, T, B.height_samples, in.feet_pos[f][0], in.feet_pos[f][1])
                                      : pf_terrain_min3(P, B.height_samples, in.feet_pos[f][0], in.feet_pos[f][1]))
                        * P.vertical_scale;
            const float fh = in.feet_pos[f][2] - below;
            cm = pf_maxf(cm, fh);
            fh_[f] = fh; cm_[f] = cm;
        }
        const bool contact = in.cf[P.feet_indices[f] * 3 + 2] > 1.0f;
        const bool cfilt = contact || lc_old_[f];
        first_contact[f] = (fat_f > 0.0f) && cfilt;
        fat_f = fat_f + dt;
        fat[f] = fat_f;
        cfilt_[f] = cfilt; contact_[f] = contact;
    }
    if (kFast && PF_NUM_FEET == 2) {
        if (P.include_feet_height_rewards) {
            pf_st2(B.feet_height + e * 2, fh_[0], fh_[1]);
            pf_st2(B.current_max_feet_height + e * 2, cm_[0], cm_[1]);
            pf_st2(B.last_max_feet_height + e * 2, lmfh[0], lmfh[1]);
        }
        pf_stb2(B.contact_filt + e * 2, cfilt_[0], cfilt_[1]);
        pf_stb2(B.last_contacts + e * 2, contact_[0], contact_[1]);
        pf_stb2(B.first_contact + e * 2, first_contact[0], first_contact[1]);
        pf_st2(B.feet_air_time + e * 2, fat[0], fat[1]);
        pf_st2(B.last_feet_air_time + e * 2, lfat[0], lfat[1]);
    } else {
#pragma unroll
        for (int f = 0; f < PF_NUM_FEET; ++f) {
            const int k = e * PF_NUM_FEET + f;
            if (P.include_feet_height_rewards) {
                B.feet_height[k] = fh_[f]; B.current_max_feet_height[k] = cm_[f]; B.last_max_feet_height[k] = lmfh[f];
            }
            B.contact_filt[k] = cfilt_[f]; B.last_contacts[k] = contact_[f]; B.first_contact[k] = first_contact[f];
            B.feet_air_time[k] = fat[f]; B.last_feet_air_time[k] = lfat[f];
        }
    }

    // ---- PF:479-493 callback: commands ----
    const int C = P.num_commands;
#pragma unroll
    for (int i = 0; i < 4; ++i) out.commands[i] = cmd_old_[i];
    const uint32_t genv = (uint32_t)(P.env_id_offset + e);
    if (P.resample_commands && (ep_len % P.resample_period) == 0) {       // LR:337-351
        const uint32_t ek = pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_RESAMPLE), genv);
        out.commands[0] = (P.cmd_range[0][1] - P.cmd_range[0][0]) * pf_rng_uniform(ek, 0) + P.cmd_range[0][0];
        out.commands[1] = (P.cmd_range[1][1] - P.cmd_range[1][0]) * pf_rng_uniform(ek, 1) + P.cmd_range[1][0];
        if (P.heading_command)
            out.commands[3] = (P.cmd_range[3][1] - P.cmd_range[3][0]) * pf_rng_uniform(ek, 2) + P.cmd_range[3][0];   // the THIRD draw of LR:346-349, heading or yaw: lane 2 either way
        else
            out.commands[2] = (P.cmd_range[2][1] - P.cmd_range[2][0]) * pf_rng_uniform(ek, 2) + P.cmd_range[2][0];
        const float keep = pf_norm2(out.commands[0], out.commands[1]) > 0.2f ? 1.0f : 0.0f;
        out.commands[0] *= keep; out.commands[1] *= keep;
    }
    if (P.heading_command) {                                              // PF:487-490
        const float fwd_b[3] = {1.0f, 0.0f, 0.0f};
        float fwd[3];
        pf_quat_apply(q, fwd_b, fwd);
        const float heading = atan2f(fwd[1], fwd[0]);
        out.commands[2] = pf_clipf(0.5f * pf_wrap_to_pi(out.commands[3] - heading), -1.0f, 1.0f);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) if (i < C) B.commands[e * C + i] = out.commands[i];

    // ---- PF:239-247 termination ----
    bool terminate = false;
#pragma unroll
    for (int i = 0; i < PF_MAX_BODIES; ++i)
        if (i < P.num_termination) {
            const float* F = in.cf + P.termination_indices[i] * 3;
            terminate = terminate || (kFast ? pf_sumsq3(F[0], F[1], F[2]) >= T.sq_gt_termination
                                            : pf_norm3(F[0], F[1], F[2]) > 1.0f);
        }
    const bool time_out = (float)ep_len > P.max_episode_length;
    const bool reset = terminate || time_out;
    B.terminate[e] = terminate; B.time_out_buf[e] = time_out; B.reset_buf[e] = reset;
    out.reset = reset;

    // ---- PF:301-318 rewards, summed in cfg (alphabetical) order ----
    PfRewIn c;
    c.lin_vel = lin_vel; c.ang_vel = out.ang_vel; c.gravity = out.gravity; c.commands = out.commands;
    c.tq = tq_; c.ldv = ldv_; c.la = la_; c.ac = out.actions; c.dof = in.dof; c.cf = in.cf;
    c.fat = fat; c.lfat = lfat; c.lmfh = lmfh; c.first_contact = first_contact;
    c.feet0 = in.feet_pos[0]; c.feet1 = in.feet_pos[1];
    c.bh = in.have_heights ? in.height_mean : in.root[2];
    c.reset = reset; c.time_out = time_out;
    float rew = 0.0f;
    if (kCanon) {
        float* es_out = B.episode_sums + e;
        // One pass over the canonical order.  Every comparison is warp uniform (constant bank), every term's code
        // is specialised, every es[] index is a compile-time constant.
#define PF_X(ID)                                                                                     \
        if (pf_term_on<kMask, ID>(T)) {                                                              \
            const float r = pf_reward_term<kD>(P, T, (ID), c) * T.scale[ID];                         \
            rew += r;                                                                                \
            const float s = es[ID] + r;                                                              \
            es_out[pf_term_off<kMask, ID>(T, N)] = s;                                                \
            if (sums_out) sums_out[pf_term_row<kMask, ID>(T)] = s;                                   \
        }
        PF_REWARD_CANON(PF_X)
#undef PF_X
    } else {
        for (int k = 0; k < R_terms; ++k) {
            const float r = pf_reward_term<0>(P, T, P.reward_ids[k], c) * P.reward_scales[k];
            rew += r;
            const float s = B.episode_sums[(size_t)k * N + e] + r;
            B.episode_sums[(size_t)k * N + e] = s;
            if (sums_out) sums_out[k] = s;
        }
    }
    if (P.only_positive_rewards) rew = pf_maxf(rew, 0.0f);
    if (P.has_termination_term) {
        const int k = R_terms;
        const float r = ((reset && !time_out) ? 1.0f : 0.0f) * P.termination_scale;
        rew += r;
        const float s = es_term + r;
        B.episode_sums[(size_t)k * N + e] = s;
        if (sums_out) sums_out[k] = s;
    }
    B.rew_buf[e] = rew;
    B.episode_length_buf[e] = ep_len;

#pragma unroll
    for (int d = 0; d < PF_MAX_DOF; ++d) {
        out.dof_pos[d] = d < D ? in.dof[2 * d] : 0.0f; out.dof_vel[d] = d < D ? in.dof[2 * d + 1] : 0.0f;
    }
    out.root_z = in.root[2];
#pragma unroll
    for (int i = 0; i < 6; ++i) out.root_vel[i] = in.root[7 + i];
}

// `mode`: 0 generic, 1 table, 2 the literal pointfoot mask (the caller checked that it applies)
PF_HD void pf_env_step(const pf_params& P, const pf_env_buffers& B, const PfRewTab& T, int e, const PfEnvIn& in,
                       uint32_t step, float* sums_out, PfEnvOut& out, int mode) {
    PfEnvPre pre;
    if (mode == 2) { pf_env_preload<PF_MASK_POINTFOOT>(P, B, T, e, pre); pf_env_step_t<PF_MASK_POINTFOOT>(P, B, T, e, in, pre, step, sums_out, out); }
    else if (mode == 1) { pf_env_preload<PF_MASK_TABLE>(P, B, T, e, pre); pf_env_step_t<PF_MASK_TABLE>(P, B, T, e, in, pre, step, sums_out, out); }
    else { pf_env_preload<PF_MASK_GENERIC>(P, B, T, e, pre); pf_env_step_t<PF_MASK_GENERIC>(P, B, T, e, in, pre, step, sums_out, out); }
}

// which statement of the step applies to P (and env overrides for tests: force the table / the generic one)
static inline int pf_step_mode(const pf_params& P, PfRewTab& T, bool force_generic, bool force_table) {
    const bool canon = pf_build_reward_table(P, T) && P.num_dof == PF_MAX_DOF;
    if (!canon || force_generic) return 0;
    if (force_table || pf_reward_mask(P) != PF_MASK_POINTFOOT) return 1;
    return 2;
}

// ------------------------------------------------------------------------------------------------
// reset_idx for ONE env (PF:249-299, 549-588, 613-636).  Writes the reset rows in place into root_states /
// dof_state (PhysX-owned), updates the curriculum, zeroes the per-env buffers and the episode sums.
// `sums_out[R]` receives the episode sums before they are zeroed (for extras["episode"], PF:276-280).
// `root_pre` = this env's root row BEFORE the reset (the curriculum measures the distance walked).
template <int kD>
PF_HD void pf_env_reset_t(const pf_params& P, const pf_env_buffers& B, int e, const float* root_pre,
                          const float cmd_xy[2], uint32_t step, bool update_curriculum,
                          float* sums_out, PfEnvOut& out) {
    const int D = kD ? kD : P.num_dof, N = P.num_envs;
    const uint32_t genv = (uint32_t)(P.env_id_offset + e);

    // ---- PF:613-636 terrain curriculum ----
    if (P.terrain_curriculum && update_curriculum) {
        const float dist = pf_norm2(root_pre[0] - B.env_origins[e * 3 + 0], root_pre[1] - B.env_origins[e * 3 + 1]);
        const bool up = dist > P.terrain_env_length / 2.0f;
        const bool down = (dist < (pf_norm2(cmd_xy[0], cmd_xy[1]) * P.max_episode_length_s) * 0.5f) && !up;
        long long lvl = B.terrain_levels[e] + (up ? 1 : 0) - (down ? 1 : 0);
        if (lvl >= P.max_terrain_level) {
            const uint32_t ek = pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_CURRICULUM), genv);
            lvl = (long long)(pf_rng_u32(ek, 0) % (uint32_t)P.max_terrain_level);
        } else if (lvl < 0) {
            lvl = 0;
        }
        B.terrain_levels[e] = lvl;
        const float* o = B.terrain_origins + ((size_t)lvl * P.num_terrain_types + B.terrain_types[e]) * 3;
        for (int i = 0; i < 3; ++i) B.env_origins[e * 3 + i] = o[i];
    }

    // ---- PF:549-559 dofs: pos = default * U(0.5, 1.5), vel = 0 ----
    {
        const uint32_t ek = pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_RESET_DOF), genv);
#pragma unroll
        for (int d = 0; d < PF_MAX_DOF; ++d)
            if (d < D) {
                const float u = (1.5f - 0.5f) * pf_rng_uniform(ek, d) + 0.5f;
                const float p = P.default_dof_pos[d] * u;
                B.dof_state[((size_t)e * D + d) * 2 + 0] = p;
                B.dof_state[((size_t)e * D + d) * 2 + 1] = 0.0f;
                out.dof_pos[d] = p; out.dof_vel[d] = 0.0f;
            }
    }
    // ---- PF:566-584 root state ----
    {
        const uint32_t ek = pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_RESET_ROOT), genv);
        float r[13];
#pragma unroll
        for (int i = 0; i < 13; ++i) r[i] = P.base_init_state[i];
#pragma unroll
        for (int i = 0; i < 3; ++i) r[i] = r[i] + B.env_origins[e * 3 + i];
        if (P.custom_origins) {
#pragma unroll
            for (int i = 0; i < 2; ++i) r[i] = r[i] + ((1.0f - (-1.0f)) * pf_rng_uniform(ek, i) + (-1.0f));
        }
#pragma unroll
        for (int i = 0; i < 6; ++i) r[7 + i] = (0.5f - (-0.5f)) * pf_rng_uniform(ek, 2 + i) + (-0.5f);
#pragma unroll
        for (int i = 0; i < 13; ++i) B.root_states[(size_t)e * 13 + i] = r[i];
        out.root_z = r[2];
#pragma unroll
        for (int i = 0; i < 6; ++i) out.root_vel[i] = r[7 + i];
    }
    // ---- PF:290-299 buffers ----
#pragma unroll
    for (int d = 0; d < PF_MAX_DOF; ++d) if (d < D) { B.last_actions[e * D + d] = 0.0f; B.last_dof_vel[e * D + d] = 0.0f; }
#pragma unroll
    for (int f = 0; f < PF_NUM_FEET; ++f) {
        const int k = e * PF_NUM_FEET + f;
        B.feet_air_time[k] = 0.0f; B.last_feet_air_time[k] = 0.0f;
        B.current_max_feet_height[k] = 0.0f; B.last_max_feet_height[k] = 0.0f;
    }
    B.episode_length_buf[e] = 0;
    B.reset_buf[e] = 1;
    // ---- PF:276-280 episode sums ----
    const int R = P.num_reward_terms + (P.has_termination_term ? 1 : 0);
    for (int k = 0; k < R; ++k) {
        if (sums_out) sums_out[k] = B.episode_sums[(size_t)k * N + e];
        B.episode_sums[(size_t)k * N + e] = 0.0f;
    }
}

PF_HD void pf_env_reset(const pf_params& P, const pf_env_buffers& B, int e, const float* root_pre,
                        const float cmd_xy[2], uint32_t step, bool update_curriculum,
                        float* sums_out, PfEnvOut& out) {
    pf_env_reset_t<0>(P, B, e, root_pre, cmd_xy, step, update_curriculum, sums_out, out);
}

// Number of leading (non height-scan) observation columns: PF:376-382.
PF_HD int pf_obs_base_width(const pf_params& P) { return 6 + 3 * P.num_dof + 3; }

// Column j (< base width) of the observation row before noise, PF:376-382.
PF_HD float pf_obs_base_col(const pf_params& P, const PfEnvOut& o, const float* actions, int j) {
    const int D = P.num_dof;
    if (j < 3) return o.ang_vel[j] * P.obs_scale_ang_vel;
    if (j < 6) return o.gravity[j - 3];
    j -= 6;
    if (j < D) return (o.dof_pos[j] - P.default_dof_pos[j]) * P.obs_scale_dof_pos;
    j -= D;
    if (j < D) return o.dof_vel[j] * P.obs_scale_dof_vel;
    j -= D;
    if (j < D) return actions[j];
    j -= D;
    return o.commands[j] * P.commands_scale[j];
}

// The same row written in one go with compile-time indices (the kernel's statement: `o` stays in registers).
template <int kD>
PF_HD void pf_obs_base_row(const pf_params& P, const PfEnvOut& o, float* row) {
    const int D = kD ? kD : P.num_dof;
#pragma unroll
    for (int i = 0; i < 3; ++i) { row[i] = o.ang_vel[i] * P.obs_scale_ang_vel; row[3 + i] = o.gravity[i]; }
#pragma unroll
    for (int d = 0; d < PF_MAX_DOF; ++d)
        if (d < D) {
            row[6 + d] = (o.dof_pos[d] - P.default_dof_pos[d]) * P.obs_scale_dof_pos;
            row[6 + D + d] = o.dof_vel[d] * P.obs_scale_dof_vel;
            row[6 + 2 * D + d] = o.actions[d];
        }
#pragma unroll
    for (int i = 0; i < 3; ++i) row[6 + 3 * D + i] = o.commands[i] * P.commands_scale[i];
}

// Height-scan observation column, PF:367-369: clip(root_z - 0.5 - h, -1, 1) * scale
PF_HD float pf_obs_height_col(const pf_params& P, float root_z, int16_t raw) {
    const float h = (float)raw * P.vertical_scale;
    return pf_clipf((root_z - 0.5f) - h, -1.0f, 1.0f) * P.obs_scale_height;
}

// _add_noise_to_obs (PF:328-338) for column `lane` of either buffer, then the clip of PF:189-192.
PF_HD float pf_obs_finish(const pf_params& P, float v, uint32_t noise_env_key, int lane,
                          const float* __restrict__ noise_vec) {
    if (P.add_noise) {
        const float s = noise_vec[lane];
        if (s != 0.0f) v = v + (2.0f * pf_rng_uniform(noise_env_key, lane) - 1.0f) * s;
    }
    return pf_clipf(v, -P.clip_obs, P.clip_obs);
}

// _compute_torques for one DOF, PF:523-547 (LR:353-375)
PF_HD float pf_pd_torque(const pf_params& P, int d, float action, float q, float qd, float last_qd) {
    const float as = action * P.action_scale;
    float t;
    if (P.control_type == PF_CONTROL_P)
        t = P.p_gains[d] * ((as + P.default_dof_pos[d]) - q) - P.d_gains[d] * qd;
    else if (P.control_type == PF_CONTROL_V)
        t = P.p_gains[d] * (as - qd) - (P.d_gains[d] * (qd - last_qd)) / P.sim_dt;
    else
        t = as;
    return pf_clipf(t, -P.torque_limits[d], P.torque_limits[d]);
}

// _push_robots for one env, PF:590-611: force on body 0 = quat_rotate(base_quat, U(-F, F)^3), z * 0.5
PF_HD void pf_push_force(const pf_params& P, int e, const float* quat, float max_push_force, uint32_t step,
                         float out[3]) {
    const uint32_t ek = pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_PUSH), (uint32_t)(P.env_id_offset + e));
    float f[3];
    for (int i = 0; i < 3; ++i)
        f[i] = (max_push_force - (-max_push_force)) * pf_rng_uniform(ek, i) + (-max_push_force);
    pf_quat_rotate(quat, f, out, false);
    out[2] = out[2] * 0.5f;
}
