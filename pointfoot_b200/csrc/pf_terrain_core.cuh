// Per-pixel statements of the terrain raster (see pf_terrain.cu): host + device, so that tests/host_emu can drive the SAME
// code serially in a container without a GPU.  Heights are int16 counts of vertical_scale; the float64 expressions follow
// numpy's evaluation order (IEEE double, no contraction: compile with -fmad=false / -ffp-contract=off).
#pragma once
#include <stdint.h>
#include <math.h>
#include "../../include/pointfoot_b200.h"
#include "pf_rng.cuh"

#ifdef __CUDACC__
#define PF_THD __host__ __device__ __forceinline__
#else
#define PF_THD static inline
#endif

PF_THD int pf_t_min(int a, int b) { return a < b ? a : b; }
PF_THD int pf_t_max(int a, int b) { return a > b ? a : b; }
PF_THD int pf_t_ceil_div(int a, int b) { return a > 0 ? (a + b - 1) / b : 0; }          // len(range(0, a, b))
PF_THD int pf_t_floor_div2(int a) { return a >= 0 ? a / 2 : -((-a + 1) / 2); }           // python a // 2
PF_THD bool pf_t_in(int x, int y, const int32_t* r) { return x >= r[0] && x < r[1] && y >= r[2] && y < r[3]; }

// one uniform integer in [0, n): u32 % n, as oracle/ref_harness.py::build_reference_terrain's hook
PF_THD int pf_t_draw(uint32_t ek, uint32_t lane, int n) { return (int)(pf_rng_u32(ek, lane) % (uint32_t)n); }

// pyramid_sloped_terrain before the clip: (max_height * xx * yy).astype(int16), xx = (c - |c - x|) / c
PF_THD int pf_t_pyramid(int max_height, int W, int x, int y) {
    const int c = W / 2;                                           // int(width / 2)
    const double xx = (double)(c - abs(c - x)) / (double)c;
    const double yy = (double)(c - abs(c - y)) / (double)c;
    return (int)(int16_t)(long long)(((double)max_height * xx) * yy);   // float64 -> int16: truncation toward zero
}

// Height of pixel (x, y) of a W x W sub-terrain described by S; ek = RNG key of the sub-terrain.
PF_THD int16_t pf_terrain_pixel(const pf_subterrain& S, uint32_t ek, int W, int x, int y) {
    const int32_t* p = S.p;
    int h = 0;
    switch (S.kind) {
    case PF_SUB_SLOPE:
    case PF_SUB_NOISY_SLOPE: {
        h = pf_t_pyramid(p[0], W, x, y);
        const int c1 = W / 2 - p[1];                               // x1 = y1 = width // 2 - platform_size
        const int h0 = pf_t_pyramid(p[0], W, c1, c1);
        h = pf_t_min(pf_t_max(h, pf_t_min(h0, 0)), pf_t_max(h0, 0));   // np.clip(field, min(h0, 0), max(h0, 0))
        if (S.kind == PF_SUB_NOISY_SLOPE) {                    // random_uniform_terrain: bilinear upsampling of an nx x ny grid
            const int nx = p[5], ny = p[6];
            const double tx = (double)(x * (nx - 1)) / (double)(W - 1), ty = (double)(y * (ny - 1)) / (double)(W - 1);
            const int ix = pf_t_min((int)tx, nx - 2), iy = pf_t_min((int)ty, ny - 2);
            const double fx = tx - (double)ix, fy = ty - (double)iy;
            double g[2][2];
            for (int a = 0; a < 2; ++a)
                for (int b = 0; b < 2; ++b)
                    g[a][b] = (double)(p[2] + p[4] * pf_t_draw(ek, (uint32_t)((ix + a) * ny + (iy + b)), p[3]));
            const double z = (g[0][0] * (1.0 - fx) + g[1][0] * fx) * (1.0 - fy) + (g[0][1] * (1.0 - fx) + g[1][1] * fx) * fy;
            h = (int)(int16_t)(h + (int)(int16_t)(long long)rint(z));      // int16 += rint(z).astype(int16)
        }
    } break;
    case PF_SUB_STAIRS: {
        const int sw = p[0];
        int t = pf_t_min(pf_t_min(x / sw, (W - 1 - x) / sw), pf_t_min(y / sw, (W - 1 - y) / sw));
        t = pf_t_min(t, p[2]);
        h = t * p[1];
    } break;
    case PF_SUB_OBSTACLES: {
        const int mh = p[0], nw = pf_t_ceil_div(p[2] - p[1], 4);
        for (int r = p[3] - 1; r >= 0; --r) {                      // the LAST rectangle painted over a pixel wins
            const uint32_t l = 5u * (uint32_t)r;
            const int w = p[1] + 4 * pf_t_draw(ek, l, nw), len = p[1] + 4 * pf_t_draw(ek, l + 1, nw);
            const int si = 4 * pf_t_draw(ek, l + 2, pf_t_ceil_div(W - w, 4)), sj = 4 * pf_t_draw(ek, l + 3, pf_t_ceil_div(W - len, 4));
            if (x >= si && x < si + w && y >= sj && y < sj + len) {
                const int k = pf_t_draw(ek, l + 4, 4);             // [-mh, -mh // 2, mh // 2, mh]
                h = k == 0 ? -mh : (k == 1 ? pf_t_floor_div2(-mh) : (k == 2 ? pf_t_floor_div2(mh) : mh));
                break;
            }
        }
        if (pf_t_in(x, y, p + 4)) h = 0;
    } break;
    case PF_SUB_STONES: {
        const int ss = p[0], period = p[0] + p[1], nh = 2 * p[2] + 1, lo = -p[2] - 1;   // height_range = arange(-mh - 1, mh)
        h = p[3];
        const int band = y / period;
        if (y - band * period < ss) {
            const uint32_t l = 128u * (uint32_t)band;
            const int sx = pf_t_draw(ek, l, ss);
            if (x < pf_t_max(0, sx - p[1])) h = lo + pf_t_draw(ek, l + 1, nh);
            else if (x >= sx) {
                const int s = (x - sx) / period;
                if ((x - sx) - s * period < ss) h = lo + pf_t_draw(ek, l + 2 + (uint32_t)s, nh);
            }
        }
        if (pf_t_in(x, y, p + 4)) h = 0;
    } break;
    case PF_SUB_GAP:
        if (pf_t_in(x, y, p + 0)) h = -1000;
        if (pf_t_in(x, y, p + 4)) h = 0;
        break;
    case PF_SUB_PIT:
        if (pf_t_in(x, y, p + 4)) h = p[0];
        break;
    default: break;
    }
    return (int16_t)h;
}

// pixel (r, c) of the whole map (terrain.py:147-158): border zeros around num_rows x num_cols cells of W x W
PF_THD int16_t pf_terrain_map_pixel(const pf_terrain_params& P, const pf_subterrain* subs, int r, int c) {
    const int W = P.cell_px, x = r - P.border_px, y = c - P.border_px;
    if (x < 0 || y < 0) return 0;
    const int i = x / W, j = y / W;
    if (i >= P.num_rows || j >= P.num_cols) return 0;
    const uint32_t sub = (uint32_t)(i * P.num_cols + j);
    const uint32_t ek = pf_rng_env_key(pf_rng_key(P.seed, 0u, PF_RNG_TERRAIN), sub);
    return pf_terrain_pixel(subs[sub], ek, W, x - i * W, y - j * W);
}

// np.linspace(0, (n - 1) * scale, n)[k]
PF_THD double pf_t_linspace(int k, int n, double scale) {
    const double stop = (double)(n - 1) * scale;
    if (n == 1) return 0.0;
    if (k == n - 1) return stop;
    return (double)k * (stop / (double)(n - 1));
}

// convert_heightfield_to_trimesh: vertex (r, c) with the slope-threshold correction
PF_THD void pf_terrain_vertex(const int16_t* hf, int R, int Cn, double hs, double vs, double thr, int r, int c, float* v) {
    double x = pf_t_linspace(r, R, hs), y = pf_t_linspace(c, Cn, hs);
    const int16_t h = hf[(size_t)r * Cn + c];
    if (thr >= 0.0) {
        const double th = thr * (hs / vs);                          // slope_threshold *= horizontal_scale / vertical_scale
        auto d = [&](int r1, int c1) { return (double)(int16_t)(hf[(size_t)r1 * Cn + c1] - h); };   // int16 difference
        double mx = 0.0, my = 0.0, mc = 0.0;
        if (r < R - 1 && d(r + 1, c) > th) mx += 1.0;
        if (r > 0 && d(r - 1, c) > th) mx -= 1.0;
        if (c < Cn - 1 && d(r, c + 1) > th) my += 1.0;
        if (c > 0 && d(r, c - 1) > th) my -= 1.0;
        if (r < R - 1 && c < Cn - 1 && d(r + 1, c + 1) > th) mc += 1.0;
        if (r > 0 && c > 0 && d(r - 1, c - 1) > th) mc -= 1.0;
        x += (mx + mc * (mx == 0.0 ? 1.0 : 0.0)) * hs;
        y += (my + mc * (my == 0.0 ? 1.0 : 0.0)) * hs;
    }
    v[0] = (float)x; v[1] = (float)y; v[2] = (float)((double)h * vs);
}
