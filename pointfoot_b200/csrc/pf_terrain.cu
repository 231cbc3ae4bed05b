// Terrain generation: legged_gym/utils/terrain.py::Terrain + the isaacgym.terrain_utils primitives
// it calls, as raster kernels (include/pointfoot_b200.h, "Terrain generation").
//   pf_terrain_cells_kernel    one CTA per (map cell, 16-row strip), one thread per pixel: the sub-terrain's closed form
//                              (pyramid, stairs, gap, pit) or its random structure re-derived from the counter-based contract
//                              (noise grid nodes and stone bands per pixel, the 20 rectangles once per CTA into shared memory)
//                              -- no sub-terrain buffers, no paste step (terrain.py:147-158), int16 stores along the map's columns
//   pf_terrain_origins_kernel  one warp per sub-terrain: max over the centre patch -> env_origins (terrain.py:160-167)
//   pf_terrain_vertices_kernel / pf_terrain_triangles_kernel   convert_heightfield_to_trimesh (terrain.py:66-70)
// Init-time work in the reference (numpy, ~2.4 s for the default 1300 x 2100 map); here HBM bound: 2 B written per pixel,
// the per-pixel arithmetic (<= 100 hashes for an obstacle cell) hides under the store stream.  Compiled with -fmad=false.
#include <cuda_runtime.h>
#include <stdio.h>
#include "pf_terrain_core.cuh"
#include "pf_launch.h"

void pf_set_error(const char* msg);

#define PF_T_STRIP 16          // rows of a cell one CTA rasterises
#define PF_T_MAX_RECTS 64     // rectangles of an obstacle cell staged in shared memory (the reference paints 20)

// One CTA per (cell, strip of PF_T_STRIP rows).  The structure shared by the cell's pixels is derived ONCE per CTA: the
// rectangles of an obstacle cell (5 draws each; a pixel then compares against <= 20 staged rectangles instead of re-deriving
// 100 draws) -- everything else goes through the same pf_terrain_pixel the host emulation runs.
__global__ void __launch_bounds__(256) pf_terrain_cells_kernel(const pf_terrain_params P, const pf_subterrain* __restrict__ subs,
                                                               int16_t* __restrict__ hf, const int strips) {
    __shared__ int s_rect[PF_T_MAX_RECTS][5];
    __shared__ pf_subterrain s_sub;
    const int sub = blockIdx.x / strips, strip = blockIdx.x - sub * strips, W = P.cell_px;
    if (threadIdx.x < 16) reinterpret_cast<int32_t*>(&s_sub)[threadIdx.x] = reinterpret_cast<const int32_t*>(subs + sub)[threadIdx.x];
    __syncthreads();
    const pf_subterrain& S = s_sub;
    const uint32_t ek = pf_rng_env_key(pf_rng_key(P.seed, 0u, PF_RNG_TERRAIN), (uint32_t)sub);
    const bool staged = S.kind == PF_SUB_OBSTACLES && S.p[3] <= PF_T_MAX_RECTS;
    if (staged && (int)threadIdx.x < S.p[3]) {
        const int r = threadIdx.x, mh = S.p[0], nw = pf_t_ceil_div(S.p[2] - S.p[1], 4);
        const uint32_t l = 5u * (uint32_t)r;
        const int w = S.p[1] + 4 * pf_t_draw(ek, l, nw), len = S.p[1] + 4 * pf_t_draw(ek, l + 1, nw);
        const int k = pf_t_draw(ek, l + 4, 4);
        s_rect[r][0] = 4 * pf_t_draw(ek, l + 2, pf_t_ceil_div(W - w, 4));
        s_rect[r][1] = 4 * pf_t_draw(ek, l + 3, pf_t_ceil_div(W - len, 4));
        s_rect[r][2] = w; s_rect[r][3] = len;
        s_rect[r][4] = k == 0 ? -mh : (k == 1 ? pf_t_floor_div2(-mh) : (k == 2 ? pf_t_floor_div2(mh) : mh));
    }
    __syncthreads();
    const int i = sub / P.num_cols, j = sub - i * P.num_cols;
    const int x0 = strip * PF_T_STRIP, nx = min(PF_T_STRIP, W - x0);
    int16_t* base = hf + (size_t)(P.border_px + i * W) * P.tot_cols + (P.border_px + j * W);
    for (int k = threadIdx.x; k < nx * W; k += blockDim.x) {
        const int x = x0 + k / W, y = k % W;
        int16_t h;
        if (staged) {
            int v = 0;
            for (int r = S.p[3] - 1; r >= 0; --r)
                if (x >= s_rect[r][0] && x < s_rect[r][0] + s_rect[r][2] && y >= s_rect[r][1] && y < s_rect[r][1] + s_rect[r][3]) { v = s_rect[r][4]; break; }
            if (pf_t_in(x, y, S.p + 4)) v = 0;
            h = (int16_t)v;
        } else {
            h = pf_terrain_pixel(S, ek, W, x, y);
        }
        base[(size_t)x * P.tot_cols + y] = h;
    }
}

// terrain.py:160-167: env_origins[i, j] = [(i + 0.5) * env_length, (j + 0.5) * env_width, max(patch) * vertical_scale]
__global__ void __launch_bounds__(256) pf_terrain_origins_kernel(const pf_terrain_params P, const int16_t* __restrict__ hf,
                                                                 double* __restrict__ origins) {
    const int lane = threadIdx.x & 31;
    const int sub = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (sub >= P.num_rows * P.num_cols) return;
    const int i = sub / P.num_cols, j = sub - i * P.num_cols, W = P.cell_px;
    const int x1 = P.origin_patch[0], x2 = P.origin_patch[1], y1 = P.origin_patch[2], y2 = P.origin_patch[3];
    const int pw = y2 - y1, np_ = (x2 - x1) * pw;
    int m = -32768;
    for (int k = lane; k < np_; k += 32) {
        const int x = x1 + k / pw, y = y1 + k % pw;
        const int v = hf[(size_t)(P.border_px + i * W + x) * P.tot_cols + (P.border_px + j * W + y)];
        m = v > m ? v : m;
    }
    for (int o = 16; o > 0; o >>= 1) { const int t = __shfl_xor_sync(0xffffffffu, m, o); m = t > m ? t : m; }
    if (lane == 0) {
        origins[sub * 3 + 0] = ((double)i + 0.5) * P.env_length;
        origins[sub * 3 + 1] = ((double)j + 0.5) * P.env_width;
        origins[sub * 3 + 2] = (double)m * P.vertical_scale;
    }
}

__global__ void __launch_bounds__(256) pf_terrain_vertices_kernel(const int16_t* __restrict__ hf, int R, int Cn, double hs, double vs,
                                                                  double thr, float* __restrict__ vertices) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)R * Cn) return;
    const int r = (int)(i / Cn), c = (int)(i - (long long)r * Cn);
    float v[3];
    pf_terrain_vertex(hf, R, Cn, hs, vs, thr, r, c, v);
    vertices[i * 3 + 0] = v[0]; vertices[i * 3 + 1] = v[1]; vertices[i * 3 + 2] = v[2];
}

// two triangles per cell (r, c): (ind0, ind3, ind1) and (ind0, ind2, ind3), ind0 = r * cols + c
__global__ void __launch_bounds__(256) pf_terrain_triangles_kernel(int R, int Cn, uint32_t* __restrict__ tri) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)(R - 1) * (Cn - 1)) return;
    const int r = (int)(i / (Cn - 1)), c = (int)(i - (long long)r * (Cn - 1));
    const uint32_t i0 = (uint32_t)(r * Cn + c), i1 = i0 + 1, i2 = i0 + (uint32_t)Cn, i3 = i2 + 1;
    uint32_t* t = tri + i * 6;
    t[0] = i0; t[1] = i3; t[2] = i1; t[3] = i0; t[4] = i2; t[5] = i3;
}

static int t_fail(int code, const char* msg) { pf_set_error(msg); return code; }
#define T_CUDA(call) do { if ((call) != cudaSuccess) return t_fail(2, #call " failed"); } while (0)

extern "C" int pf_terrain_sizeof_params(void) { return (int)sizeof(pf_terrain_params); }

extern "C" int pf_terrain_generate(const pf_terrain_params* p, const pf_subterrain* subs, int16_t* height_field, double* env_origins,
                                   void* stream) {
    if (!p || !subs || !height_field || !env_origins) return t_fail(1, "pf_terrain_generate: null argument");
    if (p->abi_version != PF_ABI_VERSION) return t_fail(1, "pf_terrain_generate: ABI version mismatch");
    if (p->num_rows <= 0 || p->num_cols <= 0 || p->cell_px < 4 || p->border_px < 0 ||
        p->tot_rows != p->num_rows * p->cell_px + 2 * p->border_px || p->tot_cols != p->num_cols * p->cell_px + 2 * p->border_px)
        return t_fail(1, "pf_terrain_generate: inconsistent map geometry");
    const int* o = p->origin_patch;
    if (o[0] < 0 || o[1] > p->cell_px || o[0] >= o[1] || o[2] < 0 || o[3] > p->cell_px || o[2] >= o[3])
        return t_fail(1, "pf_terrain_generate: origin patch outside the sub-terrain");
    cudaStream_t st = (cudaStream_t)stream;
    const int nsub = p->num_rows * p->num_cols, strips = (p->cell_px + PF_T_STRIP - 1) / PF_T_STRIP;
    // the border (terrain.py:62 np.zeros) and then the cells
    T_CUDA(cudaMemsetAsync(height_field, 0, sizeof(int16_t) * (size_t)p->tot_rows * p->tot_cols, st));
    PF_LAUNCHED(), pf_terrain_cells_kernel<<<nsub * strips, 256, 0, st>>>(*p, subs, height_field, strips);
    T_CUDA(cudaGetLastError());
    PF_LAUNCHED(), pf_terrain_origins_kernel<<<(nsub * 32 + 255) / 256, 256, 0, st>>>(*p, height_field, env_origins);
    T_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int pf_terrain_trimesh(const int16_t* height_field, int32_t rows, int32_t cols, double horizontal_scale, double vertical_scale,
                                  double slope_threshold, float* vertices, uint32_t* triangles, void* stream) {
    if (!height_field || !vertices || !triangles) return t_fail(1, "pf_terrain_trimesh: null argument");
    if (rows < 2 || cols < 2 || (long long)rows * cols > 0x7fffffffLL) return t_fail(1, "pf_terrain_trimesh: size out of range");
    cudaStream_t st = (cudaStream_t)stream;
    const long long nv = (long long)rows * cols, nc = (long long)(rows - 1) * (cols - 1);
    PF_LAUNCHED(), pf_terrain_vertices_kernel<<<(unsigned)((nv + 255) / 256), 256, 0, st>>>(height_field, rows, cols, horizontal_scale,
                                                                                            vertical_scale, slope_threshold, vertices);
    T_CUDA(cudaGetLastError());
    PF_LAUNCHED(), pf_terrain_triangles_kernel<<<(unsigned)((nc + 255) / 256), 256, 0, st>>>(rows, cols, triangles);
    T_CUDA(cudaGetLastError());
    return 0;
}
