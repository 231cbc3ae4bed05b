"""TEST INFRASTRUCTURE: CPU restatement of ``isaacgym.terrain_utils`` -- the third-party module
``legged_gym/utils/terrain.py:34`` imports for the sub-terrain primitives.

PARITY UNPINNED for this file: Isaac Gym (Preview 4, setup.py pins ``isaacgym``; a closed download) is absent from
/root/reference and from this image, so the primitives below restate the module's PUBLISHED algorithm
(python/isaacgym/terrain_utils.py of the Preview-4 package, also carried verbatim by NVIDIA's open IsaacGymEnvs repository)
from its documented behaviour; nothing here could be checked against the original file.  What IS pinned is everything the
reference itself owns: ``Terrain.__init__`` / ``curiculum`` / ``randomized_terrain`` / ``make_terrain`` /
``add_terrain_to_map`` / ``gap_terrain`` / ``pit_terrain`` (terrain.py:38-187) run UNMODIFIED over this module
(oracle/ref_harness.py::build_reference_terrain) to produce tests/golden/terrain_*.npz.

Only tests/ and tests/golden/make_golden.py may import this file.

Random draws: the original calls ``np.random.choice`` / ``np.random.randint`` on numpy's global stream.  Here every draw goes
through ``draw_index(n)`` -> an integer in [0, n); by default that is numpy's stream (the published behaviour), the harness
swaps in the counter-based contract of oracle/rng.py (purpose TERRAIN, "env" = sub-terrain index, lane = draw number), which
is what the CUDA raster kernel evaluates per pixel.
"""
import numpy as np

_draw_hook = None


def draw_index(n, lane):
    """one uniform integer in [0, n).  ``lane`` names the draw STRUCTURALLY (which rectangle / band / grid node it belongs
    to) so that a raster kernel can evaluate it per pixel; numpy's global stream (the published behaviour) ignores it"""
    if _draw_hook is not None:
        return _draw_hook(int(n), int(lane))
    return int(np.random.randint(0, n))


# structural lane numbering of the draws (shared with csrc/pf_terrain.cu)
LANE_NOISE = 0            # random_uniform_terrain: + a * ny + b for grid node (a, b)
LANE_RECT = 0             # discrete_obstacles_terrain: + 5 * r + {0 width, 1 length, 2 start_i, 3 start_j, 4 height}
LANE_BAND = 0             # stepping_stones_terrain: + 128 * band + {0 start offset, 1 first-hole height, 2 + s height of stone s}
LANE_LAYOUT = 0x10000     # Terrain.randomized_terrain (terrain.py:78-79): + {0 choice, 1 difficulty}


def set_draw_hook(fn):
    global _draw_hook
    _draw_hook = fn


class SubTerrain:
    def __init__(self, terrain_name="terrain", width=256, length=256, vertical_scale=1.0, horizontal_scale=1.0):
        self.terrain_name = terrain_name
        self.vertical_scale = vertical_scale
        self.horizontal_scale = horizontal_scale
        self.width = width
        self.length = length
        self.height_field_raw = np.zeros((self.width, self.length), dtype=np.int16)


def bilinear_upsample(grid, n_out_x, n_out_y):
    """interp2d(kind='linear') of a regular grid sampled at ``np.linspace`` points spanning the same extent: plain bilinear
    interpolation in float64 (the extent cancels: only the index ratio (n_in - 1) / (n_out - 1) enters)"""
    gx, gy = grid.shape
    out = np.empty((n_out_x, n_out_y), np.float64)
    for a in range(n_out_x):
        tx = a * (gx - 1) / (n_out_x - 1)
        ix = min(int(tx), gx - 2)
        fx = tx - ix
        for b in range(n_out_y):
            ty = b * (gy - 1) / (n_out_y - 1)
            iy = min(int(ty), gy - 2)
            fy = ty - iy
            g00, g10, g01, g11 = (float(grid[ix, iy]), float(grid[ix + 1, iy]), float(grid[ix, iy + 1]), float(grid[ix + 1, iy + 1]))
            out[a, b] = (g00 * (1 - fx) + g10 * fx) * (1 - fy) + (g01 * (1 - fx) + g11 * fx) * fy
    return out


def random_uniform_terrain(terrain, min_height, max_height, step=1, downsampled_scale=None):
    if downsampled_scale is None:
        downsampled_scale = terrain.horizontal_scale
    min_height = int(min_height / terrain.vertical_scale)
    max_height = int(max_height / terrain.vertical_scale)
    step = int(step / terrain.vertical_scale)
    heights_range = np.arange(min_height, max_height + step, step)
    nx = int(terrain.width * terrain.horizontal_scale / downsampled_scale)
    ny = int(terrain.length * terrain.horizontal_scale / downsampled_scale)
    down = np.empty((nx, ny), np.int64)
    for a in range(nx):                                   # np.random.choice(heights_range, (nx, ny)): row-major draw order
        for b in range(ny):
            down[a, b] = heights_range[draw_index(len(heights_range), LANE_NOISE + a * ny + b)]
    z = np.rint(bilinear_upsample(down, terrain.width, terrain.length))
    terrain.height_field_raw += z.astype(np.int16)
    return terrain


def pyramid_sloped_terrain(terrain, slope=1, platform_size=1.):
    x = np.arange(0, terrain.width)
    y = np.arange(0, terrain.length)
    center_x = int(terrain.width / 2)
    center_y = int(terrain.length / 2)
    xx = ((center_x - np.abs(center_x - x)) / center_x).reshape(terrain.width, 1)
    yy = ((center_y - np.abs(center_y - y)) / center_y).reshape(1, terrain.length)
    max_height = int(slope * (terrain.horizontal_scale / terrain.vertical_scale) * (terrain.width / 2))
    terrain.height_field_raw += (max_height * xx * yy).astype(terrain.height_field_raw.dtype)
    platform_size = int(platform_size / terrain.horizontal_scale / 2)
    x1 = terrain.width // 2 - platform_size
    y1 = terrain.length // 2 - platform_size
    min_h = min(terrain.height_field_raw[x1, y1], 0)
    max_h = max(terrain.height_field_raw[x1, y1], 0)
    terrain.height_field_raw = np.clip(terrain.height_field_raw, min_h, max_h)
    return terrain


def pyramid_stairs_terrain(terrain, step_width, step_height, platform_size=1.):
    step_width = int(step_width / terrain.horizontal_scale)
    step_height = int(step_height / terrain.vertical_scale)
    platform_size = int(platform_size / terrain.horizontal_scale)
    height = 0
    start_x, stop_x, start_y, stop_y = 0, terrain.width, 0, terrain.length
    while (stop_x - start_x) > platform_size and (stop_y - start_y) > platform_size:
        start_x += step_width
        stop_x -= step_width
        start_y += step_width
        stop_y -= step_width
        height += step_height
        terrain.height_field_raw[start_x: stop_x, start_y: stop_y] = height
    return terrain


def discrete_obstacles_terrain(terrain, max_height, min_size, max_size, num_rects, platform_size=1.):
    max_height = int(max_height / terrain.vertical_scale)
    min_size = int(min_size / terrain.horizontal_scale)
    max_size = int(max_size / terrain.horizontal_scale)
    platform_size = int(platform_size / terrain.horizontal_scale)
    (i, j) = terrain.height_field_raw.shape
    height_range = [-max_height, -max_height // 2, max_height // 2, max_height]
    width_range = range(min_size, max_size, 4)
    length_range = range(min_size, max_size, 4)
    for r in range(num_rects):
        width = width_range[draw_index(len(width_range), LANE_RECT + 5 * r)]
        length = length_range[draw_index(len(length_range), LANE_RECT + 5 * r + 1)]
        si = range(0, i - width, 4)
        sj = range(0, j - length, 4)
        start_i = si[draw_index(len(si), LANE_RECT + 5 * r + 2)]
        start_j = sj[draw_index(len(sj), LANE_RECT + 5 * r + 3)]
        terrain.height_field_raw[start_i:start_i + width, start_j:start_j + length] = height_range[draw_index(4, LANE_RECT + 5 * r + 4)]
    x1 = (terrain.width - platform_size) // 2
    x2 = (terrain.width + platform_size) // 2
    y1 = (terrain.length - platform_size) // 2
    y2 = (terrain.length + platform_size) // 2
    terrain.height_field_raw[x1:x2, y1:y2] = 0
    return terrain


def stepping_stones_terrain(terrain, stone_size, stone_distance, max_height, platform_size=1., depth=-10):
    stone_size = int(stone_size / terrain.horizontal_scale)
    stone_distance = int(stone_distance / terrain.horizontal_scale)
    max_height = int(max_height / terrain.vertical_scale)
    platform_size = int(platform_size / terrain.horizontal_scale)
    height_range = np.arange(-max_height - 1, max_height, step=1)
    start_x, start_y = 0, 0
    terrain.height_field_raw[:, :] = int(depth / terrain.vertical_scale)
    nh, band = len(height_range), 0
    if terrain.length >= terrain.width:
        while start_y < terrain.length:
            stop_y = min(terrain.length, start_y + stone_size)
            start_x = draw_index(stone_size, LANE_BAND + 128 * band)
            stop_x = max(0, start_x - stone_distance)                                   # the piece in front of the first stone
            terrain.height_field_raw[0: stop_x, start_y: stop_y] = height_range[draw_index(nh, LANE_BAND + 128 * band + 1)]
            s = 0
            while start_x < terrain.width:
                stop_x = min(terrain.width, start_x + stone_size)
                terrain.height_field_raw[start_x: stop_x, start_y: stop_y] = height_range[draw_index(nh, LANE_BAND + 128 * band + 2 + s)]
                start_x += stone_size + stone_distance
                s += 1
            start_y += stone_size + stone_distance
            band += 1
    else:
        while start_x < terrain.width:
            stop_x = min(terrain.width, start_x + stone_size)
            start_y = draw_index(stone_size, LANE_BAND + 128 * band)
            stop_y = max(0, start_y - stone_distance)
            terrain.height_field_raw[start_x: stop_x, 0: stop_y] = height_range[draw_index(nh, LANE_BAND + 128 * band + 1)]
            s = 0
            while start_y < terrain.length:
                stop_y = min(terrain.length, start_y + stone_size)
                terrain.height_field_raw[start_x: stop_x, start_y: stop_y] = height_range[draw_index(nh, LANE_BAND + 128 * band + 2 + s)]
                start_y += stone_size + stone_distance
                s += 1
            start_x += stone_size + stone_distance
            band += 1
    x1 = (terrain.width - platform_size) // 2
    x2 = (terrain.width + platform_size) // 2
    y1 = (terrain.length - platform_size) // 2
    y2 = (terrain.length + platform_size) // 2
    terrain.height_field_raw[x1:x2, y1:y2] = 0
    return terrain


def convert_heightfield_to_trimesh(height_field_raw, horizontal_scale, vertical_scale, slope_threshold=None):
    """vertices fp32[rows*cols, 3], triangles uint32[2*(rows-1)*(cols-1), 3]; cells steeper than the threshold get their
    vertices moved so that the face becomes vertical"""
    hf = height_field_raw
    num_rows, num_cols = hf.shape
    y = np.linspace(0, (num_cols - 1) * horizontal_scale, num_cols)
    x = np.linspace(0, (num_rows - 1) * horizontal_scale, num_rows)
    yy, xx = np.meshgrid(y, x)
    if slope_threshold is not None:
        slope_threshold *= horizontal_scale / vertical_scale
        move_x = np.zeros((num_rows, num_cols))
        move_y = np.zeros((num_rows, num_cols))
        move_corners = np.zeros((num_rows, num_cols))
        move_x[:num_rows - 1, :] += (hf[1:num_rows, :] - hf[:num_rows - 1, :] > slope_threshold)
        move_x[1:num_rows, :] -= (hf[:num_rows - 1, :] - hf[1:num_rows, :] > slope_threshold)
        move_y[:, :num_cols - 1] += (hf[:, 1:num_cols] - hf[:, :num_cols - 1] > slope_threshold)
        move_y[:, 1:num_cols] -= (hf[:, :num_cols - 1] - hf[:, 1:num_cols] > slope_threshold)
        move_corners[:num_rows - 1, :num_cols - 1] += (hf[1:num_rows, 1:num_cols] - hf[:num_rows - 1, :num_cols - 1] > slope_threshold)
        move_corners[1:num_rows, 1:num_cols] -= (hf[:num_rows - 1, :num_cols - 1] - hf[1:num_rows, 1:num_cols] > slope_threshold)
        xx += (move_x + move_corners * (move_x == 0)) * horizontal_scale
        yy += (move_y + move_corners * (move_y == 0)) * horizontal_scale
    vertices = np.zeros((num_rows * num_cols, 3), dtype=np.float32)
    vertices[:, 0] = xx.flatten()
    vertices[:, 1] = yy.flatten()
    vertices[:, 2] = hf.flatten() * vertical_scale
    triangles = -np.ones((2 * (num_rows - 1) * (num_cols - 1), 3), dtype=np.uint32)
    for i in range(num_rows - 1):
        ind0 = np.arange(0, num_cols - 1) + i * num_cols
        ind1 = ind0 + 1
        ind2 = ind0 + num_cols
        ind3 = ind2 + 1
        start = 2 * i * (num_cols - 1)
        stop = start + 2 * (num_cols - 1)
        triangles[start:stop:2, 0] = ind0
        triangles[start:stop:2, 1] = ind3
        triangles[start:stop:2, 2] = ind1
        triangles[start + 1:stop:2, 0] = ind0
        triangles[start + 1:stop:2, 1] = ind2
        triangles[start + 1:stop:2, 2] = ind3
    return vertices, triangles
