"""Host-side construction of `Hierarchical2D<Float, 0>` (include/mitsuba/core/distr_2d.h:344-470): the MIP hierarchy over the
bilinear patches of a 2-D array that `EnvironmentMapEmitter` importance-samples (src/emitters/envmap.cpp:235, 312).  Only the
*build* lives here (load time, numpy, float32 arithmetic in the reference's order); sampling and evaluation are device code
(csrc/pltb_envmap.cuh) and, for the tests, oracle code.

Flattened layout handed to the device (all float32 words, integers stored as bit patterns):
    [n_levels, res_x, res_y, 0,  then per level: offset (relative to the block start), width]  followed by the level arrays.
Level 0 is the normalised input in row-major order; levels >= 1 store 2x2 patches contiguously (`Level::index`, :772-776).
"""
import numpy as np

F32 = np.float32


def log2i_ceil(v):
    return int(v - 1).bit_length() if v > 1 else 0


def level_index(x, y, width):
    """Level::index (distr_2d.h:772-776)"""
    return ((x & 1) | (((x & ~1) | (y & 1)) << 1)) + (y & ~1) * width


def build_hierarchical(data, normalize=True):
    """data: (res_y, res_x) float32.  Returns [(width, flat float32 array)] for levels 0 .. max_level + 1."""
    data = np.ascontiguousarray(data, dtype=F32)
    H, W = data.shape
    npx, npy = W - 1, H - 1
    if npx < 1 or npy < 1:
        raise RuntimeError("Hierarchical2D: the array must be at least 2x2")
    max_level = log2i_ceil(max(npx, npy))
    sizes = [(W, H)]
    lx, ly = npx, npy
    for _ in range(max_level, -1, -1):
        lx += lx & 1; ly += ly & 1
        sizes.append((lx, ly))
        lx >>= 1; ly >>= 1
    levels = [np.zeros(w * h, F32) for w, h in sizes]
    # level 1: integrate the linear interpolant per patch (:436-446)
    avg = F32(0.25) * (((data[:-1, :-1] + data[:-1, 1:]) + data[1:, :-1]) + data[1:, 1:])
    total = float(avg.astype(np.float64).sum())
    ys, xs = np.mgrid[0:npy, 0:npx]
    levels[1][level_index(xs, ys, sizes[1][0])] = avg
    scale = F32(float(npx * npy) / total) if (normalize and total > 0) else F32(1.0)
    levels[0][:] = (data * scale).reshape(-1)
    levels[1] *= scale
    # MIP hierarchy (:458-475)
    cx, cy = npx, npy
    for level in range(2, max_level + 2):
        cx, cy = (cx + 1) >> 1, (cy + 1) >> 1
        ys, xs = np.mgrid[0:cy, 0:cx]
        src = level_index(2 * xs, 2 * ys, sizes[level - 1][0])
        l0 = levels[level - 1]
        levels[level][level_index(xs, ys, sizes[level][0])] = ((l0[src] + l0[src + 1]) + l0[src + 2]) + l0[src + 3]
    return [(sizes[i][0], levels[i]) for i in range(len(levels))]


def flatten_hierarchical(levels, res):
    """-> float32 block described in the module docstring"""
    n = len(levels)
    head = np.zeros(4 + 2 * n, np.uint32)
    head[0], head[1], head[2] = n, res[0], res[1]
    off = len(head)
    for i, (w, a) in enumerate(levels):
        head[4 + 2 * i], head[5 + 2 * i] = off, w
        off += a.size
    return np.concatenate([head.view(F32)] + [a for _, a in levels])
