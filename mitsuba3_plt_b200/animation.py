"""Frame loop of the reference's camera animation (scripts/rendering/disk_animation.py:108-134; BASELINE.json config 5:
`scenes/gratingcap` / `scenes/disk`, 1 spp per frame x 240 frames): many short renders of one resident scene whose sensor
`to_world` changes per frame.

Per frame i of n:  t = i / n,  easing = sin(pi t - pi/2) / 2 + 1/2,
    to_world_i = Transform().rotate([0, 1, 0], max_rotation * easing).translate(easing * move_vector) @ initial
(disk_animation.py:112-123; frames = 90 in the script, 240 in the config), rendered with `seed = i`.

Multi-GPU: with spp < world size there is nothing to shard inside a frame, so *frames* are dealt round-robin to
the ranks (`frames_of_rank`); every rank renders whole frames with `reduce=False` and there is no collective on the data path.
"""
import math

import numpy as np

from . import capi
from .xmlscene import Transform

MOVE_VECTOR = (-2.7, 0.0, 2.1)   # disk_animation.py:114
MAX_ROTATION = -90.0             # disk_animation.py:115


def easing(i, frames):
    t = i / frames
    return float(np.float32(math.sin(math.pi * t - math.pi / 2.0) / 2.0 + 0.5))   # mi.Float(...) == float32


def camera_to_world(initial, i, frames, move_vector=MOVE_VECTOR, max_rotation=MAX_ROTATION):
    """`to_world` of frame i (4x4 Transform) from the scene's initial sensor transform."""
    e = easing(i, frames)
    init = initial if isinstance(initial, Transform) else Transform(np.asarray(initial, np.float64).reshape(4, 4))
    return Transform().rotate([0, 1, 0], max_rotation * e).translate([e * v for v in move_vector]) @ init


def frames_of_rank(frames, rank, world):
    """Round-robin deal of frame indices: contiguous coverage, disjoint across ranks."""
    return list(range(rank, frames, world))


def render_frames(scene, integrator, frames=240, spp=1, sensor_key=None, *, ctx=None, on_frame=None,
                  move_vector=MOVE_VECTOR, max_rotation=MAX_ROTATION, indices=None):
    """Render the frames of this rank (all of them on one GPU).  Returns {frame index: (H, W, 3) float32 image}; with
    `on_frame(i, image)` given, images are handed over instead of being kept."""
    from . import context, traverse
    ctx = ctx or context()
    params = traverse(scene)
    key = sensor_key or next(k for k in params if k.endswith(".to_world"))
    initial = params[key]
    out = {}
    W, H = scene.film_size()
    staging = capi.PinnedArray((H, W, 3))   # page-locked frame buffer: the per-frame image download runs at PCIe speed
    for i in (indices if indices is not None else frames_of_rank(frames, ctx.rank, ctx.world)):
        params[key] = camera_to_world(initial, i, frames, move_vector, max_rotation)
        params.update()
        img = integrator.render(scene, spp=spp, seed=i, ctx=ctx, reduce=False, out=staging.array)
        if on_frame is not None:
            on_frame(i, img)        # valid until the next frame is rendered
        else:
            out[i] = img.copy()
    params[key] = initial
    params.update()
    return out
