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
                  move_vector=MOVE_VECTOR, max_rotation=MAX_ROTATION, indices=None, inflight=1):
    """Render the frames of this rank (all of them on one GPU).  Returns {frame index: (H, W, 3) float32 image}; with
    `on_frame(i, image)` given, images are handed over instead of being kept.

    `inflight` > 1 keeps that many frames in flight on the GPU: a 1-spp frame is a chain of ~35 dependent kernels, each bounded
    below by the latency of one path's work, so a single frame leaves most SMs idle most of the time.  Every lane has its own
    context (stream, wavefront state, film) and device copy of the scene and renders whole frames with the seeds and sensors
    of the serial loop, so each frame's image is the one the serial loop produces; only the order in which `on_frame` is
    called changes (calls are serialised, not ordered)."""
    from . import context, traverse
    ctx = ctx or context()
    params = traverse(scene)
    key = sensor_key or next(k for k in params if k.endswith(".to_world"))
    initial = params[key]
    out = {}
    W, H = scene.film_size()
    todo = list(indices if indices is not None else frames_of_rank(frames, ctx.rank, ctx.world))
    if inflight > 1 and len(todo) > 1:
        _render_frames_inflight(scene, integrator, todo, frames, spp, params._index[key], initial, ctx, on_frame, out,
                                move_vector, max_rotation, min(inflight, len(todo)))
        return out
    staging = capi.PinnedArray((H, W, 3))   # page-locked frame buffer: the per-frame image download runs at PCIe speed
    for i in todo:
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


def _render_frames_inflight(scene, integrator, todo, frames, spp, sensor_index, initial, ctx, on_frame, out, move_vector,
                            max_rotation, n_lanes):
    import copy
    import threading
    from . import Context, Scene
    W, H = scene.film_size()
    lock, errors = threading.Lock(), []
    lanes = scene.__dict__.setdefault("_anim_lanes", {})

    def lane(k):
        try:
            slot = lanes.get((ctx.device, k))   # reused across calls: a context owns its wavefront state and film
            if slot is None:
                slot = lanes[(ctx.device, k)] = (Context(ctx.device), Scene(scene.flat, scene.tree))
            lctx, lscene = slot
            integ = copy.copy(integrator)
            staging = capi.PinnedArray((H, W, 3))
            dev = lscene.device(lctx)
            for i in todo[k::n_lanes]:
                m = camera_to_world(initial, i, frames, move_vector, max_rotation).matrix
                m = np.ascontiguousarray(np.asarray(m, np.float64).reshape(16).astype(np.float32))
                capi.check(capi.lib().pltb_scene_update_sensor(dev, sensor_index, capi.fptr(m)))
                img = integ.render(lscene, spp=spp, seed=i, ctx=lctx, reduce=False, out=staging.array)
                with lock:
                    integrator._last_stats = integ._last_stats
                    if on_frame is not None:
                        on_frame(i, img)
                    else:
                        out[i] = img.copy()
        except Exception as e:   # noqa: BLE001 -- re-raised on the calling thread
            errors.append(e)

    threads = [threading.Thread(target=lane, args=(k,)) for k in range(n_lanes)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    if errors:
        raise errors[0]
