"""Procedural stand-ins for the reference's config scenes (BASELINE.json `configs`).

The reference's scene files (XML + PLY + PNG under scenes/) are not redistributed here and do not exist on the GPU box,
so tests/, smoke() and bench.py build *equivalent* scenes from scratch: same camera, same materials and plugin
parameters, same light, geometry regenerated procedurally with the same triangle counts and placement (32x16 UV spheres of
960 triangles, two-triangle quads).  Files are written in Mitsuba's own formats (XML 2.1, binary PLY, PNG) into a cache
directory and loaded through `load_file`, so the XML / PLY / bitmap loaders are exercised exactly as with the originals.
A user who has the reference tree simply calls `mi.load_file("<ref>/scenes/cbox-grating/cbox.xml")` instead.

Measured from the reference meshes (scenes/cbox*/meshes/*.ply): box = [-1,1]^3, light quad at y = 0.98624 with half-size
0.50592, spheres centred at (-0.4782,-0.64555,-0.1665) r 0.3568, (0.1512,-0.7712,0.5576) r 0.2274,
(0.4599,-0.69485,-0.1575) r 0.2988.
"""
import os
import struct
import tempfile

import numpy as np

_SPHERES = [((-0.4782, -0.64555, -0.1665), 0.3568), ((0.1512, -0.7712, 0.5576), 0.2274), ((0.4599, -0.69485, -0.1575), 0.2988)]
_LIGHT_Y, _LIGHT_H = 0.98624, 0.50592


def write_ply(path, pos, faces, nrm=None, uv=None):
    """binary_little_endian PLY with the property names the reference's meshes use (x y z [nx ny nz] [s t])"""
    pos = np.asarray(pos, np.float32)
    cols, names = [pos], ["x", "y", "z"]
    if nrm is not None:
        cols.append(np.asarray(nrm, np.float32)); names += ["nx", "ny", "nz"]
    if uv is not None:
        cols.append(np.asarray(uv, np.float32)); names += ["s", "t"]
    v = np.ascontiguousarray(np.concatenate(cols, 1), dtype="<f4")
    faces = np.asarray(faces, "<i4")
    with open(path, "wb") as f:
        hdr = ["ply", "format binary_little_endian 1.0", f"element vertex {len(v)}"]
        hdr += [f"property float {n}" for n in names]
        hdr += [f"element face {len(faces)}", "property list uchar int vertex_indices", "end_header"]
        f.write(("\n".join(hdr) + "\n").encode())
        f.write(v.tobytes())
        rec = np.zeros(len(faces), dtype=[("n", "u1"), ("v", "<i4", (3,))])
        rec["n"], rec["v"] = 3, faces
        f.write(rec.tobytes())


def uv_sphere(center, radius, segments=32, rings=16):
    """UV sphere with a duplicated seam column: (segments+1)*(rings+1) vertices, segments*(2*rings-2) triangles
    (960 for 32x16, as the reference's Sphere*.ply); smooth normals, (s,t) = (phi/2pi, 1 - theta/pi)."""
    c = np.asarray(center, np.float64)
    pos, nrm, uv = [], [], []
    for r in range(rings + 1):
        th = np.pi * r / rings
        for s in range(segments + 1):
            ph = 2 * np.pi * s / segments
            n = np.array([np.sin(th) * np.cos(ph), np.cos(th), np.sin(th) * np.sin(ph)])
            pos.append(c + radius * n); nrm.append(n); uv.append((s / segments, 1 - r / rings))
    faces = []
    idx = lambda r, s: r * (segments + 1) + s
    for r in range(rings):
        for s in range(segments):
            a, b, cc, d = idx(r, s), idx(r, s + 1), idx(r + 1, s + 1), idx(r + 1, s)
            if r > 0:
                faces.append((a, b, cc))
            if r < rings - 1:
                faces.append((a, cc, d))
    return np.array(pos, np.float32), np.array(nrm, np.float32), np.array(uv, np.float32), np.array(faces, np.int32)


def quad(p0, p1, p2, p3, uv=((0, 0), (1, 0), (1, 1), (0, 1))):
    return np.array([p0, p1, p2, p3], np.float32), np.array(uv, np.float32), np.array([[0, 1, 2], [0, 2, 3]], np.int32)


def uv_test_texture(size=512):
    """A colour-gradient checker in the spirit of a 'UV test' chart (8x8 cells, hue by column, value by row)."""
    y, x = np.mgrid[0:size, 0:size]
    cx, cy = (x * 8) // size, (y * 8) // size
    hue = cx / 8.0
    val = 0.35 + 0.6 * (cy / 7.0)
    k = lambda n: (n + hue * 6) % 6
    ch = lambda n: val - val * 0.8 * np.clip(np.minimum(k(n), 4 - k(n)), 0, 1)
    rgb = np.stack([ch(5), ch(3), ch(1)], -1)
    rgb = np.where(((cx + cy) % 2 == 0)[..., None], rgb, rgb * 0.55 + 0.1)
    return (np.clip(rgb, 0, 1) * 255 + 0.5).astype(np.uint8)


_CAMERA = """
  <sensor type="perspective" id="camera">
    <string name="fov_axis" value="x"/>
    <float name="fov" value="39.597755"/>
    <float name="near_clip" value="0.1"/>
    <float name="far_clip" value="100"/>
    <transform name="to_world">
      <rotate x="1" angle="179.999991348578"/>
      <rotate y="1" angle="5.008956130975331e-06"/>
      <rotate z="1" angle="180.00000500895632"/>
      <translate value="0 0 3.700172"/>
    </transform>
    <sampler type="independent"><integer name="sample_count" value="$spp"/></sampler>
    <film type="hdrfilm"><integer name="width" value="$resx"/><integer name="height" value="$resy"/></film>
  </sensor>
"""


def _diffuse(id_, rgb):
    return (f'  <bsdf type="twosided" id="{id_}"><bsdf type="diffuse"><rgb name="reflectance" value="{rgb}"/></bsdf></bsdf>\n')


def _mesh(file, bsdf, face_normals, extra=""):
    fn = '<boolean name="face_normals" value="true"/>' if face_normals else ""
    return f'  <shape type="ply"><string name="filename" value="meshes/{file}"/>{fn}<ref id="{bsdf}" name="bsdf"/>{extra}</shape>\n'


def write_cbox(directory, grating):
    """C1 (`scenes/cbox`) or C2 (`scenes/cbox-grating`) equivalent.  Returns the path of the XML file."""
    os.makedirs(os.path.join(directory, "meshes"), exist_ok=True)
    os.makedirs(os.path.join(directory, "textures"), exist_ok=True)
    m = lambda n: os.path.join(directory, "meshes", n)
    for i, (c, r) in enumerate(_SPHERES):
        p, n, t, f = uv_sphere(c, r)
        write_ply(m(f"sphere_{i}.ply"), p, f, n, t)
    walls = {
        "ceiling": quad((1, 1, -1), (1, 1, 1), (-1, 1, 1), (-1, 1, -1)),
        "left": quad((-1, -1, 1), (-1, -1, -1), (-1, 1, -1), (-1, 1, 1)),
        "back": quad((-1, -1, -1), (1, -1, -1), (1, 1, -1), (-1, 1, -1)),
        "right": quad((1, -1, -1), (1, -1, 1), (1, 1, 1), (1, 1, -1)),
        "floor": quad((-1, -1, -1), (-1, -1, 1), (1, -1, 1), (1, -1, -1)),
        "light": quad((-_LIGHT_H, _LIGHT_Y, -_LIGHT_H), (_LIGHT_H, _LIGHT_Y, -_LIGHT_H), (_LIGHT_H, _LIGHT_Y, _LIGHT_H),
                      (-_LIGHT_H, _LIGHT_Y, _LIGHT_H), uv=((0, 1), (1, 1), (1, 0), (0, 0))),
    }
    for name, (p, t, f) in walls.items():
        write_ply(m(f"{name}.ply"), p, f, None, t)
    x = ['<scene version="2.1.0">\n  <default name="spp" value="64"/>\n  <default name="resx" value="500"/>\n'
         '  <default name="resy" value="500"/>\n  <integrator type="plt"><integer name="max_depth" value="7"/>'
         '<integer name="rr_depth" value="50"/></integrator>\n', _CAMERA]
    x.append('  <bsdf type="dielectric" id="glass"><float name="int_ior" value="1.4"/>'
             '<rgb name="specular_transmittance" value="1 1 1"/></bsdf>\n')
    x.append('  <bsdf type="diffuse" id="black"><rgb name="reflectance" value="0"/></bsdf>\n')
    x.append(_diffuse("white", "0.8 0.8 0.8"))
    light = '<emitter type="area"><rgb name="radiance" value="10 5 5"/></emitter>'
    if grating:
        from PIL import Image
        Image.fromarray(uv_test_texture()).save(os.path.join(directory, "textures", "uvtest.png"))
        x.append('  <bsdf type="twosided" id="grating"><bsdf type="roughgrating">'
                 '<string name="lobe_type" value="sinusoidal"/><float name="height" value="0.04"/>'
                 '<float name="inv_period_x" value="0.7"/><float name="inv_period_y" value="0.4"/>'
                 '<boolean name="radial" value="true"/><integer name="lobes" value="7"/>'
                 '<float name="grating_angle" value="1.509"/><float name="multiplier" value="10"/>'
                 '<float name="alpha" value="0.02"/></bsdf></bsdf>\n')
        x.append('  <bsdf type="twosided" id="uvtest"><bsdf type="diffuse"><texture type="bitmap" name="reflectance">'
                 '<string name="filename" value="textures/uvtest.png"/><string name="wrap_mode" value="mirror"/>'
                 '</texture></bsdf></bsdf>\n')
        x.append(_mesh("light.ply", "black", True, light))
        x.append(_mesh("sphere_0.ply", "grating", False))
        x.append(_mesh("sphere_1.ply", "uvtest", False))
        x.append(_mesh("sphere_2.ply", "glass", False))
        x.append(_mesh("ceiling.ply", "white", True))
        x.append(_mesh("floor.ply", "grating", True))
    else:
        x.append('  <bsdf type="twosided" id="mirror"><bsdf type="conductor">'
                 '<rgb name="specular_reflectance" value="0.8 0.8 0.8"/></bsdf></bsdf>\n')
        x.append('  <bsdf type="twosided" id="default"><bsdf type="diffuse"/></bsdf>\n')
        x.append(_diffuse("red", "0.800474 0.049934 0.051623"))
        x.append(_diffuse("green", "0.077992 0.587736 0.073049"))
        x.append(_mesh("light.ply", "black", True, light))
        x.append(_mesh("sphere_0.ply", "mirror", False))
        x.append(_mesh("sphere_1.ply", "default", False))
        x.append(_mesh("sphere_2.ply", "glass", False))
        for name, mat in (("ceiling", "white"), ("left", "red"), ("back", "white"), ("right", "green"), ("floor", "white")):
            x.append(_mesh(f"{name}.ply", mat, True))
    x.append("</scene>\n")
    path = os.path.join(directory, "scene.xml")
    with open(path, "w") as f:
        f.write("".join(x))
    return path


_BUILDERS = {"cbox": lambda d: write_cbox(d, False), "cbox-grating": lambda d: write_cbox(d, True)}


def scene_path(name, cache_dir=None):
    """Materialise builtin scene `name` (once per process/cache dir) and return the path of its XML file."""
    if name not in _BUILDERS:
        raise RuntimeError(f'unknown builtin scene "{name}" (have: {", ".join(sorted(_BUILDERS))})')
    cache_dir = cache_dir or os.path.join(tempfile.gettempdir(), f"pltb200_scenes_{os.getuid()}")
    d = os.path.join(cache_dir, name)
    xml = os.path.join(d, "scene.xml")
    if not os.path.exists(xml):
        tmp = tempfile.mkdtemp(dir=os.path.dirname(d) if os.path.isdir(os.path.dirname(d)) else None)
        _BUILDERS[name](tmp)
        os.makedirs(cache_dir, exist_ok=True)
        try:
            os.rename(tmp, d)
        except OSError:  # another rank won the race
            pass
    return xml


def load(name, **defaults):
    from . import load_file
    return load_file(scene_path(name), **defaults)
