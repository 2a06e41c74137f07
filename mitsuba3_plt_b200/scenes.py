"""The config scenes (BASELINE.json `configs`) by name.

`load(name)` resolves "cbox", "cbox-grating", "gratingcap" and "veach-mis" to the reference's OWN scene files (XML + PLY + PNG,
shipped as input fixtures under tests/golden/scenes/, or wherever $PLTB_SCENES_DIR points) -- these are what bench.py, smoke()
and the parity tests render.  The procedural writers below remain for the scenes the reference does not hold as files
("spectrograph:*" is built by scripts/spectrograph/scene.py, "disk" lacks its envmap blob), for the ":sky" / ":envmap"
variants, and as `load(name, procedural=True)` stand-ins whose oracle counters tests/test_host.py compares with the files':
same camera, materials, plugin parameters and light, geometry regenerated with the same triangle counts and placement
(32x16 UV spheres of 960 triangles, two-triangle quads), written in Mitsuba's own formats (XML 2.1, binary PLY, PNG) into a
cache directory and loaded through `load_file`.

Measured from the reference meshes (scenes/cbox*/meshes/*.ply): box = [-1,1]^3, light quad at y = 0.98624 with half-size
0.50592, spheres centred at (-0.4782,-0.64555,-0.1665) r 0.3568, (0.1512,-0.7712,0.5576) r 0.2274,
(0.4599,-0.69485,-0.1575) r 0.2988.
"""
import math
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


def write_cbox(directory, grating, sky=None):
    """C1 (`scenes/cbox`) or C2 (`scenes/cbox-grating`) equivalent.  Returns the path of the XML file.
    `sky`: "sky" adds a `constant` environment emitter (the open cbox-grating box then sees it through its missing walls)."""
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
    if sky == "sky":
        x.append('  <emitter type="constant"><rgb name="radiance" value="0.35 0.5 0.8"/></emitter>\n')
    elif sky == "envmap":   # latitude-longitude EXR (ZIP-compressed half floats): sky gradient, ground, a small bright sun
        from . import imageio
        os.makedirs(os.path.join(directory, "textures"), exist_ok=True)
        imageio.write_exr(os.path.join(directory, "textures", "sky.exr"), procedural_sky(), half=True, compression="zip")
        x.append('  <emitter type="envmap"><string name="filename" value="textures/sky.exr"/><float name="scale" value="1.5"/>'
                 '<transform name="to_world"><rotate y="1" angle="40"/><rotate x="1" angle="-15"/></transform></emitter>\n')
    elif sky:
        raise RuntimeError(f'unknown scene option "{sky}"')
    x.append("</scene>\n")
    path = os.path.join(directory, "scene.xml")
    with open(path, "w") as f:
        f.write("".join(x))
    return path


def procedural_sky(width=96, height=48):
    """(height, width, 3) float32 environment map: blue-to-white sky above the horizon, dim brown ground, a 3x3-texel sun"""
    v = (np.arange(height, dtype=np.float32) + 0.5) / height
    u = (np.arange(width, dtype=np.float32) + 0.5) / width
    up = np.clip(1.0 - 2.0 * v, 0.0, 1.0)[:, None]
    sky = np.stack([0.35 + 0.5 * (1 - up), 0.5 + 0.4 * (1 - up), 0.9 + 0.0 * up], -1) * (0.6 + 0.4 * np.cos(2 * np.pi * u)[None, :, None])
    ground = np.array([0.12, 0.09, 0.06], np.float32) * (1.0 + 0.3 * np.sin(6 * np.pi * u))[None, :, None]
    img = np.where((v < 0.5)[:, None, None], sky, np.broadcast_to(ground, sky.shape)).astype(np.float32)
    img[height // 5:height // 5 + 3, width // 3:width // 3 + 3] = [60.0, 52.0, 40.0]
    return img


def _xf(m):
    return '<transform name="to_world"><matrix value="' + " ".join(repr(float(v)) for v in m) + '"/></transform>'


def write_veach_mis(directory):
    """C4 (`scenes/veach-mis/scene.xml`) equivalent: four tilted `cube` plates with Beckmann rough conductors of increasing
    roughness, two diffuse `rectangle` walls, three spherical area lights of radius 1 / 0.5 / 0.05, tent filter."""
    os.makedirs(directory, exist_ok=True)
    x = ['<scene version="3.0.0">\n  <default name="integrator" value="plt"/>\n  <default name="spp" value="512"/>\n'
         '  <default name="resy" value="720"/>\n  <default name="resx" value="1280"/>\n  <default name="max_depth" value="7"/>\n'
         '  <integrator type="$integrator"><integer name="max_depth" value="$max_depth"/>'
         '<integer name="rr_depth" value="50"/></integrator>\n'
         '  <sensor type="perspective"><float name="fov" value="35"/>' +
         _xf([-4.37113e-08, 0, -1, 28.2792, 0, 1, 0, 3.5, 1, 0, -4.37113e-08, 1.23612e-06, 0, 0, 0, 1]) +
         '<sampler type="independent"><integer name="sample_count" value="$spp"/></sampler>'
         '<film type="hdrfilm"><integer name="width" value="$resx"/><integer name="height" value="$resy"/>'
         '<string name="pixel_format" value="rgb"/><rfilter type="tent"/></film></sensor>\n']
    x.append(_diffuse("DiffuseBSDF", "0.5, 0.5, 0.5"))
    x.append(_diffuse("NullBSDF", "0, 0, 0"))
    plates = [("Smooth", 0.01, [0.805757, 0.0961775, 0, 0.264069, -0.673242, 0.115108, 0, 4.09801]),
              ("Glossy", 0.05, [0.972057, 0.0567134, 0, 3.06163, -0.396994, 0.138865, 0, 2.71702]),
              ("Rough", 0.1, [1.03191, 0.0277252, 0, 7.09981, -0.194077, 0.147415, 0, 1.81891]),
              ("SuperRough", 0.25, [1.04217, 0.0182831, 0, 10.6769, -0.127982, 0.148882, 0, 1.23376])]
    for name, alpha, rows in plates:
        x.append(f'  <bsdf type="twosided" id="{name}BSDF"><bsdf type="roughconductor"><float name="alpha" value="{alpha}"/>'
                 '<string name="distribution" value="beckmann"/><rgb name="specular_reflectance" value="0.3, 0.3, 0.3"/>'
                 '<rgb name="eta" value="0.200438, 0.924033, 1.10221"/><rgb name="k" value="3.91295, 2.45285, 2.14219"/>'
                 '</bsdf></bsdf>\n')
    for name, alpha, rows in plates:
        x.append(f'  <shape type="cube" id="{name}">{_xf(rows + [0, 0, 4, 0, 0, 0, 0, 1])}<ref id="{name}BSDF"/></shape>\n')
    x.append('  <shape type="rectangle">' + _xf([9.9, 0, 0, 4.9, 0, -4.32743e-07, 9.9, 0, 0, -23.76, -1.03858e-06, 0, 0, 0, 0, 1]) +
             '<ref id="DiffuseBSDF"/></shape>\n')
    x.append('  <shape type="rectangle">' +
             _xf([-4.32743e-07, -4.32743e-07, 9.9, -5, -9.9, 1.89158e-14, -4.32743e-07, 9.9, 0, -23.76, -1.03858e-06, 0, 0, 0, 0, 1]) +
             '<ref id="DiffuseBSDF"/></shape>\n')
    for r, z, L in ((1, -2.8, 7.59909), (0.5, 0, 30.3964), (0.05, 2.7, 3039.64)):
        x.append(f'  <shape type="sphere"><float name="radius" value="{r}"/><point name="center" x="0" y="6.5" z="{z}"/>'
                 f'<ref id="NullBSDF"/><emitter type="area"><rgb name="radiance" value="{L}, {L}, {L}"/></emitter></shape>\n')
    x.append("</scene>\n")
    path = os.path.join(directory, "scene.xml")
    with open(path, "w") as f:
        f.write("".join(x))
    return path


def write_gratingcap(directory):
    """C5 (`scenes/gratingcap/gratingcap.xml`) equivalent: a thin vertical grating strip (`Plane.ply`: x = 0, |y| <= 10.690928,
    |z| <= 0.18630536, uv as in the file) lit by a directional light, and a half-cylinder projection screen (`Circle.ply`:
    radius 5, 15 columns at -78.75 deg ... +78.75 deg, y in [0, 4.0269127], 28 triangles, every uv = (0, 1) as in the file)
    with a `principled` BSDF.  Camera, materials and light copied parameter by parameter (gratingcap.xml:15-79); the file's
    `path` integrator is replaced by the one the scripts pass (`-i plt`)."""
    os.makedirs(os.path.join(directory, "meshes"), exist_ok=True)
    m = lambda n: os.path.join(directory, "meshes", n)
    hy, hz = 10.690928, 0.18630536
    p, t, f = quad((0, -hy, hz), (0, hy, hz), (0, hy, -hz), (0, -hy, -hz), uv=((0, 1), (1, 1), (1, 0), (0, 0)))
    write_ply(m("Plane.ply"), p, f, None, t)
    ang = np.radians(-78.75 + 11.25 * np.arange(15))
    pos, faces = [], []
    for a in ang:
        pos += [(5 * math.cos(a), 0.0, 5 * math.sin(a)), (5 * math.cos(a), 4.0269127, 5 * math.sin(a))]
    for k in range(14):
        a0, a1, b0, b1 = 2 * k, 2 * k + 1, 2 * k + 2, 2 * k + 3
        faces += [(a0, b0, b1), (a0, b1, a1)]
    write_ply(m("Circle.ply"), pos, faces, None, np.tile(np.float32([0, 1]), (30, 1)))
    x = ['<scene version="2.1.0">\n  <default name="spp" value="512"/>\n  <default name="resx" value="800"/>\n'
         '  <default name="resy" value="600"/>\n  <integrator type="plt"><integer name="max_depth" value="7"/>'
         '<integer name="rr_depth" value="50"/></integrator>\n'
         '  <sensor type="perspective" id="elm__1"><string name="fov_axis" value="x"/><float name="fov" value="39.597755"/>'
         '<float name="principal_point_offset_x" value="0.000000"/><float name="principal_point_offset_y" value="-0.000000"/>'
         '<float name="near_clip" value="0.100000"/><float name="far_clip" value="100.000000"/>'
         '<transform name="to_world"><rotate x="1" angle="-153.5592908827048"/><rotate y="1" angle="59.25153588002377"/>'
         '<rotate z="1" angle="-179.999991348578"/><translate value="-4.074883 6.534825 4.860810"/></transform>'
         '<sampler type="independent"><integer name="sample_count" value="$spp"/></sampler>'
         '<film type="hdrfilm"><integer name="width" value="$resx"/><integer name="height" value="$resy"/></film></sensor>\n',
         '  <bsdf type="twosided" id="mat-grating"><bsdf type="roughgrating"><string name="lobe_type" value="sinusoidal"/>'
         '<float name="height" value="0.04"/><float name="inv_period" value="0.9"/><boolean name="radial" value="false"/>'
         '<integer name="lobes" value="7"/><float name="grating_angle" value="0"/><float name="multiplier" value="10"/>'
         '<float name="alpha" value="0.01"/></bsdf></bsdf>\n',
         '  <bsdf type="twosided" id="mat-screen"><bsdf type="principled"><rgb value="0.800000 0.800000 0.800000" name="base_color"/>'
         '<rgb value="1.000000 1.000000 1.000000" name="spec_tint"/><float name="spec_trans" value="0.000000"/>'
         '<float name="metallic" value="0.000000"/><float name="anisotropic" value="0.000000"/>'
         '<float name="roughness" value="0.250000"/><float name="sheen" value="0.000000"/>'
         '<rgb value="1.000000 1.000000 1.000000" name="sheen_tint"/><float name="clearcoat" value="0.000000"/>'
         '<float name="clearcoat_gloss" value="0.000900"/><float name="specular" value="0.500000"/></bsdf></bsdf>\n',
         '  <emitter type="directional" id="elm__4"><rgb value="100.000000 100.000000 100.000000" name="irradiance"/>' +
         _xf([0.382683, 0, -0.923880, 5, -0.923880, 0, -0.382683, 5, 0, 1, 0, 0, 0, 0, 0, 1]) + '</emitter>\n',
         _mesh("Plane.ply", "mat-grating", True), _mesh("Circle.ply", "mat-screen", True), "</scene>\n"]
    path = os.path.join(directory, "scene.xml")
    with open(path, "w") as f:
        f.write("".join(x))
    return path


def write_disk(directory):
    """`scenes/disk/disk.xml` equivalent (config 5, driven by disk_animation.py): a leaning optical disc with a radial
    sinusoidal `roughgrating` (inv_period 0.8/0.8, 7 lobes, alpha 0.01; disk.xml:45-56), a 56 x 56 ground plane, a metallic
    `principled` box with a bitmap base colour and four dark metallic feet (disk.xml:64-97), lit only by an `envmap` with the
    file's to_world (:107-114).  The reference's EXR (`textures/empty_play_room.exr`) is missing from its snapshot, so a
    procedural sky is written instead; geometry is regenerated with comparable triangle counts (disc 384, box 12, feet 4 x 124
    in the original -> 4 x 64 here)."""
    from PIL import Image
    from . import imageio
    os.makedirs(os.path.join(directory, "meshes"), exist_ok=True)
    os.makedirs(os.path.join(directory, "textures"), exist_ok=True)
    m = lambda n: os.path.join(directory, "meshes", n)
    # disc: annulus r in [0.3, 2.4], 96 segments x 2 rings, uv centred on (0.5, 0.5) for the radial grating direction
    seg, radii, ro = 96, (0.3, 1.35, 2.4), 2.4
    ax, ay = math.radians(62.0), math.radians(-20.0)
    rx = np.array([[1, 0, 0], [0, math.cos(ax), -math.sin(ax)], [0, math.sin(ax), math.cos(ax)]])
    ry = np.array([[math.cos(ay), 0, math.sin(ay)], [0, 1, 0], [-math.sin(ay), 0, math.cos(ay)]])
    rot, centre = ry @ rx, np.array([-0.39, 1.277, 0.273])
    pos, uv, faces = [], [], []
    for r in radii:
        for k in range(seg):
            a = 2 * math.pi * k / seg
            lx, ly = r * math.cos(a), r * math.sin(a)
            pos.append(centre + rot @ np.array([lx, 0.0, ly])); uv.append((0.5 + 0.5 * lx / ro, 0.5 + 0.5 * ly / ro))
    for ring in range(2):
        for k in range(seg):
            a0, a1 = ring * seg + k, ring * seg + (k + 1) % seg
            b0, b1 = a0 + seg, a1 + seg
            faces += [(a0, b0, b1), (a0, b1, a1)]
    write_ply(m("Circle.ply"), pos, faces, None, uv)
    p, t, f = quad((-28.283, 0, 28.283), (28.283, 0, 28.283), (28.283, 0, -28.283), (-28.283, 0, -28.283))
    write_ply(m("Plane_001.ply"), p, f, None, t)

    def box(lo, hi):
        lo, hi = np.array(lo, np.float32), np.array(hi, np.float32)
        P, N, T, F = [], [], [], []
        for axis in range(3):
            for sgn in (-1, 1):
                u, v = (axis + 1) % 3, (axis + 2) % 3
                n = np.zeros(3); n[axis] = sgn
                base = len(P)
                for (cu, cv) in ((0, 0), (1, 0), (1, 1), (0, 1)) if sgn > 0 else ((0, 0), (0, 1), (1, 1), (1, 0)):
                    q = np.zeros(3); q[axis] = hi[axis] if sgn > 0 else lo[axis]
                    q[u] = hi[u] if cu else lo[u]; q[v] = hi[v] if cv else lo[v]
                    P.append(q); N.append(n); T.append((cu, cv))
                F += [(base, base + 1, base + 2), (base, base + 2, base + 3)]
        return P, N, T, F

    P, N, T, F = box((-2.216, 0.079, -5.391), (4.521, 1.333, -1.004))
    write_ply(m("Object_4.ply"), P, F, N, T)

    def foot(cx, cz, r=0.282, y0=-0.038, y1=0.094, n=16):
        P, N, T, F = [], [], [], []
        for k in range(n):            # side wall
            a = 2 * math.pi * k / n
            d = np.array([math.cos(a), 0, math.sin(a)])
            P += [np.array([cx, y0, cz]) + r * d, np.array([cx, y1, cz]) + r * d]; N += [d, d]; T += [(k / n, 0), (k / n, 1)]
        for k in range(n):
            a0, a1, b0, b1 = 2 * k, 2 * k + 1, 2 * ((k + 1) % n), 2 * ((k + 1) % n) + 1
            F += [(a0, a1, b1), (a0, b1, b0)]
        for y, ny in ((y1, 1.0), (y0, -1.0)):   # caps (fans)
            c = len(P); P.append(np.array([cx, y, cz])); N.append(np.array([0, ny, 0])); T.append((0.5, 0.5))
            for k in range(n):
                a = 2 * math.pi * k / n
                P.append(np.array([cx + r * math.cos(a), y, cz + r * math.sin(a)])); N.append(np.array([0, ny, 0]))
                T.append((0.5 + 0.5 * math.cos(a), 0.5 + 0.5 * math.sin(a)))
            for k in range(n):
                i0, i1 = c + 1 + k, c + 1 + (k + 1) % n
                F.append((c, i1, i0) if ny > 0 else (c, i0, i1))
        return P, N, T, F

    for name, (cx, cz) in (("Object_6", (-1.728, -1.555)), ("Object_8", (-1.746, -4.887)), ("Object_10", (4.059, -1.555)),
                           ("Object_12", (4.041, -4.905))):
        P, N, T, F = foot(cx, cz)
        write_ply(m(f"{name}.ply"), P, F, N, T)
    Image.fromarray(uv_test_texture(256)).save(os.path.join(directory, "textures", "Material.005_baseColor.png"))
    imageio.write_exr(os.path.join(directory, "textures", "empty_play_room.exr"), procedural_sky(128, 64), compression="zip")
    pr = ('<rgb value="1.000000 1.000000 1.000000" name="spec_tint"/><float name="spec_trans" value="0.000000"/>'
          '<float name="metallic" value="1.000000"/><float name="anisotropic" value="0.000000"/>'
          '<float name="roughness" value="0.250000"/><float name="sheen" value="0.000000"/>'
          '<rgb value="1.000000 1.000000 1.000000" name="sheen_tint"/><float name="clearcoat" value="0.000000"/>'
          '<float name="clearcoat_gloss" value="0.000000"/><float name="specular" value="0.500000"/>')
    x = ['<scene version="2.1.0">\n  <default name="spp" value="4096"/>\n  <default name="resx" value="800"/>\n'
         '  <default name="resy" value="600"/>\n  <integrator type="plt"><integer name="max_depth" value="7"/>'
         '<integer name="rr_depth" value="50"/></integrator>\n'
         '  <sensor type="perspective" id="elm__1"><string name="fov_axis" value="x"/><float name="fov" value="39.379790"/>'
         '<float name="near_clip" value="0.100000"/><float name="far_clip" value="100.000000"/>'
         '<transform name="to_world"><rotate x="1" angle="-153.5592908827048"/><rotate y="1" angle="-18.26924314288295"/>'
         '<rotate z="1" angle="-179.999991348578"/><translate value="4.924093 7.262070 11.595866"/></transform>'
         '<sampler type="independent"><integer name="sample_count" value="$spp"/></sampler>'
         '<film type="hdrfilm"><integer name="width" value="$resx"/><integer name="height" value="$resy"/></film></sensor>\n',
         '  <bsdf type="twosided" id="mat-Material.001"><bsdf type="roughgrating"><string name="lobe_type" value="sinusoidal"/>'
         '<float name="height" value="0.04"/><float name="inv_period_x" value="0.8"/><float name="inv_period_y" value="0.8"/>'
         '<boolean name="radial" value="true"/><integer name="lobes" value="7"/><float name="grating_angle" value="0"/>'
         '<float name="multiplier" value="10"/><float name="alpha" value="0.01"/></bsdf></bsdf>\n',
         '  <bsdf type="twosided" id="default-bsdf"><bsdf type="diffuse"/></bsdf>\n',
         '  <bsdf type="twosided" id="mat-Material.005"><bsdf type="principled"><texture type="bitmap" name="base_color">'
         '<string name="filename" value="textures/Material.005_baseColor.png"/></texture>' + pr + '</bsdf></bsdf>\n',
         '  <bsdf type="twosided" id="mat-Material.009"><bsdf type="principled">'
         '<rgb value="0.049707 0.049707 0.066626" name="base_color"/>' + pr + '</bsdf></bsdf>\n',
         '  <emitter type="envmap" id="elm__7"><string name="filename" value="textures/empty_play_room.exr"/>'
         '<float name="scale" value="1.000000"/>' + _xf([0, 0, 1, 0, 0, 1, 0, 0, -1, 0, 0, 0, 0, 0, 0, 1]) + '</emitter>\n',
         _mesh("Circle.ply", "mat-Material.001", True), _mesh("Plane_001.ply", "default-bsdf", True),
         _mesh("Object_4.ply", "mat-Material.005", False)]
    x += [_mesh(f"{n}.ply", "mat-Material.009", False) for n in ("Object_6", "Object_8", "Object_10", "Object_12")]
    x.append("</scene>\n")
    path = os.path.join(directory, "scene.xml")
    with open(path, "w") as f:
        f.write("".join(x))
    return path


def write_spectrograph(directory, spectrum="bbody_3000", n=150, points=150, min_wl=400.0, max_wl=700.0, use_srfs=True):
    """C3: the scene `scripts/spectrograph/scene.py:57-175` builds -- a batch of `n` orthographic 1x1 sensors, each looking
    along the second-order diffraction direction of its wavelength with a Gaussian SRF (`gen_srfs.py`), one `rectangle`
    with a ggx sinusoidal `roughgrating`, one `directional` light whose irradiance is `spectrum`:
    "bbody_<K>" (blackbody), "d65", "prior" (uniform 1)."""
    import mitsuba3_plt_b200 as mi
    os.makedirs(os.path.join(directory, "spectra"), exist_ok=True)
    # gen_srfs (scripts/spectrograph/gen_srfs.py:35-47)
    delta = (max_wl - min_wl) / (n + 2)
    domain = np.linspace(min_wl, max_wl, points)
    measured = np.linspace(min_wl, max_wl, n + 2)[1:-1]
    # generate_scene_elements (scripts/spectrograph/scene.py:57-144), through the same public types the script uses
    grating = mi.scalar_rgb.DiffractionGrating3f(0.0, mi.ScalarVector2f(0.7, 0.7), 0.05, 5, mi.DiffractionGratingType.Sinusoidal, 1,
                                                 mi.ScalarVector2f(0.0))
    grating_frame = mi.scalar_rgb.Frame3f(mi.ScalarVector3f(0, -1, 0), mi.ScalarVector3f(0, 0, 1), mi.ScalarVector3f(1, 0, 0))
    inc = math.radians(-15) - math.pi / 2
    light = (math.sin(inc), 0.0, math.cos(inc))   # scalar_sph_to_dir(incident, 0), scripts/utils/math.py:16-21
    light_dir_local = grating_frame.to_local(light)
    x = ['<scene version="2.1.0">\n  <default name="spp" value="512"/>\n'
         '  <integrator type="plt"><integer name="max_depth" value="7"/><integer name="rr_depth" value="50"/></integrator>\n'
         '  <sensor type="batch">\n']
    for i, wl in enumerate(measured):
        std = delta / 4
        srf = np.exp(-np.square(domain - wl) / (2 * std ** 2))
        srf /= np.trapezoid(srf, domain)
        mi.spectrum_to_file(os.path.join(directory, "spectra", f"srf_{i}.spd"), domain, srf)
        local_diff_dir, _ = grating.diffract(-light_dir_local, mi.ScalarVector2i(2, 0), float(wl * 1e-3))
        diff_dir = grating_frame.to_world(local_diff_dir)
        m = mi.scalar_rgb.Transform4f().look_at(origin=np.ravel(diff_dir * 0.4), target=[0, 0, 0], up=[0, 0, 1]).matrix
        srf_xml = f'<spectrum name="srf" filename="spectra/srf_{i}.spd"/>' if use_srfs else ""
        x.append(f'    <sensor type="orthographic" name="sensor{i}">{_xf(m.reshape(-1))}<film type="hdrfilm">'
                 '<integer name="width" value="1"/><integer name="height" value="1"/>'
                 f'<boolean name="sample_border" value="false"/></film>{srf_xml}</sensor>\n')
    x.append(f'    <film type="hdrfilm"><integer name="width" value="{n}"/><integer name="height" value="1"/>'
             '<boolean name="sample_border" value="false"/></film>\n'
             '    <sampler type="independent"><integer name="sample_count" value="$spp"/></sampler>\n  </sensor>\n')
    if spectrum.startswith("bbody_"):
        irr = f'<spectrum type="blackbody" name="irradiance"><float name="temperature" value="{float(spectrum[6:])}"/></spectrum>'
    elif spectrum == "d65":
        irr = '<spectrum type="d65" name="irradiance"/>'
    elif spectrum == "prior":
        irr = '<spectrum type="uniform" name="irradiance"><float name="value" value="1.0"/></spectrum>'
    else:
        raise RuntimeError(f'spectrograph: unknown spectrum "{spectrum}"')
    x.append(f'  <emitter type="directional"><vector name="direction" x="{light[0]!r}" y="{light[1]!r}" z="{light[2]!r}"/>'
             f'{irr}</emitter>\n')
    x.append('  <shape type="rectangle"><bsdf type="twosided"><bsdf type="roughgrating">'
             '<string name="distribution" value="ggx"/><string name="lobe_type" value="sinusoidal"/>'
             '<float name="height" value="0.05"/><float name="inv_period_x" value="0.7"/><float name="inv_period_y" value="0.7"/>'
             '<boolean name="radial" value="false"/><integer name="lobes" value="5"/><float name="multiplier" value="10.0"/>'
             '<float name="grating_angle" value="0.0"/><float name="alpha" value="0.03"/></bsdf></bsdf>' +
             _xf([0, 0, 0.1, 0, -0.1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 0, 1]) + '</shape>\n</scene>\n')
    path = os.path.join(directory, "scene.xml")
    with open(path, "w") as f:
        f.write("".join(x))
    return path


_BUILDERS = {"cbox": lambda d, a=None: write_cbox(d, False, a), "cbox-grating": lambda d, a=None: write_cbox(d, True, a),
             "veach-mis": write_veach_mis, "spectrograph": write_spectrograph,
             "gratingcap": write_gratingcap, "disk": write_disk}


def scene_path(name, cache_dir=None):
    """Materialise builtin scene `name` (once per process/cache dir) and return the path of its XML file."""
    base, _, arg = name.partition(":")   # "spectrograph:d65", "spectrograph:bbody_3000", "spectrograph:prior"
    if base not in _BUILDERS:
        raise RuntimeError(f'unknown builtin scene "{name}" (have: {", ".join(sorted(_BUILDERS))})')
    cache_dir = cache_dir or os.path.join(tempfile.gettempdir(), f"pltb200_scenes_{os.getuid()}")
    d = os.path.join(cache_dir, name.replace(":", "_"))
    xml = os.path.join(d, "scene.xml")
    if not os.path.exists(xml):
        tmp = tempfile.mkdtemp(dir=os.path.dirname(d) if os.path.isdir(os.path.dirname(d)) else None)
        _BUILDERS[base](tmp, arg) if arg else _BUILDERS[base](tmp)
        os.makedirs(cache_dir, exist_ok=True)
        try:
            os.rename(tmp, d)
        except OSError:  # another rank won the race
            pass
    return xml


# the reference's own files for the named configs (relative to the scenes directory) and the -D defaults its scripts pass
_REFERENCE_FILES = {"cbox": ("cbox/cbox.xml", {}), "cbox-grating": ("cbox-grating/cbox.xml", {}),
                    "gratingcap": ("gratingcap/gratingcap.xml", {}),
                    "gratings": ("gratings/gratings.xml", {}),   # directionalspot light on a non-radial sinusoidal grating (results/gratings)
                    "veach-mis": ("veach-mis/scene.xml", {"integrator": "plt", "max_depth": 7})}


def reference_scene_path(name):
    """path of the reference's scene file for config scene `name`, or None when it is not shipped"""
    if name not in _REFERENCE_FILES:
        return None
    root = os.environ.get("PLTB_SCENES_DIR") or os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                              "tests", "golden", "scenes")
    path = os.path.join(root, _REFERENCE_FILES[name][0])
    return path if os.path.exists(path) else None


def load(name, procedural=False, **defaults):
    from . import load_file
    if name.endswith(":procedural"):
        name, procedural = name[:-len(":procedural")], True
    ref = None if procedural else reference_scene_path(name)
    if ref:
        return load_file(ref, **{**_REFERENCE_FILES[name][1], **defaults})
    return load_file(scene_path(name), **defaults)
