"""Scene description front end: Mitsuba XML (version 2.x/3.x) and `load_dict` dictionaries -> a tree of
`Properties` nodes.

Mirrors the parts of the reference loader the config scenes use (src/core/xml.cpp:1368 `load_file`,
`<default>` / `$name` substitution :951-1010, `<ref>`, transforms :1020-1140, rgb/spectrum handling
:1146-1265; `load_dict` src/core/python/xml_v.cpp:91-116).
"""
import math
import os
import re
import xml.etree.ElementTree as ET

import numpy as np


class Properties(dict):
    """Typed key/value bag with a plugin name and id (src/core/properties.cpp)."""

    def __init__(self, plugin_name="", **kw):
        super().__init__(**kw)
        self.plugin_name = plugin_name
        self.id = None
        self.cls = None          # XML tag: scene / shape / bsdf / emitter / sensor / film / ...
        self.children = []       # unnamed nested plugin nodes, in order

    def has_property(self, key):
        return key in self

    def get(self, key, def_value=None):  # mi.Properties.get(name, def_value=...)
        return super().get(key, def_value)

    def named_children(self):
        return [(k, v) for k, v in self.items() if isinstance(v, Properties)]

    def all_children(self):
        return [(None, c) for c in self.children] + self.named_children()


# ---- transforms (include/mitsuba/core/transform.h) -----------------------------------------------------------------

def translate(v):
    m = np.eye(4)
    m[:3, 3] = v
    return m


def scale(v):
    v = np.broadcast_to(np.asarray(v, dtype=np.float64), (3,))
    return np.diag([v[0], v[1], v[2], 1.0])


def rotate(axis, angle_deg):
    a = np.asarray(axis, dtype=np.float64)
    a = a / np.linalg.norm(a)
    s, c = math.sin(math.radians(angle_deg)), math.cos(math.radians(angle_deg))
    x, y, z = a
    m = np.eye(4)
    m[:3, :3] = [[c + x * x * (1 - c), x * y * (1 - c) - z * s, x * z * (1 - c) + y * s],
                 [y * x * (1 - c) + z * s, c + y * y * (1 - c), y * z * (1 - c) - x * s],
                 [z * x * (1 - c) - y * s, z * y * (1 - c) + x * s, c + z * z * (1 - c)]]
    return m


def look_at(origin, target, up):
    origin, target, up = (np.asarray(v, dtype=np.float64) for v in (origin, target, up))
    d = target - origin
    d /= np.linalg.norm(d)
    left = np.cross(up, d)
    left /= np.linalg.norm(left)
    new_up = np.cross(d, left)
    m = np.eye(4)
    m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = left, new_up, d, origin
    return m


class Transform:
    """Chainable stand-in for mi.ScalarTransform4f used in load_dict dictionaries."""

    def __init__(self, m=None):
        self.matrix = np.eye(4) if m is None else np.asarray(m, dtype=np.float64)

    def __matmul__(self, o):
        return Transform(self.matrix @ (o.matrix if isinstance(o, Transform) else np.asarray(o)))

    def translate(self, v): return self @ Transform(translate(v))
    def scale(self, v): return self @ Transform(scale(v))
    def rotate(self, axis, angle): return self @ Transform(rotate(axis, angle))
    def look_at(self, origin, target, up): return self @ Transform(look_at(origin, target, up))

    def to_frame(self, frame):
        """Transform4f.to_frame(Frame3f): columns s, t, n (include/mitsuba/core/transform.h)"""
        m = np.eye(4)
        m[:3, 0], m[:3, 1], m[:3, 2] = frame.s, frame.t, frame.n
        return self @ Transform(m)


def _floats(s):
    return [float(x) for x in re.split(r"[\s,]+", s.strip()) if x]


def _vec3(el, default=0.0):
    if "value" in el.attrib:
        v = _floats(el.attrib["value"])
        return v * 3 if len(v) == 1 else v
    return [float(el.attrib.get(k, default)) for k in "xyz"]


def _parse_transform(el):
    m = np.eye(4)
    for c in el:
        t = c.tag
        if t == "translate":
            cur = translate(_vec3(c))
        elif t == "scale":
            cur = scale(_vec3(c, 1.0))
        elif t == "rotate":
            cur = rotate([float(c.attrib.get(k, 0)) for k in "xyz"], float(c.attrib["angle"]))
        elif t in ("lookat", "look_at"):
            cur = look_at(_floats(c.attrib["origin"]), _floats(c.attrib["target"]),
                          _floats(c.attrib.get("up", "0 1 0")))
        elif t == "matrix":
            v = _floats(c.attrib["value"])
            cur = np.array(v).reshape(4, 4) if len(v) == 16 else np.block(
                [[np.array(v).reshape(3, 3), np.zeros((3, 1))], [np.zeros((1, 3)), np.ones((1, 1))]])
        else:
            raise RuntimeError(f"unsupported transform element <{t}>")
        m = cur @ m
    return m


_PLUGIN_TAGS = {"scene", "shape", "bsdf", "emitter", "sensor", "film", "sampler", "integrator", "rfilter",
                "texture", "spectrum", "medium", "phase", "volume"}


class Color(tuple):
    """<rgb value=...>; becomes an srgb / d65 texture depending on context (xml.cpp:1146-1162)."""


class SpectrumValue:
    """<spectrum value=...|filename=...>: a constant, (wavelength:value) pairs or a .spd file."""

    def __init__(self, const=None, wavelengths=None, values=None):
        self.const, self.wavelengths, self.values = const, wavelengths, values


def read_spd(path):
    """include/mitsuba/core/spectrum.h:515-534 (spectrum_from_file)"""
    wl, val = [], []
    for line in open(path):
        line = line.split("#")[0].strip()
        if not line:
            continue
        a, b = line.split()[:2]
        wl.append(float(a)); val.append(float(b))
    return wl, val


def _parse_spectrum(el, base):
    if "filename" in el.attrib:
        wl, val = read_spd(os.path.join(base, el.attrib["filename"]))
        return SpectrumValue(wavelengths=wl, values=val)
    s = el.attrib["value"]
    if ":" in s:
        pairs = [p.split(":") for p in re.split(r"[\s,]+", s.strip()) if p]
        return SpectrumValue(wavelengths=[float(a) for a, _ in pairs], values=[float(b) for _, b in pairs])
    return SpectrumValue(const=float(s))


def load_string_tree(root, base, defaults):
    params = dict(defaults)
    ids = {}

    def sub(s):
        def rep(m):
            k = m.group(1)
            if k not in params:
                raise RuntimeError(f'Undefined parameter "${k}"')
            return str(params[k])
        return re.sub(r"\$(\w+)", rep, s)

    def attrs(el):
        for k in list(el.attrib):
            el.attrib[k] = sub(el.attrib[k])

    def parse_node(el):
        attrs(el)
        p = Properties(el.attrib.get("type", ""))
        p.cls = el.tag
        p.id = el.attrib.get("id")
        p.base_path = base
        for c in el:
            attrs(c)
            t, name = c.tag, c.attrib.get("name")
            if t == "default":
                params.setdefault(c.attrib["name"], c.attrib["value"])
            elif t in ("float",):
                p[name] = float(c.attrib["value"])
            elif t == "integer":
                p[name] = int(c.attrib["value"])
            elif t == "boolean":
                p[name] = c.attrib["value"].lower() == "true"
            elif t == "string":
                p[name] = c.attrib["value"]
            elif t in ("vector", "point"):
                p[name] = np.array(_vec3(c))
            elif t == "rgb":
                v = _floats(c.attrib["value"])
                p[name] = Color(v * 3 if len(v) == 1 else v)
            elif t == "spectrum" and "type" not in c.attrib:
                p[name] = _parse_spectrum(c, base)
            elif t == "transform":
                p[name] = _parse_transform(c)
            elif t == "ref":
                rid = c.attrib["id"]
                if rid not in ids:
                    raise RuntimeError(f'reference to unknown id "{rid}"')
                if name:
                    p[name] = ids[rid]
                else:
                    p.children.append(ids[rid])
            elif t in _PLUGIN_TAGS:
                child = parse_node(c)
                if child.id:
                    ids[child.id] = child
                if name and p.cls != "scene":
                    p[name] = child
                else:
                    p.children.append(child)
            elif t in ("include", "alias", "path"):
                raise RuntimeError(f"<{t}> is not supported")
            else:
                raise RuntimeError(f"unsupported XML element <{t}>")
        return p

    if root.tag != "scene":
        raise RuntimeError("root element must be <scene>")
    # <default> elements may appear anywhere at the top level before use
    for c in root:
        if c.tag == "default":
            params.setdefault(c.attrib["name"], sub(c.attrib["value"]))
    return parse_node(root)


def load_file_tree(path, **defaults):
    tree = ET.parse(path)
    return load_string_tree(tree.getroot(), os.path.dirname(os.path.abspath(path)), defaults)


def dict_to_tree(d, base="."):
    """mi.load_dict: {'type': ..., key: value | nested dict}"""
    from .spectra import MI_CIE_MIN  # noqa: F401  (keeps import graph simple)

    def conv(node, cls_hint=None):
        p = Properties(node["type"])
        p.cls = cls_hint or _guess_cls(node["type"])
        p.id = node.get("id")
        p.base_path = base
        for k, v in node.items():
            if k in ("type", "id"):
                continue
            if isinstance(v, dict):
                if v.get("type") == "rgb":
                    val = v["value"]
                    p[k] = Color([val] * 3 if np.isscalar(val) else list(val))
                elif v.get("type") == "spectrum":
                    val = v.get("value")
                    if "filename" in v:
                        wl, vals = read_spd(v["filename"]); p[k] = SpectrumValue(wavelengths=wl, values=vals)
                    elif np.isscalar(val):
                        p[k] = SpectrumValue(const=float(val))
                    else:
                        p[k] = SpectrumValue(wavelengths=[a for a, _ in val], values=[b for _, b in val])
                elif v.get("type") == "ref":
                    raise RuntimeError("dict refs are not supported")
                else:
                    child = conv(v)
                    if p.cls == "scene" or child.cls in ("shape", "emitter", "sensor") and p.cls == "scene":
                        p.children.append(child)
                    else:
                        p[k] = child
            elif isinstance(v, Properties):
                p[k] = v
            elif isinstance(v, Transform):
                p[k] = v.matrix
            elif isinstance(v, (list, tuple, np.ndarray)) and not isinstance(v, Color):
                a = np.asarray(v, dtype=np.float64)
                p[k] = a
            else:
                p[k] = v
        return p
    return conv(d)


_CLS = {
    "scene": "scene", "ply": "shape", "serialized": "shape", "obj": "shape", "rectangle": "shape", "sphere": "shape", "cube": "shape",
    "diffuse": "bsdf", "conductor": "bsdf", "dielectric": "bsdf", "roughconductor": "bsdf", "roughgrating": "bsdf",
    "twosided": "bsdf", "principled": "bsdf", "area": "emitter", "directional": "emitter", "constant": "emitter", "envmap": "emitter",
    "perspective": "sensor", "orthographic": "sensor", "batch": "sensor", "hdrfilm": "film",
    "independent": "sampler", "gaussian": "rfilter", "tent": "rfilter", "box": "rfilter",
    "bitmap": "texture", "srgb": "texture", "d65": "texture", "uniform": "texture", "regular": "texture",
    "irregular": "texture", "blackbody": "texture", "plt": "integrator", "stokes": "integrator", "path": "integrator",
}


def _guess_cls(t):
    if t not in _CLS:
        raise RuntimeError(f'Plugin "{t}" not found')
    return _CLS[t]
