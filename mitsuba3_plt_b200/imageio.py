"""Image files either side of the render path: what the reference's scripts call through
`mi.util.write_bitmap(path, image)` / `mi.Bitmap(path)` (`scripts/rendering/main-headless.py:60-73`, `src/core/bitmap.cpp`).

OpenEXR is written and read here without the OpenEXR library (it is not in the image; the reference's `ext/openexr` submodule is
empty): single-part scanline files, FLOAT or HALF channels, compression NONE on write, NONE / ZIPS / ZIP on read -- enough for
`result.exr` outputs and for an `envmap` texture saved by common tools with ZIP compression.  PNG goes through Pillow.
"""
import struct
import zlib

import numpy as np

_MAGIC = 20000630
_PIXEL_TYPES = {1: np.dtype("<f2"), 2: np.dtype("<f4"), 0: np.dtype("<u4")}


def _attr(name, typ, payload):
    return name.encode() + b"\0" + typ.encode() + b"\0" + struct.pack("<i", len(payload)) + payload


def write_exr(path, image, channel_names=None, half=False, compression="none"):
    """image: (H, W) or (H, W, C) float array.  Default channel names: Y | R,G,B | R,G,B,A (others: ch0, ch1, ...)."""
    a = np.asarray(image)
    if a.ndim == 2:
        a = a[..., None]
    if a.ndim != 3:
        raise RuntimeError("write_exr(): expected an (H, W[, C]) array")
    H, W, Cn = a.shape
    names = list(channel_names) if channel_names else {1: ["Y"], 3: ["R", "G", "B"], 4: ["R", "G", "B", "A"]}.get(
        Cn, [f"ch{i}" for i in range(Cn)])
    if len(names) != Cn:
        raise RuntimeError("write_exr(): one name per channel")
    dt = np.dtype("<f2") if half else np.dtype("<f4")
    order = sorted(range(Cn), key=lambda i: names[i])          # channels are stored in alphabetical order
    chlist = b"".join(names[i].encode() + b"\0" + struct.pack("<iBBBBii", 1 if half else 2, 0, 0, 0, 0, 1, 1) for i in order) + b"\0"
    box = struct.pack("<iiii", 0, 0, W - 1, H - 1)
    hdr = struct.pack("<ii", _MAGIC, 2)
    hdr += _attr("channels", "chlist", chlist)
    hdr += _attr("compression", "compression", bytes([{"none": 0, "zips": 2, "zip": 3}[compression]]))
    hdr += _attr("dataWindow", "box2i", box) + _attr("displayWindow", "box2i", box)
    hdr += _attr("lineOrder", "lineOrder", b"\0") + _attr("pixelAspectRatio", "float", struct.pack("<f", 1.0))
    hdr += _attr("screenWindowCenter", "v2f", struct.pack("<ff", 0.0, 0.0))
    hdr += _attr("screenWindowWidth", "float", struct.pack("<f", 1.0)) + b"\0"
    planar = np.ascontiguousarray(a[:, :, order].transpose(0, 2, 1).astype(dt))   # (H, C, W): per row, channel after channel
    lines = {"none": 1, "zips": 1, "zip": 16}[compression]
    blocks = []
    for y in range(0, H, lines):
        raw = planar[y:y + lines].tobytes()
        blocks.append((y, raw if compression == "none" else _zip(raw)))
    offsets, pos = [], len(hdr) + 8 * len(blocks)
    for _, d in blocks:
        offsets.append(pos); pos += 8 + len(d)
    with open(path, "wb") as f:
        f.write(hdr)
        f.write(np.asarray(offsets, "<u8").tobytes())
        for y, d in blocks:
            f.write(struct.pack("<ii", y, len(d)))
            f.write(d)


def _read_header(buf):
    magic, version = struct.unpack_from("<ii", buf, 0)
    if magic != _MAGIC:
        raise RuntimeError("read_exr(): not an OpenEXR file")
    if version & 0x200 or version & 0x1000 or version & 0x800:
        raise RuntimeError("read_exr(): tiled / multi-part / deep files are not supported")
    pos, attrs = 8, {}
    while buf[pos] != 0:
        e = buf.index(b"\0", pos); name = buf[pos:e].decode(); pos = e + 1
        e = buf.index(b"\0", pos); typ = buf[pos:e].decode(); pos = e + 1
        size, = struct.unpack_from("<i", buf, pos); pos += 4
        attrs[name] = (typ, buf[pos:pos + size]); pos += size
    return attrs, pos + 1


def _unzip(block, expected):
    """ZIP / ZIPS block -> raw bytes (zlib, byte-delta predictor, even/odd byte de-interleave); blocks that did not shrink are
    stored raw."""
    if len(block) >= expected:
        return bytes(block)
    d = np.frombuffer(zlib.decompress(block), np.uint8).astype(np.int64)
    d[1:] -= 128
    t = (np.cumsum(d) & 0xFF).astype(np.uint8)
    half = (len(t) + 1) // 2
    out = np.empty(len(t), np.uint8)
    out[0::2], out[1::2] = t[:half], t[half:]
    return out.tobytes()


def _zip(data):
    """inverse of _unzip (used for compression="zip" on write)"""
    r = np.frombuffer(data, np.uint8)
    t = np.concatenate([r[0::2], r[1::2]]).astype(np.int64)
    d = t.copy()
    d[1:] = (t[1:] - t[:-1] + 128 + 256) & 0xFF
    z = zlib.compress(d.astype(np.uint8).tobytes(), 6)
    return z if len(z) < len(data) else bytes(data)


def read_exr(path):
    """Returns (array (H, W, C) float32, channel names); channels R, G, B[, A] come first in that order when present."""
    buf = open(path, "rb").read()
    attrs, pos = _read_header(buf)
    comp = attrs["compression"][1][0]
    if comp not in (0, 2, 3):
        raise RuntimeError(f"read_exr(): compression {comp} is not supported (NONE, ZIPS and ZIP are)")
    x0, y0, x1, y1 = struct.unpack("<iiii", attrs["dataWindow"][1])
    W, H = x1 - x0 + 1, y1 - y0 + 1
    ch, p, raw = [], 0, attrs["channels"][1]
    while raw[p] != 0:
        e = raw.index(b"\0", p); name = raw[p:e].decode(); p = e + 1
        ptype, _, _, _, _, xs, ys = struct.unpack_from("<iBBBBii", raw, p); p += 16
        if xs != 1 or ys != 1:
            raise RuntimeError("read_exr(): subsampled channels are not supported")
        ch.append((name, _PIXEL_TYPES[ptype]))
    lines = {0: 1, 2: 1, 3: 16}[comp]
    n_blocks = (H + lines - 1) // lines
    offs = np.frombuffer(buf, "<u8", n_blocks, pos)
    row_bytes = sum(dt.itemsize for _, dt in ch) * W
    out = np.zeros((H, W, len(ch)), np.float32)
    for o in offs:
        y, size = struct.unpack_from("<ii", buf, int(o))
        rows = min(lines, y0 + H - y)
        data = buf[int(o) + 8:int(o) + 8 + size]
        if comp:
            data = _unzip(data, rows * row_bytes)
        q = 0
        for r in range(rows):
            for c, (_, dt) in enumerate(ch):
                out[y - y0 + r, :, c] = np.frombuffer(data, dt, W, q).astype(np.float32)
                q += W * dt.itemsize
    names = [n for n, _ in ch]
    pref = [n for n in ("R", "G", "B", "A") if n in names]
    order = [names.index(n) for n in pref] + [i for i, n in enumerate(names) if n not in pref]
    return out[:, :, order], [names[i] for i in order]


def write_bitmap(path, data, write_async=False):
    """`mi.util.write_bitmap` (`src/python/python/util.py`): .exr keeps linear float32; .png / .jpg get the sRGB transfer
    curve and 8-bit quantisation (Bitmap.convert(PixelFormat.RGB, UInt8, srgb_gamma=True)); .npy as is."""
    a = np.asarray(data)
    ext = path.rsplit(".", 1)[-1].lower()
    if ext == "exr":
        return write_exr(path, a)
    if ext == "npy":
        return np.save(path, a)
    if ext in ("png", "jpg", "jpeg"):
        from PIL import Image
        x = np.nan_to_num(a.astype(np.float32), nan=0.0, posinf=1.0, neginf=0.0).clip(0.0, 1.0)
        srgb = np.where(x <= 0.0031308, x * 12.92, 1.055 * np.power(x, 1 / 2.4) - 0.055)
        return Image.fromarray((srgb * 255.0 + 0.5).astype(np.uint8).squeeze()).save(path)
    raise RuntimeError(f'write_bitmap(): unsupported file extension ".{ext}"')
