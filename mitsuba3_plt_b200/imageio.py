"""Image files either side of the render path: what the reference's scripts call through
`mi.util.write_bitmap(path, image)` / `mi.Bitmap(path)` (`scripts/rendering/main-headless.py:60-73`, `src/core/bitmap.cpp`).

OpenEXR is written and read here without the OpenEXR library (it is not in the image; the reference's `ext/openexr` submodule is
empty): single-part scanline files, FLOAT or HALF channels, compression NONE / ZIPS / ZIP on write, NONE / ZIPS / ZIP / PIZ on read (PIZ is what Mitsuba itself writes) -- enough for
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


# ---- PIZ (what Mitsuba's own Bitmap::write produces): Huffman-coded 16-bit words, 2-D Haar-like wavelet, value LUT ----------
_HUF_ENCSIZE = (1 << 16) + 1
_SHORT_ZEROCODE_RUN, _LONG_ZEROCODE_RUN, _SHORTEST_LONG_RUN = 59, 63, 6


def _huf_uncompress(comp, n_raw):
    """OpenEXR hufUncompress: canonical Huffman code over 16-bit symbols, code lengths packed in 6 bits with zero runs, symbol
    iM is a run-length escape (8-bit repeat count of the previous output)."""
    im, iM, _, n_bits, _ = struct.unpack_from("<iiiii", comp, 0)
    data = np.frombuffer(comp, np.uint8, len(comp) - 20, 20)
    # -- unpack the code-length table (MSB-first bit reader)
    pos, acc, nacc = 0, 0, 0
    lengths = np.zeros(_HUF_ENCSIZE, np.int64)
    sym = im
    while sym <= iM:
        while nacc < 6:
            acc = (acc << 8) | int(data[pos]); pos += 1; nacc += 8
        nacc -= 6
        l = (acc >> nacc) & 63
        if l == _LONG_ZEROCODE_RUN:
            while nacc < 8:
                acc = (acc << 8) | int(data[pos]); pos += 1; nacc += 8
            nacc -= 8
            sym += ((acc >> nacc) & 255) + _SHORTEST_LONG_RUN
        elif l >= _SHORT_ZEROCODE_RUN:
            sym += l - _SHORT_ZEROCODE_RUN + 2
        else:
            lengths[sym] = l
            sym += 1
        acc &= (1 << nacc) - 1
    # -- canonical codes (hufCanonicalCodeTable): within a length, codes increase with the symbol index
    count = np.bincount(lengths, minlength=59)
    start, c = [0] * 59, 0
    for i in range(58, 0, -1):
        nc = (c + int(count[i])) >> 1
        start[i] = c
        c = nc
    table = {}
    for symv in np.nonzero(lengths)[0]:
        l = int(lengths[symv])
        table[(l, start[l])] = int(symv)
        start[l] += 1
    max_len = int(lengths.max()) if table else 0
    # -- decode
    bits = np.unpackbits(data[pos:pos + (n_bits + 7) // 8])[:n_bits]
    out = np.zeros(n_raw, np.uint16)
    n_out, code, l, i = 0, 0, 0, 0
    bl = bits.tolist()
    while i < n_bits and n_out < n_raw:
        code = (code << 1) | bl[i]; i += 1; l += 1
        symv = table.get((l, code))
        if symv is None:
            if l > max_len:
                raise RuntimeError("read_exr(): corrupt PIZ Huffman stream")
            continue
        if symv == iM:   # run-length escape
            rep = 0
            for _ in range(8):
                rep = (rep << 1) | bl[i]; i += 1
            out[n_out:n_out + rep] = out[n_out - 1]
            n_out += rep
        else:
            out[n_out] = symv
            n_out += 1
        code, l = 0, 0
    if n_out != n_raw:
        raise RuntimeError("read_exr(): PIZ block decoded to the wrong size")
    return out


def _wdec14(l, h):
    ls, hs = l.astype(np.int16).astype(np.int32), h.astype(np.int16).astype(np.int32)
    ai = ls + (hs & 1) + (hs >> 1)
    return (ai & 0xFFFF).astype(np.uint16), ((ai - hs) & 0xFFFF).astype(np.uint16)


def _wdec16(l, h):
    m, d = l.astype(np.int32), h.astype(np.int32)
    bb = (m - (d >> 1)) & 0xFFFF
    aa = (d + bb - (1 << 15)) & 0xFFFF
    return aa.astype(np.uint16), bb.astype(np.uint16)


def _wav2_decode(a, max_value):
    """OpenEXR wav2Decode on a (ny, nx) uint16 array, in place (vectorised per level)."""
    ny, nx = a.shape
    dec = _wdec14 if max_value < (1 << 14) else _wdec16
    n = min(nx, ny)
    p = 1
    while p <= n:
        p <<= 1
    p >>= 1
    p2 = p
    p >>= 1
    while p >= 1:
        ys = np.arange(0, ny - p2 + 1, p2) if ny - p2 >= 0 else np.zeros(0, np.int64)
        xs = np.arange(0, nx - p2 + 1, p2) if nx - p2 >= 0 else np.zeros(0, np.int64)
        if len(ys) and len(xs):
            Y, X = np.ix_(ys, xs)
            i00, i10 = dec(a[Y, X], a[Y + p, X])
            i01, i11 = dec(a[Y, X + p], a[Y + p, X + p])
            a[Y, X], a[Y, X + p] = dec(i00, i01)
            a[Y + p, X], a[Y + p, X + p] = dec(i10, i11)
        if nx & p and len(ys):      # odd column: the block that starts after the last full one
            xo = (xs[-1] + p2) if len(xs) else 0
            v0, v1 = dec(a[ys, xo], a[ys + p, xo])
            a[ys, xo], a[ys + p, xo] = v0, v1
        if ny & p and len(xs):      # odd row
            yo = (ys[-1] + p2) if len(ys) else 0
            v0, v1 = dec(a[yo, xs], a[yo, xs + p])
            a[yo, xs], a[yo, xs + p] = v0, v1
        p2 = p
        p >>= 1


def _unpiz(block, rows, W, ch):
    """one PIZ block -> scanline-interleaved raw bytes (the layout NONE / ZIP blocks have)"""
    sizes = [dt.itemsize // 2 for _, dt in ch]
    n_words = sum(sizes) * W * rows
    min_nz, max_nz = struct.unpack_from("<HH", block, 0)
    bitmap = np.zeros(8192, np.uint8)
    pos = 4
    if min_nz <= max_nz:
        bitmap[min_nz:max_nz + 1] = np.frombuffer(block, np.uint8, max_nz - min_nz + 1, pos)
        pos += max_nz - min_nz + 1
    length, = struct.unpack_from("<i", block, pos)
    pos += 4
    words = _huf_uncompress(block[pos:pos + length], n_words)
    present = np.unpackbits(bitmap, bitorder="little").astype(bool)
    present[0] = True
    lut = np.zeros(1 << 16, np.uint16)
    values = np.nonzero(present)[0]
    lut[:len(values)] = values
    max_value = len(values) - 1
    planes, off = [], 0
    for sz in sizes:            # per channel: (rows, W * size) words, wavelet-decoded per 16-bit half
        n = rows * W * sz
        a = words[off:off + n].reshape(rows, W, sz).copy()
        for j in range(sz):
            plane = np.ascontiguousarray(a[:, :, j])
            _wav2_decode(plane, max_value)
            a[:, :, j] = plane
        planes.append(lut[a].reshape(rows, W * sz))
        off += n
    return np.concatenate(planes, axis=1).astype("<u2").tobytes()   # row by row, channel after channel


def read_exr(path):
    """Returns (array (H, W, C) float32, channel names); channels R, G, B[, A] come first in that order when present."""
    buf = open(path, "rb").read()
    attrs, pos = _read_header(buf)
    comp = attrs["compression"][1][0]
    if comp not in (0, 2, 3, 4):
        raise RuntimeError(f"read_exr(): compression {comp} is not supported (NONE, ZIPS, ZIP and PIZ are)")
    x0, y0, x1, y1 = struct.unpack("<iiii", attrs["dataWindow"][1])
    W, H = x1 - x0 + 1, y1 - y0 + 1
    ch, p, raw = [], 0, attrs["channels"][1]
    while raw[p] != 0:
        e = raw.index(b"\0", p); name = raw[p:e].decode(); p = e + 1
        ptype, _, _, _, _, xs, ys = struct.unpack_from("<iBBBBii", raw, p); p += 16
        if xs != 1 or ys != 1:
            raise RuntimeError("read_exr(): subsampled channels are not supported")
        ch.append((name, _PIXEL_TYPES[ptype]))
    lines = {0: 1, 2: 1, 3: 16, 4: 32}[comp]
    n_blocks = (H + lines - 1) // lines
    offs = np.frombuffer(buf, "<u8", n_blocks, pos)
    row_bytes = sum(dt.itemsize for _, dt in ch) * W
    out = np.zeros((H, W, len(ch)), np.float32)
    for o in offs:
        y, size = struct.unpack_from("<ii", buf, int(o))
        rows = min(lines, y0 + H - y)
        data = buf[int(o) + 8:int(o) + 8 + size]
        if comp == 4:
            data = _unpiz(data, rows, W, ch) if size < rows * row_bytes else data
        elif comp:
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
