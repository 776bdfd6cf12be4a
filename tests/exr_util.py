"""TEST INFRASTRUCTURE: checkers for the output-encode step (not imported by the product path).
half_bits_tinyexr restates tinyexr's float -> half conversion in numpy; read_exr_half decodes the EXR files
both the reference (ZIP) and the GPU encoder (uncompressed) write."""
import numpy as np


def half_bits_tinyexr(img):
    """numpy restatement of tinyexr's float_to_half_full (tinyexr.h:889-924), the conversion the reference's
    EXR output goes through: float denormals flush to zero, round half up on mantissa bit 12, the carry may
    run into the exponent.  Returns uint16 bit patterns.  (Checker for tests, not a product path.)"""
    u = np.ascontiguousarray(img, dtype=np.float32).view(np.uint32).astype(np.int64)
    sign = (u >> 16) & 0x8000
    ex = (u >> 23) & 0xff
    man = u & 0x7fffff
    newexp = ex - 127 + 15
    normal = ((newexp << 10) | (man >> 13)) + ((man & 0x1000) != 0)
    sh = np.clip(14 - newexp, 0, 62)
    mant = man | 0x800000
    den = (mant >> sh) + (((mant >> np.clip(13 - newexp, 0, 62)) & 1) != 0)
    den = np.where(14 - newexp <= 24, den, 0)
    o = np.where(newexp >= 31, 31 << 10, np.where(newexp <= 0, den, normal))
    o = np.where(ex == 255, (31 << 10) | np.where(man != 0, 0x200, 0), o)
    o = np.where(ex == 0, 0, o)
    return (sign | (o & 0x7fff)).astype(np.uint16)


def read_exr_half(data):
    """Minimal reader for single-part scanline OpenEXR files with HALF channels, uncompressed or ZIP
    (what tinyexr's SaveEXR writes): returns ({channel: uint16 bit patterns (h, w)}, attributes)."""
    import struct
    import zlib
    assert data[:4] == b"\x76\x2f\x31\x01", "not an OpenEXR file"
    pos = 8
    attrs = {}
    while data[pos] != 0:
        e = data.index(b"\0", pos); name = data[pos:e].decode(); pos = e + 1
        e = data.index(b"\0", pos); typ = data[pos:e].decode(); pos = e + 1
        size = struct.unpack_from("<i", data, pos)[0]; pos += 4
        attrs[name] = (typ, data[pos:pos + size]); pos += size
    pos += 1
    chans = []
    cl = attrs["channels"][1]; p = 0
    while cl[p] != 0:
        e = cl.index(b"\0", p); cname = cl[p:e].decode(); p = e + 1
        ptype = struct.unpack_from("<i", cl, p)[0]; p += 16
        assert ptype == 1, "HALF channels only"
        chans.append(cname)
    comp = attrs["compression"][1][0]
    x0, y0, x1, y1 = struct.unpack("<4i", attrs["dataWindow"][1])
    w, h = x1 - x0 + 1, y1 - y0 + 1
    lines_per_block = {0: 1, 2: 1, 3: 16}[comp]
    nblocks = (h + lines_per_block - 1) // lines_per_block
    offsets = struct.unpack_from("<%dQ" % nblocks, data, pos)
    out = {c: np.zeros((h, w), dtype=np.uint16) for c in chans}
    for off in offsets:
        y, size = struct.unpack_from("<ii", data, off)
        raw = data[off + 8:off + 8 + size]
        nl = min(lines_per_block, y0 + h - y)
        expect = nl * len(chans) * w * 2
        if comp != 0 and size < expect:
            t = np.frombuffer(zlib.decompress(raw), dtype=np.uint8).astype(np.int32)
            t = (np.cumsum(t - np.concatenate(([0], np.full(len(t) - 1, 128)))) & 0xff).astype(np.uint8)     # predictor
            half = (len(t) + 1) // 2
            raw = np.empty(len(t), dtype=np.uint8)
            raw[0::2] = t[:half]; raw[1::2] = t[half:]                                                      # de-interleave
            raw = raw.tobytes()
        arr = np.frombuffer(raw, dtype="<u2").reshape(nl, len(chans), w)
        for ci, c in enumerate(chans):
            out[c][y - y0:y - y0 + nl] = arr[:, ci]
    return out, attrs
