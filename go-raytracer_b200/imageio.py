"""Raster output: PNG (what the reference writes, cmd/gml/main.go:383-390,
cmd/example/main.go:21-28) and binary PPM "P6" (the ICFP contract that
RenderArgs.File names).  Dependency-free (zlib + struct) so it works
on a bare GPU box; also a small PNG reader for the golden images."""
import struct
import zlib

import numpy as np


def write_png(path, rgba):
    """rgba: uint8 array [H, W, 4] (image.RGBA.Pix layout)."""
    a = np.ascontiguousarray(rgba, dtype=np.uint8)
    h, w, c = a.shape
    assert c == 4
    raw = np.empty((h, 1 + w * 4), dtype=np.uint8)
    raw[:, 0] = 0
    raw[:, 1:] = a.reshape(h, w * 4)

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n")
        f.write(chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 6, 0, 0, 0)))
        f.write(chunk(b"IDAT", zlib.compress(raw.tobytes(), 6)))
        f.write(chunk(b"IEND", b""))


def write_ppm(path, rgba):
    a = np.ascontiguousarray(rgba, dtype=np.uint8)
    h, w, _ = a.shape
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (w, h))
        f.write(a[:, :, :3].tobytes())


def read_png(path):
    """8-bit, non-interlaced gray / RGB / RGBA / palette PNG -> uint8 [H, W, 4]."""
    with open(path, "rb") as f:
        data = f.read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n", "not a PNG"
    pos, idat, plte, trns = 8, [], None, None
    while pos < len(data):
        n, tag = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        pos += 12 + n
        if tag == b"IHDR":
            w, h, depth, ctype, _, _, interlace = struct.unpack(">IIBBBBB", body)
        elif tag == b"IDAT":
            idat.append(body)
        elif tag == b"PLTE":
            plte = np.frombuffer(body, dtype=np.uint8).reshape(-1, 3)
        elif tag == b"tRNS":
            trns = body
        elif tag == b"IEND":
            break
    assert depth == 8 and interlace == 0, "only 8-bit non-interlaced PNGs"
    ch = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
    raw = np.frombuffer(zlib.decompress(b"".join(idat)), dtype=np.uint8).reshape(h, 1 + w * ch)
    out = np.zeros((h, w * ch), dtype=np.uint8)
    prev = np.zeros(w * ch, dtype=np.int32)
    for y in range(h):
        ft = raw[y, 0]
        line = raw[y, 1:].astype(np.int32)
        if ft == 0:
            cur = line
        elif ft == 2:
            cur = (line + prev) & 255
        elif ft == 1:
            cur = line.reshape(w, ch)
            cur = (np.cumsum(cur, axis=0) & 255).reshape(-1)
        else:
            cur = np.zeros(w * ch, dtype=np.int32)
            lp = line.tolist()
            pp = prev.tolist()
            cl = [0] * (w * ch)
            for i in range(w * ch):
                a = cl[i - ch] if i >= ch else 0
                b = pp[i]
                c = pp[i - ch] if i >= ch else 0
                if ft == 3:
                    pred = (a + b) >> 1
                else:
                    p = a + b - c
                    pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
                    pred = a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)
                cl[i] = (lp[i] + pred) & 255
            cur = np.array(cl, dtype=np.int32)
        out[y] = cur
        prev = cur
    px = out.reshape(h, w, ch)
    rgba = np.empty((h, w, 4), dtype=np.uint8)
    rgba[..., 3] = 255
    if ctype == 2:
        rgba[..., :3] = px
    elif ctype == 6:
        rgba[:] = px
    elif ctype == 0:
        rgba[..., :3] = px
    elif ctype == 4:
        rgba[..., :3] = px[..., :1]
        rgba[..., 3] = px[..., 1]
    else:
        rgba[..., :3] = plte[px[..., 0]]
        if trns is not None:
            alpha = np.full(256, 255, dtype=np.uint8)
            alpha[:len(trns)] = np.frombuffer(trns, dtype=np.uint8)
            rgba[..., 3] = alpha[px[..., 0]]
    return rgba
