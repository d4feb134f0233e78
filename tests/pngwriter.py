"""A small PNG encoder for tests (the import path's decoder is checked against files written here).
Every colour type / bit depth, per-row filters, Adam7 interlace, PLTE / tRNS, split IDAT chunks and the three
deflate block kinds (stored, fixed, dynamic) can be chosen.  Written from the PNG specification; uses zlib."""
import struct
import zlib

import numpy as np

CHANNELS = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}


def _chunk(kind: bytes, body: bytes) -> bytes:
    return struct.pack(">I", len(body)) + kind + body + struct.pack(">I", zlib.crc32(kind + body) & 0xFFFFFFFF)


def _pack_rows(samples: np.ndarray, depth: int) -> list[bytes]:
    """samples: (h, w, c) integers < 2**depth -> one packed byte string per row."""
    h, w, c = samples.shape
    rows = []
    for y in range(h):
        flat = samples[y].reshape(-1).astype(np.uint32)
        if depth == 8:
            rows.append(flat.astype(np.uint8).tobytes())
        elif depth == 16:
            rows.append(flat.astype(">u2").tobytes())
        else:
            per = 8 // depth
            pad = (-len(flat)) % per
            f = np.concatenate([flat, np.zeros(pad, np.uint32)]).reshape(-1, per)
            shifts = np.arange(per - 1, -1, -1, dtype=np.uint32) * depth
            rows.append((f << shifts).sum(axis=1).astype(np.uint8).tobytes())
    return rows


def _paeth(a, b, c):
    p = a + b - c
    pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
    return a if pa <= pb and pa <= pc else (b if pb <= pc else c)


def _filter(rows: list[bytes], bpp: int, filters) -> bytes:
    out = bytearray()
    prev = bytes(len(rows[0])) if rows else b""
    for y, row in enumerate(rows):
        f = filters[y % len(filters)]
        out.append(f)
        line = bytearray(len(row))
        for i, v in enumerate(row):
            a = row[i - bpp] if i >= bpp else 0
            b = prev[i]
            c = prev[i - bpp] if i >= bpp else 0
            pred = 0 if f == 0 else a if f == 1 else b if f == 2 else (a + b) >> 1 if f == 3 else _paeth(a, b, c)
            line[i] = (v - pred) & 0xFF
        out += line
        prev = row
    return bytes(out)


def encode(samples: np.ndarray, color: int, depth: int, *, interlace=False, filters=(0,), palette=None, trns=None, level=6,
           strategy=zlib.Z_DEFAULT_STRATEGY, idat_split=0, extra_chunks=()) -> bytes:
    """samples: (h, w, channels) raw sample integers (palette indices for colour type 3)."""
    samples = np.asarray(samples)
    h, w, c = samples.shape
    assert c == CHANNELS[color]
    bpp = max(1, c * depth // 8)
    if interlace:
        x0, y0, dx, dy = (0, 4, 0, 2, 0, 1, 0), (0, 0, 4, 0, 2, 0, 1), (8, 8, 4, 4, 2, 2, 1), (8, 8, 8, 4, 4, 2, 2)
        raw = b""
        for p in range(7):
            sub = samples[y0[p]::dy[p], x0[p]::dx[p]]
            if sub.shape[0] and sub.shape[1]:
                raw += _filter(_pack_rows(sub, depth), bpp, filters)
    else:
        raw = _filter(_pack_rows(samples, depth), bpp, filters)
    comp = zlib.compressobj(level, zlib.DEFLATED, 15, 8, strategy)
    data = comp.compress(raw) + comp.flush()
    out = b"\x89PNG\r\n\x1a\n" + _chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, color, 0, 0, 1 if interlace else 0))
    for kind, body in extra_chunks:
        out += _chunk(kind, body)
    if palette is not None:
        out += _chunk(b"PLTE", np.asarray(palette, np.uint8).tobytes())
    if trns is not None:
        out += _chunk(b"tRNS", bytes(trns))
    if idat_split:
        for i in range(0, len(data), idat_split):
            out += _chunk(b"IDAT", data[i:i + idat_split])
    else:
        out += _chunk(b"IDAT", data)
    return out + _chunk(b"IEND", b"")


def encode_rgba8(rgba: np.ndarray, **kw) -> bytes:
    return encode(np.asarray(rgba, np.uint8), 6, 8, **kw)
