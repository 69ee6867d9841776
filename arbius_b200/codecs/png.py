"""Deterministic PNG writer.

The solution CID covers the PNG bytes (reference: miner/src/index.ts:867-874 writes the decoded
container output verbatim; miner/src/models.ts:43-52 hashes it), so the encoder must emit the same
bytes for the same pixels on every box and every run. We therefore fix every degree of freedom:
8-bit RGB / RGBA / grey, filter type 0 on every scanline, one IDAT chunk, no ancillary chunks, and a
deflate stream produced by our own encoder in ``codecs/deflate.py`` (stored or fixed-Huffman with a
pinned match finder) instead of whatever zlib build the host has.  SURVEY §7.3 item 3, §8 S9.
"""
from __future__ import annotations

import struct
import zlib

import numpy as np

from .deflate import deflate_fixed, deflate_stored, zlib_wrap

_SIG = b"\x89PNG\r\n\x1a\n"


def _chunk(tag: bytes, data: bytes) -> bytes:
    return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)


def encode_png(img: np.ndarray, mode: str = "fixed") -> bytes:
    """img uint8 [H, W] / [H, W, 3] / [H, W, 4] -> PNG bytes. mode: 'fixed' (default) or 'stored'."""
    a = np.ascontiguousarray(img)
    if a.dtype != np.uint8:
        raise TypeError("PNG encoder takes uint8")
    if a.ndim == 2:
        a = a[:, :, None]
    h, w, c = a.shape
    ctype = {1: 0, 3: 2, 4: 6}[c]
    raw = np.empty((h, 1 + w * c), dtype=np.uint8)
    raw[:, 0] = 0                       # filter type None on every row
    raw[:, 1:] = a.reshape(h, w * c)
    payload = raw.tobytes()
    body = deflate_stored(payload) if mode == "stored" else deflate_fixed(payload)
    idat = zlib_wrap(body, payload)
    ihdr = struct.pack(">IIBBBBB", w, h, 8, ctype, 0, 0, 0)
    return _SIG + _chunk(b"IHDR", ihdr) + _chunk(b"IDAT", idat) + _chunk(b"IEND", b"")


def decode_png(data: bytes) -> np.ndarray:
    """Minimal decoder for our own output (filter 0 only) — used by tests and the website mock."""
    assert data[:8] == _SIG
    pos, idat, hdr = 8, b"", None
    while pos < len(data):
        (n,) = struct.unpack(">I", data[pos:pos + 4])
        tag = data[pos + 4:pos + 8]
        body = data[pos + 8:pos + 8 + n]
        if tag == b"IHDR":
            hdr = struct.unpack(">IIBBBBB", body)
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    w, h, _, ctype, *_ = hdr
    c = {0: 1, 2: 3, 6: 4}[ctype]
    raw = np.frombuffer(zlib.decompress(idat), dtype=np.uint8).reshape(h, 1 + w * c)
    assert (raw[:, 0] == 0).all()
    out = raw[:, 1:].reshape(h, w, c)
    return out[:, :, 0] if c == 1 else out
