"""Pinned deflate encoders for the deterministic PNG writer.

``deflate_fixed`` = one final block, fixed Huffman codes, greedy LZ77 with a single-probe 15-bit hash
table (no lazy matching, no chain) — every choice is fixed so the bit stream is a pure function of
the input. The native twin is ``ab_deflate_fixed`` in csrc/host_codecs.cu (same algorithm bit for
bit; used when the library is built because pure Python is ~1 s/MB). ``deflate_stored`` is the
zero-compression alternative (BTYPE=00 blocks of 65535 bytes).
"""
from __future__ import annotations

import ctypes as C
import struct
import zlib

_LEN_BASE = [3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258]
_LEN_EXTRA = [0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0]
_DIST_BASE = [1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073,
              4097, 6145, 8193, 12289, 16385, 24577]
_DIST_EXTRA = [0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13]


def _rev(v: int, n: int) -> int:
    r = 0
    for _ in range(n):
        r = (r << 1) | (v & 1)
        v >>= 1
    return r


def _litlen_code(sym: int):
    if sym < 144:
        return _rev(0x30 + sym, 8), 8
    if sym < 256:
        return _rev(0x190 + sym - 144, 9), 9
    if sym < 280:
        return _rev(sym - 256, 7), 7
    return _rev(0xC0 + sym - 280, 8), 8


_LITLEN = [_litlen_code(s) for s in range(288)]
_DIST = [(_rev(d, 5), 5) for d in range(30)]


def deflate_fixed_py(data: bytes) -> bytes:
    out = bytearray()
    acc = 0
    nbits = 0

    def put(v: int, n: int):
        nonlocal acc, nbits
        acc |= v << nbits
        nbits += n
        while nbits >= 8:
            out.append(acc & 0xFF)
            acc >>= 8
            nbits -= 8

    put(1, 1)
    put(1, 2)
    n = len(data)
    table = [-1] * (1 << 15)
    i = 0
    while i < n:
        mlen = mdist = 0
        if i + 2 < n:
            key = (data[i] << 16) | (data[i + 1] << 8) | data[i + 2]
            h = ((key * 2654435761) & 0xFFFFFFFF) >> 17
            cand = table[h]
            table[h] = i
            if cand >= 0 and i - cand <= 32768 and data[cand:cand + 3] == data[i:i + 3]:
                maxl = min(n - i, 258)
                l = 3
                while l < maxl and data[cand + l] == data[i + l]:
                    l += 1
                mlen, mdist = l, i - cand
        if mlen >= 3:
            li = 28
            while _LEN_BASE[li] > mlen:
                li -= 1
            put(*_LITLEN[257 + li])
            if _LEN_EXTRA[li]:
                put(mlen - _LEN_BASE[li], _LEN_EXTRA[li])
            di = 29
            while _DIST_BASE[di] > mdist:
                di -= 1
            put(*_DIST[di])
            if _DIST_EXTRA[di]:
                put(mdist - _DIST_BASE[di], _DIST_EXTRA[di])
            i += mlen
        else:
            put(*_LITLEN[data[i]])
            i += 1
    put(*_LITLEN[256])
    if nbits:
        out.append(acc & 0xFF)
    return bytes(out)


_native = None


def _native_lib():
    global _native
    if _native is None:
        try:
            from ..ops import native
            if native.available():
                lib = C.CDLL(str(native.lib_path()))
                lib.ab_deflate_fixed.restype = C.c_size_t
                lib.ab_deflate_fixed.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p, C.c_size_t]
                _native = lib
            else:
                _native = False
        except OSError:
            _native = False
    return _native


def deflate_fixed(data: bytes) -> bytes:
    lib = _native_lib()
    if not lib:
        return deflate_fixed_py(data)
    cap = len(data) * 9 // 8 + 64
    buf = C.create_string_buffer(cap)
    n = lib.ab_deflate_fixed(data, len(data), buf, cap)
    if n == C.c_size_t(-1).value:
        raise RuntimeError("deflate output buffer too small")
    return buf.raw[:n]


def deflate_stored(data: bytes) -> bytes:
    out = bytearray()
    n = len(data)
    if n == 0:
        return b"\x01\x00\x00\xff\xff"
    for off in range(0, n, 65535):
        blk = data[off:off + 65535]
        final = 1 if off + 65535 >= n else 0
        out += struct.pack("<BHH", final, len(blk), len(blk) ^ 0xFFFF) + blk
    return bytes(out)


def zlib_wrap(deflate_body: bytes, raw: bytes) -> bytes:
    """RFC 1950 container: CMF/FLG 0x78 0x01, body, Adler-32 of the uncompressed data."""
    return b"\x78\x01" + deflate_body + struct.pack(">I", zlib.adler32(raw) & 0xFFFFFFFF)
