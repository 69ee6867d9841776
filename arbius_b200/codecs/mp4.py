"""Deterministic MP4 writer: Motion-JPEG samples in an ISO base media file.

The video templates name ``out-1.mp4`` as their output (reference: templates/zeroscopev2xl.json, damo.json,
robust_video_matting.json ``output[0]``) and the solution CID covers the file bytes, so the writer must be a pure
function of (frames, fps): no ffmpeg/OpenCV, no timestamps from the clock, no encoder heuristics. Frames are coded
with the library's own integer baseline JPEG encoder (csrc/host_jpeg.cu) and muxed here as an ``mp4v`` sample
entry whose ES descriptor declares objectTypeIndication 0x6C (JPEG) — the registered way to carry MJPEG in MP4,
played by ffmpeg/VLC/browsers with an MJPEG decoder. Layout: ftyp, mdat, moov (one video track, one chunk).
SURVEY §7.3 item 3, §8 row S10.
"""
from __future__ import annotations

import ctypes as C
import struct
from typing import List, Sequence

import numpy as np


def _box(tag: bytes, payload: bytes) -> bytes:
    return struct.pack(">I", 8 + len(payload)) + tag + payload


def _full(tag: bytes, version: int, flags: int, payload: bytes) -> bytes:
    return _box(tag, struct.pack(">I", (version << 24) | flags) + payload)


_MATRIX = struct.pack(">9I", 0x10000, 0, 0, 0, 0x10000, 0, 0, 0, 0x40000000)


def _descr(tag: int, payload: bytes) -> bytes:
    n = len(payload)
    assert n < 128 * 128
    size = bytes([0x80 | (n >> 7), n & 0x7F]) if n >= 128 else bytes([n])
    return bytes([tag]) + size + payload


def encode_jpeg(frame: np.ndarray, quality: int = 90) -> bytes:
    """uint8 [H, W, 3] -> baseline JPEG bytes through the native integer encoder."""
    from ..ops import native
    lib = C.CDLL(str(native.lib_path())) if native.available() else native.load()
    lib.ab_jpeg_encode.restype = C.c_size_t
    lib.ab_jpeg_encode.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t]
    a = np.ascontiguousarray(frame, dtype=np.uint8)
    h, w, c = a.shape
    assert c == 3
    cap = h * w * 3 + 4096
    buf = C.create_string_buffer(cap)
    n = lib.ab_jpeg_encode(a.ctypes.data, w, h, quality, buf, cap)
    if n == C.c_size_t(-1).value:
        raise RuntimeError("jpeg output buffer too small")
    return buf.raw[:n]


def mux_mjpeg_mp4(samples: Sequence[bytes], width: int, height: int, fps: int) -> bytes:
    n = len(samples)
    timescale = fps * 512                      # integer ticks per frame = 512
    delta = 512
    duration = n * delta
    ftyp = _box(b"ftyp", b"isom" + struct.pack(">I", 512) + b"isomiso2mp41")
    mdat = _box(b"mdat", b"".join(samples))
    first_off = len(ftyp) + 8
    # --- sample table ---
    dsi = b""
    dec_cfg = _descr(0x04, bytes([0x6C, 0x11]) + b"\x00\x00\x00" + struct.pack(">II", 0, 0) + dsi)
    es = _descr(0x03, struct.pack(">HB", 1, 0) + dec_cfg + _descr(0x06, b"\x02"))
    esds = _full(b"esds", 0, 0, es)
    mp4v = _box(b"mp4v", b"\x00" * 6 + struct.pack(">H", 1) + b"\x00" * 16 + struct.pack(">HH", width, height) +
                struct.pack(">II", 0x00480000, 0x00480000) + struct.pack(">I", 0) + struct.pack(">H", 1) +
                b"\x00" * 32 + struct.pack(">H", 0x18) + struct.pack(">h", -1) + esds)
    stsd = _full(b"stsd", 0, 0, struct.pack(">I", 1) + mp4v)
    stts = _full(b"stts", 0, 0, struct.pack(">III", 1, n, delta))
    stsc = _full(b"stsc", 0, 0, struct.pack(">IIII", 1, 1, n, 1))
    stsz = _full(b"stsz", 0, 0, struct.pack(">II", 0, n) + b"".join(struct.pack(">I", len(s)) for s in samples))
    stco = _full(b"stco", 0, 0, struct.pack(">II", 1, first_off))
    stbl = _box(b"stbl", stsd + stts + stsc + stsz + stco)
    dref = _full(b"dref", 0, 0, struct.pack(">I", 1) + _full(b"url ", 0, 1, b""))
    minf = _box(b"minf", _full(b"vmhd", 0, 1, b"\x00" * 8) + _box(b"dinf", dref) + stbl)
    hdlr = _full(b"hdlr", 0, 0, struct.pack(">I", 0) + b"vide" + b"\x00" * 12 + b"VideoHandler\x00")
    mdhd = _full(b"mdhd", 0, 0, struct.pack(">IIIIHH", 0, 0, timescale, duration, 0x55C4, 0))
    mdia = _box(b"mdia", mdhd + hdlr + minf)
    tkhd = _full(b"tkhd", 0, 3, struct.pack(">IIIII", 0, 0, 1, 0, duration) + b"\x00" * 8 +
                 struct.pack(">hhhH", 0, 0, 0, 0) + _MATRIX + struct.pack(">II", width << 16, height << 16))
    trak = _box(b"trak", tkhd + mdia)
    mvhd = _full(b"mvhd", 0, 0, struct.pack(">IIII", 0, 0, timescale, duration) + struct.pack(">IH", 0x10000, 0x100) +
                 b"\x00" * 10 + _MATRIX + b"\x00" * 24 + struct.pack(">I", 2))
    moov = _box(b"moov", mvhd + trak)
    return ftyp + mdat + moov


def encode_mp4(frames: np.ndarray, fps: int, quality: int = 90) -> bytes:
    """frames uint8 [F, H, W, 3] -> MP4 bytes (deterministic)."""
    f = np.asarray(frames)
    assert f.ndim == 4 and f.shape[-1] == 3 and f.dtype == np.uint8
    samples = [encode_jpeg(f[i], quality) for i in range(f.shape[0])]
    return mux_mjpeg_mp4(samples, f.shape[2], f.shape[1], int(fps))


def parse_boxes(data: bytes, start: int = 0, end: int | None = None) -> List[tuple]:
    """[(tag, payload_start, payload_end)] of the boxes in a range — used by tests and the demuxer below."""
    out, pos = [], start
    end = len(data) if end is None else end
    while pos + 8 <= end:
        size, tag = struct.unpack(">I4s", data[pos:pos + 8])
        if size < 8:
            break
        out.append((tag, pos + 8, pos + size))
        pos += size
    return out


def demux_samples(data: bytes) -> List[bytes]:
    """Recover the JPEG samples from a file written by ``mux_mjpeg_mp4`` (tests, RVM input reading)."""
    top = {t: (a, b) for t, a, b in parse_boxes(data)}

    def find(path, rng):
        a, b = rng
        for name in path:
            for t, s, e in parse_boxes(data, a, b):
                if t == name:
                    a, b = s, e
                    break
            else:
                raise ValueError(f"box {name!r} not found")
        return a, b
    stbl = find([b"trak", b"mdia", b"minf", b"stbl"], top[b"moov"])
    sz_a, _ = find([b"stsz"], stbl)
    _, n = struct.unpack(">II", data[sz_a + 4:sz_a + 12])
    sizes = struct.unpack(f">{n}I", data[sz_a + 12:sz_a + 12 + 4 * n])
    co_a, _ = find([b"stco"], stbl)
    off = struct.unpack(">I", data[co_a + 8:co_a + 12])[0]
    out = []
    for s in sizes:
        out.append(data[off:off + s])
        off += s
    return out


def probe_fps(data: bytes, default: int = 30) -> int:
    """frames per second from mdhd timescale / stts sample delta (constant-rate files only)"""
    try:
        rng = (0, len(data))
        for name in (b"moov", b"trak", b"mdia"):
            rng = next((s, e) for t, s, e in parse_boxes(data, *rng) if t == name)
        mdhd = next((s, e) for t, s, e in parse_boxes(data, *rng) if t == b"mdhd")
        timescale = struct.unpack(">I", data[mdhd[0] + 12:mdhd[0] + 16])[0]
        for name in (b"minf", b"stbl", b"stts"):
            rng = next((s, e) for t, s, e in parse_boxes(data, *rng) if t == name)
        delta = struct.unpack(">I", data[rng[0] + 12:rng[0] + 16])[0]
        return max(1, round(timescale / delta))
    except (StopIteration, struct.error, ZeroDivisionError):
        return default
