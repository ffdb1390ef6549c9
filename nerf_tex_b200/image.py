"""Image out: what Logger.render_image / write_image do after the renderer returned (network/logger.py:121-144).

``finish_image`` runs on the device (one kernel: gaussian low-pass + strided downsample of util/interpolate.py:67-82,
premultiplied -> straight alpha, 8-bit quantisation); ``encode_png`` is a small host-side PNG writer (zlib) standing
in for ``tf.io.encode_png``."""

from __future__ import annotations

import ctypes as C
import struct
import zlib
from typing import Optional, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import as_cuda, check, ptr, stream_ptr


def gaussian_taps(downsampling_factor: int, std: float = 0.5) -> np.ndarray:
    """Normalised 1-D weights of util/interpolate.py:67-75 (the 2-D kernel there is their outer product)."""
    size = int(downsampling_factor * std * 6)
    sigma = downsampling_factor * std
    x = np.linspace(-(size - 1) / 2, (size - 1) / 2, size).astype(np.float32) + np.float32(0.5 if size % 2 == 0 else 0.0)
    k = np.exp(-0.5 * (x / np.float32(sigma)) ** 2).astype(np.float32)
    return (k / k.sum()).astype(np.float32)


def finish_image(color_pred, alpha_pred, height: int, width: int, downsampling_factor: int = 1, straight_alpha: bool = True,
                 want_u8: bool = False) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """logger.py:126-137: returns (img[h,w,4] fp32, img_u8[h,w,4] or None)."""
    color = as_cuda(color_pred).reshape(-1, 3).contiguous()
    alpha = as_cuda(alpha_pred).reshape(-1).contiguous()
    assert color.shape[0] == height * width and alpha.shape[0] == height * width
    f = int(downsampling_factor)
    h, w = -(-height // f), -(-width // f)
    img = torch.empty((h, w, 4), dtype=torch.float32, device=color.device)
    u8 = torch.empty((h, w, 4), dtype=torch.uint8, device=color.device) if want_u8 else None
    taps = gaussian_taps(f) if f > 1 else np.ones(1, np.float32)
    arr = (C.c_float * len(taps))(*[float(v) for v in taps])
    with torch.cuda.device(color.device):
        check(_lib.load().ntx_image_finish(ptr(color), ptr(alpha), int(height), int(width), f, arr, len(taps),
                                           int(bool(straight_alpha)), ptr(img), ptr(u8), stream_ptr()))
    return img, u8


def encode_png(img_u8) -> bytes:
    """8-bit RGBA / RGB / grey PNG (filter type 0 on every row)."""
    a = img_u8.detach().cpu().numpy() if isinstance(img_u8, torch.Tensor) else np.asarray(img_u8)
    assert a.dtype == np.uint8 and a.ndim == 3 and a.shape[2] in (1, 3, 4)
    h, w, c = a.shape
    color_type = {1: 0, 3: 2, 4: 6}[c]
    raw = np.concatenate([np.zeros((h, 1), np.uint8), a.reshape(h, w * c)], axis=1).tobytes()

    def chunk(tag: bytes, data: bytes) -> bytes:
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    return (b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, color_type, 0, 0, 0)) +
            chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def decode_png(data: bytes) -> np.ndarray:
    """Inverse of ``encode_png`` for the files it writes (8-bit, non-interlaced, filter 0..4)."""
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, hdr = 8, b"", None
    while pos < len(data):
        n, tag = struct.unpack(">I", data[pos:pos + 4])[0], data[pos + 4:pos + 8]
        body = data[pos + 8:pos + 8 + n]
        if tag == b"IHDR":
            hdr = struct.unpack(">IIBBBBB", body)
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    w, h, depth, ctype = hdr[0], hdr[1], hdr[2], hdr[3]
    c = {0: 1, 2: 3, 6: 4}[ctype]
    assert depth == 8
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + w * c)
    out = np.zeros((h, w * c), np.uint8)
    for y in range(h):
        ft, line = raw[y, 0], raw[y, 1:].astype(np.int32)
        prev = out[y - 1].astype(np.int32) if y else np.zeros(w * c, np.int32)
        if ft == 0:
            rec = line
        elif ft == 2:
            rec = line + prev
        else:
            rec = np.zeros(w * c, np.int32)
            for i in range(w * c):
                a_ = rec[i - c] if i >= c else 0
                b_ = prev[i]
                c_ = prev[i - c] if i >= c else 0
                if ft == 1:
                    p = a_
                elif ft == 3:
                    p = (a_ + b_) // 2
                else:
                    pa, pb, pc = abs(b_ - c_), abs(a_ - c_), abs(a_ + b_ - 2 * c_)
                    p = a_ if pa <= pb and pa <= pc else (b_ if pb <= pc else c_)
                rec[i] = (line[i] + p) & 255
        out[y] = (rec & 255).astype(np.uint8)
    return out.reshape(h, w, c)


def write_image(img_path: str, color_pred, alpha_pred, height: int, width: int, downsampling_factor: int = 1) -> torch.Tensor:
    """render_image + write_image of network/logger.py:121-144 for the PNG case; returns the fp32 image."""
    img, u8 = finish_image(color_pred, alpha_pred, height, width, downsampling_factor, straight_alpha=True, want_u8=True)
    with open(img_path, "wb") as f:
        f.write(encode_png(u8))
    return img
