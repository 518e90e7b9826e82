"""TEST INFRASTRUCTURE: decoders for the two files of the output stage, written from the format descriptions
(PNG via PIL; Radiance .hdr — flat or new-style run-length scanlines — here), so that the reference's files
(stbi_write_png / stbi_write_hdr through Camera::WriteColorAttachment) and the product's can be compared by
PAYLOAD: decoded RGB8 bytes, decoded RGBE quadruples."""
from __future__ import annotations

import io

import numpy as np


def png_rgb8(data: bytes) -> np.ndarray:
    from PIL import Image
    return np.array(Image.open(io.BytesIO(data)).convert("RGB"), dtype=np.uint8)


def hdr_rgbe(data: bytes) -> np.ndarray:
    """[H, W, 4] uint8 RGBE quadruples of a Radiance picture ("-Y H +X W" orientation only)"""
    pos = 0
    while True:
        end = data.index(b"\n", pos)
        line = data[pos:end]
        pos = end + 1
        if line == b"":
            break
    end = data.index(b"\n", pos)
    tok = data[pos:end].split()
    assert tok[0] == b"-Y" and tok[2] == b"+X", tok
    h, w = int(tok[1]), int(tok[3])
    pos = end + 1
    out = np.zeros((h, w, 4), np.uint8)
    for y in range(h):
        if 8 <= w < 32768 and data[pos] == 2 and data[pos + 1] == 2 and (data[pos + 2] << 8 | data[pos + 3]) == w:
            pos += 4
            for c in range(4):  # each component of the scanline run-length encoded on its own
                x = 0
                while x < w:
                    n = data[pos]; pos += 1
                    if n > 128:
                        out[y, x:x + n - 128, c] = data[pos]; pos += 1
                        x += n - 128
                    else:
                        out[y, x:x + n, c] = np.frombuffer(data, np.uint8, n, pos); pos += n
                        x += n
        else:
            out[y] = np.frombuffer(data, np.uint8, 4 * w, pos).reshape(w, 4); pos += 4 * w
    assert pos == len(data), (pos, len(data))
    return out
