"""Radiance .hdr (RGBE) files for the sky-map tests: a writer of our own (flat and run-length scanlines) and a seeded
equirectangular test sky. Test infrastructure."""
import numpy as np


def float_to_rgbe(rgb):
    """(h, w, 3) float32 -> (h, w, 4) uint8, the classic Ward encoding"""
    rgb = np.asarray(rgb, np.float32)
    m = rgb.max(axis=-1)
    out = np.zeros(rgb.shape[:2] + (4,), np.uint8)
    ok = m > 1e-32
    mant, exp = np.frexp(m[ok])
    scale = (mant * 256.0 / m[ok])[:, None]
    out[ok, :3] = np.clip(rgb[ok] * scale, 0, 255).astype(np.uint8)
    out[ok, 3] = (exp + 128).astype(np.uint8)
    return out


def _rle_channel(values):
    """one channel of one scanline: runs (count + 128, value) of >= 3 equal bytes, dumps (count, bytes...) otherwise"""
    out, i, n = bytearray(), 0, len(values)
    while i < n:
        run = 1
        while i + run < n and run < 127 and values[i + run] == values[i]:
            run += 1
        if run >= 3:
            out += bytes([128 + run, values[i]])
            i += run
            continue
        j = i
        while j < n and j - i < 128:
            r = 1
            while j + r < n and r < 3 and values[j + r] == values[j]:
                r += 1
            if r >= 3:
                break
            j += 1
        out += bytes([j - i]) + bytes(values[i:j])
        i = j
    return bytes(out)


def write_hdr(rgbe, rle=True, identifier=b"#?RADIANCE", extra_header=(b"EXPOSURE=1.0",)):
    """(h, w, 4) uint8 RGBE -> bytes of a Radiance file (-Y h +X w)"""
    h, w = rgbe.shape[:2]
    head = identifier + b"\n" + b"".join(line + b"\n" for line in extra_header) + b"FORMAT=32-bit_rle_rgbe\n\n"
    head += b"-Y %d +X %d\n" % (h, w)
    body = bytearray()
    if not rle or w < 8 or w >= 32768:
        body += rgbe.tobytes()
    else:
        for j in range(h):
            body += bytes([2, 2, w >> 8, w & 255])
            for k in range(4):
                body += _rle_channel(rgbe[j, :, k].tolist())
    return head + bytes(body)


def make_sky(width=256, height=128, seed=7):
    """a seeded equirectangular sky: a smooth gradient, banded noise, a few very bright texels (a sun) and flat runs (so that
    the run-length encoder emits both kinds of record)"""
    g = np.random.default_rng(seed)
    v = np.linspace(0, 1, height, dtype=np.float32)[:, None, None]
    u = np.linspace(0, 1, width, dtype=np.float32)[None, :, None]
    sky = (0.2 + 0.8 * (1 - v)) * np.array([0.5, 0.7, 1.0], np.float32) + 0.1 * np.sin(12 * u) ** 2
    sky = sky + g.uniform(0, 0.2, (height, width, 3)).astype(np.float32)
    sky[height // 5: height // 5 + 3, width // 3: width // 3 + 3] = (9000.0, 8000.0, 6000.0)
    sky[-height // 8:, :] = (0.05, 0.04, 0.03)  # flat ground: long runs
    sky[0, :8] = 0.0  # black texels: exponent byte 0
    return sky.astype(np.float32)
