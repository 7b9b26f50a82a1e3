"""TEST INFRASTRUCTURE: per-pixel parity report for image gates (the reference's
docs/RenderingValidation.md: "per-pixel signed diff vs ground truth, never aggregate-only").

report(name, got, ref) returns the aggregate metrics AND the worst 8x8 tile (mean relative error of the tile, its
position), writes a signed-difference image (PNG: red = got brighter, blue = got darker, scaled to the worst pixel) and
appends one JSON line to <out>/report.jsonl. <out> is $SPK_PARITY_OUT, else gpurun_out/parity under the repo root, so
the files come back from the GPU box; the ones quoted in DESIGN.md are copied to profiles/."""
import json
import os
import struct
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TILE = 8


def out_dir():
    d = os.environ.get("SPK_PARITY_OUT", os.path.join(ROOT, "gpurun_out", "parity"))
    os.makedirs(d, exist_ok=True)
    return d


def write_png(path, rgb8):
    h, w, _ = rgb8.shape
    raw = b"".join(b"\x00" + rgb8[j].tobytes() for j in range(h))

    def chunk(tag, data):
        c = struct.pack(">I", len(data)) + tag + data
        return c + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) +
                chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def metrics(got, ref):
    g, r = got[..., :3].astype(np.float64), ref[..., :3].astype(np.float64)
    rel = np.abs(g - r).sum(-1) / np.maximum(r.sum(-1), 1e-3)
    mse = np.mean((g - r) ** 2)
    peak = max(r.max(), 1e-6)
    psnr = float("inf") if mse == 0 else float(10 * np.log10(peak * peak / mse))
    h, w = rel.shape
    th, tw = h // TILE, w // TILE
    worst = dict(worst_tile_mean_rel=0.0, worst_tile_xy=(0, 0))
    if th and tw:
        tiles = rel[:th * TILE, :tw * TILE].reshape(th, TILE, tw, TILE).mean(axis=(1, 3))
        k = int(np.argmax(tiles))
        worst = dict(worst_tile_mean_rel=float(tiles.flat[k]), worst_tile_xy=(int(k % tw) * TILE, int(k // tw) * TILE))
    signed = (g - r).sum(-1)
    lit = r.sum(-1) >= 1e-3  # pixels whose reference value is not black: the only ones a RELATIVE error is defined on
    extra = dict(rel_l1=float(np.abs(g - r).sum() / max(r.sum(), 1e-12)), frac_lit=float(lit.mean()),
                 mean_rel_lit=float(rel[lit].mean()) if lit.any() else 0.0)
    return dict(mean_rel=float(rel.mean()), **extra, max_rel=float(rel.max()), frac_gt_1e4=float((rel > 1e-4).mean()),
                frac_identical=float((got[..., :3] == ref[..., :3]).all(-1).mean()), psnr=psnr,
                mean_signed=float(signed.mean()), **worst), signed


def report(name, got, ref, extra=None):
    m, signed = metrics(got, ref)
    scale = float(np.abs(signed).max())
    img = np.full(signed.shape + (3,), 255, np.uint8)
    if scale > 0:
        a = np.clip(np.abs(signed) / scale, 0, 1)
        fade = (255 * (1 - a)).astype(np.uint8)
        pos = signed > 0
        img[..., 1] = fade
        img[..., 2] = np.where(pos, fade, 255)
        img[..., 0] = np.where(pos, 255, fade)
    d = out_dir()
    write_png(os.path.join(d, f"{name}_signed_diff.png"), img)
    line = dict(name=name, shape=list(got.shape[:2]), signed_diff_scale=scale, **m, **(extra or {}))
    with open(os.path.join(d, "report.jsonl"), "a") as f:
        f.write(json.dumps(line) + "\n")
    print("parity", json.dumps(line))
    return m
