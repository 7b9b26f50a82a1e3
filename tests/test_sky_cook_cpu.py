"""The sky map of a SkyLight on its way into the path tracer (data source of a14): Radiance .hdr -> RGBAFloat16
(Image2D::LoadFromFile's HDR branch) -> 6 cube faces (SkyLightCookJob::Execute). The host library's restatement
(sparkle_b200/host/sky_cook.cpp) against the reference's own compiled functions, bit for bit, and against a committed golden
checksum that travels without the reference."""
import ctypes as C
import hashlib

import numpy as np
import pytest

import hdr_files as H
import oracle_lib as O
from sparkle_b200 import _abi as A


def host_decode(data):
    lib = A.host_lib()
    w, h, p = A.u32(0), A.u32(0), C.c_void_p()
    if lib.spk_host_decode_hdr(data, len(data), C.byref(w), C.byref(h), C.byref(p)) != 0:
        return None
    n = w.value * h.value * 4
    out = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint16)), (n,)).reshape(h.value, w.value, 4).copy()
    lib.spk_host_free_buffer(p)
    return out


def host_cook(eq, size):
    eq = np.ascontiguousarray(eq, np.uint16)
    faces = np.zeros((6, size, size, 4), np.uint16)
    assert A.host_lib().spk_host_cook_sky_cube(eq.ctypes.data, eq.shape[1], eq.shape[0], size, faces.ctypes.data) == 0
    return faces


def files():
    rgbe = H.float_to_rgbe(H.make_sky())
    yield "rle", H.write_hdr(rgbe)
    yield "flat", H.write_hdr(rgbe, rle=False)
    yield "rgbe id, narrow (flat by rule)", H.write_hdr(rgbe[:5, :7], identifier=b"#?RGBE")
    odd = H.float_to_rgbe(H.make_sky(129, 33, seed=9))
    yield "odd size rle", H.write_hdr(odd, extra_header=(b"# comment", b"GAMMA=1"))


def test_half_conversion_matches_numpy():
    """FloatToHalf / HalfToFloat (Eigen's round-to-nearest-even) through the decoder: every exponent byte, every mantissa"""
    rgbe = np.zeros((256, 256, 4), np.uint8)
    rgbe[..., 0] = np.arange(256)[None, :]
    rgbe[..., 1] = 255 - np.arange(256)[None, :]
    rgbe[..., 2] = 128
    rgbe[..., 3] = np.arange(256)[:, None]
    got = host_decode(H.write_hdr(rgbe, rle=False))
    f = np.ldexp(np.float32(1.0), rgbe[..., 3].astype(np.int32) - 136).astype(np.float32)[..., None] * rgbe[..., :3].astype(np.float32)
    f[rgbe[..., 3] == 0] = 0.0
    want = np.clip(f, -65504.0, 65504.0).astype(np.float16).view(np.uint16)
    assert np.array_equal(got[..., :3], want)
    assert np.all(got[..., 3] == np.float16(1.0).view(np.uint16))


@pytest.mark.skipif(not O.available("reference"), reason="oracle/_ref not built")
def test_decode_equals_reference():
    ref = O.Oracle("reference")
    for name, data in files():
        want = ref.kat_decode_hdr(data)
        got = host_decode(data)
        assert want is not None and got is not None, name
        assert got.shape == want.shape and np.array_equal(got, want), name
    # what stb_image refuses, the host decoder refuses
    good = H.write_hdr(H.float_to_rgbe(H.make_sky(64, 16)))
    for name, data in [("no identifier", b"#?RADIANT\n" + good[11:]), ("no format line", good.replace(b"FORMAT=32-bit_rle_rgbe\n", b"")),
                       ("other layout", good.replace(b"-Y 16 +X 64", b"+Y 16 +X 64")),
                       ("wrong scanline length", good.replace(bytes([2, 2, 0, 64]), bytes([2, 2, 0, 63]), 1)),
                       ("decoded size beyond an int", good.replace(b"-Y 16 +X 64", b"-Y 16000 +X 16000"))]:
        assert ref.kat_decode_hdr(data) is None, name
        assert host_decode(data) is None, name
    # (a flat file of width >= 8, the "flat" case above, is read through stb_image's mid-file fallback: its first scanline
    # does not start a run-length record)


@pytest.mark.skipif(not O.available("reference"), reason="oracle/_ref not built")
def test_cook_equals_reference():
    ref = O.Oracle("reference")
    eq = host_decode(H.write_hdr(H.float_to_rgbe(H.make_sky())))
    for size in (64, 257):
        want = ref.kat_cook_sky_cube(eq, size)
        got = host_cook(eq, size)
        assert np.array_equal(got, want), (size, int((got != want).sum()))
    # the faces carry the sky: bright sun texels survive, the poles are seen by the +-Z faces
    f = got.view(np.float16).astype(np.float32)
    assert f[..., :3].max() > 1000.0 and f[5, ..., :3].mean() < f[4, ..., :3].mean()


def test_cook_golden_checksum():
    """the pin that travels: SHA-1 of the 6 x 128 x 128 faces cooked from the seeded test sky, written when the test above
    was green against the reference (tests/golden/sky_cook.sha1)"""
    import os
    eq = host_decode(H.write_hdr(H.float_to_rgbe(H.make_sky())))
    faces = host_cook(eq, 128)
    digest = hashlib.sha1(eq.tobytes() + faces.tobytes()).hexdigest()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sky_cook.sha1")
    if os.environ.get("SPK_WRITE_GOLDEN") == "1":
        ref = O.Oracle("reference")
        assert np.array_equal(ref.kat_cook_sky_cube(eq, 128), faces)
        with open(path, "w") as fh:
            fh.write(digest + "\n")
    assert open(path).read().strip() == digest
