"""USD ingestion (sparkle_b200/host/usd_loader.cpp restating libraries/source/io/scene/USDLoader.cpp) on
the CPU: (1) a hand-written .usda whose expected world positions, material and camera are derived by hand from the
loader's rules, (2) the round trip the reference guarantees (docs/USD.md): a scene written in USDExporter's layout
(sparkle_b200/usd_export.py) loads back to the same closest hits as the scene it was written from."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_lib as O
from sparkle_b200 import _abi as A
from sparkle_b200 import geometry, scenes, usd_export
from sparkle_b200.scene import SceneBuilder, orbit_camera

HERE = os.path.dirname(os.path.abspath(__file__))


def load(path, w=64, h=48):
    lib = A.host_lib()
    s = lib.spk_host_scene_load_usd(str(path).encode(), w, h)
    if not s:
        raise RuntimeError(lib.spk_host_last_error().decode())
    return s, lib.spk_host_scene_desc(s).contents, lib.spk_host_scene_camera(s).contents


def probe(path):
    out = (C.c_uint64 * 9)()
    assert A.host_lib().spk_host_usd_probe(str(path).encode(), out) == 0, A.host_lib().spk_host_last_error().decode()
    keys = ("nodes", "meshes", "skipped", "materials", "textures", "cameras", "dome_lights", "distant_lights", "triangles")
    return dict(zip(keys, [int(v) for v in out]))


def test_hand_written_fixture():
    path = os.path.join(HERE, "golden", "hand_scene.usda")
    st = probe(path)
    assert st == dict(nodes=4, meshes=1, skipped=1, materials=1, textures=0, cameras=1, dome_lights=1, distant_lights=1,
                      triangles=1)
    s, desc, cam = load(path)
    assert desc.num_primitives == 1 and desc.num_meshes == 1 and desc.num_materials == 1
    # world = root * child: p -> Rz(+90) * 2 * p + (1, 0, 0), then + (10, 20, 30)
    #   (0,0,0)   -> (11, 20, 30);  (1,0,0) -> (0,2,0)+(1,0,0)+(10,20,30) = (11, 22, 30);  (0,1,0.5) -> (-2,0,1)+... = (9, 20, 31)
    fwd = np.array(list(desc.primitives[0].fwd), np.float32).reshape(3, 4)
    pts = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0.5]], np.float32)
    world = pts @ fwd[:, :3].T + fwd[:, 3]
    assert np.array_equal(world, np.array([[11, 20, 30], [11, 22, 30], [9, 20, 31]], np.float32))
    inv = np.array(list(desc.primitives[0].inv), np.float32).reshape(3, 4)
    assert np.allclose(world @ inv[:, :3].T + inv[:, 3], pts, atol=1e-6)
    # opacity 0 -> dielectric with eta = ior (USDLoader.cpp:364, 411-415); constants as authored
    m = desc.materials[0]
    assert m.type == A.MAT_DIELECTRIC and np.float32(m.eta) == np.float32(1.6)
    assert list(m.base_color) == [0.5, 0.25, 0.125] and np.float32(m.metallic) == np.float32(0.75)
    assert np.float32(m.roughness) == np.float32(0.3) and m.base_color_tex == A.INVALID_ID
    # missing tangents: (n x major) x n with n = (0,0,1): |n.x| < 1/sqrt(3) -> major = Right -> (1, 0, 0), w = 1
    tang = np.ctypeslib.as_array(desc.meshes[0].tangents, shape=(3, 4))
    assert np.array_equal(tang, np.tile(np.array([1, 0, 0, 1], np.float32), (3, 1)))
    # dome light: colour * intensity * 2^exposure = (0.25, 0.5, 1) * 6
    assert desc.sky.has_sky == 1 and list(desc.sky.color) == [1.5, 3.0, 6.0]
    # camera: node at world (0, -5, 1), identity rotation -> forward +Y, yaw 0, pitch 0, radius = focusDistance = 5,
    # centre = (0, 0, 1); lens 50 mm / 24 mm / f8
    expect = orbit_camera(64, 48, center=(0, 0, 1), radius=5.0, pitch=0.0, yaw=0.0, focal_length=0.05, sensor_height=0.024,
                          aperture=8.0)
    for f in ("position", "front", "up", "right", "max_u", "max_v", "lower_left"):
        assert np.allclose(list(getattr(cam, f)), list(getattr(expect, f)), atol=1e-5), f
    assert np.float32(cam.aperture_radius) == np.float32(expect.aperture_radius)
    A.host_lib().spk_host_scene_free(s)


def test_malformed_files_fail_loudly(tmp_path):
    lib = A.host_lib()
    bad = tmp_path / "bad.usda"
    for text, why in [("def Xform \"root\" {}", "#usda"), ("#usda 1.0\ndef Xform \"root\" {\n def Mesh \"m\" {", "unterminated"),
                      ("#usda 1.0\ndef Xform \"r\" { def Mesh \"m\" { rel material:binding = </nope> normal3f[] normals = [(0,0,1)]*3\n"
                       "point3f[] points = [(0,0,0),(1,0,0),(0,1,0)] int[] faceVertexIndices = [0,1,2] } }", "")]:
        bad.write_text(text)
        assert not lib.spk_host_scene_load_usd(str(bad).encode(), 8, 8)
        assert why in lib.spk_host_last_error().decode()
    assert not lib.spk_host_scene_load_usd(str(tmp_path / "scene.usdc").encode(), 8, 8)
    assert "usda" in lib.spk_host_last_error().decode()


@pytest.mark.parametrize("kind", ["port"] + (["reference"] if O.available("reference") else []))
def test_round_trip_equals_the_source_scene(tmp_path, kind):
    """write TestScene-procedural (textured meshes, glass, metal, spheres as unit-sphere meshes) in the exporter's layout,
    load it back, and trace the same rays through both descriptions with the oracle"""
    orc = O.Oracle(kind)
    sb, camkw = scenes.testscene()
    path = usd_export.write_usda(sb, tmp_path / "scene.usda", camera=camkw)
    st = probe(path)
    assert st["meshes"] == len(sb.prims) and st["skipped"] == 0 and st["cameras"] == 1 and st["dome_lights"] == 1
    assert st["textures"] == 6 and st["materials"] == len({p["material"] for p in sb.prims})
    s, desc, cam = load(path, 96, 64)
    # the source scene with its analytic spheres baked to the unit-sphere mesh, as the exporter does (docs/USD.md)
    src = SceneBuilder()
    src.textures, src.materials = sb.textures, sb.materials
    sphere = None
    for p in sb.prims:
        if p["kind"] == A.PRIM_SPHERE:
            if sphere is None:
                sphere = src.add_mesh(geometry.uv_sphere(32, 1.0, "sphere"))
            r = float(np.sqrt(np.float32(p["radius2"])))
            src.add_mesh_instance(sphere, p["material"], tuple(p["fwd"][:, 3]), (0, 0, 0), (r, r, r))
        else:
            t, e, sc = p["trs"]
            mid = src.add_mesh(sb.meshes[p["mesh"]])
            src.add_mesh_instance(mid, p["material"], t, e, sc)
    src.set_sky(sb.sky["color"])
    sdesc = src.finalize()
    rays = geometry.random_rays(40000, seed=11, extent=14.0)
    rays[:, 2] = np.abs(rays[:, 2]) * 0.5 + 0.05
    a, _ = orc.trace_closest(desc, rays)
    b, _ = orc.trace_closest(sdesc, rays)
    hit = b["prim"] != A.INVALID_ID
    assert hit.mean() > 0.3
    # same primitive order, same faces; t equal up to the text round trip of the matrices (floats print exactly, the
    # inverse is recomputed from the matrix instead of from translation / rotation / scale)
    assert np.array_equal(a["prim"], b["prim"]) or (a["prim"] != b["prim"]).mean() < 1e-4
    same = a["prim"] == b["prim"]
    assert np.array_equal(a["face"][same], b["face"][same])
    # u, v and the world t (recomputed from the x component of the hit point alone, Ray::InverseAt) amplify the
    # last-bit differences of the recomputed inverse matrix (the floor is scaled 50 : 50 : 0.001): equal for nearly all rays,
    # close for the rest
    m = same & hit
    for f, tight, loose in (("u", 1e-5, 5e-3), ("v", 1e-5, 5e-3), ("t", 1e-3, 0.2)):
        err = np.abs(a[f][m] - b[f][m])
        assert np.quantile(err, 0.99) <= tight and err.max() <= loose, (f, float(np.quantile(err, 0.99)), float(err.max()))
    # materials survive: a textured input loses its factor (docs/USD.md), the rest is equal
    for k, p in enumerate(sb.prims):
        m0, m1 = sb.materials[p["material"]], desc.materials[desc.primitives[k].material]
        assert m1.type == m0["type"]
        if m0["base_color_tex"] == A.INVALID_ID:
            assert np.allclose(list(m1.base_color), m0["base_color"], atol=1e-7)
            assert np.float32(m1.metallic) == np.float32(m0["metallic"]) and np.float32(m1.roughness) == np.float32(m0["roughness"])
        else:
            assert m1.base_color_tex != A.INVALID_ID and m1.normal_tex != A.INVALID_ID and m1.metallic_roughness_tex != A.INVALID_ID
    # camera and sky
    expect = orbit_camera(96, 64, **camkw)
    for f in ("position", "front", "max_u", "max_v", "lower_left"):
        assert np.allclose(list(getattr(cam, f)), list(getattr(expect, f)), rtol=1e-5, atol=2e-5), f
    assert np.allclose(list(desc.sky.color), sb.sky["color"], atol=1e-7)
    # images of both descriptions through the CPU path tracer: equal within the tolerance of the image gates
    img_a, _ = orc.render(desc, cam, 96, 64, 0, 2, max_bounce=4)
    img_b, _ = orc.render(sdesc, expect, 96, 64, 0, 2, max_bounce=4)
    rel = np.abs(img_a[..., :3] - img_b[..., :3]).sum(-1) / np.maximum(img_b[..., :3].sum(-1), 1e-3)
    assert np.median(rel) < 1e-4
    A.host_lib().spk_host_scene_free(s)


def test_reader_accepts_the_syntax_variants_of_usda(tmp_path):
    """comments, doc strings, `custom` / `uniform` qualifiers, scientific notation, float3 tangents (w = 1), a node without
    an xformOp (identity), a quad mesh (skipped like a non-single-indexable one), unknown prim types (skipped with a log)"""
    text = '''#usda 1.0
(
    doc = """multi
line"""
    defaultPrim = "World"
)
def Xform "World" (
    kind = "component"
    customData = { string note = "braces { inside } a string" }
)
{
    def Xform "n" {
        custom float userProperties:x = 1.5e0
        def Mesh "tri" (
            prepend apiSchemas = ["MaterialBindingAPI"]
        )
        {
            uniform token subdivisionScheme = "none"
            int[] faceVertexCounts = [3]
            int[] faceVertexIndices = [0, 1, 2]
            rel material:binding = </World/Looks/m>
            normal3f[] normals = [(0, 0, 1), (0, 0, 1), (0, 0, 1)]
            point3f[] points = [(0, 0, 0), (1e0, 0, 0), (0, 2.5E-1, 0)]  # a comment after a value
            float3[] primvars:tangents = [(0, 1, 0), (0, 1, 0), (0, 1, 0)]
        }
        def Mesh "quad" {
            int[] faceVertexCounts = [4]
            int[] faceVertexIndices = [0, 1, 2, 3]
            rel material:binding = </World/Looks/m>
            normal3f[] normals = [(0, 0, 1), (0, 0, 1), (0, 0, 1), (0, 0, 1)]
            point3f[] points = [(0, 0, 0), (1, 0, 0), (1, 1, 0), (0, 1, 0)]
        }
        def SphereLight "unsupported" { float inputs:radius = 1 }
    }
    def Scope "Looks" {
        def Material "m" {
            def Shader "s" {
                uniform token info:id = "UsdPreviewSurface"
                color3f inputs:diffuseColor = (0.25, 0.5, 0.75)
            }
        }
    }
}
'''
    path = tmp_path / "variants.usda"
    path.write_text(text)
    st = probe(path)
    assert st["meshes"] == 1 and st["skipped"] == 1 and st["materials"] == 1 and st["nodes"] == 2
    s, desc, _ = load(path)
    fwd = np.array(list(desc.primitives[0].fwd), np.float32).reshape(3, 4)
    assert np.array_equal(fwd, np.eye(4, dtype=np.float32)[:3])  # no xformOp anywhere: identity
    tang = np.ctypeslib.as_array(desc.meshes[0].tangents, shape=(3, 4))
    assert np.array_equal(tang, np.tile(np.array([0, 1, 0, 1], np.float32), (3, 1)))
    verts = np.ctypeslib.as_array(desc.meshes[0].vertices, shape=(3, 3))
    assert np.array_equal(verts, np.array([[0, 0, 0], [1, 0, 0], [0, 0.25, 0]], np.float32))
    m = desc.materials[0]
    # UsdPreviewSurface fallbacks for inputs that are not authored: metallic 0, roughness 0.5, opaque -> PBR
    assert m.type == A.MAT_PBR and list(m.base_color) == [0.25, 0.5, 0.75]
    assert np.float32(m.metallic) == 0 and np.float32(m.roughness) == np.float32(0.5)
    A.host_lib().spk_host_scene_free(s)


def test_dome_light_with_a_sky_map(tmp_path):
    """DomeLight inputs:texture:file: the .hdr next to the .usda is decoded and cooked into the 6 x 1024^2 RGBAFloat16 cube
    (SkyLight::SetSkyMap -> SkyLightCookJob); the scene description carries the faces (has_map)."""
    import hdr_files as H
    sb, camkw = scenes.testscene(models=False)
    data = H.write_hdr(H.float_to_rgbe(H.make_sky()))
    path = usd_export.write_usda(sb, tmp_path / "sky.usda", camera=camkw, sky_hdr=("sky.hdr", data))
    s, desc, _ = load(path)
    assert desc.sky.has_sky == 1 and desc.sky.has_map == 1
    faces = []
    for f in range(6):
        t = desc.sky.faces[f]
        assert (t.width, t.height, t.format) == (1024, 1024, A.FMT_RGBA16F)
        faces.append(np.ctypeslib.as_array(C.cast(t.pixels, C.POINTER(C.c_uint16)), (1024, 1024, 4)).copy())
    faces = np.stack(faces)
    # the same cube through the cook entry point, and through the reference's own functions where they are built
    lib = A.host_lib()
    w, h, p = A.u32(0), A.u32(0), C.c_void_p()
    assert lib.spk_host_decode_hdr(data, len(data), C.byref(w), C.byref(h), C.byref(p)) == 0
    eq = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint16)), (h.value, w.value, 4)).copy()
    lib.spk_host_free_buffer(p)
    again = np.zeros_like(faces)
    assert lib.spk_host_cook_sky_cube(eq.ctypes.data, w.value, h.value, 1024, again.ctypes.data) == 0
    assert np.array_equal(faces, again)
    if O.available("reference"):
        assert np.array_equal(faces, O.Oracle("reference").kat_cook_sky_cube(eq, 1024))
    lib.spk_host_scene_free(s)
    # a missing or broken file fails the load, loudly
    os.remove(tmp_path / "textures" / "sky.hdr")
    with pytest.raises(RuntimeError, match="sky.hdr"):
        load(path)
    (tmp_path / "textures" / "sky.hdr").write_bytes(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n+Y 4 +X 4\n" + bytes(64))
    with pytest.raises(RuntimeError, match="decode"):
        load(path)
