"""Writes a SceneBuilder scene as ASCII USD (.usda) in the layout of the reference's exporter
(libraries/source/io/scene/USDExporter.cpp; docs/USD.md): what the packaged default scene resources/packed/TestScene.usda
is an export of, consumed here by sparkle_b200/host/usd_loader.cpp.

  stage         Z-up, metersPerUnit 1, defaultPrim = the root Xform (USDExporter.cpp:571-575)
  node          one `Xform` per primitive with a single `xformOp:transform` matrix4d, written row-vector style
                (the transpose of Transform::GetMatrix, USDExporter.cpp:53-67)
  mesh          child `Mesh`: points, vertex normals, faceVertexCounts = 3.., faceVertexIndices, primvars:st,
                primvars:tangents (float4), `rel material:binding` (:286-390)
  sphere        SpherePrimitive exports its tessellated unit sphere (docs/USD.md): written as the unit-sphere mesh under
                an Xform with translation = centre and scale = radius, so it reloads as a MESH primitive
  material      `Material` under the root scope /_materials with st_reader, UsdUVTexture shaders and a UsdPreviewSurface
                (:172-284): constants or texture connections; a textured input loses its constant factor; dielectric =
                opacity 0 + ior = eta; metallic / roughness read .b / .g of one packed image
  textures      8-bit PNG files under textures/ next to the .usda (:95-118)
  camera        `Camera` under an Xform holding the orbit camera's node transform: translation = position, rotation =
                (pitch, 0, yaw) as the world-to-camera euler angles (OrbitCameraComponent.cpp:16-30); lens attributes in
                millimetres, focusDistance = orbit radius (:392-407)
  sky           `DomeLight` with inputs:color (procedural sky, :424-433)"""
import math
import os

import numpy as np

from . import _abi as A
from .geometry import trs, uv_sphere
from .gltf_export import write_png


def _fmt(x):
    return repr(float(np.float32(x)))  # shortest text that reads back to the same float32


def _tuples(a):
    a = np.asarray(a)
    return "[" + ", ".join("(" + ", ".join(_fmt(v) for v in row) + ")" for row in a) + "]"


def _matrix(fwd):
    """3x4 rows of Transform::GetMatrix -> matrix4d text, transposed (row-vector convention)"""
    m = np.eye(4, dtype=np.float64)
    m[:3, :] = np.asarray(fwd, np.float64).reshape(3, 4)
    rows = m.T
    return "( " + ", ".join("(" + ", ".join(_fmt(v) for v in r) + ")" for r in rows) + " )"


def _valid(name):
    out = "".join(c if (c.isalnum() or c == "_") else "_" for c in name) or "prim"
    return "_" + out if out[0].isdigit() else out


def write_usda(sb, path, camera=None, width=None, height=None, sky_hdr=None):
    """sb: SceneBuilder (before or after finalize). camera: dict(center, radius, pitch, yaw[, focal_length,
    sensor_height, aperture]) or None. sky_hdr: (file name, bytes of an equirectangular Radiance .hdr) written next to the
    textures and referenced by the dome light like USDExporter.cpp:449-463 does for a SkyLight with a sky map. Returns the
    .usda path."""
    path = str(path)
    out_dir = os.path.dirname(path)
    os.makedirs(os.path.join(out_dir, "textures"), exist_ok=True)
    tex_files = {}

    def texture_file(tid):
        if tid not in tex_files:
            px, fmt = sb.textures[tid]
            if fmt not in (A.FMT_RGBA8_SRGB, A.FMT_RGBA8_UNORM):
                raise ValueError("write_usda: 8-bit textures only (float images would be .hdr)")
            rel = f"textures/{len(tex_files)}_tex{tid}.png"
            write_png(os.path.join(out_dir, rel), px)
            tex_files[tid] = (rel, fmt == A.FMT_RGBA8_SRGB)
        return tex_files[tid]

    L = ["#usda 1.0", "(", '    defaultPrim = "root"', "    metersPerUnit = 1", '    upAxis = "Z"', ")", ""]
    L += ['def Xform "root" (', '    displayName = "root"', ")", "{",
          "    matrix4d xformOp:transform = " + _matrix(np.eye(4)[:3]),
          '    uniform token[] xformOpOrder = ["xformOp:transform"]', ""]
    used_materials = {}

    def material_path(mid):
        if mid not in used_materials:
            used_materials[mid] = f"material_{len(used_materials)}"
        return "/_materials/" + used_materials[mid]

    unit_sphere = None
    n_mesh = 0
    for k, p in enumerate(sb.prims):
        if p["kind"] == A.PRIM_SPHERE:
            if unit_sphere is None:
                unit_sphere = uv_sphere(32, 1.0, "sphere")  # Mesh::GetUnitSphere
            mesh, name = unit_sphere, "sphere"
        else:
            mesh, name = sb.meshes[p["mesh"]], sb.meshes[p["mesh"]].name or "mesh"
        L += [f'    def Xform "node_{k}" (', f'        displayName = "node {k}"', "    )", "    {",
              "        matrix4d xformOp:transform = " + _matrix(p["fwd"]),
              '        uniform token[] xformOpOrder = ["xformOp:transform"]', "",
              f'        def Mesh "{_valid(name)}_{n_mesh}"', "        {",
              "            int[] faceVertexCounts = [" + ", ".join(["3"] * len(mesh.indices)) + "]",
              "            int[] faceVertexIndices = [" + ", ".join(str(int(i)) for i in mesh.indices.reshape(-1)) + "]",
              f"            rel material:binding = <{material_path(p['material'])}>",
              "            normal3f[] normals = " + _tuples(mesh.normals) + " (",
              '                interpolation = "vertex"', "            )",
              "            point3f[] points = " + _tuples(mesh.vertices)]
        if mesh.uvs is not None:
            L += ["            texCoord2f[] primvars:st = " + _tuples(mesh.uvs) + " (",
                  '                interpolation = "vertex"', "            )"]
        L += ["            float4[] primvars:tangents = " + _tuples(mesh.tangents) + " (",
              '                interpolation = "vertex"', "            )", "        }", "    }", ""]
        n_mesh += 1
    if camera is not None:
        pitch, yaw = float(camera.get("pitch", 10.0)), float(camera.get("yaw", -20.0))
        radius = float(camera.get("radius", 25.0))
        center = np.asarray(camera.get("center", (0, 0, 0)), np.float64)
        # OrbitCameraComponent::UpdateTransform: direction = Rz(-yaw) * Rx(-pitch) * Front, position = centre - dir * r,
        # node rotation = +(pitch, 0, yaw)
        rp, ry = math.radians(-pitch), math.radians(-yaw)
        d = np.array([-math.sin(ry) * math.cos(rp), math.cos(ry) * math.cos(rp), math.sin(rp)])
        pos = center - d * radius
        fwd, _ = trs(tuple(pos), (math.radians(pitch), 0.0, math.radians(yaw)), (1, 1, 1))
        L += ['    def Xform "camera_node"', "    {", "        matrix4d xformOp:transform = " + _matrix(fwd),
              '        uniform token[] xformOpOrder = ["xformOp:transform"]', "", '        def Camera "camera"', "        {",
              f"            float focalLength = {_fmt(camera.get('focal_length', 0.035) * 1000.0)}",
              f"            float focusDistance = {_fmt(radius)}", f"            float fStop = {_fmt(camera.get('aperture', 22.0))}",
              '            token projection = "perspective"',
              f"            float verticalAperture = {_fmt(camera.get('sensor_height', 0.024) * 1000.0)}", "        }", "    }", ""]
    if sb.sky.get("has_sky") or sky_hdr:
        c = sb.sky.get("color", (1.0, 1.0, 1.0))
        dome = [f"            color3f inputs:color = ({_fmt(c[0])}, {_fmt(c[1])}, {_fmt(c[2])})",
                "            float inputs:intensity = 1"]
        if sky_hdr:
            name, data = sky_hdr
            with open(os.path.join(out_dir, "textures", name), "wb") as fh:
                fh.write(data)
            dome.append(f"            asset inputs:texture:file = @textures/{name}@")
        L += ['    def Xform "sky_node"', "    {", "        matrix4d xformOp:transform = " + _matrix(np.eye(4)[:3]),
              '        uniform token[] xformOpOrder = ["xformOp:transform"]', "", '        def DomeLight "sky"', "        {"]
        L += dome + ["        }", "    }", ""]
    L += ["}", "", 'def Scope "_materials"', "{"]
    for mid, name in used_materials.items():
        m = sb.materials[mid]
        mp = "/_materials/" + name
        L += [f'    def Material "{name}" (', f'        displayName = "material {mid}"', "    )", "    {",
              f"        token outputs:surface.connect = <{mp}/surface.outputs:surface>", "",
              '        def Shader "st_reader"', "        {", '            uniform token info:id = "UsdPrimvarReader_float2"',
              '            string inputs:varname = "st"', "            float2 outputs:result", "        }", ""]

        def tex_shader(shader, tid):
            rel, srgb = texture_file(tid)
            return [f'        def Shader "{shader}"', "        {", '            uniform token info:id = "UsdUVTexture"',
                    f"            asset inputs:file = @{rel}@",
                    f'            token inputs:sourceColorSpace = "{"sRGB" if srgb else "raw"}"',
                    f"            float2 inputs:st.connect = <{mp}/st_reader.outputs:result>",
                    '            token inputs:wrapS = "repeat"', '            token inputs:wrapT = "repeat"',
                    "            float3 outputs:rgb", "        }", ""]

        S = ['        def Shader "surface"', "        {", '            uniform token info:id = "UsdPreviewSurface"']
        if m["base_color_tex"] != A.INVALID_ID:
            L += tex_shader("base_color_tex", m["base_color_tex"])
            S.append(f"            color3f inputs:diffuseColor.connect = <{mp}/base_color_tex.outputs:rgb>")
        else:
            S.append("            color3f inputs:diffuseColor = (" + ", ".join(_fmt(v) for v in m["base_color"]) + ")")
        if m["emissive_tex"] != A.INVALID_ID:
            L += tex_shader("emissive_tex", m["emissive_tex"])
            S.append(f"            color3f inputs:emissiveColor.connect = <{mp}/emissive_tex.outputs:rgb>")
        else:
            S.append("            color3f inputs:emissiveColor = (" + ", ".join(_fmt(v) for v in m["emissive"]) + ")")
        if m["normal_tex"] != A.INVALID_ID:
            L += tex_shader("normal_tex", m["normal_tex"])
            S.append(f"            normal3f inputs:normal.connect = <{mp}/normal_tex.outputs:rgb>")
        if m["metallic_roughness_tex"] != A.INVALID_ID:
            L += tex_shader("metallic_roughness_tex", m["metallic_roughness_tex"])
            S.append(f"            float inputs:metallic.connect = <{mp}/metallic_roughness_tex.outputs:b>")
            S.append(f"            float inputs:roughness.connect = <{mp}/metallic_roughness_tex.outputs:g>")
        else:
            S.append(f"            float inputs:metallic = {_fmt(m['metallic'])}")
            S.append(f"            float inputs:roughness = {_fmt(m['roughness'])}")
        if m["type"] == A.MAT_DIELECTRIC:
            S.append("            float inputs:opacity = 0")
            S.append(f"            float inputs:ior = {_fmt(m['eta'])}")
        S += ["            int inputs:useSpecularWorkflow = 0", "            token outputs:surface", "        }"]
        L += S + ["    }", ""]
    L += ["}", ""]
    with open(path, "w") as f:
        f.write("\n".join(L))
    return path
